#!/usr/bin/env python3
"""Development timing of the literal network's weight-gradient backward (freqnet_backward_kernel, three launches)
on random rows; prints the gradient checksum so that variants can be compared."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1 << 22)
    ap.add_argument("--iters", type=int, default=3)
    a = ap.parse_args()
    net = api.FrequencyNetwork(synth.mlp_weights(112, 128, 8, 16, seed=0))
    g = torch.Generator(device="cuda").manual_seed(1)
    s = torch.rand((a.rows, 5), device="cuda", generator=g) * 2 - 1
    dl = ((torch.rand((a.rows, 4), device="cuda", generator=g) - 0.5) * 1e-2).to(torch.float16)
    grads = torch.zeros(net.n_params(), dtype=torch.float32, device="cuda")
    net.backward(None, s, dl, grads, n=a.rows)
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.iters):
        grads.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        net.backward(None, s, dl, grads, n=a.rows)
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    tf = 51 * 2.0 * 128 ** 3 * (a.rows / 128) / (ms * 1e-3) / 1e12
    print("freqnet backward: %d rows, %.2f ms (%.1f ms at 33.5 M rows), %.0f TFLOP/s executed; |grad| sum %.6e, max %.4e"
          % (a.rows, ms, ms * 33.5e6 / a.rows, tf, float(grads.abs().sum()), float(grads.abs().max())))


if __name__ == "__main__":
    main()
