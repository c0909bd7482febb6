#!/usr/bin/env python3
"""Development timing of the MLP alone (pre-encoded fp16 input) and of the encode-only kernel:
separates the tensor-core part of the network stage from the hash-grid gathers."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, synth  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 40_960_000
    lv = synth.hashgrid_levels()
    net = api.Network(lv, synth.hashgrid_table(lv, seed=1), synth.mlp_weights(seed=2))
    x = (torch.rand((n, 32), device="cuda") - 0.5).half()
    rad = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    out = torch.empty((n, 16), dtype=torch.float16, device="cuda")
    ms = timeit(lambda: net.forward(None, x, radiance=rad, pre_encoded=True))
    print("MLP only, radiance out : %.3f ms  %.2f Gsamples/s  %.1f TFLOP/s  %.0f GB/s (80 B/sample)" %
          (ms, n / ms / 1e6, 14336 * n / ms / 1e9, 80 * n / ms / 1e6))
    ms = timeit(lambda: net.forward(None, x, output=out, pre_encoded=True))
    print("MLP only, out16        : %.3f ms  %.2f Gsamples/s  %.1f TFLOP/s  %.0f GB/s (96 B/sample)" %
          (ms, n / ms / 1e6, 14336 * n / ms / 1e9, 96 * n / ms / 1e6))


if __name__ == "__main__":
    main()
