#!/usr/bin/env python3
"""Development timing of the reference's literal network (Frequency + 8x128, network_freq.cu) at the
sample count of BASELINE configs[1] (one 800x800 view x 64 samples/ray = 40.96 M rows)."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, synth  # noqa: E402

FLOP_PER_ROW = 2 * (112 * 128 + 7 * 128 * 128 + 128 * 16)   # 262 144


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=800 * 800 * 64)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out16", action="store_true", help="also write the padded fp16 [N][16] output")
    a = ap.parse_args()
    net = api.FrequencyNetwork(synth.mlp_weights(112, 128, 8, 16, seed=0))
    g = torch.Generator(device="cuda").manual_seed(1)
    s = torch.rand((a.rows, 5), device="cuda", generator=g) * 2 - 1
    rad = torch.empty((a.rows, 4), dtype=torch.float32, device="cuda")
    o16 = torch.empty((a.rows, 16), dtype=torch.float16, device="cuda") if a.out16 else None
    import time
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 1.5:      # clock spin-up: a B200 out of idle needs ~1 s of load
        for _ in range(8):
            net.forward(None, s, output=o16, radiance=rad)
        torch.cuda.synchronize()
    ts = []
    for _ in range(a.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        net.forward(None, s, output=o16, radiance=rad)
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                        "MEASURED_PEAKS.json"))) if os.path.exists("MEASURED_PEAKS.json") else {}
    tf = FLOP_PER_ROW * a.rows / (ms * 1e-3) / 1e12
    pk = peaks.get("bf16_tflops", 1590.0)
    print(json.dumps(dict(kernel="freqnet_forward", rows=a.rows, ms_median=ms, ms_best=float(np.min(ts)),
                          tflops=tf, peak_tflops=pk, frac=tf / pk, samples_per_s=a.rows / (ms * 1e-3),
                          hbm_bytes_per_row=20 + 16 + (32 if a.out16 else 0))))


if __name__ == "__main__":
    main()
