#!/usr/bin/env python3
"""Development timing of each stage at BASELINE config-2 scale (not the contract bench)."""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, pipeline, synth  # noqa: E402


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
    return float(np.median(ts)), float(np.min(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--G", type=int, default=128)
    ap.add_argument("--W", type=int, default=800)
    ap.add_argument("--H", type=int, default=800)
    ap.add_argument("--S", type=int, default=4)
    ap.add_argument("--hmax", type=int, default=16)
    ap.add_argument("--occ", default="lego")
    a = ap.parse_args()
    occ = synth.occupancy_lego(a.G) if a.occ == "lego" else synth.occupancy_bernoulli(a.G, float(a.occ), 1)
    grid = api.OccupancyGrid.from_occupancy(occ, a.G)
    r = pipeline.ViewRenderer(grid, a.W, a.H, synth.focal_length(), a.W / a.H, a.S, a.hmax,
                              pipeline.CellRadiance(synth.cell_radiance(a.G)))
    pose = synth.poses(4, seed=1)[0]
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 1.5:      # clock spin-up: a B200 out of idle needs ~1 s of load
        r.render_host(pose)
    Hh = r.total_hits()
    R, N = r.R, Hh * a.S
    nh = r.hits.num_hits.cpu().numpy()
    print("occupancy %.3f  R=%d  hits=%d (%.2f/ray, %d rays hit, %d capped)  samples=%d" %
          (occ.mean(), R, Hh, Hh / R, (nh > 0).sum(), (nh >= a.hmax).sum(), N))
    peak = 6548.2
    for name, fn, byts in (
            ("march(count+scan+fill)", r.stage_march, 28 * R + 32 * Hh + a.G ** 3 / 8 + 64),
            ("sampler", r.stage_sampler, 16 * R + 24 * Hh + 24 * N),
            ("cell_radiance", r.stage_radiance, 4 * Hh + 16 * N),
            ("composite fwd", r.stage_composite, 20 * N + 8 * R + 12 * R),
            ("whole view (device)", r.render_device, 0)):
        med, best = timeit(fn)
        gbs = byts / (med * 1e-3) / 1e9
        print("%-24s median %.3f ms  best %.3f ms  %8.1f GB/s (%.1f%% of %.0f)  %.1f Mrays/s" %
              (name, med, best, gbs, 100 * gbs / peak, peak, R / med / 1e3))
    t0 = time.perf_counter()
    for _ in range(5):
        r.render_host(pose)
    print("render_host e2e: %.3f ms/view" % ((time.perf_counter() - t0) / 5 * 1e3))


if __name__ == "__main__":
    main()
