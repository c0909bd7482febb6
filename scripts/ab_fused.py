#!/usr/bin/env python3
"""Development A/B of the render kernels at BASELINE config-2 scale: times the march and the fused
sampler+network+compositing kernel, the staged network forward, and writes the pixels so that two
runs (e.g. RTXNERF_FUSED_V1=1 against the default) can be compared.  Not the contract bench."""
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
    ap.add_argument("--out", default="")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    occ = synth.occupancy_lego(a.G)
    grid = api.OccupancyGrid.from_occupancy(occ, a.G)
    lv = synth.hashgrid_levels()
    net = api.Network(lv, synth.hashgrid_table(lv, seed=3), synth.mlp_weights(seed=4))
    r = pipeline.ViewRenderer(grid, a.W, a.H, synth.focal_length(), a.W / a.H, a.S, a.hmax,
                              pipeline.NetworkRadiance(net))
    pose = synth.poses(4, seed=1)[0]
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 1.5:
        r.render_host(pose)
    N = r.total_hits() * a.S
    print("%s R=%d samples=%d" % (a.tag, r.R, N))
    for name, fn in (("march(records)", r.stage_march_fused), ("march", r.stage_march), ("fused", r.stage_fused),
                     ("sampler", r.stage_sampler),
                     ("network(staged)", r.stage_radiance), ("composite", r.stage_composite)):
        med, best = timeit(fn)
        print("%-18s median %.3f ms  best %.3f ms   %.2f Gsamples/s" % (name, med, best, N / med / 1e6))
    r.stage_fused()
    torch.cuda.synchronize()
    fused = r.pixels.cpu().numpy().copy()
    r.stage_sampler(); r.stage_radiance(); r.stage_composite()
    torch.cuda.synchronize()
    staged = r.pixels.cpu().numpy().copy()
    print("fused vs staged: max abs diff %.3g   mean pixel %.5f" % (np.abs(fused - staged).max(), fused.mean()))
    if a.out:
        np.save(a.out, fused)


if __name__ == "__main__":
    main()
