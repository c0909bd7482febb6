#!/usr/bin/env python3
"""Development timing of the ray store + batch assembly at BASELINE configs[3] scale: 2^18-ray batches
from V views of 800 x 800 (S=4 x H_max=32), batch_assemble alone (kernel by kernel share via the
achieved bandwidth) and the training step fed from the store instead of a per-step march."""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, pipeline, synth  # noqa: E402


def timed(fn, reps, flush=None):
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.add_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(np.min(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--rays", type=int, default=1 << 18)
    ap.add_argument("--S", type=int, default=4)
    ap.add_argument("--hmax", type=int, default=32)
    ap.add_argument("--group", type=int, default=32)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    G, W, H = 128, 800, 800
    occ = synth.occupancy_lego(G)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    n = a.views * W * H
    gt = torch.rand((n, 3), device="cuda")
    store = api.RayStore(a.views, W * H, n * a.hmax, pixels_gt=gt)
    poses = synth.poses(a.views, seed=5)
    t0 = time.perf_counter()
    for v in range(a.views):
        store.add_view(grid, torch.as_tensor(poses[v].reshape(16)).cuda(), synth.focal_length(), 1.0, W, H, a.hmax)
    torch.cuda.synchronize()
    print("store: %d views, %d rays, %d hits (%.2f/ray), filled in %.1f ms, %.2f GB resident"
          % (a.views, n, store.used_hits, store.used_hits / n, (time.perf_counter() - t0) * 1e3,
             (store.used_hits * 24 + n * 28) / 1e9))
    hits = api.HitBuffers(a.rays, a.rays * a.hmax, want_t=False, want_cells=False, want_origins=False)
    tgt = torch.empty((a.rays, 3), device="cuda")
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 1.5:
        api.batch_assemble(store, 1, a.group, 0, a.rays, hits, gt_pixels=tgt)
        torch.cuda.synchronize()
    for group in sorted({1, a.group}):
        k = [0]

        def one():
            k[0] += 1
            api.batch_assemble(store, 1, group, (k[0] * a.rays) % (n - a.rays + 1) // 32 * 32, a.rays, hits, gt_pixels=tgt)
        med, best = timed(one, a.reps, flush)
        T = int(hits.total_hits.item())
        # algorithmic bytes: per ray 4 (num_hits) + 4 (index) + 8 (dir) + 12 (gt) read, 4 + 4 + 8 + 12 written;
        # per hit 24 read + 24 written
        bytes_ = a.rays * 56 + T * 48
        print("batch_assemble group=%-2d: %d rays, %d hits: median %.3f ms (best %.3f) = %.0f GB/s of algorithmic bytes"
              % (group, a.rays, T, med, best, bytes_ / med / 1e6))
    med, best = timed(lambda: api.ray_permutation(n, 3, a.group), 5)
    print("ray_permutation of all %d rays: %.3f ms" % (n, med))

    lv = synth.hashgrid_levels()
    net = api.Network(lv, synth.hashgrid_table(lv, seed=1), synth.mlp_weights(seed=2))
    tr = pipeline.Trainer(grid, net, W, H, synth.focal_length(), 1.0, a.S, a.hmax, a.rays)
    k = [0]

    def step():
        k[0] += 1
        tr.step_from_store(store, 1, (k[0] * a.rays) % (n - a.rays + 1) // 32 * 32, a.group)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    for group in sorted({1, a.group}):
        a.group = group
        med, best = timed(step, a.reps)
        Nsmp = int(tr.view.hits.total_hits.item()) * a.S
        print("training step from the store (group=%d): median %.3f ms (best %.3f) -> %.2f Mrays/s, %.2f Gsamples/s, loss %.5f"
              % (group, med, best, a.rays / med / 1e3, Nsmp / med / 1e6, float(tr.loss_sum.item())))


if __name__ == "__main__":
    main()
