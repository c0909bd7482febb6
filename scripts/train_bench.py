#!/usr/bin/env python3
"""Development timing of the training step at BASELINE configs[3] scale (2^18 rays, S=4 x H_max=32)."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, pipeline, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=1 << 18)
    ap.add_argument("--S", type=int, default=4)
    ap.add_argument("--hmax", type=int, default=32)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    G, W, H = 128, 800, 800
    occ = synth.occupancy_lego(G)
    lv = synth.hashgrid_levels()
    net = api.Network(lv, synth.hashgrid_table(lv, seed=1), synth.mlp_weights(seed=2))
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    tr = pipeline.Trainer(grid, net, W, H, synth.focal_length(), 1.0, a.S, a.hmax, a.rays)
    poses = synth.poses(a.steps + 2, seed=5)
    target = torch.rand((a.rays, 3), device="cuda")
    names = ["march", "sampler", "fwd", "composite", "loss", "comp_bwd", "net_bwd", "adam"]
    v, h = tr.view, tr.view.hits

    def net_bwd():
        tr.net.mlp_backward(None, tr.encoded, tr.radiance_grad, tr.S, h.total_hits, v.N_cap, tr.dencoded,
                            tr.grads[tr.n_table:])
        tr._scatter()

    fns = [v.stage_march_fused if tr.from_hits else v.stage_march, tr._sample, v.stage_radiance, v.stage_composite,
           lambda: api.l2_loss(v.pixels, tr.target, tr.loss_scale, tr.loss_grad, loss_sum=tr.loss_sum),
           lambda: api.launch_volrender_backward_cuda(None, tr.loss_grad, v.radiance, v.t_vals, h.num_hits, h.indices,
                                                      tr.R, tr.S, tr.radiance_grad),
           net_bwd, tr.optimizer_step]
    tot = np.zeros(len(fns))
    import time
    t0 = time.perf_counter()
    v.pose_dev.copy_(torch.as_tensor(poses[0].reshape(16)).cuda())
    v.ray_begin = 0
    while time.perf_counter() - t0 < 1.5:      # clock spin-up: a B200 out of idle needs ~1 s of load
        tr.target.copy_(target)
        for f in fns:
            f()
        torch.cuda.synchronize()
    for k in range(a.steps + 2):
        v.pose_dev.copy_(torch.as_tensor(poses[k].reshape(16)).cuda())
        v.ray_begin = (k * 7919) % (W * H - a.rays)
        tr.target.copy_(target)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(fns) + 1)]
        ev[0].record()
        for i, f in enumerate(fns):
            f()
            ev[i + 1].record()
        torch.cuda.synchronize()
        if k >= 2:
            tot += [ev[i].elapsed_time(ev[i + 1]) for i in range(len(fns))]
    tot /= a.steps
    n = int(h.total_hits.item()) * a.S
    print("rays %d samples %d (%.1f/ray)" % (a.rays, n, n / a.rays))
    for nm, t in zip(names, tot):
        print("%-10s %8.3f ms" % (nm, t))
    print("step total %.3f ms -> %.2f Mrays/s, %.2f Gsamples/s; loss %.5f" %
          (tot.sum(), a.rays / tot.sum() / 1e3, n / tot.sum() / 1e6, float(tr.loss_sum.item())))


if __name__ == "__main__":
    main()
