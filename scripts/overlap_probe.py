#!/usr/bin/env python3
"""Timing probe: can mlp_backward (tensor / shared-memory bound, 8 warps per SM) and the hash-grid scatter (LSU /
L2-reduction bound) of the training step run CONCURRENTLY on one GPU?  Launches both over the full batch on two
streams (the data dependence is ignored: timing only) and compares with the serial sum."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, pipeline, synth  # noqa: E402


def main():
    G, W, H, R, S, HM = 128, 800, 800, 1 << 18, 4, 32
    lv = synth.hashgrid_levels()
    net = api.Network(lv, synth.hashgrid_table(lv, seed=1), synth.mlp_weights(seed=2))
    grid = api.OccupancyGrid.from_occupancy(synth.occupancy_lego(G), G)
    tr = pipeline.Trainer(grid, net, W, H, synth.focal_length(), 1.0, S, HM, R)
    pose = torch.as_tensor(synth.poses(1, seed=5)[0].reshape(16)).cuda()
    tgt = torch.rand((R, 3), device="cuda")
    for _ in range(3):
        tr.step(pose, 1000, tgt)
    torch.cuda.synchronize()
    v, h = tr.view, tr.view.hits
    hi = torch.cuda.Stream(priority=-1)
    lo = torch.cuda.Stream(priority=0)

    def mlp(stream):
        tr.net.mlp_backward(stream, tr.encoded, tr.radiance_grad, tr.S, h.total_hits, v.N_cap, tr.dencoded,
                            tr.grads[tr.n_table:])

    def scat(stream):
        tr.net.hashgrid_backward_rays(stream, v.samples, tr.dencoded, h.num_hits, h.indices, tr.R, tr.S,
                                      tr.grads[:tr.n_table])

    def timed(fn, n=5):
        ts = []
        for _ in range(n):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    cur = torch.cuda.current_stream()
    t_mlp = timed(lambda: mlp(None))
    t_sc = timed(lambda: scat(None))

    def both(first_hi):
        def run():
            start = torch.cuda.Event()
            start.record(cur)
            hi.wait_event(start)
            lo.wait_event(start)
            if first_hi:
                mlp(hi)
                scat(lo)
            else:
                scat(lo)
                mlp(hi)
            a, b = torch.cuda.Event(), torch.cuda.Event()
            a.record(hi)
            b.record(lo)
            cur.wait_event(a)
            cur.wait_event(b)
        return run

    print("mlp_backward alone %.3f ms, scatter alone %.3f ms, serial sum %.3f ms" % (t_mlp, t_sc, t_mlp + t_sc))
    print("both, mlp (high priority) launched first : %.3f ms" % timed(both(True)))
    print("both, scatter launched first             : %.3f ms" % timed(both(False)))


if __name__ == "__main__":
    main()
