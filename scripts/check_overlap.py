#!/usr/bin/env python3
"""torchrun --nproc-per-node 2 scripts/check_overlap.py: the piecewise all-reduce issued inside Trainer.backward
(side stream, weights first) gives the same reduced gradient as one all-reduce after the backward, and the
parameters of the ranks stay identical after Adam."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rtx_nerf_b200 import api, pipeline, synth  # noqa: E402


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    G, W, H, S, HM, R = 128, 256, 192, 4, 16, 256 * 64
    occ = synth.occupancy_lego(G)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    lv = synth.hashgrid_levels()
    poses = synth.poses(4, seed=3)
    target = torch.rand((R, 3), device="cuda", generator=torch.Generator("cuda").manual_seed(5 + rank))
    out = {}
    for overlap in (True, False):
        net = api.Network(lv, synth.hashgrid_table(lv, seed=1), synth.mlp_weights(seed=2))
        tr = pipeline.Trainer(grid, net, W, H, synth.focal_length(), W / H, S, HM, R, process_group=dist.group.WORLD)
        tr.overlap = overlap
        v = tr.view
        v.pose_dev.copy_(torch.as_tensor(poses[rank].reshape(16)).cuda())
        v.ray_begin = 256 * 40
        tr.target.copy_(target)
        tr.forward()
        tr.backward()
        tr.allreduce()
        torch.cuda.synchronize()
        g = tr.grads.clone()
        tr.optimizer_step()
        torch.cuda.synchronize()
        out[overlap] = (g, tr.master.clone())
    # two executions of the scatter differ in the order of their fp32 atomics: compare to summation noise
    gmax = float(out[False][0].abs().max())
    same_g = float((out[True][0] - out[False][0]).abs().max()) < 1e-4 * gmax
    same_p = float((out[True][1] - out[False][1]).abs().max()) < 1e-5
    other = [torch.zeros_like(out[True][1]) for _ in range(dist.get_world_size())]
    dist.all_gather(other, out[True][1])
    same_ranks = all(torch.equal(other[0], o) for o in other)
    nz = int((out[True][0] != 0).sum())
    print("rank %d: reduced gradient overlap == plain: %s; parameters: %s; identical across ranks: %s; nonzero grads %d"
          % (rank, same_g, same_p, same_ranks, nz))
    assert same_g and same_p and same_ranks and nz > 0
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
