"""The N > 1 path on CPU: two gloo ranks (world_size 2) run the data-parallel training step of
rtx_nerf_b200.parallel with the ORACLE standing in for the kernels (test infrastructure only):
each rank takes its ray shard of a batch, computes the gradient chain, all-reduces the flat
buffer, applies Adam.  Checks: shards partition the batch; summed gradient == single-process
gradient of the whole batch; parameters stay bit-identical across ranks."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from rtx_nerf_b200 import parallel, synth  # noqa: E402

G, W, H, S, HM = 32, 16, 12, 4, 6
LOSS_SCALE = 256.0


def _inputs():
    lv = synth.hashgrid_levels(n_levels=16, log2_hashmap_size=12)
    rng = np.random.default_rng(0)
    table = (rng.uniform(-1, 1, lv["n_entries"] * 2) * 0.5).astype(np.float16)
    w = synth.mlp_weights(seed=1)
    occ = synth.occupancy_bernoulli(G, 0.3, seed=2)
    pose = synth.poses(1, seed=3)[0]
    target = rng.random((W * H, 3), dtype=np.float32)
    return lv, table, w, occ, pose, target


def _gradient(orc, lv, table, w, occ, pose, target, begin, end, n_loss):
    """Oracle chain on rays [begin, end): flat fp64 gradient [table | weights]; the loss is
    normalised by the WHOLE batch (n_loss elements) so shard gradients add up to the batch's."""
    focal = synth.focal_length()
    m = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=HM, csr=True, rays=(begin, end))
    nh, idx, vd = m.num_hits[begin:end], m.indices[begin:end], m.view_dirs[begin:end]
    so, to = orc.sampler(m.start_points, m.end_points, vd, nh, idx, S)
    enc = orc.hashgrid_encode(lv, table, so)
    _, rad = orc.mlp_forward(w, 32, 64, 2, 1, enc)
    px = orc.volrender(rad, nh, idx, to, S)
    diff = px - target[begin:end]
    lg = (np.float32(LOSS_SCALE) * np.float32(2.0) * diff / np.float32(n_loss)).astype(np.float16)
    rg = orc.volrender_backward(lg, rad, to, nh, idx, S)
    de, gw = orc.mlp_backward(w, 1, enc, rg)
    gt = orc.hashgrid_backward(lv, so, de)
    return np.concatenate([gt.reshape(-1), gw])


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc
    lv, table, w, occ, pose, target = _inputs()
    begin, end = parallel.ray_shard(W * H, rank, world)
    g = _gradient(orc, lv, table, w, occ, pose, target, begin, end, W * H * 3)
    flat = torch.from_numpy(g.astype(np.float32))
    parallel.allreduce_sum_(flat)
    params = np.concatenate([table, w]).astype(np.float32)
    scale = parallel.grad_scale(LOSS_SCALE) * world      # the loss was already normalised by the whole batch
    new, p16, _, _ = orc.adam_step(params, flat.numpy(), np.zeros_like(params), np.zeros_like(params), scale, 1e-3,
                                   0.9, 0.999, 1e-8, 1)
    both = [torch.zeros(p16.size, dtype=torch.int32) for _ in range(world)]
    dist.all_gather(both, torch.from_numpy(p16.view(np.uint16).astype(np.int32)))
    assert all(torch.equal(both[0], b) for b in both), "replicated parameters diverged"
    mx = parallel.max_over_ranks([float(rank + 1)], "cpu")
    assert mx == [float(world)]
    if rank == 0:
        np.save(os.path.join(out_dir, "summed.npy"), flat.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_shards_partition():
    for n, world in ((640000, 8), (1 << 18, 3), (7, 8), (100, 1)):
        seen = []
        for r in range(world):
            b, e = parallel.ray_shard(n, r, world)
            seen += list(range(b, e)) if n < 1000 else [b, e]
            assert 0 <= b <= e <= n
        if n < 1000:
            assert seen == list(range(n))
        views = sorted(v for r in range(world) for v in parallel.view_shard(100, r, world))
        assert views == list(range(100))


@pytest.mark.timeout(300)
def test_two_rank_training_step_gloo(orc, tmp_path):
    world, port = 2, 29000 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    summed = np.load(tmp_path / "summed.npy")
    lv, table, w, occ, pose, target = _inputs()
    whole = _gradient(orc, lv, table, w, occ, pose, target, 0, W * H, W * H * 3)
    assert np.abs(whole).max() > 0
    np.testing.assert_allclose(summed, whole, rtol=0, atol=1e-5 * np.abs(whole).max())
