"""SURVEY 8 f3, second half: the reference's training-loop front end (main.cu:612-676 — shuffle, slice,
collect, scan, gather, five H2D copies) as a resident ray store + three kernels, bit-exact against the
oracle's restatement of the host loops."""
import ctypes as C

import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = pytest.mark.gpu

KAT = {(8, 0): [6, 0, 7, 5, 3, 2, 1, 4]}


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("n,group,first,count", [(1, 1, 0, 1), (2, 1, 0, 2), (5, 1, 1, 3), (1000, 1, 0, 1000),
                                                 (4096, 32, 64, 2048), (65537, 1, 0, 65537),
                                                 (1 << 18, 1, 0, 1 << 18), (64000000, 32, 63000000, 1000000),
                                                 (64000000, 1, 12345, 300001), (10, 1, 10, 0)])
def test_ray_permutation_matches_oracle(cuda_api, orc, n, group, first, count):
    for seed in (0, 2024, (1 << 64) - 1):
        got = cuda_api.ray_permutation(n, seed, group, first, count).cpu().numpy()
        assert np.array_equal(got, orc.ray_permutation(n, seed, group, first, count))
    assert cuda_api.ray_permutation(8, 0).cpu().tolist() == KAT[(8, 0)]


def test_ray_permutation_full_size_is_a_bijection(cuda_api):
    """BASELINE config 3/4 shape: 100 views x 800 x 800 = 64 M rays; every ray exactly once."""
    import torch
    n = 64000000
    p = cuda_api.ray_permutation(n, seed=4, group=32)
    seen = torch.zeros(n, dtype=torch.uint8, device="cuda")
    seen[p] = 1
    assert int(seen.sum(dtype=torch.int64)) == n
    assert int(p.sum()) == n * (n - 1) // 2
    assert bool(((p.view(-1, 32)[:, 1:] - p.view(-1, 32)[:, :-1]) == 1).all())


def _filled_store(api, G, W, H, HM, n_views, want_t, seed):
    occ = synth.occupancy_bernoulli(G, 0.25, seed=seed)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    poses = synth.poses(n_views, seed=seed + 1)
    gt = np.random.default_rng(seed + 2).random((n_views * W * H, 3)).astype(np.float32)
    store = api.RayStore(n_views, W * H, n_views * W * H * HM, want_t=want_t, pixels_gt=_t(gt))
    for v in range(n_views):
        store.add_view(grid, _t(poses[v]), synth.focal_length(), 1.0, W, H, HM)
    return occ, grid, poses, gt, store


def _host_store(store, gt):
    d = dict(rays_per_view=store.rays_per_view, view_base=store.view_base.cpu().numpy(),
             num_hits=store.num_hits.cpu().numpy(), indices=store.indices.cpu().numpy(),
             view_dirs=store.view_dirs.cpu().numpy(), start_points=store.start_points.cpu().numpy(),
             end_points=store.end_points.cpu().numpy(), pixels_gt=gt)
    if store.t_start is not None:
        d["t_start"], d["t_end"] = store.t_start.cpu().numpy(), store.t_end.cpu().numpy()
    return d


def test_ray_store_holds_the_marches_of_all_views(cuda_api, orc):
    """add_view == the oracle's march of that view, placed at the view's hit base."""
    G, W, H, HM, V = 32, 40, 24, 12, 3
    occ, grid, poses, gt, store = _filled_store(cuda_api, G, W, H, HM, V, True, seed=10)
    assert store.view_base_host == store.view_base.cpu().tolist()
    for v in range(V):
        o = orc.ray_march_dda(poses[v], synth.focal_length(), 1.0, W, H, occ, G, csr=True, max_hits=HM)
        r0, h0 = v * W * H, store.view_base_host[v]
        assert np.array_equal(store.num_hits[r0:r0 + W * H].cpu().numpy(), o.num_hits)
        assert np.array_equal(store.indices[r0:r0 + W * H].cpu().numpy(), o.indices)
        # angles: CUDA atan2f vs libm, the tolerance of the march parity tests (tests/test_gpu_parity.py)
        np.testing.assert_allclose(store.view_dirs[r0:r0 + W * H].cpu().numpy(), o.view_dirs, rtol=1e-5, atol=1e-6)
        for name in ("start_points", "end_points", "t_start", "t_end"):
            assert np.array_equal(getattr(store, name)[h0:h0 + o.total_hits].cpu().numpy(), getattr(o, name)), name
    assert store.used_hits == sum(int(store.num_hits[v * W * H:(v + 1) * W * H].sum()) for v in range(V))
    with pytest.raises(ValueError):
        store.add_view(grid, _t(poses[0]), synth.focal_length(), 1.0, W, H, HM)
    grid.destroy()


@pytest.mark.parametrize("group,first,batch,want_t", [(1, 0, 960, True), (1, 700, 1000, False), (32, 64, 1024, True),
                                                      (1, 0, 2880, True), (32, 2879 - 2879 % 32, 32, False),
                                                      (1, 17, 1, True), (1, 5, 0, True)])
def test_batch_assemble_matches_oracle(cuda_api, orc, group, first, batch, want_t):
    import torch
    api = cuda_api
    G, W, H, HM, V = 32, 40, 24, 12, 3
    occ, grid, poses, gt, store = _filled_store(api, G, W, H, HM, V, want_t, seed=20)
    hits = api.HitBuffers(max(batch, 1), max(batch, 1) * HM, want_t=want_t, want_cells=False, want_origins=False)
    d_gt = torch.full((max(batch, 1), 3), -1.0, dtype=torch.float32, device="cuda")
    d_ids = torch.full((max(batch, 1),), -1, dtype=torch.int64, device="cuda")
    api.batch_assemble(store, 77, group, first, batch, hits, gt_pixels=d_gt, ray_ids=d_ids)
    torch.cuda.synchronize()
    o = orc.batch_assemble(_host_store(store, gt), 77, group, first, batch)
    T = o["total_hits"]
    assert int(hits.total_hits.item()) == T
    assert np.array_equal(d_ids[:batch].cpu().numpy(), o["ray_ids"])
    assert np.array_equal(hits.num_hits[:batch].cpu().numpy(), o["num_hits"])
    assert np.array_equal(hits.indices[:batch].cpu().numpy(), o["indices"])
    assert np.array_equal(hits.view_dirs[:batch].cpu().numpy(), o["view_dirs"])
    assert np.array_equal(d_gt[:batch].cpu().numpy(), o["gt_pixels"])
    names = ("start_points", "end_points") + (("t_start", "t_end") if want_t else ())
    for name in names:
        assert np.array_equal(getattr(hits, name)[:T].cpu().numpy(), o[name]), name
    grid.destroy()


def test_batch_assemble_capacity_and_errors(cuda_api, orc):
    import torch
    api = cuda_api
    G, W, H, HM, V = 16, 32, 16, 8, 2
    occ, grid, poses, gt, store = _filled_store(api, G, W, H, HM, V, False, seed=30)
    ho = _host_store(store, gt)
    full = orc.batch_assemble(ho, 1, 1, 0, 512)
    cap = full["total_hits"] // 2
    hits = api.HitBuffers(512, cap + 64, want_t=False, want_cells=False, want_origins=False)
    hits.capacity = cap                                   # the 64 slots past the capacity must stay untouched
    hits.start_points.fill_(-7.0)
    hits.end_points.fill_(-7.0)
    api.batch_assemble(store, 1, 1, 0, 512, hits)
    torch.cuda.synchronize()
    assert int(hits.total_hits.item()) == full["total_hits"]          # the total stays true
    assert np.array_equal(hits.start_points[:cap].cpu().numpy(), full["start_points"][:cap])
    assert np.array_equal(hits.end_points[:cap].cpu().numpy(), full["end_points"][:cap])
    assert bool((hits.start_points[cap:] == -7.0).all()) and bool((hits.end_points[cap:] == -7.0).all())
    # argument errors: past the end (the reference slices past the end, Q6), group not dividing, gt without a source
    lib = api._lib.load()
    st, hs = store.c_struct(), hits.c_struct()
    ws = hits.workspace
    call = lambda **k: lib.rtxnerf_batch_assemble(C.byref(st), 1, k.get("group", 1), k.get("first", 0), k.get("n", 512),
                                                   C.byref(hs), k.get("gt"), None, C.c_void_p(ws.data_ptr()),
                                                   k.get("wsb", ws.numel()), None)
    assert call(first=600) == -1 and call(group=3) == -1 and call(n=-1) == -1
    assert call(wsb=16) == api._lib.load().rtxnerf_batch_assemble(C.byref(st), 1, 1, 0, 512, C.byref(hs), None, None,
                                                                  None, 0, None) != 0
    store.pixels_gt = None
    st2 = store.c_struct()
    assert lib.rtxnerf_batch_assemble(C.byref(st2), 1, 1, 0, 512, C.byref(hs), C.c_void_p(ws.data_ptr()), None,
                                      C.c_void_p(ws.data_ptr()), ws.numel(), None) == -1
    with pytest.raises(ValueError):
        api.batch_assemble(api.RayStore(2, 16, 64), 1, 1, 0, 8, hits)   # store not filled
    grid.destroy()


def test_batch_full_size_properties(cuda_api):
    """BASELINE config 4 shape: 2^18-ray batches drawn from 800 x 800 views by the seeded shuffle.  Size-
    independent properties: counts and records equal the store's at the shuffled ids (checked with device
    gathers), scan is the exclusive prefix sum, an epoch's batches cover every ray once."""
    import torch
    api = cuda_api
    G, W, H, HM, V, B = 128, 800, 800, 32, 4, 1 << 18
    occ = synth.occupancy_lego(G)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    poses = synth.poses(V, seed=1)
    n = V * W * H
    gt = torch.rand((n, 3), device="cuda")
    store = api.RayStore(V, W * H, n * HM, pixels_gt=gt)
    for v in range(V):
        store.add_view(grid, _t(poses[v]), synth.focal_length(), 1.0, W, H, HM)
    hits = api.HitBuffers(B, B * HM, want_t=False, want_cells=False, want_origins=False)
    d_gt = torch.empty((B, 3), device="cuda")
    d_ids = torch.empty(B, dtype=torch.int64, device="cuda")
    seen = torch.zeros(n, dtype=torch.int32, device="cuda")
    for b in range(n // B):
        api.batch_assemble(store, 9, 32, b * B, B, hits, gt_pixels=d_gt, ray_ids=d_ids)
        ids = d_ids
        assert torch.equal(ids, api.ray_permutation(n, 9, 32, b * B, B))
        seen[ids] += 1
        nh = store.num_hits[ids]
        assert torch.equal(hits.num_hits, nh)
        assert torch.equal(hits.indices.long(), torch.cumsum(nh.long(), 0) - nh.long())
        T = int(hits.total_hits.item())
        assert T == int(nh.sum(dtype=torch.int64)) and 0 < T <= B * HM
        assert torch.equal(hits.view_dirs, store.view_dirs[ids]) and torch.equal(d_gt, gt[ids])
        # hit j of batch ray k <- store hit view_base[view] + indices[ray] + j
        owner = torch.repeat_interleave(torch.arange(B, device="cuda"), nh.long())
        j = torch.arange(T, device="cuda") - hits.indices.long()[owner]
        src = store.view_base[ids[owner] // (W * H)] + store.indices[ids[owner]].long() + j
        assert torch.equal(hits.start_points[:T], store.start_points[src])
        assert torch.equal(hits.end_points[:T], store.end_points[src])
    assert int(seen.max()) <= 1 and int(seen.sum(dtype=torch.int64)) == (n // B) * B
    grid.destroy()


def test_store_fed_forward_equals_the_per_view_forward(cuda_api):
    """One group spanning the whole store makes the shuffle the identity, so a batch is a contiguous ray
    range of the view — exactly what Trainer.step marches itself.  Hit records, samples and composited
    pixels must then be bit-identical between the two front ends (same kernels downstream)."""
    import torch
    from rtx_nerf_b200 import pipeline
    api = cuda_api
    G, W, H, HM, S, R, first = 32, 32, 32, 8, 4, 256, 128
    occ = synth.occupancy_bernoulli(G, 0.3, seed=40)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    lv = synth.hashgrid_levels()
    net = api.Network(lv, synth.hashgrid_table(lv, seed=41), synth.mlp_weights(seed=42))
    pose = _t(synth.poses(1, seed=43)[0])
    gt = torch.rand((W * H, 3), device="cuda")
    store = api.RayStore(1, W * H, W * H * HM, pixels_gt=gt)
    store.add_view(grid, pose, synth.focal_length(), 1.0, W, H, HM)
    assert api.ray_permutation(W * H, 5, W * H).cpu().tolist() == list(range(W * H))
    tr = pipeline.Trainer(grid, net, W, H, synth.focal_length(), 1.0, S, HM, R)
    tr.from_hits = False        # this test compares the sample buffers too: keep the sampler in the step
    v = tr.view
    v.pose_dev.copy_(pose.reshape(16))
    v.ray_begin = first
    tr.forward()
    torch.cuda.synchronize()
    T = int(v.hits.total_hits.item())
    a = [x.clone() for x in (v.hits.num_hits, v.hits.indices, v.hits.view_dirs, v.hits.start_points[:T],
                             v.hits.end_points[:T], v.samples[:T * S], v.t_vals[:T * S], v.pixels)]
    for x in (v.hits.num_hits, v.hits.indices, v.hits.view_dirs, v.hits.start_points, v.hits.end_points, v.samples,
              v.t_vals, v.pixels):
        x.zero_()
    api.batch_assemble(store, 5, W * H, first, R, v.hits, gt_pixels=tr.target)
    v.stage_sampler()
    v.stage_radiance()
    v.stage_composite()
    torch.cuda.synchronize()
    assert int(v.hits.total_hits.item()) == T and T > 0
    b = (v.hits.num_hits, v.hits.indices, v.hits.view_dirs, v.hits.start_points[:T], v.hits.end_points[:T],
         v.samples[:T * S], v.t_vals[:T * S], v.pixels)
    for i, (x, y) in enumerate(zip(a, b)):
        assert torch.equal(x, y), i
    assert torch.equal(tr.target, gt[first:first + R])
    grid.destroy()
