"""GPU parity tests (run on the B200 box: pytest -m gpu).  The CUDA path is called through the
C ABI (rtx_nerf_b200.api over librtxnerf_b200.so) and compared with the CPU oracle on the same
seeded inputs, with the committed golden vectors (the reference's own outputs), and with the
reference's own kernels compiled for sm_100a (oracle/_ref/libref_cuda.so) when present.

Bars (BASELINE.json north_star): hit counts and hit-cell lists bit-exact; t-values within 1e-5
relative (they are in fact bit-exact: both sides spell the same IEEE operations); composited RGB
within 1e-3 absolute.
"""
import ctypes as C
import os

import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _t(a, dtype=None):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda() if dtype is None else \
        torch.from_numpy(np.ascontiguousarray(a)).to(dtype).cuda()


def _march_gpu(api, occ, G, pose, focal, aspect, W, H, max_hits, capacity=None, ray_begin=0, n_rays=None):
    import torch
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    R = W * H if n_rays is None else n_rays
    cap = capacity if capacity is not None else R * max_hits
    hb = api.HitBuffers(R, cap)
    api.ray_march_csr(grid, _t(pose.reshape(-1)), focal, aspect, W, H, hb, max_hits, ray_begin=ray_begin, n_rays=R)
    torch.cuda.synchronize()
    total = int(hb.total_hits.item())
    out = dict(num_hits=hb.num_hits.cpu().numpy(), indices=hb.indices.cpu().numpy(), total=total,
               view_dirs=hb.view_dirs.cpu().numpy(), ray_origins=hb.ray_origins.cpu().numpy())
    n = min(total, cap)
    for k in ("t_start", "t_end", "start_points", "end_points", "hit_cells"):
        out[k] = getattr(hb, k)[:n].cpu().numpy()
    grid.destroy()
    return out


def _assert_march_equal(gpu, ora):
    assert np.array_equal(gpu["num_hits"], ora.num_hits), "hit counts differ"
    assert np.array_equal(gpu["indices"], ora.indices)
    assert gpu["total"] == ora.total_hits
    assert np.array_equal(gpu["hit_cells"], ora.hit_prims), "hit-cell lists differ"
    for k in ("t_start", "t_end", "start_points", "end_points"):
        assert np.array_equal(gpu[k], getattr(ora, k)), k + " not bit-exact"
    assert np.array_equal(gpu["ray_origins"], ora.ray_origins)
    np.testing.assert_allclose(gpu["view_dirs"], ora.view_dirs, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("G,p,W,H,max_hits,seed", [
    (8, 1.0, 32, 32, 24, 0),          # the reference's literal grid (main.cu:394), H_max = 3G
    (16, 0.3, 40, 24, 48, 1),         # non-square view
    (32, 0.15, 64, 64, 96, 2),
    (64, 0.05, 64, 64, 16, 3),        # capped
    (128, 0.10, 64, 64, 384, 4),      # BASELINE config 1: 128^3, 4096 rays, Bernoulli(0.10)
    (128, 1.0, 32, 32, 384, 5),       # dense 128^3
])
def test_ray_march_csr_bit_exact(cuda_api, orc, G, p, W, H, max_hits, seed):
    occ = synth.occupancy_bernoulli(G, p, seed=seed) if p < 1.0 else np.ones(G ** 3, np.uint8)
    focal = synth.focal_length()
    for pose in synth.poses(2, seed=200 + seed):
        gpu = _march_gpu(cuda_api, occ, G, pose, focal, W / H, W, H, max_hits)
        ora = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=max_hits, csr=True)
        _assert_march_equal(gpu, ora)


@pytest.mark.parametrize("G,kind,W,H,max_hits,seed", [
    (128, "lego", 200, 150, 16, 0),   # bench shape (S=4 x H_max=16), 235 tiles: multi-window look-back
    (128, "lego", 96, 64, 32, 1),     # training shape (H_max=32), largest one-pass cap
    (32, 0.15, 64, 64, 8, 2),
    (64, 0.5, 50, 30, 1, 3),          # one hit per ray
    (16, 0.02, 37, 29, 5, 4),         # sparse: most rays miss, ragged last tile
])
def test_ray_march_one_pass_bit_exact(cuda_api, orc, G, kind, W, H, max_hits, seed):
    """max_hits <= 32 takes the single-pass kernel (shared-memory hit staging, decoupled look-back,
    coalesced copy-out); same bit-exact contract as the two-pass path."""
    occ = synth.occupancy_lego(G, seed=seed) if kind == "lego" else synth.occupancy_bernoulli(G, kind, seed=seed)
    focal = synth.focal_length()
    for pose in synth.poses(2, seed=300 + seed):
        gpu = _march_gpu(cuda_api, occ, G, pose, focal, W / H, W, H, max_hits)
        ora = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=max_hits, csr=True)
        _assert_march_equal(gpu, ora)


def test_ray_march_one_pass_edges(cuda_api, orc):
    focal = synth.focal_length()
    G = 16
    occ = synth.occupancy_bernoulli(G, 0.4, seed=9)
    pose = synth.poses(1, seed=8)[0]
    full = orc.ray_march_dda(pose, focal, 1.0, 24, 24, occ, G, max_hits=12, csr=True)
    # empty grid
    gpu = _march_gpu(cuda_api, np.zeros(G ** 3, np.uint8), G, pose, focal, 1.0, 16, 16, 8)
    assert gpu["total"] == 0 and gpu["num_hits"].sum() == 0 and np.all(gpu["indices"] == 0)
    # ragged ray sub-range
    gpu = _march_gpu(cuda_api, occ, G, pose, focal, 1.0, 24, 24, 12, ray_begin=100, n_rays=333)
    assert np.array_equal(gpu["num_hits"], full.num_hits[100:433])
    lo, hi = full.indices[100], full.indices[432] + full.num_hits[432]
    assert np.array_equal(gpu["hit_cells"], full.hit_prims[lo:hi])
    assert np.array_equal(gpu["start_points"], full.start_points[lo:hi])
    assert np.array_equal(gpu["end_points"], full.end_points[lo:hi])
    # capacity overflow
    gpu = _march_gpu(cuda_api, occ, G, pose, focal, 1.0, 24, 24, 12, capacity=100)
    assert gpu["total"] == full.total_hits and gpu["hit_cells"].shape[0] == 100
    assert np.array_equal(gpu["hit_cells"], full.hit_prims[:100])
    assert np.array_equal(gpu["end_points"], full.end_points[:100])
    # operands outside the certified range of the hoisted-reciprocal division: zero, denormal
    # and huge origin components, axis-parallel view (identity rotation: tiny direction components)
    for t in ((0.0, 0.0, 0.0), (1e-37, -3e-39, 0.0), (0.0, 2.5, 1e-30), (40.0, 0.0, -1e20)):
        odd = np.eye(4, dtype=np.float32)
        odd[:3, 3] = np.array(t, np.float32)
        for W, H in ((24, 24), (1, 1), (3, 1)):
            gpu = _march_gpu(cuda_api, occ, G, odd, np.float32(40.0), W / H, W, H, 12)
            ora = orc.ray_march_dda(odd, np.float32(40.0), W / H, W, H, occ, G, max_hits=12, csr=True)
            _assert_march_equal(gpu, ora)
            gpu = _march_gpu(cuda_api, occ, G, odd, np.float32(40.0), W / H, W, H, 48)   # two-pass path
            ora = orc.ray_march_dda(odd, np.float32(40.0), W / H, W, H, occ, G, max_hits=48, csr=True)
            _assert_march_equal(gpu, ora)


def test_ray_march_lego_config1(cuda_api, orc):
    G, W, H = 128, 64, 64
    occ = synth.occupancy_lego(G, seed=0)
    assert 0.02 < occ.mean() < 0.15
    focal = synth.focal_length()
    pose = synth.poses(1, seed=3)[0]
    gpu = _march_gpu(cuda_api, occ, G, pose, focal, 1.0, W, H, 3 * G)
    ora = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G, csr=True)
    _assert_march_equal(gpu, ora)


def test_ray_march_edge_cases(cuda_api, orc):
    focal = synth.focal_length()
    pose = synth.poses(1, seed=8)[0]
    G = 16
    # empty grid: no hits, total 0
    gpu = _march_gpu(cuda_api, np.zeros(G ** 3, np.uint8), G, pose, focal, 1.0, 16, 16, 8)
    assert gpu["total"] == 0 and gpu["num_hits"].sum() == 0 and np.all(gpu["indices"] == 0)
    # camera outside the cube
    occ = synth.occupancy_bernoulli(G, 0.4, seed=9)
    far = pose.copy()
    far[:3, 3] *= 10.0
    gpu = _march_gpu(cuda_api, occ, G, far, focal, 1.0, 24, 24, 48)
    ora = orc.ray_march_dda(far, focal, 1.0, 24, 24, occ, G, csr=True)
    assert ora.total_hits > 0
    _assert_march_equal(gpu, ora)
    # ray sub-range (ragged: not a multiple of the block size), the multi-GPU sharding unit
    gpu = _march_gpu(cuda_api, occ, G, pose, focal, 1.0, 24, 24, 48, ray_begin=100, n_rays=333)
    full = orc.ray_march_dda(pose, focal, 1.0, 24, 24, occ, G, csr=True)
    assert np.array_equal(gpu["num_hits"], full.num_hits[100:433])
    lo, hi = full.indices[100], full.indices[432] + full.num_hits[432]
    assert np.array_equal(gpu["hit_cells"], full.hit_prims[lo:hi])
    assert np.array_equal(gpu["t_end"], full.t_end[lo:hi])
    # capacity overflow: total still reports the truth, nothing written past capacity
    gpu = _march_gpu(cuda_api, occ, G, pose, focal, 1.0, 24, 24, 48, capacity=100)
    assert gpu["total"] == full.total_hits and gpu["hit_cells"].shape[0] == 100
    assert np.array_equal(gpu["hit_cells"], full.hit_prims[:100])


def test_ray_march_dense_params_vs_golden(cuda_api, orc):
    """The reference's own launch shape: Params block + dense [ray][3G] buffers (main.cu:481-508)."""
    import torch
    api = cuda_api
    g = np.load(os.path.join(GOLD, "march_g8_dense.npz"))
    G, W, H, arr = int(g["G"]), int(g["W"]), int(g["H"]), int(g["arr_size"])
    holder = api.RTXDataHolder()
    holder.initContext()
    holder.createModule("optixPrograms.ptx")
    holder.createProgramGroups()
    holder.linkPipeline(False)
    holder.buildSBT()
    d_aabb = holder.initAccelerationStructure(api.make_grid(G))
    R = W * H
    dev = "cuda"
    bufs = dict(start_points=torch.full((R * arr, 3), -2.0, device=dev), end_points=torch.full((R * arr, 3), -2.0, device=dev),
                t_start=torch.full((R * arr,), -2.0, device=dev), t_end=torch.full((R * arr,), -2.0, device=dev),
                num_hits=torch.zeros(R, dtype=torch.int32, device=dev), view=torch.empty((R, 2), device=dev),
                origins=torch.empty((R, 3), device=dev), look_at=_t(g["pose"].reshape(-1)))
    p = api.Params()
    p.handle = holder.gas_handle
    p.aabb = d_aabb
    p.look_at = bufs["look_at"].data_ptr()
    p.num_primitives = G ** 3
    p.intersection_arr_size = arr
    p.start_points, p.end_points = bufs["start_points"].data_ptr(), bufs["end_points"].data_ptr()
    p.t_start, p.t_end = bufs["t_start"].data_ptr(), bufs["t_end"].data_ptr()
    p.num_hits, p.viewing_direction, p.ray_origins = bufs["num_hits"].data_ptr(), bufs["view"].data_ptr(), bufs["origins"].data_ptr()
    p.focal_length, p.aspect_ratio = float(g["focal"]), float(g["aspect"])
    p.width, p.height = W, H
    api.optixLaunch(holder.pipeline_ray_march, None, p, C.sizeof(p), holder.sbt_ray_march, W, H, 1)
    torch.cuda.synchronize()
    nh = bufs["num_hits"].cpu().numpy()
    assert np.array_equal(nh, g["num_hits"])
    mask = np.arange(arr)[None, :] < nh[:, None]
    for k in ("t_start", "t_end"):
        np.testing.assert_allclose(bufs[k].cpu().numpy().reshape(R, arr)[mask], g[k].reshape(R, arr)[mask],
                                   rtol=1e-5, atol=1e-6)
        assert np.all(bufs[k].cpu().numpy().reshape(R, arr)[~mask] == -2.0)   # unhit slots untouched
    for k in ("start_points", "end_points"):
        np.testing.assert_allclose(bufs[k].cpu().numpy().reshape(R, arr, 3)[mask], g[k].reshape(R, arr, 3)[mask],
                                   rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(bufs["view"].cpu().numpy(), g["view_dirs"], rtol=1e-5, atol=1e-6)
    # device make_grid equals the host one
    class _Raw:  # wrap the raw device pointer through __cuda_array_interface__
        __cuda_array_interface__ = dict(data=(int(d_aabb), False), shape=(G ** 3, 6), typestr="<f4", version=2)
    aabb_dev = torch.as_tensor(_Raw(), device=dev)
    assert np.array_equal(aabb_dev.cpu().numpy(), orc.make_grid(G))


# ------------------------------------------------------------------------------- sampler
def _sampler_gpu(api, sp, ep, vd, nh, idx, S, st):
    import torch
    N = int(nh.sum()) * S
    t_vals = torch.full((max(N, 1),), float("nan"), device="cuda")
    samples = torch.full((max(N, 1), 5), float("nan"), device="cuda")
    api.launchSampler(_t(sp), _t(ep), _t(vd), t_vals, samples, nh.size, 0, _t(nh), _t(idx), st, None, S)
    torch.cuda.synchronize()
    return samples[:N].cpu().numpy(), t_vals[:N].cpu().numpy()


def test_sampler_vs_golden_and_oracle(cuda_api, orc):
    g = np.load(os.path.join(GOLD, "sampler_volrender_s32.npz"))
    for name, st in (("regular", 0), ("stratified", 1), ("uniform", 2)):
        s, t = _sampler_gpu(cuda_api, g["start_points"], g["end_points"], g["view_dirs"], g["num_hits"],
                            g["indices"], 32, st)
        so, to = orc.sampler(g["start_points"], g["end_points"], g["view_dirs"], g["num_hits"], g["indices"], 32, st)
        assert np.array_equal(t, to) and np.array_equal(s, so), name + ": CUDA != oracle (bit-exact expected)"
        assert np.array_equal(t, g["t_vals_" + name])
        np.testing.assert_allclose(s, g["samples_" + name], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("S", [1, 3, 4, 5, 32, 64])
def test_sampler_ragged_and_odd_s(cuda_api, orc, S):
    rng = np.random.default_rng(S)
    B = 777
    nh = rng.integers(0, 9, B).astype(np.int32)
    nh[::7] = 0
    idx, total = orc.exclusive_scan(nh)
    sp = rng.uniform(-1, 1, (total, 3)).astype(np.float32)
    ep = (sp + rng.uniform(-0.05, 0.05, (total, 3))).astype(np.float32)
    vd = rng.uniform(-3, 3, (B, 2)).astype(np.float32)
    for st in (0, 1, 2):
        s, t = _sampler_gpu(cuda_api, sp, ep, vd, nh, idx, S, st)
        so, to = orc.sampler(sp, ep, vd, nh, idx, S, st)
        assert np.array_equal(t, to), (S, st)
        assert np.array_equal(s, so), (S, st)


# ------------------------------------------------------------------------------- vol_render
def _volrender_gpu(api, rad, nh, idx, tv, S):
    import torch
    px = torch.full((nh.size, 3), float("nan"), device="cuda")
    api.launch_volrender_cuda(None, _t(rad), _t(nh), _t(idx), _t(tv), nh.size, S, px)
    torch.cuda.synchronize()
    return px.cpu().numpy()


def _volrender_bwd_gpu(api, lg, rad, tv, nh, idx, S):
    import torch
    out = torch.zeros((tv.size, 4), dtype=torch.float16, device="cuda")
    api.launch_volrender_backward_cuda(None, _t(lg), _t(rad), _t(tv), _t(nh), _t(idx), nh.size, S, out)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_volrender_vs_golden_and_oracle(cuda_api, orc):
    g = np.load(os.path.join(GOLD, "sampler_volrender_s32.npz"))
    for S, rad, tv, ref in ((32, g["radiance"], g["t_vals_regular"], g["pixels"]),
                            (4, g["radiance_s4"], g["t_vals_s4"], g["pixels_s4"])):
        px = _volrender_gpu(cuda_api, rad, g["num_hits"], g["indices"], tv, S)
        np.testing.assert_allclose(px, ref, rtol=0, atol=1e-3)          # the north_star bar
        np.testing.assert_allclose(px, ref, rtol=0, atol=2e-5)          # what it actually achieves
        np.testing.assert_allclose(px, orc.volrender(rad, g["num_hits"], g["indices"], tv, S), rtol=0, atol=2e-5)


def test_volrender_backward_vs_golden_and_oracle(cuda_api, orc):
    g = np.load(os.path.join(GOLD, "sampler_volrender_s32.npz"))
    for S, sfx in ((32, ""), (4, "_s4")):
        rad = g["radiance" + sfx]
        tv = g["t_vals_regular"] if S == 32 else g["t_vals_s4"]
        n = int(g["num_hits"].sum()) * S
        gr = _volrender_bwd_gpu(cuda_api, g["loss_gradients"], rad, tv[:n], g["num_hits"], g["indices"], S)
        ref = g["radiance_gradients" + sfx][:n]
        ora = orc.volrender_backward(g["loss_gradients"], rad, tv[:n], g["num_hits"], g["indices"], S)
        np.testing.assert_allclose(gr.astype(np.float32), ref.astype(np.float32), rtol=2e-3, atol=1e-6)
        np.testing.assert_allclose(gr.astype(np.float32), ora.astype(np.float32), rtol=2e-3, atol=1e-6)
        assert (gr.view(np.uint16) == ref.view(np.uint16)).mean() > 0.99


@pytest.mark.parametrize("S", [1, 4, 7, 32])
def test_volrender_ragged(cuda_api, orc, S):
    rng = np.random.default_rng(10 + S)
    B = 1500
    nh = rng.integers(0, 40, B).astype(np.int32)
    nh[::5] = 0
    idx, total = orc.exclusive_scan(nh)
    N = total * S
    rad = np.empty((N, 4), np.float32)
    rad[:, :3] = rng.random((N, 3), dtype=np.float32)
    rad[:, 3] = rng.random(N, dtype=np.float32) * 20
    tv = np.tile(((np.arange(S) + 1) / S).astype(np.float32), total)
    px = _volrender_gpu(cuda_api, rad, nh, idx, tv, S)
    np.testing.assert_allclose(px, orc.volrender(rad, nh, idx, tv, S), rtol=0, atol=2e-5)
    assert np.all(px[nh == 0] == 0)
    lg = (rng.standard_normal((B, 3)) * 0.5).astype(np.float16)
    gr = _volrender_bwd_gpu(cuda_api, lg, rad, tv, nh, idx, S)
    ora = orc.volrender_backward(lg, rad, tv, nh, idx, S)
    np.testing.assert_allclose(gr.astype(np.float32), ora.astype(np.float32), rtol=2e-3, atol=1e-6)


def test_volrender_linearity_full_size(cuda_api):
    """Size-independent property at BASELINE config-2 scale (640 000 rays x 64 samples): the
    composited colour is linear in the sample colours, and zero density renders black."""
    import torch
    R, S, HM = 640_000, 4, 16
    nh = torch.full((R,), HM, dtype=torch.int32, device="cuda")
    idx = (torch.arange(R, device="cuda", dtype=torch.int64) * HM).to(torch.int32)
    N = R * HM * S
    gen = torch.Generator(device="cuda").manual_seed(0)
    rad = torch.rand((N, 4), device="cuda", generator=gen)
    rad[:, 3] *= 5.0
    tv = ((torch.arange(N, device="cuda") % S + 1).float() / S).contiguous()
    px1 = torch.empty((R, 3), device="cuda")
    px2 = torch.empty((R, 3), device="cuda")
    cuda_api.launch_volrender_cuda(None, rad, nh, idx, tv, R, S, px1)
    rad2 = rad.clone()
    rad2[:, :3] *= 0.5
    cuda_api.launch_volrender_cuda(None, rad2, nh, idx, tv, R, S, px2)
    torch.cuda.synchronize()
    assert torch.allclose(px1 * 0.5, px2, atol=1e-6)
    assert torch.isfinite(px1).all() and (px1 >= 0).all() and (px1 <= 1.0 + 1e-5).all()
    rad[:, 3] = 0
    cuda_api.launch_volrender_cuda(None, rad, nh, idx, tv, R, S, px1)
    torch.cuda.synchronize()
    assert (px1 == 0).all()


# ------------------------------------------------------------------------------- the whole config-1 path
def test_config1_end_to_end(cuda_api, orc):
    """BASELINE configs[0]: 128^3 grid, 4096 rays, fixed per-cell density/RGB, S = 32, H_max = 3G:
    march -> sampler -> cell radiance -> compositing, CUDA vs oracle."""
    import torch
    api = cuda_api
    G, W, H, S = 128, 64, 64, 32
    occ = synth.occupancy_bernoulli(G, 0.10, seed=1)
    cells = synth.cell_radiance(G, seed=2)
    pose = synth.poses(1, seed=4)[0]
    focal = synth.focal_length()
    R = W * H
    ora = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G, csr=True)
    so, to = orc.sampler(ora.start_points, ora.end_points, ora.view_dirs, ora.num_hits, ora.indices, S)
    rado = orc.cell_radiance(ora.hit_prims, S, cells)
    pxo = orc.volrender(rado, ora.num_hits, ora.indices, to, S)

    grid = api.OccupancyGrid.from_occupancy(occ, G)
    cap = ora.total_hits + 16
    hb = api.HitBuffers(R, cap)
    api.ray_march_csr(grid, _t(pose.reshape(-1)), focal, 1.0, W, H, hb, 3 * G)
    N = cap * S
    t_vals = torch.empty(N, device="cuda")
    samples = torch.empty((N, 5), device="cuda")
    rad = torch.empty((N, 4), device="cuda")
    px = torch.empty((R, 3), device="cuda")
    api.launchSampler(hb.start_points, hb.end_points, hb.view_dirs, t_vals, samples, R, G, hb.num_hits, hb.indices,
                      api.SAMPLING_REGULAR, None, S)
    api.cell_radiance(hb.hit_cells, hb.total_hits, cap, S, _t(cells), rad)
    api.launch_volrender_cuda(samples, rad, hb.num_hits, hb.indices, t_vals, R, S, px)
    torch.cuda.synchronize()
    n = ora.total_hits * S
    assert int(hb.total_hits.item()) == ora.total_hits
    assert np.array_equal(hb.num_hits.cpu().numpy(), ora.num_hits)
    assert np.array_equal(hb.hit_cells[:ora.total_hits].cpu().numpy(), ora.hit_prims)
    assert np.array_equal(t_vals[:n].cpu().numpy(), to)
    s = samples[:n].cpu().numpy()
    assert np.array_equal(s[:, :3], so[:, :3])
    np.testing.assert_allclose(s[:, 3:], so[:, 3:], rtol=1e-5, atol=1e-6)
    assert np.array_equal(rad[:n].cpu().numpy(), rado)          # densities: exact
    np.testing.assert_allclose(px.cpu().numpy(), pxo, rtol=0, atol=1e-3)
    np.testing.assert_allclose(px.cpu().numpy(), pxo, rtol=0, atol=2e-5)


# ------------------------------------------------------------------------------- reference kernels on B200
def test_reference_cuda_kernels_agree(cuda_api, orc):
    """Second, independent GPU oracle: the reference's sampler.cu / vol_render.cu compiled
    unchanged for sm_100a (oracle/_ref/libref_cuda.so)."""
    import torch
    path = os.path.join(os.path.dirname(GOLD), "..", "oracle", "_ref", "libref_cuda.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    ref = C.CDLL(os.path.abspath(path))
    rng = np.random.default_rng(3)
    B, S = 2048, 32          # B multiple of 1024: the reference reads before its bounds guard (Q13)
    nh = rng.integers(0, 12, B).astype(np.int32)
    idx, total = orc.exclusive_scan(nh)
    sp = rng.uniform(-1, 1, (total, 3)).astype(np.float32)
    ep = (sp + rng.uniform(-0.05, 0.05, (total, 3))).astype(np.float32)
    vd = rng.uniform(-3, 3, (B, 2)).astype(np.float32)
    N = total * S
    d = dict(sp=_t(sp), ep=_t(ep), vd=_t(vd), nh=_t(nh), idx=_t(idx))
    t_ref = torch.empty(N, device="cuda")
    s_ref = torch.empty((N, 5), device="cuda")
    vp = lambda t: C.c_void_p(t.data_ptr())
    rc = ref.ref_cuda_launchSampler(vp(d["sp"]), vp(d["ep"]), vp(d["vd"]), vp(t_ref), vp(s_ref), B, 0, vp(d["nh"]),
                                    vp(d["idx"]), 0, None)
    assert rc == 0
    torch.cuda.synchronize()
    s, t = _sampler_gpu(cuda_api, sp, ep, vd, nh, idx, S, 0)
    assert np.array_equal(t, t_ref.cpu().numpy())
    assert np.array_equal(s, s_ref.cpu().numpy()), "new sampler != reference sampler kernel (bit-exact expected)"
    rad = np.empty((N, 4), np.float32)
    rad[:, :3] = rng.random((N, 3), dtype=np.float32)
    rad[:, 3] = rng.random(N, dtype=np.float32) * 20
    px_ref = torch.empty((B, 3), device="cuda")
    d_rad = _t(rad)
    assert ref.ref_cuda_volrender(vp(d_rad), vp(d["nh"]), vp(d["idx"]), vp(t_ref), B, S, vp(px_ref)) == 0
    torch.cuda.synchronize()
    px = _volrender_gpu(cuda_api, rad, nh, idx, t, S)
    np.testing.assert_allclose(px, px_ref.cpu().numpy(), rtol=0, atol=2e-5)
    lg = (rng.standard_normal((B, 3)) * 0.5).astype(np.float16)
    g_ref = torch.zeros((N, 4), dtype=torch.float16, device="cuda")
    d_lg = _t(lg)
    assert ref.ref_cuda_volrender_backward(vp(d_lg), vp(d_rad), vp(t_ref), vp(d["nh"]), vp(d["idx"]), B, S, vp(g_ref)) == 0
    torch.cuda.synchronize()
    gr = _volrender_bwd_gpu(cuda_api, lg, rad, t, nh, idx, S)
    assert np.array_equal(gr.view(np.uint16), g_ref.cpu().numpy().view(np.uint16)), \
        "new backward != reference backward kernel (bit-exact expected: same op sequence)"


@pytest.mark.parametrize("max_hits", [16, 48])
def test_packed_hit_records_equal_the_point_arrays(cuda_api, max_hits):
    """rtxnerf_ray_march_csr_seg writes (sx, sy, sz, ex, ey, ez, 0, 0) per hit beside — or instead of — the two
    [capacity][3] arrays: same bits, one-pass (<= 32 hits) and staged paths, and nothing else changes."""
    import torch
    api = cuda_api
    G, W, H = 64, 96, 64
    grid = api.OccupancyGrid.from_occupancy(synth.occupancy_lego(G, seed=2), G)
    pose = _t(synth.poses(1, seed=12)[0].reshape(-1))
    focal = synth.focal_length()
    R = W * H
    ref = api.HitBuffers(R, R * max_hits)
    api.ray_march_csr(grid, pose, focal, W / H, W, H, ref, max_hits)
    both = api.HitBuffers(R, R * max_hits, want_segments=True)
    api.ray_march_csr(grid, pose, focal, W / H, W, H, both, max_hits)
    only = api.HitBuffers(R, R * max_hits, want_segments=True)
    only.start_points.fill_(-5.0)
    only.end_points.fill_(-5.0)
    api.ray_march_csr(grid, pose, focal, W / H, W, H, only, max_hits, points=False)
    torch.cuda.synchronize()
    n = int(ref.total_hits.item())
    assert n > 1000 and int(both.total_hits.item()) == n and int(only.total_hits.item()) == n
    sp, ep = ref.start_points[:n].cpu().numpy(), ref.end_points[:n].cpu().numpy()
    for hb in (both, only):
        seg = hb.segments8[:n].cpu().numpy()
        assert np.array_equal(seg[:, 0:3], sp) and np.array_equal(seg[:, 3:6], ep) and not seg[:, 6:].any()
        for k in ("num_hits", "indices", "t_start", "t_end", "hit_cells"):
            assert torch.equal(getattr(hb, k)[:n] if k in ("t_start", "t_end", "hit_cells") else getattr(hb, k),
                               getattr(ref, k)[:n] if k in ("t_start", "t_end", "hit_cells") else getattr(ref, k)), k
    assert torch.equal(both.start_points[:n], ref.start_points[:n]) and torch.equal(both.end_points[:n], ref.end_points[:n])
    assert float(only.start_points.max()) == -5.0 and float(only.end_points.max()) == -5.0     # left alone
    grid.destroy()
