"""BASELINE.json's full sizes through size-independent properties (the oracle finishes only samples of
them in seconds): CSR structure, one-pass == two-pass, sub-range == full range, spot checks against the
oracle on ray sub-ranges, fused == staged render, frequency network on the full sample buffer."""
import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = pytest.mark.gpu


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _march(api, grid, pose, focal, W, H, cap, ray_begin=0, n_rays=None, want_t=True):
    import torch
    R = W * H - ray_begin if n_rays is None else n_rays
    hb = api.HitBuffers(R, R * cap, want_t=want_t)
    api.ray_march_csr(grid, _t(pose.reshape(-1)), focal, W / H, W, H, hb, cap, ray_begin=ray_begin, n_rays=R)
    torch.cuda.synchronize()
    return hb


def _csr_properties(hb, cap):
    import torch
    nh = hb.num_hits.to(torch.int64)
    total = int(hb.total_hits.item())
    assert int(nh.sum().item()) == total
    assert int(nh.max().item()) <= cap and int(nh.min().item()) >= 0
    excl = torch.cumsum(nh, 0) - nh
    assert torch.equal(excl.to(torch.int32), hb.indices)
    ts, te = hb.t_start[:total], hb.t_end[:total]
    assert bool((ts >= 0).all()) and bool((te >= ts).all()) and bool(torch.isfinite(te).all())
    ep = hb.end_points[:total]
    assert bool((ep.abs() <= 1.0 + 1e-5).all())                      # exit points lie on cell faces inside the cube
    return total


def test_config2_full_view_march_and_render(cuda_api, orc):
    """800x800, 128^3 lego grid, S=4 x H_max=16: one-pass CSR properties, oracle on three ray bands,
    fused render == staged render on all 640 000 pixels."""
    import torch
    from rtx_nerf_b200 import pipeline
    api = cuda_api
    G, W, H, S, HM = 128, 800, 800, 4, 16
    occ = synth.occupancy_lego(G, seed=0)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    pose, focal = synth.poses(3, seed=123)[2], synth.focal_length()
    hb = _march(api, grid, pose, focal, W, H, HM)
    total = _csr_properties(hb, HM)
    assert total > 0
    nh, idx = hb.num_hits.cpu().numpy(), hb.indices.cpu().numpy()
    cells = hb.hit_cells[:total].cpu().numpy()
    assert np.all(occ[cells] == 1)                                    # every reported cell is occupied
    for b in (0, 317 * W + 5, W * H - 4096):                          # first rows, middle (ragged), last rows
        o = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G, max_hits=HM, csr=True, rays=(b, b + 4096))
        assert np.array_equal(nh[b:b + 4096], o.num_hits[b:b + 4096])   # the oracle indexes by global ray
        lo, hi = idx[b], idx[b + 4095] + nh[b + 4095]
        assert np.array_equal(cells[lo:hi], o.hit_prims)
        assert np.array_equal(hb.t_end[lo:hi].cpu().numpy(), o.t_end)
        assert np.array_equal(hb.start_points[lo:hi].cpu().numpy(), o.start_points)
    # a sub-range launch (the multi-GPU shard) reproduces the same records
    sub = _march(api, grid, pose, focal, W, H, HM, ray_begin=100000, n_rays=50001)
    lo, hi = idx[100000], idx[150000] + nh[150000]
    assert torch.equal(sub.num_hits, hb.num_hits[100000:150001])
    assert torch.equal(sub.end_points[:hi - lo], hb.end_points[lo:hi])
    assert torch.equal(sub.hit_cells[:hi - lo], hb.hit_cells[lo:hi])
    del hb, sub
    lv = synth.hashgrid_levels()
    table = np.random.default_rng(1).uniform(-1, 1, lv["n_entries"] * 2).astype(np.float16)
    net = api.Network(lv, table, synth.mlp_weights(seed=2))
    r = pipeline.ViewRenderer(grid, W, H, focal, 1.0, S, HM, pipeline.NetworkRadiance(net))
    staged = r.render_host(pose, fused=False).clone()
    r.pixels.fill_(-1.0)
    fused = r.render_host(pose, fused=True)
    assert float((fused - staged).abs().max()) <= 4e-6
    assert float(staged.min()) >= 0.0 and float(staged.max()) <= 1.0 + 1e-5 and float(staged.std()) > 1e-3
    # the same three 4096-ray bands through the ORACLE's network and compositing: the benchmarked view, fused
    # and staged, against orc.hashgrid_encode -> mlp_forward -> volrender at north_star's 1e-3 (fp32 RGB)
    fused_np, staged_np = fused.numpy(), staged.numpy()
    w = synth.mlp_weights(seed=2)
    for b in (0, 317 * W + 5, W * H - 4096):
        o = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G, max_hits=HM, csr=True, rays=(b, b + 4096))
        onh, oidx = o.num_hits[b:b + 4096], o.indices[b:b + 4096]
        so, to = orc.sampler(o.start_points, o.end_points, o.view_dirs[b:b + 4096], onh, oidx, S)
        _, rad = orc.mlp_forward(w, 32, 64, 2, 1, orc.hashgrid_encode(lv, table, so))
        pxo = orc.volrender(rad, onh, oidx, to, S)
        assert np.abs(fused_np[b:b + 4096] - pxo).max() < 1e-3, "fused render differs from the oracle on band %d" % b
        assert np.abs(staged_np[b:b + 4096] - pxo).max() < 1e-3, "staged render differs from the oracle on band %d" % b


def test_config2_full_view_reference_network_bands(cuda_api, orc):
    """The same view through the reference's literal network (Frequency + 8 x 128, main.cu:47-68), ray-ordered:
    three 4096-ray bands against orc.frequency_encode -> mlp_forward -> volrender.  Tolerance 2e-3: eight
    fp16 hidden layers and a MUFU sine against the oracle's double-precision one (tests/test_gpu_freqnet.py)."""
    from rtx_nerf_b200 import pipeline
    api = cuda_api
    G, W, H, S, HM = 128, 800, 800, 4, 16
    occ = synth.occupancy_lego(G, seed=0)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    pose, focal = synth.poses(3, seed=123)[2], synth.focal_length()
    fw = synth.mlp_weights(112, 128, 8, 16, seed=4)
    fnet = api.FrequencyNetwork(fw)
    r = pipeline.ViewRenderer(grid, W, H, focal, 1.0, S, HM, pipeline.FrequencyNetworkRadiance(fnet))
    px = r.render_host(pose).numpy().copy()
    for b in (0, 317 * W + 5, W * H - 4096):
        o = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G, max_hits=HM, csr=True, rays=(b, b + 4096))
        onh, oidx = o.num_hits[b:b + 4096], o.indices[b:b + 4096]
        so, to = orc.sampler(o.start_points, o.end_points, o.view_dirs[b:b + 4096], onh, oidx, S)
        _, rad = orc.mlp_forward(fw, 112, 128, 8, 1, orc.frequency_encode(so))
        pxo = orc.volrender(rad, onh, oidx, to, S)
        assert np.abs(px[b:b + 4096] - pxo).max() < 2e-3, "reference-network render differs on band %d" % b


def test_config4_shape_one_pass_equals_two_pass_records(cuda_api, orc):
    """2^18 rays, H_max = 32 (largest one-pass cap) vs H_max = 33 (two-pass path): the first 32 hits of
    every ray are the same records."""
    import torch
    api = cuda_api
    G, W, H = 128, 800, 800
    occ = synth.occupancy_lego(G, seed=0)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    pose, focal = synth.poses(2, seed=7)[1], synth.focal_length()
    R = 1 << 18
    a = _march(api, grid, pose, focal, W, H, 32, ray_begin=7919, n_rays=R)
    b = _march(api, grid, pose, focal, W, H, 33, ray_begin=7919, n_rays=R)
    _csr_properties(a, 32)
    _csr_properties(b, 33)
    assert torch.equal(torch.minimum(b.num_hits, torch.full_like(b.num_hits, 32)), a.num_hits)
    # gather the first min(nh,32) records of the two-pass result in CSR order
    nh = a.num_hits.to(torch.int64)
    ray = torch.repeat_interleave(torch.arange(R, device="cuda"), nh)
    j = torch.arange(int(nh.sum().item()), device="cuda") - a.indices.to(torch.int64)[ray]
    src = b.indices.to(torch.int64)[ray] + j
    n = int(nh.sum().item())
    for k in ("t_start", "t_end", "start_points", "end_points", "hit_cells"):
        assert torch.equal(getattr(b, k)[src], getattr(a, k)[:n]), k


def test_config5_stress_march(cuda_api, orc):
    """1920x1080, 512^3 grid at 10 % occupancy, H_max = 64: CSR properties at full size and the oracle
    on two ray bands."""
    api = cuda_api
    G, W, H, HM = 512, 1920, 1080, 64
    occ = synth.occupancy_bernoulli(G, 0.10, seed=5)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    pose, focal = synth.poses(1, seed=11)[0], synth.focal_length()
    hb = _march(api, grid, pose, focal, W, H, HM)
    total = _csr_properties(hb, HM)
    assert total > W * H                                              # more than one hit per ray on average
    nh, idx = hb.num_hits.cpu().numpy(), hb.indices.cpu().numpy()
    for b in (0, 540 * W + 901):
        o = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=HM, csr=True, rays=(b, b + 2048))
        assert np.array_equal(nh[b:b + 2048], o.num_hits[b:b + 2048])
        lo, hi = int(idx[b]), int(idx[b + 2047] + nh[b + 2047])
        assert np.array_equal(hb.hit_cells[lo:hi].cpu().numpy(), o.hit_prims)
        assert np.array_equal(hb.t_start[lo:hi].cpu().numpy(), o.t_start)
        assert np.array_equal(hb.end_points[lo:hi].cpu().numpy(), o.end_points)
    grid.destroy()


def test_frequency_network_full_sample_buffer(cuda_api, orc):
    """The reference's literal network over a config-2-sized buffer (40.96 M rows): rows sampled across
    the whole buffer match the oracle, identical rows give identical outputs wherever they sit."""
    import torch
    api = cuda_api
    n = 800 * 800 * 64
    w = synth.mlp_weights(112, 128, 8, 16, seed=3)
    net = api.FrequencyNetwork(w)
    g = torch.Generator(device="cuda").manual_seed(5)
    s = torch.rand((n, 5), device="cuda", generator=g) * 2 - 1
    s[n - 1000:] = s[:1000]                                           # first tile rows repeated in the last tiles
    rad = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    net.forward(None, s, radiance=rad)
    torch.cuda.synchronize()
    assert torch.equal(rad[n - 1000:], rad[:1000])
    pick = torch.randint(0, n, (3000,), device="cuda", generator=g)
    enc = orc.frequency_encode(s[pick].cpu().numpy())
    _, want = orc.mlp_forward(w, 112, 128, 8, 1, enc)
    np.testing.assert_allclose(rad[pick].cpu().numpy(), want, rtol=0, atol=4e-3)
    assert bool(torch.isfinite(rad).all())


@pytest.mark.parametrize("G,kind,W,H,cap", [(64, "bernoulli", 320, 200, 64), (32, "dense", 96, 64, 96),
                                             (128, "lego", 400, 300, 40), (512, "bernoulli", 1920, 1080, 64)])
def test_staged_single_pass_march_equals_two_pass(cuda_api, G, kind, W, H, cap):
    """max_hits in 33..1024: with a workspace of rtxnerf_ray_march_workspace_bytes_for the march is ONE
    traversal (hit records staged in the workspace, persistent grid); with the base-size workspace the C ABI
    falls back to count -> scan -> fill.  Every output array must be identical."""
    import ctypes as C
    import torch
    from rtx_nerf_b200 import _lib
    api = cuda_api
    occ = {"bernoulli": lambda: synth.occupancy_bernoulli(G, 0.10 if G == 512 else 0.3, seed=5),
           "dense": lambda: synth.occupancy_bernoulli(G, 1.0, seed=1),
           "lego": lambda: synth.occupancy_lego(G, seed=0)}[kind]()
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    pose, focal = synth.poses(3, seed=13)[2], synth.focal_length()
    a = _march(api, grid, pose, focal, W, H, cap)                       # api sizes the workspace: staged path
    lib = _lib.load()
    R = W * H
    assert a.workspace.numel() >= lib.rtxnerf_ray_march_workspace_bytes_for(R, cap) > \
        lib.rtxnerf_ray_march_workspace_bytes(R)
    total = _csr_properties(a, cap)
    b = api.HitBuffers(R, R * cap)
    base = lib.rtxnerf_ray_march_workspace_bytes(R)
    pose_dev = _t(pose.reshape(-1))
    cam = _lib.Camera(pose_dev.data_ptr(), float(focal), float(W / H), W, H)
    st = b.c_struct()
    rc = lib.rtxnerf_ray_march_csr(grid._h, C.byref(cam), 0, R, cap, C.byref(st), b.workspace.data_ptr(), base, None)
    assert rc == 0
    torch.cuda.synchronize()
    assert int(b.total_hits.item()) == total
    assert torch.equal(a.num_hits, b.num_hits) and torch.equal(a.indices, b.indices)
    assert torch.equal(a.view_dirs, b.view_dirs)
    for k in ("t_start", "t_end", "start_points", "end_points", "hit_cells"):
        assert torch.equal(getattr(a, k)[:total], getattr(b, k)[:total]), k
    grid.destroy()
