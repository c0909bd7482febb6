"""GPU tests of the hash-grid encoding and the fused tcgen05 MLP through the C ABI, against
the CPU oracle and the CUDA-core check kernel.  PARITY UNPINNED w.r.t. tiny-cuda-nn (absent):
the bar is the oracle's restatement; tolerances: encoded features bit-exact (same op order),
network outputs within 2e-3 absolute (fp16 activations, tensor-core accumulation order),
composited RGB within 1e-3 absolute (north_star)."""
import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _net(api, table_scale=1.0, seed=0, n_levels=16, out_activation=1):
    lv = synth.hashgrid_levels(n_levels=n_levels)
    rng = np.random.default_rng(seed)
    table = (rng.uniform(-1, 1, lv["n_entries"] * 2) * table_scale).astype(np.float16)
    w = synth.mlp_weights(seed=seed + 1)
    return api.Network(lv, table, w, out_activation=out_activation), lv, table, w


def _samples(n, seed):
    rng = np.random.default_rng(seed)
    s = np.zeros((n, 5), np.float32)
    s[:, :3] = rng.uniform(-1, 1, (n, 3))
    s[:, 3:] = rng.uniform(-3, 3, (n, 2))
    return s


def test_hashgrid_encode_bit_exact(cuda_api, orc):
    import torch
    net, lv, table, _ = _net(cuda_api)
    s = _samples(5000, 1)
    s[:8, :3] = [[-1, -1, -1], [1, 1, 1], [0, 0, 0], [1, -1, 1], [-1, 1, -1], [0.999999, 0.5, -0.5], [-1, 0, 1], [0.25, 0.25, 0.25]]
    e = net.encode(_t(s))
    torch.cuda.synchronize()
    eo = orc.hashgrid_encode(lv, table, s)
    assert np.array_equal(e.cpu().numpy().view(np.uint16), eo.view(np.uint16)), "encoded features differ from the oracle"


@pytest.mark.parametrize("n", [1, 127, 128, 129, 4096 + 77])
def test_mlp_tcgen05_pre_encoded_vs_oracle(cuda_api, orc, n):
    """Pure MLP on the tensor cores (pre-encoded fp16 input), ragged tile counts."""
    import torch
    net, _, _, w = _net(cuda_api)
    rng = np.random.default_rng(n)
    x = rng.uniform(-1, 1, (n, 32)).astype(np.float16)
    out = torch.zeros((n, 16), dtype=torch.float16, device="cuda")
    rad = torch.zeros((n, 4), dtype=torch.float32, device="cuda")
    net.forward(None, _t(x), output=out, radiance=rad, pre_encoded=True)
    torch.cuda.synchronize()
    o16, orad = orc.mlp_forward(w, 32, 64, 2, 1, x)
    np.testing.assert_allclose(out.cpu().numpy().astype(np.float32), o16.astype(np.float32), rtol=0, atol=2e-3)
    np.testing.assert_allclose(rad.cpu().numpy(), orad, rtol=0, atol=2e-3)
    assert np.array_equal(rad.cpu().numpy(), out.cpu().numpy()[:, :4].astype(np.float32))


def test_mlp_no_output_activation_and_simt_agree(cuda_api, orc):
    import torch
    net, _, _, w = _net(cuda_api, out_activation=0)
    x = np.random.default_rng(5).uniform(-2, 2, (3000, 32)).astype(np.float16)
    a = net.forward(None, _t(x), pre_encoded=True, impl=cuda_api.IMPL_TCGEN05)
    b = net.forward(None, _t(x), pre_encoded=True, impl=cuda_api.IMPL_SIMT)
    torch.cuda.synchronize()
    o16, _ = orc.mlp_forward(w, 32, 64, 2, 0, x)
    ref = o16.astype(np.float32)
    tol = 4e-3 * max(1.0, float(np.abs(ref).max()))
    np.testing.assert_allclose(a.cpu().numpy().astype(np.float32), ref, rtol=0, atol=tol)
    np.testing.assert_allclose(b.cpu().numpy().astype(np.float32), ref, rtol=0, atol=tol)


def test_fused_encode_mlp_vs_oracle(cuda_api, orc):
    """The product kernel: sampler rows [N][5] -> hash-grid -> MLP -> fp32 (r,g,b,sigma)."""
    import torch
    net, lv, table, w = _net(cuda_api)
    n = 20000
    s = _samples(n, 7)
    rad = torch.zeros((n, 4), dtype=torch.float32, device="cuda")
    out = torch.zeros((n, 16), dtype=torch.float16, device="cuda")
    net.forward(None, _t(s), output=out, radiance=rad)
    torch.cuda.synchronize()
    enc = orc.hashgrid_encode(lv, table, s)
    o16, orad = orc.mlp_forward(w, 32, 64, 2, 1, enc)
    np.testing.assert_allclose(rad.cpu().numpy(), orad, rtol=0, atol=2e-3)
    np.testing.assert_allclose(out.cpu().numpy().astype(np.float32), o16.astype(np.float32), rtol=0, atol=2e-3)
    assert np.abs(orad - 0.5).max() > 0.1          # the test exercises the network, not sigmoid(0)
    # device-side row count: only the first 2*3000 rows may be written
    rad2 = torch.full((n, 4), -1.0, device="cuda")
    cnt = torch.tensor([3000], dtype=torch.int64, device="cuda")
    net.forward(None, _t(s), radiance=rad2, d_count=cnt, count_scale=2)
    torch.cuda.synchronize()
    r2 = rad2.cpu().numpy()
    assert np.array_equal(r2[:6000], rad.cpu().numpy()[:6000]) and np.all(r2[6000:] == -1.0)


def test_network_render_config2_small(cuda_api, orc):
    """BASELINE configs[1] in miniature: lego grid, S = 4, H_max = 16, hash-grid + 2x64 MLP,
    whole path CUDA vs oracle; RGB within 1e-3."""
    import torch
    from rtx_nerf_b200 import pipeline
    api = cuda_api
    G, W, H, S, HM = 128, 48, 48, 4, 16
    occ = synth.occupancy_lego(G)
    net, lv, table, w = _net(api, table_scale=1.0, seed=3)
    pose = synth.poses(1, seed=6)[0]
    focal = synth.focal_length()
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    r = pipeline.ViewRenderer(grid, W, H, focal, 1.0, S, HM, pipeline.NetworkRadiance(net))
    assert r.can_fuse
    px = r.render_host(pose, fused=False).numpy().copy()          # staged: sampler, network, compositing
    px_fused = r.render_host(pose, fused=True).numpy().copy()     # the same in one kernel (view_fused.cu)
    ora = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G, max_hits=HM, csr=True)
    so, to = orc.sampler(ora.start_points, ora.end_points, ora.view_dirs, ora.num_hits, ora.indices, S)
    _, rado = orc.mlp_forward(w, 32, 64, 2, 1, orc.hashgrid_encode(lv, table, so))
    pxo = orc.volrender(rado, ora.num_hits, ora.indices, to, S)
    assert r.total_hits() == ora.total_hits
    assert np.array_equal(r.hits.num_hits.cpu().numpy(), ora.num_hits)
    np.testing.assert_allclose(px, pxo, rtol=0, atol=1e-3)
    np.testing.assert_allclose(px_fused, pxo, rtol=0, atol=1e-3)
    # same per-sample arithmetic; only the order of the per-ray sums differs (sequential vs warp scan)
    np.testing.assert_allclose(px_fused, px, rtol=0, atol=2e-6)


@pytest.mark.parametrize("G,kind,W,H,S,HM", [
    (32, 0.2, 40, 24, 3, 12),       # S not a power of two: repeated-addition t (sampler.cu:65)
    (16, 0.3, 33, 17, 32, 8),       # the reference's literal S = 32 (sampler.h:4)
    (64, 0.02, 50, 30, 4, 16),      # sparse: most rays miss, ragged groups
    (16, 0.0, 20, 20, 4, 4),        # empty grid: every pixel black
])
def test_fused_render_matches_staged_and_oracle(cuda_api, orc, G, kind, W, H, S, HM):
    from rtx_nerf_b200 import pipeline
    api = cuda_api
    occ = synth.occupancy_bernoulli(G, kind, seed=5) if kind > 0 else np.zeros(G ** 3, np.uint8)
    net, lv, table, w = _net(api, table_scale=1.0, seed=7)
    pose = synth.poses(1, seed=9)[0]
    focal = synth.focal_length()
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    r = pipeline.ViewRenderer(grid, W, H, focal, W / H, S, HM, pipeline.NetworkRadiance(net))
    px = r.render_host(pose, fused=False).numpy().copy()
    r.pixels.fill_(-3.0)
    px_fused = r.render_host(pose, fused=True).numpy().copy()
    np.testing.assert_allclose(px_fused, px, rtol=0, atol=2e-6)
    ora = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=HM, csr=True)
    if ora.total_hits:
        so, to = orc.sampler(ora.start_points, ora.end_points, ora.view_dirs, ora.num_hits, ora.indices, S)
        _, rado = orc.mlp_forward(w, 32, 64, 2, 1, orc.hashgrid_encode(lv, table, so))
        pxo = orc.volrender(rado, ora.num_hits, ora.indices, to, S)
    else:
        pxo = np.zeros((W * H, 3), np.float32)
    np.testing.assert_allclose(px_fused, pxo, rtol=0, atol=1e-3)


@pytest.mark.parametrize("n_rays,S", [(1, 4), (33, 4), (700, 3), (257, 32)])
def test_ray_coherent_schedule_matches_flat_kernel(cuda_api, n_rays, S):
    """network_forward_rays (lane = ray, warp = sample slot) must write the SAME bits as the flat
    kernel on every CSR row, and leave rows outside (num_hits, indices) untouched; ragged rays
    (0 hits, 1 hit, many hits) and a ray count that does not fill a 32-ray group."""
    import torch
    net, _, _, _ = _net(cuda_api, seed=11)
    rng = np.random.default_rng(n_rays * 7 + S)
    nh = rng.integers(0, 9, n_rays).astype(np.int32)
    nh[rng.random(n_rays) < 0.2] = 0
    idx = np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32)
    n = int(nh.sum()) * S
    pad = 37                                   # rows after the last ray: must stay untouched
    s = _samples(n + pad, n_rays)
    ds = _t(s)
    flat16 = torch.zeros((n + pad, 16), dtype=torch.float16, device="cuda")
    flat4 = torch.zeros((n + pad, 4), dtype=torch.float32, device="cuda")
    net.forward(None, ds, output=flat16, radiance=flat4)
    ray16 = torch.full((n + pad, 16), -7.0, dtype=torch.float16, device="cuda")
    ray4 = torch.full((n + pad, 4), -7.0, dtype=torch.float32, device="cuda")
    net.forward_rays(None, ds, _t(nh), _t(idx), n_rays, S, output=ray16, radiance=ray4)
    torch.cuda.synchronize()
    assert np.array_equal(ray16.cpu().numpy()[:n].view(np.uint16), flat16.cpu().numpy()[:n].view(np.uint16))
    assert np.array_equal(ray4.cpu().numpy()[:n], flat4.cpu().numpy()[:n])
    assert np.all(ray4.cpu().numpy()[n:] == -7.0) and np.all(ray16.cpu().numpy()[n:].astype(np.float32) == -7.0)
    # radiance-only call (no out16): same radiance
    ray4b = torch.zeros((n + pad, 4), dtype=torch.float32, device="cuda")
    net.forward_rays(None, ds, _t(nh), _t(idx), n_rays, S, radiance=ray4b)
    torch.cuda.synchronize()
    assert np.array_equal(ray4b.cpu().numpy()[:n], flat4.cpu().numpy()[:n])


def test_batch_render_views_host_matches_single_view_calls(cuda_api):
    """ViewRenderer.render_views_host (configs[2]: a batch of views, the device->host copy of view k
    under the kernels of view k+1) returns exactly what render_host returns view by view, for the
    fused and the staged path, and leaves the renderer's pixel buffer where it was."""
    import torch
    from rtx_nerf_b200 import api, pipeline
    G, W, H, S, HM = 64, 96, 80, 4, 16
    occ = synth.occupancy_lego(G)
    lv = synth.hashgrid_levels()
    net = api.Network(lv, synth.hashgrid_table(lv, seed=1), synth.mlp_weights(seed=2))
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    r = pipeline.ViewRenderer(grid, W, H, synth.focal_length(), W / H, S, HM, pipeline.NetworkRadiance(net))
    poses = synth.poses(5, seed=21)
    for fused in (True, False):
        single = [r.render_host(p, fused=fused).clone() for p in poses]
        before = r.pixels.data_ptr()
        out = r.render_views_host(poses, fused=fused)
        assert r.pixels.data_ptr() == before
        assert tuple(out.shape) == (5, W * H, 3) and out.is_pinned()
        for k in range(5):
            assert torch.equal(out[k], single[k]), "view %d differs (fused=%s)" % (k, fused)
    assert any(float(s.abs().max()) > 0 for s in single)
    with pytest.raises(ValueError):
        r.render_views_host(poses, out=torch.empty((5, W * H, 3)))
