"""occupancy-grid rebuild from densities on the device, counter-based sampling
modes, headless image writer."""
import os

import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = pytest.mark.gpu


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("G", [2, 8, 32])
def test_grid_update_from_density(cuda_api, G):
    import torch
    api = cuda_api
    n = G ** 3
    rng = np.random.default_rng(G)
    grid = api.OccupancyGrid.from_occupancy(np.zeros(n, np.uint8), G)
    # cell centres in primitive order p = (x*G+y)*G+z, bounds of make_grid (main.cu:154-174)
    rows = grid.cell_samples().cpu().numpy()
    L = np.float32(2.0) / np.float32(G)
    p = np.arange(n)
    want = np.stack([(p // (G * G)), (p // G) % G, p % G], 1).astype(np.float32) * L - np.float32(1) + L / 2
    assert np.array_equal(rows[:, :3], want.astype(np.float32)) and np.all(rows[:, 3:] == 0)
    # plain threshold on the sigma column of a [n][4] radiance buffer
    rad = rng.uniform(0, 1, (n, 4)).astype(np.float32)
    d_rad = _t(rad)
    grid.update_from_density(d_rad[:, 3], threshold=0.6)
    torch.cuda.synchronize()
    assert np.array_equal(grid.read_bits(), synth.pack_bits((rad[:, 3] > 0.6).astype(np.uint8)))
    # running maximum with decay, two rounds
    ema = torch.zeros(n, dtype=torch.float32, device="cuda")
    s1, s2 = rng.uniform(0, 1, n).astype(np.float32), rng.uniform(0, 0.5, n).astype(np.float32)
    grid.update_from_density(_t(s1), threshold=0.4, decay=0.95, ema=ema)
    grid.update_from_density(_t(s2), threshold=0.4, decay=0.95, ema=ema)
    torch.cuda.synchronize()
    e = np.maximum(np.maximum(np.float32(0) * np.float32(0.95), s1) * np.float32(0.95), s2)
    assert np.array_equal(ema.cpu().numpy(), e)
    assert np.array_equal(grid.read_bits(), synth.pack_bits((e > 0.4).astype(np.uint8)))
    grid.destroy()


def test_grid_update_feeds_the_march(cuda_api, orc):
    """network density -> bitfield -> march: the rebuilt grid marches exactly like a grid created
    from the same bits on the host."""
    import torch
    api = cuda_api
    G, W, H, HM = 32, 32, 24, 12
    lv = synth.hashgrid_levels()
    table = np.random.default_rng(3).uniform(-1, 1, lv["n_entries"] * 2).astype(np.float16)
    net = api.Network(lv, table, synth.mlp_weights(seed=4))
    grid = api.OccupancyGrid.from_occupancy(np.ones(G ** 3, np.uint8), G)
    rows = grid.cell_samples()
    rad = torch.empty((G ** 3, 4), dtype=torch.float32, device="cuda")
    net.forward(None, rows, radiance=rad)
    sig = rad[:, 3]
    thr = float(sig.median().item())
    grid.update_from_density(sig, threshold=thr)
    torch.cuda.synchronize()
    occ = (rad[:, 3].cpu().numpy() > np.float32(thr)).astype(np.uint8)
    assert 0 < occ.sum() < G ** 3
    assert np.array_equal(grid.read_bits(), synth.pack_bits(occ))
    pose, focal = synth.poses(1, seed=2)[0], synth.focal_length()
    hb = api.HitBuffers(W * H, W * H * HM)
    api.ray_march_csr(grid, _t(pose.reshape(-1)), focal, W / H, W, H, hb, HM)
    torch.cuda.synchronize()
    ora = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=HM, csr=True)
    assert np.array_equal(hb.num_hits.cpu().numpy(), ora.num_hits)
    assert np.array_equal(hb.hit_cells[:ora.total_hits].cpu().numpy(), ora.hit_prims)


@pytest.mark.parametrize("G,max_hits", [(2, 6), (4, 12), (32, 12), (64, 40)])
def test_host_bit_update_reaches_the_brick_copy(cuda_api, orc, G, max_hits):
    """rtxnerf_grid_update replaces the linear bits; the march reads its brick-ordered copy (a 64-bit word per
    4x4x4 cells, march.cu: GridBits), which every writer of the bits must rebuild — including grids smaller than
    one brick (G = 2) and the staged path (max_hits > 32)."""
    import torch
    api = cuda_api
    W, H = 40, 30
    occ_a = synth.occupancy_bernoulli(G, 0.5, seed=11)
    occ_b = synth.occupancy_bernoulli(G, 0.3, seed=12)
    grid = api.OccupancyGrid.from_occupancy(occ_a, G)
    pose, focal = synth.poses(1, seed=5)[0], synth.focal_length()
    for occ in (occ_a, occ_b, occ_a):
        grid.update(synth.pack_bits(occ))
        hb = api.HitBuffers(W * H, W * H * max_hits)
        api.ray_march_csr(grid, _t(pose.reshape(-1)), focal, W / H, W, H, hb, max_hits)
        torch.cuda.synchronize()
        ora = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=max_hits, csr=True)
        assert np.array_equal(hb.num_hits.cpu().numpy(), ora.num_hits)
        assert np.array_equal(hb.hit_cells[:ora.total_hits].cpu().numpy(), ora.hit_prims)
        assert np.array_equal(hb.end_points[:ora.total_hits].cpu().numpy(), ora.end_points)
    grid.destroy()


@pytest.mark.parametrize("sample_type,S", [(1, 32), (2, 32), (1, 5), (2, 4)])
def test_seeded_sampling_modes(cuda_api, orc, sample_type, S):
    """STRATIFIED / UNIFORM with counter-based draws: bit-exact against the oracle, different seeds
    differ, and (unlike the reference's minstd quirk) rays do not share their draws."""
    import torch
    api = cuda_api
    rng = np.random.default_rng(S)
    nh = rng.integers(0, 6, 300).astype(np.int32)
    idx = np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32)
    Hh = int(nh.sum())
    sp = rng.uniform(-1, 1, (Hh, 3)).astype(np.float32)
    ep = (sp + rng.uniform(-0.02, 0.02, (Hh, 3))).astype(np.float32)
    vd = rng.uniform(0, 3, (300, 2)).astype(np.float32)
    out = {}
    for seed in (7, 8):
        t_vals = torch.zeros(Hh * S, device="cuda")
        samples = torch.zeros((Hh * S, 5), device="cuda")
        api.launchSampler(_t(sp), _t(ep), _t(vd), t_vals, samples, nh.size, 0, _t(nh), _t(idx), sample_type, None, S,
                          seed=seed)
        torch.cuda.synchronize()
        so, to = orc.sampler(sp, ep, vd, nh, idx, S, sample_type, seed=seed)
        assert np.array_equal(samples.cpu().numpy(), so)
        assert np.array_equal(t_vals.cpu().numpy(), to)
        out[seed] = so
    assert not np.array_equal(out[7], out[8])
    # parameters t recovered along x: two rays with the same hit count must not repeat each other
    r = np.flatnonzero(nh == nh.max())[:2]
    a = out[7][idx[r[0]] * S:(idx[r[0]] + 1) * S, 0] - sp[idx[r[0]], 0]
    b = out[7][idx[r[1]] * S:(idx[r[1]] + 1) * S, 0] - sp[idx[r[1]], 0]
    ta, tb = a / (ep[idx[r[0]], 0] - sp[idx[r[0]], 0]), b / (ep[idx[r[1]], 0] - sp[idx[r[1]], 0])
    assert np.abs(ta - tb).max() > 1e-3


def test_write_ppm(cuda_api, tmp_path):
    api = cuda_api
    W, H = 5, 3
    px = np.linspace(-0.2, 1.2, W * H * 3).astype(np.float32).reshape(W * H, 3)
    path = os.path.join(tmp_path, "view.ppm")
    api.write_ppm(path, _t(px), W, H)
    raw = open(path, "rb").read()
    head = b"P6\n5 3\n255\n"
    assert raw.startswith(head) and len(raw) == len(head) + W * H * 3
    want = (np.clip(px, 0, 1) * 255 + 0.5).astype(np.uint8).reshape(-1)
    assert np.array_equal(np.frombuffer(raw[len(head):], np.uint8), want)
