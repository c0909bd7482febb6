"""The reference's literal network (main.cu:47-68): Frequency encoding + 8 x 128 MLP on tcgen05 with
TMEM-resident activations (network_freq.cu) against the CPU oracle.  tiny-cuda-nn is absent from the
reference archive, so the oracle restates the published definitions (PARITY UNPINNED)."""
import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = pytest.mark.gpu


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _rows(n, seed):
    rng = np.random.default_rng(seed)
    s = np.empty((n, 5), np.float32)
    s[:, :3] = rng.uniform(-1, 1, (n, 3))
    s[:, 3] = rng.uniform(0, np.pi, n)
    s[:, 4] = rng.uniform(-np.pi, np.pi, n)
    return s


def _ulp_diff_f16(a, b):
    """|a - b| in units of the fp16 spacing at max(|a|, |b|) (>= 2^-14: normal range)."""
    a32, b32 = a.astype(np.float32), b.astype(np.float32)
    mag = np.maximum(np.maximum(np.abs(a32), np.abs(b32)), 2.0 ** -14)
    ulp = 2.0 ** (np.floor(np.log2(mag)) - 10)
    return np.abs(a32 - b32) / ulp


def test_frequency_encoding_matches_oracle(cuda_api, orc):
    import torch
    s = _rows(1000, 1)
    s[0] = 0.0
    s[1, :3] = (1.0, -1.0, 0.5)      # exact multiples of 1/2: zeros of sin / cos at every octave
    net = cuda_api.FrequencyNetwork(synth.mlp_weights(112, 128, 8, 16, seed=0))
    enc = net.encode(_t(s))
    torch.cuda.synchronize()
    got, want = enc.cpu().numpy(), orc.frequency_encode(s)
    assert got.shape == (1000, 112)
    assert np.all(got[:, 108:] == np.float16(1.0))           # padding is ones
    d = np.abs(got.astype(np.float32) - want.astype(np.float32))
    assert d.max() <= 2.0 ** -10                              # one fp16 ulp at 1.0 (values are in [-1, 1])
    assert np.mean(got.view(np.uint16) == want.view(np.uint16)) > 0.98


@pytest.mark.parametrize("n,seed", [(1, 0), (127, 1), (128, 2), (129, 3), (256 * 5 + 17, 4), (128 * 310, 5)])
def test_freqnet_forward_matches_oracle(cuda_api, orc, n, seed):
    """Ragged row counts (one row, tile edges, an odd number of tiles, more pairs than SMs)."""
    import torch
    w = synth.mlp_weights(112, 128, 8, 16, seed=seed)
    net = cuda_api.FrequencyNetwork(w)
    s = _rows(n, 10 + seed)
    pad = 19
    out16 = torch.full((n + pad, 16), -7.0, dtype=torch.float16, device="cuda")
    rad = torch.full((n + pad, 4), -7.0, dtype=torch.float32, device="cuda")
    net.forward(None, _t(np.concatenate([s, np.zeros((pad, 5), np.float32)])), output=out16, radiance=rad, n=n)
    torch.cuda.synchronize()
    o16, r4 = out16.cpu().numpy(), rad.cpu().numpy()
    assert np.all(r4[n:] == -7.0) and np.all(o16[n:].astype(np.float32) == -7.0)   # rows past n untouched
    m = min(n, 4096)                                          # the oracle side is ~0.3 MFLOP per row
    enc_o = orc.frequency_encode(s[:m])
    want16, want4 = orc.mlp_forward(w, 112, 128, 8, 1, enc_o)
    # the kernel's own features (<= 1 fp16 ulp from the oracle's) through the oracle MLP: isolates the MLP
    enc_k = net.encode(_t(s[:m])).cpu().numpy()
    mlp16, _ = orc.mlp_forward(w, 112, 128, 8, 1, enc_k)
    assert _ulp_diff_f16(o16[:m], mlp16).max() <= 2.0         # fp32 vs double accumulation, fp16 re-rounding
    np.testing.assert_allclose(o16[:m].astype(np.float32), want16.astype(np.float32), rtol=0, atol=4e-3)
    np.testing.assert_array_equal(r4[:m], o16[:m, :4].astype(np.float32))
    if n > m:                                                 # every later tile: same rows give same results
        idx = np.arange(0, n, 997)
        again = torch.empty((idx.size, 16), dtype=torch.float16, device="cuda")
        net.forward(None, _t(s[idx]), output=again)
        torch.cuda.synchronize()
        assert np.array_equal(again.cpu().numpy().view(np.uint16), o16[idx].view(np.uint16))


def test_freqnet_device_count_and_linear_output(cuda_api, orc):
    """n taken from the device hit total (d_count * count_scale) and out_activation = 0."""
    import torch
    w = synth.mlp_weights(112, 128, 8, 16, seed=7)
    net = cuda_api.FrequencyNetwork(w, out_activation=0)
    s = _rows(700, 3)
    cnt = torch.tensor([150], dtype=torch.int64, device="cuda")
    rad = torch.full((700, 4), -7.0, dtype=torch.float32, device="cuda")
    net.forward(None, _t(s), radiance=rad, d_count=cnt, count_scale=4)
    torch.cuda.synchronize()
    r4 = rad.cpu().numpy()
    assert np.all(r4[600:] == -7.0)
    _, want4 = orc.mlp_forward(w, 112, 128, 8, 0, net.encode(_t(s[:600])).cpu().numpy())
    np.testing.assert_allclose(r4[:600], want4, rtol=2e-3, atol=2e-3)


def test_freqnet_renders_config1_shape(cuda_api, orc):
    """march -> sampler -> the reference's literal network -> compositing vs the oracle path."""
    from rtx_nerf_b200 import pipeline
    api = cuda_api
    G, W, H, S, HM = 32, 24, 24, 8, 12
    occ = synth.occupancy_bernoulli(G, 0.2, seed=1)
    w = synth.mlp_weights(112, 128, 8, 16, seed=2)
    net = api.FrequencyNetwork(w)
    pose, focal = synth.poses(1, seed=4)[0], synth.focal_length()
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    r = pipeline.ViewRenderer(grid, W, H, focal, 1.0, S, HM, pipeline.FrequencyNetworkRadiance(net))
    px = r.render_host(pose).numpy().copy()
    ora = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G, max_hits=HM, csr=True)
    so, to = orc.sampler(ora.start_points, ora.end_points, ora.view_dirs, ora.num_hits, ora.indices, S)
    _, rado = orc.mlp_forward(w, 112, 128, 8, 1, orc.frequency_encode(so))
    pxo = orc.volrender(rado, ora.num_hits, ora.indices, to, S)
    np.testing.assert_allclose(px, pxo, rtol=0, atol=2e-3)
