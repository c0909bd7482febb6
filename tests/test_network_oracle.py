"""CPU tests of the encoding / MLP oracle (PARITY UNPINNED: tiny-cuda-nn is absent from the
reference archive, so these pin the restatement to the published algorithm's properties and to
an independent numpy evaluation, not to the reference)."""
import numpy as np

from rtx_nerf_b200 import synth


def test_level_geometry_python_equals_c(orc):
    lv = synth.hashgrid_levels()
    sc, rs, sz = orc.hashgrid_geometry(16, 2, 19, 16, lv["per_level_scale"])
    assert np.array_equal(sc, lv["scale"]) and np.array_equal(rs, lv["res"]) and np.array_equal(sz, lv["size"])
    off, total = orc.hashgrid_offsets(16, 2, 19, 16, lv["per_level_scale"])
    assert np.array_equal(off, lv["offset"]) and total == lv["n_entries"]
    assert lv["res"][0] == 16 and lv["res"][-1] == 2048          # N_min .. N_max of instant-ngp
    assert lv["size"].max() == 1 << 19


def test_encode_interpolates_dense_level(orc):
    """On a dense level, a sample exactly on a grid vertex returns that vertex's entry."""
    lv = synth.hashgrid_levels(n_levels=2, log2_hashmap_size=19, base_resolution=16)
    rng = np.random.default_rng(0)
    table = rng.uniform(-1, 1, lv["n_entries"] * 2).astype(np.float16)
    res, scale = int(lv["res"][0]), float(lv["scale"][0])
    g = np.array([3, 5, 7])
    x01 = (g - 0.5 + 1e-4) / scale * 0 + (g + 0.0) / scale - 0.5 / scale     # pos = scale*x + 0.5 = g exactly-ish
    s = np.zeros((1, 5), np.float32)
    s[0, :3] = x01 * 2 - 1
    enc = orc.hashgrid_encode(lv, table, s).astype(np.float32)[0]
    idx = g[0] + g[1] * res + g[2] * res * res
    want = table.reshape(-1, 2)[idx].astype(np.float32)
    np.testing.assert_allclose(enc[:2], want, atol=2e-2)


def test_encode_continuity_and_range(orc):
    lv = synth.hashgrid_levels()
    rng = np.random.default_rng(1)
    table = rng.uniform(-1, 1, lv["n_entries"] * 2).astype(np.float16)
    s = np.zeros((4096, 5), np.float32)
    s[:, :3] = rng.uniform(-1, 1, (4096, 3))
    e = orc.hashgrid_encode(lv, table, s).astype(np.float32)
    assert e.shape == (4096, 32) and np.all(np.abs(e) <= 1.0 + 1e-3)
    s2 = s.copy()
    s2[:, :3] += 1e-6
    e2 = orc.hashgrid_encode(lv, table, s2).astype(np.float32)
    assert np.abs(e - e2).max() < 0.02          # trilinear interpolation is continuous


def test_mlp_oracle_vs_numpy(orc):
    rng = np.random.default_rng(2)
    w = synth.mlp_weights(seed=3)
    x = rng.uniform(-1, 1, (512, 32)).astype(np.float16)
    out16, rad = orc.mlp_forward(w, 32, 64, 2, 1, x)
    wf = w.astype(np.float64)
    W0, W1, W2 = wf[:2048].reshape(64, 32), wf[2048:6144].reshape(64, 64), wf[6144:].reshape(16, 64)
    h = np.maximum(x.astype(np.float64) @ W0.T, 0).astype(np.float32).astype(np.float16).astype(np.float64)
    h = np.maximum(h @ W1.T, 0).astype(np.float32).astype(np.float16).astype(np.float64)
    o = 1.0 / (1.0 + np.exp(-(h @ W2.T)))
    np.testing.assert_allclose(out16.astype(np.float64), o, atol=1e-3)
    assert np.array_equal(rad, out16[:, :4].astype(np.float32))
    assert w.size == 7168                        # SURVEY: 7 168 MLP parameters


def test_frequency_encoding_oracle_properties(orc):
    """Composite Frequency encoding of main.cu:47-61 restated: layout, padding, known values."""
    s = np.array([[0.0, 0.5, -0.25, 1.0, 0.125], [0.3, -0.7, 0.9, 2.0, -3.0]], np.float32)
    e = orc.frequency_encode(s).astype(np.float32)
    assert e.shape == (2, 112) and np.all(e[:, 108:] == 1.0)
    # x = 0: sin = 0, cos = 1 at every octave
    assert np.all(e[0, 0:20:2] == 0.0) and np.all(e[0, 1:20:2] == 1.0)
    # y = 0.5: sin(pi/2) = 1, cos = 0; next octave sin(pi) = 0, cos = -1; then cos = 1
    assert e[0, 20] == 1.0 and e[0, 21] == 0.0 and e[0, 22] == 0.0 and e[0, 23] == -1.0 and e[0, 25] == 1.0
    # z = -0.25: sin(-pi/4)
    assert abs(e[0, 40] + np.sqrt(0.5)) < 1e-3 and abs(e[0, 41] - np.sqrt(0.5)) < 1e-3
    # theta block starts at 60 with 12 frequencies, phi at 84
    assert abs(e[0, 84] - np.sin(np.pi * 0.125)) < 1e-3 and abs(e[0, 85] - np.cos(np.pi * 0.125)) < 1e-3
    # sin^2 + cos^2 = 1 within fp16 rounding everywhere
    p = e[1, :108].reshape(54, 2)
    assert np.all(np.abs((p ** 2).sum(1) - 1.0) < 3e-3)


def test_seeded_sampler_oracle_properties(orc):
    """Counter-based modes: stratified draws stay inside their stratum, uniform draws in
    [0,1), draws differ between rays and seeds, seed 0 reproduces the reference's repeated sequence."""
    rng = np.random.default_rng(0)
    nh = np.array([3, 0, 3, 2], np.int32)
    idx = np.array([0, 3, 3, 6], np.int32)
    sp = rng.uniform(-1, 1, (8, 3)).astype(np.float32)
    ep = sp + np.float32(0.01)
    vd = np.zeros((4, 2), np.float32)
    S = 8
    s7, t7 = orc.sampler(sp, ep, vd, nh, idx, S, 1, seed=7)
    t = ((s7[:, 0] - np.repeat(sp[:, 0], S)) / np.float32(0.01)).reshape(8, S)
    lo = np.arange(S, dtype=np.float32) / S
    assert np.all(t >= lo - 1e-3) and np.all(t <= lo + 1.0 / S + 1e-3)
    assert np.array_equal(t7.reshape(8, S), np.tile(lo, (8, 1)))              # t_vals = stratum start (sampler.cu:96)
    s8, _ = orc.sampler(sp, ep, vd, nh, idx, S, 1, seed=8)
    assert not np.array_equal(s7, s8)
    u, tu = orc.sampler(sp, ep, vd, nh, idx, S, 2, seed=7)
    tt = (u[:, 0] - np.repeat(sp[:, 0], S)) / np.float32(0.01)
    assert np.all(tt > -1e-3) and np.all(tt < 1 + 1e-3) and np.all(tu == 0)    # UNIFORM stores t_initial = 0 (Q12)
    # seed 0 = the reference: rays 0 and 2 (same hit count) draw the identical sequence (Q14)
    s0, _ = orc.sampler(sp, ep, vd, nh, idx, S, 1, seed=0)
    t0 = ((s0[:, 0] - np.repeat(sp[:, 0], S)) / np.float32(0.01)).reshape(8, S)
    assert np.allclose(t0[0:3], t0[3:6], atol=2e-3)
    assert not np.allclose(t[0:3], t[3:6], atol=2e-3)
