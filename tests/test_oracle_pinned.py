"""CPU tests: the oracle (oracle/rtxnerf_oracle.c) against the golden vectors produced by the
reference's own code (tests/golden/make_golden.py), and the oracle's two marchers against each
other.  These pin the oracle; the -m gpu tests then compare the CUDA path with the oracle.

Tolerances (BASELINE.json north_star): hit counts / hit-cell lists bit-exact; t-values within
1e-5 relative; composited RGB within 1e-3 absolute.  The host build of the reference is g++
(-ffp-contract=fast), the oracle follows nvcc's PTX contraction, so values agree to a few ulp,
not bit for bit.
"""
import os

import numpy as np
import pytest

from rtx_nerf_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL_T = 1e-5


def _load(name):
    return np.load(os.path.join(GOLD, name))


def _assert_hits_match(res, g, arr):
    R = int(g["W"]) * int(g["H"])
    assert np.array_equal(res.num_hits, g["num_hits"]), "hit counts differ from the reference"
    mask = np.arange(arr)[None, :] < g["num_hits"][:, None]
    for k in ("t_start", "t_end"):
        a = getattr(res, k).reshape(R, arr)[mask]
        b = g[k].reshape(R, arr)[mask]
        np.testing.assert_allclose(a, b, rtol=RTOL_T, atol=1e-6, err_msg=k)
    for k in ("start_points", "end_points"):
        a = getattr(res, k).reshape(R, arr, 3)[mask]
        b = g[k].reshape(R, arr, 3)[mask]
        np.testing.assert_allclose(a, b, rtol=RTOL_T, atol=1e-6, err_msg=k)
    np.testing.assert_allclose(res.view_dirs, g["view_dirs"], rtol=RTOL_T, atol=1e-6)
    np.testing.assert_array_equal(res.ray_origins, g["ray_origins"])


def test_march_brute_vs_reference_dense_g8(orc):
    g = _load("march_g8_dense.npz")
    G, W, H, arr = int(g["G"]), int(g["W"]), int(g["H"]), int(g["arr_size"])
    res = orc.ray_march_brute(g["pose"], float(g["focal"]), float(g["aspect"]), W, H, orc.make_grid(G), arr_size=arr)
    assert res.capped == 0
    _assert_hits_match(res, g, arr)
    # first hit of every ray: the camera sits inside the grid (Q9) so t_start == 0
    first = res.t_start.reshape(W * H, arr)[:, 0]
    assert np.all(first == 0.0)


def test_march_dda_vs_reference_dense_g8(orc):
    g = _load("march_g8_dense.npz")
    G, W, H, arr = int(g["G"]), int(g["W"]), int(g["H"]), int(g["arr_size"])
    res = orc.ray_march_dda(g["pose"], float(g["focal"]), float(g["aspect"]), W, H, np.ones(G ** 3, np.uint8), G,
                            arr_size=arr)
    _assert_hits_match(res, g, arr)


def test_march_vs_reference_sparse_g16(orc):
    g = _load("march_g16_sparse.npz")
    G, W, H, arr = int(g["G"]), int(g["W"]), int(g["H"]), int(g["arr_size"])
    occ = g["occ"]
    prim_ids = np.flatnonzero(occ).astype(np.int32)
    aabb = orc.make_grid(G)[prim_ids]
    args = (g["pose"], float(g["focal"]), float(g["aspect"]), W, H)
    brute = orc.ray_march_brute(*args, aabb, prim_ids=prim_ids, arr_size=arr)
    dda = orc.ray_march_dda(*args, occ, G, arr_size=arr)
    _assert_hits_match(brute, g, arr)
    _assert_hits_match(dda, g, arr)


@pytest.mark.parametrize("G,p,wh,seed", [(8, 1.0, 24, 0), (16, 1.0, 16, 1), (32, 0.15, 24, 2), (32, 1.0, 8, 3),
                                         (64, 0.05, 24, 4)])
def test_dda_equals_brute_bit_exact(orc, G, p, wh, seed):
    """The interval sweep must reproduce the all-primitives semantics bit for bit."""
    occ = synth.occupancy_bernoulli(G, p, seed=seed) if p < 1.0 else np.ones(G ** 3, np.uint8)
    prim_ids = np.flatnonzero(occ).astype(np.int32)
    aabb = orc.make_grid(G)[prim_ids]
    focal = synth.focal_length()
    for pose in synth.poses(2, seed=100 + seed):
        brute = orc.ray_march_brute(pose, focal, 1.0, wh, wh, aabb, prim_ids=prim_ids, arr_size=3 * G)
        dda = orc.ray_march_dda(pose, focal, 1.0, wh, wh, occ, G)
        assert np.array_equal(brute.num_hits, dda.num_hits)
        assert np.array_equal(brute.hit_prims, dda.hit_prims)
        for k in ("t_start", "t_end", "start_points", "end_points"):
            assert np.array_equal(getattr(brute, k), getattr(dda, k), equal_nan=True), k


def test_dda_equals_brute_camera_outside_grid(orc):
    """Pose translation x10 puts the camera (after the reference's /10) outside [-1,1]^3."""
    G = 16
    occ = synth.occupancy_bernoulli(G, 0.4, seed=9)
    prim_ids = np.flatnonzero(occ).astype(np.int32)
    aabb = orc.make_grid(G)[prim_ids]
    pose = synth.poses(1, seed=77)[0].copy()
    pose[:3, 3] *= 10.0
    focal = synth.focal_length()
    brute = orc.ray_march_brute(pose, focal, 1.0, 20, 20, aabb, prim_ids=prim_ids, arr_size=3 * G)
    dda = orc.ray_march_dda(pose, focal, 1.0, 20, 20, occ, G)
    assert brute.num_hits.sum() > 0
    assert np.array_equal(brute.num_hits, dda.num_hits)
    assert np.array_equal(brute.hit_prims, dda.hit_prims)
    assert np.array_equal(brute.t_start, dda.t_start, equal_nan=True)


def test_march_csr_equals_dense_and_cap(orc):
    G = 16
    occ = synth.occupancy_bernoulli(G, 0.5, seed=3)
    pose = synth.poses(1, seed=5)[0]
    focal = synth.focal_length()
    dense = orc.ray_march_dda(pose, focal, 1.0, 16, 16, occ, G)
    csr = orc.ray_march_dda(pose, focal, 1.0, 16, 16, occ, G, csr=True)
    assert np.array_equal(dense.num_hits, csr.num_hits)
    idx, total = orc.exclusive_scan(dense.num_hits)
    assert total == csr.total_hits and np.array_equal(idx, csr.indices)
    assert np.array_equal(orc.compact_hits(dense.num_hits, idx, 3 * G, dense.t_start, 1), csr.t_start)
    assert np.array_equal(orc.compact_hits(dense.num_hits, idx, 3 * G, dense.end_points, 3), csr.end_points)
    capped = orc.ray_march_dda(pose, focal, 1.0, 16, 16, occ, G, max_hits=4, csr=True)
    assert np.array_equal(capped.num_hits, np.minimum(dense.num_hits, 4))
    assert capped.capped == int((dense.num_hits >= 4).sum())


def test_empty_grid_and_nonpow2(orc):
    G = 8
    pose = synth.poses(1, seed=5)[0]
    res = orc.ray_march_dda(pose, synth.focal_length(), 1.0, 8, 8, np.zeros(G ** 3, np.uint8), G)
    assert res.num_hits.sum() == 0
    with pytest.raises(ValueError):
        orc.ray_march_dda(pose, synth.focal_length(), 1.0, 8, 8, np.ones(1000, np.uint8), 10)
    # the brute-force marcher takes any resolution (reference semantics, main.cu:154-174)
    b = orc.ray_march_brute(pose, synth.focal_length(), 1.0, 8, 8, orc.make_grid(10), arr_size=30)
    assert b.num_hits.min() >= 1


# ------------------------------------------------------------------------------- sampler / vol_render
def test_sampler_vs_reference(orc):
    g = _load("sampler_volrender_s32.npz")
    for name, st in (("regular", 0), ("stratified", 1), ("uniform", 2)):
        s, t = orc.sampler(g["start_points"], g["end_points"], g["view_dirs"], g["num_hits"], g["indices"], S=32,
                           sample_type=st)
        np.testing.assert_array_equal(t, g["t_vals_" + name], err_msg=name)
        np.testing.assert_allclose(s, g["samples_" + name], rtol=RTOL_T, atol=1e-6, err_msg=name)
        assert np.array_equal(s[:, 3:], g["samples_" + name][:, 3:])
    # Q12: t_vals are segment-local (i+1)/32 in REGULAR mode
    t = g["t_vals_regular"].reshape(-1, 32)
    assert np.array_equal(t, np.tile((np.arange(32, dtype=np.float32) + 1) / 32, (t.shape[0], 1)))
    # Q14: every ray draws the same sequence, UNIFORM stores t_vals = 0
    assert np.all(g["t_vals_uniform"] == 0)


def test_volrender_vs_reference(orc):
    g = _load("sampler_volrender_s32.npz")
    px = orc.volrender(g["radiance"], g["num_hits"], g["indices"], g["t_vals_regular"], 32)
    np.testing.assert_allclose(px, g["pixels"], rtol=0, atol=1e-5)
    px4 = orc.volrender(g["radiance_s4"], g["num_hits"], g["indices"], g["t_vals_s4"], 4)
    np.testing.assert_allclose(px4, g["pixels_s4"], rtol=0, atol=1e-5)


def test_volrender_backward_vs_reference(orc):
    g = _load("sampler_volrender_s32.npz")
    for S, sfx in ((32, ""), (4, "_s4")):
        rad = g["radiance" + sfx]
        tv = g["t_vals_regular"] if S == 32 else g["t_vals_s4"]
        gr = orc.volrender_backward(g["loss_gradients"], rad, tv, g["num_hits"], g["indices"], S)
        ref = g["radiance_gradients" + sfx]
        n = int(g["num_hits"].sum()) * S
        a, b = gr[:n].astype(np.float32), ref[:n].astype(np.float32)
        # fp16 outputs: allow one half ulp-step of disagreement from expf differences
        np.testing.assert_allclose(a, b, rtol=2e-3, atol=1e-6)
        assert (gr[:n].view(np.uint16) == ref[:n].view(np.uint16)).mean() > 0.99


def test_live_reference_if_present(orc):
    """When the host build of the reference is present (oracle/_ref), re-run it on fresh seeds."""
    if orc.ref_host() is None:
        pytest.skip("oracle/_ref/libref_host.so not built")
    G = 8
    pose = synth.poses(5, seed=42)[3]
    focal = synth.focal_length()
    aabb = orc.make_grid(G)
    ref = orc.ref_ray_march(pose, focal, 1.0, 12, 12, G, aabb, 3 * G)
    mine = orc.ray_march_dda(pose, focal, 1.0, 12, 12, np.ones(G ** 3, np.uint8), G)
    assert np.array_equal(ref.num_hits, mine.num_hits)
    m = np.arange(3 * G)[None, :] < ref.num_hits[:, None]
    np.testing.assert_allclose(mine.t_end.reshape(144, -1)[m], ref.t_end.reshape(144, -1)[m], rtol=RTOL_T, atol=1e-6)


def test_dda_equals_brute_sparse_128_lego(orc):
    """the DDA oracle every 128^3 test trusts, against the all-primitives marcher on the
    benchmark's own sparse 128^3 lego grid (~1e5 AABBs), bit for bit."""
    G = 128
    occ = synth.occupancy_lego(G, seed=0)
    prim_ids = np.flatnonzero(occ).astype(np.int32)
    aabb = orc.make_grid(G)[prim_ids]
    focal = synth.focal_length()
    pose = synth.poses(3, seed=123)[2]
    W = H = 32
    brute = orc.ray_march_brute(pose, focal, 1.0, W, H, aabb, prim_ids=prim_ids, arr_size=3 * G)
    dda = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G)
    assert brute.num_hits.sum() > 10 * W * H / 4
    assert np.array_equal(brute.num_hits, dda.num_hits)
    assert np.array_equal(brute.hit_prims, dda.hit_prims)
    for k in ("t_start", "t_end", "start_points", "end_points"):
        assert np.array_equal(getattr(brute, k), getattr(dda, k), equal_nan=True), k


def test_live_reference_sparse_128_lego(orc):
    """The reference's OWN optixPrograms.cu (host build, oracle/_ref) marching the sparse 128^3 lego grid
    against the DDA oracle: hit counts exact; t / points to 1e-5 relative or 5e-6 absolute — g++ contracts
    (plane - o) / d differently from nvcc, an ulp of the ~0.5 coordinates (6e-8) over a direction component of
    ~0.02 is 3e-6, which is most of a t of 7e-5 across a cell corner."""
    if orc.ref_host() is None:
        pytest.skip("oracle/_ref/libref_host.so not built")
    G = 128
    occ = synth.occupancy_lego(G, seed=0)
    prim_ids = np.flatnonzero(occ).astype(np.int32)
    aabb = orc.make_grid(G)[prim_ids]
    focal = synth.focal_length()
    pose = synth.poses(3, seed=123)[2]
    W = H = 24
    arr = 3 * G
    ref = orc.ref_ray_march(pose, focal, 1.0, W, H, G, aabb, arr)
    mine = orc.ray_march_dda(pose, focal, 1.0, W, H, occ, G)
    assert ref.num_hits.sum() > 0
    assert np.array_equal(ref.num_hits, mine.num_hits)
    m = np.arange(arr)[None, :] < ref.num_hits[:, None]
    for k, c in (("t_start", 1), ("t_end", 1), ("start_points", 3), ("end_points", 3)):
        a = getattr(mine, k).reshape(W * H, arr, c)[m]
        b = getattr(ref, k).reshape(W * H, arr, c)[m]
        np.testing.assert_allclose(a, b, rtol=RTOL_T, atol=5e-6, err_msg=k)
