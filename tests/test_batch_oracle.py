"""CPU tests of the oracle's ray shuffle + batch assembly (main.cu:612-676): the seeded
permutation is a bijection for every size and group, slices agree with the whole, and the batch
gather equals the reference's host loops written out naively."""
import numpy as np
import pytest


@pytest.mark.parametrize("n,group", [(1, 1), (2, 1), (3, 1), (5, 1), (16, 1), (17, 1), (1000, 1), (4096, 32),
                                     (65536, 1), (65537, 1), (640000, 32), (96, 32), (32, 32)])
def test_permutation_is_a_bijection(orc, n, group):
    p = orc.ray_permutation(n, seed=12345, group=group)
    assert p.dtype == np.int64 and p.shape == (n,)
    assert np.array_equal(np.sort(p), np.arange(n))
    if group > 1:     # groups of consecutive rays stay together and in order
        g = p.reshape(-1, group)
        assert np.all(g[:, 0] % group == 0) and np.all(np.diff(g, axis=1) == 1)


def test_one_group_is_the_identity(orc):
    """group = n: a single group, so batches are contiguous ray ranges (the per-view front end)."""
    assert orc.ray_permutation(1024, 5, 1024).tolist() == list(range(1024))
    assert orc.ray_permutation(1024, 9, 1024, 128, 256).tolist() == list(range(128, 384))


def test_permutation_slices_and_seeds(orc):
    n = 10007
    whole = orc.ray_permutation(n, seed=7)
    assert np.array_equal(orc.ray_permutation(n, 7, 1, 1234, 4321), whole[1234:1234 + 4321])
    assert orc.ray_permutation(n, 7, 1, n, 0).shape == (0,)
    other = orc.ray_permutation(n, seed=8)
    assert not np.array_equal(whole, other)
    assert (whole == np.arange(n)).mean() < 0.01          # it does shuffle
    assert np.array_equal(whole, orc.ray_permutation(n, seed=7))   # stateless: same seed, same order


def test_permutation_known_answers(orc):
    """Pinned values of this repo's permutation (its own definition: std::random_shuffle's generator is
    implementation-defined) — the device kernel is checked against the same function."""
    assert orc._splitmix64_stream(0, 2) == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4]
    assert orc.ray_permutation(8, seed=0).tolist() == [6, 0, 7, 5, 3, 2, 1, 4]
    assert orc.ray_permutation(1 << 18, seed=2024, first=0, count=4).tolist() == [109174, 183339, 42871, 245296]
    # 100 views x 800 x 800 rays in groups of 32: the last group of the shuffled order
    assert orc.ray_permutation(64000000, 99, 32, 63999968, 32)[[0, 31]].tolist() == [56545920, 56545951]


@pytest.mark.parametrize("bad", [dict(n=0), dict(n=10, group=3), dict(n=10, first=8, count=3), dict(n=10, first=-1)])
def test_permutation_argument_errors(orc, bad):
    kw = dict(n=10, seed=1, group=1, first=0, count=None)
    kw.update(bad)
    with pytest.raises(ValueError):
        orc.ray_permutation(**kw)


def _store(rng, n_views, rpv, max_hits, want_t=True):
    n = n_views * rpv
    nh = rng.integers(0, max_hits + 1, n).astype(np.int32)
    nh[rng.random(n) < 0.3] = 0
    per_view = nh.reshape(n_views, rpv)
    ind = np.concatenate([np.cumsum(v) - v for v in per_view]).astype(np.int32)
    tot = per_view.sum(1)
    gap = 5                                              # views need not be packed back to back
    base = np.cumsum(tot + gap) - (tot + gap)
    H = int(base[-1] + tot[-1])
    st = dict(rays_per_view=rpv, view_base=base.astype(np.int64), num_hits=nh, indices=ind,
              view_dirs=rng.normal(size=(n, 2)).astype(np.float32),
              start_points=rng.normal(size=(H, 3)).astype(np.float32),
              end_points=rng.normal(size=(H, 3)).astype(np.float32),
              pixels_gt=rng.random((n, 3)).astype(np.float32))
    if want_t:
        st["t_start"] = rng.random(H).astype(np.float32)
        st["t_end"] = rng.random(H).astype(np.float32)
    return st


def test_batch_assemble_matches_naive_gather(orc):
    rng = np.random.default_rng(5)
    st = _store(rng, 3, 64, 6)
    b = orc.batch_assemble(st, seed=11, group=1, first=32, batch_rays=128)
    ids = orc.ray_permutation(192, 11, 1, 32, 128)
    assert np.array_equal(b["ray_ids"], ids)
    assert np.array_equal(b["num_hits"], st["num_hits"][ids])
    assert b["total_hits"] == int(st["num_hits"][ids].sum())
    assert np.array_equal(b["indices"], np.cumsum(b["num_hits"]) - b["num_hits"])
    assert np.array_equal(b["view_dirs"], st["view_dirs"][ids])
    assert np.array_equal(b["gt_pixels"], st["pixels_gt"][ids])
    # hit records: the reference's gather (main.cu:651-660) as a flat concatenation in batch order
    src = np.concatenate([st["view_base"][r // 64] + st["indices"][r] + np.arange(st["num_hits"][r]) for r in ids]).astype(int)
    for name in ("start_points", "end_points", "t_start", "t_end"):
        assert np.array_equal(b[name], st[name][src]), name


def test_batch_assemble_capacity_and_empty(orc):
    rng = np.random.default_rng(6)
    st = _store(rng, 2, 32, 4, want_t=False)
    full = orc.batch_assemble(st, 3, 32, 0, 64)
    cut = orc.batch_assemble(st, 3, 32, 0, 64, capacity=7)
    assert cut["total_hits"] == full["total_hits"] and cut["start_points"].shape == (7, 3)
    assert np.array_equal(cut["start_points"], full["start_points"][:7])
    assert "t_start" not in full
    none = orc.batch_assemble(st, 3, 32, 64, 0)
    assert none["total_hits"] == 0 and none["num_hits"].shape == (0,)
    st["num_hits"][:] = 0                                 # a batch of rays that all miss
    st["indices"][:] = 0
    miss = orc.batch_assemble(st, 3, 1, 0, 64)
    assert miss["total_hits"] == 0 and np.all(miss["indices"] == 0)


def test_ranks_deal_disjoint_batches(orc):
    """parallel.batch_first: over an epoch the ranks' batches are disjoint slices of one shuffle."""
    from rtx_nerf_b200 import parallel
    n, B, world = 4096, 256, 3
    seen = np.zeros(n, np.int32)
    rounds = n // (B * world)
    for step in range(rounds):
        for rank in range(world):
            epoch, first = parallel.batch_first(step, rank, world, B, n)
            assert epoch == 0 and first + B <= n
            seen[orc.ray_permutation(n, 100 + epoch, 32, first, B)] += 1
    assert seen.max() == 1 and seen.sum() == rounds * world * B
    assert parallel.batch_first(rounds, 1, world, B, n) == (1, B)
    with pytest.raises(ValueError):
        parallel.batch_first(0, 0, 8, 1024, 4096)
