"""rtxnerf_ray_march_csr_static — the march with its scratch in a device array of the library (no cudaMalloc, no
caller workspace; include/rtxnerf.h).  No counterpart in the reference.
Checked against the existing entry point rtxnerf_ray_march_csr on the same inputs — the march is
bit-exact by contract, so the tolerance is zero except for view_dirs (atan2f, same kernel: also equal) — on calls
IN A ROW, so that the second and third call find the previous call's contents in the shared array.

The file sorts last on purpose (a failure here must not hide other tests) and passes both
alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = pytest.mark.gpu

FIELDS = ("num_hits", "indices", "view_dirs", "ray_origins", "total_hits")
HIT_FIELDS = ("t_start", "t_end", "start_points", "end_points", "hit_cells")


def _snapshot(hb):
    import torch
    torch.cuda.synchronize()
    out = {k: getattr(hb, k).cpu().numpy().copy() for k in FIELDS}
    n = min(int(out["total_hits"][0]), hb.capacity)
    for k in HIT_FIELDS:
        out[k] = getattr(hb, k)[:n].cpu().numpy().copy()
    return out


def _march_both(api, grid, pose, focal, W, H, max_hits, ray_begin=0, n_rays=None):
    """(existing entry point, new entry point) on fresh, differently poisoned output buffers"""
    import torch
    from rtx_nerf_b200 import _lib
    R = W * H - ray_begin if n_rays is None else n_rays
    look_at = torch.from_numpy(np.ascontiguousarray(pose.reshape(-1))).cuda()
    outs = []
    for which in ("workspace", "static"):
        hb = api.HitBuffers(R, R * max_hits)
        for k in FIELDS + HIT_FIELDS:
            getattr(hb, k).fill_(-7 if which == "static" else -3)
        if which == "workspace":
            api.ray_march_csr(grid, look_at, focal, W / H, W, H, hb, max_hits, ray_begin=ray_begin, n_rays=R)
        else:      # straight through the C ABI: the caller passes inputs and outputs, nothing else
            cam = _lib.Camera(look_at.data_ptr(), float(focal), float(W / H), int(W), int(H))
            st = hb.c_struct()
            rc = _lib.load().rtxnerf_ray_march_csr_static(grid._h, C.byref(cam), int(ray_begin), int(R), int(max_hits),
                                                          C.byref(st), None)
            assert rc == 0, "rtxnerf_ray_march_csr_static returned %d" % rc
        outs.append(_snapshot(hb))
    return outs


def _assert_same(ref, new):
    assert int(ref["total_hits"][0]) > 0, "test scene has no hits"
    for k in FIELDS + HIT_FIELDS:
        assert ref[k].shape == new[k].shape, k
        assert np.array_equal(ref[k], new[k]), k + " differs between rtxnerf_ray_march_csr and ..._static"


def test_static_march_matches_workspace_march_on_calls_in_a_row(cuda_api):
    import torch
    api = cuda_api
    focal = synth.focal_length()
    G = 128
    grid = api.OccupancyGrid.from_occupancy(synth.occupancy_lego(G, seed=3), G)
    poses = synth.poses(3, seed=77)
    before = torch.cuda.memory_allocated()
    # call 1: one-pass kernel (max_hits <= 32), 235 tiles -> status words + ticket of the shared array in use
    _assert_same(*_march_both(api, grid, poses[0], focal, 200, 150, 16))
    # call 2, right behind it: another pose, another tile count, a cap above 32 (count -> scan -> fill reuses the
    # same bytes as block offsets / block sums) — finds call 1's contents in the array
    _assert_same(*_march_both(api, grid, poses[1], focal, 96, 64, 48))
    # call 3: the one-pass kernel again over call 2's leftovers, ragged sub-range of a view
    _assert_same(*_march_both(api, grid, poses[2], focal, 160, 120, 32, ray_begin=1000, n_rays=160 * 120 - 1777))
    # call 4: the same inputs as call 1 once more
    _assert_same(*_march_both(api, grid, poses[0], focal, 200, 150, 16))
    grid.destroy()
    assert torch.cuda.memory_allocated() <= before + (1 << 20)      # nothing of ours stayed allocated


def test_static_march_refuses_more_rays_than_its_scratch_holds(cuda_api):
    import torch
    from rtx_nerf_b200 import _lib
    api = cuda_api
    lib = _lib.load()
    limit = int(lib.rtxnerf_ray_march_csr_static_max_rays())
    assert limit == 1 << 21
    G = 8
    grid = api.OccupancyGrid.from_occupancy(np.ones(G ** 3, np.uint8), G)
    hb = api.HitBuffers(128, 128 * 4)
    look_at = torch.from_numpy(np.ascontiguousarray(synth.poses(1, seed=1)[0].reshape(-1))).cuda()
    W, H = 4096, 1024                                                 # 2^22 pixels
    cam = _lib.Camera(look_at.data_ptr(), float(synth.focal_length()), float(W / H), W, H)
    st = hb.c_struct()
    rc = lib.rtxnerf_ray_march_csr_static(grid._h, C.byref(cam), 0, limit + 1, 4, C.byref(st), None)
    assert rc == -3, "expected RTXNERF_ECAPACITY, got %d" % rc
    with pytest.raises(Exception):
        _lib.check(rc, "ray_march_csr_static")
    grid.destroy()
