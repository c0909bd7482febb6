#!/usr/bin/env python3
"""Generates the committed golden vectors under tests/golden/ by RUNNING THE REFERENCE ITSELF.

The reference has no tests or fixtures of its own, so the pin is its own code:
oracle/_ref/libref_host.so = /root/reference/{sampler/sampler.cu, vol_render/vol_render.cu,
rtx/src/optixPrograms.cu} compiled for the host by oracle/build.py (CUDA-on-host shim; the
OptiX runtime is emulated, see oracle/host_shim/optix.h).  Run in the build container (needs
/root/reference):  python tests/golden/make_golden.py
Inputs are seeded; the npz files store inputs AND the reference's outputs.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import build as obuild  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from rtx_nerf_b200 import synth  # noqa: E402


def csr_from_dense(res, arr):
    nh = res.num_hits
    idx, total = orc.exclusive_scan(nh)
    sp = orc.compact_hits(nh, idx, arr, res.start_points, 3)
    ep = orc.compact_hits(nh, idx, arr, res.end_points, 3)
    return idx, total, sp, ep


def main():
    obuild.build_all()
    assert orc.ref_host() is not None, "reference host build missing (need /root/reference)"
    focal = synth.focal_length()

    # ---- 1. the reference's literal configuration: dense 8^3 grid (main.cu:394), 16x16 view
    G, W, H = 8, 16, 16
    pose = synth.poses(3, seed=11)[1]
    aabb = orc.make_grid(G)
    r = orc.ref_ray_march(pose, focal, 1.0, W, H, G, aabb, 3 * G)
    assert r.capped == 0
    np.savez_compressed(os.path.join(HERE, "march_g8_dense.npz"), G=G, W=W, H=H, pose=pose, focal=np.float32(focal),
                        aspect=np.float32(1.0), arr_size=3 * G, num_hits=r.num_hits, t_start=r.t_start,
                        t_end=r.t_end, start_points=r.start_points, end_points=r.end_points,
                        view_dirs=r.view_dirs, ray_origins=r.ray_origins)

    # ---- 2. sparse 16^3: only occupied cells are primitives; non-square view, aspect != 1
    G2, W2, H2 = 16, 20, 12
    occ = synth.occupancy_bernoulli(G2, 0.3, seed=5)
    prim_ids = np.flatnonzero(occ).astype(np.int32)
    aabb2 = orc.make_grid(G2)[prim_ids]
    pose2 = synth.poses(4, seed=12)[2]
    r2 = orc.ref_ray_march(pose2, focal, W2 / H2, W2, H2, G2, aabb2, 3 * G2)
    assert r2.capped == 0
    np.savez_compressed(os.path.join(HERE, "march_g16_sparse.npz"), G=G2, W=W2, H=H2, pose=pose2,
                        focal=np.float32(focal), aspect=np.float32(W2 / H2), arr_size=3 * G2, occ=occ,
                        num_hits=r2.num_hits, t_start=r2.t_start, t_end=r2.t_end, start_points=r2.start_points,
                        end_points=r2.end_points, view_dirs=r2.view_dirs, ray_origins=r2.ray_origins)

    # ---- 3. sampler + vol_render on the first 64 rays of case 1 (S = 32, the reference constant)
    B = 64
    nh = r.num_hits[:B].copy()
    dense_sp = r.start_points.reshape(W * H, 3 * G, 3)[:B].reshape(-1, 3)
    dense_ep = r.end_points.reshape(W * H, 3 * G, 3)[:B].reshape(-1, 3)
    idx, total = orc.exclusive_scan(nh)
    sp = orc.compact_hits(nh, idx, 3 * G, dense_sp, 3)
    ep = orc.compact_hits(nh, idx, 3 * G, dense_ep, 3)
    vd = r.view_dirs[:B].copy()
    out = dict(B=B, num_hits=nh, indices=idx, start_points=sp, end_points=ep, view_dirs=vd)
    for name, st in (("regular", 0), ("stratified", 1), ("uniform", 2)):
        s, t = orc.ref_sampler(sp, ep, vd, nh, idx, sample_type=st, grid_res=G)
        out["samples_" + name] = s
        out["t_vals_" + name] = t
    rng = np.random.default_rng(21)
    N = total * 32
    radiance = np.empty((N, 4), np.float32)
    radiance[:, :3] = rng.random((N, 3), dtype=np.float32)
    radiance[:, 3] = rng.random(N, dtype=np.float32) * 20.0
    out["radiance"] = radiance
    out["pixels"] = orc.ref_volrender(radiance, nh, idx, out["t_vals_regular"], 32)
    lg = (rng.standard_normal((B, 3)) * 0.5).astype(np.float16)
    out["loss_gradients"] = lg
    out["radiance_gradients"] = orc.ref_volrender_backward(lg, radiance, out["t_vals_regular"], nh, idx, 32)
    # vol_render takes num_samples_per_hit at run time (vol_render.h:11): also pin S = 4 on the
    # same t values re-read with stride (first 4 samples' worth of buffers)
    N4 = total * 4
    out["radiance_s4"] = radiance[:N4].copy()
    out["t_vals_s4"] = np.tile(np.array([0.25, 0.5, 0.75, 1.0], np.float32), total)
    out["pixels_s4"] = orc.ref_volrender(out["radiance_s4"], nh, idx, out["t_vals_s4"], 4)
    out["radiance_gradients_s4"] = orc.ref_volrender_backward(lg, out["radiance_s4"], out["t_vals_s4"], nh, idx, 4)
    np.savez_compressed(os.path.join(HERE, "sampler_volrender_s32.npz"), **out)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print("%-32s %8d bytes" % (f, os.path.getsize(os.path.join(HERE, f))))


if __name__ == "__main__":
    main()
