#!/usr/bin/env python3
"""SURVEY 8(d) "second GPU baseline": the reference's own sampler.cu / vol_render.cu kernels, compiled
unchanged for sm_100a (oracle/_ref/libref_cuda.so, built by oracle/build.py from the sources where they
lie), timed on the B200 next to this repo's kernels on IDENTICAL device buffers.

The reference's sampler hard-codes 32 samples per hit cell (sampler/sampler.h:4), so the workload is
S = 32 x 2 hits per ray = 64 samples/ray over 640 000 rays (N = 40.96 M samples, config 2's count).
Development/measurement script: not on the product path, reads nothing under /root/reference.
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rtx_nerf_b200 import api  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=640000)   # multiple of 1024 (the reference reads before its guard)
    ap.add_argument("--hits", type=int, default=2)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    path = os.path.join(ROOT, "oracle", "_ref", "libref_cuda.so")
    if not os.path.exists(path):
        print("oracle/_ref/libref_cuda.so not built (python oracle/build.py where /root/reference exists)")
        return 0
    ref = C.CDLL(path)
    B, S, Hr = a.rays, 32, a.hits
    H, N = B * Hr, B * Hr * S
    g = torch.Generator(device="cuda").manual_seed(11)
    nh = torch.full((B,), Hr, dtype=torch.int32, device="cuda")
    idx = (torch.arange(B, dtype=torch.int32, device="cuda") * Hr).contiguous()
    sp = torch.rand((H, 3), generator=g, device="cuda") * 2 - 1
    ep = sp + (torch.rand((H, 3), generator=g, device="cuda") - 0.5) * 0.03
    vd = torch.rand((B, 2), generator=g, device="cuda") * 6 - 3
    rad = torch.rand((N, 4), generator=g, device="cuda")
    rad[:, 3] *= 20
    lg = (torch.randn((B, 3), generator=g, device="cuda") * 0.5).half()
    t_ref, t_new = torch.empty(N, device="cuda"), torch.empty(N, device="cuda")
    s_ref, s_new = torch.empty((N, 5), device="cuda"), torch.empty((N, 5), device="cuda")
    px_ref, px_new = torch.empty((B, 3), device="cuda"), torch.empty((B, 3), device="cuda")
    g_ref = torch.zeros((N, 4), dtype=torch.float16, device="cuda")
    g_new = torch.zeros((N, 4), dtype=torch.float16, device="cuda")
    vp = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    # the reference's launchers run on the legacy default stream; ours on torch's current stream
    legacy = None

    def chk(rc):
        if rc != 0:
            raise RuntimeError("reference kernel launch failed: %d" % rc)

    rows = []
    pairs = (
        ("sampler (REGULAR)", 16 * B + 24 * H + 24 * N,
         lambda: chk(ref.ref_cuda_launchSampler(vp(sp), vp(ep), vp(vd), vp(t_ref), vp(s_ref), B, 0, vp(nh), vp(idx), 0, legacy)),
         lambda: api.launchSampler(sp, ep, vd, t_new, s_new, B, 0, nh, idx, api.SAMPLING_REGULAR, None, S)),
        ("compositing fwd", 20 * N + 8 * B + 12 * B,
         lambda: chk(ref.ref_cuda_volrender(vp(rad), vp(nh), vp(idx), vp(t_ref), B, S, vp(px_ref))),
         lambda: api.launch_volrender_cuda(None, rad, nh, idx, t_new, B, S, px_new)),
        ("compositing bwd", 20 * N + 8 * N + 14 * B,
         lambda: chk(ref.ref_cuda_volrender_backward(vp(lg), vp(rad), vp(t_ref), vp(nh), vp(idx), B, S, vp(g_ref))),
         lambda: api.launch_volrender_backward_cuda(None, lg, rad, t_new, nh, idx, B, S, g_new)),
    )
    import time
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 1.5:      # clock spin-up: a B200 out of idle needs ~1 s of load
        pairs[0][3]()
        torch.cuda.synchronize()
    print("rays %d, %d hits/ray x S=%d -> %d samples; identical device inputs" % (B, Hr, S, N))
    for name, byts, f_ref, f_new in pairs:
        with torch.cuda.stream(torch.cuda.default_stream()):
            ms_ref = timeit(f_ref)
            ms_new = timeit(f_new)
        rows.append({"kernel": name, "reference_ms": ms_ref, "this_repo_ms": ms_new, "speedup": ms_ref / ms_new,
                     "reference_gbs": byts / ms_ref / 1e6, "this_repo_gbs": byts / ms_new / 1e6})
        print("%-20s reference %8.3f ms (%6.0f GB/s)   this repo %7.3f ms (%6.0f GB/s)   x%.1f" %
              (name, ms_ref, byts / ms_ref / 1e6, ms_new, byts / ms_new / 1e6, ms_ref / ms_new))
    same = (torch.equal(s_ref, s_new) and torch.equal(t_ref, t_new)
            and torch.equal(g_ref.view(torch.int16), g_new.view(torch.int16)))
    dpx = float((px_ref - px_new).abs().max())
    print("outputs: sampler/t/backward bit-identical = %s, pixels max |diff| = %.2e" % (same, dpx))
    if a.json:
        with open(a.json, "w") as f:
            json.dump({"rays": B, "samples": N, "rows": rows, "bit_identical": bool(same), "pixel_max_abs_diff": dpx}, f, indent=1)
    return 0 if same and dpx < 2e-5 else 1


if __name__ == "__main__":
    sys.exit(main())
