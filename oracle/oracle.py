"""numpy front-end of the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  It loads oracle/liboracle.so (our C restatement) and, when
present, oracle/_ref/libref_host.so (the reference's own sources compiled for the host).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)
_u16p = C.POINTER(C.c_uint16)
_u8p = C.POINTER(C.c_uint8)


class _MarchOut(C.Structure):
    _fields_ = [("t_start", _f32p), ("t_end", _f32p), ("start_points", _f32p), ("end_points", _f32p),
                ("hit_prims", _i32p), ("num_hits", _i32p), ("view_dirs", _f32p), ("ray_origins", _f32p),
                ("indices", _i32p), ("arr_size", C.c_int), ("max_hits", C.c_int)]


class _HashCfg(C.Structure):
    _fields_ = [("n_levels", C.c_int), ("n_features", C.c_int), ("log2_hashmap_size", C.c_int),
                ("base_resolution", C.c_int), ("per_level_scale", C.c_float)]


def _ptr(a, typ):
    return None if a is None else a.ctypes.data_as(typ)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            import importlib.util
            spec = importlib.util.spec_from_file_location("oracle_build", os.path.join(HERE, "build.py"))
            m = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(m)
            m.build_oracle()
        _lib = C.CDLL(path)
        _lib.orc_exclusive_scan.restype = C.c_longlong
        _lib.orc_hashgrid_offsets.restype = C.c_longlong
        _lib.orc_f2h.restype = C.c_uint16
        _lib.orc_h2f.restype = C.c_float
        _lib.orc_h2f.argtypes = [C.c_uint16]
        _lib.orc_f2h.argtypes = [C.c_float]
    return _lib


def ref_host():
    """The reference's own code compiled for the host, or None when not built."""
    global _ref
    if _ref is None:
        path = os.path.join(HERE, "_ref", "libref_host.so")
        if not os.path.exists(path):
            return None
        _ref = C.CDLL(path)
    return _ref


def default_threads():
    return os.cpu_count() or 1


# --------------------------------------------------------------------------- grid / rays
def make_grid(G):
    aabb = np.empty((G ** 3, 6), np.float32)
    lib().orc_make_grid(C.c_int(G), _ptr(aabb, _f32p))
    return aabb


def raygen(look_at, focal, aspect, W, H, x, y):
    look_at = np.ascontiguousarray(look_at, np.float32).reshape(16)
    o = np.empty(3, np.float32)
    d = np.empty(3, np.float32)
    vd = np.empty(2, np.float32)
    lib().orc_raygen(_ptr(look_at, _f32p), C.c_float(focal), C.c_float(aspect), C.c_uint(W), C.c_uint(H),
                     C.c_uint(x), C.c_uint(y), _ptr(o, _f32p), _ptr(d, _f32p), _ptr(vd, _f32p))
    return o, d, vd


class MarchResult:
    """Hit records of one view.  Dense ([R][arr_size]) or CSR (indices set)."""

    def __init__(self, R, slots, indices=None, arr_size=0, max_hits=0, want_records=True):
        self.num_hits = np.zeros(R, np.int32)
        self.view_dirs = np.empty((R, 2), np.float32)
        self.ray_origins = np.empty((R, 3), np.float32)
        self.indices = indices
        self.arr_size = arr_size
        self.max_hits = max_hits
        if want_records:
            self.t_start = np.full(slots, np.nan, np.float32)
            self.t_end = np.full(slots, np.nan, np.float32)
            self.start_points = np.full((slots, 3), np.nan, np.float32)
            self.end_points = np.full((slots, 3), np.nan, np.float32)
            self.hit_prims = np.full(slots, -1, np.int32)
        else:
            self.t_start = self.t_end = self.start_points = self.end_points = self.hit_prims = None
        self.capped = 0

    def c_struct(self):
        return _MarchOut(_ptr(self.t_start, _f32p), _ptr(self.t_end, _f32p), _ptr(self.start_points, _f32p),
                         _ptr(self.end_points, _f32p), _ptr(self.hit_prims, _i32p), _ptr(self.num_hits, _i32p),
                         _ptr(self.view_dirs, _f32p), _ptr(self.ray_origins, _f32p),
                         _ptr(self.indices, _i32p), self.arr_size, self.max_hits)


def _march(fn, look_at, focal, aspect, W, H, geom_args, arr_size, max_hits, csr, rays, nthreads):
    look_at = np.ascontiguousarray(look_at, np.float32).reshape(16)
    R = W * H
    rb, re = rays if rays is not None else (0, R)
    nthreads = nthreads or default_threads()
    cap = max_hits if max_hits > 0 else arr_size

    def call(res):
        st = res.c_struct()
        rc = fn(_ptr(look_at, _f32p), C.c_float(focal), C.c_float(aspect), C.c_uint(W), C.c_uint(H),
                *geom_args, C.c_uint(rb), C.c_uint(re), C.byref(st), C.c_int(nthreads))
        if rc < 0:
            raise ValueError("oracle march refused the grid (rc=%d)" % rc)
        res.capped = rc
        return res

    if not csr:
        return call(MarchResult(R, R * arr_size, None, arr_size, cap))
    counts = call(MarchResult(R, 0, None, arr_size, cap, want_records=False))
    indices = np.zeros(R, np.int32)
    total = lib().orc_exclusive_scan(_ptr(counts.num_hits, _i32p), C.c_int(R), _ptr(indices, _i32p))
    res = MarchResult(R, int(total), indices, arr_size, cap)
    call(res)
    res.total_hits = int(total)
    return res


def ray_march_brute(look_at, focal, aspect, W, H, aabb, prim_ids=None, arr_size=None, max_hits=0,
                    csr=False, rays=None, nthreads=None):
    """optixPrograms.cu:43-248 with every relaunch testing all primitives."""
    aabb = np.ascontiguousarray(aabb, np.float32).reshape(-1, 6)
    assert arr_size is not None, "arr_size (slots per ray) is required"
    if prim_ids is not None:
        prim_ids = np.ascontiguousarray(prim_ids, np.int32)
    geom = (_ptr(aabb, _f32p), _ptr(prim_ids, _i32p), C.c_int(aabb.shape[0]))
    return _march(lib().orc_ray_march_brute, look_at, focal, aspect, W, H, geom, arr_size, max_hits, csr,
                  rays, nthreads)


def ray_march_dda(look_at, focal, aspect, W, H, occ, G, arr_size=None, max_hits=0, csr=False, rays=None,
                  nthreads=None):
    """Same results by the exact interval sweep; occ = uint8 [G^3]."""
    occ = np.ascontiguousarray(occ, np.uint8).reshape(-1)
    assert occ.size == G ** 3
    if arr_size is None:
        arr_size = 3 * G
    geom = (_ptr(occ, _u8p), C.c_int(G))
    return _march(lib().orc_ray_march_dda, look_at, focal, aspect, W, H, geom, arr_size, max_hits, csr,
                  rays, nthreads)


def exclusive_scan(num_hits):
    num_hits = np.ascontiguousarray(num_hits, np.int32)
    idx = np.empty_like(num_hits)
    total = lib().orc_exclusive_scan(_ptr(num_hits, _i32p), C.c_int(num_hits.size), _ptr(idx, _i32p))
    return idx, int(total)


def compact_hits(num_hits, indices, arr_size, dense, comps):
    dense = np.ascontiguousarray(dense, np.float32)
    total = int(num_hits.sum())
    out = np.empty((total, comps) if comps > 1 else total, np.float32)
    lib().orc_compact_hits(_ptr(num_hits, _i32p), _ptr(indices, _i32p), C.c_int(num_hits.size),
                           C.c_int(arr_size), _ptr(dense, _f32p), C.c_int(comps), _ptr(out, _f32p))
    return out


# --------------------------------------------------------------------------- sampler / compositing
def sampler(start_points, end_points, view_dirs, num_hits, indices, S=32, sample_type=0, nthreads=None, seed=0):
    """seed = 0: the reference's own draws (sampler.cu:25,117); seed != 0: counter-based draws (f4)."""
    start_points = np.ascontiguousarray(start_points, np.float32)
    end_points = np.ascontiguousarray(end_points, np.float32)
    view_dirs = np.ascontiguousarray(view_dirs, np.float32)
    num_hits = np.ascontiguousarray(num_hits, np.int32)
    indices = np.ascontiguousarray(indices, np.int32)
    N = int(num_hits.sum()) * S
    samples = np.empty((N, 5), np.float32)
    t_vals = np.empty(N, np.float32)
    lib().orc_sampler_seeded(_ptr(start_points, _f32p), _ptr(end_points, _f32p), _ptr(view_dirs, _f32p),
                             _ptr(t_vals, _f32p), _ptr(samples, _f32p), C.c_int(num_hits.size),
                             _ptr(num_hits, _i32p), _ptr(indices, _i32p), C.c_int(S), C.c_int(sample_type),
                             C.c_ulonglong(seed), C.c_int(nthreads or default_threads()))
    return samples, t_vals


def cell_radiance(hit_prims, S, cell_rgbs):
    hit_prims = np.ascontiguousarray(hit_prims, np.int32)
    cell_rgbs = np.ascontiguousarray(cell_rgbs, np.float32)
    out = np.empty((hit_prims.size * S, 4), np.float32)
    lib().orc_cell_radiance(_ptr(hit_prims, _i32p), C.c_longlong(hit_prims.size), C.c_int(S),
                            _ptr(cell_rgbs, _f32p), _ptr(out, _f32p))
    return out


def volrender(radiance, num_hits, indices, t_vals, S, nthreads=None):
    radiance = np.ascontiguousarray(radiance, np.float32)
    t_vals = np.ascontiguousarray(t_vals, np.float32)
    num_hits = np.ascontiguousarray(num_hits, np.int32)
    indices = np.ascontiguousarray(indices, np.int32)
    pixels = np.empty((num_hits.size, 3), np.float32)
    lib().orc_volrender(_ptr(radiance, _f32p), _ptr(num_hits, _i32p), _ptr(indices, _i32p),
                        _ptr(t_vals, _f32p), C.c_int(num_hits.size), C.c_int(S), _ptr(pixels, _f32p),
                        C.c_int(nthreads or default_threads()))
    return pixels


def volrender_backward(loss_grad_h, radiance, t_vals, num_hits, indices, S, nthreads=None):
    """loss_grad_h: float16 [B][3]; returns float16 [N][4]."""
    lg = np.ascontiguousarray(loss_grad_h, np.float16).view(np.uint16)
    radiance = np.ascontiguousarray(radiance, np.float32)
    t_vals = np.ascontiguousarray(t_vals, np.float32)
    num_hits = np.ascontiguousarray(num_hits, np.int32)
    indices = np.ascontiguousarray(indices, np.int32)
    out = np.zeros((t_vals.size, 4), np.uint16)
    lib().orc_volrender_backward(_ptr(lg, _u16p), _ptr(radiance, _f32p), _ptr(t_vals, _f32p),
                                 _ptr(num_hits, _i32p), _ptr(indices, _i32p), C.c_int(num_hits.size),
                                 C.c_int(S), _ptr(out, _u16p), C.c_int(nthreads or default_threads()))
    return out.view(np.float16)


# --------------------------------------------------------------------------- encoding + MLP
def hashgrid_offsets(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale):
    cfg = _HashCfg(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale)
    off = (C.c_longlong * (n_levels + 1))()
    total = lib().orc_hashgrid_offsets(C.byref(cfg), off)
    return np.array(list(off), np.int64), int(total)


def hashgrid_geometry(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale):
    cfg = _HashCfg(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale)
    sc = np.empty(n_levels, np.float32)
    rs = np.empty(n_levels, np.uint32)
    sz = np.empty(n_levels, np.uint32)
    lib().orc_hashgrid_geometry(C.byref(cfg), _ptr(sc, _f32p), _ptr(rs, C.POINTER(C.c_uint32)),
                                _ptr(sz, C.POINTER(C.c_uint32)))
    return sc, rs, sz


def hashgrid_encode(levels, table_h, samples, nthreads=None):
    """levels: dict from rtx_nerf_b200.synth.hashgrid_levels (scale/res/size/offset arrays)."""
    L, F = levels["n_levels"], levels["n_features"]
    sc = np.ascontiguousarray(levels["scale"], np.float32)
    rs = np.ascontiguousarray(levels["res"], np.uint32)
    sz = np.ascontiguousarray(levels["size"], np.uint32)
    off = np.ascontiguousarray(levels["offset"], np.int64)
    table = np.ascontiguousarray(table_h, np.float16).view(np.uint16)
    samples = np.ascontiguousarray(samples, np.float32)
    n = samples.shape[0]
    out = np.empty((n, L * F), np.uint16)
    u32p = C.POINTER(C.c_uint32)
    lib().orc_hashgrid_encode(C.c_int(L), C.c_int(F), _ptr(sc, _f32p), _ptr(rs, u32p), _ptr(sz, u32p),
                              _ptr(off, C.POINTER(C.c_longlong)), _ptr(table, _u16p), _ptr(samples, _f32p),
                              C.c_int(samples.shape[1]), C.c_longlong(n), _ptr(out, _u16p),
                              C.c_int(nthreads or default_threads()))
    return out.view(np.float16)


def frequency_encode(samples, n_freq_xyz=10, n_freq_dir=12, padded_width=112, nthreads=None):
    """Composite Frequency encoding of [n][5] (x,y,z,theta,phi) rows -> fp16 [n][padded_width]."""
    samples = np.ascontiguousarray(samples, np.float32)
    n = samples.shape[0]
    out = np.empty((n, padded_width), np.uint16)
    lib().orc_frequency_encode(_ptr(samples, _f32p), C.c_int(samples.shape[1]), C.c_longlong(n), C.c_int(n_freq_xyz),
                               C.c_int(n_freq_dir), C.c_int(padded_width), _ptr(out, _u16p),
                               C.c_int(nthreads or default_threads()))
    return out.view(np.float16)


def mlp_forward(weights_h, in_dim, width, n_hidden, out_activation, x_h, nthreads=None):
    w = np.ascontiguousarray(weights_h, np.float16).view(np.uint16)
    x = np.ascontiguousarray(x_h, np.float16).view(np.uint16)
    n = x.shape[0]
    out16 = np.empty((n, 16), np.uint16)
    rad = np.empty((n, 4), np.float32)
    lib().orc_mlp_forward(_ptr(w, _u16p), C.c_int(in_dim), C.c_int(width), C.c_int(n_hidden),
                          C.c_int(out_activation), _ptr(x, _u16p), C.c_longlong(n), _ptr(out16, _u16p),
                          _ptr(rad, _f32p), C.c_int(nthreads or default_threads()))
    return out16.view(np.float16), rad


def mlp_backward(weights_h, out_activation, enc_h, dl_dradiance_h):
    """-> (dEnc fp16 [n][32], grad_w float64 [7168]) for the 32->64->64->16 network."""
    w = np.ascontiguousarray(weights_h, np.float16).view(np.uint16)
    enc = np.ascontiguousarray(enc_h, np.float16).view(np.uint16)
    g = np.ascontiguousarray(dl_dradiance_h, np.float16).view(np.uint16)
    n = enc.shape[0]
    d_enc = np.zeros((n, 32), np.uint16)
    gw = np.zeros(w.size, np.float64)
    lib().orc_mlp_backward(_ptr(w, _u16p), C.c_int(out_activation), _ptr(enc, _u16p), _ptr(g, _u16p),
                           C.c_longlong(n), _ptr(d_enc, _u16p), _ptr(gw, C.POINTER(C.c_double)))
    return d_enc.view(np.float16), gw


def freqnet_backward(weights_h, out_activation, enc_h, dl_dradiance_h):
    """-> grad_w float64 [131072] for the Frequency + 8 x 128 network (enc: fp16 [n][112])."""
    w = np.ascontiguousarray(weights_h, np.float16).view(np.uint16)
    enc = np.ascontiguousarray(enc_h, np.float16).view(np.uint16)
    g = np.ascontiguousarray(dl_dradiance_h, np.float16).view(np.uint16)
    gw = np.zeros(w.size, np.float64)
    lib().orc_freqnet_backward(_ptr(w, _u16p), C.c_int(out_activation), _ptr(enc, _u16p), _ptr(g, _u16p),
                               C.c_longlong(enc.shape[0]), _ptr(gw, C.POINTER(C.c_double)))
    return gw


def hashgrid_backward(levels, samples, d_enc_h):
    """-> grad_table float64 [n_entries][2]."""
    L = levels["n_levels"]
    sc = np.ascontiguousarray(levels["scale"], np.float32)
    rs = np.ascontiguousarray(levels["res"], np.uint32)
    sz = np.ascontiguousarray(levels["size"], np.uint32)
    off = np.ascontiguousarray(levels["offset"], np.int64)
    samples = np.ascontiguousarray(samples, np.float32)
    d_enc = np.ascontiguousarray(d_enc_h, np.float16).view(np.uint16)
    gt = np.zeros((int(levels["n_entries"]), 2), np.float64)
    u32p = C.POINTER(C.c_uint32)
    lib().orc_hashgrid_backward(C.c_int(L), _ptr(sc, _f32p), _ptr(rs, u32p), _ptr(sz, u32p),
                                _ptr(off, C.POINTER(C.c_longlong)), _ptr(samples, _f32p),
                                C.c_int(samples.shape[1]), C.c_longlong(samples.shape[0]), _ptr(d_enc, _u16p),
                                _ptr(gt, C.POINTER(C.c_double)))
    return gt


def l2_loss(pred, target, loss_scale=1.0):
    """tcnn L2 loss as called at main.cu:755-760: values, fp16 gradients, scalar sum."""
    pred = np.asarray(pred, np.float32).reshape(-1)
    target = np.asarray(target, np.float32).reshape(-1)
    n = np.float32(pred.size)
    diff = pred - target
    values = diff * diff / n
    grads = (np.float32(loss_scale) * np.float32(2.0) * diff / n).astype(np.float16)
    return values, grads, float(values.astype(np.float64).sum())


def adam_step(master, grads, m1, m2, grad_scale, lr, beta1, beta2, eps, step, n_encoding=0):
    """tcnn Adam as configured at main.cu:40-46; returns (master, fp16 params, m1, m2).  The first n_encoding
    parameters are encoding parameters: tiny-cuda-nn's adam_step returns early for them when their gradient is
    exactly zero (weight and both moments untouched)."""
    g_raw = np.asarray(grads, np.float32)
    g = g_raw * np.float32(grad_scale)
    n1 = np.float32(beta1) * m1 + np.float32(1 - beta1) * g
    n2 = np.float32(beta2) * m2 + np.float32(1 - beta2) * g * g
    lr_t = np.float32(lr * np.sqrt(1.0 - beta2 ** step) / (1.0 - beta1 ** step))
    new = (master - lr_t * n1 / (np.sqrt(n2) + np.float32(eps))).astype(np.float32)
    skip = np.zeros(g_raw.shape, bool)
    skip[:n_encoding] = g_raw[:n_encoding] == 0
    new = np.where(skip, np.asarray(master, np.float32), new).astype(np.float32)
    n1 = np.where(skip, m1, n1).astype(np.float32)
    n2 = np.where(skip, m2, n2).astype(np.float32)
    return new, new.astype(np.float16), n1, n2


# --------------------------------------------------------------------------- the reference itself, on the host
# ---------------------------------------------------------------------------------------
# Ray shuffle + batch assembly.  Follows the reference's training-loop front end,
# main.cu:612-676: shuffle the rays (:615), slice a batch (:621), collect num_hits / gt pixel /
# view_dir (:622-626), exclusive scan (:631-637), gather start/end points in batch order
# (:649-660).  std::random_shuffle's generator is implementation-defined, so the permutation
# itself is this repo's definition (4-round Feistel, cycle-walked); everything after it is the
# reference's gather, restated with plain loops over rays.
# ---------------------------------------------------------------------------------------
_M64 = (1 << 64) - 1


def _splitmix64_stream(seed, count):
    s, out = seed & _M64, []
    for _ in range(count):
        s = (s + 0x9E3779B97F4A7C15) & _M64
        z = s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
        out.append(z ^ (z >> 31))
    return out


def _mix32(h):
    h = h.astype(np.uint64)
    h ^= h >> np.uint64(16)
    h = (h * np.uint64(0x85EBCA6B)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(13)
    h = (h * np.uint64(0xC2B2AE35)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(16)
    return h


def ray_permutation(n, seed, group=1, first=0, count=None):
    """Ray ids at positions [first, first+count) of the seeded shuffle of n rays (int64)."""
    if count is None:
        count = n - first
    if n <= 0 or group <= 0 or n % group or first < 0 or count < 0 or first + count > n:
        raise ValueError("ray_permutation: bad arguments")
    ng = n // group
    bits = 2
    while bits < 64 and (1 << bits) < ng:
        bits += 1
    bits += bits & 1
    hb = np.uint64(bits // 2)
    mask = np.uint64((1 << (bits // 2)) - 1)
    keys = [np.uint64(k >> 32) for k in _splitmix64_stream(seed, 4)]
    pos = np.arange(first, first + count, dtype=np.uint64)
    x = pos // np.uint64(group)
    todo = np.ones(count, dtype=bool)
    while todo.any():
        xi = x[todo]
        L, R = (xi >> hb) & mask, xi & mask
        for k in keys:
            F = _mix32(R ^ k) & mask
            L, R = R, L ^ F
        xi = (L << hb) | R
        x[todo] = xi
        todo[todo] = xi >= np.uint64(ng)
    return (x * np.uint64(group) + pos % np.uint64(group)).astype(np.int64)


def batch_assemble(store, seed, group, first, batch_rays, capacity=None):
    """store: dict(rays_per_view, view_base[n_views] int64, num_hits, indices (view-local), view_dirs [n][2],
    start_points / end_points [H][3], optional t_start / t_end [H], pixels_gt [n][3]).  Returns a dict
    with the batch CSR (main.cu:622-660) in shuffled order."""
    n = len(store["num_hits"])
    ids = ray_permutation(n, seed, group, first, batch_rays)
    nh = store["num_hits"][ids].astype(np.int32)
    ind, total = exclusive_scan(nh)
    cap = total if capacity is None else min(total, int(capacity))
    out = dict(ray_ids=ids, num_hits=nh, indices=ind, total_hits=total,
               view_dirs=store["view_dirs"][ids].astype(np.float32),
               start_points=np.zeros((cap, 3), np.float32), end_points=np.zeros((cap, 3), np.float32))
    if store.get("pixels_gt") is not None:
        out["gt_pixels"] = store["pixels_gt"][ids].astype(np.float32)
    for name in ("t_start", "t_end"):
        if store.get(name) is not None:
            out[name] = np.zeros(cap, np.float32)
    for k, r in enumerate(ids):                      # the host gather loop of main.cu:651-660
        src = int(store["view_base"][r // store["rays_per_view"]]) + int(store["indices"][r])
        for j in range(int(nh[k])):
            dst = int(ind[k]) + j
            if dst >= cap:
                break
            out["start_points"][dst] = store["start_points"][src + j]
            out["end_points"][dst] = store["end_points"][src + j]
            for name in ("t_start", "t_end"):
                if name in out:
                    out[name][dst] = store[name][src + j]
    return out


def ref_ray_march(look_at, focal, aspect, W, H, G, aabb, arr_size, rays=None):
    """The reference's own OptiX programs over the emulated runtime (oracle/host_shim/optix.h)."""
    r = ref_host()
    look_at = np.ascontiguousarray(look_at, np.float32).reshape(16)
    aabb = np.ascontiguousarray(aabb, np.float32).reshape(-1, 6)
    R = W * H
    rb, re = rays if rays is not None else (0, R)
    res = MarchResult(R, R * arr_size, None, arr_size, arr_size)
    res.hit_prims = None
    rc = r.ref_optix_ray_march(_ptr(look_at, _f32p), C.c_float(focal), C.c_float(aspect), C.c_uint(W),
                               C.c_uint(H), C.c_int(G), _ptr(aabb, _f32p), C.c_int(aabb.shape[0]),
                               C.c_int(arr_size), _ptr(res.t_start, _f32p), _ptr(res.t_end, _f32p),
                               _ptr(res.start_points, _f32p), _ptr(res.end_points, _f32p),
                               _ptr(res.num_hits, _i32p), _ptr(res.view_dirs, _f32p),
                               _ptr(res.ray_origins, _f32p), C.c_uint(rb), C.c_uint(re))
    res.capped = rc
    return res


def ref_sampler(start_points, end_points, view_dirs, num_hits, indices, sample_type=0, grid_res=0):
    r = ref_host()
    S = r.ref_host_samples_per_segment()
    start_points = np.ascontiguousarray(start_points, np.float32)
    end_points = np.ascontiguousarray(end_points, np.float32)
    view_dirs = np.ascontiguousarray(view_dirs, np.float32)
    num_hits = np.ascontiguousarray(num_hits, np.int32)
    indices = np.ascontiguousarray(indices, np.int32)
    N = int(num_hits.sum()) * S
    samples = np.empty((N, 5), np.float32)
    t_vals = np.empty(N, np.float32)
    r.ref_host_launchSampler(_ptr(start_points, _f32p), _ptr(end_points, _f32p), _ptr(view_dirs, _f32p),
                             _ptr(t_vals, _f32p), _ptr(samples, _f32p), C.c_int(num_hits.size),
                             C.c_int(grid_res), _ptr(num_hits, _i32p), _ptr(indices, _i32p),
                             C.c_int(sample_type))
    return samples, t_vals


def ref_volrender(radiance, num_hits, indices, t_vals, S):
    r = ref_host()
    radiance = np.ascontiguousarray(radiance, np.float32)
    t_vals = np.ascontiguousarray(t_vals, np.float32)
    num_hits = np.ascontiguousarray(num_hits, np.int32)
    indices = np.ascontiguousarray(indices, np.int32)
    pixels = np.empty((num_hits.size, 3), np.float32)
    r.ref_host_volrender(_ptr(radiance, _f32p), _ptr(num_hits, _i32p), _ptr(indices, _i32p),
                         _ptr(t_vals, _f32p), C.c_int(num_hits.size), C.c_int(S), _ptr(pixels, _f32p))
    return pixels


def ref_volrender_backward(loss_grad_h, radiance, t_vals, num_hits, indices, S):
    r = ref_host()
    lg = np.ascontiguousarray(loss_grad_h, np.float16).view(np.uint16)
    radiance = np.ascontiguousarray(radiance, np.float32)
    t_vals = np.ascontiguousarray(t_vals, np.float32)
    num_hits = np.ascontiguousarray(num_hits, np.int32)
    indices = np.ascontiguousarray(indices, np.int32)
    out = np.zeros((t_vals.size, 4), np.uint16)
    r.ref_host_volrender_backward(_ptr(lg, _u16p), _ptr(radiance, _f32p), _ptr(t_vals, _f32p),
                                  _ptr(num_hits, _i32p), _ptr(indices, _i32p), C.c_int(num_hits.size),
                                  C.c_int(S), _ptr(out, _u16p))
    return out.view(np.float16)
