"""ctypes binding of librtxnerf_b200.so (the C ABI of include/rtxnerf.h).

There is NO fallback: if the CUDA library is missing or does not load, importing a
product entry point raises.  The library is built in-tree by rtx_nerf_b200/build.py.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librtxnerf_b200.so")

c_void_p = C.c_void_p
c_int = C.c_int
c_ll = C.c_longlong
c_float = C.c_float
c_size_t = C.c_size_t


class Params(C.Structure):
    """Mirror of rtxnerf_params == the reference's `struct Params` (rtx/include/params.h:14-42)."""
    _fields_ = [("handle", C.c_uint64), ("look_at", c_void_p), ("num_primitives", c_int),
                ("intersection_arr_size", c_int), ("aabb", c_void_p), ("start_points", c_void_p),
                ("end_points", c_void_p), ("ray_origins", c_void_p), ("t_start", c_void_p),
                ("t_end", c_void_p), ("num_hits", c_void_p), ("viewing_direction", c_void_p),
                ("focal_length", c_float), ("aspect_ratio", c_float), ("delta", c_float * 3),
                ("min_point", c_float * 3), ("max_point", c_float * 3), ("width", C.c_uint),
                ("height", C.c_uint)]


class Camera(C.Structure):
    _fields_ = [("look_at", c_void_p), ("focal_length", c_float), ("aspect_ratio", c_float),
                ("width", C.c_uint), ("height", C.c_uint)]


class Hits(C.Structure):
    _fields_ = [("num_hits", c_void_p), ("indices", c_void_p), ("view_dirs", c_void_p),
                ("ray_origins", c_void_p), ("t_start", c_void_p), ("t_end", c_void_p),
                ("start_points", c_void_p), ("end_points", c_void_p), ("hit_cells", c_void_p),
                ("total_hits", c_void_p), ("capacity", c_ll)]


class HashGrid(C.Structure):
    """rtxnerf_hashgrid (include/rtxnerf.h)."""
    _fields_ = [("table", c_void_p), ("n_levels", c_int), ("n_features", c_int), ("scale", c_float * 16),
                ("res", C.c_uint32 * 16), ("size", C.c_uint32 * 16), ("offset", C.c_uint32 * 16)]


class FreqNet(C.Structure):
    """rtxnerf_freqnet (include/rtxnerf.h)."""
    _fields_ = [("weights_packed", c_void_p), ("n_freq_xyz", c_int), ("n_freq_dir", c_int), ("width", c_int),
                ("n_hidden", c_int), ("out_activation", c_int)]


class RayStore(C.Structure):
    """rtxnerf_ray_store (include/rtxnerf.h)."""
    _fields_ = [("n_rays", c_ll), ("rays_per_view", c_ll), ("view_base", c_void_p), ("num_hits", c_void_p),
                ("indices", c_void_p), ("view_dirs", c_void_p), ("start_points", c_void_p),
                ("end_points", c_void_p), ("t_start", c_void_p), ("t_end", c_void_p), ("pixels_gt", c_void_p)]


class Mlp(C.Structure):
    """rtxnerf_mlp (include/rtxnerf.h)."""
    _fields_ = [("weights", c_void_p), ("in_dim", c_int), ("width", c_int), ("n_hidden", c_int),
                ("out_activation", c_int)]


# symbol -> (restype, argtypes); this table is also what the CPU test checks against the header
SIGNATURES = {
    "rtxnerf_abi_version": (c_int, []),
    "rtxnerf_error_string": (C.c_char_p, [c_int]),
    "rtxnerf_kernel_launches": (C.c_ulonglong, []),
    "rtxnerf_grid_create": (c_int, [c_void_p, c_int, C.POINTER(c_void_p)]),
    "rtxnerf_grid_create_from_aabbs": (c_int, [c_void_p, c_int, c_int, C.POINTER(c_void_p)]),
    "rtxnerf_grid_update": (c_int, [c_void_p, c_void_p, c_void_p]),
    "rtxnerf_grid_destroy": (c_int, [c_void_p]),
    "rtxnerf_grid_resolution": (c_int, [c_void_p]),
    "rtxnerf_grid_device_aabbs": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "rtxnerf_grid_cell_samples": (c_int, [c_int, c_void_p, c_void_p]),
    "rtxnerf_grid_update_from_density": (c_int, [c_void_p, c_void_p, c_int, c_float, c_void_p, c_float, c_void_p]),
    "rtxnerf_grid_read_bits": (c_int, [c_void_p, c_void_p]),
    "rtxnerf_write_ppm": (c_int, [C.c_char_p, c_void_p, c_int, c_int]),
    "rtxnerf_ray_march_dense": (c_int, [C.POINTER(Params), c_void_p]),
    "rtxnerf_ray_march_workspace_bytes": (c_size_t, [c_ll]),
    "rtxnerf_ray_march_workspace_bytes_for": (c_size_t, [c_ll, c_int]),
    "rtxnerf_ray_march_csr": (c_int, [c_void_p, C.POINTER(Camera), c_ll, c_ll, c_int, C.POINTER(Hits),
                                      c_void_p, c_size_t, c_void_p]),
    "rtxnerf_ray_march_csr_seg": (c_int, [c_void_p, C.POINTER(Camera), c_ll, c_ll, c_int, C.POINTER(Hits), c_void_p,
                                          c_void_p, c_size_t, c_void_p]),
    "rtxnerf_ray_march_csr_static_max_rays": (c_ll, []),
    "rtxnerf_ray_march_csr_static": (c_int, [c_void_p, C.POINTER(Camera), c_ll, c_ll, c_int, C.POINTER(Hits),
                                             c_void_p]),
    "rtxnerf_launch_sampler": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                       c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "rtxnerf_launch_sampler_seeded": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                              c_void_p, c_void_p, c_int, c_int, C.c_ulonglong, c_void_p]),
    "rtxnerf_launch_volrender": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                         c_void_p, c_void_p]),
    "rtxnerf_launch_volrender_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                                  c_int, c_int, c_void_p, c_void_p]),
    "rtxnerf_launch_volrender_backward_mode": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                                       c_int, c_int, c_void_p, c_int, c_void_p]),
    "rtxnerf_cell_radiance": (c_int, [c_void_p, c_void_p, c_ll, c_int, c_void_p, c_void_p, c_void_p]),
    "rtxnerf_hashgrid_encode": (c_int, [C.POINTER(HashGrid), c_void_p, c_int, c_void_p, c_ll, c_ll, c_void_p,
                                        c_void_p]),
    "rtxnerf_network_forward": (c_int, [C.POINTER(HashGrid), C.POINTER(Mlp), c_void_p, c_int, c_void_p, c_ll, c_ll,
                                        c_void_p, c_void_p, c_int, c_void_p]),
    "rtxnerf_network_forward_rays": (c_int, [C.POINTER(HashGrid), C.POINTER(Mlp), c_void_p, c_int, c_void_p,
                                             c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "rtxnerf_render_rays_fused": (c_int, [C.POINTER(HashGrid), C.POINTER(Mlp), c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_int, c_int, c_int, c_void_p, c_void_p]),
    "rtxnerf_render_rays_fused_seg": (c_int, [C.POINTER(HashGrid), C.POINTER(Mlp), c_void_p, c_void_p, c_void_p,
                                              c_int, c_int, c_int, c_void_p, c_void_p]),
    "rtxnerf_freqnet_param_count": (c_size_t, []),
    "rtxnerf_freqnet_packed_count": (c_size_t, []),
    "rtxnerf_freqnet_pack": (c_int, [c_void_p, c_void_p, c_void_p]),
    "rtxnerf_freqnet_encode": (c_int, [c_void_p, c_int, c_ll, c_void_p, c_void_p]),
    "rtxnerf_freqnet_forward": (c_int, [C.POINTER(FreqNet), c_void_p, c_int, c_void_p, c_ll, c_ll, c_void_p, c_void_p,
                                        c_void_p]),
    "rtxnerf_freqnet_backward": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_ll, c_ll, c_void_p,
                                         c_void_p]),
    "rtxnerf_network_backward_rays": (c_int, [C.POINTER(HashGrid), C.POINTER(Mlp), c_void_p, c_int, c_void_p,
                                              c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_ll, c_void_p,
                                              c_void_p, c_void_p, c_void_p]),
    "rtxnerf_mlp_backward": (c_int, [C.POINTER(Mlp), c_void_p, c_void_p, c_void_p, c_ll, c_ll, c_void_p, c_void_p,
                                     c_void_p]),
    "rtxnerf_hashgrid_backward_rays": (c_int, [C.POINTER(HashGrid), c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                               c_int, c_int, c_void_p, c_void_p]),
    "rtxnerf_l2_loss": (c_int, [c_void_p, c_void_p, c_ll, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "rtxnerf_adam_step": (c_int, [c_ll, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_float,
                                  c_float, c_float, c_int, c_void_p]),
    "rtxnerf_adam_step_sparse": (c_int, [c_ll, c_ll, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float,
                                         c_float, c_float, c_float, c_int, c_void_p]),
    "rtxnerf_ray_permutation": (c_int, [c_ll, C.c_ulonglong, c_int, c_ll, c_ll, c_void_p, c_void_p]),
    "rtxnerf_batch_workspace_bytes": (c_size_t, [c_ll]),
    "rtxnerf_batch_assemble": (c_int, [C.POINTER(RayStore), C.c_ulonglong, c_int, c_ll, c_ll, C.POINTER(Hits),
                                       c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "rtxnerf_batch_assemble_seg": (c_int, [C.POINTER(RayStore), C.c_ulonglong, c_int, c_ll, c_ll, C.POINTER(Hits),
                                           c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "rtxnerf_network_forward_seg": (c_int, [C.POINTER(HashGrid), C.POINTER(Mlp), c_void_p, c_void_p, c_void_p,
                                            c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "rtxnerf_hashgrid_backward_seg": (c_int, [C.POINTER(HashGrid), c_void_p, c_void_p, c_void_p, c_void_p,
                                              c_int, c_int, c_void_p, c_void_p]),
}

_lib = None


class RtxNerfError(RuntimeError):
    pass


def load():
    """Load the CUDA library (never a fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtxNerfError(
                "librtxnerf_b200.so is missing (%s): build it with `python rtx_nerf_b200/build.py` "
                "(or __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        if lib.rtxnerf_abi_version() != 1:
            raise RtxNerfError("librtxnerf_b200.so ABI version mismatch")
        _lib = lib
    return _lib


def check(code, what=""):
    if code != 0:
        msg = load().rtxnerf_error_string(code)
        raise RtxNerfError("%s failed: %s (code %d)" % (what or "rtxnerf call", msg.decode() if msg else "?", code))


def kernel_launches():
    """Kernels the library itself has launched in this process (bench.py's gpu_launches claim)."""
    return int(load().rtxnerf_kernel_launches())


def count_launches(n):
    """Kept for callers; the authoritative count is the library's own (kernel_launches())."""
    return None
