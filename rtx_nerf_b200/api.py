"""Host-side mirror of the reference's operator interface for the per-ray hot path.

Same names, argument order and meaning as the reference's C++ entry points, over the C ABI
(include/rtxnerf.h).  torch is used only for device memory and streams.

  launchSampler                    sampler/sampler.h:19-30
  launch_volrender_cuda            vol_render/vol_render.h:5-13
  launch_volrender_backward_cuda   vol_render/vol_render.h:15-25
  make_grid                        main.cu:154-174
  RTXDataHolder / Params / optixLaunch   rtx/include/rtxFunctions.h:56-100, params.h:14-42, main.cu:506-508
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import Params  # noqa: F401  (re-exported: the reference's launch-parameter block)

SAMPLING_REGULAR = 0               # sampler/sampler.h:5-9
SAMPLING_STRATIFIED_JITTERING = 1
SAMPLING_UNIFORM = 2
NUM_SAMPLES_PER_SEGMENT = 32       # sampler/sampler.h:4


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream_ptr(stream):
    if stream is None:
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if isinstance(stream, int):
        return C.c_void_p(stream)
    return C.c_void_p(stream.cuda_stream)


def _chk_dev(t, dtype, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise TypeError("%s must be a contiguous CUDA tensor of %s" % (name, dtype))


# ----------------------------------------------------------------------------- sampler
def launchSampler(d_start_points, d_end_points, d_view_dirs, d_t_vals, d_sampled_points, batch_size,
                  grid_res, d_num_hits, d_indices, sample_type=SAMPLING_REGULAR, stream=None,
                  samples_per_segment=NUM_SAMPLES_PER_SEGMENT, seed=0):
    """sampler/sampler.h:19-30.  Asynchronous on `stream`.  seed != 0: counter-based draws for the
    STRATIFIED_JITTERING / UNIFORM modes instead of the reference's repeated minstd sequence."""
    for t, dt, n in ((d_start_points, torch.float32, "d_start_points"), (d_end_points, torch.float32, "d_end_points"),
                     (d_view_dirs, torch.float32, "d_view_dirs"), (d_t_vals, torch.float32, "d_t_vals"),
                     (d_sampled_points, torch.float32, "d_sampled_points"), (d_num_hits, torch.int32, "d_num_hits"),
                     (d_indices, torch.int32, "d_indices")):
        _chk_dev(t, dt, n)
    lib = _lib.load()
    _lib.check(lib.rtxnerf_launch_sampler_seeded(_p(d_start_points), _p(d_end_points), _p(d_view_dirs),
                                                 _p(d_t_vals), _p(d_sampled_points), int(batch_size), int(grid_res),
                                                 _p(d_num_hits), _p(d_indices), int(sample_type),
                                                 int(samples_per_segment), int(seed), _stream_ptr(stream)),
               "launchSampler")
    _lib.count_launches(1 if batch_size > 0 else 0)


# ----------------------------------------------------------------------------- vol_render
def launch_volrender_cuda(network_inputs, network_outputs, num_hits, indices, ray_hit, batch_size,
                          num_samples_per_hit, pixels, stream=None):
    """vol_render/vol_render.h:5-13 (network_inputs is unused, as in the reference)."""
    for t, dt, n in ((network_outputs, torch.float32, "network_outputs"), (num_hits, torch.int32, "num_hits"),
                     (indices, torch.int32, "indices"), (ray_hit, torch.float32, "ray_hit"),
                     (pixels, torch.float32, "pixels")):
        _chk_dev(t, dt, n)
    lib = _lib.load()
    _lib.check(lib.rtxnerf_launch_volrender(_p(network_inputs), _p(network_outputs), _p(num_hits), _p(indices),
                                            _p(ray_hit), int(batch_size), int(num_samples_per_hit), _p(pixels),
                                            _stream_ptr(stream)), "launch_volrender_cuda")
    _lib.count_launches(1 if batch_size > 0 else 0)


BACKWARD_REFERENCE, BACKWARD_ANALYTIC = 0, 1


def launch_volrender_backward_cuda(loss_values, loss_gradients, sampled_points_radiance, t_hit, num_hits,
                                   indices, batch_size, num_samples_per_hit, radiance_gradients, stream=None,
                                   backward_mode=BACKWARD_REFERENCE):
    """vol_render/vol_render.h:15-25 (loss_values is unused, as in the reference).  backward_mode:
    BACKWARD_REFERENCE = the reference's own formula (vol_render.cu:118-135, the default, parity mode);
    BACKWARD_ANALYTIC = the exact gradient of the forward pass."""
    for t, dt, n in ((loss_gradients, torch.float16, "loss_gradients"),
                     (sampled_points_radiance, torch.float32, "sampled_points_radiance"),
                     (t_hit, torch.float32, "t_hit"), (num_hits, torch.int32, "num_hits"),
                     (indices, torch.int32, "indices"), (radiance_gradients, torch.float16, "radiance_gradients")):
        _chk_dev(t, dt, n)
    lib = _lib.load()
    _lib.check(lib.rtxnerf_launch_volrender_backward_mode(_p(loss_values), _p(loss_gradients),
                                                          _p(sampled_points_radiance), _p(t_hit), _p(num_hits),
                                                          _p(indices), int(batch_size), int(num_samples_per_hit),
                                                          _p(radiance_gradients), int(backward_mode),
                                                          _stream_ptr(stream)),
               "launch_volrender_backward_cuda")
    _lib.count_launches(1 if batch_size > 0 else 0)


# ----------------------------------------------------------------------------- grid + ray march
def make_grid(resolution):
    """main.cu:154-174 on the host: [res^3][6] float32 AABBs, prim = (x*res+y)*res+z."""
    L = np.float32(2.0) / np.float32(resolution)
    i = np.arange(resolution, dtype=np.float32)
    lo = (np.float32(-1.0) + i * L).astype(np.float32)
    hi = (lo + L).astype(np.float32)
    g = np.empty((resolution, resolution, resolution, 6), np.float32)
    g[..., 0] = lo[:, None, None]
    g[..., 1] = lo[None, :, None]
    g[..., 2] = lo[None, None, :]
    g[..., 3] = hi[:, None, None]
    g[..., 4] = hi[None, :, None]
    g[..., 5] = hi[None, None, :]
    return g.reshape(-1, 6)


class OccupancyGrid:
    """Device occupancy bitfield + handle.  Replaces the GAS built by
    RTXDataHolder::initAccelerationStructure (rtx/src/rtxFunctions.cpp:293-351)."""

    def __init__(self, handle, G):
        self._h = handle
        self.G = G

    @classmethod
    def from_bits(cls, occ_bits, G):
        bits = np.ascontiguousarray(occ_bits, np.uint32)
        if bits.size != (G ** 3 + 31) // 32:
            raise ValueError("occ_bits must hold G^3 bits")
        h = C.c_void_p()
        _lib.check(_lib.load().rtxnerf_grid_create(bits.ctypes.data_as(C.c_void_p), int(G), C.byref(h)), "grid_create")
        return cls(h, G)

    @classmethod
    def from_occupancy(cls, occ_u8, G):
        from .synth import pack_bits
        return cls.from_bits(pack_bits(occ_u8), G)

    @classmethod
    def from_aabbs(cls, aabbs, G):
        a = np.ascontiguousarray(aabbs, np.float32).reshape(-1, 6)
        h = C.c_void_p()
        _lib.check(_lib.load().rtxnerf_grid_create_from_aabbs(a.ctypes.data_as(C.c_void_p), a.shape[0], int(G),
                                                              C.byref(h)), "grid_create_from_aabbs")
        return cls(h, G)

    @property
    def handle(self):
        return self._h.value

    def update(self, occ_bits, stream=None):
        bits = np.ascontiguousarray(occ_bits, np.uint32)
        _lib.check(_lib.load().rtxnerf_grid_update(self._h, bits.ctypes.data_as(C.c_void_p), _stream_ptr(stream)),
                   "grid_update")

    def cell_samples(self, rows=None, stream=None):
        """[G^3][5] fp32 rows at the cell centres, primitive order: the network's input for an update."""
        if rows is None:
            rows = torch.empty((self.G ** 3, 5), dtype=torch.float32, device="cuda")
        _lib.check(_lib.load().rtxnerf_grid_cell_samples(int(self.G), _p(rows), _stream_ptr(stream)), "grid_cell_samples")
        return rows

    def update_from_density(self, sigma, threshold, decay=0.0, ema=None, stream=None):
        """Device-side rebuild: sigma is a float tensor whose element [p*stride] is the
        density of cell p — a [G^3] vector or the [:, 3] column view of the network's [G^3][4]
        radiance.  ema: optional [G^3] running maximum, ema = max(ema*decay, sigma)."""
        if sigma.dtype != torch.float32 or not sigma.is_cuda:
            raise TypeError("sigma must be a float32 CUDA tensor")
        stride = sigma.stride(0) if sigma.dim() == 1 else None
        if stride is None or sigma.shape[0] != self.G ** 3:
            raise ValueError("sigma must be a 1-D (possibly strided) view of G^3 densities")
        _lib.check(_lib.load().rtxnerf_grid_update_from_density(self._h, sigma.data_ptr(), int(stride), float(decay),
                                                                _p(ema), float(threshold), _stream_ptr(stream)),
                   "grid_update_from_density")

    def read_bits(self):
        out = np.empty((self.G ** 3 + 31) // 32, np.uint32)
        _lib.check(_lib.load().rtxnerf_grid_read_bits(self._h, out.ctypes.data_as(C.c_void_p)), "grid_read_bits")
        return out

    def device_aabbs(self):
        p = C.c_void_p()
        _lib.check(_lib.load().rtxnerf_grid_device_aabbs(self._h, C.byref(p)), "grid_device_aabbs")
        return p.value

    def destroy(self):
        if self._h is not None and self._h.value:
            _lib.load().rtxnerf_grid_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class RTXDataHolder:
    """Interface mirror of the reference's OptiX host wrapper (rtx/include/rtxFunctions.h:56-100).
    The OptiX context / module / pipeline / SBT steps have nothing to build on a GPU without RT
    cores and are kept as no-ops so the reference's call sequence (main.cu:381-399) still reads."""

    def __init__(self):
        self.gas_handle = 0
        self.pipeline_ray_march = "ray_march"
        self.sbt_ray_march = None
        self.grid = None

    def initContext(self):
        _lib.load()

    def createModule(self, ptx_filename=None):
        pass

    def createProgramGroups(self):
        pass

    def linkPipeline(self, debug=False):
        pass

    def buildSBT(self):
        pass

    def initAccelerationStructure(self, grid_aabbs, resolution=None):
        """Takes the host AABB list (std::vector<OptixAabb>), returns the DEVICE AABB array
        pointer the caller stores in Params.aabb (main.cu:399,492)."""
        a = np.ascontiguousarray(grid_aabbs, np.float32).reshape(-1, 6)
        if resolution is None:
            resolution = int(round(2.0 / float(a[0, 3] - a[0, 0])))
        self.grid = OccupancyGrid.from_aabbs(a, resolution)
        self.gas_handle = self.grid.handle
        return self.grid.device_aabbs()


def optixLaunch(pipeline, stream, params, params_size=None, sbt=None, width=None, height=None, depth=1):
    """main.cu:506-508.  `params` is the host copy of the launch block (the reference passes a
    device copy; the new kernels take their arguments by value).  Dense output layout."""
    if width is not None:
        params.width = width
    if height is not None:
        params.height = height
    _lib.check(_lib.load().rtxnerf_ray_march_dense(C.byref(params), _stream_ptr(stream)), "optixLaunch(ray_march)")
    _lib.count_launches(1)


class HitBuffers:
    """Caller-owned CSR hit records for up to n_rays rays and `capacity` hits."""

    def __init__(self, n_rays, capacity, device="cuda", want_t=True, want_cells=True, want_origins=True,
                 want_segments=False):
        dev = torch.device(device)
        self.n_rays, self.capacity = n_rays, capacity
        self.num_hits = torch.empty(n_rays, dtype=torch.int32, device=dev)
        self.indices = torch.empty(n_rays, dtype=torch.int32, device=dev)
        self.view_dirs = torch.empty((n_rays, 2), dtype=torch.float32, device=dev)
        self.ray_origins = torch.empty((n_rays, 3), dtype=torch.float32, device=dev) if want_origins else None
        self.t_start = torch.empty(capacity, dtype=torch.float32, device=dev) if want_t else None
        self.t_end = torch.empty(capacity, dtype=torch.float32, device=dev) if want_t else None
        self.start_points = torch.empty((capacity, 3), dtype=torch.float32, device=dev)
        self.end_points = torch.empty((capacity, 3), dtype=torch.float32, device=dev)
        self.hit_cells = torch.empty(capacity, dtype=torch.int32, device=dev) if want_cells else None
        self.total_hits = torch.zeros(1, dtype=torch.int64, device=dev)
        # packed (sx, sy, sz, ex, ey, ez, 0, 0) per hit for the fused render kernel (rtxnerf_ray_march_csr_seg)
        self.segments8 = torch.empty((capacity, 8), dtype=torch.float32, device=dev) if want_segments else None
        ws = _lib.load().rtxnerf_ray_march_workspace_bytes(n_rays)
        self.workspace = torch.empty(ws, dtype=torch.uint8, device=dev)

    def c_struct(self, points=True):
        return _lib.Hits(_p(self.num_hits), _p(self.indices), _p(self.view_dirs), _p(self.ray_origins),
                         _p(self.t_start), _p(self.t_end), _p(self.start_points) if points else None,
                         _p(self.end_points) if points else None, _p(self.hit_cells), _p(self.total_hits),
                         self.capacity)


def ray_march_csr(grid, look_at_dev, focal_length, aspect_ratio, width, height, hits, max_hits, ray_begin=0,
                  n_rays=None, stream=None, points=True):
    """Ray-grid intersection straight to CSR: replaces optixLaunch + D2H + thrust reduce/scan +
    host gather (main.cu:506-543,631-676).  Asynchronous.  With HitBuffers(want_segments=True) the packed 32-byte
    hit records are written too (ray_march_csr_seg); points=False then leaves start_points / end_points out."""
    _chk_dev(look_at_dev, torch.float32, "look_at")
    if n_rays is None:
        n_rays = width * height - ray_begin
    if n_rays > hits.n_rays:
        raise ValueError("HitBuffers too small for %d rays" % n_rays)
    cam = _lib.Camera(_p(look_at_dev), float(focal_length), float(aspect_ratio), int(width), int(height))
    has_seg = getattr(hits, "segments8", None) is not None
    st = hits.c_struct(points=points or not has_seg)
    need = _lib.load().rtxnerf_ray_march_workspace_bytes_for(int(n_rays), int(max_hits))
    if hits.workspace.numel() < need:      # first march with a cap above 32: room for the staged single pass
        hits.workspace = torch.empty(need, dtype=torch.uint8, device=hits.workspace.device)
    if has_seg:
        _lib.check(_lib.load().rtxnerf_ray_march_csr_seg(grid._h, C.byref(cam), int(ray_begin), int(n_rays),
                                                         int(max_hits), C.byref(st), _p(hits.segments8),
                                                         _p(hits.workspace), hits.workspace.numel(),
                                                         _stream_ptr(stream)), "ray_march_csr_seg")
    else:
        _lib.check(_lib.load().rtxnerf_ray_march_csr(grid._h, C.byref(cam), int(ray_begin), int(n_rays), int(max_hits),
                                                     C.byref(st), _p(hits.workspace), hits.workspace.numel(),
                                                     _stream_ptr(stream)), "ray_march_csr")
    _lib.count_launches(3 if n_rays > 0 else 0)


def ray_march_csr_static(grid, look_at_dev, focal_length, aspect_ratio, width, height, hits, max_hits, ray_begin=0,
                         n_rays=None, stream=None):
    """ray_march_csr through rtxnerf_ray_march_csr_static: no device allocation, the scratch is a device array
    inside the library (at most rtxnerf_ray_march_csr_static_max_rays() rays per call; calls that overlap on the
    GPU race on it).  No counterpart in the reference (include/rtxnerf.h).  `hits.workspace` is not used."""
    _chk_dev(look_at_dev, torch.float32, "look_at")
    if n_rays is None:
        n_rays = width * height - ray_begin
    if n_rays > hits.n_rays:
        raise ValueError("HitBuffers too small for %d rays" % n_rays)
    cam = _lib.Camera(_p(look_at_dev), float(focal_length), float(aspect_ratio), int(width), int(height))
    st = hits.c_struct()
    _lib.check(_lib.load().rtxnerf_ray_march_csr_static(grid._h, C.byref(cam), int(ray_begin), int(n_rays),
                                                        int(max_hits), C.byref(st), _stream_ptr(stream)),
               "ray_march_csr_static")
    _lib.count_launches(3 if n_rays > 0 else 0)


def cell_radiance(hit_cells, total_hits, capacity_hits, samples_per_segment, cell_rgbs, radiance, stream=None):
    """BASELINE config 1: fixed per-cell (r,g,b,sigma) stands in for the network."""
    _lib.check(_lib.load().rtxnerf_cell_radiance(_p(hit_cells), _p(total_hits), int(capacity_hits),
                                                 int(samples_per_segment), _p(cell_rgbs), _p(radiance),
                                                 _stream_ptr(stream)), "cell_radiance")
    _lib.count_launches(1 if capacity_hits > 0 else 0)


# ----------------------------------------------------------------------------- ray store / batches
class RayStore:
    """March results of ALL training views resident in HBM (replaces the host `ray_payloads` vector of
    main.cu:509-541): per-ray num_hits / view-local CSR index / view_dir / gt pixel, per-hit entry and
    exit points, one hit base per view.  Filled view by view with ray_march_csr (`add_view`)."""

    def __init__(self, n_views, rays_per_view, capacity_hits, device="cuda", want_t=False, pixels_gt=None):
        dev = torch.device(device)
        self.n_views, self.rays_per_view = int(n_views), int(rays_per_view)
        self.n_rays = self.n_views * self.rays_per_view
        self.capacity = int(capacity_hits)
        self.num_hits = torch.zeros(self.n_rays, dtype=torch.int32, device=dev)
        self.indices = torch.zeros(self.n_rays, dtype=torch.int32, device=dev)
        self.view_dirs = torch.zeros((self.n_rays, 2), dtype=torch.float32, device=dev)
        self.start_points = torch.empty((self.capacity, 3), dtype=torch.float32, device=dev)
        self.end_points = torch.empty((self.capacity, 3), dtype=torch.float32, device=dev)
        self.t_start = torch.empty(self.capacity, dtype=torch.float32, device=dev) if want_t else None
        self.t_end = torch.empty(self.capacity, dtype=torch.float32, device=dev) if want_t else None
        self.pixels_gt = pixels_gt
        if pixels_gt is not None:
            _chk_dev(pixels_gt, torch.float32, "pixels_gt")
            if pixels_gt.numel() != self.n_rays * 3:
                raise ValueError("pixels_gt must be [n_rays][3]")
        self.view_base = torch.zeros(self.n_views, dtype=torch.int64, device=dev)
        self.view_base_host = [0] * self.n_views
        self.used_hits = 0
        self.views_filled = 0
        self._total = torch.zeros(1, dtype=torch.int64, device=dev)
        self._workspace = torch.empty(_lib.load().rtxnerf_ray_march_workspace_bytes(self.rays_per_view),
                                      dtype=torch.uint8, device=dev)

    def add_view(self, grid, look_at_dev, focal_length, aspect_ratio, width, height, max_hits):
        """Marches the next view straight into the store's slices.  Synchronises once (reads the
        view's hit total to place the next view): dataset preparation, not the per-step path."""
        if width * height != self.rays_per_view:
            raise ValueError("view size does not match rays_per_view")
        v = self.views_filled
        if v >= self.n_views:
            raise ValueError("ray store is full")
        r0, h0 = v * self.rays_per_view, self.used_hits
        room = self.capacity - h0
        cam = _lib.Camera(_p(look_at_dev), float(focal_length), float(aspect_ratio), int(width), int(height))
        sl = lambda t, a: None if t is None else C.c_void_p(t.data_ptr() + a * t.element_size())
        st = _lib.Hits(sl(self.num_hits, r0), sl(self.indices, r0), sl(self.view_dirs, 2 * r0), None,
                       sl(self.t_start, h0), sl(self.t_end, h0), sl(self.start_points, 3 * h0),
                       sl(self.end_points, 3 * h0), None, _p(self._total), room)
        need = _lib.load().rtxnerf_ray_march_workspace_bytes_for(self.rays_per_view, int(max_hits))
        if self._workspace.numel() < need:
            self._workspace = torch.empty(need, dtype=torch.uint8, device=self._workspace.device)
        _lib.check(_lib.load().rtxnerf_ray_march_csr(grid._h, C.byref(cam), 0, self.rays_per_view, int(max_hits),
                                                     C.byref(st), _p(self._workspace), self._workspace.numel(),
                                                     _stream_ptr(None)), "ray_march_csr (ray store)")
        total = int(self._total.item())
        if total > room:
            raise _lib.RtxNerfError("ray store capacity exceeded: view %d needs %d hit slots, %d left" % (v, total, room))
        self.view_base_host[v] = h0
        self.view_base[v] = h0
        self.used_hits = h0 + total
        self.views_filled = v + 1
        _lib.count_launches(3)
        return total

    def c_struct(self):
        return _lib.RayStore(self.n_rays, self.rays_per_view, _p(self.view_base), _p(self.num_hits), _p(self.indices),
                             _p(self.view_dirs), _p(self.start_points), _p(self.end_points), _p(self.t_start),
                             _p(self.t_end), _p(self.pixels_gt))


def ray_permutation(n, seed, group=1, first=0, count=None, out=None, stream=None):
    """Ray ids at positions [first, first+count) of the seeded shuffle (int64, device): the device
    stand-in for std::random_shuffle(ray_payloads) at main.cu:615."""
    count = n - first if count is None else count
    if out is None:
        out = torch.empty(count, dtype=torch.int64, device="cuda")
    _chk_dev(out, torch.int64, "out")
    _lib.check(_lib.load().rtxnerf_ray_permutation(int(n), int(seed), int(group), int(first), int(count), _p(out),
                                                   _stream_ptr(stream)), "ray_permutation")
    _lib.count_launches(1 if count > 0 else 0)
    return out


def batch_assemble(store, seed, group, first, batch_rays, hits, gt_pixels=None, ray_ids=None, workspace=None,
                   stream=None, points=True):
    """One training batch from the ray store: replaces main.cu:621-676 (host slice + collect +
    H2D + thrust reduce/scan + host gather + H2D).  `hits` is a HitBuffers that receives the batch
    CSR (num_hits, indices, view_dirs, start/end points, t's, total_hits).  Asynchronous."""
    if store.views_filled != store.n_views:
        raise ValueError("ray store holds %d of %d views" % (store.views_filled, store.n_views))
    if batch_rays > hits.n_rays:
        raise ValueError("HitBuffers too small for %d rays" % batch_rays)
    lib = _lib.load()
    need = lib.rtxnerf_batch_workspace_bytes(int(batch_rays))
    if workspace is None:
        workspace = hits.workspace
        if workspace.numel() < need:
            workspace = hits.workspace = torch.empty(need, dtype=torch.uint8, device=hits.num_hits.device)
    st = store.c_struct()
    if getattr(hits, "segments8", None) is not None:      # packed records too (or alone: points=False)
        hs = hits.c_struct(points=points)
        _lib.check(lib.rtxnerf_batch_assemble_seg(C.byref(st), int(seed), int(group), int(first), int(batch_rays),
                                                  C.byref(hs), _p(hits.segments8), _p(gt_pixels), _p(ray_ids),
                                                  _p(workspace), workspace.numel(), _stream_ptr(stream)),
                   "batch_assemble_seg")
    else:
        hs = hits.c_struct()
        _lib.check(lib.rtxnerf_batch_assemble(C.byref(st), int(seed), int(group), int(first), int(batch_rays),
                                              C.byref(hs), _p(gt_pixels), _p(ray_ids), _p(workspace), workspace.numel(),
                                              _stream_ptr(stream)), "batch_assemble")
    _lib.count_launches(3 if batch_rays > 0 else 0)


# ----------------------------------------------------------------------------- network
IMPL_TCGEN05 = 0
IMPL_SIMT = 1


def write_ppm(path, pixels, width, height):
    """Headless image output: [height*width][3] fp32 RGB (host or device) -> binary PPM."""
    px = pixels.detach().cpu().numpy() if hasattr(pixels, "detach") else np.asarray(pixels)
    px = np.ascontiguousarray(px, np.float32).reshape(-1)
    if px.size != width * height * 3:
        raise ValueError("pixels must hold width*height RGB triples")
    _lib.check(_lib.load().rtxnerf_write_ppm(str(path).encode(), px.ctypes.data_as(C.c_void_p), int(width), int(height)),
               "write_ppm")


class FrequencyNetwork:
    """The network main.cu itself configures (main.cu:47-68): Composite Frequency encoding
    (10 frequencies on xyz, tiny-cuda-nn's default 12 on theta/phi, padded with ones to 112) +
    FullyFusedMLP 8 x 128 ReLU, sigmoid output, driven like network->forward(stream, input,
    &output) (main.cu:721).  tcgen05 kernel with the activations resident in TMEM
    (network_freq.cu)."""

    n_input_dims = 5
    n_output_dims = 4
    IN, WIDTH, N_HIDDEN = 112, 128, 8

    def __init__(self, weights_h, out_activation=1, device="cuda"):
        w = np.ascontiguousarray(weights_h, np.float16).reshape(-1)
        n = int(_lib.load().rtxnerf_freqnet_param_count())
        if w.size != n:
            raise ValueError("FrequencyNetwork needs %d fp16 weights, got %d" % (n, w.size))
        self.weights = torch.as_tensor(w).to(device)
        self.packed = torch.empty(int(_lib.load().rtxnerf_freqnet_packed_count()), dtype=torch.float16, device=device)
        self.out_activation = int(out_activation)
        self.repack()

    def repack(self, stream=None):
        """Row-major weights -> tensor-core operand layout (after construction / every optimizer step)."""
        _lib.check(_lib.load().rtxnerf_freqnet_pack(_p(self.weights), _p(self.packed), _stream_ptr(stream)),
                   "freqnet_pack")
        self._net = _lib.FreqNet(self.packed.data_ptr(), 10, 12, self.WIDTH, self.N_HIDDEN, self.out_activation)

    def padded_output_width(self):
        return 16            # main.cu:715

    def n_params(self):
        return self.weights.numel()

    def encode(self, samples, n=None, encoded=None, stream=None):
        n = samples.shape[0] if n is None else n
        if encoded is None:
            encoded = torch.empty((n, self.IN), dtype=torch.float16, device=samples.device)
        _lib.check(_lib.load().rtxnerf_freqnet_encode(_p(samples), samples.shape[1], int(n), _p(encoded),
                                                      _stream_ptr(stream)), "freqnet_encode")
        return encoded

    def forward(self, stream, input, output=None, radiance=None, n=None, d_count=None, count_scale=1):
        """input: fp32 [N][>=5] sample rows (GPUMatrix<float>(5,N)); output: fp16 [N][16]
        (GPUMatrix<half>(16,N)); radiance: fp32 [N][4] (rows 0..3 as float, main.cu:723-728)."""
        n_max = input.shape[0] if n is None else n
        if output is None and radiance is None:
            output = torch.empty((n_max, 16), dtype=torch.float16, device=input.device)
        _lib.check(_lib.load().rtxnerf_freqnet_forward(C.byref(self._net), _p(input), input.shape[1], _p(d_count),
                                                       int(count_scale), int(n_max), _p(output), _p(radiance),
                                                       _stream_ptr(stream)), "freqnet_forward")
        return output if output is not None else radiance


def _freqnet_backward_method():
    def backward(self, stream, input, dl_dradiance, grad_weights, n=None, d_count=None, count_scale=1):
        """network->backward (main.cu:781) for the Frequency + 8 x 128 network: fp32 weight gradients
        accumulated into grad_weights [n_params]; input as forward(), dl_dradiance fp16 [N][4]."""
        n_max = input.shape[0] if n is None else n
        _lib.check(_lib.load().rtxnerf_freqnet_backward(_p(self.weights), self.out_activation, _p(input),
                                                        input.shape[1], _p(dl_dradiance), _p(d_count),
                                                        int(count_scale), int(n_max), _p(grad_weights),
                                                        _stream_ptr(stream)), "freqnet_backward")

    FrequencyNetwork.backward = backward


_freqnet_backward_method()


class Network:
    """Encoding + fully fused MLP: the object the reference gets from tcnn::create_from_config
    (main.cu:325) and drives with network->forward(stream, input, &output) (main.cu:721).
    Configuration of BASELINE configs 2-5: 16-level hash grid (F=2) -> 2 x 64 ReLU -> 16 outputs,
    sigmoid output activation (main.cu:63), fp16 parameters."""

    n_input_dims = 5     # x, y, z, theta, phi (main.cu:322)
    n_output_dims = 4    # r, g, b, sigma      (main.cu:323)

    def __init__(self, levels, table_h, weights_h, in_dim=32, width=64, n_hidden=2, out_activation=1,
                 device="cuda"):
        self.levels = levels
        # one flat fp16 parameter buffer [table | weights] (the reference keeps all tcnn parameters
        # in one buffer too, main.cu:328-342): a training step all-reduces / updates it in one go
        t = np.ascontiguousarray(table_h, np.float16).reshape(-1)
        w = np.ascontiguousarray(weights_h, np.float16).reshape(-1)
        self.params = torch.as_tensor(np.concatenate([t, w])).to(device)
        self.table = self.params[:t.size]
        self.weights = self.params[t.size:]
        self.in_dim, self.width, self.n_hidden, self.out_activation = in_dim, width, n_hidden, out_activation
        hg = _lib.HashGrid()
        hg.table = self.table.data_ptr()
        hg.n_levels, hg.n_features = levels["n_levels"], levels["n_features"]
        for l in range(levels["n_levels"]):
            hg.scale[l] = float(levels["scale"][l])
            hg.res[l] = int(levels["res"][l])
            hg.size[l] = int(levels["size"][l])
            hg.offset[l] = int(levels["offset"][l])
        self._hg = hg
        self._mlp = _lib.Mlp(self.weights.data_ptr(), in_dim, width, n_hidden, out_activation)

    def padded_output_width(self):
        return 16            # main.cu:715

    def n_params(self):
        return self.weights.numel() + self.table.numel()

    def encode(self, samples, n=None, d_count=None, count_scale=1, encoded=None, stream=None):
        n_max = samples.shape[0] if n is None else n
        if encoded is None:
            encoded = torch.empty((n_max, self.levels["n_levels"] * 2), dtype=torch.float16, device=samples.device)
        _lib.check(_lib.load().rtxnerf_hashgrid_encode(C.byref(self._hg), _p(samples), samples.shape[1], _p(d_count),
                                                       int(count_scale), int(n_max), _p(encoded),
                                                       _stream_ptr(stream)), "hashgrid_encode")
        _lib.count_launches(1 if n_max > 0 else 0)
        return encoded

    def forward(self, stream, input, output=None, radiance=None, n=None, d_count=None, count_scale=1,
                impl=IMPL_TCGEN05, pre_encoded=False):
        """input: fp32 [N][5] sample rows (GPUMatrix<float>(5,N)) or, with pre_encoded, fp16 [N][32].
        output: fp16 [N][16] (GPUMatrix<half>(16,N)); radiance: fp32 [N][4] (rows 0..3 as float)."""
        n_max = input.shape[0] if n is None else n
        if output is None and radiance is None:
            output = torch.empty((n_max, 16), dtype=torch.float16, device=input.device)
        enc = None if pre_encoded else C.byref(self._hg)
        _lib.check(_lib.load().rtxnerf_network_forward(enc, C.byref(self._mlp), _p(input), input.shape[1],
                                                       _p(d_count), int(count_scale), int(n_max), _p(output),
                                                       _p(radiance), int(impl), _stream_ptr(stream)),
                   "network_forward")
        _lib.count_launches(1 if n_max > 0 else 0)
        return output if output is not None else radiance

    def forward_rays(self, stream, samples, num_hits, indices, n_rays, samples_per_segment, output=None,
                     radiance=None, encoded_out=None):
        """Ray-coherent schedule of forward() over the sampler's CSR sample buffer: same rows,
        same results, gathers coalesced across neighbouring rays (network_rays.cu)."""
        if output is None and radiance is None:
            raise ValueError("forward_rays needs an output and/or radiance buffer")
        _lib.check(_lib.load().rtxnerf_network_forward_rays(C.byref(self._hg), C.byref(self._mlp), _p(samples),
                                                            samples.shape[1], _p(num_hits), _p(indices),
                                                            int(n_rays), int(samples_per_segment), _p(output),
                                                            _p(radiance), _p(encoded_out), _stream_ptr(stream)),
                   "network_forward_rays")
        _lib.count_launches(1 if n_rays > 0 else 0)
        return output if output is not None else radiance

    def render_rays_fused(self, stream, start_points, end_points, num_hits, indices, n_rays, samples_per_segment,
                          pixels, sample_type=0):
        """launchSampler + forward + launch_volrender_cuda of the render path in one kernel
        (view_fused.cu): CSR segments in, pixels out, no sample / radiance buffers."""
        _lib.check(_lib.load().rtxnerf_render_rays_fused(C.byref(self._hg), C.byref(self._mlp), _p(start_points),
                                                         _p(end_points), _p(num_hits), _p(indices), int(n_rays),
                                                         int(samples_per_segment), int(sample_type), _p(pixels),
                                                         _stream_ptr(stream)), "render_rays_fused")
        return pixels

    def render_rays_fused_seg(self, stream, segments8, num_hits, indices, n_rays, samples_per_segment, pixels,
                              sample_type=0):
        """render_rays_fused over the packed hit records of ray_march_csr_seg ([capacity][8] floats)."""
        _lib.check(_lib.load().rtxnerf_render_rays_fused_seg(C.byref(self._hg), C.byref(self._mlp), _p(segments8),
                                                             _p(num_hits), _p(indices), int(n_rays),
                                                             int(samples_per_segment), int(sample_type), _p(pixels),
                                                             _stream_ptr(stream)), "render_rays_fused_seg")
        return pixels

    def backward_rays(self, stream, samples, encoded, dl_dradiance, num_hits, indices, n_rays,
                      samples_per_segment, d_count, n_max, dencoded, grad_table, grad_weights):
        """network->backward (main.cu:781) with the intended wiring (Q5): dl_dradiance is the fp16
        [N][4] buffer of launch_volrender_backward_cuda.  Gradients accumulate (fp32)."""
        _lib.check(_lib.load().rtxnerf_network_backward_rays(
            C.byref(self._hg), C.byref(self._mlp), _p(samples), samples.shape[1], _p(encoded), _p(dl_dradiance),
            _p(num_hits), _p(indices), int(n_rays), int(samples_per_segment), _p(d_count), int(n_max), _p(dencoded),
            _p(grad_table), _p(grad_weights), _stream_ptr(stream)), "network_backward_rays")
        _lib.count_launches(2 if n_rays > 0 and n_max > 0 else 0)


def _network_stage_methods():
    def mlp_backward(self, stream, encoded, dl_dradiance, samples_per_segment, d_count, n_max, dencoded, grad_weights):
        """First stage of backward_rays alone: MLP dgrad + weight gradients (fp32, accumulated)."""
        _lib.check(_lib.load().rtxnerf_mlp_backward(C.byref(self._mlp), _p(encoded), _p(dl_dradiance), _p(d_count),
                                                    int(samples_per_segment), int(n_max), _p(dencoded),
                                                    _p(grad_weights), _stream_ptr(stream)), "mlp_backward")

    def hashgrid_backward_rays(self, stream, samples, dencoded, num_hits, indices, n_rays, samples_per_segment,
                               grad_table):
        """Second stage of backward_rays alone: dL/d features scattered into the table gradient."""
        _lib.check(_lib.load().rtxnerf_hashgrid_backward_rays(
            C.byref(self._hg), _p(samples), samples.shape[1], _p(dencoded), _p(num_hits), _p(indices), int(n_rays),
            int(samples_per_segment), _p(grad_table), _stream_ptr(stream)), "hashgrid_backward_rays")

    def forward_seg(self, stream, segments8, num_hits, indices, n_rays, samples_per_segment, output=None,
                    radiance=None, encoded_out=None):
        """forward_rays without a sample buffer: positions rebuilt from the packed hit records ([capacity][8]) by the
        sampler's arithmetic (REGULAR sampling); same rows, same results."""
        if output is None and radiance is None:
            raise ValueError("forward_seg needs an output and/or radiance buffer")
        _lib.check(_lib.load().rtxnerf_network_forward_seg(
            C.byref(self._hg), C.byref(self._mlp), _p(segments8), _p(num_hits), _p(indices), int(n_rays),
            int(samples_per_segment), _p(output), _p(radiance), _p(encoded_out), _stream_ptr(stream)),
            "network_forward_seg")
        _lib.count_launches(1 if n_rays > 0 else 0)
        return output if output is not None else radiance

    def hashgrid_backward_seg(self, stream, segments8, dencoded, num_hits, indices, n_rays, samples_per_segment,
                              grad_table):
        """hashgrid_backward_rays with the positions rebuilt from the packed hit records."""
        _lib.check(_lib.load().rtxnerf_hashgrid_backward_seg(
            C.byref(self._hg), _p(segments8), _p(dencoded), _p(num_hits), _p(indices), int(n_rays),
            int(samples_per_segment), _p(grad_table), _stream_ptr(stream)), "hashgrid_backward_seg")

    Network.mlp_backward = mlp_backward
    Network.hashgrid_backward_rays = hashgrid_backward_rays
    Network.forward_seg = forward_seg
    Network.hashgrid_backward_seg = hashgrid_backward_seg


_network_stage_methods()


def l2_loss(pred, target, loss_scale, gradients, values=None, loss_sum=None, stream=None):
    """loss->evaluate + reduce_sum (main.cu:755-760): fp16 gradients, optional values / running sum."""
    n = pred.numel()
    _lib.check(_lib.load().rtxnerf_l2_loss(_p(pred), _p(target), int(n), float(loss_scale), _p(values), _p(gradients),
                                           _p(loss_sum), _stream_ptr(stream)), "l2_loss")
    _lib.count_launches(1 if n > 0 else 0)


def adam_step(master, params, grads, m1, m2, grad_scale, step, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8,
              stream=None, n_encoding=0):
    """optimizer->step (main.cu:787) with the Adam config of main.cu:40-46; zeroes grads.  n_encoding: the first
    n_encoding parameters are encoding (hash-table) entries, skipped in a step where their gradient is exactly
    zero (tiny-cuda-nn's Adam)."""
    n = master.numel()
    _lib.check(_lib.load().rtxnerf_adam_step_sparse(int(n), int(n_encoding), _p(master), _p(params), _p(grads), _p(m1),
                                                    _p(m2), float(grad_scale), float(lr), float(beta1), float(beta2),
                                                    float(eps), int(step), _stream_ptr(stream)), "adam_step")
    _lib.count_launches(1 if n > 0 else 0)
