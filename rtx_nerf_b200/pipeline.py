"""The per-ray hot path as one pipeline over the C ABI: ray march -> sampler -> radiance ->
compositing, in the reference's call order (main.cu:506-508 then :704-746), with every buffer
allocated once (the reference re-mallocs per iteration, main.cu:683-694).

The radiance stage is pluggable: the network (hash-grid + fused MLP, BASELINE configs 2-5) or
the fixed per-cell table of BASELINE configs[0].
"""
import ctypes as C

import numpy as np
import torch

from . import _lib, api, parallel


class ViewRenderer:
    """Renders views of W x H rays (or a contiguous ray range of them: the multi-GPU shard)."""

    def __init__(self, grid, width, height, focal_length, aspect_ratio, samples_per_segment, max_hits,
                 radiance, n_rays=None, ray_begin=0, capacity_hits=None, device="cuda", keep_t=False,
                 want_segments=None):
        self.grid, self.W, self.H = grid, width, height
        self.focal, self.aspect = float(focal_length), float(aspect_ratio)
        self.S, self.max_hits = int(samples_per_segment), int(max_hits)
        self.ray_begin = int(ray_begin)
        self.R = int(n_rays if n_rays is not None else width * height - ray_begin)
        self.cap = int(capacity_hits if capacity_hits is not None else self.R * self.max_hits)
        self.radiance_stage = radiance
        dev = torch.device(device)
        fusable = isinstance(radiance, NetworkRadiance) and radiance.fusable
        self.hits = api.HitBuffers(self.R, self.cap, device=dev, want_t=keep_t, want_cells=radiance.needs_cells,
                                   want_origins=False,
                                   want_segments=fusable if want_segments is None else bool(want_segments))
        N = self.cap * self.S
        self.N_cap = N
        self._dev = dev
        # the staged path's per-sample buffers (40 B per sample: 21 GB at config 5) are allocated on first use:
        # the fused render path never touches them
        self._t_vals = self._samples = self._radiance = None
        self.pixels = torch.empty((self.R, 3), dtype=torch.float32, device=dev)
        self.pose_dev = torch.empty(16, dtype=torch.float32, device=dev)
        self.pose_pinned = torch.empty(16, dtype=torch.float32).pin_memory()
        self.pixels_pinned = torch.empty((self.R, 3), dtype=torch.float32).pin_memory()
        march = 1 if max_hits <= 32 else 3
        self.launches_per_view = (march + 1) if self.can_fuse else (march + 1 + radiance.launches + 1)

    @property
    def t_vals(self):
        if self._t_vals is None:
            self._t_vals = torch.empty(self.N_cap, dtype=torch.float32, device=self._dev)
        return self._t_vals

    @property
    def samples(self):
        if self._samples is None:
            self._samples = torch.empty((self.N_cap, 5), dtype=torch.float32, device=self._dev)
        return self._samples

    @property
    def radiance(self):
        if self._radiance is None:
            self._radiance = torch.empty((self.N_cap, 4), dtype=torch.float32, device=self._dev)
        return self._radiance

    # --- stages (all asynchronous on the current stream) -------------------------------
    def stage_march(self):
        api.ray_march_csr(self.grid, self.pose_dev, self.focal, self.aspect, self.W, self.H, self.hits,
                          self.max_hits, ray_begin=self.ray_begin, n_rays=self.R)

    def stage_march_fused(self):
        """The march of the fused render path: packed hit records only (no start / end point arrays)."""
        api.ray_march_csr(self.grid, self.pose_dev, self.focal, self.aspect, self.W, self.H, self.hits,
                          self.max_hits, ray_begin=self.ray_begin, n_rays=self.R,
                          points=self.hits.segments8 is None)

    def stage_sampler(self):
        h = self.hits
        api.launchSampler(h.start_points, h.end_points, h.view_dirs, self.t_vals, self.samples, self.R,
                          self.grid.G, h.num_hits, h.indices, api.SAMPLING_REGULAR, None, self.S)

    def stage_radiance(self):
        self.radiance_stage(self)

    def stage_composite(self):
        h = self.hits
        # (network_inputs is unused, as in the reference: pass the sample buffer only if one exists)
        api.launch_volrender_cuda(self._samples, self.radiance, h.num_hits, h.indices, self.t_vals, self.R, self.S,
                                  self.pixels)

    def stage_fused(self):
        """Sampler + network + compositing in one kernel (render path with a NetworkRadiance)."""
        h = self.hits
        if h.segments8 is not None:
            self.radiance_stage.net.render_rays_fused_seg(None, h.segments8, h.num_hits, h.indices, self.R, self.S,
                                                          self.pixels)
            return
        self.radiance_stage.net.render_rays_fused(None, h.start_points, h.end_points, h.num_hits, h.indices,
                                                  self.R, self.S, self.pixels)

    @property
    def can_fuse(self):
        return isinstance(self.radiance_stage, NetworkRadiance) and self.radiance_stage.fusable

    def stages(self, fused=None):
        """[(name, callable)] of one view: the fused render path when available (or asked for)."""
        fused = self.can_fuse if fused is None else fused
        if fused:
            return [("march", self.stage_march_fused), ("sample+network+composite", self.stage_fused)]
        return [("march", self.stage_march), ("sampler", self.stage_sampler), ("network", self.stage_radiance),
                ("composite", self.stage_composite)]

    def render_device(self, pose_dev=None, fused=None):
        """Pose already on the device; returns the device pixel buffer."""
        if pose_dev is not None:
            self.pose_dev.copy_(pose_dev, non_blocking=True)
        for _, fn in self.stages(fused):
            fn()
        return self.pixels

    def render_host(self, pose_host, fused=None):
        """The user-facing call: host pose in, host pixels out (H2D + D2H inside)."""
        self.pose_pinned.copy_(torch.as_tensor(np.asarray(pose_host, np.float32).reshape(16)))
        self.pose_dev.copy_(self.pose_pinned, non_blocking=True)
        self.render_device(fused=fused)
        self.pixels_pinned.copy_(self.pixels, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.pixels_pinned

    def render_views_host(self, poses_host, out=None, fused=None):
        """Batch render (BASELINE configs[2], the reference's per-view loop main.cu:560-790): host
        poses in, host pixels out for every view.  The device->host copy of view k runs on a copy
        stream under the kernels of view k+1 (two device pixel buffers); returns the pinned
        [n_views][R][3] buffer after the last copy has landed."""
        n = len(poses_host)
        if out is None:
            out = torch.empty((n, self.R, 3), dtype=torch.float32, pin_memory=True)
        if tuple(out.shape) != (n, self.R, 3) or out.dtype != torch.float32 or not out.is_pinned():
            raise ValueError("render_views_host: out must be a pinned float32 [n_views][R][3] tensor")
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream()
            self._pixels_alt = torch.empty_like(self.pixels)
        poses_pinned = torch.empty((n, 16), dtype=torch.float32, pin_memory=True)
        poses_pinned.copy_(torch.as_tensor(np.asarray(poses_host, np.float32).reshape(n, 16)))
        main = torch.cuda.current_stream()
        bufs = [self.pixels, self._pixels_alt]
        copied = [None, None]
        try:
            for k in range(n):
                b = k & 1
                if copied[b] is not None:
                    main.wait_event(copied[b])          # this pixel buffer has left the device
                self.pixels = bufs[b]
                self.pose_dev.copy_(poses_pinned[k], non_blocking=True)
                self.render_device(fused=fused)
                rendered = torch.cuda.Event()
                rendered.record(main)
                self._copy_stream.wait_event(rendered)
                with torch.cuda.stream(self._copy_stream):
                    out[k].copy_(bufs[b], non_blocking=True)
                    copied[b] = torch.cuda.Event()
                    copied[b].record(self._copy_stream)
        finally:
            self.pixels = bufs[0]
        self._copy_stream.synchronize()
        main.synchronize()
        return out

    def total_hits(self):
        return int(self.hits.total_hits.item())


class CellRadiance:
    """BASELINE configs[0]: fixed random density / RGB per grid cell."""
    needs_cells = True
    launches = 1

    def __init__(self, cell_rgbs, device="cuda"):
        self.cells = torch.as_tensor(np.ascontiguousarray(cell_rgbs, np.float32)).to(device)

    def __call__(self, r):
        api.cell_radiance(r.hits.hit_cells, r.hits.total_hits, r.cap, r.S, self.cells, r.radiance)


class NetworkRadiance:
    """BASELINE configs 2-5: hash-grid encoding + fused MLP (tcgen05) evaluated on the sampler's
    [N][5] buffer, fp32 (r,g,b,sigma) written straight into the compositing input."""
    needs_cells = False
    launches = 1

    def __init__(self, network, impl=api.IMPL_TCGEN05, schedule="rays"):
        self.net, self.impl, self.schedule = network, impl, schedule
        self.fusable = schedule == "rays" and impl == api.IMPL_TCGEN05

    def __call__(self, r):
        if self.schedule == "rays" and self.impl == api.IMPL_TCGEN05:
            self.net.forward_rays(None, r.samples, r.hits.num_hits, r.hits.indices, r.R, r.S,
                                  radiance=r.radiance)
        else:
            self.net.forward(None, r.samples, radiance=r.radiance, n=r.N_cap, d_count=r.hits.total_hits,
                             count_scale=r.S, impl=self.impl)


class FrequencyNetworkRadiance:
    """The reference's own network configuration (main.cu:47-68) as the radiance stage: Frequency
    encoding + 8 x 128 MLP on the sampler's [N][5] buffer (api.FrequencyNetwork)."""
    needs_cells = False
    launches = 1
    fusable = False

    def __init__(self, network):
        self.net = network

    def __call__(self, r):
        self.net.forward(None, r.samples, radiance=r.radiance, n=r.N_cap, d_count=r.hits.total_hits,
                         count_scale=r.S)


class Trainer:
    """One data-parallel training step of the reference's loop (main.cu:612-805) on this rank's
    rays: march -> [sampler] -> network forward (features kept) -> compositing -> L2 loss ->
    compositing backward -> network backward -> [all-reduce of the flat fp32 gradient buffer]
    -> Adam.  Every buffer is allocated once; the step is 8 kernel launches plus the collective (9 with the
    sampler: the hash-grid network rebuilds its sample positions from the march's packed hit records, from_hits).

    Rays of a step are a contiguous range of one view (ray_begin, n_rays); ranks take different
    views/ranges.  Weights stay replicated: every rank applies the same Adam step to the same
    summed gradient — the all-reduce is the only exchange."""

    def __init__(self, grid, network, width, height, focal_length, aspect_ratio, samples_per_segment, max_hits,
                 n_rays, loss_scale=128.0, lr=1e-3, process_group=None, device="cuda",
                 backward_mode=api.BACKWARD_REFERENCE):
        self.net, self.S, self.R = network, int(samples_per_segment), int(n_rays)
        self.loss_scale, self.lr = float(loss_scale), float(lr)
        self.pg = process_group
        self.overlap, self._side, self._reduced = True, None, False     # piecewise all-reduce on a side stream
        self.backward_mode = int(backward_mode)     # compositing backward: the reference's formula or the exact gradient
        self.world = 1
        if process_group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(process_group)
        # the reference's literal network (Frequency + 8 x 128, main.cu:47-68) trains through the same step: its
        # backward recomputes the forward per tile, so no features are kept, and it has no encoding parameters
        self.literal = isinstance(network, api.FrequencyNetwork)
        # hash-grid network, REGULAR sampling: forward and scatter rebuild the sample positions from the packed hit
        # records (one 256-bit load per sample), so the step launches no sampler and touches no [N][5] sample
        # buffer; the t-values the compositing reads depend only on a sample's index inside its segment and are
        # written once.  from_hits = False puts the sampler back into the step.
        self.from_hits = not self.literal
        stage = FrequencyNetworkRadiance(network) if self.literal else _EncodedNetworkRadiance(self)
        self.view = ViewRenderer(grid, width, height, focal_length, aspect_ratio, samples_per_segment, max_hits,
                                 stage, n_rays=n_rays, ray_begin=0, device=device, want_segments=self.from_hits)
        v = self.view
        dev = torch.device(device)
        N = v.N_cap
        if not self.literal:
            self.encoded = torch.empty((N, 32), dtype=torch.float16, device=dev)
            self.dencoded = torch.empty((N, 32), dtype=torch.float16, device=dev)
        self.radiance_grad = torch.empty((N, 4), dtype=torch.float16, device=dev)
        self.target = torch.empty((self.R, 3), dtype=torch.float32, device=dev)
        self.loss_grad = torch.empty((self.R, 3), dtype=torch.float16, device=dev)
        self.loss_sum = torch.zeros(1, dtype=torch.float32, device=dev)
        self.params = network.weights if self.literal else network.params         # the flat fp16 parameter buffer
        n_params = self.params.numel()
        self.n_table = 0 if self.literal else network.table.numel()
        self.grads = torch.zeros(n_params, dtype=torch.float32, device=dev)      # [table | weights], all-reduced
        self.master = self.params.float()                                          # fp32 master (main.cu:331)
        self.m1 = torch.zeros(n_params, dtype=torch.float32, device=dev)
        self.m2 = torch.zeros(n_params, dtype=torch.float32, device=dev)
        self.step_count = 0
        if self.from_hits:
            # sampler.cu:54-66 (REGULAR): t_vals of sample i of a segment = (i steps of 1/S) + 1/S in fp32
            inv_s = np.float32(1.0) / np.float32(self.S)
            tv, t = np.empty(self.S, np.float32), np.float32(0.0)
            for i in range(self.S):
                t = np.float32(t + inv_s)
                tv[i] = t
            v.t_vals.copy_(torch.from_numpy(np.tile(tv, v.cap)).to(dev))

    def _sample(self):
        if not self.from_hits:
            self.view.stage_sampler()

    def forward(self):
        v = self.view
        if self.from_hits:
            v.stage_march_fused()      # packed hit records only
        else:
            v.stage_march()
        self._sample()
        v.stage_radiance()
        v.stage_composite()

    def backward(self):
        """Loss, compositing backward, network backward.  With more than one rank the all-reduce is issued HERE,
        piecewise on a side stream: the MLP weight gradients (28 KB) as soon as mlp_backward has finished, under
        the scatter; the table gradient after the scatter.  allreduce() then only joins the side stream."""
        v, h = self.view, self.view.hits
        api.l2_loss(v.pixels, self.target, self.loss_scale, self.loss_grad, loss_sum=self.loss_sum)
        api.launch_volrender_backward_cuda(None, self.loss_grad, v.radiance, v.t_vals, h.num_hits, h.indices,
                                           self.R, self.S, self.radiance_grad, backward_mode=self.backward_mode)
        if self.literal:
            self.net.backward(None, v.samples, self.radiance_grad, self.grads, n=v.N_cap, d_count=h.total_hits,
                              count_scale=self.S)
            self._reduced = False
            return
        if self.world == 1 or not self.overlap:
            self.net.mlp_backward(None, self.encoded, self.radiance_grad, self.S, h.total_hits, v.N_cap, self.dencoded,
                                  self.grads[self.n_table:])
            self._scatter()
            self._reduced = False
            return
        main = torch.cuda.current_stream()
        if self._side is None:
            self._side = torch.cuda.Stream()
        self.net.mlp_backward(None, self.encoded, self.radiance_grad, self.S, h.total_hits, v.N_cap, self.dencoded,
                              self.grads[self.n_table:])
        ev = torch.cuda.Event()
        ev.record(main)
        self._side.wait_event(ev)
        with torch.cuda.stream(self._side):
            parallel.allreduce_sum_(self.grads[self.n_table:], self.pg)
        self._scatter()
        ev2 = torch.cuda.Event()
        ev2.record(main)
        self._side.wait_event(ev2)
        with torch.cuda.stream(self._side):
            parallel.allreduce_sum_(self.grads[:self.n_table], self.pg)
        self._reduced = True

    def _scatter(self):
        v, h = self.view, self.view.hits
        if self.from_hits:
            self.net.hashgrid_backward_seg(None, h.segments8, self.dencoded, h.num_hits, h.indices, self.R, self.S,
                                           self.grads[:self.n_table])
        else:
            self.net.hashgrid_backward_rays(None, v.samples, self.dencoded, h.num_hits, h.indices, self.R, self.S,
                                            self.grads[:self.n_table])

    def allreduce(self):
        if self.world > 1:
            if getattr(self, "_reduced", False):
                torch.cuda.current_stream().wait_stream(self._side)     # issued in backward()
                self._reduced = False
            else:
                parallel.allreduce_sum_(self.grads, self.pg)

    def optimizer_step(self):
        self.step_count += 1
        api.adam_step(self.master, self.params, self.grads, self.m1, self.m2,
                      parallel.grad_scale(self.loss_scale, self.pg) if self.world > 1 else 1.0 / self.loss_scale,
                      self.step_count, lr=self.lr, n_encoding=self.n_table)
        if self.literal:
            self.net.repack()          # row-major fp16 weights -> the forward kernel's operand layout

    def step(self, pose_dev, ray_begin, target_dev):
        """pose (16 floats) and target pixels [n_rays][3] already on the device."""
        v = self.view
        v.pose_dev.copy_(pose_dev, non_blocking=True)
        v.ray_begin = int(ray_begin)
        self.target.copy_(target_dev, non_blocking=True)
        self.forward()
        self.backward()
        self.allreduce()
        self.optimizer_step()

    def step_from_store(self, store, seed, first, group=1):
        """One step on positions [first, first + n_rays) of the seeded shuffle of a resident RayStore
        (the reference's own batching, main.cu:612-676: shuffled rays of ALL views, not a range of
        one view).  batch_assemble takes the place of the march: three launches, no host work."""
        v = self.view
        api.batch_assemble(store, seed, group, first, self.R, v.hits, gt_pixels=self.target,
                           points=not self.from_hits)
        self._sample()
        v.stage_radiance()
        v.stage_composite()
        self.backward()
        self.allreduce()
        self.optimizer_step()

    @property
    def launches_per_step(self):
        # + loss, compositing bwd, network bwd (2), Adam; no sampler launch when positions come from the hit records
        return self.view.launches_per_view + 1 + 1 + 2 + 1 - (1 if self.from_hits else 0)


class _EncodedNetworkRadiance:
    """Training forward: NetworkRadiance that also keeps the encoded features."""
    needs_cells = False
    launches = 1

    def __init__(self, trainer):
        self.t = trainer

    def __call__(self, r):
        if self.t.from_hits:
            h = r.hits
            self.t.net.forward_seg(None, h.segments8, h.num_hits, h.indices, r.R, r.S, radiance=r.radiance,
                                   encoded_out=self.t.encoded)
            return
        self.t.net.forward_rays(None, r.samples, r.hits.num_hits, r.hits.indices, r.R, r.S, radiance=r.radiance,
                                encoded_out=self.t.encoded)
