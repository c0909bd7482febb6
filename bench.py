#!/usr/bin/env python3
"""bench.py — the driver's benchmark contract for the rtx_nerf per-ray hot path on B200.

  python bench.py --gpus N --steps K --warmup W            (our arm; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W

One "step" = one pass of the hot path over one 800x800 view (BASELINE.json configs[1]):
ray march over the 128^3 lego-shaped occupancy grid -> sampler (S = 4 samples per hit cell,
H_max = 16 hit cells: "64 samples/ray") -> hash-grid encoding + 2x64 fused MLP -> alpha
compositing.  Every rank renders its own view each step (views shard across GPUs with no
data-path collective): weak scaling, value = total rays of all ranks / time.
Synthetic data (seeded poses, lego-shaped grid, random-init hash table and MLP weights).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = dict(workload="configs[1]: one 800x800 view per step per GPU, 128^3 lego-shaped grid, "
                         "S=4 samples/hit x H_max=16 hits (64 samples/ray), 16-level hash-grid + 2x64 fused MLP",
                width=800, height=800, grid=128, samples_per_hit=4, max_hits=16,
                network="hashgrid L16 F2 T2^19 Nmin16 Nmax2048 -> 64 -> 64 -> 16 (sigmoid), fp16",
                parallelism="views sharded across ranks, no collective",
                l2="per-step sample/radiance buffers (1.6 GB) exceed L2; hash table and bitfield are L2-resident by design")
FLOP_PER_SAMPLE = 14336          # 2*(32*64 + 64*64 + 64*16)


def kernel_sources_sha16():
    """Hash of the CUDA sources the .so is built from: ties an ncu capture to the kernels it measured."""
    import glob
    import hashlib
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "rtx_nerf_b200", "csrc", "*.cu")) +
                    glob.glob(os.path.join(ROOT, "rtx_nerf_b200", "csrc", "*.cuh"))):
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def ncu_traffic(stage):
    """dram bytes per launch of `stage`'s kernel from the committed `ncu --set full` capture
    (profiles/ncu_traffic.json, written by scripts/update_traffic.py from a capture of THIS round).  A capture
    taken before the kernels last changed is refused loudly: (None, why)."""
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(tpath):
        return None, "no profiles/ncu_traffic.json"
    d = json.load(open(tpath))
    if d.get("sources_sha16") != kernel_sources_sha16():
        why = ("STALE: profiles/ncu_traffic.json was captured for kernel sources %s (%s), the tree is at %s — "
               "re-run scripts/update_traffic.py on a fresh capture" % (d.get("sources_sha16"), d.get("round"),
                                                                        kernel_sources_sha16()))
        sys.stderr.write("bench.py: roofline.traffic " + why + "\n")
        return None, why
    return d.get("traffic", {}).get(stage), "ncu --set full, %s" % d.get("capture")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf=d["bf16_tflops"], tf_sustained=d.get("bf16_tflops_sustained"), src="measured")
    return dict(hbm=6650.0, tf=1590.0, tf_sustained=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md: ONE
    `nvidia-smi -lms 200` process, started before and killed after — spawning a fresh nvidia-smi
    per sample re-initialises NVML over all GPUs of the box and perturbs the run it watches)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                f = [x.strip() for x in line.strip().split(",")]
                if len(f) >= 6:
                    self.rows.append(f)
        except Exception:
            pass

    def summary(self):
        if self.proc is not None:
            try:
                self.proc.terminate()          # the exact process started above
                self.proc.wait(timeout=5)
            except Exception:
                pass
        rows = []
        for r in list(self.rows):
            try:
                rows.append((float(r[0]), float(r[1]), r[2:6]))
            except ValueError:
                pass
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["unsampled"])
        sm = sorted(r[0] for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2][i].lower().startswith("active") for r in rows)]
        return dict(sm_mhz=sm[len(sm) // 2], sm_min_mhz=sm[0], sm_max_mhz=rows[0][1], reasons=reasons, samples=len(sm))


def scene():
    from rtx_nerf_b200 import synth
    G = WORKLOAD["grid"]
    occ = synth.occupancy_lego(G, seed=0)
    levels = synth.hashgrid_levels()
    table = synth.hashgrid_table(levels, seed=1)
    weights = synth.mlp_weights(seed=2)
    return occ, levels, table, weights


# ------------------------------------------------------------------------------------------- CPU arm
def cpu_path(sample_rays, poses, nthreads, repeats=1):
    """The hot path on host cores: oracle restatement (march via the exact DDA sweep, sampler, hash-grid +
    MLP) and the REFERENCE'S OWN host-compiled vol_render (oracle/_ref, any S) when it is present.  The
    sample is a band of rows through the image centre (the top rows of a view mostly miss the scene).
    Returns (rays_per_s, samples_per_s, description)."""
    from oracle import oracle as orc
    from rtx_nerf_b200 import synth
    occ, levels, table, weights = scene()
    W, H, G, S, HM = WORKLOAD["width"], WORKLOAD["height"], WORKLOAD["grid"], WORKLOAD["samples_per_hit"], WORKLOAD["max_hits"]
    focal = synth.focal_length()
    have_ref = orc.ref_host() is not None
    t_total, rays, samples = 0.0, 0, 0
    for i in range(repeats):
        pose = poses[i % len(poses)]
        t0 = time.perf_counter()
        rb = max(0, (H // 2) * W - sample_rays // 2)
        re = min(W * H, rb + sample_rays)
        m = orc.ray_march_dda(pose, focal, W / H, W, H, occ, G, max_hits=HM, csr=True, rays=(rb, re),
                              nthreads=nthreads)
        nh, idx = m.num_hits[rb:re], m.indices[rb:re]
        s, t = orc.sampler(m.start_points, m.end_points, m.view_dirs[rb:re], nh, idx, S, 0, nthreads=nthreads)
        enc = orc.hashgrid_encode(levels, table, s, nthreads=nthreads)
        _, rad = orc.mlp_forward(weights, 32, 64, 2, 1, enc, nthreads=nthreads)
        if have_ref:
            orc.ref_volrender(rad, nh, idx, t, S)      # vol_render.cu:19-73 compiled for the host
        else:
            orc.volrender(rad, nh, idx, t, S, nthreads=nthreads)
        t_total += time.perf_counter() - t0
        rays += sample_rays
        samples += s.shape[0]
    return rays / t_total, samples / t_total, ("%d rays through the centre of an 800x800 view x %d%s"
                                               % (sample_rays, repeats, "; compositing = the reference's own "
                                                  "vol_render.cu on the host" if have_ref else ""))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from rtx_nerf_b200 import synth
    nthreads = os.cpu_count() or 1
    poses = synth.poses(args.steps + args.warmup + 1, seed=123)
    sample = 16000
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_path(2000, poses, nthreads)
    t_steps = []
    rays_s, samp_s = [], []
    for k in range(args.steps):
        t0 = time.perf_counter()
        r, s, desc = cpu_path(sample, poses[k + 1:], nthreads)
        t_steps.append(time.perf_counter() - t0)
        rays_s.append(r)
        samp_s.append(s)
    v = float(np.mean(rays_s))
    line = dict(metric="rays/s", value=v, unit="rays/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * float(np.mean(t_steps)), higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", impl="reference", config=WORKLOAD, samples_per_s=float(np.mean(samp_s)),
                cpu_baseline=dict(value=v, unit="rays/s", cores=nthreads, kind="port",
                                  sample="each step = %s (CPU oracle: exact-DDA march, sampler, hash-grid + MLP); "
                                         "OptiX / tiny-cuda-nn are unavailable on this box so the reference's own "
                                         "march/MLP cannot run" % desc),
                e2e=dict(value=v, unit="rays/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from rtx_nerf_b200 import _lib, api, pipeline, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"       # the version banner goes to stdout, where ONE JSON line is expected
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.load()
    pk = peaks()

    occ, levels, table, weights = scene()
    W, H, G, S, HM = WORKLOAD["width"], WORKLOAD["height"], WORKLOAD["grid"], WORKLOAD["samples_per_hit"], WORKLOAD["max_hits"]
    focal = synth.focal_length()
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    net = api.Network(levels, table, weights)
    r = pipeline.ViewRenderer(grid, W, H, focal, W / H, S, HM, pipeline.NetworkRadiance(net))
    # every rank renders a different view each step (views round-robin across ranks)
    n_views = args.steps + args.warmup
    poses = synth.poses(n_views * world + 1, seed=123)
    my_poses = [poses[1 + k * world + rank] for k in range(n_views)]
    poses_dev = [torch.as_tensor(p.reshape(16)).cuda() for p in my_poses]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # the render path: march + ONE fused sampler/network/compositing kernel (view_fused.cu);
    # --staged times the reference-shaped four-stage path through its separate entry points
    stage_list = r.stages(fused=not args.staged)
    stage_names = [n for n, _ in stage_list]
    stage_fns = [f for _, f in stage_list]

    # ---- warm-up
    clocks = ClockSampler(local) if rank == 0 else None
    if clocks:
        clocks.start()
    for k in range(args.warmup):
        r.render_device(poses_dev[k], fused=not args.staged)
    barrier()

    # ---- per-stage durations (CUDA events on the launch stream; explains the headline)
    n_stage_steps = min(args.steps, 10)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(stage_fns) + 1)] for _ in range(n_stage_steps)]
    for k in range(n_stage_steps):
        r.pose_dev.copy_(poses_dev[args.warmup + k], non_blocking=True)
        ev[k][0].record()
        for i, fn in enumerate(stage_fns):
            fn()
            ev[k][i + 1].record()
    barrier()
    stage_ms = [float(np.mean([ev[k][i].elapsed_time(ev[k][i + 1]) for k in range(n_stage_steps)]))
                for i in range(len(stage_fns))]

    # ---- timed: EXACTLY K steps.  The launches of a view are captured once in a CUDA graph
    # (every size that depends on the hit count is read on the device), each step = pose copy +
    # one graph launch; pose already resident in HBM.
    launches0 = _lib.kernel_launches()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for fn in stage_fns:
            fn()
    launches_per_step = _lib.kernel_launches() - launches0
    graph.replay()
    # untimed clock spin-up: a B200 coming out of idle needs up to a second or two of load to reach
    # its steady SM clock; without it the first process on a cold GPU measures the ramp.  Each
    # batch is a rehearsal of the timed loop (same K poses), so consecutive batches are comparable:
    # steady = two consecutive batches within 1.5 %, never less than --spinup s, never more than 8 s.
    t_spin = time.perf_counter()
    prev = None
    n_spin = 0
    hits_total = torch.zeros(1, dtype=torch.int64, device="cuda")
    while True:
        n_spin += 1
        a_ev, b_ev = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a_ev.record()
        for k in range(args.steps):
            r.pose_dev.copy_(poses_dev[args.warmup + k], non_blocking=True)
            graph.replay()
            hits_total += r.hits.total_hits     # also loads this op's kernel before the timed region
        b_ev.record()
        torch.cuda.synchronize()
        cur = a_ev.elapsed_time(b_ev)
        el = time.perf_counter() - t_spin
        if el >= args.spinup and (args.spinup == 0 or el > 8.0 or (prev is not None and abs(cur - prev) < 0.015 * prev)):
            break
        prev = cur
    spin_info = dict(seconds=round(el, 3), rehearsals=n_spin, last_ms_per_view=cur / args.steps)
    hits_total.zero_()
    barrier()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for k in range(args.steps):
        r.pose_dev.copy_(poses_dev[args.warmup + k], non_blocking=True)
        graph.replay()
        hits_total += r.hits.total_hits
    t_end.record()
    barrier()
    launches = launches_per_step * args.steps
    ms_total = t_start.elapsed_time(t_end)
    hits = int(hits_total.item())

    # ---- e2e: the user's call, host pose in / host pixels out every step
    # (ViewRenderer.render_views_host: the batch-of-views call of configs[2]; every step copies its
    # 64-byte pose host->device from pinned memory and its 7.68 MB of pixels device->host, the
    # latter on a copy stream under the next view's kernels.  The one-view synchronous call
    # render_host is timed beside it.)
    e2e_poses = [my_poses[args.warmup + k] for k in range(args.steps)]
    e2e_out = torch.empty((args.steps, r.R, 3), dtype=torch.float32, pin_memory=True)
    r.render_views_host(e2e_poses[:2], out=e2e_out[:2], fused=not args.staged)      # untimed: stream + buffers
    barrier()
    t0 = time.perf_counter()
    r.render_views_host(e2e_poses, out=e2e_out, fused=not args.staged)
    barrier()
    e2e_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    for k in range(args.steps):
        r.render_host(my_poses[args.warmup + k], fused=not args.staged)
    e2e_sync_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if not np.array_equal(e2e_out[args.steps - 1].numpy(), r.pixels_pinned.numpy()):
        raise RuntimeError("render_views_host and render_host disagree on the last view")
    clock_summary = clocks.summary() if clocks else None

    # ---- the same view through the reference's OWN network configuration (main.cu:47-68:
    # Frequency encoding + 8 x 128 FullyFusedMLP), staged path; rank 0 only, short run
    refnet = None
    if not args.no_refnet and rank == 0:
        fnet = api.FrequencyNetwork(synth.mlp_weights(112, 128, 8, 16, seed=4))
        r2 = pipeline.ViewRenderer(grid, W, H, focal, W / H, S, HM, pipeline.FrequencyNetworkRadiance(fnet))
        for k in range(2):
            r2.render_device(poses_dev[k % len(poses_dev)])
        torch.cuda.synchronize()
        n2 = 5
        ev2 = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(n2)]
        for k in range(n2):
            r2.pose_dev.copy_(poses_dev[(args.warmup + k) % len(poses_dev)], non_blocking=True)
            ev2[k][0].record()
            for i, (_, fn) in enumerate(r2.stages(fused=False)):
                fn()
                ev2[k][i + 1].record()
        torch.cuda.synchronize()
        ms2 = [float(np.mean([ev2[k][i].elapsed_time(ev2[k][i + 1]) for k in range(n2)])) for i in range(4)]
        n_s2 = float(r2.hits.total_hits.item()) * S
        tf2 = 262144.0 * n_s2 / (ms2[2] * 1e-3) / 1e12
        refnet = dict(workload="configs[1] view with the reference's literal network: Frequency(10 / 12 frequencies) "
                               "-> 8 x 128 ReLU -> 16 (sigmoid), fp16, 262144 FLOP/sample; staged path",
                      ms_per_view=sum(ms2), rays_per_s=W * H / (sum(ms2) * 1e-3), samples_per_s=n_s2 / (sum(ms2) * 1e-3),
                      network_ms=ms2[2], network_tflops=tf2, peak_tflops=pk["tf"], frac_of_peak=tf2 / pk["tf"],
                      stages_ms=dict(march=ms2[0], sampler=ms2[1], network=ms2[2], composite=ms2[3]))
        if world == 1 and not args.no_train:
            # the training step of configs[3] through this network: 2^18 rays, S=4 x H_max=32
            TR2, TS2, THM2 = 1 << 18, 4, 32
            tr2 = pipeline.Trainer(grid, fnet, W, H, focal, W / H, TS2, THM2, TR2)
            tgt2 = torch.rand((TR2, 3), device="cuda")
            for k in range(2):
                tr2.step(poses_dev[k % len(poses_dev)], (k * 7919) % (W * H - TR2), tgt2)
            torch.cuda.synchronize()
            ea, eb, ec = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            n3, bw_ms = 3, 0.0
            ea.record()
            for k in range(n3):
                v2 = tr2.view
                v2.pose_dev.copy_(poses_dev[(2 + k) % len(poses_dev)], non_blocking=True)
                v2.ray_begin = ((2 + k) * 7919) % (W * H - TR2)
                tr2.target.copy_(tgt2, non_blocking=True)
                tr2.forward()
                tr2.backward()
                tr2.optimizer_step()
            eb.record()
            torch.cuda.synchronize()
            n_s3 = float(tr2.view.hits.total_hits.item()) * TS2
            ms3 = ea.elapsed_time(eb) / n3
            # backward alone (three launches of freqnet_backward_kernel)
            ea.record()
            tr2.net.backward(None, tr2.view.samples, tr2.radiance_grad, tr2.grads, n=tr2.view.N_cap,
                             d_count=tr2.view.hits.total_hits, count_scale=TS2)
            ec.record()
            torch.cuda.synchronize()
            bw_ms = ea.elapsed_time(ec)
            tf3 = 51 * 2.0 * 128 * 128 * 128 * (n_s3 / 128) / (bw_ms * 1e-3) / 1e12    # 51 layer products per tile (27 + 15 + 9)
            refnet["train_step"] = dict(
                workload="configs[3] shape through the literal network: 2^18 rays, S=4 x H_max=32, forward, loss, "
                         "compositing backward, weight-gradient backward (forward recomputed per launch, 3 launches, 51 layer products per tile for 27 minimal), Adam, repack",
                ms_per_step=ms3, rays_per_s=TR2 / (ms3 * 1e-3), samples_per_s=n_s3 / (ms3 * 1e-3),
                backward_ms=bw_ms, backward_tflops_executed=tf3, frac_of_peak=tf3 / pk["tf"])
            del tr2
        del r2, fnet

    # ---- BASELINE configs[4] (stress): 1920x1080 view, 512^3 grid at 10 % occupancy, up to 256 samples/ray
    # (S = 4 x H_max = 64), through the network on the fused render path; N = 1 only, a short run
    cfg5 = None
    if not args.no_config5 and world == 1:
        G5, W5, H5, S5, HM5 = 512, 1920, 1080, 4, 64
        occ5 = synth.occupancy_bernoulli(G5, 0.10, seed=1)
        grid5 = api.OccupancyGrid.from_occupancy(occ5, G5)
        r5 = pipeline.ViewRenderer(grid5, W5, H5, focal, W5 / H5, S5, HM5, pipeline.NetworkRadiance(net))
        for k in range(2):
            r5.render_device(poses_dev[k % len(poses_dev)])
        torch.cuda.synchronize()
        n5 = 4
        st5 = r5.stages(fused=True)
        ev5 = [[torch.cuda.Event(enable_timing=True) for _ in range(len(st5) + 1)] for _ in range(n5)]
        hits5 = 0
        for k in range(n5):
            r5.pose_dev.copy_(poses_dev[(args.warmup + k) % len(poses_dev)], non_blocking=True)
            ev5[k][0].record()
            for i, (_, fn) in enumerate(st5):
                fn()
                ev5[k][i + 1].record()
            torch.cuda.synchronize()
            hits5 += r5.total_hits()
        ms5 = [float(np.mean([ev5[k][i].elapsed_time(ev5[k][i + 1]) for k in range(n5)])) for i in range(len(st5))]
        R5, H5h = W5 * H5, hits5 / n5
        N5 = H5h * S5
        cfg5 = dict(workload="configs[4]: one 1920x1080 view, 512^3 Bernoulli(0.10) grid, S=4 x H_max=64 (<= 256 "
                             "samples/ray), 16-level hash-grid + 2x64 fused MLP, fused render path",
                    ms_per_view=sum(ms5), rays_per_s=R5 / (sum(ms5) * 1e-3), samples_per_s=N5 / (sum(ms5) * 1e-3),
                    rays=R5, hits_per_ray=H5h / R5, samples=N5,
                    stages={nme: dict(ms=ms) for (nme, _), ms in zip(st5, ms5)})
        gb5 = (28 * R5 + 32 * H5h + G5 ** 3 / 8 + 64) / (ms5[0] * 1e-3) / 1e9
        cfg5["stages"]["march"].update(bound="hbm", achieved=gb5, peak=pk["hbm"], unit="GB/s", frac=gb5 / pk["hbm"])
        tf5 = FLOP_PER_SAMPLE * N5 / (ms5[1] * 1e-3) / 1e12
        cfg5["stages"][st5[1][0]].update(bound="tensor", achieved=tf5, peak=pk["tf"], unit="TFLOP/s",
                                         frac=tf5 / pk["tf"], samples_per_s=N5 / (ms5[1] * 1e-3))
        del r5, grid5, occ5
        torch.cuda.empty_cache()

    # ---- training step (BASELINE configs[3]): 2^18 rays x (S=4, H_max=32), fwd + bwd + Adam,
    # NCCL all-reduce of the flat fp32 gradient buffer when N > 1; a short run, reported beside
    # the headline
    train = None
    if not args.no_train:
        del graph
        TR, TS, THM, t_warm, t_steps = 1 << 18, 4, 32, 2, 5
        tr = pipeline.Trainer(grid, net, W, H, focal, W / H, TS, THM, TR,
                              process_group=dist.group.WORLD if world > 1 else None)
        target = torch.rand((TR, 3), device="cuda")
        tposes = poses_dev[:t_warm + t_steps]
        for k in range(t_warm):
            tr.step(tposes[k % len(tposes)], (k * 7919) % (W * H - TR), target)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c, d = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ar_ms = 0.0
        a.record()
        for k in range(t_steps):
            v = tr.view
            v.pose_dev.copy_(tposes[(t_warm + k) % len(tposes)], non_blocking=True)
            v.ray_begin = ((t_warm + k) * 7919) % (W * H - TR)
            tr.forward()
            tr.backward()
            c.record()
            tr.allreduce()
            d.record()
            tr.optimizer_step()
        b.record()
        barrier()
        ar_ms = c.elapsed_time(d)   # the last step's collective, INCLUDING the wait for the slowest rank's backward
        ar_alone = 0.0
        if world > 1:               # the collective alone: all ranks enter together, 5 back-to-back calls
            barrier()
            c.record()
            for _ in range(5):
                tr.allreduce()
            d.record()
            barrier()
            ar_alone = c.elapsed_time(d) / 5
        train = dict(ms=a.elapsed_time(b) / t_steps, allreduce_ms=ar_ms, allreduce_alone_ms=ar_alone, rays=TR,
                     samples=float(tr.view.hits.total_hits.item()) * TS, grad_bytes=tr.grads.numel() * 4,
                     launches_per_step=tr.launches_per_step, loss=float(tr.loss_sum.item()) / (t_warm + t_steps))
        # the same step fed the reference's way (main.cu:612-676): 2^18 rays drawn from ALL of this rank's views
        # by the seeded shuffle of a resident ray store (batch_assemble instead of a per-step march)
        SV = 4
        store = api.RayStore(SV, W * H, SV * W * H * THM, pixels_gt=torch.rand((SV * W * H, 3), device="cuda"))
        for k in range(SV):
            store.add_view(grid, poses_dev[k % len(poses_dev)], focal, W / H, W, H, THM)
        n_batches = store.n_rays // TR
        for k in range(t_warm):
            tr.step_from_store(store, 1 + rank, (k % n_batches) * TR, 32)
        barrier()
        a.record()
        for k in range(t_steps):
            tr.step_from_store(store, 1 + rank, ((t_warm + k) % n_batches) * TR, 32)
        b.record()
        barrier()
        train.update(shuffled_ms=a.elapsed_time(b) / t_steps, shuffled_views=SV,
                     shuffled_samples=float(tr.view.hits.total_hits.item()) * TS)
        del store

    # ---- max over ranks
    tv = [train["ms"], train["allreduce_ms"], train["samples"]] if train else [0.0, 0.0, 0.0]
    ar_alone_all = 0.0
    if train and world > 1:
        t_ar = torch.tensor([train["allreduce_alone_ms"], train["shuffled_ms"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(t_ar, op=dist.ReduceOp.MAX)
        ar_alone_all, train["shuffled_ms"] = float(t_ar[0]), float(t_ar[1])
    vals = torch.tensor([ms_total, e2e_s * 1e3, float(hits)] + tv + stage_ms, dtype=torch.float64, device="cuda")
    per_rank = [[ms_total / args.steps, spin_info["last_ms_per_view"]]]
    if world > 1:
        mine = torch.tensor(per_rank[0], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [[float(x) for x in t] for t in allr]
        mx = vals.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vals.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms_total, e2e_ms = float(mx[0]), float(mx[1])
        hits_all = float(sm[2])
        if train:
            train["ms"], train["allreduce_ms"], train["samples_all"] = float(mx[3]), float(mx[4]), float(sm[5])
        stage_ms = [float(x) for x in mx[6:]]
    else:
        e2e_ms, hits_all = e2e_s * 1e3, float(hits)
        if train:
            train["samples_all"] = train["samples"]

    if rank == 0:
        R = W * H
        rays_all = R * args.steps * world
        samples_all = hits_all * S
        value = rays_all / (ms_total * 1e-3)
        # per-launch algorithmic work of this rank's kernels, per step
        Hh = hits / args.steps
        Ns = Hh * S
        byts = {"march": 28 * R + 32 * Hh + G ** 3 / 8 + 64, "sampler": 16 * R + 24 * Hh + 24 * Ns,
                "composite": 20 * Ns + 8 * R + 12 * R}
        stages = {}
        for nme, ms in zip(stage_names, stage_ms):
            if "network" in nme:
                tf = FLOP_PER_SAMPLE * Ns / (ms * 1e-3) / 1e12
                stages[nme] = dict(ms=ms, bound="tensor", achieved=tf, peak=pk["tf"], unit="TFLOP/s", frac=tf / pk["tf"],
                                   samples_per_s=Ns / (ms * 1e-3))
            else:
                gb = byts[nme] / (ms * 1e-3) / 1e9
                stages[nme] = dict(ms=ms, bound="hbm", achieved=gb, peak=pk["hbm"], unit="GB/s", frac=gb / pk["hbm"])
        dom = max(stages, key=lambda k: stages[k]["ms"])
        d = stages[dom]
        # dram bytes per launch of the dominant kernel from the committed ncu --set full capture
        traffic, traffic_src = ncu_traffic(dom)
        roofline = dict(kernel=dom, bound=d["bound"], achieved=d["achieved"], peak=d["peak"], unit=d["unit"],
                        frac=d["frac"], traffic=traffic, traffic_source=traffic_src, peak_source=pk["src"],
                        note="achieved = algorithmic work per launch (SURVEY 8d) / mean CUDA-event duration of that "
                             "stage in a separate eager (un-graphed) pass over %d views just before the timed "
                             "region; kernel share of that pass = %.2f" % (n_stage_steps, d["ms"] / sum(stage_ms)))
        line = dict(metric="rays/s", value=value, unit="rays/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms_total / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                    dtype="f16 tensor-core MLP (fp32 accumulate), f32 march/sampler/compositing", data="synthetic",
                    config=dict(WORKLOAD, path="staged: march, sampler, network, compositing kernels" if args.staged
                                else "march kernel + one fused sampler/encoding/MLP/compositing kernel"),
                    samples_per_s=samples_all / (ms_total * 1e-3),
                    rays_per_step_per_gpu=R, samples_per_step_per_gpu=Ns,
                    e2e=dict(value=rays_all / (e2e_ms * 1e-3), unit="rays/s", h2d_bytes_per_step=64,
                             d2h_bytes_per_step=R * 12, ms_per_step=e2e_ms / args.steps,
                             call="ViewRenderer.render_views_host (K views, D2H of view k under view k+1)",
                             one_view_sync_call_ms=e2e_sync_ms),
                    gpu_launches=launches, roofline=roofline, stages=stages, clocks=clock_summary, spinup=spin_info,
                    per_rank_ms=dict(timed=[round(x[0], 4) for x in per_rank], last_rehearsal=[round(x[1], 4) for x in per_rank]))
        if refnet:
            line["reference_network"] = refnet
        if cfg5:
            line["config5"] = cfg5
        if train:
            # top level, so that the driver's per-N table carries the path that HAS a collective
            line["train_step_ms"] = train["ms"]
            line["train_step_store_fed_ms"] = train["shuffled_ms"]
            line["train_rays_per_s"] = train["rays"] * world / (train["shuffled_ms"] * 1e-3)
            line["train_step"] = dict(
                workload="configs[3]: 2^18 rays per GPU per step, S=4 x H_max=32 (128 samples/ray), network fwd+bwd, "
                         "compositing fwd+bwd, L2 loss, Adam; NCCL all-reduce(sum) of the fp32 gradient buffer",
                ms_per_step=train["ms"], rays_per_s=train["rays"] * world / (train["ms"] * 1e-3),
                samples_per_s=train["samples_all"] / (train["ms"] * 1e-3), allreduce_ms=train["allreduce_ms"], allreduce_alone_ms=ar_alone_all,
                allreduce_bytes=train["grad_bytes"], launches_per_step=train["launches_per_step"],
                mean_loss=train["loss"], scaling="weak",
                shuffled=dict(workload="same step, rays drawn from %d resident views per rank by the seeded shuffle "
                                       "(RayStore + batch_assemble, groups of 32 adjacent pixels)" % train["shuffled_views"],
                              ms_per_step=train["shuffled_ms"], rays_per_s=train["rays"] * world / (train["shuffled_ms"] * 1e-3),
                              samples_last_step=train["shuffled_samples"]))
        if world == 1 and not args.no_cpu:
            nthreads = os.cpu_count() or 1
            cpu_path(1000, my_poses, nthreads)
            rs, ss, desc = cpu_path(16000, my_poses, nthreads)
            line["cpu_baseline"] = dict(value=rs, unit="rays/s", cores=nthreads, kind="port", samples_per_s=ss,
                                        sample=desc + " (CPU oracle, all host threads)")
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-train", action="store_true", help="skip the training-step leg (profiling runs)")
    ap.add_argument("--no-refnet", action="store_true", help="skip the Frequency + 8x128 network leg")
    ap.add_argument("--no-config5", action="store_true", help="skip the 1920x1080 / 512^3 stress-configuration leg")
    ap.add_argument("--spinup", type=float, default=1.5, help="seconds of untimed load before the timed steps")
    ap.add_argument("--staged", action="store_true",
                    help="render through the separate sampler / network / compositing entry points instead of the "
                         "fused kernel")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
