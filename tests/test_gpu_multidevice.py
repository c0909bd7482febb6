"""SURVEY 8(b) "Threading": the C-ABI entry points hold no per-process device state, so two host threads of ONE
process can each drive their own GPU (function attributes such as the dynamic shared-memory limit are per
device: librtxnerf_b200.so keeps them per device, common.cuh RtxPerDeviceOnce).  Needs two visible GPUs."""
import threading

import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def test_two_threads_two_devices_render_and_train_forward(cuda_api):
    import torch
    from rtx_nerf_b200 import pipeline
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    api = cuda_api
    G, W, H, S, HM = 64, 96, 64, 4, 16
    occ = synth.occupancy_lego(G, seed=0)
    lv = synth.hashgrid_levels()
    table = np.random.default_rng(1).uniform(-1, 1, lv["n_entries"] * 2).astype(np.float16)
    weights = synth.mlp_weights(seed=2)
    fw = synth.mlp_weights(112, 128, 8, 16, seed=4)
    pose, focal = synth.poses(2, seed=9)[1], synth.focal_length()
    out, err = {}, {}

    def work(dev):
        try:
            torch.cuda.set_device(dev)
            with torch.cuda.device(dev):
                d = "cuda:%d" % dev
                grid = api.OccupancyGrid.from_occupancy(occ, G)
                net = api.Network(lv, table, weights, device=d)
                r = pipeline.ViewRenderer(grid, W, H, focal, W / H, S, HM, pipeline.NetworkRadiance(net), device=d)
                res = []
                for _ in range(3):          # > 48 KB dynamic shared memory kernels: march, fused render, staged
                    res.append(r.render_host(pose, fused=True).numpy().copy())
                    res.append(r.render_host(pose, fused=False).numpy().copy())
                fnet = api.FrequencyNetwork(fw, device=d)
                r2 = pipeline.ViewRenderer(grid, W, H, focal, W / H, S, HM, pipeline.FrequencyNetworkRadiance(fnet),
                                           device=d)
                res.append(r2.render_host(pose).numpy().copy())
                torch.cuda.synchronize(dev)
                out[dev] = res
        except Exception as e:      # surfaced in the main thread
            err[dev] = e

    th = [threading.Thread(target=work, args=(i,)) for i in (0, 1)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not err, err
    for a, b in zip(out[0], out[1]):
        assert np.array_equal(a, b), "the two devices disagree"
    assert np.abs(out[0][0] - out[0][1]).max() < 4e-6      # fused == staged on each
