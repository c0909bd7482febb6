"""The ANALYTIC compositing backward (rtxnerf_launch_volrender_backward_mode, RTXNERF_BACKWARD_ANALYTIC): an
explicit, flagged mode beside the reference's own formula (vol_render.cu:118-135 is not the gradient of
vol_render.cu:55-67).  Checked against torch.autograd of the forward formula in fp64, and by a
training run whose PSNR has to rise — with the reference-mode backward the step is only shown to step."""
import numpy as np
import pytest

from rtx_nerf_b200 import synth

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _forward_f64(rad, t_vals, num_hits, indices, S):
    """vol_render.cu:55-67 per ray in torch fp64 (delta carried across segments, tau includes the sample)."""
    import torch
    px = []
    for r in range(len(num_hits)):
        n, b = int(num_hits[r]) * S, int(indices[r]) * S
        if n == 0:
            px.append(torch.zeros(3, dtype=torch.float64))
            continue
        c, sig, t = rad[b:b + n, :3], rad[b:b + n, 3], t_vals[b:b + n]
        delta = (t - torch.cat([torch.zeros(1, dtype=torch.float64), t[:-1]])).abs()
        ds = delta * sig
        tau = torch.cumsum(ds, 0)
        w = torch.exp(-tau) * (1.0 - torch.exp(-ds))
        px.append((w[:, None] * c).sum(0))
    return torch.stack(px)


@pytest.mark.parametrize("S", [4, 32, 5])
def test_analytic_backward_vs_autograd(cuda_api, S):
    import torch
    api = cuda_api
    rng = np.random.default_rng(S)
    nh = rng.integers(0, 7, 97).astype(np.int32)
    nh[5] = 0
    nh[11] = 40 if S <= 5 else 3                      # a ray longer than one 32-sample chunk
    idx = (np.cumsum(nh) - nh).astype(np.int32)
    n = int(nh.sum()) * S
    rad = np.concatenate([rng.uniform(0, 1, (n, 3)), rng.uniform(0, 20, (n, 1))], 1).astype(np.float32)
    tv = np.tile((np.arange(S, dtype=np.float32) + 1) / S, int(nh.sum()))            # sampler.cu:65-66 t_vals
    g = rng.uniform(-1, 1, (97, 3)).astype(np.float16)
    out = torch.zeros((n, 4), dtype=torch.float16, device="cuda")
    api.launch_volrender_backward_cuda(None, _t(g), _t(rad), _t(tv), _t(nh), _t(idx), 97, S, out,
                                       backward_mode=api.BACKWARD_ANALYTIC)
    torch.cuda.synchronize()
    r64 = torch.tensor(rad.astype(np.float64), requires_grad=True)
    px = _forward_f64(r64, torch.tensor(tv.astype(np.float64)), nh, idx, S)
    (px * torch.tensor(g.astype(np.float64))).sum().backward()
    ref = r64.grad.numpy()
    got = out.cpu().numpy().astype(np.float64)
    # fp32 arithmetic + one fp16 rounding of the result: 1e-3 relative, 1e-4 of the largest entry absolute
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=1e-4 * np.abs(ref).max())
    # and it is NOT what the reference's formula gives (otherwise the mode would be pointless)
    out_ref = torch.zeros((n, 4), dtype=torch.float16, device="cuda")
    api.launch_volrender_backward_cuda(None, _t(g), _t(rad), _t(tv), _t(nh), _t(idx), 97, S, out_ref)
    torch.cuda.synchronize()
    assert np.abs(out_ref.cpu().numpy().astype(np.float64) - ref).max() > 1e-2 * np.abs(ref).max()


def _psnr(a, b):
    return float(-10.0 * np.log10(np.mean((a - b) ** 2)))


def test_training_with_analytic_backward_raises_psnr(cuda_api):
    """200 Adam steps on one 64x48 view of the lego grid against a synthetic target image (colour ramp +
    stripes): the PSNR over the pixels that hit the grid must rise by more than 6 dB.  This is the training step
    shown to train; the reference-mode backward run beside it is reported, not asserted."""
    import torch
    from rtx_nerf_b200 import pipeline
    api = cuda_api
    G, W, H, S, HM = 128, 64, 48, 4, 16
    R = W * H
    occ = synth.occupancy_lego(G)
    grid = api.OccupancyGrid.from_occupancy(occ, G)
    focal = synth.focal_length()
    pose = synth.poses(1, seed=4)[0]
    lv = synth.hashgrid_levels()
    # target: a smooth colour ramp with a stripe pattern over the pixels whose rays hit the grid (a ray without
    # hits composites to 0 whatever the parameters)
    probe = pipeline.ViewRenderer(grid, W, H, focal, W / H, S, HM,
                                  pipeline.NetworkRadiance(api.Network(lv, synth.hashgrid_table(lv, seed=1),
                                                                       synth.mlp_weights(seed=2))))
    probe.render_host(pose)
    hit = probe.hits.num_hits.cpu().numpy() > 0
    assert hit.mean() > 0.3
    yy, xx = np.mgrid[0:H, 0:W]
    target = np.stack([xx / W, yy / H, 0.5 + 0.4 * np.sin(xx / 4.0)], -1).reshape(-1, 3).astype(np.float32)
    target[~hit] = 0.0

    def run(mode):
        student = api.Network(lv, synth.hashgrid_table(lv, seed=1), synth.mlp_weights(seed=2))
        tr = pipeline.Trainer(grid, student, W, H, focal, W / H, S, HM, R, loss_scale=1024.0, lr=1e-2,
                              backward_mode=mode)
        rs = pipeline.ViewRenderer(grid, W, H, focal, W / H, S, HM, pipeline.NetworkRadiance(student))
        p0 = _psnr(rs.render_host(pose).numpy()[hit], target[hit])
        pose_dev, tgt = _t(pose.reshape(16)), _t(target)
        for _ in range(200):
            tr.step(pose_dev, 0, tgt)
        torch.cuda.synchronize()
        assert torch.isfinite(tr.master).all()
        return p0, _psnr(rs.render_host(pose).numpy()[hit], target[hit])

    p0, p1 = run(api.BACKWARD_ANALYTIC)
    q0, q1 = run(api.BACKWARD_REFERENCE)
    print("PSNR analytic backward %.2f -> %.2f dB; reference-formula backward %.2f -> %.2f dB" % (p0, p1, q0, q1))
    assert p1 > p0 + 6.0, (p0, p1)
