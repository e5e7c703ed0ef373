"""Pullbacks of the fine-grained operators (SURVEY 8b export list: mdb_warp_bwd, mdb_prediction_loss_bwd,
mdb_ssim_bwd, mdb_apply_mask_bwd, mdb_sample_bwd, mdb_plane_volume_rendering_bwd) against autograd of the float64
oracle, and the closure of `slow_depth` (src/simple_depth.jl:25-41) differentiated function by function -- as Zygote
does in the reference -- against the fused one-pass gradient.  Gradient gate: 1e-4 of the tensor's max-abs."""
import numpy as np
import pytest
import torch

import parity as P

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(got, ref):
    ref = ref.double().cpu()
    return float((got.double().cpu() - ref).abs().max() / ref.abs().max())


def test_ssim_bwd_vs_oracle():
    import mdb200
    from oracle import torch_ref as T
    g = torch.Generator().manual_seed(5)
    x = torch.rand((2, 3, 21, 37), generator=g)
    y = (x + 0.05 * torch.randn(x.shape, generator=g)).clamp(0, 1)
    x[:, :, :6, :9] = 0.3          # a flat patch: the float32 cancellation case of E[x^2] - mu^2
    y[:, :, :6, :9] = 0.31
    cot = torch.randn(x.shape, generator=g)
    x64 = x.double().requires_grad_()
    (T.ssim(x64, y.double()) * cot.double()).sum().backward()
    xg = x.to(DEV).requires_grad_()
    out = mdb200.SSIM()(xg, y.to(DEV))
    (out * cot.to(DEV)).sum().backward()
    assert _rel(xg.grad, x64.grad) < P.GRAD_RTOL


@pytest.mark.parametrize("C,S", [(3, 2), (1, 2), (3, 1)])
def test_prediction_loss_bwd_vs_oracle(C, S):
    import mdb200
    from oracle import torch_ref as T
    g = torch.Generator().manual_seed(6)
    tgt = torch.rand((2, C, 18, 29), generator=g)
    preds = [(tgt + 0.1 * torch.randn(tgt.shape, generator=g)).clamp(0, 1) for _ in range(S)]
    cot = torch.rand((2, 1, 18, 29), generator=g)
    p64 = [p.double().requires_grad_() for p in preds]
    (T.prediction_loss(p64, tgt.double()) * cot.double()).sum().backward()
    pg = [p.to(DEV).requires_grad_() for p in preds]
    out = mdb200.prediction_loss(mdb200.SSIM(), pg, tgt.to(DEV))
    assert (out.double().cpu() - T.prediction_loss([p.double() for p in preds], tgt.double())).abs().max() < 1e-5
    (out * cot.to(DEV)).sum().backward()
    for a, b in zip(pg, p64):
        assert _rel(a.grad, b.grad) < P.GRAD_RTOL


def test_apply_mask_bwd():
    import mdb200
    g = torch.Generator().manual_seed(7)
    m, w, cot = (torch.rand((2, 1, 9, 11), generator=g) for _ in range(3))
    w[0, 0, 0, :3] = m[0, 0, 0, :3]                       # ties: the mask keeps them
    mg, wg = m.to(DEV).requires_grad_(), w.to(DEV).requires_grad_()
    (mdb200._apply_mask(mg, wg) * cot.to(DEV)).sum().backward()
    take_w = (w < m)
    assert torch.equal(wg.grad.cpu(), torch.where(take_w, cot, torch.zeros_like(cot)))
    assert torch.equal(mg.grad.cpu(), torch.where(take_w, torch.zeros_like(cot), cot))


def test_warp_bwd_vs_oracle():
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    inp = synth.make_loss_inputs(72, 40, 2, seed=9, scales=(1.0,))
    gen = torch.Generator().manual_seed(8)
    x64 = torch.tensor(inp.x, dtype=torch.float64)
    K64, iK64 = torch.tensor(inp.K, dtype=torch.float64), torch.tensor(inp.invK, dtype=torch.float64)
    d64 = torch.tensor(inp.disparities[0], dtype=torch.float64, requires_grad=True)
    Ps64 = []
    for r, t, sid in zip(inp.rvecs, inp.tvecs, inp.source_ids):
        R, tt = T.composeT_m(torch.tensor(r, dtype=torch.float64), torch.tensor(t, dtype=torch.float64), sid < 2)
        Ps64.append((R.detach().requires_grad_(), tt.detach().requires_grad_()))
    cots = [torch.randn((2, 3, 40, 72), generator=gen) for _ in range(2)]
    w64 = T.warp(d64, x64, Ps64, iK64, K64, min_depth=0.1, max_depth=100.0, source_ids=inp.source_ids)
    sum((w * c.double()).sum() for w, c in zip(w64, cots)).backward()
    # device side: (R, t) in the Julia layouts the C ABI takes: R (N,3,3) indexed [n, j, i], t (N,1,3)
    x = torch.tensor(inp.x, device=DEV)
    dg = torch.tensor(inp.disparities[0], device=DEV).requires_grad_()
    Ps = [(R.detach().transpose(-1, -2).contiguous().float().to(DEV).requires_grad_(),
           tt.detach().reshape(2, 1, 3).float().to(DEV).requires_grad_()) for R, tt in Ps64]
    w = mdb200.warp(dg, x, Ps, None, None, inp.invK, inp.K, min_depth=0.1, max_depth=100.0, source_ids=inp.source_ids)
    for a, b in zip(w, w64):
        assert (a.double().cpu() - b.detach()).abs().max() < 1e-5
    sum((a * c.to(DEV)).sum() for a, c in zip(w, cots)).backward()
    assert _rel(dg.grad, d64.grad) < P.GRAD_RTOL
    for (Rg, tg), (R64, t64) in zip(Ps, Ps64):
        assert _rel(Rg.grad.transpose(-1, -2), R64.grad) < P.GRAD_RTOL
        assert _rel(tg.grad.reshape(2, 3), t64.grad.reshape(2, 3)) < P.GRAD_RTOL


def test_slow_depth_closure_function_by_function():
    """composeT -> warp -> prediction_loss -> mean, + smooth_loss (src/simple_depth.jl:25-41) through the separate
    pullbacks == the fused mdb_train_loss_fwdbwd gradient."""
    import mdb200
    from mdb200 import synth
    inp = synth.make_loss_inputs(96, 40, 1, seed=11, scales=(1.0,))
    x = torch.tensor(inp.x, device=DEV)
    disp = torch.tensor(inp.disparities[0], device=DEV).requires_grad_()
    poses = [mdb200.Pose(torch.tensor(r, device=DEV).requires_grad_(), torch.tensor(t, device=DEV).requires_grad_())
             for r, t in zip(inp.rvecs, inp.tvecs)]
    ssim = mdb200.SSIM()
    target = x[:, 1].contiguous()
    Ps = [mdb200.composeT(p.rvec, p.tvec, sid < 2) for p, sid in zip(poses, inp.source_ids)]
    warped = mdb200.warp(disp, x, Ps, None, None, inp.invK, inp.K, min_depth=0.1, max_depth=100.0, source_ids=inp.source_ids)
    loss = mdb200.prediction_loss(ssim, warped, target).mean() + mdb200.smooth_loss(disp[:, 0], target)
    loss.backward()
    plan = mdb200.simple_depth_plan(x, inp.K, inp.invK)
    res = plan(x, [disp.detach()], [mdb200.Pose(p.rvec.detach(), p.tvec.detach()) for p in poses])
    torch.cuda.synchronize()
    assert abs(float(loss) - float(res.loss)) <= 1e-5 * abs(float(res.loss))
    assert _rel(disp.grad, res.d_disparities[0]) < P.GRAD_RTOL
    for p, g in zip(poses, res.d_poses):
        assert _rel(p.rvec.grad, g.rvec) < P.GRAD_RTOL
        assert _rel(p.tvec.grad, g.tvec) < P.GRAD_RTOL


def test_plane_volume_rendering_bwd_vs_oracle():
    import mdb200
    from oracle import torch_ref as T
    g = torch.Generator().manual_seed(12)
    B, Np, H, W = 2, 7, 10, 13
    rgb = torch.rand((B, Np, 3, H, W), generator=g)
    sigma = torch.rand((B, Np, 1, H, W), generator=g)
    z = torch.sort(torch.rand((B, Np, H, W, 1), generator=g) * 5 + 1, dim=1)[0]
    xyz = torch.cat([torch.randn((B, Np, H, W, 2), generator=g), z], -1)
    cots = [torch.randn((B, 3, H, W), generator=g), torch.randn((B, Np, 1, H, W), generator=g), torch.randn((B, Np, 1, H, W), generator=g)]
    leaves64 = [t.double().requires_grad_() for t in (rgb, sigma, xyz)]
    outs64 = T.plane_volume_rendering(*leaves64)
    sum((o * c.double()).sum() for o, c in zip(outs64, cots)).backward()
    leaves = [t.to(DEV).requires_grad_() for t in (rgb, sigma, xyz)]
    outs = mdb200.plane_volume_rendering(*leaves)
    for o, o64 in zip(outs, outs64):
        assert (o.double().cpu() - o64.detach()).abs().max() < 1e-5
    sum((o * c.to(DEV)).sum() for o, c in zip(outs, cots)).backward()
    for a, b in zip(leaves, leaves64):
        assert _rel(a.grad, b.grad) < P.GRAD_RTOL


def test_sample_bwd_vs_oracle():
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    ri = synth.make_render_inputs(W=24, H=16, Np=4, B=2)
    g = torch.Generator().manual_seed(13)
    src = torch.rand((2, 4, 5, 16, 24), generator=g)
    depth = 1.0 / torch.tensor(ri.disparity_src)
    cot = torch.randn(src.shape, generator=g)
    s64 = src.double().requires_grad_()
    pose64 = T.Pose(torch.tensor(ri.rvec, dtype=torch.float64), torch.tensor(ri.tvec, dtype=torch.float64))
    tgt64, _ = T.sample(s64, depth.double(), pose64, torch.tensor(ri.K, dtype=torch.float64), torch.tensor(ri.K_inv, dtype=torch.float64))
    (tgt64 * cot.double()).sum().backward()
    sg = src.to(DEV).requires_grad_()
    pose = mdb200.Pose(torch.tensor(ri.rvec, device=DEV), torch.tensor(ri.tvec, device=DEV))
    tgt, _ = mdb200.sample(sg, depth.to(DEV).contiguous(), pose, ri.K, ri.K_inv)
    assert (tgt.double().cpu() - tgt64.detach()).abs().max() < 1e-5
    (tgt * cot.to(DEV)).sum().backward()
    assert _rel(sg.grad, s64.grad) < P.GRAD_RTOL
