"""Shared helpers for the GPU parity tests: run the oracle (float64, autograd) and the CUDA
path on identical inputs and compare under the tolerances BASELINE.json states
(loss 1e-5 relative; disparity / pose gradients 1e-4 relative to the tensor's max-abs)."""
import numpy as np
import torch

from oracle import torch_ref as T

LOSS_RTOL = 1e-5
GRAD_RTOL = 1e-4
# The pullback has measure-zero discontinuities (arg-min ties of min(), sign(0) of the L1 term,
# bilinear cell edges).  The kernel takes the geometric decisions in double so they agree with
# the float64 oracle; an exact float32 tie of two per-source losses can still pick the other
# arg-min, so a handful of entries (never more than MAX_OUTLIERS) may disagree.
MAX_OUTLIERS = 2


def oracle_train_loss(inp, dtype=torch.float64, normalize=True, smooth_mode=0, n_global=None):
    """float64 oracle value + gradients for a synth.LossInputs."""
    x = torch.tensor(inp.x, dtype=dtype)
    disps = [torch.tensor(d, dtype=dtype, requires_grad=True) for d in inp.disparities]
    poses = [T.Pose(torch.tensor(r, dtype=dtype, requires_grad=True), torch.tensor(t, dtype=dtype, requires_grad=True))
             for r, t in zip(inp.rvecs, inp.tvecs)]
    K = torch.tensor(inp.K.astype(np.float64), dtype=dtype)
    invK = torch.tensor(inp.invK.astype(np.float64), dtype=dtype)
    auto = None
    if inp.automasking:
        auto = T.automasking_loss(x, x[:, inp.target_id - 1], inp.source_ids)
    if normalize and smooth_mode == 0:
        P = T.Params(automasking=inp.automasking, target_size=(inp.W, inp.H))
        loss = T.train_loss(x, disps, poses, auto, K, invK, P, target_id=inp.target_id,
                            source_ids=inp.source_ids, scales=inp.scales)
    else:
        assert len(disps) == 1
        loss = T.simple_depth_loss(x, disps[0], poses, K, invK, target_id=inp.target_id, source_ids=inp.source_ids)
    gs = torch.autograd.grad(loss, disps + [p.rvec for p in poses] + [p.tvec for p in poses])
    nd = len(disps)
    S = len(poses)
    return dict(loss=float(loss.detach()), d_disp=list(gs[:nd]), d_rvec=list(gs[nd:nd + S]), d_tvec=list(gs[nd + S:]),
                auto=auto)


def grad_report(name, got, ref, rtol=GRAD_RTOL, scale_floor=0.0):
    got = got.detach().double().cpu().reshape(-1)
    ref = ref.detach().double().cpu().reshape(-1)
    scale = max(ref.abs().max().item(), scale_floor)
    err = (got - ref).abs() / max(scale, 1e-300)
    n_out = int((err > rtol).sum())
    inl = err[err <= rtol]
    return dict(name=name, scale=scale, max_err=float(err.max()), n_out=n_out, numel=err.numel(),
                max_inlier=float(inl.max()) if inl.numel() else 0.0,
                rel_l2=float((got - ref).norm() / max(ref.norm().item(), 1e-300)))


def assert_dense_grad(name, got, ref, rtol=GRAD_RTOL, scale_floor=0.0):
    r = grad_report(name, got, ref, rtol, scale_floor)
    assert r["n_out"] <= MAX_OUTLIERS, f"{r}"
    if scale_floor == 0.0:
        assert r["rel_l2"] < 1e-3, f"{r}"
    return r


def assert_small_grad(name, got, ref, rtol=GRAD_RTOL):
    r = grad_report(name, got, ref, rtol)
    assert r["n_out"] == 0, f"{r}"
    return r
