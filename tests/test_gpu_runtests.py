"""The reference's own test suite (/root/reference/test/runtests.jl) re-stated against the CUDA
path through the C ABI -- same scenarios, same tolerances -- plus oracle comparisons of every
fine-grained entry point."""
import numpy as np
import pytest
import torch
from scipy.spatial.transform import Rotation

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _t(a):
    return torch.tensor(np.asarray(a, dtype=np.float32), device=DEV).contiguous()


def test_rotations():  # runtests.jl:14-29
    import mdb200
    rng = np.random.default_rng(0)
    v = rng.random((6, 3))
    R = mdb200.so3_exp_map(_t(v)).cpu().numpy()          # Julia layout: [n, j, i]
    for n in range(6):
        assert np.allclose(R[n].T, Rotation.from_rotvec(v[n]).as_matrix(), atol=1e-5)


def test_hat_rrule():  # runtests.jl:21
    import mdb200
    from oracle import torch_ref as T
    v = torch.rand(5, 3); d = torch.rand(5, 3, 3)
    assert torch.allclose(mdb200.hat(v.to(DEV)).cpu(), T.hat(v.double()).float())
    assert torch.allclose(mdb200.hat_pullback(d.to(DEV)).cpu(), T.hat_pullback(d.double()).float(), atol=1e-6)


def test_transformation():  # runtests.jl:31-50
    import mdb200
    rng = np.random.default_rng(1)
    rvec, tvec, p = rng.random((1, 3)), rng.random((1, 1, 3)), rng.random(3)
    R, t = mdb200.composeT(_t(rvec), _t(tvec), False)
    R, t = R.cpu().numpy()[0].T, t.cpu().numpy()[0, 0]
    Rs = Rotation.from_rotvec(rvec[0]).as_matrix()
    npnt = R @ p + t
    assert np.allclose(npnt, Rs @ p + tvec[0, 0], atol=1e-6)
    Ri, ti = mdb200.composeT(_t(rvec), _t(tvec), True)
    Ri, ti = Ri.cpu().numpy()[0].T, ti.cpu().numpy()[0, 0]
    assert np.allclose(Ri @ npnt + ti, p, atol=1e-6)


def test_composeT_pullback_vs_oracle():
    import mdb200
    from oracle import torch_ref as T
    for invert in (False, True):
        rv = torch.randn(4, 3) * 0.3; tv = torch.randn(4, 1, 3)
        dR = torch.randn(4, 3, 3); dt = torch.randn(4, 1, 3)
        r64 = rv.double().requires_grad_(); t64 = tv.double().requires_grad_()
        R, t = T.composeT(r64, t64, invert)
        gr, gt = torch.autograd.grad((R * dR.double()).sum() + (t * dt.double()).sum(), (r64, t64))
        dr, dtv = mdb200.composeT_pullback(rv.to(DEV), tv.to(DEV), invert, dR.to(DEV), dt.to(DEV))
        assert torch.allclose(dr.cpu().double(), gr, atol=1e-5, rtol=1e-5)
        assert torch.allclose(dtv.cpu().double(), gt, atol=1e-5, rtol=1e-5)


def test_ssim():  # runtests.jl:52-68
    import mdb200
    ssim = mdb200.SSIM()
    one = torch.ones(1, 1, 2, 2, device=DEV)
    assert torch.allclose(ssim(one, one), torch.zeros_like(one), atol=1e-7)
    assert torch.allclose(ssim(one, torch.zeros_like(one)), torch.full_like(one, 0.5), atol=1e-1)
    a, b = torch.rand(2, 1, 2, 2, device=DEV), torch.rand(2, 1, 2, 2, device=DEV)
    assert torch.allclose(ssim(a, b), ssim(b, a), atol=1e-6)


def test_ssim_vs_oracle_including_flat_regions():
    import mdb200
    from oracle import torch_ref as T
    g = torch.Generator().manual_seed(0)
    x = torch.rand(2, 3, 37, 50, generator=g); y = torch.rand(2, 3, 37, 50, generator=g)
    x[:, :, :12] = 0.6 + 1e-3 * x[:, :, :12]   # nearly flat band: E[x^2]-mu^2 cancels in naive fp32
    y[:, :, :12] = 0.6 + 1e-3 * y[:, :, :12]
    got = mdb200.SSIM()(x.to(DEV), y.to(DEV)).cpu().double()
    assert torch.allclose(got, T.ssim(x.double(), y.double()), atol=2e-6)


def test_smooth_loss():  # runtests.jl:70-83
    import mdb200
    disp = _t(np.array([0.0, 0.2, 0.1, 0.3]).reshape(1, 2, 2))   # Julia (W,H,N) column-major data
    image = torch.ones(1, 1, 2, 2, device=DEV)
    sl = float(mdb200.smooth_loss(disp, image))
    d = disp.cpu()
    tl = float((d[:, :, :-1] - d[:, :, 1:]).abs().mean() + (d[:, :-1] - d[:, 1:]).abs().mean())
    assert abs(sl - tl) < 1e-6
    image = _t(np.array([0.1, 0.3, 0.2, 0.4]).reshape(1, 1, 2, 2))
    assert abs(float(mdb200.smooth_loss(disp, image)) - 0.2542) < 1e-4


@pytest.mark.parametrize("N,C,H,W", [(3, 3, 19, 45),      # W % 4 != 0: one column per thread
                                     (2, 3, 37, 48),      # four columns per thread, three row blocks, partial warp
                                     (1, 1, 9, 1160),     # wider than one CTA's 1024 columns, Gray image
                                     (2, 3, 24, 136)])    # warp seams inside the row
def test_smooth_loss_grad_vs_oracle(N, C, H, W):
    import mdb200
    from oracle import torch_ref as T
    g = torch.Generator().manual_seed(1)
    d = torch.rand(N, H, W, generator=g); im = torch.rand(N, C, H, W, generator=g)
    d64 = d.double().requires_grad_()
    l64 = T.smooth_loss(d64, im.double())
    (g64,) = torch.autograd.grad(l64, d64)
    loss, grad = mdb200.smooth_loss(d.to(DEV), im.to(DEV), with_grad=True)
    assert abs(float(loss) - float(l64)) < 1e-5 * abs(float(l64))
    assert (grad.cpu().double() - g64).abs().max() < 1e-4 * g64.abs().max()


def test_disparity_to_depth():  # runtests.jl:85-92
    import mdb200
    d = mdb200.disparity_to_depth(torch.rand(2, 32, 32, device=DEV), 0.1, 100.0)
    assert d.min() >= 0.1 - 1e-6 and d.max() <= 100.0 + 1e-3


def test_identity_warp():  # runtests.jl:94-122
    import mdb200
    res, N = 16, 2
    g = torch.Generator().manual_seed(0)
    image = torch.rand(N, 1, res, res, generator=g).to(DEV)
    # the reference draws depth ~ U(0,1); a depth below ~1e-3 makes the `+ 1e-7` of the projective
    # divide (src/utils.jl:101) worth more than the test's 1e-3 tolerance, so keep it off zero
    depth = (0.05 + 0.95 * torch.rand(N, res * res, 1, generator=g)).to(DEV)
    K_m = np.array([[910.0, 0, res / 2], [0, 910.0, res / 2], [0, 0, 1]])
    K, invK = K_m.T.copy(), np.linalg.inv(K_m).T.copy()       # Julia layout
    R = mdb200.so3_exp_map(torch.zeros(N, 3, device=DEV)); t = torch.zeros(N, 1, 3, device=DEV)
    pts = mdb200.Backproject(width=res, height=res)(depth, invK)
    uv = mdb200.Project(width=res, height=res)(pts, K, R, t)
    sampled = mdb200.grid_sample(image, uv.reshape(N, res, res, 2))
    assert torch.allclose(image, sampled, atol=1e-3)


def test_warp_and_losses_vs_oracle():
    """warp (P13), photometric_loss, prediction_loss, _apply_mask against the oracle."""
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    inp = synth.make_loss_inputs(104, 32, 2, seed=13)
    x = _t(inp.x); disp = _t(inp.disparities[-1])
    Ps, Ps64 = [], []
    for s, sid in enumerate(inp.source_ids):
        Ps.append(mdb200.composeT(_t(inp.rvecs[s]), _t(inp.tvecs[s]), sid < 2))
        Ps64.append(T.composeT_m(torch.tensor(inp.rvecs[s], dtype=torch.float64), torch.tensor(inp.tvecs[s], dtype=torch.float64), sid < 2))
    warped = mdb200.warp(disp, x, Ps, None, None, inp.invK, inp.K, min_depth=0.1, max_depth=100.0, source_ids=inp.source_ids)
    x64 = torch.tensor(inp.x, dtype=torch.float64)
    w64 = T.warp(torch.tensor(inp.disparities[-1], dtype=torch.float64), x64, Ps64,
                 torch.tensor(inp.invK.astype(np.float64)), torch.tensor(inp.K.astype(np.float64)),
                 min_depth=0.1, max_depth=100.0, source_ids=inp.source_ids)
    for a, b in zip(warped, w64):
        assert (a.cpu().double() - b).abs().max() < 1e-5
    ssim = mdb200.SSIM()
    tgt = x[:, 1].contiguous()
    pl = mdb200.photometric_loss(ssim, warped[0], tgt)
    assert (pl.cpu().double() - T.photometric_loss(w64[0], x64[:, 1])).abs().max() < 1e-5
    pr = mdb200.prediction_loss(ssim, warped, tgt)
    pr64 = T.prediction_loss(w64, x64[:, 1])
    assert (pr.cpu().double() - pr64).abs().max() < 1e-5
    auto = mdb200.automasking_loss(ssim, x, x[:, 1], source_ids=inp.source_ids)
    am = mdb200._apply_mask(auto, pr)
    assert (am.cpu().double() - T.apply_mask(T.automasking_loss(x64, x64[:, 1], inp.source_ids), pr64)).abs().max() < 1e-5


def test_error_convention():
    """Status codes instead of exceptions across the ABI; the glue turns them into errors."""
    import mdb200
    from mdb200 import _lib
    with pytest.raises(TypeError):
        mdb200.so3_exp_map(torch.zeros(2, 3))             # host tensor: rejected by the glue
    with pytest.raises(_lib.MdbError) as e:
        mdb200.SSIM()(torch.zeros(1, 1, 1, 4, device=DEV), torch.zeros(1, 1, 1, 4, device=DEV))  # H < 2
    assert e.value.code == -2 and "W,H >= 2" in str(e.value)
    with pytest.raises(_lib.MdbError):
        mdb200.FusedLoss(W=64, H=32, C=2, L=3, N=1, disp_sizes=[(64, 32)], scales=[1.0], target_id=2,
                         source_ids=[1, 3], K=np.eye(3), invK=np.eye(3))
