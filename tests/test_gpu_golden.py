"""The CUDA path against the committed fixtures alone (tests/golden/*.npz): no oracle code is
imported here, so this test also runs where oracle/ is absent.  Tolerances are BASELINE.json's."""
import numpy as np
import pytest
import torch

import golden_io as G
import parity as P

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("name", G.LOSS_CASES)
def test_fused_loss_vs_golden(name):
    import mdb200
    inp, gold, simple = G.load_loss_case(name)
    dev = torch.device(DEV)
    x = torch.tensor(inp.x, device=dev)
    disps = [torch.tensor(d, device=dev) for d in inp.disparities]
    poses = [mdb200.Pose(torch.tensor(r, device=dev), torch.tensor(t, device=dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
    plan = mdb200.FusedLoss(W=inp.W, H=inp.H, C=inp.C, L=3, N=inp.N,
                            disp_sizes=[(d.shape[-1], d.shape[-2]) for d in inp.disparities], scales=inp.scales,
                            target_id=inp.target_id, source_ids=inp.source_ids, K=inp.K, invK=inp.invK,
                            automasking=inp.automasking, normalize_disparity=not simple,
                            smooth_weight_mode=1 if simple else 0)
    al = None if inp.auto_loss is None else torch.tensor(inp.auto_loss, device=dev, dtype=torch.float32).contiguous()
    res = plan(x, disps, poses, al)
    torch.cuda.synchronize()
    assert abs(float(res.loss) - gold["loss"]) <= P.LOSS_RTOL * abs(gold["loss"])
    for k, (g, r) in enumerate(zip(res.d_disparities, gold["d_disp"])):
        P.assert_dense_grad(f"d_disp[{k}]", g, torch.tensor(r))
    for s in range(2):
        P.assert_small_grad(f"d_rvec[{s}]", res.d_poses[s].rvec, torch.tensor(gold["d_rvec"][s]))
        P.assert_small_grad(f"d_tvec[{s}]", res.d_poses[s].tvec, torch.tensor(gold["d_tvec"][s]))


def test_render_vs_golden():
    import mdb200
    z = G.load_render_case()
    dv = lambda a: torch.tensor(a, device=DEV).contiguous()
    ins = [dv(z[k]).requires_grad_() for k in ("rgb", "sigma", "disparity_src", "xyz_tgt", "rvec", "tvec")]
    rg, dg, mg = mdb200.render_tgt_rgb_depth(ins[0], ins[1], ins[2], ins[3], mdb200.Pose(ins[4], ins[5]), z["K_inv"], z["K"])
    assert np.abs(rg.detach().double().cpu().numpy() - z["rgb_out"]).max() < 1e-5
    assert np.abs(dg.detach().double().cpu().numpy() - z["depth_out"]).max() < 1e-5
    assert np.array_equal(mg.cpu().long().numpy(), z["mask_out"])
    ((rg * dv(z["w_rgb"]).float()).sum() + (dg * dv(z["w_depth"]).float()).sum()).backward()
    for t, k in zip(ins, ("rgb", "sigma", "disparity_src", "xyz_tgt", "rvec", "tvec")):
        ref = z["d_" + k]
        err = np.abs(t.grad.double().cpu().numpy() - ref).max() / np.abs(ref).max()
        assert err < 1e-4, (k, float(err))


def test_pose_derivative_scenario():
    """test/runtests.jl:124-142 (prints, asserts nothing): composeT forward and pullback through the C ABI
    on the scenario's numbers against the recorded oracle values."""
    import mdb200
    pd = G.known_answers()["pose_derivative"]
    r = torch.tensor([pd["r"]], device=DEV, dtype=torch.float32)
    t = torch.zeros(1, 1, 3, device=DEV)
    x = torch.tensor(pd["x"], device=DEV, dtype=torch.float32)
    target = torch.tensor(pd["target"], device=DEV, dtype=torch.float32)
    R, tt = mdb200.composeT(r, t, False)                      # Julia layout: R[n, j, i]
    res = R[0].T @ x + tt[0, 0] - target
    l = torch.sqrt((res ** 2).sum())
    assert abs(float(l) - pd["loss"]) < 1e-5 * pd["loss"]
    dRm = torch.outer(res / l, x)                              # math-natural d l / d R[i][j]
    dr, dt = mdb200.composeT_pullback(r, t, False, dRm.T.contiguous()[None], (res / l).reshape(1, 1, 3).contiguous())
    assert np.allclose(dr.cpu().numpy()[0], pd["d_r"], atol=1e-4 * max(map(abs, pd["d_r"])))
    assert np.allclose(dt.cpu().numpy().reshape(3), pd["d_t"], atol=1e-5)
