"""Holds the two CPU oracles against each other: oracle/torch_ref.py (float64 torch, autograd)
and oracle/c_port/mdb_oracle.c (plain C, every pullback written by hand), and the C one against
the known answers of /root/reference/test/runtests.jl that concern it.  Two independent
restatements only agree to 1e-9 if the forward and every derivative agree."""
import numpy as np
import pytest
import torch
from scipy.spatial.transform import Rotation

from mdb200 import synth
from oracle import c_ref
from oracle import torch_ref as T
from parity import oracle_train_loss


@pytest.fixture(scope="module", autouse=True)
def _build():
    c_ref.build()


def _close(name, got, ref, tol):
    got = np.asarray(got, dtype=np.float64).reshape(-1)
    ref = ref.detach().double().numpy().reshape(-1)
    scale = max(np.abs(ref).max(), 1e-300)
    err = np.abs(got - ref).max() / scale
    assert err < tol, f"{name}: max err {err:.3e} (scale {scale:.3e})"


def _c_loss(inp, auto, dtype=np.float64, **kw):
    return c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, auto, inp.K, inp.invK,
                            target_id=inp.target_id, source_ids=inp.source_ids, scales=inp.scales,
                            automasking=inp.automasking, dtype=dtype, **kw)


@pytest.mark.parametrize("W,H,N,auto,seed", [(48, 20, 2, False, 11), (64, 24, 3, True, 12), (33, 17, 1, True, 13)])
def test_train_loss_two_oracles_agree(W, H, N, auto, seed):
    inp = synth.make_loss_inputs(W, H, N, seed=seed, automasking=auto)
    ref = oracle_train_loss(inp)
    a = None if ref["auto"] is None else ref["auto"].numpy()
    got = _c_loss(inp, a, threads=2)
    assert abs(got["loss"] - ref["loss"]) < 1e-11 * abs(ref["loss"])
    for k in range(len(inp.disparities)):
        _close(f"d_disp[{k}]", got["d_disp"][k], ref["d_disp"][k], 1e-9)
    for s in range(2):
        _close(f"d_rvec[{s}]", got["d_rvec"][s], ref["d_rvec"][s], 1e-9)
        _close(f"d_tvec[{s}]", got["d_tvec"][s], ref["d_tvec"][s], 1e-9)


@pytest.mark.parametrize("W,H,scales", [(48, 20, (0.125, 0.25, 0.5, 1.0)), (75, 37, (0.25, 0.5, 1.0))])
def test_half_pixel_upsampling_two_oracles_agree(W, H, scales):
    """`upsample_bilinear` with half-pixel centres (align_corners = false): the C restatement against torch's
    F.interpolate(align_corners=False) + autograd; and it is a different function from the default."""
    from oracle import torch_ref as T
    inp = synth.make_loss_inputs(W, H, 2, seed=21, scales=scales, automasking=True)
    ref_align = oracle_train_loss(inp)
    T.UPSAMPLE_ALIGN_CORNERS = False
    try:
        ref = oracle_train_loss(inp)
    finally:
        T.UPSAMPLE_ALIGN_CORNERS = True
    got = _c_loss(inp, ref["auto"].numpy(), threads=2, upsample_align_corners=False)
    assert abs(got["loss"] - ref["loss"]) < 1e-11 * abs(ref["loss"])
    assert abs(ref_align["loss"] - ref["loss"]) > 1e-5 * abs(ref["loss"])
    for k in range(len(inp.disparities)):
        _close(f"d_disp[{k}]", got["d_disp"][k], ref["d_disp"][k], 1e-9)
    for s in range(2):
        _close(f"d_rvec[{s}]", got["d_rvec"][s], ref["d_rvec"][s], 1e-9)
        _close(f"d_tvec[{s}]", got["d_tvec"][s], ref["d_tvec"][s], 1e-9)


def test_simple_depth_two_oracles_agree():  # src/simple_depth.jl:25-41
    inp = synth.make_loss_inputs(40, 16, 1, seed=14, scales=(1.0,))
    ref = oracle_train_loss(inp, normalize=False, smooth_mode=1)
    got = _c_loss(inp, None, normalize=False, smooth_mode=1)
    assert abs(got["loss"] - ref["loss"]) < 1e-11 * abs(ref["loss"])
    _close("d_disp", got["d_disp"][0], ref["d_disp"][0], 1e-9)
    for s in range(2):
        _close(f"d_rvec[{s}]", got["d_rvec"][s], ref["d_rvec"][s], 1e-9)
        _close(f"d_tvec[{s}]", got["d_tvec"][s], ref["d_tvec"][s], 1e-9)


def test_large_motion_border_samples():
    """poses large enough to push samples over the border: clamp multiplier and tap bounds"""
    inp = synth.make_loss_inputs(48, 20, 2, seed=15)
    inp.rvecs = [r * 6 for r in inp.rvecs]
    inp.tvecs = [t * 8 for t in inp.tvecs]
    ref = oracle_train_loss(inp)
    got = _c_loss(inp, None)
    assert abs(got["loss"] - ref["loss"]) < 1e-11 * abs(ref["loss"])
    for k in range(4):
        _close(f"d_disp[{k}]", got["d_disp"][k], ref["d_disp"][k], 1e-9)
    for s in range(2):
        _close(f"d_rvec[{s}]", got["d_rvec"][s], ref["d_rvec"][s], 1e-9)
        _close(f"d_tvec[{s}]", got["d_tvec"][s], ref["d_tvec"][s], 1e-9)


def test_float32_variant_close_to_float64():
    """the fp32 build (the one the CPU baseline times) is the same algorithm"""
    inp = synth.make_loss_inputs(64, 24, 2, seed=16)
    a = _c_loss(inp, None)
    b = _c_loss(inp, None, dtype=np.float32, threads=2)
    assert abs(a["loss"] - b["loss"]) < 1e-4 * abs(a["loss"])
    g64, g32 = a["d_disp"][3].reshape(-1), b["d_disp"][3].reshape(-1).astype(np.float64)
    assert np.linalg.norm(g64 - g32) < 5e-2 * np.linalg.norm(g64)   # argmin/sign flips allowed, not a new algorithm


def test_threads_do_not_change_the_result():
    inp = synth.make_loss_inputs(40, 16, 5, seed=17)
    a, b = _c_loss(inp, None, threads=1), _c_loss(inp, None, threads=4)
    assert a["loss"] == b["loss"]
    assert all(np.array_equal(x, y) for x, y in zip(a["d_disp"], b["d_disp"]))


# ---- known answers of test/runtests.jl that the C file restates ---------------------------
def test_rotations_c():  # runtests.jl:14-29
    rng = np.random.default_rng(0)
    v = rng.random((5, 3))
    R = c_ref.so3_exp_map(v)
    for n in range(5):
        assert np.allclose(R[n].T, Rotation.from_rotvec(v[n]).as_matrix(), atol=1e-5)


def test_transformation_c():  # runtests.jl:31-50
    rng = np.random.default_rng(1)
    rvec, tvec, p = rng.random((1, 3)), rng.random((1, 1, 3)), rng.random(3)
    R, t = c_ref.compose_T(rvec, tvec, False)
    Ri, ti = c_ref.compose_T(rvec, tvec, True)
    npnt = R[0].T @ p + t[0, 0]
    assert np.allclose(npnt, Rotation.from_rotvec(rvec[0]).as_matrix() @ p + tvec[0, 0], atol=1e-6)
    assert np.allclose(Ri[0].T @ npnt + ti[0, 0], p, atol=1e-6)


def test_compose_T_pullback_c():
    rng = np.random.default_rng(2)
    for invert in (False, True):
        r, t = 0.3 * rng.standard_normal((3, 3)), rng.standard_normal((3, 1, 3))
        dR, dt = rng.standard_normal((3, 3, 3)), rng.standard_normal((3, 1, 3))
        rt, tt = torch.tensor(r, requires_grad=True), torch.tensor(t, requires_grad=True)
        Rj, tj = T.composeT(rt, tt, invert)
        gr, gt = torch.autograd.grad((Rj * torch.tensor(dR)).sum() + (tj * torch.tensor(dt)).sum(), [rt, tt])
        cr, ct = c_ref.compose_T_bwd(r, t, invert, dR, dt)
        assert np.allclose(cr, gr.numpy(), atol=1e-12) and np.allclose(ct, gt.numpy().reshape(-1, 3), atol=1e-12)


def test_smooth_loss_known_answer_c():  # runtests.jl:70-83 -> 0.2542
    """The reference's own vectors.  smooth_loss is reached through the slow_depth closure
    (src/simple_depth.jl:25-41): with all three frames equal and an identity pose the photometric
    term vanishes (SSIM(x,x) = 0, the warp is the identity up to Project's 1e-7)."""
    disp = np.array([[0.0, 0.2], [0.1, 0.3]]).reshape(1, 1, 2, 2)       # Julia column-major 0,.2,.1,.3 as (H,W)
    K, invK = synth.intrinsics(2, 2)
    ident = dict(rvecs=[np.full((1, 3), 1e-9)] * 2, tvecs=[np.zeros((1, 1, 3))] * 2)

    def sl(image):
        x = np.broadcast_to(image.reshape(1, 1, 1, 2, 2), (1, 3, 1, 2, 2)).copy()
        return c_ref.train_loss(x, [disp], ident["rvecs"], ident["tvecs"], None, K, invK, scales=(1.0,), automasking=False,
                                normalize=False, smooth_mode=1, grad=False)["loss"]

    tl = np.abs(disp[..., :, :-1] - disp[..., :, 1:]).mean() + np.abs(disp[..., :-1, :] - disp[..., 1:, :]).mean()
    assert abs(sl(np.ones((2, 2))) - tl) < 1e-6
    assert abs(sl(np.array([[0.1, 0.3], [0.2, 0.4]])) - 0.2542) < 1e-4


# ---- render path: the C restatement (hand-written pullback) against the torch one (autograd) ----------
def _render_case(W, H, Np, B, seed, pose_scale, rand_xyz):
    ri = synth.make_render_inputs(W, H, Np, B, seed=seed, pose_scale=pose_scale)
    rng = np.random.default_rng(seed)
    if rand_xyz:      # test/test_render_tgt_rgb_depth.jl style: xyz_tgt ~ rand, some negative sigma
        ri.xyz_tgt = (rng.random(ri.xyz_tgt.shape) + np.arange(Np)[None, :, None, None, None] * 0.5).astype(np.float32)
        ri.sigma = (rng.random(ri.sigma.shape) - 0.2).astype(np.float32)
    return ri


@pytest.mark.parametrize("W,H,Np,B,pose_scale,rand_xyz", [(24, 16, 6, 2, 2.0, False), (40, 24, 5, 1, 20.0, True), (17, 9, 3, 2, 1.0, True)])
def test_render_two_oracles_agree(W, H, Np, B, pose_scale, rand_xyz):
    ri = _render_case(W, H, Np, B, 70 + W, pose_scale, rand_xyz)
    f = lambda a: torch.tensor(a, dtype=torch.float64, requires_grad=True)
    rgb, sig, dsp, xyz, rv, tv = f(ri.rgb), f(ri.sigma), f(ri.disparity_src), f(ri.xyz_tgt), f(ri.rvec), f(ri.tvec)
    ro, do, mo = T.render_tgt_rgb_depth(rgb, sig, dsp, xyz, T.Pose(rv, tv), torch.tensor(ri.K_inv.astype(np.float64)),
                                        torch.tensor(ri.K.astype(np.float64)))
    g = torch.Generator().manual_seed(1)
    wr, wd = torch.rand(ro.shape, generator=g, dtype=torch.float64), torch.rand(do.shape, generator=g, dtype=torch.float64)
    refs = torch.autograd.grad((ro * wr).sum() + (do * wd).sum(), (rgb, sig, dsp, xyz, rv, tv))
    got = c_ref.render_tgt_rgb_depth(ri.rgb, ri.sigma, ri.disparity_src, ri.xyz_tgt, ri.rvec, ri.tvec, ri.K_inv, ri.K,
                                     g_rgb=wr.numpy(), g_depth=wd.numpy(), threads=3)
    assert np.abs(got["rgb"] - ro.detach().numpy()).max() < 1e-12
    assert np.abs(got["depth"] - do.detach().numpy()).max() < 1e-12
    assert np.array_equal(got["mask"], mo.numpy())
    for name, r in zip(["d_rgb", "d_sigma", "d_disparity_src", "d_xyz_tgt", "d_rvec", "d_tvec"], refs):
        _close(name, got[name], r, 1e-9)


def test_render_c_oracle_reproduces_golden():
    import golden_io as G
    z = G.load_render_case()
    got = c_ref.render_tgt_rgb_depth(z["rgb"], z["sigma"], z["disparity_src"], z["xyz_tgt"], z["rvec"], z["tvec"], z["K_inv"],
                                     z["K"], g_rgb=z["w_rgb"], g_depth=z["w_depth"])
    assert np.abs(got["rgb"] - z["rgb_out"]).max() < 1e-12 and np.abs(got["depth"] - z["depth_out"]).max() < 1e-12
    assert np.array_equal(got["mask"], z["mask_out"])
    for k in ("rgb", "sigma", "disparity_src", "xyz_tgt", "rvec", "tvec"):
        ref = z["d_" + k]
        assert np.abs(got["d_" + k] - ref).max() < 1e-9 * np.abs(ref).max(), k


# ---- the variants (MINE grid / mask, expected depth, producers folded in) -----------
@pytest.mark.parametrize("mine_grid,mine_mask,fused,rand_xyz", [(True, True, False, True), (True, False, True, False),
                                                                (False, True, True, False), (False, False, False, True)])
def test_render_variants_two_oracles_agree(mine_grid, mine_mask, fused, rand_xyz):
    W, H, Np, B = 24, 16, 5, 2
    ri = _render_case(W, H, Np, B, 90, 3.0, rand_xyz and not fused)
    f = lambda a: torch.tensor(a, dtype=torch.float64, requires_grad=True)
    rgb, sig, dsp, xyz, rv, tv = f(ri.rgb), f(ri.sigma), f(ri.disparity_src), f(ri.xyz_tgt), f(ri.rvec), f(ri.tvec)
    ro, do, mo, zo = T.render_tgt_rgb_depth(rgb, sig, dsp, None if fused else xyz, T.Pose(rv, tv),
                                            torch.tensor(ri.K_inv.astype(np.float64)), torch.tensor(ri.K.astype(np.float64)),
                                            mine_grid=mine_grid, mine_mask=mine_mask, depth_eps=1e-5)
    g = torch.Generator().manual_seed(2)
    wr, wd, wz = (torch.rand(t.shape, generator=g, dtype=torch.float64) for t in (ro, do, zo))
    leaves = (rgb, sig, dsp, rv, tv) if fused else (rgb, sig, dsp, xyz, rv, tv)
    refs = torch.autograd.grad((ro * wr).sum() + (do * wd).sum() + (zo * wz).sum(), leaves)
    got = c_ref.render_tgt_rgb_depth(ri.rgb, ri.sigma, ri.disparity_src, None if fused else ri.xyz_tgt, ri.rvec, ri.tvec,
                                     ri.K_inv, ri.K, g_rgb=wr.numpy(), g_depth=wd.numpy(), g_expected_depth=wz.numpy(),
                                     mine_grid=mine_grid, mine_mask=mine_mask, depth_eps=1e-5, threads=3)
    assert np.abs(got["rgb"] - ro.detach().numpy()).max() < 1e-12
    assert np.abs(got["depth"] - do.detach().numpy()).max() < 1e-12
    assert np.abs(got["expected_depth"] - zo.detach().numpy()).max() < 1e-9 * np.abs(zo.detach().numpy()).max()
    assert np.array_equal(got["mask"], mo.numpy())
    names = ["d_rgb", "d_sigma", "d_disparity_src", "d_rvec", "d_tvec"] if fused else \
            ["d_rgb", "d_sigma", "d_disparity_src", "d_xyz_tgt", "d_rvec", "d_tvec"]
    for name, r in zip(names, refs):
        _close(name, got[name], r, 1e-9)


def test_plane_xyz_producers_match_torch():
    ri = synth.make_render_inputs(20, 12, 4, 2, seed=5)
    xs, xt = c_ref.plane_xyz(20, 12, ri.disparity_src, ri.rvec, ri.tvec, ri.K_inv)
    f = lambda a: torch.tensor(a, dtype=torch.float64)
    ts = T.get_src_xyz_from_plane_disparity(T.create_meshgrid(12, 20), f(ri.disparity_src), f(ri.K_inv))
    tt = T.get_tgt_xyz_from_plane_disparity(ts, T.Pose(f(ri.rvec), f(ri.tvec)))
    assert np.abs(xs - ts.numpy()).max() < 1e-12 * np.abs(xs).max()
    assert np.abs(xt - tt.numpy()).max() < 1e-12 * np.abs(xt).max()
