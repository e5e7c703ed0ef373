"""GPU parity of the plane-disparity rendering path (R1-R7, src/render.jl) against the float64
oracle.  BASELINE.json tolerance: rendered rgb / "depth" within 1e-5; gradients 1e-4 relative.
Shapes follow the reference's own scripts (test/test_render_tgt_rgb_depth.jl: B,N,H,W = 2,32,100,200,
inputs ~ rand) and BASELINE config 4 (416x128, 32 planes, B=2)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _dev(a):
    return torch.tensor(a, device=DEV).contiguous()


def _oracle_inputs(ri, grad=False):
    from oracle import torch_ref as T
    f = lambda a: torch.tensor(a, dtype=torch.float64, requires_grad=grad)
    return (f(ri.rgb), f(ri.sigma), f(ri.disparity_src), f(ri.xyz_tgt), T.Pose(f(ri.rvec), f(ri.tvec)),
            torch.tensor(ri.K_inv.astype(np.float64)), torch.tensor(ri.K.astype(np.float64)))


def _rand_inputs(W, H, Np, B, seed, pose_scale):
    """test/test_render_tgt_rgb_depth.jl style: everything ~ rand, pose ~ pose_scale * randn."""
    from mdb200 import synth
    ri = synth.make_render_inputs(W, H, Np, B, seed=seed, pose_scale=pose_scale)
    rng = np.random.default_rng(seed)
    ri.xyz_tgt = rng.random(ri.xyz_tgt.shape).astype(np.float32)
    ri.sigma = (rng.random(ri.sigma.shape) - 0.2).astype(np.float32)   # some negative: exercises sigma .* (sigma .>= 0)
    return ri


@pytest.mark.parametrize("maker", ["config4", "rand_small_pose", "rand_big_pose"])
def test_render_forward(maker):
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    if maker == "config4":
        ri = synth.make_render_inputs(416, 128, 32, 2)
    elif maker == "rand_small_pose":
        ri = _rand_inputs(200, 100, 32, 2, 1, 1.0)
    else:
        ri = _rand_inputs(64, 40, 8, 2, 2, 20.0)     # rvec ~ 0.2, tvec ~ 1: strongly non-identity homographies
    rgb64, sig64, dsp64, xyz64, pose64, Ki64, K64 = _oracle_inputs(ri)
    ro, do, mo = T.render_tgt_rgb_depth(rgb64, sig64, dsp64, xyz64, pose64, Ki64, K64)
    rg, dg, mg = mdb200.render_tgt_rgb_depth(_dev(ri.rgb), _dev(ri.sigma), _dev(ri.disparity_src), _dev(ri.xyz_tgt),
                                             mdb200.Pose(_dev(ri.rvec), _dev(ri.tvec)), ri.K_inv, ri.K)
    assert (rg.double().cpu() - ro).abs().max() < 1e-5
    assert (dg.double().cpu() - do).abs().max() < 1e-5
    assert torch.equal(mg.cpu().long(), mo)


def test_render_backward_vs_autograd():
    import mdb200
    from oracle import torch_ref as T
    ri = _rand_inputs(96, 48, 12, 2, 3, 2.0)
    ri.xyz_tgt = (ri.xyz_tgt + np.arange(12, dtype=np.float32)[None, :, None, None, None] * 0.5).astype(np.float32)
    rgb64, sig64, dsp64, xyz64, pose64, Ki64, K64 = _oracle_inputs(ri, grad=True)
    ro, do, _ = T.render_tgt_rgb_depth(rgb64, sig64, dsp64, xyz64, pose64, Ki64, K64)
    g = torch.Generator().manual_seed(0)
    wr = torch.rand(ro.shape, generator=g, dtype=torch.float64); wd = torch.rand(do.shape, generator=g, dtype=torch.float64)
    loss64 = (ro * wr).sum() + (do * wd).sum()
    refs = torch.autograd.grad(loss64, (rgb64, sig64, dsp64, xyz64, pose64.rvec, pose64.tvec))
    ins = [_dev(a).requires_grad_() for a in (ri.rgb, ri.sigma, ri.disparity_src, ri.xyz_tgt, ri.rvec, ri.tvec)]
    rg, dg, _ = mdb200.render_tgt_rgb_depth(ins[0], ins[1], ins[2], ins[3], mdb200.Pose(ins[4], ins[5]), ri.K_inv, ri.K)
    ((rg * wr.to(DEV, torch.float32)).sum() + (dg * wd.to(DEV, torch.float32)).sum()).backward()
    names = ["d_rgb", "d_sigma", "d_disparity_src", "d_xyz_tgt", "d_rvec", "d_tvec"]
    for nm, t, r in zip(names, ins, refs):
        err = (t.grad.double().cpu() - r).abs().max() / r.abs().max()
        assert err < 1e-4, (nm, float(err))


def test_render_primitives():
    """inv, create_meshgrid, plane xyz producers, sample, plane_volume_rendering vs the oracle."""
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    ri = synth.make_render_inputs(64, 32, 6, 2, seed=5, pose_scale=3.0)
    A = torch.rand(5, 3, 3) + 2 * torch.eye(3)
    assert torch.allclose(mdb200.inv(A.to(DEV)).cpu(), torch.linalg.inv(A.transpose(1, 2)).transpose(1, 2), atol=1e-5)
    mesh = mdb200.create_meshgrid(32, 64)
    assert torch.equal(mesh.cpu().double(), T.create_meshgrid(32, 64))
    pose = mdb200.Pose(_dev(ri.rvec), _dev(ri.tvec))
    pose64 = T.Pose(torch.tensor(ri.rvec, dtype=torch.float64), torch.tensor(ri.tvec, dtype=torch.float64))
    Ki64, K64 = torch.tensor(ri.K_inv.astype(np.float64)), torch.tensor(ri.K.astype(np.float64))
    xs = mdb200.get_src_xyz_from_plane_disparity(mesh, _dev(ri.disparity_src), ri.K_inv)
    xs64 = T.get_src_xyz_from_plane_disparity(T.create_meshgrid(32, 64), torch.tensor(ri.disparity_src, dtype=torch.float64), Ki64)
    assert torch.allclose(xs.cpu().double(), xs64, rtol=1e-5, atol=1e-5)
    xt = mdb200.get_tgt_xyz_from_plane_disparity(xs, pose)
    assert torch.allclose(xt.cpu().double(), T.get_tgt_xyz_from_plane_disparity(xs64, pose64), rtol=1e-5, atol=1e-4)
    src = torch.rand(2, 6, 7, 32, 64)
    depth = 1.0 / torch.tensor(ri.disparity_src)
    tg, vm = mdb200.sample(src.to(DEV), depth.to(DEV), pose, ri.K, ri.K_inv)
    tg64, vm64 = T.sample(src.double(), depth.double(), pose64, K64, Ki64)
    assert (tg.cpu().double() - tg64).abs().max() < 1e-5 and torch.equal(vm.cpu(), vm64)
    rgb, sig, xyz = torch.rand(2, 6, 3, 32, 64), torch.rand(2, 6, 1, 32, 64), torch.rand(2, 6, 32, 64, 3)
    outs = mdb200.plane_volume_rendering(rgb.to(DEV), sig.to(DEV), xyz.to(DEV))
    outs64 = T.plane_volume_rendering(rgb.double(), sig.double(), xyz.double())
    for a, b in zip(outs, outs64):
        assert (a.cpu().double() - b).abs().max() < 1e-5


def test_render_plan_matches_autograd_wrapper():
    """The pre-planned pair of C-ABI calls (what bench.py times) gives the same numbers as the autograd path."""
    import mdb200
    ri = _rand_inputs(96, 48, 12, 2, 5, 2.0)
    ins = [_dev(a).requires_grad_() for a in (ri.rgb, ri.sigma, ri.disparity_src, ri.xyz_tgt, ri.rvec, ri.tvec)]
    rg, dg, mg = mdb200.render_tgt_rgb_depth(ins[0], ins[1], ins[2], ins[3], mdb200.Pose(ins[4], ins[5]), ri.K_inv, ri.K)
    g = torch.Generator(device=DEV).manual_seed(1)
    wr, wd = torch.rand(rg.shape, device=DEV, generator=g), torch.rand(dg.shape, device=DEV, generator=g)
    torch.autograd.backward([rg, dg], [wr, wd])
    plan = mdb200.RenderPlan(96, 48, 12, 2, ri.K_inv, ri.K)
    det = [t.detach() for t in ins]
    r2, d2, m2 = plan.forward(det[0], det[1], det[2], det[3], mdb200.Pose(det[4], det[5]))
    assert torch.equal(r2, rg.detach()) and torch.equal(d2, dg.detach()) and torch.equal(m2, mg)
    grads = plan.backward(wr, wd)
    for got, t in zip(grads, [ins[0], ins[1], ins[2], ins[3], ins[4], ins[5]]):
        ref = t.grad
        assert (got - ref).abs().max() <= 1e-5 * ref.abs().max()      # RED order differs run to run


def test_render_config4_full_size_vs_c_oracle():
    """BASELINE config 4 at full geometry (416x128, 32 planes, B=2), forward AND pullback, against the plain-C
    float64 oracle with its hand-written pullback (oracle/c_port)."""
    import mdb200
    from mdb200 import synth
    from oracle import c_ref
    ri = synth.make_render_inputs(416, 128, 32, 2)
    rng = np.random.default_rng(8)
    wr = rng.random((2, 3, 128, 416)).astype(np.float32)
    wd = (0.1 * rng.random((2, 32, 1, 128, 416))).astype(np.float32)
    ref = c_ref.render_tgt_rgb_depth(ri.rgb, ri.sigma, ri.disparity_src, ri.xyz_tgt, ri.rvec, ri.tvec, ri.K_inv, ri.K,
                                     g_rgb=wr, g_depth=wd, threads=8)
    plan = mdb200.RenderPlan(416, 128, 32, 2, ri.K_inv, ri.K)
    rg, dg, mg = plan.forward(_dev(ri.rgb), _dev(ri.sigma), _dev(ri.disparity_src), _dev(ri.xyz_tgt),
                              mdb200.Pose(_dev(ri.rvec), _dev(ri.tvec)))
    assert np.abs(rg.double().cpu().numpy() - ref["rgb"]).max() < 1e-5
    assert np.abs(dg.double().cpu().numpy() - ref["depth"]).max() < 1e-5
    assert np.array_equal(mg.cpu().numpy(), ref["mask"])
    grads = plan.backward(_dev(wr), _dev(wd))
    for got, name in zip(grads, ["d_rgb", "d_sigma", "d_disparity_src", "d_xyz_tgt", "d_rvec", "d_tvec"]):
        r = ref[name]
        if np.abs(r).max() < 1e-30:      # e.g. d_sigma of a single plane: exp(-1000 sigma) underflows float32 (denormal in float64)
            assert float(got.abs().max()) < 1e-30, name
            continue
        err = np.abs(got.double().cpu().numpy() - r).max() / np.abs(r).max()
        assert err < 1e-4, (name, float(err))


# ---- MINE grid / mask, expected depth, producers folded into the kernels -------------------
def _check_variant(ri, W, H, Np, B, mine_grid, mine_mask, fused, seed=11, threads=8):
    import mdb200
    from oracle import c_ref
    rng = np.random.default_rng(seed)
    wr = rng.random((B, 3, H, W)).astype(np.float32)
    wd = (0.1 * rng.random((B, Np, 1, H, W))).astype(np.float32)
    wz = (0.01 * rng.random((B, 1, H, W))).astype(np.float32)
    ref = c_ref.render_tgt_rgb_depth(ri.rgb, ri.sigma, ri.disparity_src, None if fused else ri.xyz_tgt, ri.rvec, ri.tvec,
                                     ri.K_inv, ri.K, g_rgb=wr, g_depth=wd, g_expected_depth=wz, mine_grid=mine_grid,
                                     mine_mask=mine_mask, depth_eps=1e-5, threads=threads)
    plan = mdb200.RenderPlan(W, H, Np, B, ri.K_inv, ri.K, mine_grid=mine_grid, mine_mask=mine_mask, expected_depth=True,
                             depth_eps=1e-5, fused_xyz=fused)
    rg, dg, mg, zg = plan.forward(_dev(ri.rgb), _dev(ri.sigma), _dev(ri.disparity_src), None if fused else _dev(ri.xyz_tgt),
                                  mdb200.Pose(_dev(ri.rvec), _dev(ri.tvec)))
    assert np.abs(rg.double().cpu().numpy() - ref["rgb"]).max() < 1e-5
    assert np.abs(dg.double().cpu().numpy() - ref["depth"]).max() < 1e-5
    zr = ref["expected_depth"]
    assert np.abs(zg.double().cpu().numpy() - zr).max() < 1e-5 * np.abs(zr).max()
    assert np.array_equal(mg.cpu().numpy(), ref["mask"])
    grads = plan.backward(_dev(wr), _dev(wd), _dev(wz))
    for got, name in zip(grads, ["d_rgb", "d_sigma", "d_disparity_src", "d_xyz_tgt", "d_rvec", "d_tvec"]):
        if fused and name == "d_xyz_tgt":
            assert got is None
            continue
        r = ref[name]
        if np.abs(r).max() < 1e-30:      # e.g. d_sigma of a single plane: exp(-1000 sigma) underflows float32 (denormal in float64)
            assert float(got.abs().max()) < 1e-30, name
            continue
        err = np.abs(got.double().cpu().numpy() - r).max() / np.abs(r).max()
        assert err < 1e-4, (name, float(err))
    return ref


@pytest.mark.parametrize("mine_grid,mine_mask,fused", [(False, False, True), (True, True, False), (True, False, True),
                                                       (False, True, False), (True, True, True)])
def test_render_variants_vs_c_oracle(mine_grid, mine_mask, fused):
    from mdb200 import synth
    W, H, Np, B = 96, 48, 12, 2
    ri = synth.make_render_inputs(W, H, Np, B, seed=21, pose_scale=3.0)
    ri.sigma = (np.random.default_rng(4).random(ri.sigma.shape) - 0.2).astype(np.float32)
    _check_variant(ri, W, H, Np, B, mine_grid, mine_mask, fused)


def test_render_variants_config4_full_size():
    """Config-4 geometry with everything switched on: the sampler MINE intends, its mask, the expected depth and
    no materialised xyz_tgt."""
    from mdb200 import synth
    ri = synth.make_render_inputs(416, 128, 32, 2)
    ref = _check_variant(ri, 416, 128, 32, 2, True, True, True)
    assert ref["mask"].max() == 32 and ref["mask"].min() < 32   # the MINE mask counts the planes that hit the image


def test_render_fused_xyz_equals_materialised_producers():
    """xyz_tgt=None is the same function as feeding the producers' output (src/render.jl:21-30, 51-62), through the
    autograd wrapper: the gradient that would land in d_xyz_tgt flows on into disparity_src and the pose."""
    import mdb200
    from mdb200 import synth
    W, H, Np, B = 64, 40, 8, 2
    ri = synth.make_render_inputs(W, H, Np, B, seed=3, pose_scale=2.0)
    mk = lambda: [_dev(a).requires_grad_() for a in (ri.rgb, ri.sigma, ri.disparity_src, ri.rvec, ri.tvec)]
    a, b = mk(), mk()
    mesh = mdb200.create_meshgrid(H, W, device=torch.device(DEV))
    g = torch.Generator(device=DEV).manual_seed(2)
    outs_a = mdb200.render_tgt_rgb_depth(a[0], a[1], a[2], None, mdb200.Pose(a[3], a[4]), ri.K_inv, ri.K, depth_eps=1e-5)
    w = [torch.rand(o.shape, device=DEV, generator=g) for o in (outs_a[0], outs_a[1], outs_a[3])]
    w[2] *= 0.01
    (outs_a[0] * w[0]).sum().add((outs_a[1] * w[1]).sum()).add((outs_a[3] * w[2]).sum()).backward()
    # materialised: d_xyz_tgt chained by hand through the producers in float64 on the host
    xyz_src = mdb200.get_src_xyz_from_plane_disparity(mesh, b[2].detach(), ri.K_inv)
    xyz_tgt = mdb200.get_tgt_xyz_from_plane_disparity(xyz_src, mdb200.Pose(b[3].detach(), b[4].detach())).requires_grad_()
    outs_b = mdb200.render_tgt_rgb_depth(b[0], b[1], b[2], xyz_tgt, mdb200.Pose(b[3], b[4]), ri.K_inv, ri.K, depth_eps=1e-5)
    for oa, ob in zip((outs_a[0], outs_a[1], outs_a[3]), (outs_b[0], outs_b[1], outs_b[3])):
        assert (oa - ob).abs().max() <= 2e-5 * ob.abs().max()
    assert torch.equal(outs_a[2], outs_b[2])
    (outs_b[0] * w[0]).sum().add((outs_b[1] * w[1]).sum()).add((outs_b[3] * w[2]).sum()).backward()
    assert (a[0].grad - b[0].grad).abs().max() <= 1e-4 * b[0].grad.abs().max()
    assert (a[1].grad - b[1].grad).abs().max() <= 1e-4 * b[1].grad.abs().max()
    # d tvec: the fused path adds sum(d_xyz_tgt) to what the homography path gives
    gt = b[4].grad + xyz_tgt.grad.sum(dim=(1, 2, 3))
    assert (a[4].grad - gt).abs().max() <= 1e-4 * gt.abs().max()


@pytest.mark.parametrize("mine_grid", [False, True])
@pytest.mark.parametrize("pose_scale,expect_fast", [(1.0, True), (3.0, True), (20.0, None), (60.0, False)])   # 20: short path, but every sample lands outside the image
def test_render_forward_float_path_vs_double_path(mine_grid, pose_scale, expect_fast):
    """The forward (and the pullback) sample a plane through the short displacement form wherever the homography's
    denominator 1 + h20 u + h21 v stays within 1 +- 1/4 over the image (render_prep_kernel decides per plane), else through
    the general double-precision homography.  Same inputs through both (test hook "render_precise"): the outputs agree
    far inside the 1e-5 parity gate, masks are equal, and with poses of ~0.6 rad (denominator far from 1 on every plane)
    the two runs are bit-identical: the short form was never taken."""
    import mdb200
    ri = _rand_inputs(416, 128, 16, 2, 31, pose_scale)
    plan = mdb200.RenderPlan(416, 128, 16, 2, ri.K_inv, ri.K, mine_grid=mine_grid, mine_mask=mine_grid)
    args = (_dev(ri.rgb), _dev(ri.sigma), _dev(ri.disparity_src), _dev(ri.xyz_tgt), mdb200.Pose(_dev(ri.rvec), _dev(ri.tvec)))
    a = [t.clone() for t in plan.forward(*args)[:3]]
    mdb200.debug_option("render_precise", 1)
    try:
        b = [t.clone() for t in plan.forward(*args)[:3]]
    finally:
        mdb200.debug_option("render_precise", 0)
    torch.cuda.synchronize()
    d_rgb, d_t = float((a[0] - b[0]).abs().max()), float((a[1] - b[1]).abs().max())
    assert d_rgb < 6e-6 and d_t < 6e-6, (d_rgb, d_t)
    assert torch.equal(a[2], b[2])
    same = torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    if expect_fast is True:
        assert not same          # the float path ran (and differs in the last bits)
    elif expect_fast is False:
        assert same


@pytest.mark.parametrize("seed", range(12))
def test_render_random_geometries(seed):
    """Seeded random shapes (odd sizes, plane counts that do not divide into the forward's four plane groups, B = 1..3), random
    variant flags and pose scales on both sides of the short-path / double-path switch: forward, expected depth, masks
    and the whole pullback against the float64 C oracle."""
    from mdb200 import synth
    rng = np.random.default_rng(500 + seed)
    W, H = int(rng.integers(9, 140)), int(rng.integers(6, 70))
    Np, B = int(rng.integers(1, 19)), int(rng.integers(1, 4))
    pose_scale = float(rng.choice([0.5, 2.0, 6.0, 20.0]))
    mine_grid, mine_mask, fused = bool(rng.integers(0, 2)), bool(rng.integers(0, 2)), bool(rng.integers(0, 2))
    ri = synth.make_render_inputs(W, H, Np, B, seed=600 + seed, pose_scale=pose_scale)
    ri.sigma = (rng.random(ri.sigma.shape) * 0.8 + 0.1).astype(np.float32)    # away from 0: exp(-1000 sigma) of the last plane amplifies 1e-7
    _check_variant(ri, W, H, Np, B, mine_grid, mine_mask, fused, seed=700 + seed, threads=4)
