"""The pullback of `abs.()` at exactly 0 is a convention of the AD stack (Zygote, unpinned): ChainRules' scalar rule
gives sign(0) = 0, ForwardDiff's Dual rule -- which Zygote's broadcast uses on CuArrays -- gives +1 at +0.0.  It is
NOT measure-zero for `slow_depth`: disp starts as fill(0.5) (src/simple_depth.jl:8), so every smoothness difference
of iteration 1 is exactly +0.0 (src/utils.jl:169-172).  Both settings exist in the kernel (mdb_loss_params
.abs_subgradient), in mdb_smooth_loss_ex and in both oracles; these tests hold them against each other."""
import numpy as np
import pytest
import torch


def _flat_inputs(W=48, H=20, N=1):
    from mdb200 import synth
    inp = synth.make_loss_inputs(W, H, N, seed=5, scales=(1.0,))
    d = np.full((N, 1, H, W), 0.5, np.float32)
    d[:, :, H // 2:, : W // 2] = 0.4            # two plateaus: ties inside, a real edge between them
    inp.disparities = [d]
    return inp


@pytest.mark.parametrize("mode", [0, 1])
def test_c_oracle_matches_torch_oracle(mode):
    from oracle import c_ref
    from oracle import torch_ref as T
    import parity as P
    inp = _flat_inputs()
    T.ABS_SUBGRADIENT = mode
    try:
        ref = P.oracle_train_loss(inp, normalize=False, smooth_mode=1)
    finally:
        T.ABS_SUBGRADIENT = 0
    got = c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, None, inp.K, inp.invK, scales=(1.0,),
                           automasking=False, normalize=False, smooth_mode=1, abs_subgradient=mode)
    assert abs(got["loss"] - ref["loss"]) < 1e-11 * abs(ref["loss"])
    g, r = got["d_disp"][0], ref["d_disp"][0].numpy()
    assert np.abs(g - r).max() <= 1e-9 * np.abs(r).max()


def test_modes_differ_on_plateaus():
    """On a constant disparity the sign(0) = 0 rule yields no smoothness gradient at all; ForwardDiff's rule yields
    w_x(p) - w_x(p-1) + w_y(p) - w_y(p-1) per pixel."""
    from oracle import c_ref
    inp = _flat_inputs()
    inp.disparities = [np.full_like(inp.disparities[0], 0.5)]
    inp.tvecs = [np.zeros_like(t) for t in inp.tvecs]      # no parallax: the photometric gradient vanishes too
    kw = dict(scales=(1.0,), automasking=False, normalize=False, smooth_mode=1)
    a = c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, None, inp.K, inp.invK, abs_subgradient=0, **kw)
    b = c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, None, inp.K, inp.invK, abs_subgradient=1, **kw)
    assert np.abs(a["d_disp"][0]).max() < 1e-7      # only the 1e-7 of the projective divide leaves a trace (~1e-8)
    assert np.abs(b["d_disp"][0]).max() > 1e-5
    assert abs(a["loss"] - b["loss"]) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1])
def test_kernel_matches_oracle(mode):
    import mdb200
    import parity as P
    from oracle import c_ref
    inp = _flat_inputs(96, 40)
    ref = c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, None, inp.K, inp.invK, scales=(1.0,),
                           automasking=False, normalize=False, smooth_mode=1, abs_subgradient=mode)
    dev = torch.device("cuda:0")
    x = torch.tensor(inp.x, device=dev)
    plan = mdb200.simple_depth_plan(x, inp.K, inp.invK, abs_subgradient=mode)
    res = plan(x, [torch.tensor(inp.disparities[0], device=dev)],
               [mdb200.Pose(torch.tensor(r, device=dev), torch.tensor(t, device=dev)) for r, t in zip(inp.rvecs, inp.tvecs)])
    torch.cuda.synchronize()
    assert abs(float(res.loss) - ref["loss"]) <= P.LOSS_RTOL * abs(ref["loss"])
    P.assert_dense_grad("d_disp", res.d_disparities[0], torch.tensor(ref["d_disp"][0]))
    # the stand-alone operator (src/utils.jl:163-177) with the same switch
    tgt = x[:, 1].contiguous()
    d = torch.tensor(inp.disparities[0][:, 0], device=dev)
    loss, grad = mdb200.smooth_loss(d, tgt, with_grad=True, abs_subgradient=mode)
    from oracle import torch_ref as T
    T.ABS_SUBGRADIENT = mode
    try:
        d64 = torch.tensor(inp.disparities[0][:, 0], dtype=torch.float64, requires_grad=True)
        l64 = T.smooth_loss(d64, torch.tensor(inp.x[:, 1], dtype=torch.float64))
        (g64,) = torch.autograd.grad(l64, d64)
    finally:
        T.ABS_SUBGRADIENT = 0
    assert abs(float(loss) - float(l64)) <= 1e-5 * abs(float(l64))
    assert (grad.double().cpu() - g64).abs().max() <= 1e-4 * g64.abs().max()


@pytest.mark.gpu
def test_slow_depth_first_step_depends_on_the_convention():
    """Iteration 1 of slow_depth from fill(0.5): nothing moves the disparity under sign(0) = 0; under ForwardDiff's
    rule ADAM moves (almost) every pixel by eta.  Both against the float64 oracle loop."""
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    inp = synth.make_loss_inputs(64, 24, 1, seed=41, scales=(1.0,))
    x64 = torch.tensor(inp.x, dtype=torch.float64)
    K64, iK64 = torch.tensor(inp.K, dtype=torch.float64), torch.tensor(inp.invK, dtype=torch.float64)
    for mode in (0, 1):
        T.ABS_SUBGRADIENT = mode
        try:
            disp_ref, _, hist_ref = T.slow_depth(x64, K64, iK64, iters=1)
        finally:
            T.ABS_SUBGRADIENT = 0
        sd = mdb200.SlowDepth(torch.tensor(inp.x, device="cuda:0"), inp.K, inp.invK, max_iters=1, abs_subgradient=mode)
        sd.run(1)
        torch.cuda.synchronize()
        moved_ref = (disp_ref - 0.5).abs()
        moved = (sd.disp.double().cpu() - 0.5).abs()
        if mode == 0:
            # the photometric gradient is ~1e-8 noise (the 1e-7 of the projective divide), so ADAM's sign step is
            # not comparable entry by entry; the smoothness term contributes exactly nothing on either side
            assert abs(float(sd.loss_history[0]) - hist_ref[0]) <= 1e-5 * abs(hist_ref[0])
        else:
            big = moved_ref > 0.9 * 3e-4                      # pixels whose smoothness gradient dominates
            assert big.float().mean() > 0.5
            agree = ((sd.disp.double().cpu() - 0.5).sign() == (disp_ref - 0.5).sign())[big]
            assert agree.float().mean() > 0.99, float(agree.float().mean())
            assert float(moved[big].mean()) > 0.9 * 3e-4
        sd.close()
