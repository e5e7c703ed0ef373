"""the device-resident optimiser loop of `slow_depth`
(src/simple_depth.jl:1-62) against the float64 oracle loop."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ETA = 3e-4


def test_adam_update_vs_oracle():
    """mdb_adam_update == Flux.Optimise.update!(ADAM) on a list of tensors of unequal sizes, 7 steps,
    gradients spanning 1e-9 .. 1 (below and above eps)."""
    import mdb200
    from oracle import torch_ref as T
    g = torch.Generator().manual_seed(3)
    shapes = [(2, 1, 13, 31), (2, 3), (2, 1, 3), (5,)]
    p64 = [torch.randn(s, generator=g, dtype=torch.float64) for s in shapes]
    pg = [p.to(DEV, torch.float32).contiguous() for p in p64]
    p64 = [p.float().double() for p in p64]
    opt, state = mdb200.ADAM(ETA), {}
    for it in range(7):
        grads = [torch.randn(s, generator=g, dtype=torch.float64) * 10.0 ** (-(it * 3) % 10) for s in shapes]
        grads = [a.float().double() for a in grads]
        T.flux_adam_step(p64, grads, state, eta=ETA)
        opt.update(pg, [a.to(DEV, torch.float32).contiguous() for a in grads])
        assert opt.step == it + 1
        for a, b in zip(pg, p64):
            # one step moves an entry by at most eta; float32 state keeps 1e-6 of that, and the float32
            # parameter itself is re-rounded every step (half an ulp)
            assert (a.double().cpu() - b).abs().max() < (2e-6 * ETA + 1.2e-7 * b.abs().max()) * (it + 1)


@pytest.mark.parametrize("W,H,iters", [(64, 24, 12), (416, 128, 5)])
def test_slow_depth_loop_vs_oracle(W, H, iters):
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    inp = synth.make_loss_inputs(W, H, 1, seed=41, scales=(1.0,))
    x64 = torch.tensor(inp.x, dtype=torch.float64)
    K64, iK64 = torch.tensor(inp.K, dtype=torch.float64), torch.tensor(inp.invK, dtype=torch.float64)
    disp_ref, poses_ref, hist_ref = T.slow_depth(x64, K64, iK64, iters=iters)
    sd = mdb200.SlowDepth(torch.tensor(inp.x, device=DEV), inp.K, inp.invK, max_iters=iters)
    before = mdb200.launch_count()
    sd.run(iters)
    torch.cuda.synchronize()
    assert mdb200.launch_count() - before >= 3 * iters           # the replays are counted
    hist = sd.loss_history.cpu().double().numpy()
    # The first two losses see (almost) identical parameters: BASELINE's 1e-5.  After that the two
    # trajectories separate slowly -- ADAM divides by |g|, and d loss / d disp starts at ~0 (no parallax at
    # tvec = 0), so float32 noise in those near-zero gradients is amplified into the early disparity steps
    # (the single-step behaviour is pinned exactly in test_adam_update_vs_oracle).
    assert np.allclose(hist[:2], np.array(hist_ref[:2]), rtol=1e-5, atol=0), (hist, hist_ref)
    assert np.allclose(hist, np.array(hist_ref), rtol=1e-3, atol=0), (hist, hist_ref)
    # ADAM divides by |g|: where the gradient is ~0 its sign is float32 noise and the entry random-walks by
    # +-eta per step, so only the bulk of the entries is compared here; the exact per-step behaviour is
    # pinned by test_slow_depth_single_step_from_oracle_state below.
    dd = (sd.disp.double().cpu() - disp_ref).abs()
    assert max(p.tvec.abs().max().item() for p in poses_ref) > 0.5 * iters * ETA      # the loop did move
    assert dd.max().item() <= 2.0 * iters * ETA
    assert dd.mean().item() < 0.02 * iters * ETA, dd.mean().item()
    for pg, pr in zip(sd.poses, poses_ref):
        assert (pg.rvec.double().cpu() - pr.rvec).abs().max() < 2e-2 * iters * ETA
        assert (pg.tvec.double().cpu() - pr.tvec).abs().max() < 2e-2 * iters * ETA
    sd.close()


def test_slow_depth_single_step_from_oracle_state():
    """Teacher forcing: load the oracle's (theta, m, v, t) after 8 iterations into the device loop, run ONE
    graph replay, compare with the oracle's 9th iteration.  With the moments given, the current gradient
    enters the step with weight 1 - beta1 = 0.1, so BASELINE's 1e-4 gradient tolerance bounds the step
    error by 1e-5 x max|g| / (sqrt(v) + eps) x eta -- checked as 2 % of eta at the worst entry."""
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    inp = synth.make_loss_inputs(96, 40, 1, seed=43, scales=(1.0,))
    x64 = torch.tensor(inp.x, dtype=torch.float64)
    K64, iK64 = torch.tensor(inp.K, dtype=torch.float64), torch.tensor(inp.invK, dtype=torch.float64)
    n, h, w = 1, inp.H, inp.W
    disp = torch.full((n, 1, h, w), 0.5, dtype=torch.float64)
    poses = [T.Pose(torch.tensor([[0.0, 0.0, 0.01]], dtype=torch.float64), torch.zeros(n, 1, 3, dtype=torch.float64)) for _ in range(2)]
    state = {}

    def one_step():
        leaves = [disp.requires_grad_()] + [t.requires_grad_() for p in poses for t in (p.rvec, p.tvec)]
        loss = T.simple_depth_loss(x64, disp.reshape(n, h, w, 1), poses, K64, iK64)
        grads = torch.autograd.grad(loss, leaves)
        with torch.no_grad():
            for t in leaves:
                t.requires_grad_(False)
            T.flux_adam_step(leaves, grads, state, eta=ETA)
        return float(loss.detach()), leaves

    for _ in range(8):
        one_step()
    # round the checkpoint to float32 (what the device holds) and continue the oracle FROM THAT
    for t in [disp] + [q for p in poses for q in (p.rvec, p.tvec)] + state["m"] + state["v"]:
        t.copy_(t.float().double())
    before = [t.clone() for t in [disp] + [q for p in poses for q in (p.rvec, p.tvec)]]
    sd = mdb200.SlowDepth(torch.tensor(inp.x, device=DEV), inp.K, inp.invK, max_iters=16)
    sd.load_state(before, state["m"], state["v"], state["t"])
    loss_ref, after = one_step()
    sd.run(1)
    torch.cuda.synchronize()
    assert abs(float(sd.loss) - loss_ref) < 1e-5 * loss_ref
    assert abs(float(sd.loss_history[8]) - loss_ref) < 1e-5 * loss_ref          # recorded at index step
    for got, b, a in zip(sd.parameters(), before, after):
        step_ref = (a - b).reshape(-1)
        step_got = (got.double().cpu() - b).reshape(-1)
        assert (step_got - step_ref).abs().max() < 0.02 * ETA, float((step_got - step_ref).abs().max())
        assert (step_got - step_ref).abs().mean() < 2e-3 * ETA
    sd.close()


def test_slow_depth_reference_signature_and_logging():
    """slow_depth(x, ssim, backprojections, projections, invKs, Ks, transfer; ...) with the log hook at
    iteration 1, 5, 10, ... (src/simple_depth.jl:23, 44-60) gives the same trajectory as one run."""
    import mdb200
    from mdb200 import synth
    inp = synth.make_loss_inputs(96, 32, 1, seed=42, scales=(1.0,))
    x = torch.tensor(inp.x, device=DEV)
    seen = []
    d1, p1, h1 = mdb200.slow_depth(x, mdb200.SSIM(), None, None, inp.invK, inp.K, None, target_id=2, source_ids=[1, 3],
                                   min_depth=0.1, max_depth=100.0, iters=20,
                                   on_log=lambda it, disp, poses, loss: seen.append((it, float(loss))))
    assert [s[0] for s in seen] == [1, 5, 10, 15, 20]
    d2, p2, h2 = mdb200.slow_depth(x, mdb200.SSIM(), None, None, inp.invK, inp.K, None, target_id=2, source_ids=[1, 3],
                                   min_depth=0.1, max_depth=100.0, iters=20)
    assert torch.equal(d1, d2) and torch.equal(h1, h2)             # deterministic, chunking does not matter
    assert float(h1[-1]) < float(h1[0])
    assert abs(seen[-1][1] - float(h1[-1])) == 0.0


def test_error_codes_of_the_new_entry_points():
    """mdb_adam_update / mdb_slow_depth_create / mdb_ingest_triplet_u8 return status codes, never throw."""
    import ctypes as ct
    import mdb200
    from mdb200 import _lib
    h = mdb200.handle()
    lib = _lib.lib
    ap = _lib.AdamParams(3e-4, 0.9, 0.999, 1e-8)
    assert lib.mdb_adam_update(h, 0, None, None, None, None, None, ct.byref(ap), None, None, None, None, 0, None) == -1
    assert b"tensors" in lib.mdb_last_error(h)
    io = _lib.LossIO()
    io.W, io.H, io.C, io.L, io.N, io.S, io.Ls = 64, 32, 3, 3, 1, 2, 2          # two scales: not slow_depth's shape
    io.dw[0], io.dh[0], io.dw[1], io.dh[1] = 32, 16, 64, 32
    p = _lib.LossParams()
    lib.mdb_loss_params_default(ct.byref(p))
    out = ct.c_void_p()
    assert lib.mdb_slow_depth_create(h, ct.byref(io), ct.byref(p), ct.byref(ap), None, 0, None, 0, None, 0, ct.byref(out)) == -2
    io.Ls, io.dw[0], io.dh[0] = 1, 64, 32
    assert lib.mdb_slow_depth_state_bytes(ct.byref(io)) >= 2 * 4 * (64 * 32 + 12)
    assert lib.mdb_slow_depth_create(h, ct.byref(io), ct.byref(p), ct.byref(ap), None, 0, None, 0, None, 0, ct.byref(out)) == -3
    assert lib.mdb_slow_depth_run(None, 1, None) == -1
    assert lib.mdb_slow_depth_destroy(None) == 0
    x = torch.empty(8, device=DEV)
    assert lib.mdb_ingest_triplet_u8(h, None, None, 4, 4, 1, 0, x.data_ptr(), None) == -1
    assert lib.mdb_ingest_triplet_u8(h, x.data_ptr(), None, 0, 4, 1, 0, x.data_ptr(), None) == -2
    with pytest.raises(ValueError):
        mdb200.ingest_triplet_u8(torch.zeros(1, 4, 12, 3, device=DEV), 4)       # float strips: rejected by the glue


def test_slow_depth_500_iterations_vs_oracle():
    """BASELINE config 1 as the reference runs it (src/simple_depth.jl:22: 500 iterations of ADAM(3e-4) from
    disp = 0.5, rvec = [0,0,0.01], tvec = 0) against the float64 oracle loop: the loss history, the final loss and the
    final mean disparity.  The trajectories are not comparable entry by entry (ADAM's sign steps amplify float32 noise
    of near-zero gradients), the functionals are: tolerances stated below."""
    import mdb200
    from mdb200 import synth
    from oracle import torch_ref as T
    inp = synth.make_config("config1")
    iters = 500
    x64 = torch.tensor(inp.x, dtype=torch.float64)
    K64, iK64 = torch.tensor(inp.K, dtype=torch.float64), torch.tensor(inp.invK, dtype=torch.float64)
    disp_ref, poses_ref, hist_ref = T.slow_depth(x64, K64, iK64, iters=iters)
    sd = mdb200.SlowDepth(torch.tensor(inp.x, device=DEV), inp.K, inp.invK, max_iters=iters)
    sd.run(iters)
    torch.cuda.synchronize()
    hist = sd.loss_history.cpu().double().numpy()
    href = np.array(hist_ref)
    rel = np.abs(hist - href) / href
    print(f"slow_depth 500: loss {hist[0]:.6f} -> {hist[-1]:.6f} (oracle {href[0]:.6f} -> {href[-1]:.6f}), "
          f"max rel dev of the history {rel.max():.2e}, final {rel[-1]:.2e}; mean disp {float(sd.disp.mean()):.6f} "
          f"(oracle {float(disp_ref.mean()):.6f})")
    # measured: final 0.9 %, median of the history 0.3 %; single iterations in the steep middle part deviate by up
    # to 17 % (the two trajectories pass the same loss level a few iterations apart)
    assert rel[-1] < 2e-2, rel[-1]                         # final loss
    assert np.median(rel) < 1e-2, np.median(rel)           # the history as a whole
    assert rel[:20].max() < 1e-3, rel[:20].max()           # and its start, before ADAM's sign noise separates them
    assert hist[-1] < 0.2 * hist[0]                        # the loop optimises (0.1987 -> ~0.018)
    assert abs(float(sd.disp.mean()) - float(disp_ref.mean())) < 1e-3 * float(disp_ref.mean())
    for pg, pr in zip(sd.poses, poses_ref):     # the synthetic triplet has little parallax: the poses stay ~3e-4
        assert (pg.tvec.double().cpu() - pr.tvec).abs().max() < 1e-4
        assert (pg.rvec.double().cpu() - pr.rvec).abs().max() < 1e-4
    sd.close()


def test_slow_depth_create_rejects_unsafe_params():
    """The graph forks the disparity's ADAM kernel off after the fused kernel, which is only safe without the
    mean-normalisation constant (added to d_disparity later): mdb_slow_depth_create refuses anything else."""
    import ctypes as ct
    import mdb200
    from mdb200 import synth, _lib
    inp = synth.make_loss_inputs(64, 24, 1, seed=41, scales=(1.0,))
    sd = mdb200.SlowDepth(torch.tensor(inp.x, device=DEV), inp.K, inp.invK, max_iters=2)
    sd.plan.params.normalize_disparity = 1
    h2 = ct.c_void_p()
    rc = _lib.lib.mdb_slow_depth_create(mdb200.handle(), ct.byref(sd.plan.io), ct.byref(sd.plan.params), ct.byref(sd.adam),
                                        ct.c_void_p(sd.state.data_ptr()), sd.state.numel(), ct.c_void_p(sd.plan.ws.data_ptr()),
                                        sd.plan.ws_bytes, ct.c_void_p(sd.loss_history.data_ptr()), 2, ct.byref(h2))
    assert rc == -1 and b"normalize_disparity" in _lib.lib.mdb_last_error(mdb200.handle())
    sd.close()


def test_slow_depth_batch_of_two():
    """N > 1 takes finalize_item_kernel's done-counter branch (last CTA sums the items, advances the ADAM step):
    two copies of the same triplet must follow the single-item trajectory, with the loss the mean over the batch."""
    import mdb200
    from mdb200 import synth
    inp = synth.make_loss_inputs(64, 24, 1, seed=41, scales=(1.0,))
    x1 = torch.tensor(inp.x, device=DEV)
    x2 = torch.cat([x1, x1], 0).contiguous()
    a = mdb200.SlowDepth(x1, inp.K, inp.invK, max_iters=6)
    b = mdb200.SlowDepth(x2, inp.K, inp.invK, max_iters=6)
    a.run(6); b.run(6)
    torch.cuda.synchronize()
    ha, hb = a.loss_history.cpu().numpy(), b.loss_history.cpu().numpy()
    print("batch of two:", ha, hb)
    # as in test_slow_depth_loop_vs_oracle: the first loss sees identical parameters, later ones feel ADAM's
    # sign steps on near-zero gradients (whose float32 noise differs between the two launch geometries)
    assert np.allclose(ha[:1], hb[:1], rtol=2e-6), (ha, hb)
    assert np.allclose(ha, hb, rtol=1e-3), (ha, hb)
    for n in range(2):
        assert (b.disp[n] - a.disp[0]).abs().max() <= 2.5 * 6 * ETA      # ADAM's +-eta random walk where g ~ 0
        assert (b.disp[n] - a.disp[0]).abs().mean() < 0.05 * 6 * ETA
    a.close(); b.close()
