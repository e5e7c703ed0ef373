"""The oracle's restatement of Flux's ADAM (third-party, unpinned) against torch.optim.Adam, which
implements the same published update, and a short slow_depth run as a sanity check of the loop."""
import torch

from mdb200 import synth
from oracle import torch_ref as T


def test_flux_adam_matches_torch_adam():
    torch.manual_seed(0)
    p0 = [torch.randn(7, 5, dtype=torch.float64), torch.randn(3, dtype=torch.float64)]
    a = [p.clone() for p in p0]
    b = [p.clone().requires_grad_() for p in p0]
    opt = torch.optim.Adam(b, lr=3e-4, betas=(0.9, 0.999), eps=1e-8)
    state = {}
    for it in range(6):
        grads = [torch.randn_like(p) * (10.0 ** -it) for p in p0]
        T.flux_adam_step(a, grads, state, eta=3e-4)
        for q, g in zip(b, grads):
            q.grad = g.clone()
        opt.step()
        for u, w in zip(a, b):
            assert torch.allclose(u, w.detach(), rtol=0, atol=1e-14)


def test_slow_depth_loss_decreases():
    inp = synth.make_loss_inputs(48, 20, 1, seed=31, scales=(1.0,))
    x = torch.tensor(inp.x, dtype=torch.float64)
    K, invK = torch.tensor(inp.K, dtype=torch.float64), torch.tensor(inp.invK, dtype=torch.float64)
    disp, poses, hist = T.slow_depth(x, K, invK, iters=30)
    assert hist[-1] < hist[0]
    assert abs(float(disp.mean()) - 0.5) < 30 * 3e-4 + 1e-9          # ADAM moves each entry by <= eta per step
    assert all(torch.isfinite(p.rvec).all() and torch.isfinite(p.tvec).all() for p in poses)
