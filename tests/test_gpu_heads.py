"""PoseDecoder tail (src/pose_decoder.jl:27-32) and find_static's statistic
(src/dtk.jl:51-69) against the float64 oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("N,h,w", [(12, 4, 13), (1, 1, 1), (3, 10, 32)])
def test_pose_head_forward_and_pullback(N, h, w):
    import mdb200
    from oracle import torch_ref as T
    g = torch.Generator().manual_seed(N * 100 + w)
    y64 = torch.randn(N, 6, h, w, generator=g, dtype=torch.float64).float().double().requires_grad_()
    pr = T.pose_head(y64)
    wr, wt = torch.randn(N, 3, generator=g, dtype=torch.float64), torch.randn(N, 1, 3, generator=g, dtype=torch.float64)
    (gy,) = torch.autograd.grad((pr.rvec * wr).sum() + (pr.tvec * wt).sum(), y64)
    y = y64.detach().float().to(DEV).requires_grad_()
    pg = mdb200.pose_head(y)
    assert pg.rvec.shape == (N, 3) and pg.tvec.shape == (N, 1, 3)
    assert (pg.rvec.double().cpu() - pr.rvec.detach()).abs().max() < 1e-6 * max(pr.rvec.abs().max().item(), 1e-3)
    assert (pg.tvec.double().cpu() - pr.tvec.detach()).abs().max() < 1e-6 * max(pr.tvec.abs().max().item(), 1e-3)
    ((pg.rvec * wr.float().to(DEV)).sum() + (pg.tvec * wt.float().to(DEV)).sum()).backward()
    assert (y.grad.double().cpu() - gy).abs().max() < 1e-6 * gy.abs().max()


def test_find_static_statistic_and_threshold():
    import mdb200
    from mdb200 import synth
    from oracle import ingest_ref as I
    from oracle import torch_ref as T
    inp = synth.make_loss_inputs(104, 32, 6, seed=51)
    x = inp.x.copy()
    x[1, 0] = x[1, 1]; x[1, 2] = x[1, 1]                 # a static sample: all three frames equal -> loss 0
    strips = synth.to_u8_strips(x)
    xq = I.batch(strips, inp.W)                          # what the device sees after ingestion
    ref = T.find_static_means(torch.tensor(xq, dtype=torch.float64)).numpy()
    xd = mdb200.ingest_triplet_u8(torch.tensor(strips, device=DEV), inp.W)
    alpha = float(np.sort(ref)[2] + np.sort(ref)[3]) / 2   # splits the batch
    keep, means = mdb200.find_static(xd, alpha)
    assert np.allclose(means.cpu().numpy(), ref, rtol=1e-5, atol=1e-7)
    assert means[1].item() < 1e-6 and not keep[1].item()        # SSIM(x, x) = 0 up to the approximate reciprocal
    assert np.array_equal(keep.cpu().numpy(), ref > alpha)
    assert torch.equal(means, mdb200.find_static(xd, alpha)[1])      # deterministic


@pytest.mark.parametrize("N,per_item", [(3, 1), (2, 7), (5, 1031), (4, 64 * 24), (2, 320 * 1024 + 3)])
def test_item_mean_cluster_reduction(N, per_item):
    """8-CTA cluster + distributed shared memory: odd sizes, unaligned slabs, fewer elements than CTAs."""
    import mdb200
    g = torch.Generator().manual_seed(per_item)
    m = torch.randn(N, per_item, generator=g, dtype=torch.float32)
    got = mdb200.item_mean(m.to(DEV))
    ref = m.double().mean(dim=1)
    assert (got.double().cpu() - ref).abs().max() < 2e-7 * max(1.0, m.abs().max().item())
    assert torch.equal(got, mdb200.item_mean(m.to(DEV)))                  # deterministic
    # an unaligned base pointer
    pad = torch.zeros(N * per_item + 1, device=DEV)
    pad[1:] = m.reshape(-1).to(DEV)
    assert torch.equal(mdb200.item_mean(pad[1:].view(N, per_item)), got)


@pytest.mark.parametrize("sizes", [[(52, 16), (104, 32), (208, 64), (416, 128)], [(13, 7), (27, 15)]])
def test_disparity_heads_vs_torch(sizes):
    """DepthDecoder output heads (src/depth_decoder.jl:46, 52-68): σ of every level's conv output in one launch, and
    the pullback g σ (1 - σ); against float64 torch.sigmoid (the second shape has element counts that are not multiples
    of four: the scalar path)."""
    import mdb200
    g = torch.Generator().manual_seed(17)
    N = 3
    logits = [torch.randn((N, 1, h, w), generator=g) * 6 for (w, h) in sizes]     # |x| up to ~25: the overflow-safe branch
    cots = [torch.randn((N, 1, h, w), generator=g) for (w, h) in sizes]
    l64 = [t.double().requires_grad_() for t in logits]
    sum((torch.sigmoid(t) * c.double()).sum() for t, c in zip(l64, cots)).backward()
    lg = [t.to("cuda:0").requires_grad_() for t in logits]
    before = mdb200.launch_count()
    outs = mdb200.disparity_heads(lg)
    assert mdb200.launch_count() - before == 1
    for o, t in zip(outs, l64):
        assert (o.double().cpu() - torch.sigmoid(t.detach())).abs().max() < 2e-7
        assert float(o.min()) >= 0.0 and float(o.max()) <= 1.0
    sum((o * c.to("cuda:0")).sum() for o, c in zip(outs, cots)).backward()
    for a, b in zip(lg, l64):
        assert (a.grad.double().cpu() - b.grad).abs().max() <= 1e-6 * b.grad.abs().max() + 1e-9
