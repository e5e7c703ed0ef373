"""sample ingestion (src/dtk.jl:29-46, src/kitty.jl:45-61).  CPU part: the numpy restatement on
hand-checkable inputs.  GPU part: mdb_ingest_triplet_u8 is bit-exact against it, and the host-buffer loss
fed uint8 strips equals the one fed the Float32 stack the oracle makes of the same strips."""
import numpy as np
import pytest
import torch

from mdb200 import synth
from oracle import ingest_ref as I

DEV = "cuda:0"


def test_oracle_layout_and_values():
    W, H = 4, 2
    strip = np.arange(H * 3 * W * 3, dtype=np.uint8).reshape(H, 3 * W, 3)
    out = I.depth10k_getindex(strip, W)
    assert out.shape == (3, 3, H, W) and out.dtype == np.float32
    # frame j, channel c, row h, column w  <-  strip[h, j W + w, c] / 255
    assert out[2, 1, 1, 3] == np.float32(strip[1, 2 * W + 3, 1]) / np.float32(255)
    fl = I.depth10k_getindex(strip, W, flip=True)
    assert np.array_equal(fl, out[..., ::-1])
    g = I.depth10k_getindex(strip, W, grayscale=True)
    r, gg, b = (np.float32(v) / np.float32(255) for v in strip[0, 5])
    assert g.shape == (3, 1, H, W)
    assert g[1, 0, 0, 1] == (np.float32(0.299) * r + np.float32(0.587) * gg) + np.float32(0.114) * b
    assert I.n0f8_to_f32(np.array([255], dtype=np.uint8))[0] == 1.0


def test_synth_strips_round_trip():
    inp = synth.make_loss_inputs(24, 10, 2, seed=5)
    s = synth.to_u8_strips(inp.x)
    back = I.batch(s, 24)
    assert back.shape == inp.x.shape and np.abs(back - inp.x).max() <= 0.5 / 255 + 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("W,H,N,gray", [(416, 128, 3, False), (416, 128, 2, True), (13, 7, 4, False), (1, 1, 1, True),
                                        (1024, 320, 2, False)])
def test_ingest_bit_exact(W, H, N, gray):
    import mdb200
    rng = np.random.default_rng(W * 7 + H)
    strips = rng.integers(0, 256, size=(N, H, 3 * W, 3), dtype=np.uint8)
    flips = (rng.random(N) < 0.5).astype(np.uint8)
    flips[0] = 1
    ref = I.batch(strips, W, flips, grayscale=gray)
    got = mdb200.ingest_triplet_u8(torch.tensor(strips, device=DEV), W, torch.tensor(flips, device=DEV), grayscale=gray)
    assert np.array_equal(got.cpu().numpy(), ref)                      # bit-exact
    got2 = mdb200.ingest_triplet_u8(torch.tensor(strips, device=DEV), W, None, grayscale=gray)
    assert np.array_equal(got2.cpu().numpy(), I.batch(strips, W, None, grayscale=gray))
    # an unaligned base pointer takes the byte path
    pad = torch.zeros(strips.size + 3, dtype=torch.uint8, device=DEV)
    pad[3:] = torch.tensor(strips, device=DEV).reshape(-1)
    got3 = mdb200.ingest_triplet_u8(pad[3:].view(N, H, 3 * W, 3), W, torch.tensor(flips, device=DEV), grayscale=gray)
    assert np.array_equal(got3.cpu().numpy(), ref)


@pytest.mark.gpu
def test_host_loss_from_u8_strips_matches_f32_path():
    import mdb200
    inp = synth.make_loss_inputs(104, 32, 5, seed=9)
    strips = synth.to_u8_strips(inp.x)
    flips = np.array([0, 1, 0, 1, 1], dtype=np.uint8)
    x_ref = I.batch(strips, inp.W, flips)
    kw = dict(W=inp.W, H=inp.H, C=3, L=3, N=inp.N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in inp.disparities],
              scales=inp.scales, target_id=2, source_ids=[1, 3], K=inp.K, invK=inp.invK, chunk=2)
    pin = lambda a: torch.tensor(a).pin_memory()
    d, r, t = [pin(a) for a in inp.disparities], [pin(a) for a in inp.rvecs], [pin(a) for a in inp.tvecs]
    a = mdb200.HostLoss(**kw)(pin(x_ref), d, r, t)
    # the u8 pipeline ramps its chunk sizes up (1, 2, ...), so the chunk losses are summed in another order and an item may
    # sit in a chunk of another size (another row segmentation): equal up to float32 summation order
    b = mdb200.HostLoss(x_format="u8_strip", **kw)(pin(strips), d, r, t, flips=pin(flips))
    assert abs(a.loss - b.loss) <= 1e-6 * abs(a.loss)
    for u, v in zip(a.d_disparities + a.d_rvec + a.d_tvec, b.d_disparities + b.d_rvec + b.d_tvec):
        assert torch.allclose(u, v, rtol=1e-5, atol=1e-6 * float(u.abs().max()))
    plan = mdb200.HostLoss(x_format="u8_strip", **kw)
    assert plan.h2d_bytes < mdb200.HostLoss(**kw).h2d_bytes * 0.45


def test_dump_for_a_julia_run_round_trips(tmp_path):
    """the synthetic inputs on disk as raw little-endian float32 in Julia column-major order."""
    import json
    inp = synth.make_loss_inputs(24, 10, 2, seed=6)
    meta = json.load(open(synth.dump_loss_inputs(inp, str(tmp_path))))
    assert meta["julia_shapes"]["x"] == [24, 10, 3, 3, 2] and meta["julia_shapes"]["disp3"] == [24, 10, 1, 2]
    x = np.fromfile(tmp_path / "x.f32", dtype="<f4")
    # Julia's x[w,h,c,l,n] (1-based, column-major) is the flat element (w-1) + W((h-1) + H((c-1) + C((l-1) + L(n-1))))
    w, h, c, l, n = 5, 3, 2, 1, 2
    assert x[(w - 1) + 24 * ((h - 1) + 10 * ((c - 1) + 3 * ((l - 1) + 3 * (n - 1))))] == inp.x[n - 1, l - 1, c - 1, h - 1, w - 1]
    assert np.array_equal(np.fromfile(tmp_path / "rvec1.f32", dtype="<f4").reshape(2, 3), inp.rvecs[1])


# ---- KittyDataset.getindex (src/kitty.jl:45-61) ---------------------------------------------------------------------
def test_kitti_oracle_hand_checked():
    # 2x down-sampling of a horizontal ramp: output pixel i sits at 2 i - 0.5 (1-based): the mean of two neighbours
    g = np.tile((np.arange(8) * 10).astype(np.uint8), (8, 1))
    assert np.array_equal(I.imresize_gray_n0f8(g, 4, 4)[0], [5, 25, 45, 65])
    # the same size is a copy; a constant image stays constant under any resize
    assert np.array_equal(I.imresize_gray_n0f8(g, 8, 8), g)
    c = np.full((7, 9), 201, np.uint8)
    assert np.unique(I.imresize_gray_n0f8(c, 3, 5)).tolist() == [201]
    # up-sampling clamps the half-pixel positions at both ends (first output = first input, last = last)
    up = I.imresize_gray_n0f8(g, 8, 16)[0]
    assert up[0] == 0 and up[-1] == 70 and np.all(np.diff(up.astype(int)) >= 0)
    # layout: (3, 1, H, W) float32 = u8 / 255, FlipX reverses the width axis of all three frames
    fr = np.random.default_rng(3).integers(0, 256, size=(3, 20, 31), dtype=np.uint8)
    out = I.kitti_getindex(fr, 10, 16)
    assert out.shape == (3, 1, 10, 16) and out.dtype == np.float32
    assert out[1, 0, 4, 7] == np.float32(I.imresize_gray_n0f8(fr[1], 10, 16)[4, 7]) / np.float32(255)
    assert np.array_equal(I.kitti_getindex(fr, 10, 16, flip=True), out[..., ::-1])


@pytest.mark.gpu
@pytest.mark.parametrize("Wo,Ho,W,H,N", [(1241, 376, 416, 128, 3),     # KITTI odometry image_0 -> config 2's size
                                         (1241, 376, 640, 192, 2), (1226, 370, 1024, 320, 1),
                                         (832, 256, 416, 128, 2),      # integer factor: every sample an exact .5 tie candidate
                                         (416, 128, 416, 128, 2),      # same size: a copy
                                         (37, 11, 64, 24, 4),          # up-sampling, clamped ends
                                         (5, 3, 1, 1, 1)])
def test_ingest_kitti_bit_exact(Wo, Ho, W, H, N):
    import mdb200
    rng = np.random.default_rng(Wo + 3 * W)
    frames = rng.integers(0, 256, size=(N, 3, Ho, Wo), dtype=np.uint8)
    flips = (rng.random(N) < 0.5).astype(np.uint8)
    flips[0] = 1
    ref = I.kitti_batch(frames, H, W, flips)
    got = mdb200.ingest_kitti_u8(torch.tensor(frames, device=DEV), H, W, torch.tensor(flips, device=DEV))
    assert got.shape == (N, 3, 1, H, W)
    assert np.array_equal(got.cpu().numpy(), ref)                      # bit-exact
    got2 = mdb200.ingest_kitti_u8(torch.tensor(frames, device=DEV), H, W)
    assert np.array_equal(got2.cpu().numpy(), I.kitti_batch(frames, H, W))


@pytest.mark.gpu
def test_kitti_frames_feed_the_loss():
    """C = 1 stack from the KITTI ingest -> fused loss == the loss on the oracle's stack of the same frames."""
    import mdb200
    rng = np.random.default_rng(11)
    base = synth.make_loss_inputs(104, 32, 2, C=1, seed=12)
    big = np.clip(np.kron(base.x[:, :, 0], np.ones((2, 3))) * 255, 0, 255).astype(np.uint8)     # (N,3,64,312)
    big = np.ascontiguousarray(big + rng.integers(0, 3, size=big.shape, dtype=np.uint8))
    ref_x = I.kitti_batch(big, 32, 104)
    x = mdb200.ingest_kitti_u8(torch.tensor(big, device=DEV), 32, 104)
    assert np.array_equal(x.cpu().numpy(), ref_x)
    sizes = [(d.shape[-1], d.shape[-2]) for d in base.disparities]
    plan = mdb200.FusedLoss(W=104, H=32, C=1, L=3, N=2, disp_sizes=sizes, scales=base.scales, target_id=base.target_id,
                            source_ids=base.source_ids, K=base.K, invK=base.invK)
    poses = [mdb200.Pose(torch.tensor(r, device=DEV), torch.tensor(t, device=DEV)) for r, t in zip(base.rvecs, base.tvecs)]
    disps = [torch.tensor(d, device=DEV) for d in base.disparities]
    a = float(plan(x, disps, poses).loss)
    b = float(plan(torch.tensor(ref_x, device=DEV), disps, poses).loss)
    assert a == b and np.isfinite(a)


# ---- uint8 frames read by the fused kernel itself (mdb_loss_io.x_format = 1) --------------------------------------
def test_n0f8_conversion_identities():
    """What csrc/photo_strip.inl's cvx() does, in numpy, for all 256 values: (2^23 + u) - 2^23 == u, and
    fma(f, r1, f r2) with r1 = fl(1/255), r2 = fl(1/255 - r1) == u / 255 correctly rounded -- whereas the one-multiply
    shortcut u * (1/255f) is NOT (126 of 256 values differ), so that shortcut would not be Float32(::N0f8)."""
    u = np.arange(256, dtype=np.float32)
    exact = I.n0f8_to_f32(np.arange(256, dtype=np.uint8))
    g = (np.uint32(0x4B000000) | np.arange(256, dtype=np.uint32)).view(np.float32)
    assert np.array_equal(g - np.float32(8388608.0), u)
    r1 = np.float32(0.003921568859368562698)
    r2 = np.float32(-2.3191758e-10)
    assert r1 == np.float32(1.0 / 255.0) and r2 == np.float32(1.0 / 255.0 - float(r1))
    t = (u * r2).astype(np.float32)
    fma = (u.astype(np.float64) * np.float64(r1) + t.astype(np.float64)).astype(np.float32)   # exact product, one rounding
    assert np.array_equal(fma, exact)
    assert int(((u * r1) != exact).sum()) == 126


@pytest.mark.gpu
@pytest.mark.parametrize("W,H,N,auto,scales", [(416, 128, 3, False, (0.125, 0.25, 0.5, 1.0)), (100, 36, 2, True, (0.25, 1.0)),
                                               (61, 45, 2, False, (1.0,))])
def test_fused_loss_reads_u8_frames_bit_identically(W, H, N, auto, scales):
    """Strips -> planar uint8 stack -> FusedLoss(x uint8)  ==  strips -> Float32 stack -> FusedLoss(x float32), bit for bit:
    loss, every disparity gradient, every pose gradient."""
    import mdb200
    inp = synth.make_loss_inputs(W, H, N, seed=17, automasking=auto, scales=scales)
    rng = np.random.default_rng(5)
    strips = torch.tensor(synth.to_u8_strips(inp.x), device=DEV)
    flips = torch.tensor((rng.random(N) < 0.5).astype(np.uint8), device=DEV)
    xf = mdb200.ingest_triplet_u8(strips, W, flips)
    xu = mdb200.ingest_triplet_u8_planar(strips, W, flips)
    assert xu.dtype == torch.uint8 and tuple(xu.shape) == tuple(xf.shape)
    assert np.array_equal(I.n0f8_to_f32(xu.cpu().numpy()), xf.cpu().numpy())
    sizes = [(d.shape[-1], d.shape[-2]) for d in inp.disparities]
    kw = dict(W=W, H=H, C=3, L=3, N=N, disp_sizes=sizes, scales=inp.scales, target_id=inp.target_id, source_ids=inp.source_ids,
              K=inp.K, invK=inp.invK, automasking=auto)
    poses = [mdb200.Pose(torch.tensor(r, device=DEV), torch.tensor(t, device=DEV)) for r, t in zip(inp.rvecs, inp.tvecs)]
    disps = [torch.tensor(d, device=DEV) for d in inp.disparities]
    al = mdb200.automasking_loss(mdb200.SSIM(), xf, xf[:, inp.target_id - 1], source_ids=inp.source_ids) if auto else None
    pf, pu = mdb200.FusedLoss(**kw), mdb200.FusedLoss(**kw)
    a, b = pf(xf, disps, poses, al), pu(xu, disps, poses, al)
    torch.cuda.synchronize()
    assert float(a.loss) == float(b.loss) and np.isfinite(float(a.loss))
    for u, v in zip(a.d_disparities, b.d_disparities):
        assert torch.equal(u, v)
    for u, v in zip(a.d_poses, b.d_poses):
        assert torch.equal(u.rvec, v.rvec) and torch.equal(u.tvec, v.tvec)
    # the same plan serves both element types from one call to the next
    c = pu(xf, disps, poses, al)
    torch.cuda.synchronize()
    assert float(c.loss) == float(a.loss)


@pytest.mark.gpu
def test_u8_frames_rejected_where_not_served():
    import mdb200
    inp = synth.make_loss_inputs(64, 32, 1, seed=3)
    xu = torch.zeros((1, 3, 3, 32, 64), dtype=torch.uint8, device=DEV)
    sizes = [(d.shape[-1], d.shape[-2]) for d in inp.disparities]
    plan = mdb200.FusedLoss(W=64, H=32, C=3, L=3, N=1, disp_sizes=sizes, scales=inp.scales, target_id=2, source_ids=[1, 3],
                            K=inp.K, invK=inp.invK, upsample_align_corners=False)
    poses = [mdb200.Pose(torch.tensor(r, device=DEV), torch.tensor(t, device=DEV)) for r, t in zip(inp.rvecs, inp.tvecs)]
    from mdb200 import _lib
    with pytest.raises(_lib.MdbError, match="x_format = 1"):
        plan(xu, [torch.tensor(d, device=DEV) for d in inp.disparities], poses)
