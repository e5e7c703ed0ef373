"""mdb_smooth_loss_noalloc -- the entry point that allocates nothing (its scratch is a static __device__ array inside
the library, no counterpart in the reference) -- against the existing
entry mdb_smooth_loss_ex with a caller-provided workspace, through the C ABI, on two calls in a row.

Collected last on purpose.  Passes alone in a fresh process and after the other tests:
it relies on no state but the library's own, and the static array needs no clearing between calls."""
import ctypes as ct

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LOSS_RTOL, GRAD_RTOL = 1e-5, 1e-4       # BASELINE.json's gates, as tests/parity.py uses them


def _call_pair(lib, h, disp, img, mode):
    """(loss, grad) from the existing entry (own workspace) and from the no-alloc entry, same inputs."""
    N, C, H, W = img.shape
    stream = ct.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: ct.c_void_p(t.data_ptr())
    wsb = lib.mdb_smooth_loss_workspace_bytes(W, H, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
    l_ref, g_ref = torch.full((1,), -7.0, device=DEV), torch.full_like(disp, -7.0)
    l_new, g_new = torch.full((1,), -9.0, device=DEV), torch.full_like(disp, -9.0)
    assert lib.mdb_smooth_loss_ex(h, p(disp), p(img), W, H, C, N, mode, p(l_ref), p(g_ref), p(ws), wsb, stream) == 0
    rc = lib.mdb_smooth_loss_noalloc(h, p(disp), p(img), W, H, C, N, mode, p(l_new), p(g_new), stream)
    assert rc == 0, lib.mdb_last_error(h)
    torch.cuda.synchronize()
    return (float(l_ref), g_ref), (float(l_new), g_new)


def _check(ref, new):
    (l_ref, g_ref), (l_new, g_new) = ref, new
    assert abs(l_new - l_ref) <= LOSS_RTOL * abs(l_ref), (l_new, l_ref)
    scale = float(g_ref.abs().max())
    assert float((g_new - g_ref).abs().max()) <= GRAD_RTOL * scale
    assert l_ref > 0 and scale > 0          # the comparison is not between two untouched buffers


def test_noalloc_matches_workspace_entry_on_two_calls_in_a_row():
    import mdb200
    from mdb200 import _lib
    lib, h = _lib.lib, mdb200.handle()
    g = torch.Generator().manual_seed(23)
    # call 1: the vectorised kernel (W % 4 == 0), 12 x 416x128 like config 2
    d1 = torch.rand((12, 128, 416), generator=g).to(DEV)
    i1 = torch.rand((12, 3, 128, 416), generator=g).to(DEV)
    _check(*_call_pair(lib, h, d1, i1, 0))
    # call 2, straight after, another shape (fewer partials than call 1 left behind, scalar kernel: W % 4 != 0),
    # plateaus in the disparity and the other abs convention: stale scratch contents must not leak in
    d2 = torch.rand((3, 37, 101), generator=g)
    d2[:, :10, :40] = 0.5
    i2 = torch.rand((3, 1, 37, 101), generator=g)
    ref, new = _call_pair(lib, h, d2.to(DEV), i2.to(DEV), 1)
    _check(ref, new)
    # and against the float64 oracle, so that "equal" is not "equally wrong"
    from oracle import torch_ref as T
    T.ABS_SUBGRADIENT = 1
    try:
        d64 = d2.double().requires_grad_()
        l64 = T.smooth_loss(d64, i2.double())
        (g64,) = torch.autograd.grad(l64, d64)
    finally:
        T.ABS_SUBGRADIENT = 0
    assert abs(new[0] - float(l64)) <= LOSS_RTOL * abs(float(l64))
    assert float((new[1].double().cpu() - g64).abs().max()) <= GRAD_RTOL * float(g64.abs().max())
    # call 1's inputs once more: same answer as the first time, bit for bit (deterministic, nothing carried over)
    a = _call_pair(lib, h, d1, i1, 0)[1]
    b = _call_pair(lib, h, d1, i1, 0)[1]
    assert a[0] == b[0] and torch.equal(a[1], b[1])


def test_noalloc_rejects_shapes_beyond_the_static_scratch():
    import mdb200
    from mdb200 import _lib
    lib, h = _lib.lib, mdb200.handle()
    # N * ceil(H / 8) = 4097 * 8 = 32776 > 32768: one row block too many.  The inputs are never touched (the call is refused
    # before any launch), so small dummies stand in for them.
    dummy = torch.zeros(16, device=DEV)
    out = torch.full((1,), 3.0, device=DEV)
    p = lambda t: ct.c_void_p(t.data_ptr())
    rc = lib.mdb_smooth_loss_noalloc(h, p(dummy), p(dummy), 8, 64, 1, 4097, 0, p(out), None, None)
    assert rc == -3                                           # MDB_E_WORKSPACE
    msg = lib.mdb_last_error(h).decode()
    assert "exceeds the built-in scratch" in msg and "mdb_smooth_loss_ex" in msg, msg
    torch.cuda.synchronize()
    assert float(out) == 3.0
    # the largest shape that fits is accepted: N * ceil(H / 8) = 32768
    N, H, W = 4096, 64, 8
    d = torch.rand((N, H, W), device=DEV)
    im = torch.rand((N, 1, H, W), device=DEV)
    _check(*_call_pair(lib, h, d, im, 0))
