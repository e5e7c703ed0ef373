"""Host-side mirror of the reference's Julia API for the view-synthesis loss path.

Same names, argument order and meaning as the reference functions (file:line cited per
function), implemented as thin calls into libmdb200.so through the C ABI.  Julia is not
installed on the build or GPU boxes, so this Python layer plays the role of the Julia glue
(monodepth2.jl_b200/julia/); `torch` is used only for device memory and
streams.

Array convention: a Julia array of size (d1,...,dk) is a contiguous CUDA float32 tensor of
shape (dk,...,d1) -- identical bytes, no transposes anywhere.  Images (W,H,C,N) are
(N,C,H,W); rvec (3,N) is (N,3); tvec (3,1,N) is (N,1,3); rotation stacks (3,3,N) are
(N,3,3) indexed [n, j, i]; K (3,3) is K_math.T.
"""
from __future__ import annotations

import ctypes as ct
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import lib, check

_handles = {}


def handle(device: Optional[int] = None):
    """One mdb_handle per CUDA device (created on first use; fails without a B200)."""
    if not torch.cuda.is_available():
        raise RuntimeError("mdb200 needs a CUDA device (sm_100); there is no CPU fallback")
    dev = torch.cuda.current_device() if device is None else device
    if dev not in _handles:
        h = ct.c_void_p()
        rc = lib.mdb_create(ct.byref(h), dev)
        if rc != 0:
            raise _lib.MdbError(rc, "mdb_create failed (no sm_100 device?)")
        _handles[dev] = h
    return _handles[dev]


def launch_count() -> int:
    return int(lib.mdb_launch_count(handle()))


def profile_begin(slots: int) -> None:
    """Record per-segment CUDA events inside the next `slots` fused-loss calls (0 = off)."""
    check(handle(), lib.mdb_profile_begin(handle(), int(slots)))


def measure_fp32_peak(repeats: int = 5):
    """-> (TFLOP/s, ms): the FP32 FMA peak of the current device, measured (bench roofline denominator)."""
    h = handle()
    scratch = torch.empty(148 * 4 * 8 * 256, dtype=torch.float32, device="cuda")
    tf, ms = ct.c_double(0.0), ct.c_double(0.0)
    check(h, lib.mdb_measure_fp32_peak(h, _p(scratch), scratch.numel() * 4, int(repeats), ct.byref(tf), ct.byref(ms), _stream()))
    return tf.value, ms.value


def debug_option(name: str, value: int) -> None:
    """Test hooks of the library ("strip_segs", "upsample_split"); 0 restores the default."""
    rc = lib.mdb_debug_option(name.encode(), int(value))
    if rc != 0:
        raise ValueError(f"unknown debug option {name!r}")


def profile_read(slot: int):
    """-> [prep+stats, fused kernel, finalize, upsample transpose] durations of that call, ms."""
    out = (ct.c_float * 4)()
    check(handle(), lib.mdb_profile_read(handle(), int(slot), out))
    return list(out)


def _stream():
    return ct.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]):
    return ct.c_void_p(t.data_ptr()) if t is not None else ct.c_void_p(0)


def _chk(t: torch.Tensor, name: str) -> torch.Tensor:
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
        raise TypeError(f"{name}: expected a contiguous CUDA float32 tensor "
                        f"(got {type(t).__name__}, {getattr(t, 'dtype', None)}, "
                        f"cuda={getattr(t, 'is_cuda', None)})")
    return t


def _k9(K) -> "ct.Array":
    """K / invK as 9 host floats in Julia (column-major) order."""
    k = torch.as_tensor(K, dtype=torch.float32, device="cpu").reshape(9)
    return (ct.c_float * 9)(*k.tolist())


# --------------------------------------------------------------------------------------
# types -- src/pose_decoder.jl:1-5, src/Monodepth.jl:37-60
# --------------------------------------------------------------------------------------
@dataclass
class Pose:
    rvec: torch.Tensor  # (N,3)     Julia (3,N)
    tvec: torch.Tensor  # (N,1,3)   Julia (3,1,N)   [render: (B,3) Julia (3,B)]


@dataclass
class Params:
    min_depth: float = 0.1
    max_depth: float = 100.0
    disparity_smoothness: float = 1e-3
    frame_ids: List[int] = field(default_factory=lambda: [1, 2, 3])
    automasking: bool = True
    target_size: Tuple[int, int] = (416, 128)  # (width, height)
    batch_size: int = 1


class SSIM:
    """`SSIM()` -- src/utils.jl:17-43 (3x3 mean pool over a reflect-padded image)."""

    def __init__(self, c1: float = 0.01 ** 2, c2: float = 0.03 ** 2):
        self.c1, self.c2 = c1, c2

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Differentiable w.r.t. x (mdb_ssim_bwd), like the reference's SSIM under Zygote."""
        return _SsimFn.apply(x, y, self.c1, self.c2)


def _ssim_ws(W, H, Cc, N, dev):
    n = int(lib.mdb_ssim_bwd_workspace_bytes(W, H, Cc, N))
    return torch.empty(n, dtype=torch.uint8, device=dev), n


class _SsimFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, c1, c2):
        x, y = x.detach(), y.detach()
        _chk(x, "x"); _chk(y, "y")
        N, Cc, H, W = x.shape
        out = torch.empty_like(x)
        check(handle(), lib.mdb_ssim(handle(), _p(x), _p(y), W, H, Cc, N, c1, c2, _p(out), _stream()))
        ctx.save_for_backward(x, y)
        ctx.c = (c1, c2)
        return out

    @staticmethod
    def backward(ctx, g):
        x, y = ctx.saved_tensors
        N, Cc, H, W = x.shape
        ws, n = _ssim_ws(W, H, Cc, N, x.device)
        dx = torch.empty_like(x)
        check(handle(), lib.mdb_ssim_bwd(handle(), _p(x), _p(y), _p(g.contiguous()), W, H, Cc, N, ctx.c[0], ctx.c[1], _p(dx),
                                         _p(ws), n, _stream()))
        return dx, None, None, None


class Backproject:
    """`Backproject(; width, height)` -- src/utils.jl:45-69."""

    def __init__(self, *, width: int, height: int):
        self.width, self.height = width, height

    def __call__(self, depth: torch.Tensor, invK) -> torch.Tensor:
        """depth (N,WH,1) [Julia (1,WH,N)] -> points (N,WH,3) [Julia (3,WH,N)]."""
        _chk(depth, "depth")
        N = depth.shape[0]
        pts = torch.empty((N, self.width * self.height, 3), device=depth.device, dtype=torch.float32)
        check(handle(), lib.mdb_backproject(handle(), _p(depth), _k9(invK), self.width, self.height, N, _p(pts), _stream()))
        return pts


class Project:
    """`Project(; width, height)` -- src/utils.jl:71-103."""

    def __init__(self, *, width: int, height: int):
        self.width, self.height = width, height

    def __call__(self, points: torch.Tensor, K, R: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
        """points (N,WH,3), R (N,3,3), t (N,1,3) -> normalised uv (N,WH,2) [Julia (2,WH,N)]."""
        _chk(points, "points"); _chk(R, "R"); _chk(t, "t")
        N = points.shape[0]
        uv = torch.empty((N, self.width * self.height, 2), device=points.device, dtype=torch.float32)
        check(handle(), lib.mdb_project(handle(), _p(points), _k9(K), _p(R), _p(t), self.width, self.height, N, _p(uv), _stream()))
        return uv


@dataclass
class TrainCache:
    """`TrainCache` -- src/Monodepth.jl:49-60."""
    ssim: SSIM
    backprojections: Backproject
    projections: Project
    K: torch.Tensor      # (3,3) Julia layout, host or device
    invK: torch.Tensor
    target_id: int
    source_ids: List[int]
    scales: List[float]


# --------------------------------------------------------------------------------------
# so3 / pose -- src/utils.jl:106-145, 185-192
# --------------------------------------------------------------------------------------
def hat(rvec: torch.Tensor) -> torch.Tensor:
    _chk(rvec, "rvec")
    N = rvec.shape[0]
    S = torch.empty((N, 3, 3), device=rvec.device, dtype=torch.float32)
    check(handle(), lib.mdb_hat(handle(), _p(rvec), N, _p(S), _stream()))
    return S


def hat_pullback(d: torch.Tensor) -> torch.Tensor:
    _chk(d, "d")
    N = d.shape[0]
    g = torch.empty((N, 3), device=d.device, dtype=torch.float32)
    check(handle(), lib.mdb_hat_pullback(handle(), _p(d), N, _p(g), _stream()))
    return g


def so3_exp_map(rvec: torch.Tensor) -> torch.Tensor:
    _chk(rvec, "rvec")
    N = rvec.shape[0]
    R = torch.empty((N, 3, 3), device=rvec.device, dtype=torch.float32)
    check(handle(), lib.mdb_so3_exp_map(handle(), _p(rvec), N, _p(R), _stream()))
    return R


class _ComposeTFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rvec, t, invert):
        rvec, t = rvec.detach(), t.detach()
        _chk(rvec, "rvec"); _chk(t, "t")
        N = rvec.shape[0]
        R = torch.empty((N, 3, 3), device=rvec.device, dtype=torch.float32)
        to = torch.empty((N, 1, 3), device=rvec.device, dtype=torch.float32)
        check(handle(), lib.mdb_compose_T(handle(), _p(rvec), _p(t), int(bool(invert)), N, _p(R), _p(to), _stream()))
        ctx.save_for_backward(rvec, t)
        ctx.invert = bool(invert)
        return R, to

    @staticmethod
    def backward(ctx, dR, dt):
        rvec, t = ctx.saved_tensors
        dR = torch.zeros((rvec.shape[0], 3, 3), device=rvec.device) if dR is None else dR.contiguous()
        dt = torch.zeros_like(t) if dt is None else dt.contiguous()
        dr, dtv = composeT_pullback(rvec, t, ctx.invert, dR, dt)
        return dr, dtv, None


def composeT(rvec: torch.Tensor, t: torch.Tensor, invert: bool):
    """src/utils.jl:185-192 (differentiable: mdb_compose_T_bwd)."""
    return _ComposeTFn.apply(rvec, t, invert)


def composeT_pullback(rvec, t, invert, dR, dt):
    for a, nm in ((rvec, "rvec"), (t, "t"), (dR, "dR"), (dt, "dt")):
        _chk(a, nm)
    N = rvec.shape[0]
    dr = torch.empty_like(rvec)
    dtv = torch.empty_like(t)
    check(handle(), lib.mdb_compose_T_bwd(handle(), _p(rvec), _p(t), int(bool(invert)), N, _p(dR), _p(dt), _p(dr), _p(dtv), _stream()))
    return dr, dtv


# --------------------------------------------------------------------------------------
# elementwise / sampling primitives
# --------------------------------------------------------------------------------------
def disparity_to_depth(disparity: torch.Tensor, min_depth: float, max_depth: float) -> torch.Tensor:
    """src/utils.jl:179-183."""
    _chk(disparity, "disparity")
    out = torch.empty_like(disparity)
    check(handle(), lib.mdb_disparity_to_depth(handle(), _p(disparity), disparity.numel(), float(min_depth), float(max_depth), _p(out), _stream()))
    return out


def grid_sample(inp: torch.Tensor, grid: torch.Tensor, padding_mode: str = "border") -> torch.Tensor:
    """NNlib `grid_sample(input, grid; padding_mode=:border)`; inp (N,C,H,W), grid (N,Ho,Wo,2)."""
    if padding_mode != "border":
        raise ValueError("only padding_mode=:border is on the reference's path")
    _chk(inp, "input"); _chk(grid, "grid")
    N, Cc, H, W = inp.shape
    Ho, Wo = grid.shape[1:3]
    out = torch.empty((N, Cc, Ho, Wo), device=inp.device, dtype=torch.float32)
    check(handle(), lib.mdb_grid_sample(handle(), _p(inp), _p(grid), W, H, Cc, N, Wo, Ho, _p(out), _stream()))
    return out


class _WarpFn(torch.autograd.Function):
    """warp forward (mdb_warp) and pullback (mdb_warp_bwd) w.r.t. disp and every (R, t)."""

    @staticmethod
    def forward(ctx, disp, x, Ks, invKs, min_depth, max_depth, source_ids, *Rt):
        disp, x = disp.detach(), x.detach()
        Rt = [t.detach() for t in Rt]
        _chk(disp, "disp"); _chk(x, "x")
        N, L, Cc, H, W = x.shape
        S = len(Rt) // 2
        outs = [torch.empty((N, Cc, H, W), device=x.device, dtype=torch.float32) for _ in range(S)]
        Rp = (ct.c_void_p * S)(*[_chk(Rt[2 * s], "R").data_ptr() for s in range(S)])
        tp = (ct.c_void_p * S)(*[_chk(Rt[2 * s + 1], "t").data_ptr() for s in range(S)])
        op = (ct.c_void_p * S)(*[o.data_ptr() for o in outs])
        ids = (ct.c_int32 * S)(*[int(s) for s in source_ids])
        check(handle(), lib.mdb_warp(handle(), _p(disp), _p(x), Rp, tp, ids, S, _k9(Ks), _k9(invKs), W, H, Cc, L, N,
                                     float(min_depth), float(max_depth), op, _stream()))
        ctx.save_for_backward(disp, x, *Rt)
        ctx.meta = (Ks, invKs, float(min_depth), float(max_depth), list(source_ids))
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gs):
        disp, x, *Rt = ctx.saved_tensors
        Ks, invKs, mind, maxd, source_ids = ctx.meta
        N, L, Cc, H, W = x.shape
        S = len(Rt) // 2
        gs = [torch.zeros((N, Cc, H, W), device=x.device) if g is None else g.contiguous() for g in gs]
        d_disp = torch.empty_like(disp)
        dR = [torch.empty_like(Rt[2 * s]) for s in range(S)]
        dt = [torch.empty_like(Rt[2 * s + 1]) for s in range(S)]
        arr = lambda ts: (ct.c_void_p * S)(*[t.data_ptr() for t in ts])
        ids = (ct.c_int32 * S)(*[int(s) for s in source_ids])
        check(handle(), lib.mdb_warp_bwd(handle(), _p(disp), _p(x), arr([Rt[2 * s] for s in range(S)]),
                                         arr([Rt[2 * s + 1] for s in range(S)]), ids, S, _k9(Ks), _k9(invKs), W, H, Cc, L, N,
                                         mind, maxd, arr(gs), _p(d_disp), arr(dR), arr(dt), _stream()))
        flat = [t for s in range(S) for t in (dR[s], dt[s])]
        return (d_disp, None, None, None, None, None, None, *flat)


def warp(disp: torch.Tensor, x: torch.Tensor, Ps, backprojections, projections, invKs, Ks, *,
         min_depth: float, max_depth: float, source_ids: Sequence[int]) -> List[torch.Tensor]:
    """`warp(...)` as called at src/simple_depth.jl:30-32 (body: src/training.jl:48-57).
    disp (N,H,W,1) or (N,1,H,W); x (N,L,C,H,W); Ps = [(R (N,3,3), t (N,1,3))].  Differentiable w.r.t. disp and the
    (R, t) pairs (mdb_warp_bwd), so the closure of `slow_depth` can be differentiated function by function."""
    flat = [t for P in Ps for t in (P[0], P[1])]
    return list(_WarpFn.apply(disp, x, Ks, invKs, min_depth, max_depth, tuple(source_ids), *flat))


def photometric_loss(ssim: SSIM, predicted: torch.Tensor, target: torch.Tensor, alpha: float = 0.85):
    """src/training.jl:1-7 -> (N,1,H,W)."""
    _chk(predicted, "predicted"); _chk(target, "target")
    N, Cc, H, W = predicted.shape
    out = torch.empty((N, 1, H, W), device=predicted.device, dtype=torch.float32)
    check(handle(), lib.mdb_photometric_loss(handle(), _p(predicted), _p(target), W, H, Cc, N, alpha, ssim.c1, ssim.c2, _p(out), _stream()))
    return out


class _PredLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, target, alpha, c1, c2, abs_subgradient, *preds):
        target = target.detach()
        preds = [p.detach() for p in preds]
        _chk(target, "target")
        N, Cc, H, W = target.shape
        S = len(preds)
        pp = (ct.c_void_p * S)(*[_chk(p, "prediction").data_ptr() for p in preds])
        out = torch.empty((N, 1, H, W), device=target.device, dtype=torch.float32)
        check(handle(), lib.mdb_prediction_loss(handle(), pp, S, _p(target), W, H, Cc, N, alpha, c1, c2, _p(out), _stream()))
        ctx.save_for_backward(target, *preds)
        ctx.meta = (alpha, c1, c2, int(abs_subgradient))
        return out

    @staticmethod
    def backward(ctx, g):
        target, *preds = ctx.saved_tensors
        alpha, c1, c2, mode = ctx.meta
        N, Cc, H, W = target.shape
        S = len(preds)
        ws, n = _ssim_ws(W, H, Cc, N, target.device)
        d = [torch.empty_like(p) for p in preds]
        arr = lambda ts: (ct.c_void_p * S)(*[t.data_ptr() for t in ts])
        check(handle(), lib.mdb_prediction_loss_bwd(handle(), arr(preds), S, _p(target), _p(g.contiguous()), W, H, Cc, N, alpha, c1, c2,
                                                    mode, arr(d), _p(ws), n, _stream()))
        return (None, None, None, None, None, *d)


def prediction_loss(ssim: SSIM, predictions: Sequence[torch.Tensor], target: torch.Tensor, alpha: float = 0.85,
                    abs_subgradient: int = 0):
    """src/training.jl:13-15.  Differentiable w.r.t. the predictions (mdb_prediction_loss_bwd: first arg-min)."""
    return _PredLossFn.apply(target, alpha, ssim.c1, ssim.c2, abs_subgradient, *predictions)


def automasking_loss(ssim: SSIM, inputs: torch.Tensor, target: torch.Tensor, *, source_ids: Sequence[int], alpha: float = 0.85):
    """`automasking_loss(ssim, inputs, target; source_ids)` -- src/training.jl:9-11, the reference's own
    signature: `inputs` (N,L,C,H,W), `target` (N,C,H,W) the target ARRAY (callers pass
    `x[:, :, :, target_id, :]`, src/dtk.jl:59-61).  A dense tensor or the view `inputs[:, target_id - 1]`
    (frames dense, items strided: no copy) are both taken as they are."""
    _chk(inputs, "inputs")
    N, L, Cc, H, W = inputs.shape
    if not (isinstance(target, torch.Tensor) and target.is_cuda and target.dtype == torch.float32
            and tuple(target.shape) == (N, Cc, H, W) and target[0].is_contiguous() and (N == 1 or target.stride(0) >= Cc * H * W)):
        raise TypeError("target: expected a CUDA float32 (N,C,H,W) tensor whose items are dense (a contiguous tensor or "
                        "inputs[:, target_id - 1])")
    S = len(source_ids)
    ids = (ct.c_int32 * S)(*[int(s) for s in source_ids])
    out = torch.empty((N, 1, H, W), device=inputs.device, dtype=torch.float32)
    check(handle(), lib.mdb_automasking_loss(handle(), _p(inputs), _p(target), int(target.stride(0)) if N > 1 else 0, ids, S,
                                             W, H, Cc, L, N, alpha, ssim.c1, ssim.c2, _p(out), _stream()))
    return out


class _ApplyMaskFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mask, warp_loss):
        mask, warp_loss = mask.detach(), warp_loss.detach()
        _chk(mask, "mask"); _chk(warp_loss, "warp_loss")
        out = torch.empty_like(warp_loss)
        check(handle(), lib.mdb_apply_mask(handle(), _p(mask), _p(warp_loss), warp_loss.numel(), _p(out), _stream()))
        ctx.save_for_backward(mask, warp_loss)
        return out

    @staticmethod
    def backward(ctx, g):
        mask, wl = ctx.saved_tensors
        dm, dw = torch.empty_like(mask), torch.empty_like(wl)
        check(handle(), lib.mdb_apply_mask_bwd(handle(), _p(mask), _p(wl), _p(g.contiguous()), wl.numel(), _p(dm), _p(dw), _stream()))
        return dm, dw


def _apply_mask(mask: torch.Tensor, warp_loss: torch.Tensor) -> torch.Tensor:
    """src/training.jl:17-19 (differentiable: the mask keeps ties)."""
    return _ApplyMaskFn.apply(mask, warp_loss)


def _smooth_raw(disparity, image, with_grad, abs_subgradient):
    _chk(disparity, "disparity"); _chk(image, "image")
    N, Cc, H, W = image.shape
    wsb = lib.mdb_smooth_loss_workspace_bytes(W, H, N)
    ws = torch.empty(wsb, dtype=torch.uint8, device=image.device)
    loss = torch.empty((), dtype=torch.float32, device=image.device)
    grad = torch.empty_like(disparity) if with_grad else None
    check(handle(), lib.mdb_smooth_loss_ex(handle(), _p(disparity), _p(image), W, H, Cc, N, int(abs_subgradient), _p(loss), _p(grad),
                                           _p(ws), wsb, _stream()))
    return loss, grad


class _SmoothFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, disparity, image, abs_subgradient):
        loss, grad = _smooth_raw(disparity.detach(), image.detach(), True, abs_subgradient)
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None, None


def smooth_loss(disparity: torch.Tensor, image: torch.Tensor, with_grad: bool = False, abs_subgradient: int = 0):
    """src/utils.jl:163-177.  disparity (N,H,W), image (N,C,H,W) -> 0-dim loss tensor
    (and d loss / d disparity when `with_grad`; `abs_subgradient`: see mdb_loss_params).  When `disparity` requires
    grad the returned loss is differentiable (value and gradient come from the same kernel launch)."""
    if disparity.requires_grad and not with_grad:
        return _SmoothFn.apply(disparity, image, abs_subgradient)
    loss, grad = _smooth_raw(disparity.detach(), image, with_grad, abs_subgradient)
    return (loss, grad) if with_grad else loss


# --------------------------------------------------------------------------------------
# the fused hot path: train_loss forward + backward in one pass
# --------------------------------------------------------------------------------------
@dataclass
class LossResult:
    loss: torch.Tensor                 # 0-dim, loss / Ls over the local items
    d_disparities: List[torch.Tensor]  # native resolution, same shapes as the inputs
    d_poses: List[Pose]                # (d_rvec (N,3), d_tvec (N,1,3)) per source
    vis_warp_loss: Optional[torch.Tensor] = None


class FusedLoss:
    """Pre-planned call of `mdb_train_loss_fwdbwd` for fixed shapes.  Owns the workspace and
    the gradient buffers so the per-step path allocates nothing (the caller-owns-everything
    rule of the ABI; the Julia glue does the same in its `rrule`)."""

    def __init__(self, *, W: int, H: int, C: int, L: int, N: int, disp_sizes: Sequence[Tuple[int, int]],
                 scales: Sequence[float], target_id: int, source_ids: Sequence[int], K, invK,
                 min_depth: float = 0.1, max_depth: float = 100.0, disparity_smoothness: float = 1e-3,
                 automasking: bool = False, normalize_disparity: bool = True, smooth_weight_mode: int = 0,
                 n_global: int = 0, alpha: float = 0.85, want_vis: bool = False, device=None,
                 out: Optional[dict] = None, abs_subgradient: int = 0, ssim: Optional["SSIM"] = None,
                 upsample_align_corners: bool = True):
        """`out` (optional): caller-provided output tensors {"loss": 0-dim or 1-element, "d_disp": [...],
        "d_rvec": [...], "d_tvec": [...]} of the right shapes -- e.g. row slices of whole-batch tensors --
        instead of buffers owned by the plan."""
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.shape = (N, L, C, H, W)
        self.disp_sizes = [(int(w), int(h)) for (w, h) in disp_sizes]
        self.S = len(source_ids)
        self.Ls = len(scales)
        p = _lib.LossParams()
        lib.mdb_loss_params_default(ct.byref(p))
        p.min_depth, p.max_depth, p.disparity_smoothness, p.alpha = min_depth, max_depth, disparity_smoothness, alpha
        p.automasking = int(automasking)
        p.normalize_disparity = int(normalize_disparity)
        p.smooth_weight_mode = int(smooth_weight_mode)
        p.n_global = int(n_global)
        p.abs_subgradient = int(abs_subgradient)
        p.upsample_align_corners = 1 if upsample_align_corners else 0
        if ssim is not None:
            p.ssim_c1, p.ssim_c2 = float(ssim.c1), float(ssim.c2)
        self.params = p
        io = _lib.LossIO()
        io.W, io.H, io.C, io.L, io.N = W, H, C, L, N
        io.target_id = int(target_id)
        io.S = self.S
        for s, sid in enumerate(source_ids):
            io.source_ids[s] = int(sid)
            io.invert[s] = int(sid < target_id)  # src/training.jl:28
        io.Ls = self.Ls
        for k, ((w, h), sc) in enumerate(zip(self.disp_sizes, scales)):
            io.dw[k], io.dh[k], io.scale[k] = w, h, float(sc)
        kk, ik = _k9(K), _k9(invK)
        for i in range(9):
            io.K[i], io.invK[i] = kk[i], ik[i]
        self.io = io
        f32 = dict(device=self.dev, dtype=torch.float32)
        out = dict(out or {})   # any subset of the outputs may be caller-provided; the plan owns the rest
        self.loss = out["loss"] if "loss" in out else torch.zeros((), **f32)
        self.d_disp = list(out["d_disp"]) if "d_disp" in out else [torch.empty((N, 1, h, w), **f32) for (w, h) in self.disp_sizes]
        self.d_rvec = list(out["d_rvec"]) if "d_rvec" in out else [torch.empty((N, 3), **f32) for _ in range(self.S)]
        self.d_tvec = list(out["d_tvec"]) if "d_tvec" in out else [torch.empty((N, 1, 3), **f32) for _ in range(self.S)]
        for t, shp in [(self.loss, None)] + [(t, (N, 1, h, w)) for t, (w, h) in zip(self.d_disp, self.disp_sizes)] + \
                [(t, (N, 3)) for t in self.d_rvec] + [(t, (N, 1, 3)) for t in self.d_tvec]:
            _chk(t, "out")
            if shp is not None and tuple(t.shape) != shp:
                raise ValueError(f"output buffer has shape {tuple(t.shape)}, expected {shp}")
        if self.loss.numel() != 1:
            raise ValueError("out['loss'] must hold one float")
        self.vis = torch.empty((N, 1, H, W), **f32) if want_vis else None
        io.loss = self.loss.data_ptr()
        for k in range(self.Ls):
            io.d_disparity[k] = self.d_disp[k].data_ptr()
        for s in range(self.S):
            io.d_rvec[s] = self.d_rvec[s].data_ptr()
            io.d_tvec[s] = self.d_tvec[s].data_ptr()
        io.vis_warp_loss = self.vis.data_ptr() if want_vis else None
        self.ws_bytes = int(lib.mdb_train_loss_workspace_bytes(ct.byref(io)))
        if self.ws_bytes == 0:
            raise _lib.MdbError(_lib.lib.mdb_version() * 0 - 2, "unsupported shape for the fused loss")
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=self.dev)

    def strip_geometry(self) -> dict:
        """Launch decomposition of the fused kernel for this plan (mdb_train_loss_strip_geometry)."""
        g = (ct.c_int32 * 4)()
        lib.mdb_train_loss_strip_geometry(ct.byref(self.io), g)
        return dict(strips=g[0], segments=g[1], rows_per_segment=g[2], warps=g[3])

    def __call__(self, x: torch.Tensor, disparities: Sequence[torch.Tensor], poses: Sequence[Pose],
                 auto_loss: Optional[torch.Tensor] = None) -> LossResult:
        io = self.io
        if x.dtype == torch.uint8:      # N0f8 frames, converted in the kernel (mdb_loss_io.x_format = 1)
            if not x.is_cuda or not x.is_contiguous():
                raise ValueError("x must be a contiguous CUDA tensor")
            io.x_format = 1
        else:
            _chk(x, "x")
            io.x_format = 0
        if tuple(x.shape) != self.shape:
            raise ValueError(f"x has shape {tuple(x.shape)}, planned for {self.shape}")
        io.x = x.data_ptr()
        for k, d in enumerate(disparities):
            _chk(d, f"disparities[{k}]")
            if d.shape[-1] != self.disp_sizes[k][0] or d.shape[-2] != self.disp_sizes[k][1]:
                raise ValueError(f"disparities[{k}] has shape {tuple(d.shape)}, planned {self.disp_sizes[k]}")
            io.disparity[k] = d.data_ptr()
        for s, pz in enumerate(poses):
            io.rvec[s] = _chk(pz.rvec, "rvec").data_ptr()
            io.tvec[s] = _chk(pz.tvec, "tvec").data_ptr()
        io.auto_loss = _chk(auto_loss, "auto_loss").data_ptr() if auto_loss is not None else None
        # the plan's device, not whatever device happens to be current: handle, stream and launches all belong to it
        with torch.cuda.device(self.dev):
            h = handle(self.dev.index)
            check(h, lib.mdb_train_loss_fwdbwd(h, ct.byref(io), ct.byref(self.params), _p(self.ws), self.ws_bytes, _stream()))
        return LossResult(self.loss, self.d_disp, [Pose(r, t) for r, t in zip(self.d_rvec, self.d_tvec)], self.vis)


class _TrainLossFn(torch.autograd.Function):
    """The Python twin of the Julia `rrule`: forward runs the fused fwd+bwd kernels with unit
    seed and caches the gradients; backward scales them by the incoming scalar cotangent."""

    @staticmethod
    def forward(ctx, plan: FusedLoss, x, auto_loss, n_disp, *tensors):
        disps = tensors[:n_disp]
        rest = tensors[n_disp:]
        S = len(rest) // 2
        poses = [Pose(rest[2 * s].detach(), rest[2 * s + 1].detach()) for s in range(S)]
        res = plan(x, [d.detach() for d in disps], poses, auto_loss)
        ctx.saved = [g.clone() for g in res.d_disparities] + \
            [t.clone() for pz in res.d_poses for t in (pz.rvec, pz.tvec)]
        return res.loss.clone()

    @staticmethod
    def backward(ctx, delta):
        return (None, None, None, None) + tuple(g * delta for g in ctx.saved)


def train_loss(model: Callable, x: torch.Tensor, auto_loss: Optional[torch.Tensor], cache: TrainCache,
               parameters: Params, do_visualization: bool = False, _plans={}):
    """`train_loss(model, x, auto_loss, cache, parameters, do_visualization)` --
    src/training.jl:21-78.  `model(x, source_ids, target_id)` returns (disparities, poses).
    Returns (loss, vis_disparity, vis_warped, vis_loss) like the reference; the loss is
    differentiable w.r.t. the disparities and poses through a custom autograd Function."""
    _chk(x, "x")
    N, L, Cc, H, W = x.shape
    disparities, poses = model(x, cache.source_ids, cache.target_id)
    sizes = tuple((d.shape[-1], d.shape[-2]) for d in disparities)
    # everything that is baked into the plan at construction is part of the key -- the intrinsics and the SSIM constants
    # too (two datasets with different K in one process must not share a plan)
    k9t = lambda K: tuple(torch.as_tensor(K, dtype=torch.float32, device="cpu").reshape(9).tolist())
    key = (N, L, Cc, H, W, sizes, tuple(cache.scales), cache.target_id, tuple(cache.source_ids),
           parameters.automasking, parameters.min_depth, parameters.max_depth, parameters.disparity_smoothness,
           bool(do_visualization), x.device.index, k9t(cache.K), k9t(cache.invK), float(cache.ssim.c1), float(cache.ssim.c2))
    if key not in _plans:
        _plans[key] = FusedLoss(W=W, H=H, C=Cc, L=L, N=N, disp_sizes=sizes, scales=cache.scales,
                                target_id=cache.target_id, source_ids=cache.source_ids, K=cache.K, invK=cache.invK,
                                min_depth=parameters.min_depth, max_depth=parameters.max_depth,
                                disparity_smoothness=parameters.disparity_smoothness,
                                automasking=parameters.automasking, want_vis=bool(do_visualization), device=x.device,
                                ssim=cache.ssim)
    plan = _plans[key]
    flat = list(disparities) + [t for pz in poses for t in (pz.rvec, pz.tvec)]
    loss = _TrainLossFn.apply(plan, x, auto_loss if parameters.automasking else None, len(disparities), *flat)
    vis_disparity = vis_warped = vis_loss = None
    if do_visualization:
        vis_disparity = disparities[-1].detach().cpu()
        vis_loss = plan.vis.detach().cpu()
        Ps = [composeT(pz.rvec.detach(), pz.tvec.detach(), sid < cache.target_id)
              for pz, sid in zip(poses, cache.source_ids)]
        full = disparities[-1].detach()
        vis_warped = [w.cpu() for w in warp(full, x, Ps, cache.backprojections, cache.projections, cache.invK,
                                            cache.K, min_depth=parameters.min_depth, max_depth=parameters.max_depth,
                                            source_ids=cache.source_ids)]
    return loss, vis_disparity, vis_warped, vis_loss


def simple_depth_plan(x: torch.Tensor, K, invK, *, target_id: int = 2, source_ids: Sequence[int] = (1, 3),
                      min_depth: float = 0.1, max_depth: float = 100.0, abs_subgradient: int = 0) -> FusedLoss:
    """The closure `slow_depth` differentiates (src/simple_depth.jl:25-41): one full-res scale,
    `mean(prediction_loss) + smooth_loss(disp, target)` with un-normalised disparity and
    smoothness weight 1."""
    N, L, Cc, H, W = x.shape
    return FusedLoss(W=W, H=H, C=Cc, L=L, N=N, disp_sizes=[(W, H)], scales=[1.0], target_id=target_id,
                     source_ids=source_ids, K=K, invK=invK, min_depth=min_depth, max_depth=max_depth,
                     automasking=False, normalize_disparity=False, smooth_weight_mode=1, device=x.device,
                     abs_subgradient=abs_subgradient)


# --------------------------------------------------------------------------------------
# the optimiser loop of slow_depth, resident on the device
# --------------------------------------------------------------------------------------
class ADAM:
    """`Flux.Optimise.ADAM(eta, (beta1, beta2))` -- the optimiser of src/simple_depth.jl:16.
    `update(params, grads)` is `Flux.Optimise.update!(opt, params, grads)` in one kernel launch for the
    whole parameter list; the step counter lives on the device."""

    def __init__(self, eta: float = 1e-3, beta=(0.9, 0.999), eps: float = 1e-8):
        self.p = _lib.AdamParams(float(eta), float(beta[0]), float(beta[1]), float(eps))
        self.state = None

    def update(self, params: Sequence[torch.Tensor], grads: Sequence[torch.Tensor]):
        n = len(params)
        for t, g in zip(params, grads):
            _chk(t, "param"); _chk(g, "grad")
            if t.numel() != g.numel():
                raise ValueError("parameter / gradient size mismatch")
        if self.state is None:
            dev = params[0].device
            self.state = dict(m=[torch.zeros_like(t) for t in params], v=[torch.zeros_like(t) for t in params],
                              counters=torch.zeros(2, dtype=torch.int32, device=dev))
        st = self.state
        arr = lambda ts: (ct.c_void_p * n)(*[t.data_ptr() for t in ts])
        sizes = (ct.c_int64 * n)(*[t.numel() for t in params])
        h = handle()
        c = st["counters"]
        check(h, lib.mdb_adam_update(h, n, arr(params), arr(grads), arr(st["m"]), arr(st["v"]), sizes, ct.byref(self.p),
                                     c.data_ptr(), c.data_ptr() + 4, None, None, 0, _stream()))

    @property
    def step(self) -> int:
        return 0 if self.state is None else int(self.state["counters"][0])


class SlowDepth:
    """The loop of `slow_depth` (src/simple_depth.jl:1-62): parameters `disp (1,W,H,1)` = 0.5, one
    `Pose([0,0,0.01], 0)` per source (:8-13), ADAM(3e-4) (:16), and per iteration the gradient of
    `mean(prediction_loss(warp(...))) + smooth_loss(disp, target)` (:25-41) followed by `update!` (:42).
    One iteration is a CUDA graph built once; `run(iters)` replays it with no host round trip.
    The reference's logging every 5 iterations (:44-60) is the caller's business: call `run(5)` and
    read `disp` / `poses` between the calls."""

    def __init__(self, x: torch.Tensor, K, invK, *, target_id: int = 2, source_ids: Sequence[int] = (1, 3),
                 min_depth: float = 0.1, max_depth: float = 100.0, eta: float = 3e-4, max_iters: int = 500,
                 disp0: float = 0.5, rvec0=(0.0, 0.0, 0.01), abs_subgradient: int = 0):
        _chk(x, "x")
        N, L, Cc, H, W = x.shape
        self.x = x
        dev = x.device
        f32 = dict(device=dev, dtype=torch.float32)
        self.plan = simple_depth_plan(x, K, invK, target_id=target_id, source_ids=source_ids,
                                      min_depth=min_depth, max_depth=max_depth, abs_subgradient=abs_subgradient)
        S = self.plan.S
        self.disp = torch.full((N, 1, H, W), float(disp0), **f32)
        self.poses = [Pose(torch.tensor([list(rvec0)] * N, **f32), torch.zeros((N, 1, 3), **f32)) for _ in range(S)]
        self.loss_history = torch.zeros(max_iters, **f32)
        io = self.plan.io
        io.x = x.data_ptr()
        io.disparity[0] = self.disp.data_ptr()
        for s, pz in enumerate(self.poses):
            io.rvec[s], io.tvec[s] = pz.rvec.data_ptr(), pz.tvec.data_ptr()
        io.auto_loss = None
        self.adam = _lib.AdamParams(float(eta), 0.9, 0.999, 1e-8)
        nbytes = int(lib.mdb_slow_depth_state_bytes(ct.byref(io)))
        self.state = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        self._h = ct.c_void_p()
        h = handle()
        check(h, lib.mdb_slow_depth_create(h, ct.byref(io), ct.byref(self.plan.params), ct.byref(self.adam),
                                           _p(self.state), nbytes, _p(self.plan.ws), self.plan.ws_bytes,
                                           _p(self.loss_history), max_iters, ct.byref(self._h)))
        self.iters_done = 0

    def run(self, iters: int) -> None:
        check(handle(), lib.mdb_slow_depth_run(self._h, int(iters), _stream()))
        self.iters_done += int(iters)

    def parameters(self) -> List[torch.Tensor]:
        """theta in the order of the optimiser state: disp, then (rvec, tvec) per source."""
        return [self.disp] + [t for pz in self.poses for t in (pz.rvec, pz.tvec)]

    def _state_views(self):
        """(m, v) float32 views into the state buffer per parameter, and the int32 counters [step, ticket, step', -, c...]
        (layout of mdb_slow_depth_create: per tensor m then v, counters at the next 256-byte boundary)."""
        f = self.state.view(torch.float32)
        off, ms, vs = 0, [], []
        for t in self.parameters():
            n = t.numel()
            ms.append(f[off:off + n].view(t.shape)); vs.append(f[off + n:off + 2 * n].view(t.shape))
            off += 2 * n
        c = (off * 4 + 255) // 256 * 256
        return ms, vs, self.state[c:c + 32].view(torch.int32)

    def load_state(self, params: Sequence[torch.Tensor], m: Sequence[torch.Tensor], v: Sequence[torch.Tensor], step: int):
        """Resume from a checkpoint of (theta, ADAM moments, step count)."""
        ms, vs, counters = self._state_views()
        for dst, src in zip(self.parameters() + ms + vs, list(params) + list(m) + list(v)):
            dst.copy_(src.to(dst.device, torch.float32).reshape(dst.shape))
        counters[0] = int(step)     # the ADAM kernel's step (disparity map)
        counters[1] = 0
        counters[2] = int(step)     # the finalize kernel's step (poses, loss history)
        counters[4:8] = 0           # bias corrections handed on between iterations: re-formed on the next step
        self.iters_done = int(step)

    @property
    def loss(self) -> torch.Tensor:
        return self.plan.loss

    def close(self) -> None:
        if self._h:
            lib.mdb_slow_depth_destroy(self._h)
            self._h = ct.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def slow_depth(x: torch.Tensor, ssim=None, backprojections=None, projections=None, invKs=None, Ks=None,
               transfer=None, *, target_id: int, source_ids: Sequence[int], min_depth: float, max_depth: float,
               log_dir: Optional[str] = None, iters: int = 500, log_step: int = 5, on_log: Optional[Callable] = None):
    """`slow_depth(x, ssim, backprojections, projections, invKs, Ks, transfer; target_id, source_ids,
    min_depth, max_depth, log_dir)` -- src/simple_depth.jl:1-62, same argument list (`ssim`,
    `backprojections`, `projections`, `transfer` are accepted and unused: the fused kernels carry
    them).  `on_log(iter, disp, poses, loss)` replaces the PNG / println block (:44-60) and, when given,
    is called at iteration 1 and every `log_step` iterations like the reference.  Returns
    (disp, poses, loss_history)."""
    sd = SlowDepth(x, Ks, invKs, target_id=target_id, source_ids=source_ids, min_depth=min_depth,
                   max_depth=max_depth, max_iters=iters)
    if on_log is None:
        sd.run(iters)
    else:
        done = 0
        while done < iters:
            nxt = 1 if done == 0 else min(iters, (done // log_step + 1) * log_step)
            sd.run(nxt - done)
            done = nxt
            on_log(done, sd.disp, sd.poses, sd.loss)
    out = (sd.disp, sd.poses, sd.loss_history[:iters])
    torch.cuda.current_stream().synchronize()
    sd.close()
    return out


# --------------------------------------------------------------------------------------
# ingestion of the dataset's own format
# --------------------------------------------------------------------------------------
def ingest_triplet_u8(strips: torch.Tensor, width: int, flips: Optional[torch.Tensor] = None, grayscale: bool = False,
                      out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`Depth10k.getindex` for a batch, on the device (src/dtk.jl:29-46): strips uint8 (N,H,3W,3) decoded
    rasters -> x float32 (N,3,C,H,W) [Julia (W,H,C,3,N)]; `flips` uint8 (N,) marks samples FlipX applies to."""
    if strips.dtype != torch.uint8 or not strips.is_cuda or not strips.is_contiguous():
        raise ValueError("strips must be a contiguous CUDA uint8 tensor (N,H,3W,3)")
    N, H, W3, ch = strips.shape
    if ch != 3 or W3 != 3 * width:
        raise ValueError(f"strips has shape {tuple(strips.shape)}, expected (N,H,{3 * width},3)")
    if flips is not None and (flips.dtype != torch.uint8 or not flips.is_cuda or flips.numel() != N):
        raise ValueError("flips must be a CUDA uint8 tensor of N entries")
    C = 1 if grayscale else 3
    if out is None:
        out = torch.empty((N, 3, C, H, width), device=strips.device, dtype=torch.float32)
    elif tuple(out.shape) != (N, 3, C, H, width):
        raise ValueError("out has the wrong shape")
    _chk(out, "out")
    h = handle()
    check(h, lib.mdb_ingest_triplet_u8(h, _p(strips), _p(flips), width, H, N, int(grayscale), _p(out), _stream()))
    return out


def ingest_triplet_u8_planar(strips: torch.Tensor, width: int, flips: Optional[torch.Tensor] = None,
                             out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """The split / FlipX / transpose of `ingest_triplet_u8` without the conversion: strips uint8 (N,H,3W,3) -> uint8
    (N,3,3,H,W), the stack `FusedLoss` reads directly when handed a uint8 `x` (Float32(::N0f8) happens in the kernel)."""
    if strips.dtype != torch.uint8 or not strips.is_cuda or not strips.is_contiguous():
        raise ValueError("strips must be a contiguous CUDA uint8 tensor (N,H,3W,3)")
    N, H, W3, ch = strips.shape
    if ch != 3 or W3 != 3 * width:
        raise ValueError(f"strips has shape {tuple(strips.shape)}, expected (N,H,{3 * width},3)")
    if flips is not None and (flips.dtype != torch.uint8 or not flips.is_cuda or flips.numel() != N):
        raise ValueError("flips must be a CUDA uint8 tensor of N entries")
    if out is None:
        out = torch.empty((N, 3, 3, H, width), device=strips.device, dtype=torch.uint8)
    elif tuple(out.shape) != (N, 3, 3, H, width) or out.dtype != torch.uint8 or not out.is_contiguous():
        raise ValueError("out must be a contiguous uint8 tensor (N,3,3,H,W)")
    h = handle()
    check(h, lib.mdb_ingest_triplet_u8_planar(h, _p(strips), _p(flips), width, H, N, _p(out), _stream()))
    return out


def ingest_kitti_u8(frames: torch.Tensor, height: int, width: int, flips: Optional[torch.Tensor] = None,
                    out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`KittyDataset.getindex` for a batch, on the device (src/kitty.jl:45-61): frames uint8 (N,3,Ho,Wo), the three
    consecutive gray `image_0` rasters of every sample -> `imresize` to (height, width) -> x float32 (N,3,1,H,W)
    [Julia (W,H,1,3,N)]; `flips` uint8 (N,) marks the samples FlipX applies to."""
    if frames.dtype != torch.uint8 or not frames.is_cuda or not frames.is_contiguous() or frames.dim() != 4 or frames.shape[1] != 3:
        raise ValueError("frames must be a contiguous CUDA uint8 tensor (N,3,Ho,Wo)")
    N, _, Ho, Wo = frames.shape
    if flips is not None and (flips.dtype != torch.uint8 or not flips.is_cuda or flips.numel() != N):
        raise ValueError("flips must be a CUDA uint8 tensor of N entries")
    if out is None:
        out = torch.empty((N, 3, 1, height, width), device=frames.device, dtype=torch.float32)
    elif tuple(out.shape) != (N, 3, 1, height, width):
        raise ValueError("out has the wrong shape")
    _chk(out, "out")
    h = handle()
    check(h, lib.mdb_ingest_kitti_u8(h, _p(frames), _p(flips), Wo, Ho, width, height, N, _p(out), _stream()))
    return out


# --------------------------------------------------------------------------------------
# the operators on the network side of the path
# --------------------------------------------------------------------------------------
class _PoseHeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, scale):
        N, six, hh, w = y.shape
        rvec = torch.empty((N, 3), device=y.device, dtype=torch.float32)
        tvec = torch.empty((N, 1, 3), device=y.device, dtype=torch.float32)
        check(handle(), lib.mdb_pose_head(handle(), _p(y), w, hh, N, float(scale), _p(rvec), _p(tvec), _stream()))
        ctx.shape, ctx.scale = (N, six, hh, w), float(scale)
        return rvec, tvec

    @staticmethod
    def backward(ctx, g_r, g_t):
        N, _, hh, w = ctx.shape
        dev = (g_r if g_r is not None else g_t).device
        g_r = torch.zeros((N, 3), device=dev) if g_r is None else g_r.contiguous()
        g_t = torch.zeros((N, 1, 3), device=dev) if g_t is None else g_t.contiguous()
        d_y = torch.empty(ctx.shape, device=dev, dtype=torch.float32)
        check(handle(), lib.mdb_pose_head_bwd(handle(), _p(g_r), _p(g_t), w, hh, N, ctx.scale, _p(d_y), _stream()))
        return d_y, None


def pose_head(conv_out: torch.Tensor, scale: float = 1e-2) -> Pose:
    """The tail of `(decoder::PoseDecoder)(features)` -- src/pose_decoder.jl:27-32: spatial mean of the
    (N,6,h,w) [Julia (w,h,6,N)] conv output, times 1e-2, split into `Pose(rvec (N,3), tvec (N,1,3))`.
    Differentiable w.r.t. `conv_out`."""
    _chk(conv_out, "conv_out")
    if conv_out.dim() != 4 or conv_out.shape[1] != 6:
        raise ValueError(f"conv_out has shape {tuple(conv_out.shape)}, expected (N,6,h,w)")
    return Pose(*_PoseHeadFn.apply(conv_out, scale))


class _DispHeadsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, *logits):
        logits = [_chk(t.detach(), "logits") for t in logits]
        Ls, N = len(logits), logits[0].shape[0]
        outs = [torch.empty_like(t) for t in logits]
        arr = lambda ts: (ct.c_void_p * Ls)(*[t.data_ptr() for t in ts])
        dw = (ct.c_int32 * Ls)(*[t.shape[-1] for t in logits]); dh = (ct.c_int32 * Ls)(*[t.shape[-2] for t in logits])
        check(handle(), lib.mdb_disparity_heads(handle(), arr(logits), dw, dh, Ls, N, arr(outs), _stream()))
        ctx.save_for_backward(*outs)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gs):
        outs = ctx.saved_tensors
        Ls, N = len(outs), outs[0].shape[0]
        gs = [torch.zeros_like(o) if g is None else g.contiguous() for g, o in zip(gs, outs)]
        d = [torch.empty_like(o) for o in outs]
        arr = lambda ts: (ct.c_void_p * Ls)(*[t.data_ptr() for t in ts])
        dw = (ct.c_int32 * Ls)(*[t.shape[-1] for t in outs]); dh = (ct.c_int32 * Ls)(*[t.shape[-2] for t in outs])
        check(handle(), lib.mdb_disparity_heads_bwd(handle(), arr(outs), arr(gs), dw, dh, Ls, N, arr(d), _stream()))
        return tuple(d)


def disparity_heads(logits: Sequence[torch.Tensor]) -> List[torch.Tensor]:
    """The σ of the DepthDecoder's output convolutions (src/depth_decoder.jl:46, 52-68), all scale levels in one launch:
    logits[k] (N,1,h_k,w_k) -> disparities[k], differentiable (pullback g σ (1 - σ) in one launch too)."""
    return list(_DispHeadsFn.apply(*logits))


def item_mean(m: torch.Tensor) -> torch.Tensor:
    """mean over everything but the batch axis, one value per item (deterministic)."""
    _chk(m, "map")
    N = m.shape[0]
    out = torch.empty((N,), device=m.device, dtype=torch.float32)
    check(handle(), lib.mdb_item_mean(handle(), _p(m), m.numel() // N, N, _p(out), _stream()))
    return out


def find_static(x: torch.Tensor, alpha: float, *, target_id: int = 2, source_ids: Sequence[int] = (1, 3)):
    """`find_static(dataset, alpha)` -- src/dtk.jl:51-69, for a batch at once: a sample is kept as
    NON-static when `mean(automasking_loss(ssim, x, target; source_ids)) > alpha`.  x (N,L,C,H,W) on the
    device (e.g. from `ingest_triplet_u8`).  Returns (non_static bool (N,), per-item means (N,))."""
    means = item_mean(automasking_loss(SSIM(), x, x[:, target_id - 1], source_ids=source_ids))
    return means > alpha, means


# --------------------------------------------------------------------------------------
# host-buffer entry: the call a user with HOST arrays makes (bench "e2e")
# --------------------------------------------------------------------------------------
@dataclass
class HostLossResult:
    loss: float
    d_disparities: List[torch.Tensor]   # pinned host tensors, native resolution
    d_rvec: List[torch.Tensor]          # pinned host (N,3) per source
    d_tvec: List[torch.Tensor]          # pinned host (N,1,3) per source


class HostLoss:
    """`train_loss` value + gradients for inputs that live in (pinned) HOST memory: the Python face of the C entry
    `mdb_train_loss_host` (csrc/host_pipeline.cu), which cuts the batch into chunks and pipelines host->device
    copies, the fused loss and the device->host drain over three streams.  Every chunk is one
    `mdb_train_loss_fwdbwd` with `n_global = N`, the contract the multi-GPU shards use."""

    def __init__(self, *, W, H, C, L, N, disp_sizes, scales, target_id, source_ids, K, invK, chunk=None,
                 automasking=False, n_global=0, device=None, x_format="f32", packed=None, min_depth=0.1, max_depth=100.0,
                 disparity_smoothness=1e-3, normalize_disparity=True, smooth_weight_mode=0, alpha=0.85, abs_subgradient=0):
        """x_format "f32": x is the Float32 (N,L,C,H,W) stack.  "u8_strip": x is the dataset's own format,
        uint8 (N,H,3W,3) decoded triplet strips (src/dtk.jl:29-46) -- a quarter of the PCIe bytes; the
        device converts with mdb_ingest_triplet_u8 (C = 1: Gray luma).
        packed: a `mdb200.dist.PackedGrads` of a batch sharded over ranks -- the call then ends with the path's
        all-reduce and returns the GLOBAL loss and pose gradients (disparity gradients stay this rank's shard)."""
        if x_format not in ("f32", "u8_strip"):
            raise ValueError("x_format must be 'f32' or 'u8_strip'")
        if x_format == "u8_strip" and L != 3:
            raise ValueError("triplet strips hold L = 3 frames")
        self.x_format = x_format
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.N, self.L, self.C, self.H, self.W = N, L, C, H, W
        self.S = len(source_ids)
        self.disp_sizes = [(int(w), int(h)) for (w, h) in disp_sizes]
        self.automasking = automasking
        p = _lib.LossParams()
        lib.mdb_loss_params_default(ct.byref(p))
        p.min_depth, p.max_depth, p.disparity_smoothness, p.alpha = min_depth, max_depth, disparity_smoothness, alpha
        p.automasking, p.normalize_disparity, p.smooth_weight_mode = int(automasking), int(normalize_disparity), int(smooth_weight_mode)
        p.n_global = int(packed.N if packed is not None else (n_global if n_global > 0 else N))
        p.abs_subgradient = int(abs_subgradient)
        io = _lib.LossIO()
        io.W, io.H, io.C, io.L, io.N = W, H, C, L, N
        io.target_id, io.S, io.Ls = int(target_id), self.S, len(scales)
        for s, sid in enumerate(source_ids):
            io.source_ids[s], io.invert[s] = int(sid), int(sid < target_id)
        for k, ((w, h), sc) in enumerate(zip(self.disp_sizes, scales)):
            io.dw[k], io.dh[k], io.scale[k] = w, h, float(sc)
        kk, ik = _k9(K), _k9(invK)
        for i in range(9):
            io.K[i], io.invK[i] = kk[i], ik[i]
        self.params, self.io = p, io
        self._h = ct.c_void_p()
        with torch.cuda.device(self.dev):
            hd = handle(self.dev.index)
            check(hd, lib.mdb_host_plan_create(hd, ct.byref(io), ct.byref(p), int(chunk or 0), 1 if x_format == "u8_strip" else 0,
                                               int(packed.lo) if packed is not None else -1, ct.byref(self._h)))
        info = (ct.c_int64 * 4)()
        lib.mdb_host_plan_info(self._h, info)
        self.n_chunks, self.h2d_bytes, self.d2h_bytes, self.n_pose = int(info[0]), int(info[1]), int(info[2]), int(info[3])
        pin = dict(dtype=torch.float32, pin_memory=True)
        self.h_loss = torch.empty((1,), **pin)
        self.h_dd = [torch.empty((N, 1, h, w), **pin) for (w, h) in self.disp_sizes]
        self.h_dr = [torch.empty((self.n_pose, 3), **pin) for _ in range(self.S)]
        self.h_dt = [torch.empty((self.n_pose, 1, 3), **pin) for _ in range(self.S)]

    @staticmethod
    def default_chunk(W, H, C, L, N, disp_sizes, automasking=False, x_format="f32") -> int:
        """Items per pipeline chunk the C entry picks for chunk_items <= 0: an eighth of the batch, but never chunks so
        small that the stream operations a chunk costs on the host outweigh the copy they overlap (>= ~16 MB)."""
        per_item = (4 * L * C * H * W if x_format == "f32" else 9 * H * W) \
            + 4 * (sum(w * h for (w, h) in disp_sizes) + (H * W if automasking else 0))
        return max(1, N // 8, min(N, 16 * 2 ** 20 // per_item + 1))

    def __call__(self, x, disparities, rvecs, tvecs, auto_loss=None, flips=None) -> HostLossResult:
        """All arguments are HOST tensors (pinned for full PCIe rate): float32, except x in the
        "u8_strip" format (uint8 (N,H,3W,3)) and its optional `flips` (uint8 (N,), FlipX per sample)."""
        hio = _lib.HostIO()
        keep = []

        def hp(t, dt=torch.float32):
            if t.is_cuda or t.dtype != dt or not t.is_contiguous():
                raise TypeError("host pipeline arguments must be contiguous HOST tensors of the documented dtype")
            keep.append(t)
            return t.data_ptr()
        hio.x = hp(x, torch.uint8 if self.x_format == "u8_strip" else torch.float32)
        hio.flip = hp(flips, torch.uint8) if flips is not None else None
        for k, d in enumerate(disparities):
            hio.disparity[k] = hp(d)
            hio.d_disparity[k] = self.h_dd[k].data_ptr()
        for s in range(self.S):
            hio.rvec[s], hio.tvec[s] = hp(rvecs[s]), hp(tvecs[s])
            hio.d_rvec[s], hio.d_tvec[s] = self.h_dr[s].data_ptr(), self.h_dt[s].data_ptr()
        hio.auto_loss = hp(auto_loss) if (self.automasking and auto_loss is not None) else None
        hio.loss = self.h_loss.data_ptr()
        with torch.cuda.device(self.dev):
            check(handle(self.dev.index), lib.mdb_train_loss_host(self._h, ct.byref(hio), _stream()))
        return HostLossResult(float(self.h_loss[0]), self.h_dd, self.h_dr, self.h_dt)

    def close(self) -> None:
        if self._h:
            lib.mdb_host_plan_destroy(self._h)
            self._h = ct.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def train_loss_host(plan: HostLoss, x, disparities, poses: Sequence[Pose], auto_loss=None, flips=None) -> HostLossResult:
    """Public host-buffer call: `train_loss` (src/training.jl:21-78) value + gradients for HOST
    inputs; `plan` is a HostLoss built once for the shapes (and for the format of `x`)."""
    return plan(x, disparities, [p.rvec for p in poses], [p.tvec for p in poses], auto_loss, flips)


# --------------------------------------------------------------------------------------
# plane-disparity volume rendering -- src/render.jl
# Layouts (reversed dims): rgb (B,Np,3,H,W), sigma (B,Np,1,H,W), xyz (B,Np,H,W,3),
# disparity_src (B,Np), pose.rvec (B,3), pose.tvec (B,3).
# --------------------------------------------------------------------------------------
def inv(x: torch.Tensor) -> torch.Tensor:
    """`Monodepth.inv` -- src/render.jl:3-17: (3,3) or batched (M,3,3) inverse."""
    _chk(x, "x")
    M = x.numel() // 9
    out = torch.empty_like(x)
    check(handle(), lib.mdb_inv3x3_batched(handle(), _p(x), M, _p(out), _stream()))
    return out


def create_meshgrid(H: int, W: int, device=None) -> torch.Tensor:
    """src/render.jl:21-23 -> (H,W,3) [Julia (3,W,H)]."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    out = torch.empty((H, W, 3), device=dev, dtype=torch.float32)
    check(handle(), lib.mdb_create_meshgrid(handle(), H, W, _p(out), _stream()))
    return out


def get_src_xyz_from_plane_disparity(meshgrid_src_homo, mpi_disparity_src, K_src_inv) -> torch.Tensor:
    """src/render.jl:25-30."""
    _chk(meshgrid_src_homo, "meshgrid"); _chk(mpi_disparity_src, "disparity")
    H, W = meshgrid_src_homo.shape[:2]
    B, Np = mpi_disparity_src.shape
    out = torch.empty((B, Np, H, W, 3), device=meshgrid_src_homo.device, dtype=torch.float32)
    check(handle(), lib.mdb_plane_xyz_src(handle(), _p(meshgrid_src_homo), _p(mpi_disparity_src), _k9(K_src_inv), W, H, Np, B, _p(out), _stream()))
    return out


def get_tgt_xyz_from_plane_disparity(xyz_src: torch.Tensor, pose: Pose) -> torch.Tensor:
    """src/render.jl:51-62."""
    _chk(xyz_src, "xyz_src"); _chk(pose.rvec, "rvec"); _chk(pose.tvec, "tvec")
    B, Np, H, W, _ = xyz_src.shape
    out = torch.empty_like(xyz_src)
    check(handle(), lib.mdb_plane_xyz_tgt(handle(), _p(xyz_src), _p(pose.rvec), _p(pose.tvec), W, H, Np, B, _p(out), _stream()))
    return out


def _render_ws(W, H, Np, B, dev):
    n = int(lib.mdb_render_workspace_bytes(W, H, Np, B))
    if n == 0:
        raise _lib.MdbError(-2, f"unsupported render shape W={W} H={H} Np={Np} B={B} (need W,H>=2, Np<=64)")
    return torch.empty(n, dtype=torch.uint8, device=dev), n


class _SampleFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, depth_src, rvec, tvec, K, K_inv):
        src = src.detach()
        _chk(src, "src"); _chk(depth_src, "depth_src")
        B, Np, Cs, H, W = src.shape
        ws, n = _render_ws(W, H, Np, B, src.device)
        tgt = torch.empty_like(src)
        valid = torch.empty((B, Np, H, W), dtype=torch.uint8, device=src.device)
        check(handle(), lib.mdb_sample(handle(), _p(src), _p(depth_src), _p(rvec), _p(tvec), _k9(K), _k9(K_inv),
                                       W, H, Cs, Np, B, _p(tgt), ct.c_void_p(valid.data_ptr()), _p(ws), n, _stream()))
        ctx.save_for_backward(depth_src, rvec, tvec)
        ctx.meta = (K, K_inv, tuple(src.shape), ws, n)
        valid = valid.bool()
        ctx.mark_non_differentiable(valid)
        return tgt, valid

    @staticmethod
    def backward(ctx, g, _gv):
        depth_src, rvec, tvec = ctx.saved_tensors
        K, K_inv, (B, Np, Cs, H, W), ws, n = ctx.meta
        d_src = torch.empty((B, Np, Cs, H, W), device=g.device, dtype=torch.float32)
        check(handle(), lib.mdb_sample_bwd(handle(), _p(g.contiguous()), _p(depth_src), _p(rvec), _p(tvec), _k9(K), _k9(K_inv),
                                           W, H, Cs, Np, B, _p(d_src), _p(ws), n, _stream()))
        return d_src, None, None, None, None, None


def sample(src: torch.Tensor, depth_src: torch.Tensor, pose: Pose, K, K_inv):
    """`sample` -- src/render.jl:64-90.  src (B,Np,Cs,H,W) [Julia (W,H,Cs,Np*B)], depth_src (B,Np)
    -> tgt (B,Np,Cs,H,W), valid_mask (B,Np,H,W) bool.  Differentiable w.r.t. src (mdb_sample_bwd)."""
    return _SampleFn.apply(src, depth_src, pose.rvec, pose.tvec, K, K_inv)


class _PvrFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rgb, sigma, xyz):
        rgb, sigma, xyz = rgb.detach(), sigma.detach(), xyz.detach()
        _chk(rgb, "rgb"); _chk(sigma, "sigma"); _chk(xyz, "xyz")
        B, Np, _, H, W = rgb.shape
        rgb_out = torch.empty((B, 3, H, W), device=rgb.device, dtype=torch.float32)
        tacc = torch.empty((B, Np, 1, H, W), device=rgb.device, dtype=torch.float32)
        wts = torch.empty_like(tacc)
        check(handle(), lib.mdb_plane_volume_rendering(handle(), _p(rgb), _p(sigma), _p(xyz), W, H, Np, B, _p(rgb_out), _p(tacc),
                                                       _p(wts), _stream()))
        ctx.save_for_backward(rgb, sigma, xyz)
        return rgb_out, tacc, wts

    @staticmethod
    def backward(ctx, g_rgb, g_tacc, g_w):
        rgb, sigma, xyz = ctx.saved_tensors
        B, Np, _, H, W = rgb.shape
        c = lambda t: None if t is None else t.contiguous()
        d_rgb, d_sigma, d_xyz = torch.empty_like(rgb), torch.empty_like(sigma), torch.empty_like(xyz)
        check(handle(), lib.mdb_plane_volume_rendering_bwd(handle(), _p(rgb), _p(sigma), _p(xyz), W, H, Np, B, _p(c(g_rgb)),
                                                           _p(c(g_tacc)), _p(c(g_w)), _p(d_rgb), _p(d_sigma), _p(d_xyz), _stream()))
        return d_rgb, d_sigma, d_xyz


def plane_volume_rendering(rgb: torch.Tensor, sigma: torch.Tensor, xyz: torch.Tensor):
    """src/render.jl:32-49 -> rgb_out (B,3,H,W), transparency_acc (B,Np,1,H,W), weights (B,Np,1,H,W).
    Differentiable (mdb_plane_volume_rendering_bwd)."""
    return _PvrFn.apply(rgb, sigma, xyz)


def _render_opts(mine_grid: bool, mine_mask: bool, depth_eps) -> "_lib.RenderOpts":
    o = _lib.RenderOpts()
    o.mine_grid, o.mine_mask = int(bool(mine_grid)), int(bool(mine_mask))
    o.depth_eps = 1e-5 if depth_eps is None else float(depth_eps)
    return o


class _RenderFn(torch.autograd.Function):
    """Forward = mdb_render_tgt_rgb_depth_fwd_ex; backward = mdb_render_tgt_rgb_depth_bwd_ex (the Julia
    glue registers the same pair as a ChainRulesCore.rrule).  `xyz_tgt` None = producers folded in;
    `depth_eps` None = no expected-depth output (the reference as written)."""

    @staticmethod
    def forward(ctx, rgb, sigma, disparity_src, xyz_tgt, rvec, tvec, K_inv, K, mine_grid, mine_mask, depth_eps):
        B, Np, _, H, W = rgb.shape
        ws, n = _render_ws(W, H, Np, B, rgb.device)
        f32 = dict(device=rgb.device, dtype=torch.float32)
        rgb_out = torch.empty((B, 3, H, W), **f32)
        depth = torch.empty((B, Np, 1, H, W), **f32)
        mask = torch.empty((B, 1, 1, H, W), device=rgb.device, dtype=torch.int32)
        want_z = depth_eps is not None
        edepth = torch.empty((B, 1, H, W), **f32) if want_z else None
        wsum = torch.empty((B, 1, H, W), **f32) if want_z else None
        opts = _render_opts(mine_grid, mine_mask, depth_eps)
        check(handle(), lib.mdb_render_tgt_rgb_depth_fwd_ex(handle(), _p(rgb), _p(sigma), _p(disparity_src), _p(xyz_tgt), _p(rvec),
                                                            _p(tvec), _k9(K_inv), _k9(K), W, H, Np, B, ct.byref(opts), _p(rgb_out),
                                                            _p(depth), ct.c_void_p(mask.data_ptr()), _p(edepth), _p(wsum), _p(ws), n,
                                                            _stream()))
        ctx.save_for_backward(rgb, sigma, disparity_src, xyz_tgt, rvec, tvec, depth, edepth, wsum)
        ctx.K, ctx.K_inv, ctx.ws, ctx.n, ctx.opts = K, K_inv, ws, n, opts
        ctx.mark_non_differentiable(mask)
        if want_z:
            return rgb_out, depth, mask, edepth
        return rgb_out, depth, mask

    @staticmethod
    def backward(ctx, g_rgb, g_depth, _g_mask, g_z=None):
        rgb, sigma, disp, xyz, rvec, tvec, depth, edepth, wsum = ctx.saved_tensors
        B, Np, _, H, W = rgb.shape
        g_rgb = torch.zeros((B, 3, H, W), device=rgb.device) if g_rgb is None else g_rgb.contiguous()
        g_depth = None if g_depth is None else g_depth.contiguous()
        g_z = None if g_z is None else g_z.contiguous()
        d_rgb, d_sigma = torch.empty_like(rgb), torch.empty_like(sigma)
        d_xyz = None if xyz is None else torch.empty_like(xyz)
        d_disp, d_r, d_t = torch.empty_like(disp), torch.empty_like(rvec), torch.empty_like(tvec)
        check(handle(), lib.mdb_render_tgt_rgb_depth_bwd_ex(handle(), _p(rgb), _p(sigma), _p(disp), _p(xyz), _p(rvec), _p(tvec),
                                                            _k9(ctx.K_inv), _k9(ctx.K), W, H, Np, B, ct.byref(ctx.opts), _p(depth),
                                                            _p(edepth), _p(wsum), _p(g_rgb), _p(g_depth), _p(g_z), _p(d_rgb),
                                                            _p(d_sigma), _p(d_xyz), _p(d_disp), _p(d_r), _p(d_t), _p(ctx.ws), ctx.n,
                                                            _stream()))
        return d_rgb, d_sigma, d_disp, d_xyz, d_r, d_t, None, None, None, None, None


class RenderPlan:
    """Pre-planned `render_tgt_rgb_depth` forward + pullback for fixed shapes: owns the workspace, the
    outputs and the gradient buffers, so a call is two C-ABI calls and nothing else (what the Julia rrule
    does with its cached buffers; `bench.py --config config4` times this).

    Variants (`mdb_render_opts`): `mine_grid`, `mine_mask`; `expected_depth=True` adds a
    fourth output sum w z / (sum w + depth_eps); `fused_xyz=True` means forward() is given `xyz_tgt=None` and
    the producers (src/render.jl:21-30, 51-62) are evaluated inside the kernels."""

    def __init__(self, W: int, H: int, Np: int, B: int, K_inv, K, device=None, *, mine_grid: bool = False,
                 mine_mask: bool = False, expected_depth: bool = False, depth_eps: float = 1e-5, fused_xyz: bool = False):
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.W, self.H, self.Np, self.B = W, H, Np, B
        self.K, self.K_inv = _k9(K), _k9(K_inv)
        self.ws, self.n = _render_ws(W, H, Np, B, self.dev)
        self.opts = _render_opts(mine_grid, mine_mask, depth_eps)
        self.fused_xyz = bool(fused_xyz)
        f32 = dict(device=self.dev, dtype=torch.float32)
        self.rgb_out = torch.empty((B, 3, H, W), **f32)
        self.depth = torch.empty((B, Np, 1, H, W), **f32)
        self.mask = torch.empty((B, 1, 1, H, W), device=self.dev, dtype=torch.int32)
        self.expected_depth = torch.empty((B, 1, H, W), **f32) if expected_depth else None
        self.weights_sum = torch.empty((B, 1, H, W), **f32) if expected_depth else None
        self.d_rgb = torch.empty((B, Np, 3, H, W), **f32)
        self.d_sigma = torch.empty((B, Np, 1, H, W), **f32)
        self.d_xyz = None if self.fused_xyz else torch.empty((B, Np, H, W, 3), **f32)
        self.d_disp = torch.empty((B, Np), **f32)
        self.d_rvec, self.d_tvec = torch.empty((B, 3), **f32), torch.empty((B, 3), **f32)
        self._in = None

    def forward(self, rgb, sigma, disparity_src, xyz_tgt, pose: Pose):
        for t, nm in ((rgb, "rgb"), (sigma, "sigma"), (disparity_src, "disparity_src"), (pose.rvec, "rvec"), (pose.tvec, "tvec")):
            _chk(t, nm)
        if self.fused_xyz != (xyz_tgt is None):
            raise ValueError("xyz_tgt must be None exactly when the plan was made with fused_xyz=True")
        if xyz_tgt is not None:
            _chk(xyz_tgt, "xyz_tgt")
        if tuple(rgb.shape) != (self.B, self.Np, 3, self.H, self.W):
            raise ValueError(f"rgb has shape {tuple(rgb.shape)}, planned for {(self.B, self.Np, 3, self.H, self.W)}")
        self._in = (rgb, sigma, disparity_src, xyz_tgt, pose.rvec, pose.tvec)
        h = handle()
        check(h, lib.mdb_render_tgt_rgb_depth_fwd_ex(h, _p(rgb), _p(sigma), _p(disparity_src), _p(xyz_tgt), _p(pose.rvec),
                                                     _p(pose.tvec), self.K_inv, self.K, self.W, self.H, self.Np, self.B,
                                                     ct.byref(self.opts), _p(self.rgb_out), _p(self.depth),
                                                     ct.c_void_p(self.mask.data_ptr()), _p(self.expected_depth),
                                                     _p(self.weights_sum), _p(self.ws), self.n, _stream()))
        if self.expected_depth is not None:
            return self.rgb_out, self.depth, self.mask, self.expected_depth
        return self.rgb_out, self.depth, self.mask

    def backward(self, g_rgb: torch.Tensor, g_depth: Optional[torch.Tensor] = None,
                 g_expected_depth: Optional[torch.Tensor] = None):
        """Pullback of the last forward: -> (d_rgb, d_sigma, d_disparity_src, d_xyz_tgt (None when fused), d_rvec, d_tvec)."""
        if self._in is None:
            raise RuntimeError("RenderPlan.backward before forward")
        rgb, sigma, disp, xyz, rvec, tvec = self._in
        _chk(g_rgb, "g_rgb")
        if g_depth is not None:
            _chk(g_depth, "g_depth")
        if g_expected_depth is not None:
            _chk(g_expected_depth, "g_expected_depth")
            if self.expected_depth is None:
                raise ValueError("g_expected_depth given but the plan was made without expected_depth=True")
        h = handle()
        check(h, lib.mdb_render_tgt_rgb_depth_bwd_ex(h, _p(rgb), _p(sigma), _p(disp), _p(xyz), _p(rvec), _p(tvec), self.K_inv,
                                                     self.K, self.W, self.H, self.Np, self.B, ct.byref(self.opts),
                                                     _p(self.depth), _p(self.expected_depth), _p(self.weights_sum), _p(g_rgb),
                                                     _p(g_depth), _p(g_expected_depth), _p(self.d_rgb), _p(self.d_sigma),
                                                     _p(self.d_xyz), _p(self.d_disp), _p(self.d_rvec), _p(self.d_tvec),
                                                     _p(self.ws), self.n, _stream()))
        return self.d_rgb, self.d_sigma, self.d_disp, self.d_xyz, self.d_rvec, self.d_tvec


def render_tgt_rgb_depth(rgb, sigma, disparity_src, xyz_tgt, pose: Pose, K_inv, K, *, mine_grid: bool = False,
                         mine_mask: bool = False, depth_eps: Optional[float] = None):
    """`render_tgt_rgb_depth` -- src/render.jl:92-114 -> (rgb (B,3,H,W), depth = transparency_acc
    (B,Np,1,H,W), mask (B,1,1,H,W) int32).  Differentiable (autograd) w.r.t. rgb, sigma,
    disparity_src, xyz_tgt and the pose.
    Keyword variants (defaults = the reference as written): `mine_grid`, `mine_mask`
    (test/test_sample.jl:123-137); `depth_eps` given -> a fourth output, the expected depth
    sum w z / (sum w + depth_eps) (B,1,H,W); `xyz_tgt=None` -> the producers (src/render.jl:21-30, 51-62) are
    evaluated inside the kernels and their gradient flows on into disparity_src and the pose."""
    for t, nm in ((rgb, "rgb"), (sigma, "sigma"), (disparity_src, "disparity_src"), (pose.rvec, "rvec"), (pose.tvec, "tvec")):
        _chk(t, nm)
    if xyz_tgt is not None:
        _chk(xyz_tgt, "xyz_tgt")
    return _RenderFn.apply(rgb, sigma, disparity_src, xyz_tgt, pose.rvec, pose.tvec, K_inv, K, mine_grid, mine_mask, depth_eps)
