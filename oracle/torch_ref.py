"""CPU oracle for the Monodepth2.jl view-synthesis loss path (TEST INFRASTRUCTURE ONLY).

This file is a float64 torch-CPU restatement, op for op, of the reference's Julia
functions on the hot path.  It is the *checker*: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it.  Nothing under ``monodepth2.jl_b200/`` imports it, and the product
path has no CPU fallback.

Pinning status
--------------
The reference is 100 % Julia and Julia is not installed on either machine, so the
reference itself cannot be run here.  The restatement is pinned against every
known-answer test the reference holds for this path (``test/runtests.jl``):
so3 vs. an independent rotation-vector implementation (:14-29), ``composeT``
round trip (:31-50), SSIM identities (:52-68), smooth-loss 0.2542 (:70-83), depth range
(:85-92), identity warp (:94-122) -> see ``tests/test_oracle_runtests.py``.
Those pin ``so3_exp_map``, ``hat`` (+pullback), ``composeT``, ``SSIM`` (weakly),
``smooth_loss``, ``disparity_to_depth``, ``Backproject``/``Project`` and the forward
of ``grid_sample`` (align-corners=true).  **Parity unpinned** (no reference test
asserts anything about them): the value and gradient of ``train_loss`` as a whole, the
``grid_sample`` gradient, ``upsample_bilinear`` alignment, tie conventions of
``minimum``/``abs``/``clamp``, and the whole render path (the reference's render
scripts compare against an external PyTorch repo and assert nothing).

Array convention
----------------
Julia arrays are column-major.  A Julia array of size ``(d1, ..., dk)`` is represented
here by a contiguous torch tensor of shape ``(dk, ..., d1)`` -- the *same bytes*.  So
images ``(W,H,C,N)`` are ``(N,C,H,W)``, ``rvec (3,N)`` is ``(N,3)``, ``tvec (3,1,N)`` is
``(N,1,3)``, rotation stacks ``(3,3,N)`` are ``(N,3,3)`` indexed ``[n, j, i]`` (i.e. each
3x3 is stored transposed), ``K (3,3)`` is ``K_math.T``.  Pixel coordinates are 1-based
inside the math, as in the reference.

Third-party semantics (NNlib / Zygote, not under /root/reference, versions unpinned --
no ``[compat]`` in Project.toml, Manifest.toml git-ignored) are taken from torch's
equivalents: ``grid_sample(bilinear, border, align_corners=True)``,
``pad(mode="reflect")``, ``avg_pool2d(3, 1)``, ``interpolate(bilinear,
align_corners=True)``, ``min(dim)`` pullback to one arg-min, ``abs`` -> ``sign``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

DT = torch.float64

# Pullback of `abs` at +-0 (Zygote, unpinned).  0: sign(0) = 0 (ChainRules' scalar rule).  1: ForwardDiff's
# Dual rule -- what Zygote's broadcast differentiates `abs.(...)` with on CuArrays: abs(d) = signbit(value) ? -d : d,
# i.e. +1 at +0.0 and -1 at -0.0.  Not measure-zero for `slow_depth`: disp starts as fill(0.5), so every
# smoothness difference of iteration 1 is exactly +0.0 (src/simple_depth.jl:8,39).
ABS_SUBGRADIENT = 0


class _Abs(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return x.abs()

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        s = torch.sign(x)
        if ABS_SUBGRADIENT:
            s = torch.where(x == 0, torch.where(torch.signbit(x), -torch.ones_like(x), torch.ones_like(x)), s)
        return g * s


def _abs(x: torch.Tensor) -> torch.Tensor:
    return _Abs.apply(x)


# --------------------------------------------------------------------------------------
# helpers for the reversed-dims convention
# --------------------------------------------------------------------------------------
def _m(a: torch.Tensor) -> torch.Tensor:
    """(…,3,3) Julia-layout stack  <->  math-natural [.., i, j] (an involution)."""
    return a.transpose(-1, -2)


@dataclass
class Pose:
    """`struct Pose{R,T}` -- src/pose_decoder.jl:1-5.  rvec (N,3) [Julia (3,N)],
    tvec (N,1,3) [Julia (3,1,N)]; render code passes tvec (B,3) [Julia (3,B)]."""
    rvec: torch.Tensor
    tvec: torch.Tensor


@dataclass
class Params:
    """`Params` -- src/Monodepth.jl:37-47 (fields used by the loss path)."""
    min_depth: float = 0.1
    max_depth: float = 100.0
    disparity_smoothness: float = 1e-3
    automasking: bool = True
    target_size: Tuple[int, int] = (416, 128)  # (width, height)


# --------------------------------------------------------------------------------------
# geometry primitives -- src/utils.jl
# --------------------------------------------------------------------------------------
def hat_m(rvec: torch.Tensor) -> torch.Tensor:
    """Math-natural skew matrices S[n,i,j] -- src/utils.jl:123-132."""
    n = rvec.shape[0]
    S = torch.zeros(n, 3, 3, dtype=rvec.dtype)
    r1, r2, r3 = rvec[:, 0], rvec[:, 1], rvec[:, 2]
    # Julia 1-based S[2,1]=r3, S[1,2]=-r3, S[3,1]=-r2, S[1,3]=r2, S[3,2]=r1, S[2,3]=-r1
    S = S.clone()
    S[:, 1, 0] = r3
    S[:, 0, 1] = -r3
    S[:, 2, 0] = -r2
    S[:, 0, 2] = r2
    S[:, 2, 1] = r1
    S[:, 1, 2] = -r1
    return S


def hat(rvec: torch.Tensor) -> torch.Tensor:
    """`hat(rvec)` in Julia layout: (N,3) -> (N,3,3) [Julia (3,3,N)]."""
    return _m(hat_m(rvec)).contiguous()


def hat_pullback(d: torch.Tensor) -> torch.Tensor:
    """The hand-written `rrule(::typeof(hat), v)` -- src/utils.jl:134-145.
    `d` is the cotangent in Julia layout (N,3,3) = [n, j, i]; returns (N,3)."""
    dm = _m(d)  # dm[n,i,j] = d[i,j,n]
    g1 = dm[:, 2, 1] - dm[:, 1, 2]
    g2 = -dm[:, 2, 0] + dm[:, 0, 2]
    g3 = dm[:, 1, 0] - dm[:, 0, 1]
    return torch.stack([g1, g2, g3], dim=1)


def so3_exp_map_m(rvec: torch.Tensor) -> torch.Tensor:
    """Rodrigues, math-natural R[n,i,j] -- src/utils.jl:106-121."""
    S = hat_m(rvec)
    S2 = S @ S
    theta = torch.sqrt((rvec * rvec).sum(dim=1))  # NaN gradient at 0, as the reference
    theta_inv = 1.0 / torch.clamp(theta, min=1e-4)  # max.(θ, 1e-4)
    f1 = (theta_inv * torch.sin(theta)).reshape(-1, 1, 1)
    f2 = (theta_inv * theta_inv * (1.0 - torch.cos(theta))).reshape(-1, 1, 1)
    eye = torch.eye(3, dtype=rvec.dtype).unsqueeze(0)
    return f1 * S + f2 * S2 + eye


def so3_exp_map(rvec: torch.Tensor) -> torch.Tensor:
    """`so3_exp_map(rvec)`: (N,3) -> (N,3,3) Julia layout."""
    return _m(so3_exp_map_m(rvec)).contiguous()


def composeT_m(rvec: torch.Tensor, t: torch.Tensor, invert: bool):
    """`composeT` -- src/utils.jl:185-192; math-natural R (N,3,3), t (N,3)."""
    R = so3_exp_map_m(rvec)
    t = t.reshape(-1, 3)
    if invert:
        R = R.transpose(1, 2)
        t = (R @ (-t).unsqueeze(-1)).squeeze(-1)
    return R, t


def composeT(rvec: torch.Tensor, tvec: torch.Tensor, invert: bool):
    """Julia-layout wrapper: returns R (N,3,3) [n,j,i], t (N,1,3)."""
    R, t = composeT_m(rvec, tvec, invert)
    return _m(R).contiguous(), t.reshape(-1, 1, 3)


def disparity_to_depth(disparity: torch.Tensor, min_depth: float, max_depth: float):
    """src/utils.jl:179-183."""
    min_disp = 1.0 / max_depth
    max_disp = 1.0 / min_depth
    return 1.0 / (disparity * (max_disp - min_disp) + min_disp)


def backproject_coordinates(width: int, height: int, dtype=DT) -> torch.Tensor:
    """`Backproject` ctor -- src/utils.jl:49-57: (3, W*H) = [w, h, 1], 1-based,
    linear index w + (h-1) W.  Returned math-natural as (3, H*W)."""
    hh, ww = torch.meshgrid(
        torch.arange(1, height + 1, dtype=dtype), torch.arange(1, width + 1, dtype=dtype),
        indexing="ij")
    return torch.stack([ww.reshape(-1), hh.reshape(-1), torch.ones(height * width, dtype=dtype)], 0)


def backproject(depth: torch.Tensor, invK_m: torch.Tensor, width: int, height: int):
    """`(b::Backproject)(depth, invK)` -- src/utils.jl:67-69.
    depth (N, H*W) ; returns math-natural points (N, 3, H*W)."""
    coords = backproject_coordinates(width, height, depth.dtype)
    rays = invK_m @ coords  # (3, HW)
    return depth.unsqueeze(1) * rays.unsqueeze(0)


def project(points: torch.Tensor, K_m: torch.Tensor, R_m: torch.Tensor, t_m: torch.Tensor,
            width: int, height: int) -> torch.Tensor:
    """`(p::Project)(points, K, R, t)` + `normalize` -- src/utils.jl:83-85, 99-103.
    points (N,3,HW), R (N,3,3) math, t (N,3) -> normalised grid (N, HW, 2) in [-1,1]."""
    cam = K_m.unsqueeze(0) @ (R_m @ points + t_m.unsqueeze(-1))
    denom = 1.0 / (cam[:, 2:3, :] + 1e-7)
    px = cam[:, 0:2, :] * denom  # (N,2,HW) 1-based pixel coordinates
    normalizer = torch.tensor([width - 1.0, height - 1.0], dtype=points.dtype).reshape(1, 2, 1)
    uv = (((px - 1.0) / normalizer) - 0.5) * 2.0
    return uv.permute(0, 2, 1)


def grid_sample(inp: torch.Tensor, grid: torch.Tensor) -> torch.Tensor:
    """NNlib `grid_sample(input, grid; padding_mode=:border)` (align-corners=true;
    pinned by test/runtests.jl:94-117).  inp (N,C,H,W), grid (N,Ho,Wo,2)."""
    return F.grid_sample(inp, grid, mode="bilinear", padding_mode="border", align_corners=True)


UPSAMPLE_ALIGN_CORNERS = True     # NNlib's default; False = half-pixel centres (nothing in the reference pins it)


def upsample_bilinear(x: torch.Tensor, size_wh: Tuple[int, int]) -> torch.Tensor:
    """NNlib `upsample_bilinear(x; size=(W,H))`, align-corners=true (UNPINNED: `UPSAMPLE_ALIGN_CORNERS`)."""
    w, h = size_wh
    return F.interpolate(x, size=(h, w), mode="bilinear", align_corners=UPSAMPLE_ALIGN_CORNERS)


# --------------------------------------------------------------------------------------
# SSIM / photometric losses -- src/utils.jl:17-43, src/training.jl:1-19
# --------------------------------------------------------------------------------------
def ssim(x: torch.Tensor, y: torch.Tensor, c1: float = 0.01 ** 2, c2: float = 0.03 ** 2):
    """`(ssim::SSIM)(x, y)` -- src/utils.jl:29-43.  x, y (N,C,H,W)."""
    xr = F.pad(x, (1, 1, 1, 1), mode="reflect")
    yr = F.pad(y, (1, 1, 1, 1), mode="reflect")
    pool = lambda a: F.avg_pool2d(a, 3, 1)
    mx, my = pool(xr), pool(yr)
    sx = pool(xr * xr) - mx * mx
    sy = pool(yr * yr) - my * my
    sxy = pool(xr * yr) - mx * my
    n = (2.0 * mx * my + c1) * (2.0 * sxy + c2)
    d = (mx * mx + my * my + c1) * (sx + sy + c2)
    return torch.clamp((1.0 - n / d) * 0.5, 0.0, 1.0)


def photometric_loss(pred: torch.Tensor, target: torch.Tensor, alpha: float = 0.85):
    """src/training.jl:1-7 -> (N,1,H,W)."""
    l1 = _abs(target - pred).mean(dim=1, keepdim=True)
    s = ssim(pred, target).mean(dim=1, keepdim=True)
    return alpha * s + (1.0 - alpha) * l1


def _min_first(stack: torch.Tensor) -> torch.Tensor:
    """`minimum(cat(...; dims=3); dims=3)` with the pullback routed to the FIRST
    arg-min (oracle convention).  stack (N,S,H,W) -> (N,1,H,W)."""
    idx = torch.argmin(stack, dim=1, keepdim=True)  # torch.argmin returns the first minimum
    return torch.gather(stack, 1, idx)


def automasking_loss(x: torch.Tensor, target: torch.Tensor, source_ids: Sequence[int]):
    """src/training.jl:9-11.  x (N,L,C,H,W), 1-based source ids."""
    return _min_first(torch.cat([photometric_loss(x[:, i - 1], target) for i in source_ids], 1))


def prediction_loss(predictions: Sequence[torch.Tensor], target: torch.Tensor):
    """src/training.jl:13-15."""
    return _min_first(torch.cat([photometric_loss(p, target) for p in predictions], 1))


def apply_mask(mask: torch.Tensor, warp_loss: torch.Tensor):
    """`_apply_mask` -- src/training.jl:17-19 (mask precedes warp_loss)."""
    return _min_first(torch.cat([mask, warp_loss], 1))


def smooth_loss(disparity: torch.Tensor, image: torch.Tensor):
    """src/utils.jl:163-177.  disparity (N,H,W) [Julia (W,H,N)], image (N,C,H,W)."""
    ddx = _abs(disparity[:, :, :-1] - disparity[:, :, 1:])
    ddy = _abs(disparity[:, :-1, :] - disparity[:, 1:, :])
    dix = (image[:, :, :, :-1] - image[:, :, :, 1:]).abs().mean(dim=1)
    diy = (image[:, :, :-1, :] - image[:, :, 1:, :]).abs().mean(dim=1)
    return (ddx * torch.exp(-dix)).mean() + (ddy * torch.exp(-diy)).mean()


# --------------------------------------------------------------------------------------
# warp + multi-scale loss assembly -- src/training.jl:21-78, src/simple_depth.jl:25-41
# --------------------------------------------------------------------------------------
def _Km(K: torch.Tensor) -> torch.Tensor:
    """Julia-layout (3,3) -> math-natural."""
    return K.reshape(3, 3).transpose(0, 1)


def warp(disp: torch.Tensor, x: torch.Tensor, Ps, invK: torch.Tensor, K: torch.Tensor, *,
         min_depth: float, max_depth: float, source_ids: Sequence[int]) -> List[torch.Tensor]:
    """`warp(...)` -- called at src/simple_depth.jl:30-32 but undefined in the fork; body
    restated from the inlined copy at src/training.jl:48-57.
    disp (N,1,H,W) full resolution, x (N,L,C,H,W), Ps = [(R_m (N,3,3) math, t_m (N,3))]."""
    n, _, h, w = disp.shape
    depth = disparity_to_depth(disp, min_depth, max_depth)
    coords = backproject(depth.reshape(n, h * w), _Km(invK), w, h)
    out = []
    for (R_m, t_m), sid in zip(Ps, source_ids):
        uvs = project(coords, _Km(K), R_m, t_m, w, h).reshape(n, h, w, 2)
        out.append(grid_sample(x[:, sid - 1], uvs))
    return out


def train_loss(x: torch.Tensor, disparities: Sequence[torch.Tensor], poses: Sequence[Pose],
               auto_loss: Optional[torch.Tensor], K: torch.Tensor, invK: torch.Tensor,
               params: Params, *, target_id: int = 2, source_ids: Sequence[int] = (1, 3),
               scales: Sequence[float] = (0.125, 0.25, 0.5, 1.0), return_parts: bool = False):
    """`train_loss` -- src/training.jl:21-78 with the network's outputs
    (`disparities`, `poses`) supplied by the caller.
    x (N,L,C,H,W); disparities[k] (N,1,h_k,w_k); auto_loss (N,1,H,W) or None."""
    target = x[:, target_id - 1]
    Ps = [composeT_m(p.rvec, p.tvec, sid < target_id) for p, sid in zip(poses, source_ids)]
    width, height = params.target_size
    loss = torch.zeros((), dtype=x.dtype)
    parts = []
    for disparity, scale in zip(disparities, scales):
        if disparity.shape[-1] != width or disparity.shape[-2] != height:
            disparity = upsample_bilinear(disparity, (width, height))
        warped = warp(disparity, x, Ps, invK, K, min_depth=params.min_depth,
                      max_depth=params.max_depth, source_ids=source_ids)
        warp_loss = prediction_loss(warped, target)
        if params.automasking:
            warp_loss = apply_mask(auto_loss, warp_loss)
        norm_disp = (disparity / (disparity.mean(dim=(2, 3), keepdim=True) + 1e-7))[:, 0]
        disp_loss = smooth_loss(norm_disp, target) * params.disparity_smoothness * scale
        loss = loss + warp_loss.mean() + disp_loss
        parts.append((warp_loss, disp_loss, warped))
    loss = loss / len(scales)
    return (loss, parts) if return_parts else loss


def simple_depth_loss(x: torch.Tensor, disp: torch.Tensor, poses: Sequence[Pose],
                      K: torch.Tensor, invK: torch.Tensor, *, target_id: int = 2,
                      source_ids: Sequence[int] = (1, 3), min_depth: float = 0.1,
                      max_depth: float = 100.0):
    """The closure differentiated in `slow_depth` -- src/simple_depth.jl:25-41.
    disp (1,H,W,1) [Julia (1,W,H,1)] is reinterpreted as (N=1,1,H,W)."""
    n = x.shape[0]
    h, w = x.shape[-2:]
    d4 = disp.reshape(n, 1, h, w)
    target = x[:, target_id - 1]
    Ps = [composeT_m(p.rvec, p.tvec, sid < target_id) for p, sid in zip(poses, source_ids)]
    warped = warp(d4, x, Ps, invK, K, min_depth=min_depth, max_depth=max_depth,
                  source_ids=source_ids)
    warp_loss = prediction_loss(warped, target).mean()
    depth_loss = smooth_loss(d4[:, 0], target)
    return warp_loss + depth_loss


# --------------------------------------------------------------------------------------
# plane-disparity rendering -- src/render.jl (quirks F8 reproduced as written)
# --------------------------------------------------------------------------------------
def create_meshgrid(H: int, W: int, dtype=DT) -> torch.Tensor:
    """src/render.jl:21-23: Julia (3,W,H) = [w,h,1] 1-based -> here (H,W,3)."""
    hh, ww = torch.meshgrid(torch.arange(1, H + 1, dtype=dtype),
                            torch.arange(1, W + 1, dtype=dtype), indexing="ij")
    return torch.stack([ww, hh, torch.ones_like(ww)], dim=-1)


def get_src_xyz_from_plane_disparity(meshgrid: torch.Tensor, disparity: torch.Tensor,
                                     K_inv: torch.Tensor) -> torch.Tensor:
    """src/render.jl:25-30.  meshgrid (H,W,3), disparity (B,Np) [Julia (Np,B)],
    K_inv Julia layout -> xyz (B,Np,H,W,3) [Julia (3,W,H,Np,B)]."""
    rays = meshgrid @ _Km(K_inv).transpose(0, 1)  # (H,W,3): K_inv_m @ [w,h,1]
    depth = (1.0 / disparity)[:, :, None, None, None]
    return rays[None, None] * depth


def get_tgt_xyz_from_plane_disparity(xyz_src: torch.Tensor, pose: Pose) -> torch.Tensor:
    """src/render.jl:51-62: R(rvec) xyz + tvec, no invert."""
    R = so3_exp_map_m(pose.rvec)  # (B,3,3)
    t = pose.tvec.reshape(-1, 3)
    return torch.einsum("bij,bnhwj->bnhwi", R, xyz_src) + t[:, None, None, None, :]


def homography_src_tgt(depth_src: torch.Tensor, pose: Pose, K: torch.Tensor,
                       K_inv: torch.Tensor) -> torch.Tensor:
    """src/render.jl:66-74: H_src_tgt = inv(K (R - t n^T / (-depth)) K_inv), n=[0,0,1].
    depth_src (B,Np) -> (B,Np,3,3) math-natural."""
    R = so3_exp_map_m(pose.rvec)  # (B,3,3)
    t = pose.tvec.reshape(-1, 3)
    n = torch.tensor([0.0, 0.0, 1.0], dtype=R.dtype)
    tn = t[:, :, None] * n[None, None, :]  # (B,3,3) outer t n^T
    temp = tn[:, None] / (-depth_src)[:, :, None, None]
    H_tgt_src = _Km(K) @ (R[:, None] - temp) @ _Km(K_inv)
    return torch.linalg.inv(H_tgt_src)


def sample(src: torch.Tensor, depth_src: torch.Tensor, pose: Pose, K: torch.Tensor,
           K_inv: torch.Tensor, *, mine_grid: bool = False, mine_mask: bool = False):
    """`sample` -- src/render.jl:64-90.  src (B*Np? ,C,H,W) given as (B,Np,C,H,W);
    returns tgt (B,Np,C,H,W) and valid_mask (B,Np,H,W) bool.
    `mine_grid` / `mine_mask`: what the reference's comparison script quotes from MINE
    (test/test_sample.jl:123-137) instead of the code as written."""
    B, Np, C, H, W = src.shape
    Hm = homography_src_tgt(depth_src, pose, K, K_inv)  # (B,Np,3,3)
    mesh = create_meshgrid(H, W, src.dtype) - torch.tensor([1.0, 1.0, 0.0], dtype=src.dtype)
    ms = torch.einsum("bnij,hwj->bnhwi", Hm, mesh)
    xy = ms[..., 0:2] / ms[..., 2:3]
    gx, gy = xy[..., 0], xy[..., 1]
    if mine_mask:   # test/test_sample.jl:123-125
        valid = ((gx < W) & (gx > -1)) & ((gy < H) & (gy > -1))
    else:           # Julia parses `a .< W .* a .>= 0` as (a < W a) & (W a >= 0)  
        valid = ((gx < W * gx) & (W * gx >= 0)) & ((gy < H * gy) & (H * gy >= 0))
    if mine_grid:   # test/test_sample.jl:131-137, torch's own sampler
        grid = torch.stack([(gx + 0.5) / (W * 0.5) - 1, (gy + 0.5) / (H * 0.5) - 1], dim=-1)
        tgt = torch.nn.functional.grid_sample(src.reshape(B * Np, C, H, W), grid.reshape(B * Np, H, W, 2),
                                              mode="bilinear", padding_mode="border", align_corners=False)
    else:
        grid = torch.stack([(gx + 0.5) / (W / 2), (gy + 0.5) / (H / 2)], dim=-1)  # no "-1" (F8a)
        tgt = grid_sample(src.reshape(B * Np, C, H, W), grid.reshape(B * Np, H, W, 2))
    return tgt.reshape(B, Np, C, H, W), valid


def plane_volume_rendering(rgb: torch.Tensor, sigma: torch.Tensor, xyz: torch.Tensor):
    """src/render.jl:32-49.  rgb (B,Np,3,H,W), sigma (B,Np,1,H,W), xyz (B,Np,H,W,3)
    -> rgb_out (B,3,H,W), transparency_acc (B,Np,1,H,W), weights (B,Np,1,H,W)."""
    B, Np, _, H, W = rgb.shape
    diff = xyz[:, 1:] - xyz[:, :-1]
    dist = torch.sqrt((diff * diff).sum(dim=-1)).unsqueeze(2)  # (B,Np-1,1,H,W)
    dist = torch.cat([dist, torch.full((B, 1, 1, H, W), 1e3, dtype=rgb.dtype)], dim=1)
    transparency = torch.exp(-dist * sigma)
    alpha = 1.0 - transparency
    tacc = torch.cumprod(transparency + 1e-6, dim=1)
    tacc = torch.cat([torch.ones((B, 1, 1, H, W), dtype=rgb.dtype), tacc[:, :-1]], dim=1)
    weights = tacc * alpha
    rgb_out = (weights * rgb).sum(dim=1)
    return rgb_out, tacc, weights


def render_tgt_rgb_depth(rgb, sigma, disparity_src, xyz_tgt, pose: Pose, K_inv, K, *, mine_grid: bool = False,
                         mine_mask: bool = False, depth_eps=None):
    """src/render.jl:92-114.  rgb (B,Np,3,H,W), sigma (B,Np,1,H,W), disparity_src (B,Np),
    xyz_tgt (B,Np,H,W,3) -> rgb (B,3,H,W), "depth" = transparency_acc (B,Np,1,H,W) (F8e),
    mask (B,1,1,H,W) integer count of valid planes.
    Variants: `xyz_tgt=None` derives it from the producers (src/render.jl:21-30, 51-62);
    `depth_eps` given -> a fourth output, the expected depth sum w z / (sum w + depth_eps), (B,1,H,W)
    ("TODO: return depth?", src/render.jl:47)."""
    depth_src = 1.0 / disparity_src
    if xyz_tgt is None:
        H, W = rgb.shape[-2:]
        xyz_tgt = get_tgt_xyz_from_plane_disparity(
            get_src_xyz_from_plane_disparity(create_meshgrid(H, W, rgb.dtype), disparity_src, K_inv), pose)
    packed = torch.cat([rgb, sigma, xyz_tgt.permute(0, 1, 4, 2, 3)], dim=2)  # 7 channels
    tgt, valid = sample(packed, depth_src, pose, K, K_inv, mine_grid=mine_grid, mine_mask=mine_mask)
    rgb_s, sigma_s, xyz_s = tgt[:, :, 0:3], tgt[:, :, 3:4], tgt[:, :, 4:7]
    sigma_s = sigma_s * (sigma_s >= 0)
    rgb_o, depth, weights = plane_volume_rendering(rgb_s, sigma_s, xyz_s.permute(0, 1, 3, 4, 2))
    mask = valid.sum(dim=1, keepdim=True).unsqueeze(1).to(torch.int64)  # (B,1,1,H,W)
    if depth_eps is None:
        return rgb_o, depth, mask
    dsyn = (weights * xyz_s[:, :, 2:3]).sum(dim=1) / (weights.sum(dim=1) + depth_eps)
    return rgb_o, depth, mask, dsyn


# --------------------------------------------------------------------------------------
# the optimiser loop of slow_depth -- src/simple_depth.jl:8-42  
# --------------------------------------------------------------------------------------
def flux_adam_step(params, grads, state, eta: float = 3e-4, beta=(0.9, 0.999), eps: float = 1e-8):
    """`Flux.Optimise.update!(ADAM(eta, beta), params, grads)` (third-party Flux.jl <= 0.13,
    `apply!(o::ADAM, x, Δ)`; unpinned): mt = b1 mt + (1-b1) Δ; vt = b2 vt + (1-b2) Δ^2;
    Δ' = mt / (1 - b1^t) / (sqrt(vt / (1 - b2^t)) + eps) * eta; x .-= Δ'.  `state` is a dict that
    this function fills on first use.  In place on `params`."""
    if not state:
        state.update(m=[torch.zeros_like(p) for p in params], v=[torch.zeros_like(p) for p in params], t=0)
    state["t"] += 1
    t = state["t"]
    for p, g, m, v in zip(params, grads, state["m"], state["v"]):
        m.mul_(beta[0]).add_(g, alpha=1 - beta[0])
        v.mul_(beta[1]).addcmul_(g, g, value=1 - beta[1])
        p.sub_(m / (1 - beta[0] ** t) / (torch.sqrt(v / (1 - beta[1] ** t)) + eps) * eta)


def slow_depth(x: torch.Tensor, K: torch.Tensor, invK: torch.Tensor, *, iters: int, target_id: int = 2,
               source_ids: Sequence[int] = (1, 3), min_depth: float = 0.1, max_depth: float = 100.0,
               eta: float = 3e-4):
    """`slow_depth` -- src/simple_depth.jl:1-62 without its logging.  x (N,L,C,H,W).
    Returns (disp (N,1,H,W), poses, [loss per iteration])."""
    n, _, _, h, w = x.shape
    dt = x.dtype
    disp = torch.full((n, 1, h, w), 0.5, dtype=dt)                                   # :8
    poses = [Pose(torch.tensor([[0.0, 0.0, 0.01]] * n, dtype=dt), torch.zeros(n, 1, 3, dtype=dt))
             for _ in source_ids]                                                    # :9-13
    state, hist = {}, []
    for _ in range(iters):                                                           # :22
        leaves = [disp.requires_grad_()] + [t.requires_grad_() for p in poses for t in (p.rvec, p.tvec)]
        loss = simple_depth_loss(x, disp.reshape(n, h, w, 1), poses, K, invK, target_id=target_id,
                                 source_ids=source_ids, min_depth=min_depth, max_depth=max_depth)   # :25-41
        grads = torch.autograd.grad(loss, leaves)
        hist.append(float(loss.detach()))
        with torch.no_grad():
            for t in leaves:
                t.requires_grad_(False)
            flux_adam_step(leaves, grads, state, eta=eta)                            # :42
    return disp, poses, hist


# --------------------------------------------------------------------------------------
# network-side boundary -- src/pose_decoder.jl:27-32, src/dtk.jl:51-69  
# --------------------------------------------------------------------------------------
def pose_head(conv_out: torch.Tensor, scale: float = 1e-2) -> Pose:
    """Tail of `(decoder::PoseDecoder)(features)` -- src/pose_decoder.jl:29-31: `mean(...; dims=(1,2))`,
    `1e-2 .* reshape(pose, (6,1,N))`, `Pose(pose[1:3,1,:], pose[4:6,:,:])`.  conv_out (N,6,h,w)."""
    pose = scale * conv_out.mean(dim=(2, 3))              # (N,6)
    return Pose(pose[:, 0:3], pose[:, 3:6].reshape(-1, 1, 3))


def find_static_means(x: torch.Tensor, *, target_id: int = 2, source_ids: Sequence[int] = (1, 3)) -> torch.Tensor:
    """The statistic `find_static` thresholds -- src/dtk.jl:59-62: mean of the automasking loss per sample."""
    return automasking_loss(x, x[:, target_id - 1], source_ids).mean(dim=(1, 2, 3))
