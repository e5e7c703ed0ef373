"""ctypes front end of the plain-C oracle ``oracle/c_port/mdb_oracle.c`` (TEST INFRASTRUCTURE ONLY).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; nothing under ``monodepth2.jl_b200/`` does.

The C file is a second restatement of the reference's loss path (src/training.jl:21-78 and the
functions of src/utils.jl it calls) with every pullback written by hand; ``oracle/torch_ref.py`` is
the first one (torch autograd).  ``tests/test_oracle_c.py`` holds them against each other.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "c_port")
MAX_S = 4
MAX_SCALES = 8


class Desc(C.Structure):
    """``mdbo_desc`` of mdb_oracle.c."""
    _fields_ = [
        ("W", C.c_int32), ("H", C.c_int32), ("C", C.c_int32), ("L", C.c_int32), ("N", C.c_int32),
        ("target_id", C.c_int32), ("S", C.c_int32),
        ("source_ids", C.c_int32 * MAX_S),
        ("Ls", C.c_int32),
        ("dw", C.c_int32 * MAX_SCALES), ("dh", C.c_int32 * MAX_SCALES),
        ("scales", C.c_double * MAX_SCALES),
        ("min_depth", C.c_double), ("max_depth", C.c_double), ("smoothness", C.c_double),
        ("alpha", C.c_double), ("c1", C.c_double), ("c2", C.c_double),
        ("automask", C.c_int32), ("normalize", C.c_int32), ("smooth_mode", C.c_int32), ("n_global", C.c_int32),
        ("abs_subgradient", C.c_int32), ("upsample_half_pixel", C.c_int32),
    ]


def build(force: bool = False) -> None:
    """Compile both variants in place (gcc, a second or two)."""
    if force:
        subprocess.run(["make", "-C", _DIR, "clean"], check=True, capture_output=True)
    subprocess.run(["make", "-C", _DIR], check=True, capture_output=True)


_libs = {}


def lib(dtype=np.float64):
    key = np.dtype(dtype).name
    if key not in _libs:
        name = {"float64": "libmdb_oracle_f64.so", "float32": "libmdb_oracle_f32.so"}[key]
        path = os.path.join(_DIR, name)
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        assert L.mdbo_sizeof_real() == np.dtype(dtype).itemsize
        L.mdbo_train_loss.restype = C.c_int
        L.mdbo_set_decisions.restype = None
        L.mdbo_set_decisions.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _libs[key] = L
    return _libs[key]


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def train_loss(x: np.ndarray, disparities: Sequence[np.ndarray], rvecs: Sequence[np.ndarray],
               tvecs: Sequence[np.ndarray], auto_loss: Optional[np.ndarray], K: np.ndarray, invK: np.ndarray, *,
               target_id: int = 2, source_ids: Sequence[int] = (1, 3),
               scales: Sequence[float] = (0.125, 0.25, 0.5, 1.0), min_depth: float = 0.1, max_depth: float = 100.0,
               smoothness: float = 1e-3, alpha: float = 0.85, automasking: bool = True, normalize: bool = True,
               smooth_mode: int = 0, n_global: int = 0, grad: bool = True, dtype=np.float64, threads: int = 1,
               abs_subgradient: int = 0, force_decisions: Optional[np.ndarray] = None, dump_decisions: bool = False,
               upsample_align_corners: bool = True):
    """Value and gradients of ``train_loss`` (or of the slow_depth closure with ``normalize=False,
    smooth_mode=1`` and one scale).  Arrays are in the repo's reversed-dims convention:
    x (N,L,C,H,W), disparities[k] (N,1,h,w), rvecs[s] (N,3), tvecs[s] (N,1,3) or (N,3), auto_loss
    (N,1,H,W) or None, K / invK the 9 numbers Julia stores.
    abs_subgradient: pullback of abs at +-0 (0: sign(0) = 0; 1: ForwardDiff's rule, +1 at +0.0).
    upsample_align_corners: alignment of `upsample_bilinear` (src/training.jl:45): True = NNlib's default.
    force_decisions: int8 (Ls,N,2C+2,H,W) -- every discrete choice of the pullback (layout in mdb_oracle.c) taken
    from this array instead of the oracle's own arithmetic; dump_decisions: also return the oracle's own choices
    (`decisions`) and the arg-min margins (`margin` (Ls,N,H,W))."""
    L = lib(dtype)
    x = np.ascontiguousarray(x, dtype=dtype)
    N, Lf, Cc, H, W = x.shape
    S, Ls = len(rvecs), len(disparities)
    d = Desc()
    d.W, d.H, d.C, d.L, d.N, d.target_id, d.S, d.Ls = W, H, Cc, Lf, N, target_id, S, Ls
    for i, sid in enumerate(source_ids):
        d.source_ids[i] = int(sid)
    disp = [np.ascontiguousarray(a, dtype=dtype) for a in disparities]
    for k, a in enumerate(disp):
        d.dh[k], d.dw[k] = a.shape[-2], a.shape[-1]
        d.scales[k] = float(scales[k])
    d.min_depth, d.max_depth, d.smoothness, d.alpha = min_depth, max_depth, smoothness, alpha
    d.c1, d.c2 = 0.01 ** 2, 0.03 ** 2                       # src/utils.jl:23
    d.automask, d.normalize, d.smooth_mode, d.n_global = int(automasking), int(normalize), smooth_mode, n_global
    d.abs_subgradient = int(abs_subgradient)
    d.upsample_half_pixel = 0 if upsample_align_corners else 1
    rv = np.ascontiguousarray(np.stack([np.asarray(r, dtype=dtype).reshape(N, 3) for r in rvecs]))
    tv = np.ascontiguousarray(np.stack([np.asarray(t, dtype=dtype).reshape(N, 3) for t in tvecs]))
    au = None if (auto_loss is None or not automasking) else np.ascontiguousarray(auto_loss, dtype=dtype)
    Kc = np.ascontiguousarray(np.asarray(K, dtype=dtype).reshape(9))
    iKc = np.ascontiguousarray(np.asarray(invK, dtype=dtype).reshape(9))
    loss = C.c_double(0.0)
    disp_ptrs = (C.c_void_p * Ls)(*[a.ctypes.data for a in disp])
    if grad:
        g_disp = [np.zeros_like(a) for a in disp]
        g_ptrs = (C.c_void_p * Ls)(*[a.ctypes.data for a in g_disp])
        g_rv, g_tv = np.zeros_like(rv), np.zeros_like(tv)
    else:
        g_disp, g_ptrs, g_rv, g_tv = None, None, None, None
    force = None
    if force_decisions is not None:
        force = np.ascontiguousarray(force_decisions, dtype=np.int8)
        assert force.shape == (Ls, N, 2 * Cc + 2, H, W), force.shape
    dump = np.zeros((Ls, N, 2 * Cc + 2, H, W), dtype=np.int8) if dump_decisions else None
    margin = np.zeros((Ls, N, H, W), dtype=np.float32) if dump_decisions else None
    L.mdbo_set_decisions(_ptr(force), _ptr(dump), _ptr(margin))
    try:
        rc = L.mdbo_train_loss(C.byref(d), _ptr(x), _ptr(rv), _ptr(tv), disp_ptrs, _ptr(au), _ptr(Kc), _ptr(iKc),
                               C.byref(loss), g_ptrs, _ptr(g_rv), _ptr(g_tv), C.c_int(threads))
    finally:
        L.mdbo_set_decisions(None, None, None)
    if rc != 0:
        raise RuntimeError(f"mdbo_train_loss failed ({rc})")
    out = dict(loss=loss.value)
    if dump_decisions:
        out.update(decisions=dump, margin=margin)
    if grad:
        out.update(d_disp=g_disp, d_rvec=[g_rv[s] for s in range(S)], d_tvec=[g_tv[s] for s in range(S)])
    return out


def so3_exp_map(rvec: np.ndarray) -> np.ndarray:
    """(N,3) -> (N,3,3) in Julia layout ([n, j, i]), like torch_ref.so3_exp_map."""
    L = lib(np.float64)
    r = np.ascontiguousarray(rvec, dtype=np.float64)
    out = np.zeros((r.shape[0], 3, 3))
    L.mdbo_so3_exp_map(_ptr(r), C.c_int(r.shape[0]), _ptr(out))
    return out


def compose_T(rvec: np.ndarray, tvec: np.ndarray, invert: bool):
    L = lib(np.float64)
    r = np.ascontiguousarray(rvec, dtype=np.float64)
    t = np.ascontiguousarray(np.asarray(tvec, dtype=np.float64).reshape(-1, 3))
    R = np.zeros((r.shape[0], 3, 3))
    to = np.zeros_like(t)
    L.mdbo_compose_T(_ptr(r), _ptr(t), C.c_int(int(invert)), C.c_int(r.shape[0]), _ptr(R), _ptr(to))
    return R, to.reshape(-1, 1, 3)


def compose_T_bwd(rvec, tvec, invert, dR, dt):
    L = lib(np.float64)
    r = np.ascontiguousarray(rvec, dtype=np.float64)
    t = np.ascontiguousarray(np.asarray(tvec, dtype=np.float64).reshape(-1, 3))
    dR = np.ascontiguousarray(dR, dtype=np.float64)
    dt = np.ascontiguousarray(np.asarray(dt, dtype=np.float64).reshape(-1, 3))
    gr, gt = np.zeros_like(r), np.zeros_like(t)
    L.mdbo_compose_T_bwd(_ptr(r), _ptr(t), C.c_int(int(invert)), C.c_int(r.shape[0]), _ptr(dR), _ptr(dt), _ptr(gr), _ptr(gt))
    return gr, gt


RENDER_MINE_GRID, RENDER_MINE_MASK, RENDER_FUSED_XYZ = 1, 2, 4


def render_tgt_rgb_depth(rgb, sigma, disparity_src, xyz_tgt, rvec, tvec, K_inv, K, *, g_rgb=None, g_depth=None,
                         g_expected_depth=None, mine_grid=False, mine_mask=False, depth_eps=1e-5,
                         dtype=np.float64, threads: int = 1):
    """`render_tgt_rgb_depth` (src/render.jl:92-114) forward and, when `g_rgb` is given, the hand-written
    pullback of  sum(g_rgb * rgb_out) + sum(g_depth * depth_out) + sum(g_expected_depth * expected_depth).
    Arrays in the repo's reversed-dims convention: rgb (B,Np,3,H,W), sigma (B,Np,1,H,W), disparity_src (B,Np),
    xyz_tgt (B,Np,H,W,3) or None (derived from the producers, src/render.jl:21-30, 51-62), rvec / tvec (B,3).
    Returns dict(rgb, depth, mask, expected_depth[, d_rgb, d_sigma, d_disparity_src, d_xyz_tgt, d_rvec, d_tvec])."""
    L = lib(dtype)
    L.mdbo_render_tgt_rgb_depth_ex.restype = C.c_int
    c = lambda a: None if a is None else np.ascontiguousarray(a, dtype=dtype)
    rgb, sigma, disparity_src, xyz_tgt, rvec, tvec = map(c, (rgb, sigma, disparity_src, xyz_tgt, rvec, tvec))
    B, Np, _, H, W = rgb.shape
    fused = xyz_tgt is None
    flags = (RENDER_MINE_GRID if mine_grid else 0) | (RENDER_MINE_MASK if mine_mask else 0) | (RENDER_FUSED_XYZ if fused else 0)
    Ki, Kk = c(np.asarray(K_inv).reshape(9)), c(np.asarray(K).reshape(9))
    out = np.zeros((B, 3, H, W), dtype=dtype)
    depth = np.zeros((B, Np, 1, H, W), dtype=dtype)
    mask = np.zeros((B, 1, 1, H, W), dtype=np.int32)
    dsyn = np.zeros((B, 1, H, W), dtype=dtype)
    g_rgb, g_depth, g_expected_depth = c(g_rgb), c(g_depth), c(g_expected_depth)
    grads = [None] * 6
    if g_rgb is not None:
        grads = [np.zeros_like(rgb), np.zeros_like(sigma), np.zeros_like(disparity_src),
                 None if fused else np.zeros_like(xyz_tgt), np.zeros_like(rvec), np.zeros_like(tvec)]
    rc = L.mdbo_render_tgt_rgb_depth_ex(C.c_int(W), C.c_int(H), C.c_int(Np), C.c_int(B), _ptr(rgb), _ptr(sigma), _ptr(disparity_src),
                                        _ptr(xyz_tgt), _ptr(rvec), _ptr(tvec), _ptr(Ki), _ptr(Kk), C.c_int(flags),
                                        C.c_double(depth_eps), _ptr(out), _ptr(depth), mask.ctypes.data_as(C.c_void_p), _ptr(dsyn),
                                        _ptr(g_rgb), _ptr(g_depth), _ptr(g_expected_depth), _ptr(grads[0]), _ptr(grads[1]),
                                        _ptr(grads[2]), _ptr(grads[3]), _ptr(grads[4]), _ptr(grads[5]), C.c_int(threads))
    if rc != 0:
        raise RuntimeError(f"mdbo_render_tgt_rgb_depth_ex failed ({rc})")
    res = dict(rgb=out, depth=depth, mask=mask, expected_depth=dsyn)
    if g_rgb is not None:
        res.update(d_rgb=grads[0], d_sigma=grads[1], d_disparity_src=grads[2], d_xyz_tgt=grads[3], d_rvec=grads[4], d_tvec=grads[5])
    return res


def plane_xyz(W, H, disparity_src, rvec, tvec, K_inv, dtype=np.float64):
    """create_meshgrid -> get_src_xyz_from_plane_disparity -> get_tgt_xyz_from_plane_disparity
    (src/render.jl:21-30, 51-62): returns (xyz_src, xyz_tgt), each (B,Np,H,W,3)."""
    L = lib(dtype)
    c = lambda a: np.ascontiguousarray(a, dtype=dtype)
    disparity_src, rvec, tvec = c(disparity_src), c(rvec), c(tvec)
    B, Np = disparity_src.shape
    Ki = c(np.asarray(K_inv).reshape(9))
    xs, xt = np.zeros((B, Np, H, W, 3), dtype=dtype), np.zeros((B, Np, H, W, 3), dtype=dtype)
    rc = L.mdbo_plane_xyz(C.c_int(W), C.c_int(H), C.c_int(Np), C.c_int(B), _ptr(disparity_src), _ptr(rvec), _ptr(tvec), _ptr(Ki),
                          _ptr(xs), _ptr(xt))
    if rc != 0:
        raise RuntimeError("mdbo_plane_xyz failed")
    return xs, xt
