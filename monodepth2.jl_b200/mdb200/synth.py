"""Synthetic inputs for the view-synthesis loss path.

Everything is generated with ``numpy.random.Generator(PCG64(seed))`` so the GPU path, the
CPU oracle and (where present) a Julia run read identical bytes.  Arrays are returned as
float32 numpy arrays in the reversed-dims convention: a Julia ``(W,H,C,L,N)`` array is a
C-contiguous ``(N,L,C,H,W)`` array here (same memory).

Shapes follow the reference's data path: frame stacks ``(W,H,C,3,N)`` with
``target_id=2, source_ids=[1,3]`` (src/dtk.jl:15-26), disparities in the sigmoid range
(src/depth_decoder.jl:46) at ``scales=[1/8,1/4,1/2,1]`` (scripts/script.jl:61-62),
pose magnitudes O(1e-2) (src/pose_decoder.jl:30), intrinsics ``f = 2648/4.63461538462``
at 416 px width, ``cx=W/2, cy=H/2`` (src/dtk.jl:16-21).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

FOCAL_416 = 2648.0 / 4.63461538462  # src/dtk.jl:16


@dataclass
class LossInputs:
    W: int
    H: int
    C: int
    N: int
    x: np.ndarray                       # (N,3,C,H,W)
    disparities: List[np.ndarray]       # [(N,1,h_k,w_k)]
    scales: List[float]
    rvecs: List[np.ndarray]             # per source (N,3)
    tvecs: List[np.ndarray]             # per source (N,1,3)
    K: np.ndarray                       # (3,3) Julia layout (= K_math.T), float32
    invK: np.ndarray
    target_id: int = 2
    source_ids: List[int] = field(default_factory=lambda: [1, 3])
    automasking: bool = False
    auto_loss: Optional[np.ndarray] = None  # (N,1,H,W), filled by the caller (oracle / GPU)


def intrinsics(W: int, H: int):
    """K, invK in Julia layout (column-major 3x3 == K_math.T), float64."""
    f = FOCAL_416 * W / 416.0
    K_m = np.array([[f, 0, W / 2.0], [0, f, H / 2.0], [0, 0, 1.0]], dtype=np.float64)
    invK_m = np.linalg.inv(K_m)
    return K_m.T.copy(), invK_m.T.copy()


def _box3(a: np.ndarray) -> np.ndarray:
    """3x3 mean with edge replication over the last two axes."""
    p = np.pad(a, [(0, 0)] * (a.ndim - 2) + [(1, 1), (1, 1)], mode="edge")
    h, w = a.shape[-2:]
    out = np.zeros_like(a)
    for dy in range(3):
        for dx in range(3):
            out += p[..., dy:dy + h, dx:dx + w]
    return out / 9.0


def make_images(W: int, H: int, C: int, N: int, rng: np.random.Generator) -> np.ndarray:
    """Frame stacks (N,3,C,H,W): a smooth multi-frequency texture, the two source frames
    displaced by a sub-pixel affine (about +-2 px), plus low-passed noise shared by the
    frames (shifted by whole pixels) and a little independent noise per frame."""
    w = np.arange(1, W + 1, dtype=np.float32)[None, None, None, None, :]
    h = np.arange(1, H + 1, dtype=np.float32)[None, None, None, :, None]
    c = np.arange(C, dtype=np.float32)[None, None, :, None, None]
    n = np.arange(N, dtype=np.float32)[:, None, None, None, None]
    # per-frame displacement: frame 1 and 3 are the sources, frame 2 the target
    dx = np.array([-1.7, 0.0, 1.3], dtype=np.float32)[None, :, None, None, None]
    dy = np.array([0.6, 0.0, -0.4], dtype=np.float32)[None, :, None, None, None]
    sx = np.array([1.004, 1.0, 0.997], dtype=np.float32)[None, :, None, None, None]
    u = (w * sx + dx) / np.float32(W)
    v = (h + dy) / np.float32(H)
    two_pi = np.float32(2 * np.pi)
    img = (0.5
           + 0.22 * np.sin(two_pi * (3 * u + 0.7 * c + 0.13 * n)) * np.cos(two_pi * (2 * v + 0.3 * n))
           + 0.08 * np.sin(two_pi * (17 * u + 5 * v + 0.31 * c))
           + 0.05 * np.cos(two_pi * (41 * u - 9 * v + 0.11 * n)))
    img = np.broadcast_to(img, (N, 3, C, H, W)).astype(np.float32)
    base = _box3(rng.uniform(-1.0, 1.0, size=(N, 1, C, H, W)).astype(np.float32)) * np.float32(0.3)
    shared = np.concatenate([np.roll(base, 2, axis=-1), base, np.roll(base, -1, axis=-1)], axis=1)
    indep = _box3(rng.uniform(-1.0, 1.0, size=(N, 3, C, H, W)).astype(np.float32)) * np.float32(0.06)
    return np.clip(img + shared + indep, 0.0, 1.0).astype(np.float32)


def make_disparities(W: int, H: int, N: int, scales, rng: np.random.Generator, noise: float = 0.05):
    out = []
    for s in scales:
        ws, hs = max(int(round(W * s)), 2), max(int(round(H * s)), 2)
        w = np.arange(1, ws + 1, dtype=np.float32)[None, None, None, :] / np.float32(ws)
        h = np.arange(1, hs + 1, dtype=np.float32)[None, None, :, None] / np.float32(hs)
        d = (0.3 + 0.2 * np.sin(2 * np.pi * w) * np.cos(2 * np.pi * h)
             + noise * rng.uniform(-1, 1, size=(N, 1, hs, ws)))
        out.append(np.clip(d, 0.02, 0.98).astype(np.float32))
    return out


def make_poses(N: int, S: int, rng: np.random.Generator):
    rvecs, tvecs = [], []
    for _ in range(S):
        r = 0.01 * rng.standard_normal((N, 3))
        small = np.linalg.norm(r, axis=1) < 1e-3  # stay away from the theta<1e-4 branch
        r[small] += 2e-3
        rvecs.append(r.astype(np.float32))
        tvecs.append((0.05 * rng.standard_normal((N, 1, 3))).astype(np.float32))
    return rvecs, tvecs


def make_loss_inputs(W: int, H: int, N: int, *, C: int = 3, seed: int = 2,
                     scales=(0.125, 0.25, 0.5, 1.0), automasking: bool = False, disp_noise: float = 0.05) -> LossInputs:
    """`disp_noise`: amplitude of the per-pixel uniform noise on the disparities (SURVEY 8d's inputs use 0.05, i.e.
    +-35 px of parallax jitter at config 5's focal length -- a worst case for the bilinear gathers)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    x = make_images(W, H, C, N, rng)
    disps = make_disparities(W, H, N, scales, rng, noise=disp_noise)
    rvecs, tvecs = make_poses(N, 2, rng)
    K, invK = intrinsics(W, H)
    return LossInputs(W=W, H=H, C=C, N=N, x=x, disparities=disps, scales=list(scales),
                      rvecs=rvecs, tvecs=tvecs, K=K.astype(np.float32),
                      invK=invK.astype(np.float32), automasking=automasking)


# Named configurations of BASELINE.json ("configs"), sizes from SURVEY 8.
CONFIGS = {
    "config1": dict(W=416, H=128, N=1, scales=(1.0,), automasking=False, seed=1),
    "config2": dict(W=416, H=128, N=12, scales=(0.125, 0.25, 0.5, 1.0), automasking=False, seed=2),
    "config3": dict(W=640, H=192, N=32, scales=(0.125, 0.25, 0.5, 1.0), automasking=True, seed=3),
    "config5": dict(W=1024, H=320, N=64, scales=(0.125, 0.25, 0.5, 1.0), automasking=False, seed=5),
}


def make_config(name: str, **over) -> LossInputs:
    kw = dict(CONFIGS[name]); kw.update(over)
    return make_loss_inputs(**kw)


@dataclass
class RenderInputs:
    W: int
    H: int
    Np: int
    B: int
    rgb: np.ndarray            # (B,Np,3,H,W)
    sigma: np.ndarray          # (B,Np,1,H,W)
    disparity_src: np.ndarray  # (B,Np)
    xyz_tgt: np.ndarray        # (B,Np,H,W,3)
    rvec: np.ndarray           # (B,3)
    tvec: np.ndarray           # (B,3)
    K: np.ndarray
    K_inv: np.ndarray


def make_render_inputs(W: int = 416, H: int = 128, Np: int = 32, B: int = 2, seed: int = 4,
                       pose_scale: float = 1.0) -> RenderInputs:
    """Config 4: rgb, sigma ~ U(0,1); disparity bins from 1 to 0.001 as
    src/model.jl:17-21; xyz_tgt = R4(R3(meshgrid)) computed in float64 and cast."""
    rng = np.random.Generator(np.random.PCG64(seed))
    rgb = rng.uniform(0, 1, size=(B, Np, 3, H, W)).astype(np.float32)
    sigma = rng.uniform(0, 1, size=(B, Np, 1, H, W)).astype(np.float32)
    edges = np.linspace(1.0, 0.001, Np + 1)
    disp = (edges[:-1] + (edges[1:] - edges[:-1]) * rng.uniform(0, 1, size=(B, Np))).astype(np.float32)
    rvec = (0.01 * pose_scale * rng.standard_normal((B, 3))).astype(np.float32)
    tvec = (0.05 * pose_scale * rng.standard_normal((B, 3))).astype(np.float32)
    K, K_inv = intrinsics(W, H)
    # xyz_src = (K_inv [w,h,1]) / disparity ; xyz_tgt = R xyz_src + t   (src/render.jl:25-30,51-62)
    ww, hh = np.meshgrid(np.arange(1, W + 1, dtype=np.float64), np.arange(1, H + 1, dtype=np.float64))
    mesh = np.stack([ww, hh, np.ones_like(ww)], axis=-1)  # (H,W,3)
    rays = mesh @ K_inv  # K_inv is Julia layout = K_inv_math.T, so row-vector @ K_inv == K_inv_m @ v
    xyz_src = rays[None, None] / disp.astype(np.float64)[:, :, None, None, None]
    from scipy.spatial.transform import Rotation  # host-side input generation only
    R = Rotation.from_rotvec(rvec.astype(np.float64)).as_matrix()  # (B,3,3)
    xyz_tgt = np.einsum("bij,bnhwj->bnhwi", R, xyz_src) + tvec.astype(np.float64)[:, None, None, None, :]
    return RenderInputs(W=W, H=H, Np=Np, B=B, rgb=rgb, sigma=sigma, disparity_src=disp,
                        xyz_tgt=xyz_tgt.astype(np.float32), rvec=rvec, tvec=tvec,
                        K=K.astype(np.float32), K_inv=K_inv.astype(np.float32))


def to_u8_strips(x: np.ndarray) -> np.ndarray:
    """Float frame stacks (N,3,3,H,W) in [0,1] -> the dataset's own format: uint8 decoded triplet strips
    (N,H,3W,3), the three frames side by side (src/dtk.jl:29-36).  Quantises to 8 bits."""
    n, l, c, h, w = x.shape
    assert l == 3 and c == 3
    q = np.clip(np.rint(x * 255.0), 0, 255).astype(np.uint8)          # (N,L,C,H,W)
    return np.ascontiguousarray(q.transpose(0, 3, 1, 4, 2).reshape(n, h, 3 * w, 3))


# --------------------------------------------------------------------------------------
# on-disk form: raw little-endian float32 in Julia column-major order + a JSON sidecar,
# so that a machine WITH Julia can feed the reference itself the same bytes
# (monodepth2.jl_b200/julia/run_reference_on_dump.jl) and pin the parity this box cannot.
# --------------------------------------------------------------------------------------
def dump_loss_inputs(inp: LossInputs, out_dir: str) -> str:
    """Writes x.f32, disp<k>.f32, rvec<s>.f32, tvec<s>.f32, K.f32, invK.f32 and inputs.json.  A C-contiguous
    (N,L,C,H,W) array IS the Julia (W,H,C,L,N) array, so the bytes are written as they are."""
    import json
    import os
    os.makedirs(out_dir, exist_ok=True)

    def put(name, a):
        np.ascontiguousarray(a, dtype="<f4").tofile(os.path.join(out_dir, name))

    put("x.f32", inp.x)
    meta = {"julia_shapes": {"x": [inp.W, inp.H, inp.C, 3, inp.N]}, "dtype": "float32 little-endian, Julia column-major",
            "target_id": inp.target_id, "source_ids": list(inp.source_ids), "scales": [float(s) for s in inp.scales],
            "automasking": bool(inp.automasking), "min_depth": 0.1, "max_depth": 100.0, "disparity_smoothness": 1e-3}
    for k, d in enumerate(inp.disparities):
        put(f"disp{k}.f32", d)
        meta["julia_shapes"][f"disp{k}"] = [d.shape[-1], d.shape[-2], 1, inp.N]
    for s, (r, t) in enumerate(zip(inp.rvecs, inp.tvecs)):
        put(f"rvec{s}.f32", r); put(f"tvec{s}.f32", t)
        meta["julia_shapes"][f"rvec{s}"] = [3, inp.N]
        meta["julia_shapes"][f"tvec{s}"] = [3, 1, inp.N]
    put("K.f32", inp.K); put("invK.f32", inp.invK)
    meta["julia_shapes"]["K"] = [3, 3]; meta["julia_shapes"]["invK"] = [3, 3]
    path = os.path.join(out_dir, "inputs.json")
    with open(path, "w") as f:
        json.dump(meta, f, indent=1)
    return path


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser(description="dump a BASELINE configuration's synthetic inputs for a Julia run of the reference")
    ap.add_argument("config", choices=sorted(CONFIGS))
    ap.add_argument("out_dir")
    ap.add_argument("--batch", type=int, default=0)
    a = ap.parse_args()
    print(dump_loss_inputs(make_config(a.config, **({"N": a.batch} if a.batch else {})), a.out_dir))
