"""numpy restatement of the reference's sample ingestion (TEST INFRASTRUCTURE ONLY -- see
oracle/torch_ref.py for who may import this).

`Depth10k.getindex` -- src/dtk.jl:29-46: one decoded image holds the three frames side by side;
`frames[:, (width*j+1):(width*(j+1))]` splits it, `augmentations(frames)` (FlipX, Augmentations.jl --
third-party, absent: restated as "reverse the width axis of every frame"), `channelview`, `cat(dims=4)`,
`permutedims((3,2,1,4))`, `Float32.`.  `Gray{Float32}.(frames)` (:33-35) is Colors.jl's Rec.601 luma
`0.299f0 r + 0.587f0 g + 0.114f0 b` (third-party, unpinned), evaluated left to right in float32 without FMA.
`Float32(::N0f8)` is `u / 255f0` correctly rounded (FixedPointNumbers).  Parity unpinned: the reference has
no test for its datasets."""
import numpy as np


def n0f8_to_f32(u8: np.ndarray) -> np.ndarray:
    return u8.astype(np.float32) / np.float32(255.0)


def depth10k_getindex(strip: np.ndarray, width: int, *, flip: bool = False, grayscale: bool = False) -> np.ndarray:
    """strip: uint8 (H, 3*width, 3) decoded raster -> float32 (3, C, H, W), i.e. Julia (W,H,C,3)."""
    H = strip.shape[0]
    img = n0f8_to_f32(strip)                                             # load(): Matrix{RGB{N0f8}}
    if grayscale:                                                        # :33-35
        r, g, b = img[..., 0], img[..., 1], img[..., 2]
        img = ((np.float32(0.299) * r + np.float32(0.587) * g) + np.float32(0.114) * b)[..., None]
    frames = [img[:, width * j:width * (j + 1)] for j in range(3)]       # :36
    if flip:                                                             # :37-39 FlipX
        frames = [f[:, ::-1] for f in frames]
    out = np.stack([np.transpose(f, (2, 0, 1)) for f in frames], axis=0)  # channelview + cat + permutedims
    assert out.shape == (3, 1 if grayscale else 3, H, width)
    return np.ascontiguousarray(out, dtype=np.float32)


def batch(strips: np.ndarray, width: int, flips=None, grayscale: bool = False) -> np.ndarray:
    """strips (N,H,3W,3) uint8 -> (N,3,C,H,W) float32 (the `DataLoader` batch, Flux.unsqueeze + cat)."""
    n = strips.shape[0]
    return np.stack([depth10k_getindex(strips[i], width, flip=bool(flips[i]) if flips is not None else False,
                                       grayscale=grayscale) for i in range(n)], axis=0)


# ---- KittyDataset.getindex -- src/kitty.jl:45-61 ------------------------------------------------------------------------
# `load` of an `image_0` PNG gives Matrix{Gray{N0f8}}; `imresize(x, (height, width))` is ImageTransformations.jl
# (third-party, not vendored, no [compat] pin in Project.toml: restated from its published algorithm, v0.9 / v0.10
# `imresize!(resized, interpolate(original, BSpline(Linear())))`):
#   * per axis, scale s = m / n (original / new length) and offset = 0.5 - 0.5 s: output index i (1-based) samples the
#     original at  p = s i + offset  ("array of boxes": half-pixel centres), clamped to [1, m];
#   * linear B-spline: f = floor(p), lowered by one where f > m - 1, d = p - f, weights (1 - d, d);
#   * tensor product, first axis (rows) outermost:  wy0 (wx0 a00 + wx1 a01) + wy1 (wx0 a10 + wx1 a11)  in Float64,
#     with Float64(::N0f8) = u / 255;
#   * the result array keeps the element type Gray{N0f8}: N0f8(v) = round(255 v), ties to even (FixedPointNumbers).
# Then `Float32.(channelview(x))`, `permutedims((2, 1))`, `Flux.unsqueeze(3)`, `cat(dims = 4)`: (W, H, 1, 3).
# Parity unpinned (no Julia here, package version not fixed by the reference); exact .5 ties of `round` -- which only integer
# scale factors produce in numbers -- follow this restatement's operation order.

def _axis_taps(n_new: int, m_old: int):
    s = np.float64(m_old) / np.float64(n_new)
    off = np.float64(0.5) - np.float64(0.5) * s
    i = np.arange(1, n_new + 1, dtype=np.float64)
    p = np.clip(s * i + off, 1.0, np.float64(m_old))
    f = np.floor(p)
    f = f - (f > m_old - 1)
    d = p - f
    i0 = f.astype(np.int64) - 1                      # 0-based
    return i0, np.minimum(i0 + 1, m_old - 1), np.float64(1.0) - d, d


def imresize_gray_n0f8(img: np.ndarray, height: int, width: int) -> np.ndarray:
    """uint8 (Ho, Wo) -> uint8 (height, width): `imresize(::Matrix{Gray{N0f8}}, (height, width))`."""
    Ho, Wo = img.shape
    if (Ho, Wo) == (height, width):
        return img.copy()
    a = img.astype(np.float64) / np.float64(255.0)
    y0, y1, wy0, wy1 = _axis_taps(height, Ho)
    x0, x1, wx0, wx1 = _axis_taps(width, Wo)
    top = wx0[None, :] * a[y0][:, x0] + wx1[None, :] * a[y0][:, x1]
    bot = wx0[None, :] * a[y1][:, x0] + wx1[None, :] * a[y1][:, x1]
    v = wy0[:, None] * top + wy1[:, None] * bot
    return np.rint(np.float64(255.0) * v).astype(np.uint8)


def kitti_getindex(frames: np.ndarray, height: int, width: int, *, flip: bool = False) -> np.ndarray:
    """frames uint8 (3, Ho, Wo): the three consecutive `image_0` rasters of a sample -> float32 (3, 1, H, W),
    i.e. Julia (W, H, 1, 3).  `flip`: FlipX from `d.augmentations` (:53-55), applied to the resized images."""
    out = []
    for j in range(3):
        r = imresize_gray_n0f8(frames[j], height, width)
        if flip:
            r = r[:, ::-1]
        out.append(n0f8_to_f32(r)[None])
    return np.ascontiguousarray(np.stack(out, axis=0), dtype=np.float32)


def kitti_batch(frames: np.ndarray, height: int, width: int, flips=None) -> np.ndarray:
    """frames (N, 3, Ho, Wo) uint8 -> (N, 3, 1, H, W) float32."""
    return np.stack([kitti_getindex(frames[i], height, width, flip=bool(flips[i]) if flips is not None else False)
                     for i in range(frames.shape[0])], axis=0)
