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
