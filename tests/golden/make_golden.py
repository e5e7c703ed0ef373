"""Generates the golden fixtures of this directory.  Run from the repo root:

    python tests/golden/make_golden.py

Provenance.  The reference is Julia and cannot run here or on the GPU box, so NONE of these files
are outputs of the reference itself.  There are two kinds:

* ``runtests_known_answers.json`` -- the known answers the reference's own test file holds for the
  path (/root/reference/test/runtests.jl, cited per entry).  These pin the oracle.
* ``*.npz`` -- inputs (mdb200.synth, fixed seeds) and outputs of the float64 torch oracle
  (oracle/torch_ref.py).  They pin the ORACLE's behaviour over time (both restatements must keep
  reproducing them) and give the CUDA path a fixed target that does not depend on the oracle code
  being importable.  Parity against the reference itself stays "unpinned" where the header of
  oracle/torch_ref.py says so.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from mdb200 import synth  # noqa: E402
from oracle import torch_ref as T  # noqa: E402
import parity as P  # noqa: E402

DT = torch.float64


def loss_case(name, W, H, N, seed, auto, scales=(0.125, 0.25, 0.5, 1.0), simple=False):
    inp = synth.make_loss_inputs(W, H, N, seed=seed, automasking=auto, scales=scales)
    ref = P.oracle_train_loss(inp, normalize=not simple, smooth_mode=1 if simple else 0)
    out = dict(x=inp.x, K=inp.K, invK=inp.invK, scales=np.asarray(inp.scales, dtype=np.float64),
               source_ids=np.asarray(inp.source_ids), target_id=np.asarray(inp.target_id),
               automasking=np.asarray(int(auto)), simple=np.asarray(int(simple)), loss=np.asarray(ref["loss"]))
    for k, d in enumerate(inp.disparities):
        out[f"disp{k}"] = d
        out[f"d_disp{k}"] = ref["d_disp"][k].numpy()
    for s in range(2):
        out[f"rvec{s}"], out[f"tvec{s}"] = inp.rvecs[s], inp.tvecs[s]
        out[f"d_rvec{s}"], out[f"d_tvec{s}"] = ref["d_rvec"][s].numpy(), ref["d_tvec"][s].numpy()
    if auto:
        out["auto_loss"] = ref["auto"].numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", ref["loss"])


def render_case(name, W, H, Np, B, seed):
    ri = synth.make_render_inputs(W, H, Np, B, seed=seed, pose_scale=2.0)
    f = lambda a: torch.tensor(a, dtype=DT, requires_grad=True)
    rgb, sig, dsp, xyz, rv, tv = f(ri.rgb), f(ri.sigma), f(ri.disparity_src), f(ri.xyz_tgt), f(ri.rvec), f(ri.tvec)
    ro, do, mo = T.render_tgt_rgb_depth(rgb, sig, dsp, xyz, T.Pose(rv, tv), torch.tensor(ri.K_inv.astype(np.float64)),
                                        torch.tensor(ri.K.astype(np.float64)))
    g = torch.Generator().manual_seed(seed)
    wr, wd = torch.rand(ro.shape, generator=g, dtype=DT), torch.rand(do.shape, generator=g, dtype=DT)
    gs = torch.autograd.grad((ro * wr).sum() + (do * wd).sum(), (rgb, sig, dsp, xyz, rv, tv))
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), rgb=ri.rgb, sigma=ri.sigma, disparity_src=ri.disparity_src, xyz_tgt=ri.xyz_tgt,
        rvec=ri.rvec, tvec=ri.tvec, K=ri.K, K_inv=ri.K_inv, rgb_out=ro.detach().numpy(), depth_out=do.detach().numpy(),
        mask_out=mo.numpy(), w_rgb=wr.numpy(), w_depth=wd.numpy(),
        **{"d_" + n: t.numpy() for n, t in zip(["rgb", "sigma", "disparity_src", "xyz_tgt", "rvec", "tvec"], gs)})
    print(name, "rgb mean", float(ro.mean()))


def known_answers():
    # runtests.jl:124-142 asserts nothing; the oracle's values are recorded as new goldens
    r = torch.tensor([[1.0, 0.0, 0.0]], dtype=DT, requires_grad=True)
    t = torch.zeros(3, dtype=DT, requires_grad=True)
    x, target = torch.tensor([3.0, 2.0, 1.0], dtype=DT), torch.tensor([1.0, 2.0, 3.0], dtype=DT)
    l = torch.sqrt(((T.so3_exp_map_m(r)[0] @ x + t - target) ** 2).sum())
    gr, gt = torch.autograd.grad(l, (r, t))
    ka = {
        "smooth_loss": {"cite": "test/runtests.jl:70-83", "disp_column_major": [0.0, 0.2, 0.1, 0.3],
                        "image_column_major": [0.1, 0.3, 0.2, 0.4], "value": 0.2542, "atol": 1e-4},
        "ssim": {"cite": "test/runtests.jl:52-68", "ones_vs_ones": 0.0, "ones_vs_zeros": 0.5, "atol_ones_vs_zeros": 0.1,
                 "symmetric": True},
        "depth_range": {"cite": "test/runtests.jl:85-92", "min_depth": 0.1, "max_depth": 100.0},
        "so3_vs_rotation_vector": {"cite": "test/runtests.jl:14-29", "atol": 1e-5},
        "composeT_round_trip": {"cite": "test/runtests.jl:31-50", "atol": 1e-6},
        "identity_warp": {"cite": "test/runtests.jl:94-122", "res": 16, "focal": 910.0, "centre": 8.0, "atol": 1e-3},
        "pose_derivative": {"cite": "test/runtests.jl:124-142 (prints, asserts nothing): oracle values recorded",
                            "x": [3, 2, 1], "target": [1, 2, 3], "r": [1, 0, 0], "t": [0, 0, 0],
                            "loss": float(l.detach()), "d_r": gr[0].tolist(), "d_t": gt.tolist()},
    }
    with open(os.path.join(HERE, "runtests_known_answers.json"), "w") as f:
        json.dump(ka, f, indent=1)


if __name__ == "__main__":
    known_answers()
    loss_case("loss_48x20_b2_automask", 48, 20, 2, 21, True)
    loss_case("loss_40x24_b1", 40, 24, 1, 22, False)
    loss_case("simple_depth_32x16", 32, 16, 1, 23, False, scales=(1.0,), simple=True)
    render_case("render_24x16_p6_b2", 24, 16, 6, 2, 24)
