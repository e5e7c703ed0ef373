"""Debug aid: render forward through the short (float) path and through the double path on the same inputs; prints where they differ."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch, mdb200
from mdb200 import synth
DEV = "cuda:0"
d = lambda a: torch.tensor(a, device=DEV).contiguous()
for (W, H, Np, B, seed, ps, mine) in [(416, 128, 16, 2, 31, 1.0, True), (416, 128, 16, 2, 31, 1.0, False), (96, 48, 12, 2, 21, 3.0, False)]:
    ri = synth.make_render_inputs(W, H, Np, B, seed=seed, pose_scale=ps)
    rng = np.random.default_rng(seed)
    ri.xyz_tgt = rng.random(ri.xyz_tgt.shape).astype(np.float32)
    plan = mdb200.RenderPlan(W, H, Np, B, ri.K_inv, ri.K, mine_grid=mine, mine_mask=mine)
    args = (d(ri.rgb), d(ri.sigma), d(ri.disparity_src), d(ri.xyz_tgt), mdb200.Pose(d(ri.rvec), d(ri.tvec)))
    a = [t.clone() for t in plan.forward(*args)[:3]]
    mdb200.debug_option("render_precise", 1)
    b = [t.clone() for t in plan.forward(*args)[:3]]
    mdb200.debug_option("render_precise", 0)
    torch.cuda.synchronize()
    dr = (a[0] - b[0]).abs(); dt = (a[1] - b[1]).abs()
    print(W, H, Np, ps, mine, "rgb max", float(dr.max()), "tacc max", float(dt.max()), "mask equal", bool(torch.equal(a[2], b[2])),
          "n rgb > 1e-5:", int((dr > 1e-5).sum()), "n tacc > 1e-5:", int((dt > 1e-5).sum()))
    idx = torch.nonzero(dt > 1e-5)[:8]
    print("  tacc idx", idx.tolist())
    idx = torch.nonzero(dr > 1e-5)[:8]
    print("  rgb idx", idx.tolist())
