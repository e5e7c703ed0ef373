"""Diagnostic (not a test): where does the CUDA gradient leave the 1e-4 band at a full-size config, and
is each such place a near-tie of a discrete decision (arg-min over sources, sign of t - w, SSIM clamp)?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
import mdb200
from mdb200 import synth
from oracle import c_ref, torch_ref as T

name, N = sys.argv[1], int(sys.argv[2])
inp = synth.make_config(name, N=N)
ref = c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, None, inp.K, inp.invK, scales=inp.scales,
                       automasking=False, threads=8)
dev = torch.device("cuda:0")
x = torch.tensor(inp.x, device=dev)
disps = [torch.tensor(d, device=dev) for d in inp.disparities]
poses = [mdb200.Pose(torch.tensor(r, device=dev), torch.tensor(t, device=dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
plan = mdb200.FusedLoss(W=inp.W, H=inp.H, C=3, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in disps],
                        scales=inp.scales, target_id=2, source_ids=[1, 3], K=inp.K, invK=inp.invK, automasking=False)
res = plan(x, disps, poses, None)
torch.cuda.synchronize()
for s in range(2):
    g, r = res.d_poses[s].rvec.double().cpu().numpy(), ref["d_rvec"][s]
    e = np.abs(g - r) / np.abs(r).max()
    print("d_rvec", s, "err per item", np.round(e.max(axis=1) * 1e4, 2))
# per-scale outliers and the decision margins at those pixels (torch oracle parts, float64)
x64 = torch.tensor(inp.x, dtype=torch.float64)
for k in range(4):
    g, r = res.d_disparities[k].double().cpu().numpy(), ref["d_disp"][k]
    e = np.abs(g - r) / np.abs(r).max()
    idx = np.argwhere(e > 1e-4)
    print(f"scale {k}: {len(idx)} outliers of {e.size}; max {e.max():.2e}")
    items = sorted(set(int(i[0]) for i in idx))
    for n in items[:4]:
        d = [torch.tensor(a[n:n + 1], dtype=torch.float64) for a in inp.disparities]
        P_ = T.Params(automasking=False, target_size=(inp.W, inp.H))
        ps = [T.Pose(torch.tensor(rr[n:n + 1], dtype=torch.float64), torch.tensor(tt[n:n + 1], dtype=torch.float64))
              for rr, tt in zip(inp.rvecs, inp.tvecs)]
        _, parts = T.train_loss(x64[n:n + 1], d, ps, None, torch.tensor(inp.K.astype(np.float64)),
                                torch.tensor(inp.invK.astype(np.float64)), P_, scales=inp.scales, return_parts=True)
        warped = parts[k][2]
        tgt = x64[n:n + 1, 1]
        pl = torch.cat([T.photometric_loss(w, tgt) for w in warped], 1)[0]        # (S,H,W)
        margin = (pl[0] - pl[1]).abs().numpy()
        l1m = torch.stack([(tgt - w).abs().amin(dim=1)[0] for w in warped]).amin(0).numpy()   # min |t-w| over c, s
        f = inp.W // g.shape[-1]
        for i in [i for i in idx if i[0] == n][:6]:
            yy, xx = int(i[2]) * f, int(i[3]) * f
            sl = (slice(max(yy - 2, 0), yy + f + 2), slice(max(xx - 2, 0), xx + f + 2))
            print(f"   item {n} px ({i[2]},{i[3]}) err {e[tuple(i)]:.2e}: min |pl0-pl1| nearby {margin[sl].min():.2e}, "
                  f"min |t-w| nearby {l1m[sl].min():.2e}")
