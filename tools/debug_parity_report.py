"""Scratch driver used during bring-up: prints the parity report instead of asserting."""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")]
import torch
import parity as P
from mdb200 import synth
import test_gpu_loss as TL

for (W, H, N, C, auto) in [(104, 32, 2, 3, False), (104, 32, 2, 3, True), (100, 36, 1, 3, False), (72, 40, 2, 1, True), (416, 128, 2, 3, False)]:
    inp = synth.make_loss_inputs(W, H, N, C=C, seed=7, automasking=auto)
    ref = P.oracle_train_loss(inp)
    res, _ = TL._run_gpu(inp, auto=ref["auto"])
    print(W, H, N, C, auto, "loss", float(res.loss), ref["loss"], abs(float(res.loss) - ref["loss"]) / abs(ref["loss"]))
    for k, (g, r) in enumerate(zip(res.d_disparities, ref["d_disp"])):
        print("  ", P.grad_report(f"d_disp[{k}]", g, r))
    for s in range(2):
        print("  ", P.grad_report(f"d_rvec[{s}]", res.d_poses[s].rvec, ref["d_rvec"][s]))
        print("  ", P.grad_report(f"d_tvec[{s}]", res.d_poses[s].tvec, ref["d_tvec"][s]))

# flip diagnostics for the last case: are the dense-gradient outliers next to near-ties of min()?
from oracle import torch_ref as T
import numpy as np
x64 = torch.tensor(inp.x, dtype=torch.float64)
K = torch.tensor(inp.K.astype(np.float64)); invK = torch.tensor(inp.invK.astype(np.float64))
poses = [T.Pose(torch.tensor(r, dtype=torch.float64), torch.tensor(t, dtype=torch.float64)) for r, t in zip(inp.rvecs, inp.tvecs)]
_, parts = T.train_loss(x64, [torch.tensor(d, dtype=torch.float64) for d in inp.disparities], poses, None, K, invK,
                        T.Params(automasking=False, target_size=(inp.W, inp.H)), scales=inp.scales, return_parts=True)
warped = parts[-1][2]
pl = [T.photometric_loss(w, x64[:, 1]) for w in warped]
gap = (pl[0] - pl[1]).abs()[:, 0]
g = res.d_disparities[-1].double().cpu()[:, 0]; r = ref["d_disp"][-1][:, 0]
err = (g - r).abs() / r.abs().max()
idx = (err > 1e-4).nonzero()
for (n, y, xx) in idx.tolist():
    y0, y1, x0, x1 = max(0, y - 1), min(inp.H, y + 2), max(0, xx - 1), min(inp.W, xx + 2)
    print("outlier", n, y, xx, "err", float(err[n, y, xx]), "min |l0-l1| in 3x3:", float(gap[n, y0:y1, x0:x1].min()))
vis_err = (res.vis_warp_loss.double().cpu() - parts[-1][0]).abs().max()
print("vis max abs err", float(vis_err))
