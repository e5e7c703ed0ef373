import sys, os, copy
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")]
import torch, numpy as np
import parity as P
from mdb200 import synth
import test_gpu_loss as TL

base = synth.make_loss_inputs(104, 32, 2, seed=7)
print("rvecs", base.rvecs, "tvecs", base.tvecs)
for n in range(2):
    for k in range(4):
        inp = copy.deepcopy(base)
        inp.N = 1
        inp.x = base.x[n:n+1].copy()
        inp.disparities = [base.disparities[k][n:n+1].copy()]
        inp.scales = [base.scales[k]]
        inp.rvecs = [r[n:n+1].copy() for r in base.rvecs]; inp.tvecs = [t[n:n+1].copy() for t in base.tvecs]
        ref = P.oracle_train_loss(inp)
        res, _ = TL._run_gpu(inp)
        for s in range(2):
            g = res.d_poses[s].rvec.double().cpu(); r = ref["d_rvec"][s]
            gt = res.d_poses[s].tvec.double().cpu(); rt = ref["d_tvec"][s]
            print(f"item {n} scale {k} src {s}: d_rvec gpu {g.numpy().ravel()} ref {r.numpy().ravel()} | d_tvec gpu {gt.numpy().ravel()} ref {rt.numpy().ravel()}")
        rep = P.grad_report("dd", res.d_disparities[0], ref["d_disp"][0])
        print("    d_disp", rep["max_err"], rep["n_out"])
