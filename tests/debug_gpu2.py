import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")]
import torch, numpy as np
import parity as P
from mdb200 import synth
import test_gpu_loss as TL

def show(tag, inp):
    ref = P.oracle_train_loss(inp)
    res, _ = TL._run_gpu(inp, auto=ref["auto"])
    nout = sum(P.grad_report("d", g, r)["n_out"] for g, r in zip(res.d_disparities, ref["d_disp"]))
    pe = [P.grad_report("r", res.d_poses[s].rvec, ref["d_rvec"][s])["max_err"] for s in range(2)]
    te = [P.grad_report("t", res.d_poses[s].tvec, ref["d_tvec"][s])["max_err"] for s in range(2)]
    print(tag, "loss rel", abs(float(res.loss) - ref["loss"]) / abs(ref["loss"]), "dense outliers", nout,
          "rvec err", ["%.2e" % v for v in pe], "tvec err", ["%.2e" % v for v in te])

for tid, sids in [(2, [1, 3]), (2, [3, 1]), (1, [2, 3]), (3, [1, 2])]:
    inp = synth.make_loss_inputs(104, 32, 2, seed=7)
    inp.target_id = tid; inp.source_ids = sids
    show(f"tgt={tid} src={sids}", inp)
# swap the poses between the sources
inp = synth.make_loss_inputs(104, 32, 2, seed=7)
inp.rvecs = inp.rvecs[::-1]; inp.tvecs = inp.tvecs[::-1]
show("poses swapped", inp)
for seed in range(20, 26):
    show(f"416x128 N=1 seed={seed}", synth.make_loss_inputs(416, 128, 1, seed=seed))
