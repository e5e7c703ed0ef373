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
