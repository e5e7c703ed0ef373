"""Experiment 20: fused-kernel time with tapered row segments (mdb_debug_option "strip_taper" = k: k segments of lengths
~ k, k-1, ..., 1, long ones first in launch order) against the library's equal segments, per batch size, L2 flushed between
launches (as bench.py does below 2 x L2 of input).  Also checks the results against the equal-segment run.
python tools/sweep_taper.py [batches...]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
import torch, mdb200
from mdb200 import synth
dev = torch.device("cuda:0")
batches = [int(a) for a in sys.argv[1:]] or [8, 16, 32]
full = synth.make_config("config5", N=max(batches))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for N in batches:
    x = torch.tensor(full.x[:N], device=dev)
    disps = [torch.tensor(d[:N], device=dev) for d in full.disparities]
    poses = [mdb200.Pose(torch.tensor(r[:N], device=dev), torch.tensor(t[:N], device=dev)) for r, t in zip(full.rvecs, full.tvecs)]
    row, base = {}, None
    for k in [-1, 0, 3, 4, 5, 6, 8]:      # -1: equal segments (the cost model's count), 0: the library's choice
        mdb200.debug_option("strip_taper", k)
        plan = mdb200.FusedLoss(W=full.W, H=full.H, C=3, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in disps],
                                scales=full.scales, target_id=2, source_ids=[1, 3], K=full.K, invK=full.invK, n_global=64)
        for _ in range(3):
            res = plan(x, disps, poses)
        steps = 20
        mdb200.profile_begin(steps)
        for _ in range(steps):
            flush.zero_()
            res = plan(x, disps, poses)
        torch.cuda.synchronize()
        ms = sum(mdb200.profile_read(i)[1] for i in range(steps)) / steps
        mdb200.profile_begin(0)
        sig = (float(res.loss), float(res.d_disparities[-1].abs().sum()), float(res.d_poses[0].rvec.abs().sum()))
        if base is None:
            base = sig
        ok = all(abs(a - b) <= 1e-5 * abs(b) for a, b in zip(sig, base))
        row[k] = (round(ms, 4), plan.strip_geometry()["segments"], ok)
        del plan
    mdb200.debug_option("strip_taper", 0)
    print(N, json.dumps(row), flush=True)
