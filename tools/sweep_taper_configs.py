"""Experiment 20c: is the taper rule (fitted at 1024x320) right at the other BASELINE shapes?  Fused-kernel ms of configs 2 and 3 for
forced segment counts k (-1 = equal segments by the cost model, 0 = the library's rule), L2 flushed.  python tools/sweep_taper_configs.py"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
import torch, mdb200
from mdb200 import synth
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name, N, ks in (("config3", 32, [0, -1, 2, 3, 4, 5, 6, 8]), ("config2", 12, [0, -1, 2, 3, 4, 6, 8]), ("config3", 8, [0, -1, 2, 3, 4, 6])):
    inp = synth.make_config(name, N=N)
    x = torch.tensor(inp.x, device=dev)
    disps = [torch.tensor(d, device=dev) for d in inp.disparities]
    poses = [mdb200.Pose(torch.tensor(r, device=dev), torch.tensor(t, device=dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
    auto = mdb200.automasking_loss(mdb200.SSIM(), x, x[:, 1], source_ids=inp.source_ids) if inp.automasking else None
    row = {}
    for k in ks:
        mdb200.debug_option("strip_taper", k)
        plan = mdb200.FusedLoss(W=inp.W, H=inp.H, C=3, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in disps],
                                scales=inp.scales, target_id=2, source_ids=[1, 3], K=inp.K, invK=inp.invK, automasking=inp.automasking)
        for _ in range(3):
            plan(x, disps, poses, auto)
        steps = 30
        mdb200.profile_begin(steps)
        for _ in range(steps):
            flush.zero_()
            plan(x, disps, poses, auto)
        torch.cuda.synchronize()
        row[f"k{k}"] = round(sum(mdb200.profile_read(i)[1] for i in range(steps)) / steps, 4)
        mdb200.profile_begin(0)
        del plan
    mdb200.debug_option("strip_taper", 0)
    print(name, N, json.dumps(row), flush=True)
