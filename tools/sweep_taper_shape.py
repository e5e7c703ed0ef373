"""Experiment 20b: shape of the tapered segments -- lengths ~ (k - i)^p -- per batch size.  python tools/sweep_taper_shape.py [batches...]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
import torch, mdb200
from mdb200 import synth
dev = torch.device("cuda:0")
batches = [int(a) for a in sys.argv[1:]] or [8, 64]
full = synth.make_config("config5", N=max(batches))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for N in batches:
    x = torch.tensor(full.x[:N], device=dev)
    disps = [torch.tensor(d[:N], device=dev) for d in full.disparities]
    poses = [mdb200.Pose(torch.tensor(r[:N], device=dev), torch.tensor(t[:N], device=dev)) for r, t in zip(full.rvecs, full.tvecs)]
    row = {}
    ks = [int(k) for k in os.environ.get('KS', '').split(',') if k] or ([3, 4] if N >= 32 else [5, 6, 7])
    for k in ks:
        for p10 in [int(q) for q in os.environ.get('PS', '').split(',') if q] or [5, 10, 15, 20]:
            mdb200.debug_option("strip_taper", k)
            mdb200.debug_option("strip_taper_pow10", p10)
            plan = mdb200.FusedLoss(W=full.W, H=full.H, C=3, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in disps],
                                    scales=full.scales, target_id=2, source_ids=[1, 3], K=full.K, invK=full.invK, n_global=64)
            for _ in range(3):
                plan(x, disps, poses)
            steps = 20
            mdb200.profile_begin(steps)
            for _ in range(steps):
                flush.zero_()
                plan(x, disps, poses)
            torch.cuda.synchronize()
            row[f"k{k}p{p10}"] = round(sum(mdb200.profile_read(i)[1] for i in range(steps)) / steps, 4)
            mdb200.profile_begin(0)
            del plan
    mdb200.debug_option("strip_taper", 0); mdb200.debug_option("strip_taper_pow10", 0)
    print(N, json.dumps(row), flush=True)
