"""Fused-kernel time against the number of row segments per strip (mdb_debug_option "strip_segs"), per batch size:
calibrates strip_geom's cost model.  python tools/sweep_segs.py [batches...]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
import torch, mdb200
from mdb200 import synth
dev = torch.device("cuda:0")
batches = [int(a) for a in sys.argv[1:]] or [8, 16, 32, 64]
full = synth.make_config("config5", N=max(batches))
out = {}
for N in batches:
    x = torch.tensor(full.x[:N], device=dev)
    disps = [torch.tensor(d[:N], device=dev) for d in full.disparities]
    poses = [mdb200.Pose(torch.tensor(r[:N], device=dev), torch.tensor(t[:N], device=dev)) for r, t in zip(full.rvecs, full.tvecs)]
    row = {}
    for segs in [0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 13, 16, 20]:
        mdb200.debug_option("strip_segs", segs)
        plan = mdb200.FusedLoss(W=full.W, H=full.H, C=3, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in disps],
                                scales=full.scales, target_id=2, source_ids=[1, 3], K=full.K, invK=full.invK, n_global=64)
        for _ in range(3):
            plan(x, disps, poses)
        mdb200.profile_begin(10)
        for _ in range(10):
            plan(x, disps, poses)
        torch.cuda.synchronize()
        ms = sum(mdb200.profile_read(i)[1] for i in range(10)) / 10
        mdb200.profile_begin(0)
        row[segs] = round(ms, 4)
        del plan
    mdb200.debug_option("strip_segs", 0)
    out[N] = row
    print(N, json.dumps(row), flush=True)
