"""Is the spread of the per-rank loss-kernel times at 8 GPUs (0.63 .. 0.69 ms) the items or the GPUs?  Times the fused kernel
on ONE GPU for each of the eight 8-item shards of config 5.  python tools/shard_cost.py [--strided]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
import torch, mdb200
from mdb200 import synth
dev = torch.device("cuda:0")
inp = synth.make_config("config5")
sizes = [(d.shape[-1], d.shape[-2]) for d in inp.disparities]
plan = mdb200.FusedLoss(W=inp.W, H=inp.H, C=3, L=3, N=8, n_global=64, disp_sizes=sizes, scales=inp.scales, target_id=2, source_ids=[1, 3],
                        K=inp.K, invK=inp.invK, automasking=False)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
import numpy as np
strided = "--strided" in sys.argv      # item i -> shard i mod 8 (what torch's DistributedSampler does) instead of contiguous slabs
for k in range(8):
    idx = np.arange(k, 64, 8) if strided else np.arange(8 * k, 8 * k + 8)
    x = torch.tensor(inp.x[idx], device=dev)
    disps = [torch.tensor(d[idx], device=dev) for d in inp.disparities]
    poses = [mdb200.Pose(torch.tensor(r[idx], device=dev), torch.tensor(t[idx], device=dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
    for _ in range(3):
        plan(x, disps, poses)
    steps = 20
    mdb200.profile_begin(steps)
    for _ in range(steps):
        flush.zero_()
        plan(x, disps, poses)
    torch.cuda.synchronize()
    seg = [mdb200.profile_read(i) for i in range(steps)]
    mdb200.profile_begin(0)
    print("shard", k, "fused ms %.4f  step ms %.4f" % (sum(s[1] for s in seg) / steps, sum(sum(s) for s in seg) / steps))
