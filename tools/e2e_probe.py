"""Host-buffer step time of HostLoss for a few pipeline shapes (config 5)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200")):
    sys.path.insert(0, p)
import torch, numpy as np
import mdb200
from mdb200 import synth
inp = synth.make_config("config5")
N = inp.N
pin = lambda a: torch.tensor(a).pin_memory()
hd, hr, ht = [pin(d) for d in inp.disparities], [pin(r) for r in inp.rvecs], [pin(t) for t in inp.tvecs]
hu8 = pin(synth.to_u8_strips(inp.x)); hx = pin(inp.x)
common = dict(W=inp.W, H=inp.H, C=3, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in inp.disparities],
              scales=inp.scales, target_id=2, source_ids=[1, 3], K=inp.K, invK=inp.invK)
for fmt, x in (("u8_strip", hu8), ("f32", hx)):
    for chunk in (8, 16):
        for ramp in (False, True):
            plan = mdb200.HostLoss(chunk=chunk, x_format=fmt, ramp_up=ramp, **common)
            poses = [mdb200.Pose(r, t) for r, t in zip(hr, ht)]
            for _ in range(3): mdb200.train_loss_host(plan, x, hd, poses)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for _ in range(10): mdb200.train_loss_host(plan, x, hd, poses)
            torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
            print(fmt, "chunk", chunk, "ramp", ramp, [b - a for a, b in plan.chunks], f"{dt*1e3:.2f} ms")
            del plan; torch.cuda.empty_cache()
