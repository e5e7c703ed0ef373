"""Times the fused kernel of every library under monodepth2.jl_b200/variants/ (tools/build_variants.sh) on the GPU:
config 5 (64 x 1024x320) and config 3 (32 x 640x192, automask), CUDA events inside the C-ABI call, and checks the loss
against the standard build.  One subprocess per library (MDB200_LIB).  python tools/time_variants.py [names...]"""
import glob, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, json
sys.path.insert(0, os.path.join(%r, "monodepth2.jl_b200"))
import torch, mdb200
from mdb200 import synth
dev = torch.device("cuda:0")
out = {}
for name, N in (("config5", int(os.environ.get("VAR_BATCH", "64"))), ("config3", 32)):
    inp = synth.make_config(name, N=N)
    x = torch.tensor(inp.x, device=dev)
    disps = [torch.tensor(d, device=dev) for d in inp.disparities]
    poses = [mdb200.Pose(torch.tensor(r, device=dev), torch.tensor(t, device=dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
    auto = mdb200.automasking_loss(mdb200.SSIM(), x, x[:, 1], source_ids=inp.source_ids) if inp.automasking else None
    plan = mdb200.FusedLoss(W=inp.W, H=inp.H, C=3, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in disps],
                            scales=inp.scales, target_id=2, source_ids=[1, 3], K=inp.K, invK=inp.invK, automasking=inp.automasking)
    for _ in range(3):
        res = plan(x, disps, poses, auto)
    steps = 10
    mdb200.profile_begin(steps)
    for _ in range(steps):
        res = plan(x, disps, poses, auto)
    torch.cuda.synchronize()
    ms = sum(mdb200.profile_read(i)[1] for i in range(steps)) / steps
    mdb200.profile_begin(0)
    out[name] = dict(fused_ms=round(ms, 4), loss=float(res.loss), g=float(res.d_disparities[-1].abs().sum()), r=float(res.d_poses[0].rvec.abs().sum()))
    del plan, x, disps
print("RESULT " + json.dumps(out))
''' % ROOT
names = sys.argv[1:]
libs = [("standard", os.path.join(ROOT, "monodepth2.jl_b200", "mdb200", "libmdb200.so"))]
for f in sorted(glob.glob(os.path.join(ROOT, "monodepth2.jl_b200", "variants", "libmdb200_*.so"))):
    n = os.path.basename(f)[len("libmdb200_"):-3]
    if not names or n in names:
        libs.append((n, f))
base = None
for n, f in libs:
    env = dict(os.environ, MDB200_LIB=f)
    p = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
    if not line:
        print(n, "FAILED", p.stderr[-400:]); continue
    r = json.loads(line[0][7:])
    if base is None:
        base = r
    ok = all(abs(r[c]["loss"] - base[c]["loss"]) <= 2e-6 * abs(base[c]["loss"]) and abs(r[c]["g"] - base[c]["g"]) <= 1e-4 * base[c]["g"]
             and abs(r[c]["r"] - base[c]["r"]) <= 1e-4 * base[c]["r"] for c in r)
    print(f"{n:28s} config5 {r['config5']['fused_ms']:.4f} ms   config3 {r['config3']['fused_ms']:.4f} ms   same result: {ok}", flush=True)
