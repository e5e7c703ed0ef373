"""Holds the reference's own outputs (written by monodepth2.jl_b200/julia/run_reference_on_dump.jl on a machine
with Julia) against the float64 oracle -- the step that turns "parity unpinned" into "pinned".

    python -m mdb200.synth config2 /tmp/dump --batch 2     # here or anywhere
    julia --project=<Monodepth2.jl> monodepth2.jl_b200/julia/run_reference_on_dump.jl /tmp/dump
    python tools/compare_reference_dump.py config2 /tmp/dump --batch 2
"""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from mdb200 import synth  # noqa: E402
from oracle import c_ref  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("config"); ap.add_argument("dir"); ap.add_argument("--batch", type=int, default=0)
a = ap.parse_args()
inp = synth.make_config(a.config, **({"N": a.batch} if a.batch else {}))
ref = c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, None, inp.K, inp.invK, target_id=inp.target_id,
                       source_ids=inp.source_ids, scales=inp.scales, automasking=False, threads=os.cpu_count() or 1)
rd = lambda n: np.fromfile(os.path.join(a.dir, n), dtype="<f4").astype(np.float64)
loss = rd("ref_loss.f32")[0]
print(f"loss: reference {loss:.8f}  oracle {ref['loss']:.8f}  rel {abs(loss - ref['loss']) / abs(ref['loss']):.2e}")
ok = abs(loss - ref["loss"]) <= 1e-5 * abs(ref["loss"])
for k, g in enumerate(ref["d_disp"]):
    r = rd(f"ref_d_disp{k}.f32")
    e = np.abs(r - g.reshape(-1)).max() / np.abs(g).max()
    print(f"d_disp[{k}]: max err {e:.2e} of max-abs"); ok &= e <= 1e-3      # the reference computes in Float32
for s in range(2):
    for nm, g in (("rvec", ref["d_rvec"][s]), ("tvec", ref["d_tvec"][s])):
        r = rd(f"ref_d_{nm}{s}.f32")
        e = np.abs(r - g.reshape(-1)).max() / np.abs(g).max()
        print(f"d_{nm}[{s}]: max err {e:.2e} of max-abs"); ok &= e <= 1e-2
print("PINNED" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
