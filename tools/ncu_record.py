"""gpurun_out/r2_strip.ncu-rep (tools/capture_profiles.sh) -> profiles/r2_strip_ncu.json: the metrics the docs and bench.py's
roofline quote, stamped with the git blob hash of csrc/photo_strip.inl (bench.py prints `traffic: null` when the kernel source
has changed since).  python tools/ncu_record.py [rep] [out]"""
import csv, hashlib, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "r2_strip.ncu-rep")
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "r2_strip_ncu.json")
def blob(rel):
    d = open(os.path.join(ROOT, rel), "rb").read()
    return hashlib.sha1(b"blob %d\0" % len(d) + d).hexdigest()
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[-1]
m, u = dict(zip(hdr, vals)), dict(zip(hdr, units))
def nbytes(k):
    return float(m[k]) * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u[k].lower()]
keep = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
units_wpx = 167772160   # config 5: 64 items x 4 scales x 2 sources x 1024 x 320
rec = {"kernel": m.get("Kernel Name", "photo_strip_kernel<3,2,false,true>"), "config": "config5", "items": 64,
       "command": "ncu --set full --clock-control none --import-source on -k regex:photo_strip -s 3 -c 1 python bench.py --steps 2 "
                  "--warmup 3 --no-e2e --no-cpu-baseline --no-aux-configs   (tools/capture_profiles.sh)",
       "dram_bytes_read": nbytes("dram__bytes_read.sum"), "dram_bytes_write": nbytes("dram__bytes_write.sum"),
       "thread_instr_per_warped_px": float(m["smsp__inst_executed.sum"]) * 32 / units_wpx,
       "metrics": {k: float(m[k]) for k in keep if k in m},
       "stalls_per_issued_instruction": {k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): float(m[k])
                                         for k in m if k.startswith("smsp__average_warps_issue_stalled") and float(m[k] or 0) > 0.02},
       "source_hash": {"monodepth2.jl_b200/csrc/photo_strip.inl": blob("monodepth2.jl_b200/csrc/photo_strip.inl")}}
json.dump(rec, open(out, "w"), indent=1)
print("wrote", out, "traffic", rec["dram_bytes_read"] + rec["dram_bytes_write"], "instr/wpx", round(rec["thread_instr_per_warped_px"], 1))
