"""gpurun_out/r2_render.ncu-rep (ncu --set full -k regex:render_(fwd|bwd)_kernel of `bench.py --config config4`) ->
profiles/r2_render_ncu.json: per kernel instantiation the metrics DESIGN.md / profiles/r2_strip_summary.md quote.
python tools/ncu_record_render.py [rep] [out]"""
import csv, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "r2_render.ncu-rep")
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "r2_render_ncu.json")
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, units = rows[0], rows[1]
keep = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_requests_srcunit_tex_op_red.sum"]
stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
plane_px = 416 * 128 * 32 * 2
recs, seen = [], set()
for r in rows[2:]:
    m = dict(zip(hdr, r))
    name = m["Kernel Name"]
    if name in seen:
        continue
    seen.add(name)
    rec = {"kernel": name, "metrics": {k: (float(m[k]), dict(zip(hdr, units))[k]) for k in keep if k in m and m[k] not in ("", "n/a")},
           "warp_instructions_per_warp_sample": float(m["smsp__inst_executed.sum"]) / (plane_px / 32),
           "stalls_per_issued_instruction": {k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): round(float(m[k]), 3)
                                             for k in stall if float(m[k]) >= 0.05}}
    recs.append(rec)
json.dump({"config": "config4: 416x128, 32 planes, B = 2", "command": "ncu --set full --clock-control none -k regex:render_(fwd|bwd)_kernel "
           "python bench.py --config config4 --steps 1 --warmup 1", "kernels": recs}, open(out, "w"), indent=1)
print("wrote", out, [(r["kernel"][:48], round(r["warp_instructions_per_warp_sample"])) for r in recs])
