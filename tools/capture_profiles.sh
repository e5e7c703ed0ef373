#!/bin/bash
# Runs ON THE GPU BOX (gpurun): the evidence files of the round.  Every profiled command is first run plain (must exit 0).
#   1. ncu launch list of one bench step                     -> gpurun_out/r2_launches.csv
#   2. ncu --set full of the strip kernel (config 5, batch 64) -> gpurun_out/r2_strip.ncu-rep (+ raw csv)
set -u
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-aux-configs"
$BENCH > gpurun_out/r2_plain.json 2> gpurun_out/r2_plain.err || { echo "plain bench failed"; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -c 40 --csv \
    --log-file gpurun_out/r2_launches.csv $BENCH > gpurun_out/r2_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:photo_strip -s 3 -c 1 -f -o gpurun_out/r2_strip $BENCH > gpurun_out/r2_ncu_full.log 2>&1
ncu -i gpurun_out/r2_strip.ncu-rep --page raw --csv > gpurun_out/r2_strip_raw.csv 2>/dev/null
# compute-sanitizer is CLOSED on this GPU pool (the wrapper answers "compute-sanitizer is closed on this pool and stays
# closed: runs under it have left GPUs needing a reset", rc 86; recorded in profiles/r2_sanitizer.md), so no memcheck /
# racecheck / initcheck pass exists; tests/test_gpu_guard.py stands in with guard bands and poisoned buffers.
