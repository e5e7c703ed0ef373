#!/bin/bash
# Runs ON THE GPU BOX (gpurun): the evidence files of the round.  Every profiled command is first run plain (must exit 0).
#   1. ncu launch list of one bench step                     -> gpurun_out/r2_launches.csv
#   2. ncu --set full of the strip kernel (config 5, batch 64) -> gpurun_out/r2_strip.ncu-rep (+ raw csv)
#   3. compute-sanitizer memcheck / racecheck / initcheck over the geometry / segment / render-backward tests
set -u
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-aux-configs"
$BENCH > gpurun_out/r2_plain.json 2> gpurun_out/r2_plain.err || { echo "plain bench failed"; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -c 40 --csv \
    --log-file gpurun_out/r2_launches.csv $BENCH > gpurun_out/r2_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:photo_strip -s 3 -c 1 -f -o gpurun_out/r2_strip $BENCH > gpurun_out/r2_ncu_full.log 2>&1
ncu -i gpurun_out/r2_strip.ncu-rep --page raw --csv > gpurun_out/r2_strip_raw.csv 2>/dev/null
SEL='geometr or segments or render_backward or bwd_vs_oracle'
python -m pytest tests -m gpu -x -q -k "$SEL" > gpurun_out/r2_sanitizer_plain.log 2>&1 || { echo "plain tests failed"; tail -5 gpurun_out/r2_sanitizer_plain.log; }
for tool in memcheck racecheck initcheck; do
  timeout 1500 compute-sanitizer --tool $tool --log-file gpurun_out/r2_sanitizer_$tool.log --print-limit 20 \
      python -m pytest tests -m gpu -x -q -k "$SEL" > gpurun_out/r2_sanitizer_${tool}_pytest.log 2>&1
  echo "$tool rc=$? $(tail -1 gpurun_out/r2_sanitizer_${tool}_pytest.log) :: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/r2_sanitizer_$tool.log | tail -1)"
done
