# usage: bash tools/sweep_c1.sh "<EXTRA flag variants separated by ;>"  (GPU box: rebuild + config-1 and config-5 bench per variant)
IFS=';' read -ra VARS <<< "$1"
i=0
for v in "${VARS[@]}"; do
  make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 EXTRA="$v" >/dev/null 2>&1
  echo "$v" > gpurun_out/sweepc1_$i.log
  python bench.py --config config1 --steps 5 --warmup 2 --no-cpu-baseline 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('c1 us/iter', d['us_per_iteration'], 'eager', d['eager_us_per_iteration'])" >> gpurun_out/sweepc1_$i.log
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('c5 ms', d['ms_per_step'], d['roofline']['segments_ms'])" >> gpurun_out/sweepc1_$i.log
  i=$((i+1))
done
make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 >/dev/null 2>&1
