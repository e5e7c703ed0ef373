# usage: bash tools/sweep.sh "<EXTRA flag variants separated by ;>"   (runs on the GPU box: rebuild + short bench per variant)
IFS=';' read -ra VARS <<< "$1"
i=0
for v in "${VARS[@]}"; do
  make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 EXTRA="$v" >/dev/null 2>&1
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/sweep_$i.log 2>&1
  echo "$v" >> gpurun_out/sweep_$i.log
  i=$((i+1))
done
make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 >/dev/null 2>&1
