# usage: [BENCH_ARGS="--config config3"] bash tools/sweep_up.sh "<EXTRA flag variants separated by ;>"  -- like sweep.sh, prints all segments
IFS=';' read -ra VARS <<< "$1"
i=0
for v in "${VARS[@]}"; do
  make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 EXTRA="$v" >/dev/null 2>&1
  python bench.py $BENCH_ARGS --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/sweep_$i.log 2>&1
  echo "$v" >> gpurun_out/sweep_$i.log
  i=$((i+1))
done
make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 >/dev/null 2>&1
python - <<'P'
import json, glob
for f in sorted(glob.glob('gpurun_out/sweep_*.log')):
    lines = open(f).read().strip().splitlines()
    for l in lines:
        if l.startswith('{'):
            d = json.loads(l); print(f, lines[-1], '| step', round(d['ms_per_step'], 3), {k: round(v, 4) for k, v in d['roofline']['segments_ms'].items()})
P
