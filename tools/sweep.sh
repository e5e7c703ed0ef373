set -x
for mb in 2 3 4; do
  make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 EXTRA=-DSTRIP_MINB=$mb >/dev/null 2>&1
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/sweep_mb$mb.log 2>&1
done
make -C monodepth2.jl_b200 clean >/dev/null; make -C monodepth2.jl_b200 -j8 >/dev/null 2>&1
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/t2.log
