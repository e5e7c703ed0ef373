# times config 4 (render) for every library under monodepth2.jl_b200/variants/ and the standard build
for lib in monodepth2.jl_b200/mdb200/libmdb200.so $(ls monodepth2.jl_b200/variants/libmdb200_*.so 2>/dev/null); do
  echo -n "$(basename $lib): "; MDB200_LIB=$PWD/$lib python bench.py --config config4 --steps 20 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fwd %.1f bwd %.1f |' % (1e3*d['roofline']['fwd']['ms'], 1e3*d['roofline']['bwd']['ms']), ' | '.join('%.1f/%.1f' % (1e3*v['fwd_ms'], 1e3*v['bwd_ms']) for v in d['variants'].values()))"
done
