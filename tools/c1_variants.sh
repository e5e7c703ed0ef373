for n in cap152 cap184 cap216 pcap160 pcap200; do
  echo -n "$n: "; MDB200_LIB=$PWD/monodepth2.jl_b200/variants/libmdb200_$n.so python bench.py --config config1 --steps 10 --warmup 3 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['us_per_iteration'], d.get('loss_first_last'))"
done
