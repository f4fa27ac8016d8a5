# A/B of librtcuda builds on the bench workload: tools/ab.sh cur head ... (build_<name>/librtcuda.so under csrc; cur = in-tree)
for v in "$@"; do
  if [ "$v" = cur ]; then unset RTCUDA_LIB; else export RTCUDA_LIB=$PWD/myraytracer_b200/csrc/build_$v/librtcuda.so; fi
  python bench.py --steps 20 --warmup 3 --no-cpu $BENCH_ARGS 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', round(d['value'],1), round(d['ms_per_step'],4), 'spatial', d['config'].get('ms_per_step_spatial_order'), 'e2e', round(d['e2e']['value'],1))"
done
