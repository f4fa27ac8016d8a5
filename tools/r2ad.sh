# round 2: flat against nested walk of the general kernel on mesh scenes (general kernel forced: --variant 2)
mkdir -p gpurun_out
for c in c4 c5 c2; do
  for tag in flat nested; do
    lib=$PWD/myraytracer_b200/librtcuda.so; [ $tag = nested ] && lib=$PWD/myraytracer_b200/librtcuda_nested.so
    RTCUDA_LIB=$lib python bench.py --config $c --variant 2 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$c general kernel $tag', round(d['ms_per_step'],3), d['frame_hash'])"
  done
done
