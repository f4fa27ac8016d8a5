# round 2: where a slow first engine creation goes (bench line of r2ak showed build_ms 1984 against 95 in r2f) — phase
# traces of rt_scene_create and of the device SAH builder, three fresh processes and one process creating five engines
mkdir -p gpurun_out
for i in 1 2 3; do
  RTCUDA_TRACE=1 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r2al_bench_$i.json 2> gpurun_out/r2al_bench_$i.err; echo "bench $i rc=$?"
  grep -h "rtcuda trace\|scene ready" gpurun_out/r2al_bench_$i.err | cut -c1-260
done
RTCUDA_TRACE=1 python tools/build_bench.py c4 3 > gpurun_out/r2al_build.json 2> gpurun_out/r2al_build.err; echo "build rc=$?"
grep -h "rtcuda trace" gpurun_out/r2al_build.err | grep "sah:\|BLAS\|total\|upload" | cut -c1-200
