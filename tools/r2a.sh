# round 2, call A: roofs (tools/microbench.cu), builders side by side, baselines of every config at HEAD of round 1
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/microbench.cu -o /tmp/mb && /tmp/mb > gpurun_out/r2a_microbench.json
cat gpurun_out/r2a_microbench.json
nproc
python tools/build_bench.py c4 2 > gpurun_out/r2a_build_bench_c4.json 2> gpurun_out/r2a_build_bench_c4.err
cat gpurun_out/r2a_build_bench_c4.json
for c in c4 c1 c2 c3 c5; do
  python bench.py --config $c --steps 10 --warmup 3 --cpu-seconds 8 > gpurun_out/r2a_bench_$c.json 2> gpurun_out/r2a_bench_$c.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/r2a_bench_$c.json").read())
print("$c", d["ms_per_step"], d["value"], d["config"].get("ms_per_step_spatial_order"), d["e2e"]["value"], d.get("cpu_baseline", {}).get("value"), d.get("cpu_baseline", {}).get("cores"))
PY
done
