# round 2, call D: wavefront secondary rays — GPU suite + C3/C5/C4 bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2d_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2d_gpu_tests.log
for c in c3 c5 c4 c1; do
  python bench.py --config $c --steps 10 --warmup 3 --no-cpu > gpurun_out/r2d_bench_$c.json 2> gpurun_out/r2d_bench_$c.err; echo "bench $c rc=$?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/r2d_bench_$c.json").read())
print("$c", d["ms_per_step"], d["value"], d["detail"]["ms_first_frame"], d["e2e"]["ms_per_step"])
PY
done
