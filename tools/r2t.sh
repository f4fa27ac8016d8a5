# round 2: auto build mode (big meshes through the device SAH builder), GPU suite with the beam tests, builders side by side, default bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2t_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r2t_gpu_tests.log
python tools/build_bench.py c4 3 > gpurun_out/r2t_build_bench_c4.json 2> gpurun_out/r2t_build_bench_c4.err; echo "build bench rc=$?"; cat gpurun_out/r2t_build_bench_c4.json
python bench.py > gpurun_out/r2t_bench_c4.json 2> gpurun_out/r2t_bench_c4.err; echo "bench rc=$?"; python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2t_bench_c4.json").read())
print({k: d[k] for k in ("value", "ms_per_step", "build_ms", "upload_ms", "bvh_build", "gpu_launches")}, d["e2e"]["value"], d["roofline"]["bound"], d["roofline"]["frac"], d.get("parity", {}).get("pass"), d.get("cpu_baseline", {}).get("value"))
PY
