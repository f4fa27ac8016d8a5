# round 2: confirmation of the committed tree (builder arenas, threaded flatten) on a fresh box — GPU suite, smoke, default bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2aq_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r2aq_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2aq_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2aq_smoke.log
python bench.py > gpurun_out/r2aq_bench_c4.json 2> gpurun_out/r2aq_bench_c4.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2aq_bench_c4.json").read())
print({k: d[k] for k in ("value", "ms_per_step", "build_ms", "upload_ms", "bvh_build", "gpu_launches")}, "e2e", d["e2e"]["value"], d["roofline"]["bound"], d["roofline"]["frac"], "parity", d.get("parity", {}).get("pass"), "cpu", d.get("cpu_baseline", {}).get("value"), d.get("speedup_measured"))
PY
