# round 2: scene arrays assembled on the device (device-built BLASes never visit the host): GPU suite, engine creation
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2ai_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r2ai_gpu_tests.log
RTCUDA_TRACE=1 python tools/build_bench.py c4 3 > gpurun_out/r2ai_build_bench.json 2> gpurun_out/r2ai_trace.log; echo "rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2ai_build_bench.json").read())
for k, v in d["builders"].items(): print(k, "create", [round(x, 1) for x in v["create_ms_all"]], "build", [round(x, 1) for x in v["build_ms_all"]], "frame", round(v["frame_ms_min"], 3), v["image_equals_host"])
PY
RTCUDA_BUILD_RESIDENT=0 python tools/build_bench.py c4 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
for k, v in d['builders'].items(): print('resident=0', k, 'create', [round(x, 1) for x in v['create_ms_all']], 'build', [round(x, 1) for x in v['build_ms_all']])"
