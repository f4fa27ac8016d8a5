# round 2: engine creation after the device-resident assembly and the uninitialised host arrays
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2aj_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -2 gpurun_out/r2aj_gpu_tests.log
RTCUDA_TRACE=1 python tools/build_bench.py c4 4 > gpurun_out/r2aj_build_bench.json 2> gpurun_out/r2aj_trace.log; echo "rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2aj_build_bench.json").read())
for k, v in d["builders"].items(): print(k, "create", [round(x, 1) for x in v["create_ms_all"]], "build", [round(x, 1) for x in v["build_ms_all"]], "frame", round(v["frame_ms_min"], 3), v["image_equals_host"])
PY
grep "rtcuda trace" gpurun_out/r2aj_trace.log | tail -8
