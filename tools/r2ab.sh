# round 2: 16-warp instantiation of the queue kernel for scenes without big meshes; engine creation after the host-copy trims
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2ab_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2ab_gpu_tests.log
bash tools/ab_libs.sh "c1 c2 c3 c5" "base"
RTCUDA_SMALL_GEOMETRY=0 bash tools/ab_libs.sh "c3" "base" | sed 's/^/32-warp /'
RTCUDA_TRACE=1 python tools/build_bench.py c4 2 > gpurun_out/r2ab_build_bench.json 2> gpurun_out/r2ab_trace.log; echo "trace rc=$?"; grep "rtcuda trace" gpurun_out/r2ab_trace.log | sed -n '9,16p'
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2ab_build_bench.json").read())
for k, v in d["builders"].items(): print(k, "create", round(v["create_ms_min"], 1), "build", round(v["build_ms_min"], 1), "frame", round(v["frame_ms_min"], 3))
PY
