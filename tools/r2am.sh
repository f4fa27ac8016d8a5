# round 2: device SAH builder on two arenas instead of ~220 cudaMalloc calls — traces, build bench, builder parity tests
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "build or sah or ingest or resident" > gpurun_out/r2am_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2am_tests.log
for i in 1 2; do
  RTCUDA_TRACE=1 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r2am_bench_$i.json 2> gpurun_out/r2am_bench_$i.err; echo "bench $i rc=$?"
  grep -h "rtcuda trace\|scene ready" gpurun_out/r2am_bench_$i.err | grep "sah:\|BLAS\|total\|upload" | cut -c1-200
done
RTCUDA_TRACE=1 python tools/build_bench.py c4 3 > gpurun_out/r2am_build.json 2> gpurun_out/r2am_build.err; echo "build rc=$?"
grep -h "rtcuda trace" gpurun_out/r2am_build.err | grep "sah:\|BLAS\|build phase" | tail -12 | cut -c1-200
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2am_build.json"))
for k, v in d["builders"].items():
    print(k, "create", [round(x, 1) for x in v["create_ms_all"]], "build", [round(x, 1) for x in v["build_ms_all"]], "frame", round(v["frame_ms_min"], 3), v["image_equals_host"], v["blasNodes"])
PY
