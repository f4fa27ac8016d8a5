# round 2 (third pass: pages touched on all threads, 2 MB pages requested): flatten of a device-ingested mesh without the one-thread zero fill (unwritten vectors, pages touched and
# per-triangle loops on all threads) — traces, build bench, ingest / flatten parity tests
mkdir -p gpurun_out
nproc
python -m pytest tests -m gpu -x -q -k "build or sah or ingest or resident or flatten or ply or anim" > gpurun_out/r2ap_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2ap_tests.log
for i in 1 2; do
  RTCUDA_TRACE=1 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r2ap_bench_$i.json 2> gpurun_out/r2ap_bench_$i.err; echo "bench $i rc=$?"
  grep -h "rtcuda trace\|scene ready" gpurun_out/r2ap_bench_$i.err | cut -c1-120
done
RTCUDA_TRACE=1 python tools/build_bench.py c4 3 > gpurun_out/r2ap_build.json 2> gpurun_out/r2ap_build.err; echo "build rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2ap_build.json"))
for k, v in d["builders"].items():
    print(k, "create", [round(x, 1) for x in v["create_ms_all"]], "build", [round(x, 1) for x in v["build_ms_all"]], "frame", round(v["frame_ms_min"], 3), v["image_equals_host"], v["blasNodes"])
PY
