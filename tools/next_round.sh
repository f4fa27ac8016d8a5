# First GPU call of the next round: what was written after round 1's B200 minutes were spent, measured.
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash tools/next_round.sh'
# 1. the pending GPU tests (device binned-SAH builder), verbosely, so that XPASS / XFAIL is visible
# 2. builders side by side on C4: build time, tree size, frame time (tools/build_bench.py)
# 3. the bench line with the device-built SAH tree next to the default
mkdir -p gpurun_out
python -m pytest tests/test_zz_pending_gpu.py -m gpu -q -rxX -s > gpurun_out/pending_gpu_tests.log 2>&1
tail -5 gpurun_out/pending_gpu_tests.log
python tools/build_bench.py c4 3 > gpurun_out/build_bench_c4.json 2> gpurun_out/build_bench_c4.err
cat gpurun_out/build_bench_c4.json
python bench.py --steps 10 --warmup 3 --no-cpu --build device-sah > gpurun_out/bench_device_sah.json 2>/dev/null
python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_host_sah.json 2>/dev/null
cat gpurun_out/bench_device_sah.json gpurun_out/bench_host_sah.json | python -c "
import sys, json
for line in sys.stdin:
    d = json.loads(line); print(d['config'].get('parallelism'), d.get('bvh_build'), d['ms_per_step'], d['value'], d.get('build_ms'))"
