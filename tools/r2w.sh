# round 2: first-frame tile probe on one GPU — GPU suite, bench with and without the probe
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2w_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2w_gpu_tests.log
for p in 1 0; do
  RTCUDA_PROBE=$p python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2w_bench_probe$p.json 2> gpurun_out/r2w_bench_probe$p.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/r2w_bench_probe$p.json").read())
print("probe $p", "ms", round(d["ms_per_step"], 3), "first", round(d["detail"]["ms_first_frame"], 3), "spatial", round(d["detail"]["ms_per_step_spatial_order"], 3), d["frame_hash"])
PY
done
