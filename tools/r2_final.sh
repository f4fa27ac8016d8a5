# round 2, final evidence for the committed build: GPU suite, default bench, launch list of the bench command, full ncu
# captures of the three render kernels (C4 lean fast kernel, C3 queue kernel at 16 warps, C3 glossy general kernel)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2f_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r2f_gpu_tests.log
python bench.py > gpurun_out/r2f_bench_c4.json 2> gpurun_out/r2f_bench_c4.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2f_bench_plain.json 2> gpurun_out/r2f_bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2f_launches_c4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2f_ncu_bench.log 2>&1
echo "launch list rc=$?"
for c in c4 c3 c3glossy; do
  python tools/render_once.py $c 3 > gpurun_out/r2f_plain_$c.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 2 -c 1 -o gpurun_out/r2f_$c python tools/render_once.py $c 3 > gpurun_out/r2f_ncu_$c.log 2>&1
  echo "$c rc=$?"; cat gpurun_out/r2f_plain_$c.log
done
for c in c1 c2 c3 c5 c3glossy c3area; do
  python bench.py --config $c --steps 10 --warmup 3 --no-cpu > gpurun_out/r2f_bench_$c.json 2> gpurun_out/r2f_bench_$c.err
done
python - <<'PY'
import json
for c in ("c1", "c2", "c3", "c4", "c5", "c3glossy", "c3area"):
    d = json.loads(open(f"gpurun_out/r2f_bench_{c}.json").read())
    print(c, "ms", round(d["ms_per_step"], 3), "Mrays/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["frame_hash"], d.get("roofline", {}).get("frac"))
PY
