# round 2: evidence for the current build — launch list of the bench command, full captures of the three render kernels
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2o_bench_plain.json 2> gpurun_out/r2o_bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2o_launches_c4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2o_ncu_bench.log 2>&1
echo "launch list rc=$?"
for c in c4 c3 c3glossy; do
  python tools/render_once.py $c 3 > gpurun_out/r2o_plain_$c.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 2 -c 1 -o gpurun_out/r2o_$c python tools/render_once.py $c 3 > gpurun_out/r2o_ncu_$c.log 2>&1
  echo "$c rc=$?"; cat gpurun_out/r2o_plain_$c.log
done
