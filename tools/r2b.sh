# round 2, call B: GPU test suite after the ABI v2 work, new bench line, GPM probe
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2b_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2b_gpu_tests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2b_bench_c4.json 2> gpurun_out/r2b_bench_c4.err; echo "bench rc=$?"; tail -3 gpurun_out/r2b_bench_c4.err
python tools/gpm_probe.py > gpurun_out/r2b_gpm.log 2>&1; cat gpurun_out/r2b_gpm.log
