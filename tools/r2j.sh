mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs -k "parity_vs_oracle or variants or golden" > gpurun_out/r2j_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j_gpu_tests.log
RTCUDA_FORCE_RECURSE=1 python bench.py --config c4 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c4 forced recurse', d['ms_per_step'])"
bash tools/ab_libs.sh "c3 c5 c4" "base"
