mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2m_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2m_gpu_tests.log
bash tools/ab_libs.sh "c3glossy c3area" "base"
python bench.py --config c3 --variant 2 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c3 general kernel', d['ms_per_step'])"
python bench.py --config c4 --variant 2 --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c4 general kernel', d['ms_per_step'])"
