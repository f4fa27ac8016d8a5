# round 2: the general kernel's scene walk with both levels in one loop (flat) against the nested loops; resident blocks
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2ac_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r2ac_gpu_tests.log
bash tools/ab_libs.sh "c3glossy c3area" "base nested flat3 flat4"
python bench.py --config c3 --variant 2 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c3 general kernel flat', round(d['ms_per_step'],3))"
RTCUDA_LIB=$PWD/myraytracer_b200/librtcuda_nested.so python bench.py --config c3 --variant 2 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c3 general kernel nested', round(d['ms_per_step'],3))"
