mkdir -p gpurun_out
RTCUDA_FORCE_RECURSE=1 python bench.py --config c4 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c4 forced recurse', d['ms_per_step'])"
RTCUDA_FORCE_RECURSE=1 RTCUDA_LIB=$PWD/myraytracer_b200/librtcuda_noinc.so python bench.py --config c4 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c4 forced recurse noinc(old cold)', d['ms_per_step'])"
bash tools/ab_libs.sh "c3 c5" "base"
