mkdir -p gpurun_out
RTCUDA_FORCE_RECURSE=1 python bench.py --config c4 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c4 forced recurse', d['ms_per_step'])"
python tools/render_once.py c5 3 > gpurun_out/r2i_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:renderKernelFast -s 2 -c 1 -o gpurun_out/r2i_c5 python tools/render_once.py c5 3 > gpurun_out/r2i_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2i_plain.log
