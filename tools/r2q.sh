# round 2: beam walk (two-plane hull + visit budget) on the B200 — frame times of the build variants, then ncu of the C4 frame
mkdir -p gpurun_out
echo "== binary packet walk only (RTCUDA_WIDE=0)"
RTCUDA_WIDE=0 bash tools/ab_libs.sh "c4" "base" 2>&1 | sed 's/^/wide0 /'
echo "== beam walk"
bash tools/ab_libs.sh "c4 c2 c5 c1 c3" "base"
bash tools/ab_libs.sh "c4 c5" "b24 b16 w32 w32b16 bud32 bud256"
python tools/render_once.py c4 3 > gpurun_out/r2q_plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 2 -c 1 -o gpurun_out/r2q_c4 python tools/render_once.py c4 3 > gpurun_out/r2q_ncu_c4.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2q_plain_c4.log
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2q_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2q_gpu_tests.log
