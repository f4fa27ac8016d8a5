# round 2: beam walk, third form (both halves of the warp test a child: axis A / axis B)
mkdir -p gpurun_out
echo "== binary packet walk only"
bash tools/ab_libs.sh "c4 c5 c2" "base" 2>&1 | sed 's/^/wide0 /'
echo "== beam walk (RTCUDA_WIDE=1)"
export RTCUDA_WIDE=1
bash tools/ab_libs.sh "c4 c5 c2 c3" "base"
bash tools/ab_libs.sh "c4 c5" "b24 bud48 bud192"
python tools/render_once.py c4 3 > gpurun_out/r2s_plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 2 -c 1 -o gpurun_out/r2s_c4 python tools/render_once.py c4 3 > gpurun_out/r2s_ncu_c4.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2s_plain_c4.log
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -rs > gpurun_out/r2s_gpu_tests.log 2>&1; echo "suite (beam on) rc=$?"; tail -3 gpurun_out/r2s_gpu_tests.log
