# round 2: beam walk of the wide BVH on the B200 — GPU parity suite, then frame times of the build variants
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2p_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2p_gpu_tests.log
echo "== binary packet walk only (RTCUDA_WIDE=0)"
RTCUDA_WIDE=0 bash tools/ab_libs.sh "c4 c2 c5" "base" 2>&1 | sed 's/^/wide0 /'
echo "== beam walk"
bash tools/ab_libs.sh "c4 c2 c5 c1 c3" "base"
bash tools/ab_libs.sh "c4 c2 c5" "b24 b16 w32 w32b24 w8"
