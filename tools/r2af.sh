# round 2: counting pass of an area-light frame on the fast kernel; GPU suite
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2af_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r2af_gpu_tests.log
bash tools/ab_libs.sh "c3area c3 c4" "base"
RTCUDA_AREA_COUNT_GENERAL=1 bash tools/ab_libs.sh "c3area" "base" | sed 's/^/general-count /'
