mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2h_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2h_gpu_tests.log
bash tools/ab_libs.sh "c3 c5 c4 c1 c2" "base"
