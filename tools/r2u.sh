# round 2: FP64 division / sqrt / pow out of line (RT_NOINLINE_MATH=1) against inlined, every bench config; GPU suite
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2u_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r2u_gpu_tests.log
bash tools/ab_libs.sh "c1 c2 c3 c4 c5 c3glossy c3area" "base nim"
