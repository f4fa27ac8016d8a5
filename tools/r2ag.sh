# round 2: two-launch frames (depth 0 in the lean instantiation, units with child rays deferred to the queue instantiation)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2ag_gpu_tests.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r2ag_gpu_tests.log
for m in 0 1; do RTCUDA_SPLIT=$m bash tools/ab_libs.sh "c5 c3" "base" | sed "s/^/split=$m /"; done
bash tools/ab_libs.sh "c5 c3 c4" "base" | sed "s/^/default /"
