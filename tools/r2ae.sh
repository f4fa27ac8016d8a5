# round 2: beam walk for closest-hit packets only (RTCUDA_WIDE=2) / any-hit packets only (=3) / both (=1) / off
mkdir -p gpurun_out
for m in 0 1 2 3; do RTCUDA_WIDE=$m bash tools/ab_libs.sh "c4 c5 c2" "base" | sed "s/^/wide=$m /"; done
