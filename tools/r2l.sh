mkdir -p gpurun_out
python tools/render_once.py c4 3 > gpurun_out/r2l_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:renderKernelFast -s 2 -c 1 -o gpurun_out/r2l_c4_lean python tools/render_once.py c4 3 > gpurun_out/r2l_ncu1.log 2>&1
RTCUDA_FORCE_RECURSE=1 ncu --set full --clock-control none --import-source on -k regex:renderKernelFast -s 2 -c 1 -o gpurun_out/r2l_c4_rec python tools/render_once.py c4 3 > gpurun_out/r2l_ncu2.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2l_plain.log
