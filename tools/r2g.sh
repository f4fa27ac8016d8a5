mkdir -p gpurun_out
export RTCUDA_LIB=$PWD/myraytracer_b200/librtcuda_b1.so
python tools/render_once.py c3 3 > gpurun_out/r2g_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:renderKernelFast -s 2 -c 1 -o gpurun_out/r2g_c3 python tools/render_once.py c3 3 > gpurun_out/r2g_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2g_plain.log
