# round 2: where engine creation spends its time (RTCUDA_TRACE), occupancy variants on the kernels with recursion, memcheck of the beam walk
mkdir -p gpurun_out
RTCUDA_TRACE=1 python tools/build_bench.py c4 2 > gpurun_out/r2aa_build_bench.json 2> gpurun_out/r2aa_trace.log; echo "trace rc=$?"; grep "rtcuda trace" gpurun_out/r2aa_trace.log | tail -40
bash tools/ab_libs.sh "c3 c5" "base b24 b16"
bash tools/ab_libs.sh "c3glossy c3area" "base gen3 gen1"
cat > /tmp/small_beam.py <<'PY'
import os, sys, tempfile
sys.path.insert(0, os.getcwd())
os.environ["RTCUDA_WIDE"] = "1"
from myraytracer_b200 import RayTracerEngine, scenegen
d = tempfile.mkdtemp()
for path in (scenegen.config_c2(d, n=64, width=160, height=96), scenegen.config_c5(d, n=32, width=128, height=72, samples=4, meshes=3)):
    e = RayTracerEngine(path)
    r = e.render(0)
    print(path, e.scene_info()["wideNodes"], r.stats["rays"])
PY
compute-sanitizer --tool memcheck --error-exitcode 7 python /tmp/small_beam.py > gpurun_out/r2aa_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -4 gpurun_out/r2aa_memcheck.log
