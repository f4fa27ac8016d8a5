# A/B of kernel build variants (librtcuda_<tag>.so next to librtcuda.so): frame time of the given configs per variant
#   bash tools/ab_libs.sh "c3 c5 c4" "b1 b2 b1n b4n"
mkdir -p gpurun_out
for tag in $2; do
  lib=$PWD/myraytracer_b200/librtcuda_$tag.so
  [ "$tag" = base ] && lib=$PWD/myraytracer_b200/librtcuda.so
  for c in $1; do
    RTCUDA_LIB=$lib python bench.py --config $c --steps 10 --warmup 3 --no-cpu > gpurun_out/ab_${tag}_$c.json 2> gpurun_out/ab_${tag}_$c.err
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ab_${tag}_$c.json").read())
    print("$tag", "$c", round(d["ms_per_step"], 3), round(d["detail"]["ms_first_frame"], 3), round(d["e2e"]["ms_per_step"], 3), d["frame_hash"], "wide", d.get("wide_nodes"))
except Exception as e:
    print("$tag", "$c", "FAILED", e)
PY
  done
done
