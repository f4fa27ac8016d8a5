# round 2 (8-GPU box): every frame ordered by probe rays (RTCUDA_PROBE=2: what a first frame gets, warm), and N=4
mkdir -p gpurun_out
run() {  # tag, nproc, bench args
  tag=$1; n=$2; shift; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $n --steps 20 --warmup 5 --no-cpu "$@" > gpurun_out/r2y_$tag.json 2> gpurun_out/r2y_$tag.err
  echo "$tag rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2y_$tag.json").read())
    print("$tag", "ms", round(d["ms_per_step"], 4), "Mrays/s", round(d["value"]), "first", round(d["detail"]["ms_first_frame"], 4), "spatial", round(d["detail"]["ms_per_step_spatial_order"], 4),
          "e2e ms", round(d["e2e"]["ms_per_step"], 4), "e2e Mrays/s", round(d["e2e"]["value"]), d["frame_hash"], d["frame_hash_ranks_path"])
except Exception as e:
    print("$tag FAILED", e)
PY
}
RTCUDA_PROBE=2 run n8_probe_always 8
RTCUDA_PROBE=2 run n4_probe_always 4
run n4 4
