# round 2 (8 GPUs): final build, C5 with the two-launch frames and C4
mkdir -p gpurun_out
run() {
  tag=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 "$@" > gpurun_out/r2ah_$tag.json 2> gpurun_out/r2ah_$tag.err
  echo "$tag rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2ah_$tag.json").read())
    print("$tag", "ms", round(d["ms_per_step"], 4), "Mrays/s", round(d["value"]), "first", round(d["detail"]["ms_first_frame"], 4), "spatial", round(d["detail"]["ms_per_step_spatial_order"], 4),
          "e2e ms", round(d["e2e"]["ms_per_step"], 4), "e2e Mrays/s", round(d["e2e"]["value"]), d["frame_hash"], d["frame_hash_ranks_path"])
except Exception as e:
    print("$tag FAILED", e)
PY
}
run c5 --config c5 --steps 5 --warmup 3 --no-cpu
run c4 --steps 20 --warmup 5 --no-cpu
