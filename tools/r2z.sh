# round 2 (8 GPUs): every frame ordered by the budgeted probe rays (RTCUDA_PROBE=2: what a first frame gets, warm)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu > gpurun_out/r2z_n8_probe_always.json 2> gpurun_out/r2z_n8_probe_always.err
echo "rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r2z_n8_probe_always.json").read())
print("n8 probe-ordered", "ms", round(d["ms_per_step"], 4), "Mrays/s", round(d["value"]), "first", round(d["detail"]["ms_first_frame"], 4), "spatial", round(d["detail"]["ms_per_step_spatial_order"], 4),
      "e2e ms", round(d["e2e"]["ms_per_step"], 4), d["frame_hash"], d["frame_hash_ranks_path"])
PY
