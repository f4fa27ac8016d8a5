# round 2 (2 GPUs): the multi-device tests of the GPU suite, bench at N=2 under torchrun, reference arm
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs -k "multi" > gpurun_out/r2v_gpu_tests_2gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2v_gpu_tests_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2v_bench_n2.json 2> gpurun_out/r2v_bench_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/r2v_bench_n2.err
python - <<PY
import json
d = json.loads(open("gpurun_out/r2v_bench_n2.json").read())
print(d["ms_per_step"], d["value"], d["detail"]["ms_first_frame"], d["detail"].get("ms_per_step_spatial_order"), json.dumps(d["e2e"])[:600], d["frame_hash"], d["frame_hash_ranks_path"])
PY
