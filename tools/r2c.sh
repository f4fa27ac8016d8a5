# round 2, call C (2 GPUs): the rest of the GPU suite incl. the multi-device paths, bench at N=2 under torchrun
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs -k "cancel or multi or c_abi or non_finite or edge or error" > gpurun_out/r2c_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2c_gpu_tests.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2c_bench_n2.json 2> gpurun_out/r2c_bench_n2.err; echo "bench rc=$?"; tail -4 gpurun_out/r2c_bench_n2.err
python - <<PY
import json
d = json.loads(open("gpurun_out/r2c_bench_n2.json").read())
print(d["ms_per_step"], d["value"], d["detail"]["ms_first_frame"], json.dumps(d["e2e"])[:900], d["frame_hash"], d["frame_hash_ranks_path"])
PY
