# round 2: the two-rank bench line after the builder / flatten changes (one process per GPU, torchrun)
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2ar_bench_n2.json 2> gpurun_out/r2ar_bench_n2.err; echo "bench n2 rc=$?"
grep -h "scene ready" gpurun_out/r2ar_bench_n2.err | cut -c1-100
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2ar_bench_n2.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step", "n_gpus", "build_ms", "gpu_launches", "scaling")}, "e2e", d["e2e"]["value"], "parity", d.get("parity", {}).get("pass"))
PY
