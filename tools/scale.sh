# tools/scale.sh N [extra bench args]: one torchrun bench at N GPUs, prints the headline numbers and the per-rank lines
N=$1; shift
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus $N --steps 30 --warmup 5 "$@" 2>gpurun_out/n$N.err | tee gpurun_out/n$N.json | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('n$N $*', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'spatial', d['config'].get('ms_per_step_spatial_order'), 'e2e', round(d['e2e']['value'],1), 'kernel', round(d['roofline']['kernel_ms'],4))"
grep "^\[rank .\] step" gpurun_out/n$N.err | sort
