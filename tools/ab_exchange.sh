N=$1
for ex in peer nccl; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 3 --no-cpu --exchange $ex 2>gpurun_out/ab_$ex.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$ex', 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'decode', round(d['stages']['decode_ms'],4), 'frame', round(d['stages']['march_ms'],4), 'e2e', round(d['e2e']['value'],1))"
done
