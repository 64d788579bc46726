N=$1
for f in "" "--no-pipeline"; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 3 --no-cpu $f 2>gpurun_out/pipe_n.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('[$f]', 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'pipelined', d.get('pipelined'), 'graph', d['cuda_graph'], 'decode', round(d['stages']['decode_ms'],4), 'frame', round(d['stages']['march_ms'],4))"
grep -i "pipelined steps\|Error" gpurun_out/pipe_n.err | head -3
done
