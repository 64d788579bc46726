for f in "" "--no-pipeline" "--no-graph"; do
timeout 300 python bench.py --steps 100 --warmup 3 --no-cpu $f 2>gpurun_out/pipe.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('[$f]', 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'pipelined', d.get('pipelined'), 'graph', d['cuda_graph'], 'decode', round(d['stages']['decode_ms'],4), 'frame', round(d['stages']['march_ms'],4))"
grep -i "pipelined\|error" gpurun_out/pipe.err | head -3
done
