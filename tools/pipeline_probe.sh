#!/bin/bash
# Where does the pipelined step (decode of slab t+1 beside the frame of slab t) lose its overlap?  Timing probes of bench.py:
# decode_only=1 (parse units only) / 2 (constant groups only), split launch on / off, pipelined / sequential.
out=gpurun_out/r2_pipeline_probe.jsonl
for dbg in "$@"; do
  for pl in "" "--no-pipeline"; do
    python bench.py --steps 100 --no-cpu --no-extras $dbg $pl 2>>gpurun_out/probe_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(json.dumps({'debug':d['debug_knobs'],'pipelined':d['pipelined'],'ms_per_step':d['ms_per_step'],'decode_ms':d['stages']['decode_ms'],'march_ms':d['stages']['march_ms']}))" >> $out
  done
done
cat $out
