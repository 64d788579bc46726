#!/bin/bash
cd /root/repo
cp tools/ab/libvr_b200_new.so volumerenderer_b200/libvr_b200.so
timeout 900 python -m pytest tests/test_gpu_decode.py tests/test_gpu_lazy.py tests/test_gpu_configs.py tests/test_gpu_indexed.py -x -q -m gpu > gpurun_out/st_tests.log 2>&1; tail -3 gpurun_out/st_tests.log
for v in new old new old; do
  cp tools/ab/libvr_b200_$v.so volumerenderer_b200/libvr_b200.so
  timeout 300 python tools/decode_ab.py --workloads slab,config3,noise --kernels persistent --fill-mode 2 --iters 30 > gpurun_out/st_$v.$RANDOM.jsonl 2> gpurun_out/st_$v.err || echo "FAIL $v"
done
cp tools/ab/libvr_b200_new.so volumerenderer_b200/libvr_b200.so
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/st_bench.json 2> gpurun_out/st_bench.err; cut -c1-300 gpurun_out/st_bench.json
for f in gpurun_out/st_new.*.jsonl gpurun_out/st_old.*.jsonl; do echo $f; python - "$f" <<'PY'
import json,sys
for l in open(sys.argv[1]):
    try: d=json.loads(l)
    except Exception: continue
    if 'ms' in d: print(' ', d['workload'], d['part'], round(d['ms'],4), round(d['min_ms'],4))
PY
done
