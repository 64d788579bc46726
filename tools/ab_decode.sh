timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_skipping.py -x -q 2>&1 | tail -2
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['stages']['decode_ms'],4), round(d['ms_per_step'],4), round(d['roofline_decode']['frac'],3))"; }
for o in 0 1 2; do VR_DECODE_ORDER=$o run order$o; done
timeout 200 python tools/quick_decode_bench.py 16 2>&1 | tail -4 | cut -c1-160
