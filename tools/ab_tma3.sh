# TMA fill throughput by box size (VR_TMA_ROWS) with the parsers switched off (VR_DECODE_ONLY=3), then the full launch
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'march', round(d['stages']['march_ms'],4), 'step', round(d['ms_per_step'],4), 'frac', round(d['roofline_decode']['frac'],3))"; }
VR_DECODE_ONLY=3 VR_FILL_TMA=0 run fills_plain_noparse
VR_DECODE_ONLY=3 VR_FILL_TMA=1 VR_TMA_ROWS=8 run fills_tma8_noparse
VR_DECODE_ONLY=3 VR_FILL_TMA=1 VR_TMA_ROWS=32 run fills_tma32_noparse
VR_FILL_TMA=1 VR_TMA_ROWS=32 run tma32
VR_FILL_TMA=1 VR_TMA_ROWS=16 run tma16
VR_FILL_TMA=1 VR_TMA_ROWS=32 timeout 200 python -m pytest tests/test_gpu_decode.py -x -q -p no:cacheprovider 2>&1 | tail -1
