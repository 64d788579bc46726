# A/B: 128-bit stream refills with a register queue (VR_DECODE_VARIANT=12) vs the 32-bit window (default)
VR_DECODE_VARIANT=12 timeout 200 python -m pytest tests/test_gpu_decode.py tests/test_gpu_lazy.py -x -q -p no:cacheprovider 2>&1 | tail -1
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'march', round(d['stages']['march_ms'],4), 'step', round(d['ms_per_step'],4), 'frac', round(d['roofline_decode']['frac'],3))"; }
VR_DECODE_VARIANT=0 run win32
VR_DECODE_VARIANT=12 run win128
VR_DECODE_VARIANT=12 VR_DECODE_ONLY=1 run win128_parse_only
VR_DECODE_VARIANT=12 VR_FILL_LATE=1 run win128_alt
