timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_skipping.py -x -q 2>&1 | tail -2
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', round(d['stages']['decode_ms'],4), round(d['ms_per_step'],4), round(d['roofline_decode']['frac'],3))"; }
VR_FILL_ATTACH=2 VR_FILL_LATE=1 run a2_alt
VR_FILL_ATTACH=2 VR_FILL_LATE=0 run a2_early
VR_FILL_ATTACH=2 VR_FILL_LATE=2 run a2_late
VR_FILL_ATTACH=4 VR_FILL_LATE=1 run a4_alt
VR_FILL_ATTACH=3 VR_FILL_LATE=1 run a3_alt
VR_FILL_ATTACH=1 VR_FILL_LATE=1 run a1_alt
