# A/B: two-level chain table in shared memory (VR_DECODE_VARIANT=13), alone and with staged streams / TMA-carried fills
VR_DECODE_VARIANT=13 VR_STAGE_KB=16 timeout 40 python -m pytest tests/test_gpu_decode.py -x -q -p no:cacheprovider 2>&1 | tail -1
run() { timeout 40 python bench.py --steps 30 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'march', round(d['stages']['march_ms'],4), 'step', round(d['ms_per_step'],4), 'frac', round(d['roofline_decode']['frac'],3))"; }
VR_DECODE_VARIANT=13 VR_STAGE_KB=16 VR_FILL_TMA=1 run lut2_stage16_tma
VR_DECODE_VARIANT=13 run lut2
VR_DECODE_VARIANT=13 VR_STAGE_KB=12 run lut2_stage12
VR_DECODE_VARIANT=13 VR_DECODE_ONLY=1 run lut2_parse_only
