run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'march', round(d['stages']['march_ms'],4), 'step', round(d['ms_per_step'],4), 'frac', round(d['roofline_decode']['frac'],3))"; }
VR_FILL_TMA=1 VR_FILL_ATTACH=2 VR_STAGE_KB=8 run tma2_stage8
VR_FILL_TMA=1 VR_FILL_ATTACH=2 VR_STAGE_KB=16 run tma2_stage16
VR_FILL_TMA=1 VR_FILL_ATTACH=3 VR_STAGE_KB=16 run tma3_stage16
VR_FILL_TMA=1 VR_FILL_ATTACH=1 VR_STAGE_KB=0 run tma1_stage0
