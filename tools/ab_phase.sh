# parse-only / fill-only decode launches (VR_DECODE_ONLY = 1 / 2; the slab is then incomplete: timing only)
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu --no-pipeline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'march', round(d['stages']['march_ms'],4), 'step', round(d['ms_per_step'],4))"; }
VR_DECODE_ONLY=0 VR_STAGE_KB=0 run all_stage0
VR_DECODE_ONLY=1 VR_STAGE_KB=0 run parse_stage0
VR_DECODE_ONLY=1 VR_STAGE_KB=8 run parse_stage8
VR_DECODE_ONLY=1 VR_STAGE_KB=16 run parse_stage16
VR_DECODE_ONLY=2 VR_STAGE_KB=0 run fill_only
VR_DECODE_ONLY=1 VR_STAGE_KB=0 VR_DECODE_VARIANT=11 run parse_win96
