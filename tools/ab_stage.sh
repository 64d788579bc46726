# A/B of the shared-memory stream staging (VR_STAGE_KB) x fill order (VR_FILL_LATE) on the GPU box
timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_skipping.py tests/test_gpu_lazy.py -x -q -p no:cacheprovider 2>&1 | tail -2
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'step', round(d['ms_per_step'],4), 'frac', round(d['roofline_decode']['frac'],3))"; }
VR_STAGE_KB=0 VR_FILL_LATE=0 run stage0_early
VR_STAGE_KB=16 VR_FILL_LATE=0 run stage16_early
VR_STAGE_KB=16 VR_FILL_LATE=1 run stage16_alt
VR_STAGE_KB=16 VR_FILL_LATE=2 run stage16_late
VR_STAGE_KB=8 VR_FILL_LATE=0 run stage8_early
VR_STAGE_KB=8 VR_FILL_LATE=1 run stage8_alt
VR_STAGE_KB=0 python tools/quick_decode_bench.py 16 2>/dev/null | cut -c1-160
VR_STAGE_KB=16 python tools/quick_decode_bench.py 16 2>/dev/null | cut -c1-160
