# A/B: constant groups attached to parse CTAs stored by the TMA unit (VR_FILL_TMA=1) vs ordinary stores
VR_FILL_TMA=1 timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_skipping.py tests/test_gpu_lazy.py tests/test_gpu_configs.py -x -q -p no:cacheprovider 2>&1 | tail -3
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'march', round(d['stages']['march_ms'],4), 'step', round(d['ms_per_step'],4), 'frac', round(d['roofline_decode']['frac'],3))"; }
VR_FILL_TMA=0 run plain_attach2
VR_FILL_TMA=1 VR_FILL_ATTACH=2 run tma_attach2
VR_FILL_TMA=1 VR_FILL_ATTACH=3 run tma_attach3
VR_FILL_TMA=1 VR_FILL_ATTACH=4 run tma_attach4
VR_FILL_TMA=1 VR_FILL_ATTACH=3 run tma_attach3_nopipe --no-pipeline
