# A/B: streams prefetched into L2 (VR_L2_PREFETCH=1) and constant fills as streaming stores (library built with -DVR_FILL_CS=1)
run() { timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu $2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', 'decode', round(d['stages']['decode_ms'],4), 'march', round(d['stages']['march_ms'],4), 'step', round(d['ms_per_step'],4), 'frac', round(d['roofline_decode']['frac'],3))"; }
VR_L2_PREFETCH=0 run base
VR_L2_PREFETCH=1 run prefetch
VR_L2_PREFETCH=1 VR_DECODE_ONLY=1 run prefetch_parse_only
cp volumerenderer_b200/libvr_b200.so /tmp/libvr_default.so
cp tools/libvr_b200_cs.so.bin volumerenderer_b200/libvr_b200.so
VR_L2_PREFETCH=0 run cs
VR_L2_PREFETCH=1 run cs_prefetch
VR_L2_PREFETCH=1 VR_FILL_LATE=1 run cs_prefetch_alt
VR_L2_PREFETCH=0 VR_DECODE_ONLY=2 run cs_fill_only
VR_L2_PREFETCH=1 timeout 200 python -m pytest tests/test_gpu_decode.py -x -q -p no:cacheprovider 2>&1 | tail -2
cp /tmp/libvr_default.so volumerenderer_b200/libvr_b200.so
