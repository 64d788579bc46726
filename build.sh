#!/bin/bash
# Builds libvr_b200.so (sm_100a) in-tree.  Usage: ./build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")/volumerenderer_b200"
SRC="csrc/vr_capi.cu csrc/vr_capi_device.cu csrc/kd_decode.cu csrc/kd_build.cu csrc/kd_stream.cpp csrc/kd_kdx.cpp csrc/kd_mrt.cpp csrc/kd_compress.cpp csrc/render_march.cu csrc/render_sortlast.cu csrc/vr_capi_render.cu csrc/vr_capi_sortlast.cu csrc/vr_capi_peer.cu csrc/vr_capi_frame.cu csrc/occupancy.cu csrc/vr_capi_occupancy.cu csrc/vr_capi_io.cu csrc/vr_capi_noalloc.cu"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-O3,-pthread \
     -ccbin /usr/bin/g++ -shared -o libvr_b200.so $SRC "$@"
echo "built $(pwd)/libvr_b200.so"
