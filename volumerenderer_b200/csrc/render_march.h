// render_march.h -- launch interface of the ray-marching kernels (render_march.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vr {
// Same field order as the oracle's Params (oracle/render_oracle.cpp): camera basis precomputed on the host.
struct RenderParams {
  float eye[3], s[3], u[3], f[3];
  float tw, th, znear, zfar;
  float step[3];
  float iso;
  int W, H, mode;  // mode 0 composite, 1 isosurface
};
cudaError_t render_launch(const RenderParams& P, const uint8_t* vol, int X, int Y, int Z, float4* out, int* hitIndex,
                          unsigned long long* sampleCount, cudaStream_t stream);
}  // namespace vr
