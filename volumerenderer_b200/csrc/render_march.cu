// render_march.cu -- stage 2 of the hot path: front-to-back ray marching over the decoded slab.
// One thread per pixel, a warp covers an 8x4 pixel tile so its rays stay coherent in the volume.
//
// Replaces the reference's GLSL fragment shaders and the fixed-function work around them
// (paths relative to /root/reference/volume_renderer/):
//   composite   raycaster.frag:18-86 (march :37-57, compositing :60-72, early termination :77, colour map :82-85)
//   isosurface  isosurface.frag:77-159 (crossing test :126, Bisection :23-42, GetGradient :47-62, PhongLighting :69-75)
//   ray set-up  raycaster.vert:14,20 + UnitBrick.h:54-75 + main.cpp:367-369,392-397 (nearest cube surface point,
//               near/far clip, clear colour white) -- done analytically per pixel instead of rasterising a cube
//   sampler     GL_LINEAR on R8 UNORM with clamp-to-edge (VolumeReader.h:120-127), done manually in fp32
// Arithmetic contract (oracle/render_oracle.cpp): fp32 throughout, every add/mul/
// div/sqrt individually rounded (the __f*_rn intrinsics below are never contracted into FMAs), fused
// multiply-add only in the three lerps of the sampler.  This makes the march bit-reproducible against
// the scalar oracle, so early-termination and iso-crossing decisions agree exactly.
#include "render_common.cuh"

namespace vr {

using namespace dev;

namespace {

template <int MODE, bool OCC>
__global__ void __launch_bounds__(256) march_kernel(const __grid_constant__ RenderParams P, const __grid_constant__ OccView occ,
                                                    const uint8_t* __restrict__ vol, int X, int Y, int Z, float4* __restrict__ out,
                                                    int* __restrict__ hitIndex, unsigned long long* __restrict__ sampleCount) {
  // block = 32x8 pixels = 8 warps; warp w covers an 8x4 tile
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int px = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
  const int py = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
  unsigned n = 0;
  if (px < P.W && py < P.H) {
    Vol v = make_vol(vol, X, Y, Z);
    float4 rgba = make_float4(1.0f, 1.0f, 1.0f, 1.0f);  // clear colour / iso miss (white)
    int hit = -1;
    V3 vUV;
    if (ray_setup(P, px, py, &vUV)) {
      V3 dir, dirStep;
      ray_dir(P, vUV, &dir, &dirStep);
      const V3 inv = recip(dirStep);
      V3 p = vUV;
      if (MODE == 0) {
        float C = 0.0f, A = 0.0f;
        for (int i = 0; i < 300; ++i) {
          p = add(p, dirStep);
          if (!inside(p)) break;
          if (OCC) {  // void cell: the sample is exactly 0 and leaves (C, A) unchanged
            const int k = occ_skip<false>(occ, v, p, dirStep, inv, 299 - i);
            if (k >= 0) {
              for (int j = 0; j < k; ++j) p = add(p, dirStep);
              i += k; n += (unsigned)((1 + k) & P.skipCountMask);
              continue;
            }
          }
          float smp = tex<false>(v, p);
          ++n;
          float prev_alpha = fsub(smp, fmul(smp, A));
          C = fadd(fmul(prev_alpha, smp), C);
          A = fadd(A, fmul(prev_alpha, 0.6f));
          if (A > 0.99f) break;
        }
        rgba = make_float4(clamp01(fsub(1.0f, C)), clamp01(fsub(1.0f, C)), 1.0f, clamp01(A));
      } else {
        for (int i = 0; i < 300; ++i) {
          p = add(p, dirStep);
          if (!inside(p)) break;
          const V3 pn = add(p, dirStep);
          if (OCC) {  // every tap of the look-ahead sample is below the iso value: no crossing can be detected here
            const int k = occ_skip<false>(occ, v, pn, dirStep, inv, 299 - i);
            if (k >= 0) {
              for (int j = 0; j < k; ++j) p = add(p, dirStep);
              i += k; n += (unsigned)((1 + k) & P.skipCountMask);
              continue;
            }
          }
          float smp = tex<false>(v, p);
          float smp2 = tex<false>(v, pn);
          ++n;
          if (fsub(smp, P.iso) < 0.0f && fsub(smp2, P.iso) >= 0.0f) {
            rgba = iso_shade<false>(v, P, p, dir, dirStep);
            hit = i;
            break;
          }
        }
      }
    }
    out[(long long)py * P.W + px] = rgba;
    if (hitIndex) hitIndex[(long long)py * P.W + px] = hit;
  }
  if (sampleCount) {
    unsigned tot = __reduce_add_sync(0xFFFFFFFFu, n);
    if (lane == 0 && tot) atomicAdd(sampleCount, (unsigned long long)tot);
  }
}

}  // namespace

cudaError_t render_launch(const RenderParams& P, const OccView& occ, const uint8_t* vol, int X, int Y, int Z, float4* out,
                          int* hitIndex, unsigned long long* sampleCount, cudaStream_t stream) {
  dim3 grid((P.W + 31) / 32, (P.H + 7) / 8), block(256);
  if (P.mode == 0) {
    if (occ.pyr) march_kernel<0, true><<<grid, block, 0, stream>>>(P, occ, vol, X, Y, Z, out, hitIndex, sampleCount);
    else march_kernel<0, false><<<grid, block, 0, stream>>>(P, occ, vol, X, Y, Z, out, hitIndex, sampleCount);
  } else {
    if (occ.pyr) march_kernel<1, true><<<grid, block, 0, stream>>>(P, occ, vol, X, Y, Z, out, hitIndex, sampleCount);
    else march_kernel<1, false><<<grid, block, 0, stream>>>(P, occ, vol, X, Y, Z, out, hitIndex, sampleCount);
  }
  return cudaGetLastError();
}

}  // namespace vr
