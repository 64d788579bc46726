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
#include <cuda_runtime.h>
#include <math_constants.h>

#include "render_march.h"

namespace vr {

namespace {

struct V3 { float x, y, z; };
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ V3 add(V3 a, V3 b) { return {fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)}; }
__device__ __forceinline__ V3 sub(V3 a, V3 b) { return {fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)}; }
__device__ __forceinline__ V3 mul(V3 a, V3 b) { return {fmul(a.x, b.x), fmul(a.y, b.y), fmul(a.z, b.z)}; }
__device__ __forceinline__ V3 scale(V3 a, float s) { return {fmul(a.x, s), fmul(a.y, s), fmul(a.z, s)}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z)); }
__device__ __forceinline__ V3 normalize(V3 a) {
  float l = __fsqrt_rn(dot(a, a));
  return {fdiv(a.x, l), fdiv(a.y, l), fdiv(a.z, l)};
}
__device__ __forceinline__ float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }

struct Vol {
  const uint8_t* __restrict__ d;
  int X, Y, Z;
  long long sy, sz;  // strides
  float fX, fY, fZ;
};

// texture(volume, p).r : GL_LINEAR, clamp-to-edge, normalised after filtering
__device__ __forceinline__ float tex(const Vol& v, V3 p) {
  float cx = fsub(fmul(p.x, v.fX), 0.5f), cy = fsub(fmul(p.y, v.fY), 0.5f), cz = fsub(fmul(p.z, v.fZ), 0.5f);
  float fx = floorf(cx), fy = floorf(cy), fz = floorf(cz);
  float wx = fsub(cx, fx), wy = fsub(cy, fy), wz = fsub(cz, fz);
  int x0 = (int)fx, y0 = (int)fy, z0 = (int)fz;
  int x1 = min(x0 + 1, v.X - 1), y1 = min(y0 + 1, v.Y - 1), z1 = min(z0 + 1, v.Z - 1);
  x0 = max(x0, 0); y0 = max(y0, 0); z0 = max(z0, 0);
  x0 = min(x0, v.X - 1); y0 = min(y0, v.Y - 1); z0 = min(z0, v.Z - 1);
  x1 = max(x1, 0); y1 = max(y1, 0); z1 = max(z1, 0);
  const uint8_t* r00 = v.d + (long long)z0 * v.sz + (long long)y0 * v.sy;
  const uint8_t* r10 = v.d + (long long)z0 * v.sz + (long long)y1 * v.sy;
  const uint8_t* r01 = v.d + (long long)z1 * v.sz + (long long)y0 * v.sy;
  const uint8_t* r11 = v.d + (long long)z1 * v.sz + (long long)y1 * v.sy;
  float c000 = (float)__ldg(r00 + x0), c100 = (float)__ldg(r00 + x1);
  float c010 = (float)__ldg(r10 + x0), c110 = (float)__ldg(r10 + x1);
  float c001 = (float)__ldg(r01 + x0), c101 = (float)__ldg(r01 + x1);
  float c011 = (float)__ldg(r11 + x0), c111 = (float)__ldg(r11 + x1);
  float c00 = __fmaf_rn(wx, fsub(c100, c000), c000), c10 = __fmaf_rn(wx, fsub(c110, c010), c010);
  float c01 = __fmaf_rn(wx, fsub(c101, c001), c001), c11 = __fmaf_rn(wx, fsub(c111, c011), c011);
  float c0 = __fmaf_rn(wy, fsub(c10, c00), c00), c1 = __fmaf_rn(wy, fsub(c11, c01), c01);
  return fmul(__fmaf_rn(wz, fsub(c1, c0), c0), 1.0f / 255.0f);
}

__device__ __forceinline__ bool inside(V3 p) {
  return p.x > 0.0f && p.x < 1.0f && p.y > 0.0f && p.y < 1.0f && p.z > 0.0f && p.z < 1.0f;
}

// nearest visible point of the cube [-0.5,0.5]^3 along the pixel's view ray
__device__ __forceinline__ bool ray_setup(const RenderParams& P, int px, int py, V3* vUV) {
  V3 eye = {P.eye[0], P.eye[1], P.eye[2]}, s = {P.s[0], P.s[1], P.s[2]}, u = {P.u[0], P.u[1], P.u[2]},
     f = {P.f[0], P.f[1], P.f[2]};
  float nx = fsub(fdiv(fmul(2.0f, fadd((float)px, 0.5f)), (float)P.W), 1.0f);
  float ny = fsub(fdiv(fmul(2.0f, fadd((float)py, 0.5f)), (float)P.H), 1.0f);
  V3 r = normalize(add(add(scale(s, fdiv(nx, P.tw)), scale(u, fdiv(ny, P.th))), f));
  float tn = -CUDART_INF_F, tf = CUDART_INF_F;
  const float e[3] = {eye.x, eye.y, eye.z}, d[3] = {r.x, r.y, r.z};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float t1 = fdiv(fsub(-0.5f, e[a]), d[a]), t2 = fdiv(fsub(0.5f, e[a]), d[a]);
    tn = fmaxf(tn, fminf(t1, t2));
    tf = fminf(tf, fmaxf(t1, t2));
  }
  if (!(tn <= tf) || tf < 0.0f) return false;
  float cosv = dot(r, f);
  float t = tn;
  float depth = fmul(tn, cosv);
  if (!(depth >= P.znear && depth <= P.zfar)) {
    t = tf;
    depth = fmul(tf, cosv);
    if (!(depth >= P.znear && depth <= P.zfar)) return false;
  }
  V3 q = add(eye, scale(r, t));
  *vUV = add(q, V3{0.5f, 0.5f, 0.5f});
  return true;
}

template <int MODE>
__global__ void __launch_bounds__(256) march_kernel(const __grid_constant__ RenderParams P, const uint8_t* __restrict__ vol,
                                                    int X, int Y, int Z, float4* __restrict__ out,
                                                    int* __restrict__ hitIndex, unsigned long long* __restrict__ sampleCount) {
  // block = 32x8 pixels = 8 warps; warp w covers an 8x4 tile
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int px = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
  const int py = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
  unsigned n = 0;
  if (px < P.W && py < P.H) {
    Vol v = {vol, X, Y, Z, (long long)X, (long long)X * Y, (float)X, (float)Y, (float)Z};
    float4 rgba = make_float4(1.0f, 1.0f, 1.0f, 1.0f);  // clear colour / iso miss (white)
    int hit = -1;
    V3 vUV;
    if (ray_setup(P, px, py, &vUV)) {
      V3 eye = {P.eye[0], P.eye[1], P.eye[2]};
      V3 dir = normalize(sub(sub(vUV, V3{0.5f, 0.5f, 0.5f}), eye));
      V3 dirStep = mul(dir, V3{P.step[0], P.step[1], P.step[2]});
      V3 p = vUV;
      if (MODE == 0) {
        float C = 0.0f, A = 0.0f;
        for (int i = 0; i < 300; ++i) {
          p = add(p, dirStep);
          if (!inside(p)) break;
          float smp = tex(v, p);
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
          float smp = tex(v, p);
          float smp2 = tex(v, add(p, dirStep));
          ++n;
          if (fsub(smp, P.iso) < 0.0f && fsub(smp2, P.iso) >= 0.0f) {
            V3 left = p, right = add(p, dirStep);
#pragma unroll 1
            for (int b = 0; b < 4; ++b) {
              V3 mid = scale(add(right, left), 0.5f);
              if (tex(v, mid) < P.iso) left = mid; else right = mid;
            }
            V3 tc = scale(add(right, left), 0.5f);
            const float D = 0.01f;
            V3 s1 = {tex(v, sub(tc, V3{D, 0, 0})), tex(v, sub(tc, V3{0, D, 0})), tex(v, sub(tc, V3{0, 0, D}))};
            V3 s2 = {tex(v, add(tc, V3{D, 0, 0})), tex(v, add(tc, V3{0, D, 0})), tex(v, add(tc, V3{0, 0, D}))};
            V3 g = sub(s1, s2);
            V3 N = normalize(V3{fdiv(g.x, 2.0f), fdiv(g.y, 2.0f), fdiv(g.z, 2.0f)});
            V3 Vv = {-dir.x, -dir.y, -dir.z};
            V3 L = Vv;
            float diffuse = fmaxf(dot(L, N), 0.0f);
            V3 halfVec = normalize(add(L, Vv));
            float specular = powf(fmaxf(0.00001f, dot(halfVec, N)), 250.0f);
            rgba = make_float4(clamp01(fadd(fmul(diffuse, 0.39f), specular)), clamp01(fadd(fmul(diffuse, 0.58f), specular)),
                               clamp01(fadd(fmul(diffuse, 0.93f), specular)), 1.0f);
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

cudaError_t render_launch(const RenderParams& P, const uint8_t* vol, int X, int Y, int Z, float4* out, int* hitIndex,
                          unsigned long long* sampleCount, cudaStream_t stream) {
  dim3 grid((P.W + 31) / 32, (P.H + 7) / 8), block(256);
  if (P.mode == 0)
    march_kernel<0><<<grid, block, 0, stream>>>(P, vol, X, Y, Z, out, hitIndex, sampleCount);
  else
    march_kernel<1><<<grid, block, 0, stream>>>(P, vol, X, Y, Z, out, hitIndex, sampleCount);
  return cudaGetLastError();
}

}  // namespace vr
