// render_common.cuh -- device helpers shared by the ray-marching kernels (render_march.cu,
// render_sortlast.cu): fp32 vector ops with one rounding per operation, the manual GL_LINEAR sampler
// and the analytic replacement of the rasterised unit cube.
//
// Arithmetic contract (oracle/render_oracle.cpp): fp32 throughout, every add/mul/
// div/sqrt individually rounded (the __f*_rn intrinsics are never contracted into FMAs), fused
// multiply-add only in the three lerps of the sampler.  This makes the march bit-reproducible against
// the scalar oracle, so early-termination and iso-crossing decisions agree exactly.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "occupancy.h"
#include "render_march.h"

namespace vr {
namespace dev {

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

// The texture: G = extent of the whole volume (what the unit cube maps to).  A device slab stores the
// box [o, o+S) of it (o = 0, S = G for a whole volume).  own = [lo, hi) is the box of base texels this
// slab answers for in a sort-last frame.
struct Vol {
  const uint8_t* __restrict__ d;
  int GX, GY, GZ;
  int ox, oy, oz;
  int SX, SY, SZ;
  long long sy, sz;  // strides of the stored array
  float fX, fY, fZ;
  int lox, loy, loz, hix, hiy, hiz;
};

__device__ __forceinline__ Vol make_vol(const uint8_t* d, int X, int Y, int Z) {
  Vol v;
  v.d = d; v.GX = X; v.GY = Y; v.GZ = Z; v.ox = v.oy = v.oz = 0; v.SX = X; v.SY = Y; v.SZ = Z;
  v.sy = X; v.sz = (long long)X * Y; v.fX = (float)X; v.fY = (float)Y; v.fZ = (float)Z;
  v.lox = v.loy = v.loz = 0; v.hix = X; v.hiy = Y; v.hiz = Z;
  return v;
}
__device__ __forceinline__ Vol make_vol(const SlabGeom& g, const uint8_t* d) {
  Vol v;
  v.d = d; v.GX = g.G[0]; v.GY = g.G[1]; v.GZ = g.G[2]; v.ox = g.o[0]; v.oy = g.o[1]; v.oz = g.o[2];
  v.SX = g.S[0]; v.SY = g.S[1]; v.SZ = g.S[2];
  v.sy = g.S[0]; v.sz = (long long)g.S[0] * g.S[1]; v.fX = (float)g.G[0]; v.fY = (float)g.G[1]; v.fZ = (float)g.G[2];
  v.lox = g.lo[0]; v.loy = g.lo[1]; v.loz = g.lo[2]; v.hix = g.hi[0]; v.hiy = g.hi[1]; v.hiz = g.hi[2];
  return v;
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// texture(volume, p).r : GL_LINEAR, clamp-to-edge, normalised after filtering.  SLAB: taps are
// translated into the stored box (and clamped to it, which only matters if the halo is too thin).
template <bool SLAB>
__device__ __forceinline__ float tex(const Vol& v, V3 p) {
  float cx = fsub(fmul(p.x, v.fX), 0.5f), cy = fsub(fmul(p.y, v.fY), 0.5f), cz = fsub(fmul(p.z, v.fZ), 0.5f);
  float fx = floorf(cx), fy = floorf(cy), fz = floorf(cz);
  float wx = fsub(cx, fx), wy = fsub(cy, fy), wz = fsub(cz, fz);
  int x0 = (int)fx, y0 = (int)fy, z0 = (int)fz;
  int x1 = clampi(x0 + 1, 0, v.GX - 1), y1 = clampi(y0 + 1, 0, v.GY - 1), z1 = clampi(z0 + 1, 0, v.GZ - 1);
  x0 = clampi(x0, 0, v.GX - 1); y0 = clampi(y0, 0, v.GY - 1); z0 = clampi(z0, 0, v.GZ - 1);
  if (SLAB) {
    x0 = clampi(x0 - v.ox, 0, v.SX - 1); x1 = clampi(x1 - v.ox, 0, v.SX - 1);
    y0 = clampi(y0 - v.oy, 0, v.SY - 1); y1 = clampi(y1 - v.oy, 0, v.SY - 1);
    z0 = clampi(z0 - v.oz, 0, v.SZ - 1); z1 = clampi(z1 - v.oz, 0, v.SZ - 1);
  }
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

// Sort-last ownership: the sample at p belongs to the slab whose box holds its base texel
// (floor(p*G - 0.5), clamped to the texture).  Every sample of the lattice has exactly one owner.
__device__ __forceinline__ bool owned(const Vol& v, V3 p) {
  int x0 = clampi((int)floorf(fsub(fmul(p.x, v.fX), 0.5f)), 0, v.GX - 1);
  int y0 = clampi((int)floorf(fsub(fmul(p.y, v.fY), 0.5f)), 0, v.GY - 1);
  int z0 = clampi((int)floorf(fsub(fmul(p.z, v.fZ), 0.5f)), 0, v.GZ - 1);
  return x0 >= v.lox && x0 < v.hix && y0 >= v.loy && y0 < v.hiy && z0 >= v.loz && z0 < v.hiz;
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


// ---- empty-space skipping (occupancy.h) -------------------------------------------------------------
// q = position whose 8 taps decide the sample (composite: the sample itself; isosurface: the look-ahead
// p + dirStep).  Returns -1 when the cell's bound does not prove the sample void; otherwise the number
// k >= 0 of FURTHER lattice steps q + j*d (j = 1..k) that provably stay inside the largest void cube of cells
// around q's cell (and inside the texture), capped by `budget`.  inv = 1/d per axis.  The estimate uses
// real-number geometry with a margin far above the accumulated rounding of the lattice (300 additions of
// magnitude <= 1: < 2e-5); the positions themselves are still produced by the caller's repeated additions.
template <bool SLAB>
__device__ __forceinline__ int occ_skip(const OccView& o, const Vol& v, V3 q, V3 d, V3 inv, int budget) {
  int bx = clampi((int)floorf(fsub(fmul(q.x, v.fX), 0.5f)), 0, v.GX - 1);
  int by = clampi((int)floorf(fsub(fmul(q.y, v.fY), 0.5f)), 0, v.GY - 1);
  int bz = clampi((int)floorf(fsub(fmul(q.z, v.fZ), 0.5f)), 0, v.GZ - 1);
  if (SLAB) { bx = clampi(bx - v.ox, 0, v.SX - 1); by = clampi(by - v.oy, 0, v.SY - 1); bz = clampi(bz - v.oz, 0, v.SZ - 1); }
  const int cx = (bx + o.sx) >> VR_OCC_LX, cy = (by + o.sy) >> VR_OCC_LY, cz = (bz + o.sz) >> VR_OCC_LZ;
  const int R = (int)__ldg(o.lvl + (cz * o.ny + cy) * o.nx + cx);  // radius (cells) of the void cube of cells around the cell
  if (R == VR_OCC_SOLID) return -1;
  if (R < 1) return 0;  // a single cell: about one step, not worth the geometry
  // the cube of cells [c - R, c + R] as a box of base texels in global voxel coordinates, clipped to the stored box
  const float kMargin = 1.0e-4f;
  float t = 1.0e9f;
  const int cc[3] = {cx, cy, cz}, sh[3] = {o.sx, o.sy, o.sz}, lg[3] = {VR_OCC_LX, VR_OCC_LY, VR_OCC_LZ};
  const int S[3] = {v.SX, v.SY, v.SZ}, org[3] = {SLAB ? v.ox : 0, SLAB ? v.oy : 0, SLAB ? v.oz : 0}, G[3] = {v.GX, v.GY, v.GZ};
  const float qq[3] = {q.x, q.y, q.z}, dd[3] = {d.x, d.y, d.z}, ii[3] = {inv.x, inv.y, inv.z}, rG[3] = {1.0f / v.fX, 1.0f / v.fY, 1.0f / v.fZ};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const int lo = max(((cc[a] - R) << lg[a]) - sh[a], 0) + org[a];
    const int hi = min(((cc[a] + R + 1) << lg[a]) - sh[a], S[a]) + org[a];
    const float plo = lo == 0 ? 0.0f : ((float)lo + 0.5f) * rG[a];
    const float phi = hi >= G[a] ? 1.0f : ((float)hi + 0.5f) * rG[a];
    if (dd[a] > 0.0f) t = fminf(t, (phi - kMargin - qq[a]) * ii[a]);
    else if (dd[a] < 0.0f) t = fminf(t, (plo + kMargin - qq[a]) * ii[a]);
    else if (!(qq[a] > plo + kMargin && qq[a] < phi - kMargin)) t = 0.0f;
  }
  const int k = (int)floorf(t) - 1;
  return min(max(k, 0), budget);
}
__device__ __forceinline__ V3 recip(V3 d) { return {1.0f / d.x, 1.0f / d.y, 1.0f / d.z}; }


// Sort-last: conservative index of the first lattice sample a slab can own.  Samples i < first_owned() are
// certainly not owned (real-number geometry with a margin far above the lattice's accumulated rounding, and
// one extra step), so the caller replays them with the three additions only -- no ownership or inside
// test: once a coordinate has left (0,1) it never returns, so the regular loop's inside test still ends
// the ray correctly when it resumes.  300 = the ray never meets the slab.
__device__ __forceinline__ int first_owned(const Vol& v, V3 p0, V3 d) {
  const float kMargin = 1.0e-4f;
  const float pp[3] = {p0.x, p0.y, p0.z}, dd[3] = {d.x, d.y, d.z}, fG[3] = {v.fX, v.fY, v.fZ};
  const int lo[3] = {v.lox, v.loy, v.loz}, hi[3] = {v.hix, v.hiy, v.hiz}, G[3] = {v.GX, v.GY, v.GZ};
  float first = 0.0f;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    if (lo[a] <= 0 && hi[a] >= G[a]) continue;  // the slab spans this axis
    const float a0 = lo[a] <= 0 ? -1.0f : ((float)lo[a] + 0.5f) / fG[a];   // owned  <=>  a0 <= p < a1 (up to rounding)
    const float a1 = hi[a] >= G[a] ? 2.0f : ((float)hi[a] + 0.5f) / fG[a];
    if (dd[a] > 0.0f) {
      if (pp[a] > a1 + kMargin) return 300;
      first = fmaxf(first, (a0 - kMargin - pp[a]) / dd[a]);
    } else if (dd[a] < 0.0f) {
      if (pp[a] < a0 - kMargin) return 300;
      first = fmaxf(first, (a1 + kMargin - pp[a]) / dd[a]);
    } else if (pp[a] < a0 - kMargin || pp[a] > a1 + kMargin) {
      return 300;
    }
  }
  // sample i sits at p0 + (i + 1) * d
  const int i = (int)floorf(fminf(first, 400.0f)) - 2;
  return min(max(i, 0), 300);
}

// dir and dirStep of the fragment (raycaster.frag:27,31 / isosurface.frag:88,92)
__device__ __forceinline__ void ray_dir(const RenderParams& P, V3 vUV, V3* dir, V3* dirStep) {
  V3 eye = {P.eye[0], P.eye[1], P.eye[2]};
  *dir = normalize(sub(sub(vUV, V3{0.5f, 0.5f, 0.5f}), eye));
  *dirStep = mul(*dir, V3{P.step[0], P.step[1], P.step[2]});
}

// isosurface.frag:126-157 once a crossing between p and p + dirStep has been detected:
// Bisection (:23-42), GetGradient (:47-62), PhongLighting (:69-75)
template <bool SLAB>
__device__ __forceinline__ float4 iso_shade(const Vol& v, const RenderParams& P, V3 p, V3 dir, V3 dirStep) {
  V3 left = p, right = add(p, dirStep);
#pragma unroll 1
  for (int b = 0; b < 4; ++b) {
    V3 mid = scale(add(right, left), 0.5f);
    if (tex<SLAB>(v, mid) < P.iso) left = mid; else right = mid;
  }
  V3 tc = scale(add(right, left), 0.5f);
  const float D = 0.01f;
  V3 s1 = {tex<SLAB>(v, sub(tc, V3{D, 0, 0})), tex<SLAB>(v, sub(tc, V3{0, D, 0})), tex<SLAB>(v, sub(tc, V3{0, 0, D}))};
  V3 s2 = {tex<SLAB>(v, add(tc, V3{D, 0, 0})), tex<SLAB>(v, add(tc, V3{0, D, 0})), tex<SLAB>(v, add(tc, V3{0, 0, D}))};
  V3 g = sub(s1, s2);
  V3 N = normalize(V3{fdiv(g.x, 2.0f), fdiv(g.y, 2.0f), fdiv(g.z, 2.0f)});
  V3 Vv = {-dir.x, -dir.y, -dir.z};
  V3 L = Vv;
  float diffuse = fmaxf(dot(L, N), 0.0f);
  V3 halfVec = normalize(add(L, Vv));
  float specular = powf(fmaxf(0.00001f, dot(halfVec, N)), 250.0f);
  return make_float4(clamp01(fadd(fmul(diffuse, 0.39f), specular)), clamp01(fadd(fmul(diffuse, 0.58f), specular)),
                     clamp01(fadd(fmul(diffuse, 0.93f), specular)), 1.0f);
}

}  // namespace dev
}  // namespace vr
