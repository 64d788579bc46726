// render_oracle.cpp -- TEST INFRASTRUCTURE ONLY (oracle).  Not part of the product: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may load it.
//
// Scalar C++ restatement of the reference's GLSL render path (no CPU renderer exists in the
// reference; its GL path is unavailable on this box), following SURVEY Appendix B:
//   ray set-up   raycaster.vert:14,20 + UnitBrick.h:54-75 + main.cpp:367-369,392-397 (GLM lookAt /
//                perspectiveFov, depth test GL_LESS, no culling, clear colour white)
//   march        raycaster.frag:21-57 / isosurface.frag:82-117
//   composite    raycaster.frag:60-85
//   isosurface   isosurface.frag:23-42 (Bisection), :47-62 (GetGradient), :69-75 (PhongLighting), :120-157
//   sampler      GL_LINEAR on an R8 UNORM 3-D texture, clamp-to-edge (VolumeReader.h:120-127; SURVEY C-6)
// PARITY UNPINNED: the reference holds no golden images and cannot be run here (GLSL + GL driver);
// this file is the contract the CUDA renderer is compared with.
// Pinned choices (SURVEY C-5..C-8): vFragColor starts at 0; clamp-to-edge; normalisation by 1/255 is
// applied after filtering; NaN normals go through fmaxf (=> black); colours are clamped to [0,1] on
// write.  All arithmetic is fp32, compiled with -ffp-contract=off; fused multiply-adds appear only
// where written explicitly (the three lerps of the sampler), so a GPU can reproduce it bit for bit.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct V3 { float x, y, z; };
static inline V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 mul(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 scale(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline V3 normalize(V3 a) { float l = std::sqrt(dot(a, a)); return {a.x / l, a.y / l, a.z / l}; }

struct Params {            // mirrors vr_render_params of the product ABI (plain floats)
  float eye[3], s[3], u[3], f[3];  // camera position and GLM lookAt basis (right, up, forward)
  float tw, th;                    // perspectiveFov [0][0], [1][1]
  float znear, zfar;
  float step[3];                   // step_size uniform
  float iso;                       // isoValue uniform
  int W, H;
  int mode;                        // 0 composite, 1 isosurface
};

struct Vol { const uint8_t* d; int64_t X, Y, Z; };

static inline float texel(const Vol& v, int64_t x, int64_t y, int64_t z) {
  x = x < 0 ? 0 : (x > v.X - 1 ? v.X - 1 : x);
  y = y < 0 ? 0 : (y > v.Y - 1 ? v.Y - 1 : y);
  z = z < 0 ? 0 : (z > v.Z - 1 ? v.Z - 1 : z);
  return (float)v.d[x + v.X * (y + v.Y * z)];
}

// texture(volume, p).r
static inline float tex(const Vol& v, V3 p) {
  float cx = p.x * (float)v.X - 0.5f, cy = p.y * (float)v.Y - 0.5f, cz = p.z * (float)v.Z - 0.5f;
  float fx = std::floor(cx), fy = std::floor(cy), fz = std::floor(cz);
  float wx = cx - fx, wy = cy - fy, wz = cz - fz;
  int64_t x0 = (int64_t)fx, y0 = (int64_t)fy, z0 = (int64_t)fz;
  float c000 = texel(v, x0, y0, z0), c100 = texel(v, x0 + 1, y0, z0);
  float c010 = texel(v, x0, y0 + 1, z0), c110 = texel(v, x0 + 1, y0 + 1, z0);
  float c001 = texel(v, x0, y0, z0 + 1), c101 = texel(v, x0 + 1, y0, z0 + 1);
  float c011 = texel(v, x0, y0 + 1, z0 + 1), c111 = texel(v, x0 + 1, y0 + 1, z0 + 1);
  float c00 = std::fmaf(wx, c100 - c000, c000), c10 = std::fmaf(wx, c110 - c010, c010);
  float c01 = std::fmaf(wx, c101 - c001, c001), c11 = std::fmaf(wx, c111 - c011, c011);
  float c0 = std::fmaf(wy, c10 - c00, c00), c1 = std::fmaf(wy, c11 - c01, c01);
  return std::fmaf(wz, c1 - c0, c0) * (1.0f / 255.0f);
}

static inline bool inside(V3 p) {  // dot(sign(p-0), sign(1-p)) == 3  <=>  0 < p < 1 on every axis
  return p.x > 0.0f && p.x < 1.0f && p.y > 0.0f && p.y < 1.0f && p.z > 0.0f && p.z < 1.0f;
}

static inline float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }

// B.1: nearest visible point of the cube [-0.5,0.5]^3 along the pixel's view ray; false = no fragment
static bool ray_setup(const Params& P, int px, int py, V3* vUV) {
  V3 eye = {P.eye[0], P.eye[1], P.eye[2]}, s = {P.s[0], P.s[1], P.s[2]}, u = {P.u[0], P.u[1], P.u[2]},
     f = {P.f[0], P.f[1], P.f[2]};
  float nx = (2.0f * ((float)px + 0.5f)) / (float)P.W - 1.0f;
  float ny = (2.0f * ((float)py + 0.5f)) / (float)P.H - 1.0f;
  V3 r = normalize(add(add(scale(s, nx / P.tw), scale(u, ny / P.th)), f));
  float tn = -INFINITY, tf = INFINITY;
  const float e[3] = {eye.x, eye.y, eye.z}, d[3] = {r.x, r.y, r.z};
  for (int a = 0; a < 3; ++a) {
    float t1 = (-0.5f - e[a]) / d[a], t2 = (0.5f - e[a]) / d[a];
    float lo = std::fmin(t1, t2), hi = std::fmax(t1, t2);
    tn = std::fmax(tn, lo);
    tf = std::fmin(tf, hi);
  }
  if (!(tn <= tf) || tf < 0.0f) return false;
  float cosv = dot(r, f);
  float t = tn;
  float depth = tn * cosv;
  if (!(depth >= P.znear && depth <= P.zfar)) {  // front face clipped (or eye inside): back face
    t = tf;
    depth = tf * cosv;
    if (!(depth >= P.znear && depth <= P.zfar)) return false;
  }
  V3 q = add(eye, scale(r, t));
  *vUV = add(q, V3{0.5f, 0.5f, 0.5f});
  return true;
}

static void shade_pixel(const Params& P, const Vol& vol, int px, int py, float* rgba, int64_t* samples,
                        int32_t* hitIndex) {
  V3 vUV;
  if (hitIndex) *hitIndex = -1;
  if (!ray_setup(P, px, py, &vUV)) {  // clear colour (main.cpp:392, clamped)
    rgba[0] = rgba[1] = rgba[2] = rgba[3] = 1.0f;
    return;
  }
  V3 eye = {P.eye[0], P.eye[1], P.eye[2]};
  V3 dir = normalize(sub(sub(vUV, V3{0.5f, 0.5f, 0.5f}), eye));
  V3 dirStep = mul(dir, V3{P.step[0], P.step[1], P.step[2]});
  V3 p = vUV;
  int64_t n = 0;
  if (P.mode == 0) {
    float C = 0.0f, A = 0.0f;
    for (int i = 0; i < 300; ++i) {
      p = add(p, dirStep);
      if (!inside(p)) break;
      float smp = tex(vol, p);
      ++n;
      float prev_alpha = smp - (smp * A);
      C = prev_alpha * smp + C;
      A += prev_alpha * 0.6f;
      if (A > 0.99f) break;
    }
    rgba[0] = clamp01(1.0f - C);
    rgba[1] = clamp01(1.0f - C);
    rgba[2] = 1.0f;  // 255 clamped
    rgba[3] = clamp01(A);
  } else {
    rgba[0] = rgba[1] = rgba[2] = rgba[3] = 1.0f;  // vec4(255,255,255,1) clamped
    for (int i = 0; i < 300; ++i) {
      p = add(p, dirStep);
      if (!inside(p)) break;
      float smp = tex(vol, p);
      float smp2 = tex(vol, add(p, dirStep));
      ++n;
      if ((smp - P.iso) < 0.0f && (smp2 - P.iso) >= 0.0f) {
        V3 left = p, right = add(p, dirStep);
        for (int b = 0; b < 4; ++b) {
          V3 mid = scale(add(right, left), 0.5f);
          if (tex(vol, mid) < P.iso) left = mid; else right = mid;
        }
        V3 tc = scale(add(right, left), 0.5f);
        const float D = 0.01f;
        V3 s1 = {tex(vol, sub(tc, V3{D, 0, 0})), tex(vol, sub(tc, V3{0, D, 0})), tex(vol, sub(tc, V3{0, 0, D}))};
        V3 s2 = {tex(vol, add(tc, V3{D, 0, 0})), tex(vol, add(tc, V3{0, D, 0})), tex(vol, add(tc, V3{0, 0, D}))};
        V3 g = sub(s1, s2);
        V3 N = normalize(V3{g.x / 2.0f, g.y / 2.0f, g.z / 2.0f});
        V3 Vv = {-dir.x, -dir.y, -dir.z};
        V3 L = Vv;
        float diffuse = std::fmax(dot(L, N), 0.0f);
        V3 halfVec = normalize(add(L, Vv));
        float specular = std::pow(std::fmax(0.00001f, dot(halfVec, N)), 250.0f);
        rgba[0] = clamp01(diffuse * 0.39f + specular);
        rgba[1] = clamp01(diffuse * 0.58f + specular);
        rgba[2] = clamp01(diffuse * 0.93f + specular);
        rgba[3] = 1.0f;
        if (hitIndex) *hitIndex = i;
        break;
      }
    }
  }
  if (samples) *samples += n;
}

}  // namespace

extern "C" {
// params: 29 floats then W,H,mode as int32 (see Params).  out: H*W*4 floats, row 0 = bottom row (GL).
// hit (optional): H*W int32 sample index of the iso hit, -1 = none.  Returns samples fetched (S).
int64_t oracle_render_rows(const uint8_t* volume, int64_t X, int64_t Y, int64_t Z, const void* params, float* out,
                           int32_t* hit, int threads, int rowStep);
int64_t oracle_render(const uint8_t* volume, int64_t X, int64_t Y, int64_t Z, const void* params, float* out,
                      int32_t* hit, int threads) {
  return oracle_render_rows(volume, X, Y, Z, params, out, hit, threads, 1);
}
// Bounded sample for the timed CPU baseline: only rows with py % rowStep == 0 are shaded.
int64_t oracle_render_rows(const uint8_t* volume, int64_t X, int64_t Y, int64_t Z, const void* params, float* out,
                           int32_t* hit, int threads, int rowStep) {
  Params P;
  std::memcpy(&P, params, sizeof(P));
  Vol vol = {volume, X, Y, Z};
  if (threads < 1) threads = 1;
  std::vector<int64_t> counts((size_t)threads, 0);
  std::vector<std::thread> pool;
  for (int w = 0; w < threads; ++w)
    pool.emplace_back([&, w]() {
      int64_t n = 0;
      for (int py = w * rowStep; py < P.H; py += threads * rowStep)
        for (int px = 0; px < P.W; ++px)
          shade_pixel(P, vol, px, py, out + 4 * ((int64_t)py * P.W + px), &n, hit ? hit + (int64_t)py * P.W + px : nullptr);
      counts[(size_t)w] = n;
    });
  for (auto& t : pool) t.join();
  int64_t S = 0;
  for (auto c : counts) S += c;
  return S;
}
int oracle_render_params_size(void) { return (int)sizeof(Params); }
}

// ---------------------------------------------------------------------------------------------------
// Sort-last restatement (no reference counterpart: the reference renders one texture on one GPU).
// CPU stand-in for the product's slab kernels, used by the world_size-2 gloo tests of the host-side
// exchange logic; ground truth for the composited frame stays oracle_render on the whole volume.
namespace {
struct Slab { const uint8_t* d; int G[3], o[3], S[3], lo[3], hi[3]; };
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline float slab_texel(const Slab& v, int x, int y, int z) {
  x = clampi(clampi(x, 0, v.G[0] - 1) - v.o[0], 0, v.S[0] - 1);
  y = clampi(clampi(y, 0, v.G[1] - 1) - v.o[1], 0, v.S[1] - 1);
  z = clampi(clampi(z, 0, v.G[2] - 1) - v.o[2], 0, v.S[2] - 1);
  return (float)v.d[x + (int64_t)v.S[0] * (y + (int64_t)v.S[1] * z)];
}
static inline float slab_tex(const Slab& v, V3 p) {
  float cx = p.x * (float)v.G[0] - 0.5f, cy = p.y * (float)v.G[1] - 0.5f, cz = p.z * (float)v.G[2] - 0.5f;
  float fx = std::floor(cx), fy = std::floor(cy), fz = std::floor(cz);
  float wx = cx - fx, wy = cy - fy, wz = cz - fz;
  int x0 = (int)fx, y0 = (int)fy, z0 = (int)fz;
  float c000 = slab_texel(v, x0, y0, z0), c100 = slab_texel(v, x0 + 1, y0, z0);
  float c010 = slab_texel(v, x0, y0 + 1, z0), c110 = slab_texel(v, x0 + 1, y0 + 1, z0);
  float c001 = slab_texel(v, x0, y0, z0 + 1), c101 = slab_texel(v, x0 + 1, y0, z0 + 1);
  float c011 = slab_texel(v, x0, y0 + 1, z0 + 1), c111 = slab_texel(v, x0 + 1, y0 + 1, z0 + 1);
  float c00 = std::fmaf(wx, c100 - c000, c000), c10 = std::fmaf(wx, c110 - c010, c010);
  float c01 = std::fmaf(wx, c101 - c001, c001), c11 = std::fmaf(wx, c111 - c011, c011);
  float c0 = std::fmaf(wy, c10 - c00, c00), c1 = std::fmaf(wy, c11 - c01, c01);
  return std::fmaf(wz, c1 - c0, c0) * (1.0f / 255.0f);
}
static inline bool slab_owned(const Slab& v, V3 p) {
  int b[3] = {clampi((int)std::floor(p.x * (float)v.G[0] - 0.5f), 0, v.G[0] - 1),
              clampi((int)std::floor(p.y * (float)v.G[1] - 0.5f), 0, v.G[1] - 1),
              clampi((int)std::floor(p.z * (float)v.G[2] - 0.5f), 0, v.G[2] - 1)};
  for (int a = 0; a < 3; ++a) if (b[a] < v.lo[a] || b[a] >= v.hi[a]) return false;
  return true;
}
static Slab make_slab(const uint8_t* d, const int32_t* geom) {
  Slab v; v.d = d;
  for (int a = 0; a < 3; ++a) { v.G[a] = geom[a]; v.o[a] = geom[3 + a]; v.S[a] = geom[6 + a]; v.lo[a] = geom[9 + a]; v.hi[a] = geom[12 + a]; }
  return v;
}
static void segment(const Slab& v, V3 p, V3 dirStep, float* C, float* A) {
  bool started = false;
  for (int i = 0; i < 300; ++i) {
    p = add(p, dirStep);
    if (!inside(p)) break;
    if (!slab_owned(v, p)) { if (started) break; continue; }
    started = true;
    float smp = slab_tex(v, p);
    float prev_alpha = smp - (smp * *A);
    *C = prev_alpha * smp + *C;
    *A += prev_alpha * 0.6f;
    if (*A > 0.99f) break;
  }
}
static void ray_dirs(const Params& P, V3 vUV, V3* dir, V3* dirStep) {
  V3 eye = {P.eye[0], P.eye[1], P.eye[2]};
  *dir = normalize(sub(sub(vUV, V3{0.5f, 0.5f, 0.5f}), eye));
  *dirStep = mul(*dir, V3{P.step[0], P.step[1], P.step[2]});
}
template <class F> static void for_pixels(const Params& P, int threads, F f) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int w = 0; w < threads; ++w)
    pool.emplace_back([&, w]() { for (int py = w; py < P.H; py += threads) for (int px = 0; px < P.W; ++px) f(px, py); });
  for (auto& t : pool) t.join();
}
}  // namespace

extern "C" {
// geom: 15 int32 = G[3], o[3], S[3], lo[3], hi[3]
void oracle_partial_composite(const uint8_t* slab, const int32_t* geom, const void* params, float* cseg, float* tseg, int threads) {
  Params P; std::memcpy(&P, params, sizeof(P));
  Slab v = make_slab(slab, geom);
  for_pixels(P, threads, [&](int px, int py) {
    float C = 0.0f, A = 0.0f; V3 vUV;
    if (ray_setup(P, px, py, &vUV)) { V3 dir, ds; ray_dirs(P, vUV, &dir, &ds); segment(v, vUV, ds, &C, &A); }
    cseg[(int64_t)py * P.W + px] = C; tseg[(int64_t)py * P.W + px] = 1.0f - A;
  });
}
void oracle_partial_iso(const uint8_t* slab, const int32_t* geom, const void* params, float* rgba, int32_t* hit, int threads) {
  Params P; std::memcpy(&P, params, sizeof(P));
  Slab v = make_slab(slab, geom);
  for_pixels(P, threads, [&](int px, int py) {
    float* o = rgba + 4 * ((int64_t)py * P.W + px);
    o[0] = o[1] = o[2] = o[3] = 0.0f;
    int32_t h = 0x7FFFFFFF; V3 vUV;
    if (ray_setup(P, px, py, &vUV)) {
      V3 dir, dirStep; ray_dirs(P, vUV, &dir, &dirStep);
      V3 p = vUV; bool started = false;
      for (int i = 0; i < 300; ++i) {
        p = add(p, dirStep);
        if (!inside(p)) break;
        if (!slab_owned(v, p)) { if (started) break; continue; }
        started = true;
        float smp = slab_tex(v, p), smp2 = slab_tex(v, add(p, dirStep));
        if ((smp - P.iso) < 0.0f && (smp2 - P.iso) >= 0.0f) {
          V3 left = p, right = add(p, dirStep);
          for (int b = 0; b < 4; ++b) { V3 mid = scale(add(right, left), 0.5f); if (slab_tex(v, mid) < P.iso) left = mid; else right = mid; }
          V3 tc = scale(add(right, left), 0.5f);
          const float D = 0.01f;
          V3 s1 = {slab_tex(v, sub(tc, V3{D, 0, 0})), slab_tex(v, sub(tc, V3{0, D, 0})), slab_tex(v, sub(tc, V3{0, 0, D}))};
          V3 s2 = {slab_tex(v, add(tc, V3{D, 0, 0})), slab_tex(v, add(tc, V3{0, D, 0})), slab_tex(v, add(tc, V3{0, 0, D}))};
          V3 g = sub(s1, s2);
          V3 N = normalize(V3{g.x / 2.0f, g.y / 2.0f, g.z / 2.0f});
          V3 L = {-dir.x, -dir.y, -dir.z};
          float diffuse = std::fmax(dot(L, N), 0.0f);
          V3 halfVec = normalize(add(L, L));
          float specular = std::pow(std::fmax(0.00001f, dot(halfVec, N)), 250.0f);
          o[0] = clamp01(diffuse * 0.39f + specular); o[1] = clamp01(diffuse * 0.58f + specular);
          o[2] = clamp01(diffuse * 0.93f + specular); o[3] = 1.0f;
          h = i;
          break;
        }
      }
    }
    hit[(int64_t)py * P.W + px] = h;
  });
}
// order: 8 x n visibility orders; contrib: W*H*2 floats
void oracle_resolve_composite(const uint8_t* slab, const int32_t* geom, const void* params, const float* tsegAll, int n, int rank,
                              const int32_t* order, const float* cseg, float* contrib, int threads) {
  Params P; std::memcpy(&P, params, sizeof(P));
  Slab v = make_slab(slab, geom);
  const int64_t npix = (int64_t)P.W * P.H;
  for_pixels(P, threads, [&](int px, int py) {
    const int64_t pix = (int64_t)py * P.W + px;
    float tC = 0.0f, tA = 0.0f; V3 vUV;
    if (ray_setup(P, px, py, &vUV)) {
      V3 dir, ds; ray_dirs(P, vUV, &dir, &ds);
      const int oct = (ds.x < 0.0f ? 1 : 0) | (ds.y < 0.0f ? 2 : 0) | (ds.z < 0.0f ? 4 : 0);
      const int32_t* ord = order + oct * n;
      float T = 1.0f, Tin = 1.0f; int myPos = -1, crossPos = -1;
      for (int k = 0; k < n; ++k) {
        int r = ord[k];
        if (r == rank) { myPos = k; Tin = T; }
        T = T * tsegAll[(int64_t)r * npix + pix];
        if (crossPos < 0 && (1.0f - T) > 0.99f) crossPos = k;
      }
      if (crossPos >= 0 && crossPos < myPos) {
      } else if (crossPos == myPos && Tin == 1.0f) {
        tC = cseg[pix]; tA = 1.0f - tsegAll[(int64_t)rank * npix + pix];
      } else if (crossPos == myPos) {
        float C = 0.0f, A = 1.0f - Tin;
        segment(v, vUV, ds, &C, &A);
        tC = C; tA = A;
      } else {
        tC = Tin * cseg[pix];
        if (crossPos < 0 && myPos == n - 1) tA = 1.0f - T;
      }
    }
    contrib[2 * pix] = tC; contrib[2 * pix + 1] = tA;
  });
}
void oracle_finish_composite(const void* params, const float* sum, float* out, int threads) {
  Params P; std::memcpy(&P, params, sizeof(P));
  for_pixels(P, threads, [&](int px, int py) {
    const int64_t pix = (int64_t)py * P.W + px;
    float* o = out + 4 * pix; V3 vUV;
    o[0] = o[1] = o[2] = o[3] = 1.0f;
    if (ray_setup(P, px, py, &vUV)) { o[0] = o[1] = clamp01(1.0f - sum[2 * pix]); o[2] = 1.0f; o[3] = clamp01(sum[2 * pix + 1]); }
  });
}
}
