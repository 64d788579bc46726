// vr_capi_render.cu -- renderer half of the C ABI: camera set-up (GLM lookAt / perspectiveFov as used
// at main.cpp:396-397), shader-selectable mode (main.cpp:71-74), headless RGBA output.
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>

#include "../../include/vr_b200.h"
#include "kd_cull.h"
#include "occupancy.h"
#include "render_march.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

namespace vr {
int g_count_fetched = 0;  // vr_debug_set("count_fetched")
int make_render_params(const vr_render_params* in, RenderParams* P) {
  if (in->width <= 0 || in->height <= 0) return VR_ERR_ARG;
  double f[3] = {in->front[0], in->front[1], in->front[2]}, up[3] = {in->up[0], in->up[1], in->up[2]};
  double fl = std::sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
  if (!(fl > 0)) return VR_ERR_ARG;
  for (double& c : f) c /= fl;
  double s[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
  double sl = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
  if (!(sl > 0)) return VR_ERR_ARG;
  for (double& c : s) c /= sl;
  double u[3] = {s[1] * f[2] - s[2] * f[1], s[2] * f[0] - s[0] * f[2], s[0] * f[1] - s[1] * f[0]};
  const double kPi = 3.14159265358979323846;
  double th = 1.0 / std::tan((double)in->vfov_deg * kPi / 180.0 / 2.0);
  double tw = th * (double)in->height / (double)in->width;
  for (int i = 0; i < 3; ++i) {
    P->eye[i] = in->eye[i]; P->s[i] = (float)s[i]; P->u[i] = (float)u[i]; P->f[i] = (float)f[i];
    P->step[i] = in->step_size[i];
  }
  P->tw = (float)tw; P->th = (float)th; P->znear = in->znear; P->zfar = in->zfar;
  P->iso = in->isoValue;
  P->W = in->width; P->H = in->height;
  P->mode = in->mode == VR_MODE_COMPOSITE ? 0 : 1;
  P->skipCountMask = g_count_fetched ? 0 : -1;
  return VR_OK;
}
// Largest voxel value m whose sample can never reach the mode's threshold (occupancy.h): composite skips
// exactly-zero cells; isosurface skips cells where even the maximum tap stays below isoValue
// (smp2 <= m/255 in the kernel's own arithmetic, so `smp2 - iso >= 0` of isosurface.frag:126 cannot hold).
int skip_threshold(const RenderParams& P) {
  if (P.mode == 0) return 0;
  int thr = -1;
  for (int m = 0; m < 256; ++m) {
    volatile float s = (float)m * (1.0f / 255.0f);
    volatile float d = s - P.iso;
    if (d < 0.0f) thr = m; else break;
  }
  return thr;
}
int skip_threshold_of(const vr_render_params* params, int* thr, int* mode) {
  RenderParams P;
  if (make_render_params(params, &P) != VR_OK) return VR_ERR_ARG;
  *thr = skip_threshold(P);
  *mode = P.mode;
  return VR_OK;
}

int render_volume(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_dev, int32_t* hit_dev,
                  uint64_t* samples_dev, bool allowPartial, const uint8_t* solidMaskDev) {
  if (!ctx || !vol || !params || !out_rgba_dev) return fail(VR_ERR_ARG, "vr_render: null argument");
  if (params->mode != VR_MODE_COMPOSITE && params->mode != VR_MODE_ISOSURFACE && params->mode != VR_MODE_ISOSURFACE_COMPRESSED)
    return fail(VR_ERR_ARG, "vr_render: unknown mode");
  if (vol->is_partial() && !allowPartial)
    return fail(VR_ERR_STATE, "vr_render: the volume was decoded lazily for one frame (vr_render_compressed_lazy); run its plan in full first");
  RenderParams P;
  if (make_render_params(params, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  if (vol->SX > 0x7FFFFFFF || vol->SY > 0x7FFFFFFF || vol->SZ > 0x7FFFFFFF) return fail(VR_ERR_UNSUPPORTED, "volume extent too large");
  CU(cudaSetDevice(ctx->device));
  OccView occ;
  { int rc = volume_occ_view(ctx, vol, skip_threshold(P), &occ); if (rc != VR_OK) return rc; }
  if (solidMaskDev && occ.lvl && P.mode != 0) {  // this frame only: the cached per-cell levels are recomputed for the next one
    CU(occ_merge_solid_launch(const_cast<uint8_t*>(occ.lvl), solidMaskDev, occ.nx * occ.ny * occ.nz, ctx->stream));
    const_cast<vr_volume*>(vol)->occLvlThr = -2;
  }
  CU(render_launch(P, occ, vol->dev, (int)vol->SX, (int)vol->SY, (int)vol->SZ, reinterpret_cast<float4*>(out_rgba_dev), hit_dev,
                   reinterpret_cast<unsigned long long*>(samples_dev), ctx->stream));
  return VR_OK;
}
}  // namespace vr

extern "C" {

int vr_render_defaults(vr_render_params* p) {
  if (!p) return fail(VR_ERR_ARG, "null");
  std::memset(p, 0, sizeof(*p));
  p->eye[2] = -0.75f;                 // camPos   main.cpp:33
  p->front[2] = 1.0f;                 // camFront main.cpp:34
  p->up[1] = 1.0f;                    // camUp    main.cpp:35
  p->vfov_deg = 50.0f;                // fov      main.cpp:40
  p->znear = 0.1f; p->zfar = 100.0f;  // main.cpp:397
  p->step_size[0] = 1.0f / 256.0f; p->step_size[1] = 1.0f / 256.0f; p->step_size[2] = 1.0f / 128.0f;  // main.cpp:330-331
  p->isoValue = 40.0f / 255.0f;       // main.cpp:52,334
  p->width = 1600; p->height = 1200;  // main.cpp:27
  p->mode = VR_MODE_COMPOSITE;
  return VR_OK;
}

int vr_camera_basis(const vr_render_params* in, float out[20]) {
  if (!in || !out) return fail(VR_ERR_ARG, "null");
  RenderParams P;
  if (make_render_params(in, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  std::memcpy(out, &P, 20 * sizeof(float));
  return VR_OK;
}

int vr_render(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_dev,
              int32_t* hit_dev, uint64_t* samples_dev) {
  return render_volume(ctx, vol, params, out_rgba_dev, hit_dev, samples_dev, false);
}

int vr_render_to_host(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_host,
                      int32_t* hit_host, uint64_t* samples_host) {
  if (!ctx || !params || !out_rgba_host) return fail(VR_ERR_ARG, "vr_render_to_host: null argument");
  if (params->width <= 0 || params->height <= 0) return fail(VR_ERR_ARG, "vr_render_to_host: empty frame");
  CU(cudaSetDevice(ctx->device));
  const size_t px = (size_t)params->width * (size_t)params->height;
  // the device-side frame lives in the context and only grows (one Draw() per frame, main.cpp:404, allocates nothing)
  if (px > ctx->framePixels) {
    cudaFree(ctx->frameDev); ctx->frameDev = nullptr; ctx->framePixels = 0;
    CU(cudaMalloc((void**)&ctx->frameDev, px * 16));
    ctx->framePixels = px;
  }
  if (hit_host && px > ctx->frameHitPixels) {
    cudaFree(ctx->frameHitDev); ctx->frameHitDev = nullptr; ctx->frameHitPixels = 0;
    CU(cudaMalloc((void**)&ctx->frameHitDev, px * 4));
    ctx->frameHitPixels = px;
  }
  if (samples_host && !ctx->frameCountDev) CU(cudaMalloc((void**)&ctx->frameCountDev, 8));
  float* d = ctx->frameDev;
  int32_t* dh = hit_host ? ctx->frameHitDev : nullptr;
  uint64_t* ds = samples_host ? ctx->frameCountDev : nullptr;
  if (ds) CU(cudaMemsetAsync(ds, 0, 8, ctx->stream));
  int rc = vr_render(ctx, vol, params, d, dh, ds);
  if (rc != VR_OK) return rc;
  cudaError_t e = cudaMemcpyAsync(out_rgba_host, d, px * 16, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && hit_host) e = cudaMemcpyAsync(hit_host, dh, px * 4, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && samples_host) e = cudaMemcpyAsync(samples_host, ds, 8, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) return fail_cuda(e, "render download");
  return VR_OK;
}

int vr_render_compressed(vr_context* ctx, vr_decode_plan* plan, const vr_render_params* params, float* out_rgba_dev,
                         int32_t* hit_dev, uint64_t* samples_dev) {
  if (!plan || !params) return fail(VR_ERR_ARG, "null plan");
  if (!plan->dst) return fail(VR_ERR_STATE, "vr_render_compressed: the plan's destination volume was destroyed");
  int rc = vr_decode_plan_run(plan);
  if (rc != VR_OK) return rc;
  // the stream index bounds every tile from both sides, so the isosurface march can also skip look-up cells in which
  // every reachable voxel is >= isoValue (kd_cull.h); the mask is cached in the plan per (isoValue, step_size)
  const uint8_t* solid = nullptr;
  vr_volume* d = plan->dst;
  int thr = 0, mode = 0;
  if (ctx == plan->ctx && skip_threshold_of(params, &thr, &mode) == VR_OK && mode != 0 && thr >= 0 && plan->occFused && d->skipping &&
      d->occRaw && d->occEpoch == plan->occEpoch) {
    const int64_t G[3] = {d->SX, d->SY, d->SZ};
    int reach[3], stepReach[3];
    cull_reach(mode, params->step_size, G, reach);
    cull_step_reach(params->step_size, G, stepReach);
    if (plan_lazy_prepare(plan, mode, thr, reach, stepReach) == VR_OK) solid = plan->lazy.solidDev;
  }
  return render_volume(ctx, d, params, out_rgba_dev, hit_dev, samples_dev, false, solid);
}

int vr_render_compressed_lazy(vr_context* ctx, vr_decode_plan* plan, const vr_render_params* params, float* out_rgba_dev,
                              int32_t* hit_dev, uint64_t* samples_dev, int64_t stats[4]) {
  if (!ctx || !plan || !params || !out_rgba_dev) return fail(VR_ERR_ARG, "vr_render_compressed_lazy: null argument");
  if (ctx != plan->ctx) return fail(VR_ERR_ARG, "vr_render_compressed_lazy: the plan belongs to another context");
  if (!plan->dst) return fail(VR_ERR_STATE, "vr_render_compressed_lazy: the plan's destination volume was destroyed");
  vr_volume* d = plan->dst;
  int thr = 0, mode = 0;
  if (skip_threshold_of(params, &thr, &mode) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  const bool lazy = plan->occFused && d->skipping && d->occRaw && d->occEpoch == plan->occEpoch && thr >= 0;
  if (!lazy) {  // tiles do not coincide with occupancy cells (odd placement) or skipping is off: every tile is needed
    int rc = vr_decode_plan_run(plan);
    if (rc != VR_OK) return rc;
    if (stats) { stats[0] = stats[1] = (int64_t)plan->n << plan->geom.Dt; stats[2] = plan->bytesVoxels; stats[3] = plan->bytesStream; }
    return render_volume(ctx, d, params, out_rgba_dev, hit_dev, samples_dev, false);
  }
  const int64_t G[3] = {d->SX, d->SY, d->SZ};
  int reach[3], stepReach[3];
  cull_reach(mode, params->step_size, G, reach);
  cull_step_reach(params->step_size, G, stepReach);
  int rc = plan_lazy_prepare(plan, mode, thr, reach, stepReach);
  if (rc != VR_OK) return rc;
  if (!d->occValid) {
    if (plan->bytesVoxels == d->SX * d->SY * d->SZ) {
      d->occValid = true;  // the plan covers the volume: after this pass every cell carries an exact maximum
    } else {  // cells outside the plan: bound what is stored there now, before this pass leaves stale voxels behind
      OccView tmp;
      rc = volume_occ_view(ctx, d, thr, &tmp);
      if (rc != VR_OK) return rc;
    }
  }
  rc = plan_lazy_run(plan);
  if (rc != VR_OK) return rc;
  if (stats) { stats[0] = plan->lazy.tilesTotal; stats[1] = plan->lazy.tilesWritten; stats[2] = plan->lazy.bytesVoxels; stats[3] = plan->lazy.bytesStream; }
  return render_volume(ctx, d, params, out_rgba_dev, hit_dev, samples_dev, true, plan->lazy.solidDev);
}

/* host only: the tiles a frame of `params` can read (kd_cull.h), for bricks placed on the cell grid of a volume of dims[] */
int vr_lazy_tile_mask(vr_stream* const* streams, int32_t n, const int64_t* origins, const int64_t dims[3], const vr_render_params* params,
                      uint8_t* needed_out, int32_t* tile_boxes_out, int32_t* tiles_per_brick) {
  if (!streams || n <= 0 || !origins || !dims || !params) return fail(VR_ERR_ARG, "vr_lazy_tile_mask: null argument");
  int thr = 0, mode = 0;
  if (skip_threshold_of(params, &thr, &mode) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  for (int i = 0; i < n; ++i) {
    if (!streams[i]) return fail(VR_ERR_ARG, "null stream");
    if (!streams[i]->kd.index.built) { int rc = kd_prepare(&streams[i]->kd); if (rc != VR_OK) return fail(rc, "prepare"); }
  }
  const vr_kd_geom& g = streams[0]->kd.geom;
  const uint32_t nt = 1u << g.Dt;
  if (tiles_per_brick) *tiles_per_brick = (int32_t)nt;
  if (!needed_out) return VR_OK;
  if (g.tx != (1 << VR_OCC_LX) || g.ty != (1 << VR_OCC_LY) || g.tz != (1 << VR_OCC_LZ))
    return fail(VR_ERR_UNSUPPORTED, "tiles do not coincide with occupancy cells for this brick shape");
  CullGrid cg;
  cg.lg[0] = VR_OCC_LX; cg.lg[1] = VR_OCC_LY; cg.lg[2] = VR_OCC_LZ;
  for (int a = 0; a < 3; ++a) cg.nc[a] = (int)((dims[a] - 1) >> cg.lg[a]) + 1;
  std::vector<const KdIndex*> idx((size_t)n);
  std::vector<int64_t> base((size_t)n);
  for (int i = 0; i < n; ++i) {
    const vr_kd_header& h = streams[i]->kd.hdr;
    if (h.X != g.X || h.Y != g.Y || h.Z != g.Z) return fail(VR_ERR_ARG, "all streams must share one brick shape");
    const int64_t* o = origins + 3 * i;
    if ((o[0] & 31) || (o[1] & 31) || (o[2] & 15) || o[0] < 0 || o[1] < 0 || o[2] < 0 || o[0] + h.X > dims[0] || o[1] + h.Y > dims[1] ||
        o[2] + h.Z > dims[2])
      return fail(VR_ERR_ARG, "brick placement off the cell grid or outside the volume");
    idx[(size_t)i] = &streams[i]->kd.index;
    base[(size_t)i] = (o[0] >> 5) + (int64_t)cg.nc[0] * ((o[1] >> 5) + (int64_t)cg.nc[1] * (o[2] >> 4));
  }
  std::vector<int32_t> tc(nt);
  for (uint32_t t = 0; t < nt; ++t) {
    int ox, oy, oz;
    vr_tile_origin(g, t, &ox, &oy, &oz);
    tc[t] = (ox >> 5) + cg.nc[0] * ((oy >> 5) + cg.nc[1] * (oz >> 4));
    if (tile_boxes_out)
      for (int i = 0; i < n; ++i) {
        int32_t* b = tile_boxes_out + ((size_t)i * nt + t) * 3;
        b[0] = (int32_t)origins[3 * i] + ox; b[1] = (int32_t)origins[3 * i + 1] + oy; b[2] = (int32_t)origins[3 * i + 2] + oz;
      }
  }
  int reach[3], stepReach[3];
  cull_reach(mode, params->step_size, dims, reach);
  cull_step_reach(params->step_size, dims, stepReach);
  cull_tiles(cg, n, nt, idx.data(), [&](int b, uint32_t t) -> int64_t { return base[(size_t)b] + tc[t]; }, mode, thr, reach, stepReach,
             needed_out, nullptr);
  return VR_OK;
}

/* exact bounds of the decoded voxels per tile, from the stream index (no decode): 2^Dt entries each */
int vr_stream_tile_bounds(vr_stream* s, uint8_t* mins, uint8_t* maxs, int64_t cap, int32_t tile_dims[3]) {
  if (!s) return fail(VR_ERR_ARG, "null");
  if (!s->kd.index.built) { int rc = kd_prepare(&s->kd); if (rc != VR_OK) return fail(rc, "prepare"); }
  const int64_t nt = int64_t(1) << s->kd.geom.Dt;
  if (tile_dims) { tile_dims[0] = s->kd.geom.tx; tile_dims[1] = s->kd.geom.ty; tile_dims[2] = s->kd.geom.tz; }
  if (!mins && !maxs) return VR_OK;
  if (cap < nt) return fail(VR_ERR_ARG, "buffer too small");
  if (mins) std::memcpy(mins, s->kd.index.tileMin.data(), (size_t)nt);
  if (maxs) std::memcpy(maxs, s->kd.index.tileMax.data(), (size_t)nt);
  return VR_OK;
}

}  // extern "C"
