// vr_capi_sortlast.cu -- C ABI of the sort-last renderer (include/vr_b200.h, "sort-last rendering"):
// slab geometry and visibility orders on the host, kernels of render_sortlast.cu on the device.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "../../include/vr_b200.h"
#include "occupancy.h"
#include "render_march.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)


namespace vr { int make_render_params(const vr_render_params* in, RenderParams* P); int skip_threshold(const RenderParams& P); }

static int slab_geom(const vr_sortlast* sl, const vr_volume* v, SlabGeom* g) {
  const vr_slab_desc& d = sl->slabs[(size_t)sl->rank];
  const int64_t S[3] = {v->SX, v->SY, v->SZ};
  for (int a = 0; a < 3; ++a) {
    if (d.global[a] > 0x7FFFFFFF) return fail(VR_ERR_UNSUPPORTED, "volume extent too large");
    if (d.origin[a] < 0 || d.origin[a] + S[a] > d.global[a]) return fail(VR_ERR_ARG, "slab does not lie inside the global volume");
    if (d.own_lo[a] < d.origin[a] || d.own_hi[a] > d.origin[a] + S[a]) return fail(VR_ERR_ARG, "owned box is not inside the device slab");
    g->G[a] = (int)d.global[a]; g->o[a] = (int)d.origin[a]; g->S[a] = (int)S[a];
    g->lo[a] = (int)d.own_lo[a]; g->hi[a] = (int)d.own_hi[a];
  }
  return VR_OK;
}

extern "C" {

int vr_sortlast_create(vr_context* ctx, const vr_slab_desc* slabs, int32_t n, int32_t rank, vr_sortlast** out) {
  if (!slabs || !out || n < 1 || n > VR_MAX_RANKS || rank < 0 || rank >= n) return fail(VR_ERR_ARG, "vr_sortlast_create: bad argument");
  for (int r = 0; r < n; ++r)
    for (int a = 0; a < 3; ++a) {
      if (slabs[r].global[a] != slabs[0].global[a] || slabs[r].global[a] <= 0) return fail(VR_ERR_ARG, "slabs disagree on the global extent");
      if (slabs[r].own_lo[a] < 0 || slabs[r].own_hi[a] > slabs[r].global[a] || slabs[r].own_lo[a] >= slabs[r].own_hi[a])
        return fail(VR_ERR_ARG, "empty or out-of-range owned box");
    }
  // the owned boxes must tile the texture: equal total volume and pairwise disjoint
  long double vol = 0, tot = (long double)slabs[0].global[0] * slabs[0].global[1] * slabs[0].global[2];
  for (int r = 0; r < n; ++r) {
    vol += (long double)(slabs[r].own_hi[0] - slabs[r].own_lo[0]) * (slabs[r].own_hi[1] - slabs[r].own_lo[1]) * (slabs[r].own_hi[2] - slabs[r].own_lo[2]);
    for (int q = 0; q < r; ++q) {
      bool overlap = true;
      for (int a = 0; a < 3; ++a)
        overlap = overlap && slabs[r].own_lo[a] < slabs[q].own_hi[a] && slabs[q].own_lo[a] < slabs[r].own_hi[a];
      if (overlap) return fail(VR_ERR_ARG, "owned boxes overlap");
    }
  }
  if (vol != tot) return fail(VR_ERR_ARG, "owned boxes do not cover the texture");
  // ... and they must form a GRID: per axis one set of cut planes shared by all boxes, every box one cell of it, n = product
  // of the cell counts.  Only then is the lexicographic order below a visibility order for every ray (three boxes
  // [0,2]x[1,2], [1,2]x[0,1], [0,1]x[0,1] tile a square, but a +x+y ray can leave the second and enter the first).
  {
    int64_t cells = 1;
    for (int a = 0; a < 3; ++a) {
      std::vector<int64_t> cuts;
      for (int r = 0; r < n; ++r) { cuts.push_back(slabs[r].own_lo[a]); cuts.push_back(slabs[r].own_hi[a]); }
      std::sort(cuts.begin(), cuts.end());
      cuts.erase(std::unique(cuts.begin(), cuts.end()), cuts.end());
      for (int r = 0; r < n; ++r) {
        const auto lo = std::lower_bound(cuts.begin(), cuts.end(), slabs[r].own_lo[a]);
        if (*(lo + 1) != slabs[r].own_hi[a])
          return fail(VR_ERR_ARG, "owned boxes are not a grid partition (a box spans a cut plane of another): no front-to-back order exists for every ray");
      }
      cells *= (int64_t)cuts.size() - 1;
    }
    if (cells != n) return fail(VR_ERR_ARG, "owned boxes are not a grid partition: no front-to-back order exists for every ray");
  }
  vr_sortlast* sl = new vr_sortlast;
  sl->device = ctx ? ctx->device : -1; sl->n = n; sl->rank = rank;
  sl->slabs.assign(slabs, slabs + n);
  // Front-to-back order per direction octant.  For a grid partition any lexicographic order of the box
  // corners, each axis taken in the ray's direction of travel, is a visibility order: two boxes are
  // separated by an axis-aligned plane that spans the volume, and the ray crosses it in that sense.
  for (int oct = 0; oct < 8; ++oct) {
    std::vector<int> idx((size_t)n);
    for (int r = 0; r < n; ++r) idx[(size_t)r] = r;
    std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) {
      for (int ax = 0; ax < 3; ++ax) {
        int64_t ka = (oct >> ax) & 1 ? -slabs[a].own_lo[ax] : slabs[a].own_lo[ax];
        int64_t kb = (oct >> ax) & 1 ? -slabs[b].own_lo[ax] : slabs[b].own_lo[ax];
        if (ka != kb) return ka < kb;
      }
      return false;
    });
    for (int r = 0; r < n; ++r) sl->order[oct][r] = idx[(size_t)r];
  }
  *out = sl;  // the device copy of the orders is made by the first vr_resolve_composite (ctx may be NULL here)
  return VR_OK;
}

static int order_on_device(vr_context* ctx, vr_sortlast* sl) {
  if (sl->orderDev && sl->device == ctx->device) return VR_OK;
  if (sl->orderDev) return fail(VR_ERR_STATE, "vr_sortlast used with two devices");
  std::vector<int> flat((size_t)8 * sl->n);
  for (int oct = 0; oct < 8; ++oct) std::memcpy(&flat[(size_t)oct * sl->n], sl->order[oct], sizeof(int) * (size_t)sl->n);
  CU(cudaSetDevice(ctx->device));
  CU(cudaMalloc((void**)&sl->orderDev, flat.size() * sizeof(int)));
  CU(cudaMemcpy(sl->orderDev, flat.data(), flat.size() * sizeof(int), cudaMemcpyHostToDevice));
  sl->device = ctx->device;
  return VR_OK;
}

int vr_sortlast_destroy(vr_sortlast* sl) {
  if (!sl) return VR_OK;
  if (sl->orderDev) { cudaSetDevice(sl->device); cudaFree(sl->orderDev); }
  delete sl;
  return VR_OK;
}

int vr_sortlast_order(const vr_sortlast* sl, int32_t octant, int32_t* out) {
  if (!sl || !out || octant < 0 || octant > 7) return fail(VR_ERR_ARG, "vr_sortlast_order: bad argument");
  for (int r = 0; r < sl->n; ++r) out[r] = sl->order[octant][r];
  return VR_OK;
}

int vr_sortlast_halo(const vr_render_params* p, const int64_t global[3], int64_t halo[3]) {
  if (!p || !global || !halo) return fail(VR_ERR_ARG, "null");
  for (int a = 0; a < 3; ++a) {
    if (p->mode == VR_MODE_COMPOSITE) { halo[a] = 1; continue; }
    // base texel -> taps of tex(p + dirStep) and of the gradient probes tc +- 0.01 (tc within the step)
    double reach = (std::fabs((double)p->step_size[a]) + 0.01) * (double)global[a];
    halo[a] = (int64_t)std::ceil(reach) + 2;
  }
  return VR_OK;
}

#define SL_PROLOGUE(need_slab)                                                                      \
  if (!ctx || !sl || !params) return fail(VR_ERR_ARG, "null argument");                             \
  RenderParams P;                                                                                   \
  if (make_render_params(params, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size"); \
  SlabGeom g;                                                                                       \
  std::memset(&g, 0, sizeof(g));                                                                    \
  if (need_slab) { if (!slab) return fail(VR_ERR_ARG, "null slab"); int rc_ = slab_geom(sl, slab, &g); if (rc_ != VR_OK) return rc_; } \
  if (need_slab && slab->is_partial()) return fail(VR_ERR_STATE, "the slab was decoded lazily for one frame; run its plan in full first"); \
  CU(cudaSetDevice(ctx->device));                                                                   \
  OccView occ;                                                                                      \
  std::memset(&occ, 0, sizeof(occ));                                                                \
  if (need_slab) { int rc_ = volume_occ_view(ctx, slab, skip_threshold(P), &occ); if (rc_ != VR_OK) return rc_; }

int vr_partial_composite(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                         float* cseg, float* tseg, uint64_t* samples) {
  if (!cseg || !tseg) return fail(VR_ERR_ARG, "null output");
  SL_PROLOGUE(true)
  CU(partial_composite_launch(P, g, occ, slab->dev, cseg, tseg, reinterpret_cast<unsigned long long*>(samples), ctx->stream));
  return VR_OK;
}

int vr_partial_iso(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                   float* rgba, int32_t* hit, uint64_t* samples) {
  if (!rgba || !hit) return fail(VR_ERR_ARG, "null output");
  SL_PROLOGUE(true)
  CU(partial_iso_launch(P, g, occ, slab->dev, reinterpret_cast<float4*>(rgba), hit, reinterpret_cast<unsigned long long*>(samples), ctx->stream));
  return VR_OK;
}

int vr_resolve_composite(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                         const float* tseg_all, const float* cseg, float* contrib, uint64_t* samples) {
  if (!tseg_all || !cseg || !contrib) return fail(VR_ERR_ARG, "null buffer");
  SL_PROLOGUE(true)
  { int rc_ = order_on_device(ctx, const_cast<vr_sortlast*>(sl)); if (rc_ != VR_OK) return rc_; }
  CU(resolve_composite_launch(P, g, occ, slab->dev, tseg_all, sl->n, sl->rank, sl->orderDev, cseg, reinterpret_cast<float2*>(contrib),
                              reinterpret_cast<unsigned long long*>(samples), ctx->stream));
  return VR_OK;
}

int vr_finish_composite(vr_context* ctx, const vr_render_params* params, const float* sum, float* out) {
  if (!ctx || !params || !sum || !out) return fail(VR_ERR_ARG, "null argument");
  RenderParams P;
  if (make_render_params(params, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  CU(cudaSetDevice(ctx->device));
  CU(finish_composite_launch(P, reinterpret_cast<const float2*>(sum), reinterpret_cast<float4*>(out), ctx->stream));
  return VR_OK;
}

int vr_resolve_iso(vr_context* ctx, const vr_sortlast* sl, const vr_render_params* params, const int32_t* hit_mine,
                   const int32_t* hit_min, float* rgba) {
  if (!ctx || !sl || !params || !hit_mine || !hit_min || !rgba) return fail(VR_ERR_ARG, "null argument");
  if (params->width <= 0 || params->height <= 0) return fail(VR_ERR_ARG, "bad image size");
  CU(cudaSetDevice(ctx->device));
  CU(resolve_iso_launch(params->width, params->height, sl->rank, hit_mine, hit_min, reinterpret_cast<float4*>(rgba), ctx->stream));
  return VR_OK;
}

static int box_copy(vr_context* ctx, const vr_volume* v, const int64_t o[3], const int64_t e[3], uint8_t* dense, bool pack) {
  if (!ctx || !v || !o || !e || !dense) return fail(VR_ERR_ARG, "null argument");
  const int64_t S[3] = {v->SX, v->SY, v->SZ};
  for (int a = 0; a < 3; ++a)
    if (o[a] < 0 || e[a] <= 0 || o[a] + e[a] > S[a]) return fail(VR_ERR_ARG, "box outside the slab");
  if (pack && v->is_partial()) return fail(VR_ERR_STATE, "vr_volume_pack_box: the volume was decoded lazily for one frame; run its plan in full first");
  CU(cudaSetDevice(ctx->device));
  cudaMemcpy3DParms p;
  std::memset(&p, 0, sizeof(p));
  cudaPitchedPtr slab = make_cudaPitchedPtr(v->dev, (size_t)v->SX, (size_t)v->SX, (size_t)v->SY);
  cudaPitchedPtr buf = make_cudaPitchedPtr(dense, (size_t)e[0], (size_t)e[0], (size_t)e[1]);
  cudaPos at = make_cudaPos((size_t)o[0], (size_t)o[1], (size_t)o[2]), zero = make_cudaPos(0, 0, 0);
  p.srcPtr = pack ? slab : buf; p.srcPos = pack ? at : zero;
  p.dstPtr = pack ? buf : slab; p.dstPos = pack ? zero : at;
  p.extent = make_cudaExtent((size_t)e[0], (size_t)e[1], (size_t)e[2]);
  p.kind = cudaMemcpyDeviceToDevice;
  CU(cudaMemcpy3DAsync(&p, ctx->stream));
  if (!pack) volume_occ_written(ctx, const_cast<vr_volume*>(v), o, e);
  return VR_OK;
}
int vr_volume_pack_box(vr_context* ctx, const vr_volume* v, const int64_t origin[3], const int64_t extent[3], uint8_t* dst) {
  return box_copy(ctx, v, origin, extent, dst, true);
}
int vr_volume_unpack_box(vr_context* ctx, vr_volume* v, const int64_t origin[3], const int64_t extent[3], const uint8_t* src) {
  return box_copy(ctx, v, origin, extent, const_cast<uint8_t*>(src), false);
}

// ---- composite passes fused with their exchange step over peer windows (render_march.h, PeerTargets) --------------
int vr_peer_partial_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_sortlast* sl, const vr_volume* slab,
                              const vr_render_params* params, float* cseg, int64_t gather_off, uint64_t* samples) {
  if (!cseg || !peer) return fail(VR_ERR_ARG, "null argument");
  SL_PROLOGUE(true)
  PeerTargets t;
  int64_t wb = 0;
  { int rc_ = peer_targets(peer, slot, &t, &wb); if (rc_ != VR_OK) return rc_; }
  const int64_t npix = (int64_t)P.W * P.H;
  if (t.n != sl->n || t.rank != sl->rank || gather_off < 0 || (gather_off & 15) || gather_off + 4 * npix * t.n > wb)
    return fail(VR_ERR_ARG, "vr_peer_partial_composite: window too small or ranks differ");
  CU(partial_composite_peer_launch(P, g, occ, slab->dev, cseg, t, gather_off, reinterpret_cast<unsigned long long*>(samples), ctx->stream));
  return VR_OK;
}

int vr_peer_resolve_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_sortlast* sl, const vr_volume* slab,
                              const vr_render_params* params, int64_t gather_off, const float* cseg, int64_t red_off, uint64_t* samples) {
  if (!cseg || !peer) return fail(VR_ERR_ARG, "null argument");
  SL_PROLOGUE(true)
  { int rc_ = order_on_device(ctx, const_cast<vr_sortlast*>(sl)); if (rc_ != VR_OK) return rc_; }
  PeerTargets t;
  int64_t wb = 0;
  { int rc_ = peer_targets(peer, slot, &t, &wb); if (rc_ != VR_OK) return rc_; }
  const int64_t npix = (int64_t)P.W * P.H;
  if (t.n != sl->n || t.rank != sl->rank || npix % t.n || ((8 * npix / t.n) & 15) || gather_off < 0 || red_off < 0 || (red_off & 15) ||
      gather_off + 4 * npix * t.n > wb || red_off + 8 * npix > wb)
    return fail(VR_ERR_ARG, "vr_peer_resolve_composite: window too small, ranks differ or image not divisible into stripes");
  const float* tsegAll = reinterpret_cast<const float*>(t.base[t.rank] + gather_off);
  CU(resolve_composite_peer_launch(P, g, occ, slab->dev, tsegAll, sl->orderDev, cseg, t, red_off,
                                   reinterpret_cast<unsigned long long*>(samples), ctx->stream));
  return VR_OK;
}

int vr_peer_tin_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_sortlast* sl, const vr_render_params* params,
                          int64_t gather_off, int64_t tin_off) {
  if (!peer) return fail(VR_ERR_ARG, "null argument");
  const vr_volume* slab = nullptr;
  SL_PROLOGUE(false)
  { int rc_ = order_on_device(ctx, const_cast<vr_sortlast*>(sl)); if (rc_ != VR_OK) return rc_; }
  PeerTargets t;
  int64_t wb = 0;
  { int rc_ = peer_targets(peer, slot, &t, &wb); if (rc_ != VR_OK) return rc_; }
  const int64_t npix = (int64_t)P.W * P.H;
  if (t.n != sl->n || t.rank != sl->rank || npix % t.n || ((4 * npix / t.n) & 15) || gather_off < 0 || tin_off < 0 || (gather_off & 15) ||
      (tin_off & 15) || gather_off + 4 * npix > wb || tin_off + 4 * npix > wb)
    return fail(VR_ERR_ARG, "vr_peer_tin_composite: window too small, ranks differ or image not divisible into stripes");
  CU(tin_stripe_peer_launch(P, sl->orderDev, t, gather_off, tin_off, ctx->stream));
  return VR_OK;
}

int vr_resolve_composite_tin(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                             const float* tin, const float* tseg, const float* cseg, float* contrib, uint64_t* samples) {
  if (!tin || !tseg || !cseg || !contrib) return fail(VR_ERR_ARG, "null buffer");
  SL_PROLOGUE(true)
  { int rc_ = order_on_device(ctx, const_cast<vr_sortlast*>(sl)); if (rc_ != VR_OK) return rc_; }
  CU(resolve_composite_tin_launch(P, g, occ, slab->dev, tin, tseg, sl->n, sl->rank, sl->orderDev, cseg, reinterpret_cast<float2*>(contrib),
                                  reinterpret_cast<unsigned long long*>(samples), ctx->stream));
  return VR_OK;
}

int vr_peer_finish_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_render_params* params, int64_t red_off, int64_t frame_off) {
  return vr_peer_finish_composite_to(ctx, peer, slot, params, red_off, frame_off, -1);
}
int vr_peer_finish_composite_to(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_render_params* params, int64_t red_off, int64_t frame_off,
                                int32_t dst_rank) {
  if (!ctx || !peer || !params) return fail(VR_ERR_ARG, "null argument");
  RenderParams P;
  if (make_render_params(params, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  CU(cudaSetDevice(ctx->device));
  PeerTargets t;
  int64_t wb = 0;
  { int rc_ = peer_targets(peer, slot, &t, &wb); if (rc_ != VR_OK) return rc_; }
  const int64_t npix = (int64_t)P.W * P.H;
  if (npix % t.n || red_off < 0 || frame_off < 0 || (frame_off & 15) || red_off + 8 * npix > wb || frame_off + 16 * npix > wb)
    return fail(VR_ERR_ARG, "vr_peer_finish_composite: window too small or image not divisible into stripes");
  if (dst_rank >= t.n) return fail(VR_ERR_ARG, "vr_peer_finish_composite_to: no such rank");
  CU(finish_composite_peer_launch(P, t, red_off, frame_off, dst_rank < 0 ? -1 : dst_rank, ctx->stream));
  return VR_OK;
}

}  // extern "C"
