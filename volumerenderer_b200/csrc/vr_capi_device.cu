// vr_capi_device.cu -- device-side half of the C ABI: stream upload, volumes, decode plans, levelCut.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "../../include/vr_b200.h"
#include "kd_build.h"
#include "kd_decode.h"
#include "kd_cull.h"
#include "kd_worklist.h"
#include "occupancy.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

static inline size_t align16(size_t v) { return (v + 15) & ~size_t(15); }

void vr_stream::release() {
  if (devImage) { if (device >= 0) cudaSetDevice(device); cudaFree(devImage); devImage = nullptr; }
  if (hostImage) { if (pinned) cudaFreeHost(hostImage); else delete[] hostImage; hostImage = nullptr; }
  pinned = false;
}

namespace vr {
int stream_build_image(vr_stream* s) {
  if (s->hostImage) return VR_OK;
  if (!s->kd.index.built) {
    int rc = kd_prepare(&s->kd);
    if (rc != VR_OK) return fail(rc, "vr_stream_prepare: stream does not parse");
  }
  const KdIndex& ix = s->kd.index;
  s->wordsBytes = align16((size_t)s->kd.streamBytes() + 16);
  s->offWords = 0;
  s->offTile = s->wordsBytes;
  s->offLeafOff = s->offTile + align16(ix.tileLeafBlock.size() * 4);
  s->offTileVal = s->offLeafOff + align16(ix.leafOff.size() * 4);
  s->offLeafVal = s->offTileVal + align16(ix.tileVal.size());
  s->imageBytes = s->offLeafVal + align16(ix.leafVal.size());
  s->hostImage = new (std::nothrow) uint8_t[s->imageBytes];
  if (!s->hostImage) return fail(VR_ERR_ARG, "out of host memory");
  std::memset(s->hostImage, 0, s->imageBytes);
  std::memcpy(s->hostImage + s->offWords, s->kd.bits.data(), (size_t)s->kd.streamBytes());
  std::memcpy(s->hostImage + s->offTile, ix.tileLeafBlock.data(), ix.tileLeafBlock.size() * 4);
  if (!ix.leafOff.empty()) std::memcpy(s->hostImage + s->offLeafOff, ix.leafOff.data(), ix.leafOff.size() * 4);
  std::memcpy(s->hostImage + s->offTileVal, ix.tileVal.data(), ix.tileVal.size());
  if (!ix.leafVal.empty()) std::memcpy(s->hostImage + s->offLeafVal, ix.leafVal.data(), ix.leafVal.size());
  return VR_OK;
}
}  // namespace vr

namespace vr {
int context_chain_lut(vr_context* ctx, uint64_t chainDist, const uint32_t** out) {
  auto it = ctx->chainLuts.find(chainDist);
  if (it != ctx->chainLuts.end()) { *out = it->second; return VR_OK; }
  std::vector<uint32_t> host((1u << 14) + VR_LUT2_ENTRIES);
  for (uint32_t i = 0; i < (1u << 14); ++i) host[i] = vr_chain_lut_entry(i, chainDist);
  for (uint32_t i = 0; i < 1024u; ++i) host[VR_LUT2_OFF + i] = vr_chain_lut_entry(i | 0x3C00u, chainDist);         // symbols 0..4
  for (uint32_t i = 0; i < 16u; ++i) host[VR_LUT2_OFF + 1024u + i] = vr_chain_lut_entry(i | 0x30u, chainDist >> 40);  // symbols 5..6
  uint32_t* dev = nullptr;
  CU(cudaMalloc((void**)&dev, host.size() * 4));
  cudaError_t e = cudaMemcpyAsync(dev, host.data(), host.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { cudaFree(dev); return fail_cuda(e, "upload chain table"); }
  ctx->chainLuts[chainDist] = dev;
  *out = dev;
  return VR_OK;
}
}  // namespace vr

extern "C" {

int vr_stream_prepare(vr_stream* s) {
  if (!s) return fail(VR_ERR_ARG, "null stream");
  return stream_build_image(s);
}
int vr_stream_device_bytes(const vr_stream* s, int64_t out[2]) {
  if (!s || !out) return fail(VR_ERR_ARG, "null");
  if (!s->hostImage) return fail(VR_ERR_STATE, "stream not prepared");
  out[0] = (int64_t)s->wordsBytes;
  out[1] = (int64_t)(s->imageBytes - s->wordsBytes);
  return VR_OK;
}
int vr_stream_pin(vr_stream* s) {
  if (!s) return fail(VR_ERR_ARG, "null stream");
  int rc = stream_build_image(s);
  if (rc != VR_OK) return rc;
  if (s->pinned) return VR_OK;
  uint8_t* p = nullptr;
  CU(cudaHostAlloc((void**)&p, s->imageBytes, cudaHostAllocDefault));
  std::memcpy(p, s->hostImage, s->imageBytes);
  delete[] s->hostImage;
  s->hostImage = p;
  s->pinned = true;
  return VR_OK;
}
int vr_stream_upload(vr_context* ctx, vr_stream* s) {
  if (!ctx || !s) return fail(VR_ERR_ARG, "null");
  int rc = stream_build_image(s);
  if (rc != VR_OK) return rc;
  CU(cudaSetDevice(ctx->device));
  if (s->devImage && s->device != ctx->device) { cudaSetDevice(s->device); cudaFree(s->devImage); s->devImage = nullptr; cudaSetDevice(ctx->device); }
  if (!s->devImage) { CU(cudaMalloc((void**)&s->devImage, s->imageBytes)); s->device = ctx->device; }
  CU(cudaMemcpyAsync(s->devImage, s->hostImage, s->imageBytes, cudaMemcpyHostToDevice, ctx->stream));
  return VR_OK;
}

// ---- volume ---------------------------------------------------------------------------------------
int vr_volume_create(vr_context* ctx, int64_t SX, int64_t SY, int64_t SZ, vr_volume** out) {
  if (!ctx || !out || SX <= 0 || SY <= 0 || SZ <= 0) return fail(VR_ERR_ARG, "vr_volume_create: bad argument");
  CU(cudaSetDevice(ctx->device));
  vr_volume* v = new vr_volume;
  v->SX = SX; v->SY = SY; v->SZ = SZ; v->owned = true; v->device = ctx->device;
  cudaError_t e = cudaMalloc((void**)&v->dev, (size_t)(SX * SY * SZ));
  if (e != cudaSuccess) { delete v; return fail_cuda(e, "cudaMalloc(volume)"); }
  *out = v;
  return VR_OK;
}
int vr_volume_wrap(vr_context* ctx, void* p, int64_t SX, int64_t SY, int64_t SZ, vr_volume** out) {
  if (!ctx || !out || !p || SX <= 0 || SY <= 0 || SZ <= 0) return fail(VR_ERR_ARG, "vr_volume_wrap: bad argument");
  vr_volume* v = new vr_volume;
  v->dev = (uint8_t*)p; v->SX = SX; v->SY = SY; v->SZ = SZ; v->owned = false; v->device = ctx->device;
  *out = v;
  return VR_OK;
}
int vr_volume_destroy(vr_volume* v) {
  if (!v) return VR_OK;
  v->occ_free();
  if (v->owned && v->dev) { cudaSetDevice(v->device); cudaFree(v->dev); }
  delete v;
  return VR_OK;
}
int vr_volume_device_ptr(const vr_volume* v, void** p) {
  if (!v || !p) return fail(VR_ERR_ARG, "null");
  *p = v->dev;
  return VR_OK;
}
int vr_volume_dims(const vr_volume* v, int64_t d[3]) {
  if (!v || !d) return fail(VR_ERR_ARG, "null");
  d[0] = v->SX; d[1] = v->SY; d[2] = v->SZ;
  return VR_OK;
}
int vr_volume_download(vr_context* ctx, const vr_volume* v, uint8_t* host) {
  if (!ctx || !v || !host) return fail(VR_ERR_ARG, "null");
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemcpyAsync(host, v->dev, (size_t)(v->SX * v->SY * v->SZ), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return VR_OK;
}
int vr_volume_upload_brick(vr_context* ctx, vr_volume* v, const uint8_t* brick, int64_t BX, int64_t BY, int64_t BZ,
                           int64_t ox, int64_t oy, int64_t oz) {
  if (!ctx || !v || !brick) return fail(VR_ERR_ARG, "null");
  if (ox < 0 || oy < 0 || oz < 0 || ox + BX > v->SX || oy + BY > v->SY || oz + BZ > v->SZ)
    return fail(VR_ERR_ARG, "brick does not fit inside the volume");
  CU(cudaSetDevice(ctx->device));
  cudaMemcpy3DParms p;
  std::memset(&p, 0, sizeof(p));
  p.srcPtr = make_cudaPitchedPtr((void*)brick, (size_t)BX, (size_t)BX, (size_t)BY);
  p.dstPtr = make_cudaPitchedPtr((void*)v->dev, (size_t)v->SX, (size_t)v->SX, (size_t)v->SY);
  p.dstPos = make_cudaPos((size_t)ox, (size_t)oy, (size_t)oz);
  p.extent = make_cudaExtent((size_t)BX, (size_t)BY, (size_t)BZ);
  p.kind = cudaMemcpyHostToDevice;
  CU(cudaMemcpy3DAsync(&p, ctx->stream));
  const int64_t lo[3] = {ox, oy, oz}, ext[3] = {BX, BY, BZ};
  volume_occ_written(ctx, v, lo, ext);
  return VR_OK;
}

}  // extern "C"

// ---- stage 1: decode plans ----------------------------------------------------------------------------
void vr_decode_plan::lazy_free() {
  if (lazy.workDev || lazy.fillJobsDev || lazy.occSetDev || lazy.solidDev || lazy.rangesDev) cudaSetDevice(ctx->device);
  cudaFree(lazy.workDev); cudaFree(lazy.fillJobsDev); cudaFree(lazy.occSetDev); cudaFree(lazy.solidDev); cudaFree(lazy.rangesDev);
  lazy = Lazy();
}

extern "C" { static int env_int(const char* name, int dflt); }

namespace vr {
int plan_lazy_prepare(vr_decode_plan* p, int mode, int thr, const int reach[3], const int stepReach[3]) {
  vr_decode_plan::Lazy& lz = p->lazy;
  if (lz.built && lz.mode == mode && lz.thr == thr && lz.reach[0] == reach[0] && lz.reach[1] == reach[1] && lz.reach[2] == reach[2] &&
      lz.stepReach[0] == stepReach[0] && lz.stepReach[1] == stepReach[1] && lz.stepReach[2] == stepReach[2])
    return VR_OK;
  if (!p->occFused) return fail(VR_ERR_STATE, "lazy decode needs tiles that coincide with occupancy cells");
  p->lazy_free();
  const vr_kd_geom& g = p->geom;
  const vr_volume* d = p->dst;
  const uint32_t nt = 1u << g.Dt;
  const int n = p->n;
  std::vector<const KdIndex*> idx((size_t)n);
  for (int i = 0; i < n; ++i) idx[(size_t)i] = &p->streams[(size_t)i]->kd.index;
  CullGrid cg;
  for (int a = 0; a < 3; ++a) cg.nc[a] = d->nc[a];
  cg.lg[0] = VR_OCC_LX; cg.lg[1] = VR_OCC_LY; cg.lg[2] = VR_OCC_LZ;
  std::vector<uint8_t> needed((size_t)n * nt);
  const std::vector<int64_t>& ob = p->occBase;
  const std::vector<int32_t>& tc = p->tileCellHost;
  std::vector<uint8_t> solid;
  cull_tiles(cg, n, nt, idx.data(), [&](int b, uint32_t t) -> int64_t { return ob[(size_t)b] < 0 ? -1 : ob[(size_t)b] + tc[t]; }, mode, thr,
             reach, stepReach, needed.data(), &solid);
  std::vector<vr_kd_work> work;
  std::vector<vr_fill_job> fills;
  std::vector<vr_occ_set> culled;
  std::vector<vr_stage_range> ranges;
  kd_build_worklist(g, n, idx.data(), p->dstOffset.data(), d->SX, d->SX * d->SY, p->vec16, env_int("VR_DECODE_ORDER", 0), &work, &fills,
                    std::min(env_int("VR_FILL_ATTACH", 2), p->tma ? VR_TMA_MAX_ATTACH : 64), p->origins.data(), env_int("VR_FILL_LATE", 0), ob.data(), tc.data(), needed.data(), &culled,
                    &ranges, p->stageBytes);
  lz.tilesTotal = (int64_t)n * nt;
  lz.tilesWritten = lz.tilesTotal - (int64_t)culled.size();
  lz.bytesVoxels = lz.tilesWritten * ((int64_t)g.tx * g.ty * g.tz);
  lz.bytesStream = 0;
  for (const vr_kd_work& w : work)
    if (w.L == VR_WORK_PARSE) {
      const KdIndex& ix = *idx[(size_t)(w.id >> g.Dt)];
      const uint32_t blk = ix.tileLeafBlock[w.id & (nt - 1u)];
      lz.bytesStream += ((int64_t)ix.blockSymEnd[blk] - (int64_t)ix.blockSymBegin[blk] + 3) / 4;
    }
  if (cudaSetDevice(p->ctx->device) != cudaSuccess) return fail(VR_ERR_CUDA, "cudaSetDevice");
  cudaError_t e = cudaSuccess;
  lz.nWork = (int)work.size(); lz.nFill = (int)fills.size(); lz.nOccSet = (int)culled.size();
  if (lz.nWork) {
    e = cudaMalloc((void**)&lz.workDev, sizeof(vr_kd_work) * work.size());
    if (e == cudaSuccess) e = cudaMemcpy(lz.workDev, work.data(), sizeof(vr_kd_work) * work.size(), cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess && lz.nFill) {
    e = cudaMalloc((void**)&lz.fillJobsDev, sizeof(vr_fill_job) * fills.size());
    if (e == cudaSuccess) e = cudaMemcpy(lz.fillJobsDev, fills.data(), sizeof(vr_fill_job) * fills.size(), cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess && p->stageBytes > 0 && !ranges.empty()) {
    e = cudaMalloc((void**)&lz.rangesDev, sizeof(vr_stage_range) * ranges.size());
    if (e == cudaSuccess) e = cudaMemcpy(lz.rangesDev, ranges.data(), sizeof(vr_stage_range) * ranges.size(), cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess && lz.nOccSet) {
    static_assert(sizeof(vr_occ_set) == 16, "pairs of int64");
    e = cudaMalloc((void**)&lz.occSetDev, sizeof(vr_occ_set) * culled.size());
    if (e == cudaSuccess) e = cudaMemcpy(lz.occSetDev, culled.data(), sizeof(vr_occ_set) * culled.size(), cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess && !solid.empty()) {
    e = cudaMalloc((void**)&lz.solidDev, solid.size());
    if (e == cudaSuccess) e = cudaMemcpy(lz.solidDev, solid.data(), solid.size(), cudaMemcpyHostToDevice);
  }
  if (e != cudaSuccess) { p->lazy_free(); return fail_cuda(e, "upload lazy work list"); }
  lz.mode = mode; lz.thr = thr;
  for (int a = 0; a < 3; ++a) { lz.reach[a] = reach[a]; lz.stepReach[a] = stepReach[a]; }
  lz.built = true;
  return VR_OK;
}

// One lazy pass: decode what the frame can read, record the maxima of the rest.
int plan_lazy_run(vr_decode_plan* p) {
  vr_decode_plan::Lazy& lz = p->lazy;
  vr_volume* d = p->dst;
  if (!lz.built) return fail(VR_ERR_STATE, "lazy work list not built");
  if (!(d->occRaw && d->occEpoch == p->occEpoch && d->skipping)) return fail(VR_ERR_STATE, "occupancy grid of the destination changed");
  cudaError_t e = cudaSetDevice(p->ctx->device);
  if (e == cudaSuccess)
    e = kd_decode_launch(p->geom, p->jobsDev, lz.workDev, lz.nWork, d->dev, d->SX, d->SX * d->SY, p->vec16, p->ctx->stream, d->occRaw,
                         p->tileCellDev, lz.fillJobsDev, lz.rangesDev, p->stageBytes, p->tma ? &p->tmaMaps : nullptr, -1, 0, nullptr,
                         p->tmaRows);
  if (e == cudaSuccess) e = occ_set_cells_launch(d->occRaw, lz.occSetDev, lz.nOccSet, p->ctx->stream);
  if (e != cudaSuccess) return fail_cuda(e, "lazy decode launch");
  d->occDilDirty = true;
  if (lz.nOccSet) d->partialOwner = p;
  return VR_OK;
}
}  // namespace vr

extern "C" {

static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

int vr_decode_plan_create(vr_context* ctx, vr_stream* const* streams, int32_t n, const int64_t* origins,
                          vr_volume* dst, vr_decode_plan** out) {
  if (!ctx || !streams || n <= 0 || !origins || !dst || !out) return fail(VR_ERR_ARG, "vr_decode_plan_create: bad argument");
  CU(cudaSetDevice(ctx->device));
  const vr_kd_geom& g = streams[0]->kd.geom;
  std::vector<vr_kd_job> jobs((size_t)n);
  bool vec16 = (g.tx >= 16) && (dst->SX % 16 == 0) && (((uintptr_t)dst->dev) % 16 == 0);
  vr_decode_plan* p = new vr_decode_plan;
  p->ctx = ctx; p->geom = g; p->n = n; p->dst = dst;
  // the kernel keeps the occupancy grid current when every decode tile is exactly one cell (32x32x16)
  bool occFused = dst->skipping && g.fastPath && g.tx == 32 && g.ty == 32 && g.tz == 16 && volume_occ_alloc(dst) == VR_OK;
  for (int i = 0; i < n; ++i) {
    vr_stream* s = streams[i];
    if (!s || !s->devImage || s->device != ctx->device) { delete p; return fail(VR_ERR_STATE, "stream not uploaded to this device"); }
    const vr_kd_header& h = s->kd.hdr;
    if (h.X != g.X || h.Y != g.Y || h.Z != g.Z) { delete p; return fail(VR_ERR_ARG, "all streams of a plan must share one brick shape"); }
    const int64_t ox = origins[3 * i], oy = origins[3 * i + 1], oz = origins[3 * i + 2];
    if (ox < 0 || oy < 0 || oz < 0 || ox + h.X > dst->SX || oy + h.Y > dst->SY || oz + h.Z > dst->SZ) {
      delete p; return fail(VR_ERR_ARG, "brick placement outside the destination volume");
    }
    if (ox % 16) vec16 = false;
    vr_kd_job& j = jobs[(size_t)i];
    j.words = reinterpret_cast<const uint32_t*>(s->devImage + s->offWords);
    j.tileLeafBlock = reinterpret_cast<const uint32_t*>(s->devImage + s->offTile);
    j.leafOff = reinterpret_cast<const uint32_t*>(s->devImage + s->offLeafOff);
    j.tileVal = s->devImage + s->offTileVal;
    j.leafVal = s->devImage + s->offLeafVal;
    j.treeDist = 0; j.chainDist = 0;
    for (int hh = 0; hh <= g.leafBits && hh < 8; ++hh) j.treeDist |= (uint64_t)s->kd.distanceMap[(size_t)(g.D0 - hh)] << (8 * hh);
    for (int k = 1; k <= VR_ADD_LEVELS; ++k) j.chainDist |= (uint64_t)s->kd.distanceMap[(size_t)(g.D0 + k)] << (8 * (k - 1));
    j.dstOffset = (oz * dst->SY + oy) * dst->SX + ox;
    j.occBase = -1;
    if (occFused) {
      const int64_t l[3] = {ox + dst->shift[0], oy + dst->shift[1], oz + dst->shift[2]};
      if ((l[0] & 31) || (l[1] & 31) || (l[2] & 15)) occFused = false;
      else j.occBase = (l[0] >> 5) + (int64_t)dst->nc[0] * ((l[1] >> 5) + (int64_t)dst->nc[1] * (l[2] >> 4));
    }
    { int rc = context_chain_lut(ctx, j.chainDist, &j.chainLut); if (rc != VR_OK) { delete p; return rc; } }
    int chainSum = 0;  // the table packs the net offset in 8 signed bits: 64+32+...+1 = 127 for every reference stream
    for (int k = 1; k <= VR_ADD_LEVELS; ++k) chainSum += s->kd.distanceMap[(size_t)(g.D0 + k)];
    if (chainSum > 127) p->geom.fastPath = 0;
    p->bytesStream += s->kd.streamBytes() + 88 + h.maxTreeDepth + 1;
    p->bytesVoxels += h.X * h.Y * h.Z;
    p->bytesIndex += (int64_t)(s->imageBytes - s->wordsBytes);
  }
  p->vec16 = vec16;
  p->occFused = occFused && vec16 && p->geom.fastPath;
  p->occEpoch = dst->occEpoch;
  if (!p->occFused) for (auto& j : jobs) j.occBase = -1;
  std::vector<vr_stream*> jobStream(streams, streams + n);
  std::vector<int32_t> tileCellHost;
  if (p->occFused) {
    std::vector<int32_t>& tc = tileCellHost;
    tc.resize(size_t(1) << g.Dt);
    for (uint32_t t = 0; t < tc.size(); ++t) {
      int ox, oy, oz;
      vr_tile_origin(g, t, &ox, &oy, &oz);
      tc[t] = (ox >> 5) + dst->nc[0] * ((oy >> 5) + dst->nc[1] * (oz >> 4));
    }
    cudaError_t et = cudaMalloc((void**)&p->tileCellDev, tc.size() * 4);
    if (et == cudaSuccess) et = cudaMemcpy(p->tileCellDev, tc.data(), tc.size() * 4, cudaMemcpyHostToDevice);
    if (et != cudaSuccess) { cudaFree(p->tileCellDev); delete p; return fail_cuda(et, "upload tile cells"); }
  }
  cudaError_t e;
  {
    std::vector<const KdIndex*> idx((size_t)n);
    std::vector<int64_t> offs((size_t)n);
    for (int i = 0; i < n; ++i) { idx[(size_t)i] = &jobStream[(size_t)i]->kd.index; offs[(size_t)i] = jobs[(size_t)i].dstOffset; }
    std::vector<vr_kd_work> work;
    std::vector<vr_fill_job> fills;
    std::vector<vr_stage_range> ranges;
    // experiment, off by default (VR_STAGE_KB = 8 / 16 turns it on): parse CTAs copy their tile's stream segment into shared
    // memory first (VrWinStage).  Measured on the 120-brick slab: 8 KB 0.263 ms, 16 KB 0.275 ms against 0.267 ms without --
    // the parser is bound by instruction issue, not by the latency of its stream loads, and the larger shared-memory
    // footprint costs L1 hits of the chain table (profiles/r1_ab_stage.txt)
    p->stageBytes = (vec16 && p->geom.fastPath && (kd_decode_smem_bytes(g) % 16) == 0) ? 1024 * env_int("VR_STAGE_KB", 0) : 0;
    // experiment VR_FILL_TMA = 1: the constant groups attached to parse CTAs are stored by the TMA unit
    p->tmaRows = env_int("VR_TMA_ROWS", 8);
    p->tma = env_int("VR_FILL_TMA", 0) != 0 && vec16 && p->geom.fastPath &&
             kd_fill_maps_create(dst->dev, dst->SX, dst->SY, dst->SZ, g.tx, g.ty, p->tmaRows, &p->tmaMaps);
    std::vector<int64_t> ob((size_t)n);
    for (int i = 0; i < n; ++i) ob[(size_t)i] = jobs[(size_t)i].occBase;
    p->persistFillEvery = (vec16 && p->geom.fastPath) ? env_int("VR_DECODE_PERSIST", 0) : 0;
    const bool persist = p->persistFillEvery > 0;  // parse entries first, no attached groups: two queues
    kd_build_worklist(g, n, idx.data(), offs.data(), dst->SX, dst->SX * dst->SY, vec16, persist ? 1 : env_int("VR_DECODE_ORDER", 0), &work, &fills,
                      persist ? 0 : std::min(env_int("VR_FILL_ATTACH", 2), p->tma ? VR_TMA_MAX_ATTACH : 64), origins, env_int("VR_FILL_LATE", 0), p->occFused ? ob.data() : nullptr,
                      p->occFused ? tileCellHost.data() : nullptr, nullptr, nullptr, &ranges, p->stageBytes,
                      env_int("VR_DECODE_ONLY", 0));
    p->nWork = (int)work.size();
    p->nFill = (int)fills.size();
    p->nParse = 0;
    for (const vr_kd_work& w : work) p->nParse += w.L == VR_WORK_PARSE ? 1 : 0;
    if (persist && cudaMalloc((void**)&p->queueDev, 2 * sizeof(uint32_t)) != cudaSuccess) { (void)cudaGetLastError(); p->persistFillEvery = 0; }
    p->streams = jobStream; p->origins.assign(origins, origins + 3 * (size_t)n); p->dstOffset = offs; p->occBase = ob;
    p->tileCellHost = tileCellHost;
    e = cudaMalloc((void**)&p->workDev, sizeof(vr_kd_work) * std::max<size_t>(work.size(), 1));
    if (e == cudaSuccess) e = cudaMemcpy(p->workDev, work.data(), sizeof(vr_kd_work) * work.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && p->nFill) {
      e = cudaMalloc((void**)&p->fillJobsDev, sizeof(vr_fill_job) * fills.size());
      if (e == cudaSuccess) e = cudaMemcpy(p->fillJobsDev, fills.data(), sizeof(vr_fill_job) * fills.size(), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess && p->stageBytes > 0 && !ranges.empty()) {
      e = cudaMalloc((void**)&p->rangesDev, sizeof(vr_stage_range) * ranges.size());
      if (e == cudaSuccess) e = cudaMemcpy(p->rangesDev, ranges.data(), sizeof(vr_stage_range) * ranges.size(), cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) { cudaFree(p->workDev); cudaFree(p->fillJobsDev); cudaFree(p->tileCellDev); cudaFree(p->rangesDev); delete p; return fail_cuda(e, "upload work list"); }
  }
  if (env_int("VR_L2_PREFETCH", 0)) {  // experiment: pull the plan's streams into L2 (evict-last) before every decode launch
    std::vector<const uint8_t*> ptrs;
    std::vector<uint32_t> lines;
    for (vr_stream* s : jobStream) {
      bool seen = false;
      for (const uint8_t* q : ptrs) seen = seen || q == s->devImage;
      if (!seen) { ptrs.push_back(s->devImage); lines.push_back((uint32_t)((s->imageBytes + 127) / 128)); }
    }
    p->nPrefetch = (int)ptrs.size();
    e = cudaMalloc((void**)&p->pfPtrsDev, sizeof(void*) * ptrs.size());
    if (e == cudaSuccess) e = cudaMemcpy(p->pfPtrsDev, ptrs.data(), sizeof(void*) * ptrs.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->pfLinesDev, 4 * lines.size());
    if (e == cudaSuccess) e = cudaMemcpy(p->pfLinesDev, lines.data(), 4 * lines.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { p->nPrefetch = 0; (void)cudaGetLastError(); }
  }
  e = cudaMalloc((void**)&p->jobsDev, sizeof(vr_kd_job) * (size_t)n);
  if (e != cudaSuccess) { cudaFree(p->workDev); cudaFree(p->fillJobsDev); cudaFree(p->tileCellDev); delete p; return fail_cuda(e, "cudaMalloc(jobs)"); }
  e = cudaMemcpyAsync(p->jobsDev, jobs.data(), sizeof(vr_kd_job) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // jobs is a stack-lifetime buffer
  if (e != cudaSuccess) { cudaFree(p->jobsDev); cudaFree(p->workDev); cudaFree(p->fillJobsDev); cudaFree(p->tileCellDev); delete p; return fail_cuda(e, "upload jobs"); }
  *out = p;
  return VR_OK;
}
int vr_decode_plan_run(vr_decode_plan* p) {
  if (!p) return fail(VR_ERR_ARG, "null plan");
  CU(cudaSetDevice(p->ctx->device));
  vr_volume* d = p->dst;
  const bool occ = p->occFused && d->occRaw && d->occEpoch == p->occEpoch && d->skipping;
  if (p->nPrefetch) CU(kd_prefetch_l2_launch(p->pfPtrsDev, p->pfLinesDev, p->nPrefetch, p->ctx->stream));
  CU(kd_decode_launch(p->geom, p->jobsDev, p->workDev, p->nWork, d->dev, d->SX, d->SX * d->SY, p->vec16, p->ctx->stream,
                      occ ? d->occRaw : nullptr, occ ? p->tileCellDev : nullptr, p->fillJobsDev, p->rangesDev, p->stageBytes, p->tma ? &p->tmaMaps : nullptr, p->nParse,
                      p->queueDev ? p->persistFillEvery : 0, p->queueDev, p->tmaRows));
  if (occ) d->occDilDirty = true;  // cells of the decoded bricks are current; validity of the rest is unchanged
  else d->occValid = false;
  if (d->partialOwner == p) d->partialOwner = nullptr;  // every tile of the plan holds its voxels again
  return VR_OK;
}
int vr_decode_plan_destroy(vr_decode_plan* p) {
  if (!p) return VR_OK;
  cudaSetDevice(p->ctx->device);
  cudaFree(p->jobsDev);
  cudaFree(p->workDev);
  cudaFree(p->fillJobsDev);
  cudaFree(p->tileCellDev);
  cudaFree(p->rangesDev);
  cudaFree(p->pfPtrsDev); cudaFree(p->pfLinesDev); cudaFree(p->queueDev);
  p->lazy_free();
  delete p;
  return VR_OK;
}
int vr_decode_plan_bytes(const vr_decode_plan* p, int64_t out[3]) {
  if (!p || !out) return fail(VR_ERR_ARG, "null");
  out[0] = p->bytesStream; out[1] = p->bytesVoxels; out[2] = p->bytesIndex;
  return VR_OK;
}

int vr_kd_levelcut(vr_context* ctx, vr_stream* s, int32_t cutDepth, uint8_t* out_host) {
  if (!ctx || !s || !out_host) return fail(VR_ERR_ARG, "null");
  if (cutDepth != s->kd.hdr.maxTreeDepth)
    return fail(VR_ERR_UNSUPPORTED, "levelCut: only cutDepth == maxTreeDepth is decodable (the reference crashes otherwise)");
  int rc = vr_stream_upload(ctx, s);
  if (rc != VR_OK) return rc;
  vr_volume* v = nullptr;
  rc = vr_volume_create(ctx, s->kd.hdr.X, s->kd.hdr.Y, s->kd.hdr.Z, &v);
  if (rc != VR_OK) return rc;
  const int64_t origin[3] = {0, 0, 0};
  vr_decode_plan* p = nullptr;
  vr_stream* arr[1] = {s};
  rc = vr_decode_plan_create(ctx, arr, 1, origin, v, &p);
  if (rc == VR_OK) rc = vr_decode_plan_run(p);
  if (rc == VR_OK) rc = vr_volume_download(ctx, v, out_host);
  vr_decode_plan_destroy(p);
  vr_volume_destroy(v);
  return rc;
}

// ---- GPU stream producer (kd_build.cu) ---------------------------------------------------------------
static int build_common(vr_context* ctx, const uint8_t* volDev, int64_t X, int64_t Y, int64_t Z, int64_t rowPitch, int64_t slicePitch,
                        int tol, int epochs, vr_stream** out, vr_build_stats* stats, int64_t* iters) {
  if (!ctx->builder) ctx->builder = kd_builder_create();
  vr_stream* s = new vr_stream;
  KdBuildStats st;
  cudaError_t ce = cudaSuccess;
  int rc = kd_build_device(ctx->builder, volDev, X, Y, Z, rowPitch, slicePitch, tol, epochs, ctx->stream, &s->kd, &st, &ce, iters);
  if (rc != VR_OK) {
    delete s;
    if (rc == VR_ERR_CUDA) return fail_cuda(ce, "vr_kd_compress_device");
    return fail(rc, "vr_kd_compress_device: extents must be powers of two");
  }
  if (stats) {
    stats->maxErrorBeforeGrowth = st.maxErrorBeforeGrowth;
    stats->maxError = st.maxError;
    stats->meanL1 = st.meanL1;
    stats->meanL2 = st.meanL2;
    stats->zeroRunRewrites = st.zeroRunRewrites;
  }
  *out = s;
  return VR_OK;
}

int vr_kd_compress_device(vr_context* ctx, const uint8_t* brick_host, int64_t X, int64_t Y, int64_t Z, int tol, int epochs,
                          vr_stream** out, vr_build_stats* stats) {
  if (!ctx || !brick_host || !out || X <= 0 || Y <= 0 || Z <= 0) return fail(VR_ERR_ARG, "vr_kd_compress_device: bad argument");
  CU(cudaSetDevice(ctx->device));
  uint8_t* dev = nullptr;
  const size_t bytes = (size_t)X * (size_t)Y * (size_t)Z;
  CU(cudaMalloc((void**)&dev, bytes));
  cudaError_t e = cudaMemcpyAsync(dev, brick_host, bytes, cudaMemcpyHostToDevice, ctx->stream);
  if (e != cudaSuccess) { cudaFree(dev); return fail_cuda(e, "upload brick"); }
  int rc = build_common(ctx, dev, X, Y, Z, X, X * Y, tol, epochs, out, stats, nullptr);
  cudaFree(dev);
  return rc;
}

int vr_kd_compress_volume(vr_context* ctx, const vr_volume* vol, const int64_t origin[3], int64_t X, int64_t Y, int64_t Z, int tol,
                          int epochs, vr_stream** out, vr_build_stats* stats, int64_t* fixed_point_iterations) {
  if (!ctx || !vol || !origin || !out) return fail(VR_ERR_ARG, "vr_kd_compress_volume: bad argument");
  if (origin[0] < 0 || origin[1] < 0 || origin[2] < 0 || X <= 0 || Y <= 0 || Z <= 0 || origin[0] + X > vol->SX || origin[1] + Y > vol->SY ||
      origin[2] + Z > vol->SZ)
    return fail(VR_ERR_ARG, "vr_kd_compress_volume: brick outside the volume");
  CU(cudaSetDevice(ctx->device));
  const uint8_t* p = vol->dev + (origin[2] * vol->SY + origin[1]) * vol->SX + origin[0];
  return build_common(ctx, p, X, Y, Z, vol->SX, vol->SX * vol->SY, tol, epochs, out, stats, fixed_point_iterations);
}

}  // extern "C"
