// vr_capi_device.cu -- device-side half of the C ABI: stream upload, volumes, decode plans, levelCut.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
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

vr::ArenaMem::~ArenaMem() {
  if (host) cudaFreeHost(host);
  if (dev) { if (device >= 0) cudaSetDevice(device); cudaFree(dev); }
}
void vr_stream::drop_host() {
  if (hostImage && hostOwned) { if (pinned) cudaFreeHost(hostImage); else delete[] hostImage; }
  hostImage = nullptr; pinned = false; hostOwned = true;
  if (devOwned || !devImage) arena.reset();
}
void vr_stream::drop_dev() {
  if (devImage && devOwned) { if (device >= 0) cudaSetDevice(device); cudaFree(devImage); }
  devImage = nullptr; devOwned = true;
  if (hostOwned || !hostImage) arena.reset();
}
void vr_stream::release() {
  drop_dev();
  drop_host();
}

namespace vr {
int stream_image_layout(vr_stream* s) {
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
  return VR_OK;
}
void stream_image_write(const vr_stream* s, uint8_t* dst) {
  const KdIndex& ix = s->kd.index;
  // zero only the gaps: the pad behind the stream and the alignment tails of the sections
  auto put = [&](size_t off, const void* src, size_t n, size_t end) {
    if (n) std::memcpy(dst + off, src, n);
    std::memset(dst + off + n, 0, end - (off + n));
  };
  put(s->offWords, s->kd.bits.data(), (size_t)s->kd.streamBytes(), s->offTile);
  put(s->offTile, ix.tileLeafBlock.data(), ix.tileLeafBlock.size() * 4, s->offLeafOff);
  put(s->offLeafOff, ix.leafOff.data(), ix.leafOff.size() * 4, s->offTileVal);
  put(s->offTileVal, ix.tileVal.data(), ix.tileVal.size(), s->offLeafVal);
  put(s->offLeafVal, ix.leafVal.data(), ix.leafVal.size(), s->imageBytes);
}
int stream_build_image(vr_stream* s) {
  if (s->hostImage) return VR_OK;
  int rc = stream_image_layout(s);
  if (rc != VR_OK) return rc;
  s->hostImage = new (std::nothrow) uint8_t[s->imageBytes];
  if (!s->hostImage) return fail(VR_ERR_ARG, "out of host memory");
  s->hostOwned = true;
  stream_image_write(s, s->hostImage);
  return VR_OK;
}
}  // namespace vr

namespace vr {
int context_chain_lut(vr_context* ctx, uint64_t chainDist, const uint32_t** out) {
  auto it = ctx->chainLuts.find(chainDist);
  if (it != ctx->chainLuts.end()) { *out = it->second; return VR_OK; }
  std::vector<uint32_t> host(1u << 14);
  for (uint32_t i = 0; i < (1u << 14); ++i) host[i] = vr_chain_lut_entry(i, chainDist);
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
  s->drop_host();
  s->hostImage = p;
  s->pinned = true;
  s->hostOwned = true;
  return VR_OK;
}
int vr_stream_upload(vr_context* ctx, vr_stream* s) {
  if (!ctx || !s) return fail(VR_ERR_ARG, "null");
  int rc = stream_build_image(s);
  if (rc != VR_OK) return rc;
  CU(cudaSetDevice(ctx->device));
  if (s->devImage && s->device != ctx->device) { s->drop_dev(); cudaSetDevice(ctx->device); }
  if (!s->devImage) { CU(cudaMalloc((void**)&s->devImage, s->imageBytes)); s->device = ctx->device; s->devOwned = true; }
  CU(cudaMemcpyAsync(s->devImage, s->hostImage, s->imageBytes, cudaMemcpyHostToDevice, ctx->stream));
  return VR_OK;
}

// ---- stream arenas: the images of a timestep in one pinned buffer and one device buffer, ONE copy per upload ----------
struct vr_stream_arena {
  std::shared_ptr<vr::ArenaMem> mem;
  size_t used = 0;
  int nStreams = 0;
};
int vr_stream_arena_create(vr_context* ctx, vr_stream* const* streams, int32_t n, vr_stream_arena** out) {
  if (!ctx || !streams || n <= 0 || !out) return fail(VR_ERR_ARG, "vr_stream_arena_create: bad argument");
  std::vector<vr_stream*> uniq;
  for (int i = 0; i < n; ++i) {
    if (!streams[i]) return fail(VR_ERR_ARG, "vr_stream_arena_create: null stream");
    if (std::find(uniq.begin(), uniq.end(), streams[i]) == uniq.end()) uniq.push_back(streams[i]);
  }
  size_t total = 0;
  std::vector<size_t> off(uniq.size());
  for (size_t i = 0; i < uniq.size(); ++i) {
    int rc = stream_image_layout(uniq[i]);
    if (rc != VR_OK) return rc;
    off[i] = total;
    total += (uniq[i]->imageBytes + 255) & ~size_t(255);
  }
  CU(cudaSetDevice(ctx->device));
  auto mem = std::make_shared<vr::ArenaMem>();
  mem->device = ctx->device; mem->bytes = total;
  CU(cudaHostAlloc((void**)&mem->host, total, cudaHostAllocDefault));
  CU(cudaMalloc((void**)&mem->dev, total));
  for (size_t i = 0; i < uniq.size(); ++i) {
    vr_stream* s = uniq[i];
    if (s->hostImage) std::memcpy(mem->host + off[i], s->hostImage, s->imageBytes);
    else stream_image_write(s, mem->host + off[i]);
    s->drop_host();
    s->drop_dev();
    s->hostImage = mem->host + off[i]; s->pinned = true; s->hostOwned = false;
    s->devImage = mem->dev + off[i]; s->device = ctx->device; s->devOwned = false;
    s->arena = mem;
  }
  vr_stream_arena* a = new vr_stream_arena;
  a->mem = mem; a->used = total; a->nStreams = (int)uniq.size();
  *out = a;
  return VR_OK;
}
int vr_stream_arena_upload(vr_context* ctx, vr_stream_arena* a) {
  if (!ctx || !a) return fail(VR_ERR_ARG, "vr_stream_arena_upload: null argument");
  if (ctx->device != a->mem->device) return fail(VR_ERR_ARG, "vr_stream_arena_upload: the arena lives on another device");
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemcpyAsync(a->mem->dev, a->mem->host, a->used, cudaMemcpyHostToDevice, ctx->stream));
  return VR_OK;
}
int vr_stream_arena_bytes(const vr_stream_arena* a, int64_t* bytes) {
  if (!a || !bytes) return fail(VR_ERR_ARG, "null");
  *bytes = (int64_t)a->used;
  return VR_OK;
}
int vr_stream_arena_destroy(vr_stream_arena* a) {
  delete a;  // the buffers live until the last stream that points into them is destroyed
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
  for (vr_decode_plan* p : v->plans) p->dst = nullptr;
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
namespace { int g_read_partial = 0; }  // vr_debug_set("read_partial", 1): tests look at the tiles a lazy frame left alone
int vr_volume_download(vr_context* ctx, const vr_volume* v, uint8_t* host) {
  if (!ctx || !v || !host) return fail(VR_ERR_ARG, "null");
  if (v->is_partial() && !g_read_partial)
    return fail(VR_ERR_STATE, "vr_volume_download: the volume was decoded lazily for one frame; run its plan in full first");
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
void vr_decode_plan::lazy_free(Lazy* which) {
  Lazy& lz = which ? *which : lazy;
  if (lz.workDev || lz.fillJobsDev || lz.occSetDev || lz.solidDev || lz.parseDev) cudaSetDevice(device);
  cudaFree(lz.workDev); cudaFree(lz.fillJobsDev); cudaFree(lz.occSetDev); cudaFree(lz.solidDev); cudaFree(lz.parseDev);
  lz = Lazy();
}

// Debug knobs (vr_debug_set): explicit, process-wide, read when a plan is created / run.  Nothing in the library reads the
// environment.  decode_kernel: 0 = persistent kernel where the brick shape allows it (default), 1 = classic kernel always.
// decode_only: 0 = everything, 1 = parse tiles only, 2 = constant groups only (timing probes: the slab is then incomplete).
namespace {
struct DebugKnobs { int decodeKernel = 0, decodeOnly = 0, fillAttach = 2, fillMode = 2, tmaRows = 8, parseOrder = 0, unitStore = 0, decodeSplit = 0, producerIndex = 1; } g_dbg;
}
extern "C" int vr_debug_set(const char* key, int value) {
  if (!key) return fail(VR_ERR_ARG, "vr_debug_set: null key");
  const std::string k(key);
  if (k == "decode_kernel") g_dbg.decodeKernel = value;
  else if (k == "decode_only") g_dbg.decodeOnly = value;
  else if (k == "fill_attach") g_dbg.fillAttach = value;
  else if (k == "fill_mode") g_dbg.fillMode = value;
  else if (k == "tma_rows") g_dbg.tmaRows = value;
  else if (k == "parse_order") g_dbg.parseOrder = value;
  else if (k == "unit_store") g_dbg.unitStore = value;  // 0 / 1 = one tensor store per unit (default), 2 = 16-byte stores of the lanes
  else if (k == "decode_split") g_dbg.decodeSplit = value;  // 1: constant groups as a kernel of their own on a side stream
  else if (k == "producer_index") g_dbg.producerIndex = value;  // 0: vr_kd_compress_device leaves the index to the host walk
  else if (k == "count_fetched") vr::g_count_fetched = value;  // samples_dev counts fetching march steps only (plain march)
  else if (k == "read_partial") g_read_partial = value;  // vr_volume_download of a partially decoded volume (tests)
  else return fail(VR_ERR_ARG, "vr_debug_set: unknown key");
  return VR_OK;
}
extern "C" int vr_debug_get(const char* key, int* value) {
  if (!key || !value) return fail(VR_ERR_ARG, "vr_debug_get: null");
  const std::string k(key);
  if (k == "decode_kernel") *value = g_dbg.decodeKernel;
  else if (k == "decode_only") *value = g_dbg.decodeOnly;
  else if (k == "fill_attach") *value = g_dbg.fillAttach;
  else if (k == "fill_mode") *value = g_dbg.fillMode;
  else if (k == "tma_rows") *value = g_dbg.tmaRows;
  else if (k == "parse_order") *value = g_dbg.parseOrder;
  else if (k == "unit_store") *value = g_dbg.unitStore;
  else if (k == "decode_split") *value = g_dbg.decodeSplit;
  else if (k == "producer_index") *value = g_dbg.producerIndex;
  else if (k == "count_fetched") *value = vr::g_count_fetched;
  else if (k == "read_partial") *value = g_read_partial;
  else return fail(VR_ERR_ARG, "vr_debug_get: unknown key");
  return VR_OK;
}

namespace vr {
int context_chain_lut6(vr_context* ctx, uint64_t chainDist, const uint32_t** out) {
  auto it = ctx->chainLuts6.find(chainDist);
  if (it != ctx->chainLuts6.end()) { *out = it->second; return VR_OK; }
  std::vector<uint32_t> host(VR6_LUT_ENTRIES);
  for (uint32_t i = 0; i < VR6_LUT_ENTRIES; ++i) host[i] = vr6::chain_entry(i, chainDist);
  uint32_t* dev = nullptr;
  CU(cudaMalloc((void**)&dev, host.size() * 4));
  cudaError_t e = cudaMemcpyAsync(dev, host.data(), host.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { cudaFree(dev); return fail_cuda(e, "upload chain table"); }
  ctx->chainLuts6[chainDist] = dev;
  *out = dev;
  return VR_OK;
}

template <typename T>
static cudaError_t upload_vec(const std::vector<T>& v, T** dev) {
  *dev = nullptr;
  if (v.empty()) return cudaSuccess;
  cudaError_t e = cudaMalloc((void**)dev, sizeof(T) * v.size());
  if (e == cudaSuccess) e = cudaMemcpy(*dev, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { cudaFree(*dev); *dev = nullptr; }
  return e;
}

// The work lists of one launch over the plan's bricks (all of them, or -- needed != null -- only the tiles a frame can read:
// kd_cull.h).  Persistent kernel: parse items + constant groups; classic kernel: CTA list + attached groups.
static int plan_build_lists(vr_decode_plan* p, const uint8_t* needed, std::vector<vr_occ_set>* culled, vr_parse_item** parseDev, int* nParse,
                            vr_kd_work** workDev, int* nWork, vr_fill_job** fillsDev, int* nFill, int* tmaOut) {
  const vr_kd_geom& g = p->geom;
  const vr_volume* d = p->dst;
  const int n = p->n;
  std::vector<const KdIndex*> idx((size_t)n);
  for (int i = 0; i < n; ++i) idx[(size_t)i] = &p->streams[(size_t)i]->kd.index;
  const int64_t* ob = p->occFused ? p->occBase.data() : nullptr;
  const int32_t* tc = p->occFused ? p->tileCellHost.data() : nullptr;
  std::vector<vr_fill_job> fills;
  *parseDev = nullptr; *workDev = nullptr; *fillsDev = nullptr; *nParse = 0; *nWork = 0; *nFill = 0; *tmaOut = 0;
  cudaError_t e = cudaSuccess;
  if (p->persistent6) {
    std::vector<vr_parse_item> parse;
    std::vector<uint32_t> brickOf;
    if (!kd_build_worklist6(g, n, idx.data(), p->dstOffset.data(), d->SX, d->SX * d->SY, p->origins.data(), ob, tc, needed, culled,
                            kd_decode6_stage_chunks(), p->p6.pg, &parse, &brickOf, &fills, g_dbg.parseOrder != 0))
      return fail(VR_ERR_FORMAT, "a unit of the stream is longer than the grammar allows");
    for (size_t i = 0; i < parse.size(); ++i) {  // device addresses of the brick's prepared stream
      vr_parse_item& it = parse[i];
      const vr_stream* s = p->streams[(size_t)brickOf[i]];
      it.arena = s->devImage;
      for (int r = 0; r < 4; ++r) it.run[r].segOff += (uint32_t)s->offWords;
      it.leafOffOff = (uint32_t)(s->offLeafOff + 4ull * it.leaf0);
      it.leafValOff = (uint32_t)(s->offLeafVal + (uint64_t)it.leaf0);
      it.treeDist = p->treeDist[(size_t)brickOf[i]];
    }
    // a unit's voxels leave as one tensor store (kd_decode.cu: parse_warp); measured faster than 16-byte stores of the lanes on
    // all three workloads (slab 0.247 / 0.268 ms, config-3 layer 0.670 / 0.700, noise 0.265 / 0.288: profiles/r2_decode.md)
    *tmaOut = g_dbg.unitStore == 2 ? 0 : 1;
    e = upload_vec(parse, parseDev);
    *nParse = (int)parse.size();
  } else {
    std::vector<vr_kd_work> work;
    kd_build_worklist(g, n, idx.data(), p->dstOffset.data(), d->SX, d->SX * d->SY, p->vec16, 0, &work, &fills, std::min(g_dbg.fillAttach, 64),
                      p->origins.data(), 0, ob, tc, needed, culled, needed ? 0 : g_dbg.decodeOnly);
    e = upload_vec(work, workDev);
    *nWork = (int)work.size();
  }
  if (e == cudaSuccess) e = upload_vec(fills, fillsDev);
  *nFill = (int)fills.size();
  if (e != cudaSuccess) { cudaFree(*parseDev); cudaFree(*workDev); *parseDev = nullptr; *workDev = nullptr; return fail_cuda(e, "upload work list"); }
  return VR_OK;
}

static cudaError_t plan_launch(vr_decode_plan* p, const vr_parse_item* parseDev, int nParse, const vr_kd_work* workDev, int nWork,
                               const vr_fill_job* fillsDev, int nFill, bool occ, int only, int tmaOut) {
  vr_volume* d = p->dst;
  if (p->persistent6)
    return kd_decode6_launch(p->geom, p->p6, p->jobsDev, parseDev, nParse, fillsDev, nFill, p->queueDev, p->lut6Dev, d->dev, d->SX, d->SX * d->SY,
                             occ ? d->occRaw : nullptr, only, p->ctx->stream, g_dbg.fillMode, tmaOut,
                             g_dbg.decodeSplit && p->ctx->side.stream ? &p->ctx->side : nullptr);
  return kd_decode_launch(p->geom, p->jobsDev, workDev, nWork, d->dev, d->SX, d->SX * d->SY, p->vec16, p->ctx->stream, occ ? d->occRaw : nullptr,
                          occ ? p->tileCellDev : nullptr, fillsDev);
}

// Work lists of the tiles in `needed` (n * 2^Dt bytes) plus the (cell, max) pairs of the others, into *lzp.
// Streams opened with a tile subset (kd_kdx.cpp) must hold every live tile the lists parse.
static int plan_build_culled(vr_decode_plan* p, const uint8_t* neededPtr, const std::vector<uint8_t>& solid, vr_decode_plan::Lazy* lzp) {
  vr_decode_plan::Lazy& lz = *lzp;
  const vr_kd_geom& g = p->geom;
  const uint32_t nt = 1u << g.Dt;
  const int n = p->n;
  std::vector<const KdIndex*> idx((size_t)n);
  for (int i = 0; i < n; ++i) idx[(size_t)i] = &p->streams[(size_t)i]->kd.index;
  for (int b = 0; b < n; ++b) {
    const std::vector<uint8_t>& res = p->streams[(size_t)b]->kd.residentTiles;
    if (res.empty()) continue;
    for (uint32_t t = 0; t < nt; ++t)
      if (neededPtr[(size_t)b * nt + t] && !res[t] && idx[(size_t)b]->tileLeafBlock[t] != VR_PRUNED)
        return fail(VR_ERR_STATE, "a tile this launch must parse is not resident (stream opened with a tile subset)");
  }
  struct NeededView { const uint8_t* p; const uint8_t* data() const { return p; } const uint8_t& operator[](size_t i) const { return p[i]; } } needed{neededPtr};
  if (cudaSetDevice(p->ctx->device) != cudaSuccess) return fail(VR_ERR_CUDA, "cudaSetDevice");
  std::vector<vr_occ_set> culled;
  int rc = plan_build_lists(p, needed.data(), &culled, &lz.parseDev, &lz.nParse, &lz.workDev, &lz.nWork, &lz.fillJobsDev, &lz.nFill, &lz.tmaOut);
  if (rc != VR_OK) return rc;
  lz.tilesTotal = (int64_t)n * nt;
  lz.tilesWritten = lz.tilesTotal - (int64_t)culled.size();
  lz.bytesVoxels = lz.tilesWritten * ((int64_t)g.tx * g.ty * g.tz);
  lz.bytesStream = 0;
  for (int b = 0; b < n; ++b)
    for (uint32_t t = 0; t < nt; ++t) {
      const KdIndex& ix = *idx[(size_t)b];
      const uint32_t blk = ix.tileLeafBlock[t];
      if (blk == VR_PRUNED || !needed[(size_t)b * nt + t]) continue;
      lz.bytesStream += ((int64_t)ix.blockSymEnd[blk] - (int64_t)ix.blockSymBegin[blk] + 3) / 4;
    }
  cudaError_t e = cudaSuccess;
  lz.nOccSet = (int)culled.size();
  if (lz.nOccSet) {
    static_assert(sizeof(vr_occ_set) == 16, "pairs of int64");
    e = cudaMalloc((void**)&lz.occSetDev, sizeof(vr_occ_set) * culled.size());
    if (e == cudaSuccess) e = cudaMemcpy(lz.occSetDev, culled.data(), sizeof(vr_occ_set) * culled.size(), cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess && !solid.empty()) {
    e = cudaMalloc((void**)&lz.solidDev, solid.size());
    if (e == cudaSuccess) e = cudaMemcpy(lz.solidDev, solid.data(), solid.size(), cudaMemcpyHostToDevice);
  }
  if (e != cudaSuccess) { p->lazy_free(&lz); return fail_cuda(e, "upload lazy work list"); }
  return VR_OK;
}

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
  int rc = plan_build_culled(p, needed.data(), solid, &lz);
  if (rc != VR_OK) return rc;
  lz.mode = mode; lz.thr = thr;
  for (int a = 0; a < 3; ++a) { lz.reach[a] = reach[a]; lz.stepReach[a] = stepReach[a]; }
  lz.built = true;
  return VR_OK;
}

// One lazy pass: decode what the frame can read, record the maxima of the rest.
int plan_lazy_run(vr_decode_plan* p, vr_decode_plan::Lazy* which) {
  vr_decode_plan::Lazy& lz = which ? *which : p->lazy;
  vr_volume* d = p->dst;
  if (!lz.built) return fail(VR_ERR_STATE, "lazy work list not built");
  if (!(d->occRaw && d->occEpoch == p->occEpoch && d->skipping)) return fail(VR_ERR_STATE, "occupancy grid of the destination changed");
  cudaError_t e = cudaSetDevice(p->ctx->device);
  if (e == cudaSuccess) e = plan_launch(p, lz.parseDev, lz.nParse, lz.workDev, lz.nWork, lz.fillJobsDev, lz.nFill, true, 0, lz.tmaOut);
  if (e == cudaSuccess) e = occ_set_cells_launch(d->occRaw, lz.occSetDev, lz.nOccSet, p->ctx->stream);
  if (e != cudaSuccess) return fail_cuda(e, "lazy decode launch");
  d->occDilDirty = true;
  if (lz.nOccSet) d->partial.insert(p);
  return VR_OK;
}
}  // namespace vr

extern "C" {

int vr_decode_plan_create(vr_context* ctx, vr_stream* const* streams, int32_t n, const int64_t* origins,
                          vr_volume* dst, vr_decode_plan** out) {
  if (!ctx || !streams || n <= 0 || !origins || !dst || !out) return fail(VR_ERR_ARG, "vr_decode_plan_create: bad argument");
  for (int i = 0; i < n; ++i)
    if (!streams[i]) return fail(VR_ERR_ARG, "vr_decode_plan_create: null stream");
  CU(cudaSetDevice(ctx->device));
  const vr_kd_geom& g = streams[0]->kd.geom;
  std::vector<vr_kd_job> jobs((size_t)n);
  bool vec16 = (g.tx >= 16) && (dst->SX % 16 == 0) && (((uintptr_t)dst->dev) % 16 == 0);
  vr_decode_plan* p = new vr_decode_plan;
  p->ctx = ctx; p->device = ctx->device; p->geom = g; p->n = n; p->dst = dst;
  dst->plans.insert(p);
  // the kernel keeps the occupancy grid current when every decode tile is exactly one cell (32x32x16)
  bool occFused = dst->skipping && g.fastPath && g.tx == 32 && g.ty == 32 && g.tz == 16 && volume_occ_alloc(dst) == VR_OK;
  bool sameChain = true;
  uint64_t chain0 = 0;
  p->treeDist.resize((size_t)n);
  for (int i = 0; i < n; ++i) {
    vr_stream* s = streams[i];
    if (!s->devImage || s->device != ctx->device) { vr_decode_plan_destroy(p); return fail(VR_ERR_STATE, "stream not uploaded to this device"); }
    const vr_kd_header& h = s->kd.hdr;
    if (h.X != g.X || h.Y != g.Y || h.Z != g.Z) { vr_decode_plan_destroy(p); return fail(VR_ERR_ARG, "all streams of a plan must share one brick shape"); }
    const int64_t ox = origins[3 * i], oy = origins[3 * i + 1], oz = origins[3 * i + 2];
    if (ox < 0 || oy < 0 || oz < 0 || ox + h.X > dst->SX || oy + h.Y > dst->SY || oz + h.Z > dst->SZ) {
      vr_decode_plan_destroy(p); return fail(VR_ERR_ARG, "brick placement outside the destination volume");
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
    p->treeDist[(size_t)i] = j.treeDist;
    if (i == 0) chain0 = j.chainDist;
    else if (j.chainDist != chain0) sameChain = false;
    j.dstOffset = (oz * dst->SY + oy) * dst->SX + ox;
    j.occBase = -1;
    if (occFused) {
      const int64_t l[3] = {ox + dst->shift[0], oy + dst->shift[1], oz + dst->shift[2]};
      if ((l[0] & 31) || (l[1] & 31) || (l[2] & 15)) occFused = false;
      else j.occBase = (l[0] >> 5) + (int64_t)dst->nc[0] * ((l[1] >> 5) + (int64_t)dst->nc[1] * (l[2] >> 4));
    }
    { int rc = context_chain_lut(ctx, j.chainDist, &j.chainLut); if (rc != VR_OK) { vr_decode_plan_destroy(p); return rc; } }
    int chainSum = 0;  // the tables pack the net offset in 8 signed bits: 64+32+...+1 = 127 for every reference stream
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
  p->streams.assign(streams, streams + n);
  p->origins.assign(origins, origins + 3 * (size_t)n);
  p->dstOffset.resize((size_t)n); p->occBase.resize((size_t)n);
  for (int i = 0; i < n; ++i) { p->dstOffset[(size_t)i] = jobs[(size_t)i].dstOffset; p->occBase[(size_t)i] = jobs[(size_t)i].occBase; }
  if (p->occFused) {
    std::vector<int32_t>& tc = p->tileCellHost;
    tc.resize(size_t(1) << g.Dt);
    for (uint32_t t = 0; t < tc.size(); ++t) {
      int ox, oy, oz;
      vr_tile_origin(g, t, &ox, &oy, &oz);
      tc[t] = (ox >> 5) + dst->nc[0] * ((oy >> 5) + dst->nc[1] * (oz >> 4));
    }
    cudaError_t et = upload_vec(tc, &p->tileCellDev);
    if (et != cudaSuccess) { vr_decode_plan_destroy(p); return fail_cuda(et, "upload tile cells"); }
  }
  // which kernel: the persistent one wherever the brick shape, the destination and the streams allow it
  p->persistent6 = g_dbg.decodeKernel == 0 && vec16 && p->geom.fastPath && sameChain &&
                   kd_decode6_plan(p->geom, dst->dev, dst->SX, dst->SY, dst->SZ, &p->p6, g_dbg.tmaRows);
  if (p->persistent6) {
    int rc = context_chain_lut6(ctx, chain0, &p->lut6Dev);
    if (rc == VR_OK && cudaMalloc((void**)&p->queueDev, VR6_QUEUE_WORDS * sizeof(uint32_t)) != cudaSuccess) rc = fail(VR_ERR_CUDA, "cudaMalloc(queue)");
    if (rc == VR_OK && !ctx->side.stream) {  // split launch (kd_decode.h): one side stream + two events per context
      cudaError_t es = cudaStreamCreateWithFlags(&ctx->side.stream, cudaStreamNonBlocking);
      if (es == cudaSuccess) es = cudaEventCreateWithFlags(&ctx->side.fork, cudaEventDisableTiming);
      if (es == cudaSuccess) es = cudaEventCreateWithFlags(&ctx->side.join, cudaEventDisableTiming);
      if (es != cudaSuccess) rc = fail_cuda(es, "side stream of the decode launch");
    }
    if (rc != VR_OK) { vr_decode_plan_destroy(p); return rc; }
  }
  int rc = plan_build_lists(p, nullptr, nullptr, &p->parseDev, &p->nParse, &p->workDev, &p->nWork, &p->fillJobsDev, &p->nFill, &p->tmaOut);
  if (rc != VR_OK) { vr_decode_plan_destroy(p); return rc; }
  cudaError_t e = cudaMalloc((void**)&p->jobsDev, sizeof(vr_kd_job) * (size_t)n);
  if (e == cudaSuccess) e = cudaMemcpyAsync(p->jobsDev, jobs.data(), sizeof(vr_kd_job) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // jobs is a stack-lifetime buffer
  if (e != cudaSuccess) { vr_decode_plan_destroy(p); return fail_cuda(e, "upload jobs"); }
  *out = p;
  return VR_OK;
}
int vr_decode_plan_run(vr_decode_plan* p) {
  if (!p) return fail(VR_ERR_ARG, "null plan");
  if (!p->dst) return fail(VR_ERR_STATE, "vr_decode_plan_run: the plan's destination volume was destroyed");
  CU(cudaSetDevice(p->ctx->device));
  vr_volume* d = p->dst;
  for (const vr_stream* s : p->streams)
    if (!s->kd.residentTiles.empty())
      return fail(VR_ERR_STATE, "vr_decode_plan_run: a stream holds a tile subset only (vr_stream_open_indexed with a mask); use vr_decode_plan_run_tiles");
  const bool occ = p->occFused && d->occRaw && d->occEpoch == p->occEpoch && d->skipping;
  CU(plan_launch(p, p->parseDev, p->nParse, p->workDev, p->nWork, p->fillJobsDev, p->nFill, occ, g_dbg.decodeOnly, p->tmaOut));
  if (occ) d->occDilDirty = true;  // cells of the decoded bricks are current; validity of the rest is unchanged
  else d->occValid = false;
  d->partial.erase(p);  // every tile of the plan holds its voxels again
  return VR_OK;
}
int vr_decode_plan_run_tiles(vr_decode_plan* p, const uint8_t* needed, int64_t stats[4]) {
  if (!p || !needed) return fail(VR_ERR_ARG, "vr_decode_plan_run_tiles: null argument");
  if (!p->dst) return fail(VR_ERR_STATE, "vr_decode_plan_run_tiles: the plan's destination volume was destroyed");
  if (!p->occFused) return fail(VR_ERR_STATE, "vr_decode_plan_run_tiles needs 32x32x16 tiles that coincide with the destination's occupancy cells");
  CU(cudaSetDevice(p->ctx->device));
  p->lazy_free(&p->subset);
  int rc = plan_build_culled(p, needed, std::vector<uint8_t>(), &p->subset);
  if (rc != VR_OK) return rc;
  p->subset.built = true;
  rc = plan_lazy_run(p, &p->subset);
  if (rc != VR_OK) return rc;
  if (stats) { stats[0] = p->subset.tilesTotal; stats[1] = p->subset.tilesWritten; stats[2] = p->subset.bytesVoxels; stats[3] = p->subset.bytesStream; }
  return VR_OK;
}
int vr_decode_plan_destroy(vr_decode_plan* p) {
  if (!p) return VR_OK;
  cudaSetDevice(p->device);  // not through p->ctx: the host may have closed the context first
  cudaFree(p->jobsDev);
  cudaFree(p->workDev);
  cudaFree(p->parseDev);
  cudaFree(p->fillJobsDev);
  cudaFree(p->tileCellDev);
  cudaFree(p->queueDev);
  p->lazy_free();
  p->lazy_free(&p->subset);
  if (p->dst) {
    if (p->dst->partial.erase(p)) p->dst->partialOrphan = true;  // its stale tiles stay stale
    p->dst->plans.erase(p);
  }
  delete p;
  return VR_OK;
}
int vr_decode_plan_kernel(const vr_decode_plan* p, int32_t* which) {
  if (!p || !which) return fail(VR_ERR_ARG, "null");
  *which = p->persistent6 ? 0 : 1;
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
  int rc = kd_build_device(ctx->builder, volDev, X, Y, Z, rowPitch, slicePitch, tol, epochs, ctx->stream, &s->kd, &st, &ce, iters, g_dbg.producerIndex != 0);
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
  if (vol->is_partial()) return fail(VR_ERR_STATE, "vr_kd_compress_volume: the volume was decoded lazily for one frame; run its plan in full first");
  CU(cudaSetDevice(ctx->device));
  const uint8_t* p = vol->dev + (origin[2] * vol->SY + origin[1]) * vol->SX + origin[0];
  return build_common(ctx, p, X, Y, Z, vol->SX, vol->SX * vol->SY, tol, epochs, out, stats, fixed_point_iterations);
}

}  // extern "C"
