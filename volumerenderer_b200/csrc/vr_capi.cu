// vr_capi.cu -- the C ABI (include/vr_b200.h) over the C++/CUDA implementation.  No CPU fallback:
// every decode/render entry point needs a usable CUDA device and reports VR_ERR_CUDA otherwise.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vr_b200.h"
#include "kd_decode.h"
#include "kd_stream.h"
#include "vr_internal.h"

using namespace vr;

static thread_local std::string g_err;
namespace vr {
int fail(int code, const char* msg) { g_err = msg ? msg : ""; return code; }
int fail_cuda(cudaError_t e, const char* where) {
  g_err = std::string(where) + ": " + cudaGetErrorString(e);
  return VR_ERR_CUDA;
}
}  // namespace vr
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

static_assert(sizeof(vr_stream_header) == sizeof(vr_kd_header) && sizeof(vr_kd_header) == 88, "container header is 88 bytes");

extern "C" {

const char* vr_last_error(void) { return g_err.c_str(); }
const char* vr_version(void) { return "vr_b200 0.1 (sm_100a)"; }

// ---- context ------------------------------------------------------------------------------------
int vr_context_create(int device, vr_context** out) {
  if (!out) return fail(VR_ERR_ARG, "vr_context_create: null out");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return fail_cuda(e, "cudaGetDeviceCount");
  if (device < 0 || device >= n) return fail(VR_ERR_CUDA, "vr_context_create: no such CUDA device");
  CU(cudaSetDevice(device));
  vr_context* c = new vr_context;
  c->device = device;
  c->ownsStream = true;
  e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete c; return fail_cuda(e, "cudaStreamCreate"); }
  *out = c;
  return VR_OK;
}
int vr_context_destroy(vr_context* c) {
  if (!c) return VR_OK;
  cudaSetDevice(c->device);
  if (c->ownsStream) cudaStreamDestroy(c->stream);
  for (auto& e : c->ev) for (auto& k : e) if (k) cudaEventDestroy(k);
  if (c->evOrder) cudaEventDestroy(c->evOrder);
  for (auto& kv : c->chainLuts) cudaFree(kv.second);
  for (auto& kv : c->chainLuts6) cudaFree(kv.second);
  if (c->side.stream) cudaStreamDestroy(c->side.stream);
  if (c->side.fork) cudaEventDestroy(c->side.fork);
  if (c->side.join) cudaEventDestroy(c->side.join);
  vr::kd_builder_destroy(c->builder);
  cudaFree(c->frameDev); cudaFree(c->frameHitDev); cudaFree(c->frameCountDev);
  if (c->stagingDone) { cudaEventSynchronize(c->stagingDone); cudaEventDestroy(c->stagingDone); }
  if (c->stagingHost) cudaFreeHost(c->stagingHost);
  delete c;
  return VR_OK;
}
int vr_context_set_stream(vr_context* c, void* s) {
  if (!c) return fail(VR_ERR_ARG, "null context");
  if (c->ownsStream) cudaStreamDestroy(c->stream);
  c->ownsStream = false;
  c->stream = (cudaStream_t)s;
  return VR_OK;
}
int vr_context_synchronize(vr_context* c) {
  if (!c) return fail(VR_ERR_ARG, "null context");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  return VR_OK;
}

int vr_context_wait_for(vr_context* c, vr_context* other) {
  if (!c || !other) return fail(VR_ERR_ARG, "null context");
  if (c == other || c->stream == other->stream) return VR_OK;  // same stream: already ordered
  CU(cudaSetDevice(other->device));
  if (!other->evOrder) CU(cudaEventCreateWithFlags(&other->evOrder, cudaEventDisableTiming));
  CU(cudaEventRecord(other->evOrder, other->stream));
  CU(cudaSetDevice(c->device));
  CU(cudaStreamWaitEvent(c->stream, other->evOrder, 0));
  return VR_OK;
}
int vr_context_download(vr_context* c, void* dst, const void* src, int64_t nbytes) {
  if (!c || !dst || !src || nbytes < 0) return fail(VR_ERR_ARG, "vr_context_download: bad argument");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpyAsync(dst, src, (size_t)nbytes, cudaMemcpyDeviceToHost, c->stream));
  return VR_OK;
}

int vr_timer_begin(vr_context* c, int slot) {
  if (!c || slot < 0 || slot >= 16) return fail(VR_ERR_ARG, "bad timer slot");
  CU(cudaSetDevice(c->device));
  for (int k = 0; k < 2; ++k)
    if (!c->ev[slot][k]) CU(cudaEventCreate(&c->ev[slot][k]));
  CU(cudaEventRecord(c->ev[slot][0], c->stream));
  return VR_OK;
}
int vr_timer_end(vr_context* c, int slot) {
  if (!c || slot < 0 || slot >= 16 || !c->ev[slot][1]) return fail(VR_ERR_ARG, "bad timer slot");
  CU(cudaEventRecord(c->ev[slot][1], c->stream));
  return VR_OK;
}
int vr_timer_elapsed_ms(vr_context* c, int slot, float* ms) {
  if (!c || !ms || slot < 0 || slot >= 16 || !c->ev[slot][1]) return fail(VR_ERR_ARG, "bad timer slot");
  CU(cudaEventSynchronize(c->ev[slot][1]));
  CU(cudaEventElapsedTime(ms, c->ev[slot][0], c->ev[slot][1]));
  return VR_OK;
}

// ---- stream container -----------------------------------------------------------------------------
int vr_stream_from_memory(const vr_stream_header* hdr, const uint8_t* dm, const uint8_t* bits, int64_t nbytes,
                          vr_stream** out) {
  if (!out) return fail(VR_ERR_ARG, "null out");
  vr_stream* s = new vr_stream;
  int rc = kd_from_memory(reinterpret_cast<const vr_kd_header*>(hdr), dm, bits, nbytes, &s->kd);
  if (rc != VR_OK) { delete s; return fail(rc, "vr_stream_from_memory: bad header or short bit array"); }
  *out = s;
  return VR_OK;
}
int vr_stream_open(const char* path, vr_stream** out) {
  if (!out) return fail(VR_ERR_ARG, "null out");
  vr_stream* s = new vr_stream;
  int rc = kd_open_any(path, &s->kd);  // VolumeKdtree::open; an indexed container (kd_kdx.cpp) is recognised by its magic
  if (rc != VR_OK) { delete s; return fail(rc, rc == VR_ERR_IO ? "Tree File does not exist or is short" : "bad tree file"); }
  *out = s;
  return VR_OK;
}
int vr_stream_save(const vr_stream* s, const char* path) {
  if (!s) return fail(VR_ERR_ARG, "null stream");
  int rc = kd_save(s->kd, path);
  return rc == VR_OK ? VR_OK : fail(rc, "vr_stream_save failed (no tree to save, or I/O error)");
}
// ---- indexed container (kd_kdx.cpp) ----------------------------------------------------------------
int vr_stream_save_indexed(vr_stream* s, const char* path) {
  if (!s || !path) return fail(VR_ERR_ARG, "vr_stream_save_indexed: null argument");
  int rc = kdx_save(&s->kd, path);
  return rc == VR_OK ? VR_OK : fail(rc, "vr_stream_save_indexed failed (no tree, tile subset, stream does not parse, or I/O error)");
}
int vr_stream_serialize_indexed(vr_stream* s, uint8_t* out, int64_t cap, int64_t* nbytes) {
  if (!s || !nbytes) return fail(VR_ERR_ARG, "vr_stream_serialize_indexed: null argument");
  std::vector<uint8_t> buf;
  int rc = kdx_serialize(&s->kd, &buf);
  if (rc != VR_OK) return fail(rc, "vr_stream_serialize_indexed failed");
  *nbytes = (int64_t)buf.size();
  if (!out) return VR_OK;
  if (cap < (int64_t)buf.size()) return fail(VR_ERR_ARG, "vr_stream_serialize_indexed: buffer too small");
  std::memcpy(out, buf.data(), buf.size());
  return VR_OK;
}
static int kdx_fail(int rc) {
  return fail(rc, rc == VR_ERR_IO ? "indexed container missing or truncated" : "bad indexed container (magic, geometry, index or checksum)");
}
int vr_stream_open_indexed(const char* path, const uint8_t* tile_mask, int32_t verify, vr_stream** out, int64_t* bytes_read) {
  if (!path || !out) return fail(VR_ERR_ARG, "vr_stream_open_indexed: null argument");
  vr_stream* s = new vr_stream;
  int rc = kdx_open(path, tile_mask, verify != 0, &s->kd, bytes_read);
  if (rc != VR_OK) { delete s; return kdx_fail(rc); }
  *out = s;
  return VR_OK;
}
int vr_stream_from_memory_indexed(const uint8_t* image, int64_t nbytes, const uint8_t* tile_mask, int32_t verify, vr_stream** out) {
  if (!image || !out) return fail(VR_ERR_ARG, "vr_stream_from_memory_indexed: null argument");
  vr_stream* s = new vr_stream;
  int rc = kdx_from_memory(image, nbytes, tile_mask, verify != 0, &s->kd);
  if (rc != VR_OK) { delete s; return kdx_fail(rc); }
  *out = s;
  return VR_OK;
}
static int ensure_index(vr_stream* s) {
  if (s->kd.index.built) return VR_OK;
  int rc = kd_prepare(&s->kd);
  return rc == VR_OK ? VR_OK : fail(rc, "stream does not parse");
}
int vr_stream_index_info(vr_stream* s, int64_t out[6]) {
  if (!s || !out) return fail(VR_ERR_ARG, "vr_stream_index_info: null argument");
  int rc = ensure_index(s);
  if (rc != VR_OK) return rc;
  const KdIndex& ix = s->kd.index;
  out[0] = (int64_t)ix.tileLeafBlock.size();
  out[1] = ix.numLiveTiles;
  out[2] = int64_t(1) << (s->kd.geom.tileBits - s->kd.geom.leafBits);
  out[3] = s->kd.residentTiles.empty() ? 0 : 1;
  out[4] = (int64_t)(ix.tileLeafBlock.size() * 7 + ix.blockSymBegin.size() * 8 + ix.leafOff.size() * 5);
  out[5] = ix.maxSymbolsPerLeaf;
  return VR_OK;
}
int vr_stream_index_section(vr_stream* s, int32_t which, const void** data, int64_t* nbytes) {
  if (!s || !data || !nbytes) return fail(VR_ERR_ARG, "vr_stream_index_section: null argument");
  int rc = ensure_index(s);
  if (rc != VR_OK) return rc;
  const KdIndex& ix = s->kd.index;
  switch (which) {
    case 1: *data = ix.tileLeafBlock.data(); *nbytes = (int64_t)ix.tileLeafBlock.size() * 4; break;
    case 2: *data = ix.tileVal.data(); *nbytes = (int64_t)ix.tileVal.size(); break;
    case 3: *data = ix.tileMin.data(); *nbytes = (int64_t)ix.tileMin.size(); break;
    case 4: *data = ix.tileMax.data(); *nbytes = (int64_t)ix.tileMax.size(); break;
    case 5: *data = ix.blockSymBegin.data(); *nbytes = (int64_t)ix.blockSymBegin.size() * 4; break;
    case 6: *data = ix.blockSymEnd.data(); *nbytes = (int64_t)ix.blockSymEnd.size() * 4; break;
    case 7: *data = ix.leafOff.data(); *nbytes = (int64_t)ix.leafOff.size() * 4; break;
    case 8: *data = ix.leafVal.data(); *nbytes = (int64_t)ix.leafVal.size(); break;
    default: return fail(VR_ERR_ARG, "vr_stream_index_section: sections are numbered 1..8");
  }
  return VR_OK;
}

int vr_stream_destroy(vr_stream* s) {
  if (!s) return VR_OK;
  s->release();
  delete s;
  return VR_OK;
}
int vr_stream_get_header(const vr_stream* s, vr_stream_header* hdr) {
  if (!s || !hdr) return fail(VR_ERR_ARG, "null");
  std::memcpy(hdr, &s->kd.hdr, sizeof(*hdr));
  return VR_OK;
}
int vr_stream_get_distance_map(const vr_stream* s, uint8_t* out, int32_t cap) {
  if (!s || !out) return fail(VR_ERR_ARG, "null");
  if (cap < (int32_t)s->kd.distanceMap.size()) return fail(VR_ERR_ARG, "distance map buffer too small");
  std::memcpy(out, s->kd.distanceMap.data(), s->kd.distanceMap.size());
  return VR_OK;
}
int vr_stream_get_bits(const vr_stream* s, const uint8_t** bits, int64_t* nbytes) {
  if (!s || !bits || !nbytes) return fail(VR_ERR_ARG, "null");
  *bits = s->kd.bits.data();
  *nbytes = (int64_t)s->kd.bits.size();
  return VR_OK;
}
int vr_stream_validate(const vr_stream* s, int64_t* nodes) {
  if (!s) return fail(VR_ERR_ARG, "null stream");
  int64_t n = 0;
  int rc = kd_validate(s->kd, &n);
  if (nodes) *nodes = n;
  return rc == VR_OK ? VR_OK : fail(rc, "stream does not parse to numActiveNodes symbols");
}

// ---- MidRangeTree 4-bit shader stream ------------------------------------------------------------------
static inline int sym2(const uint8_t* bits, int64_t i) { return (bits[i >> 2] >> (2 * (int)(i & 3))) & 3; }

int vr_nibbles_pack(const uint8_t* tree, const uint8_t* range, int64_t n, uint8_t* out, int64_t cap, int64_t* out_len) {
  if (!tree || !range || n < 0 || !out_len) return fail(VR_ERR_ARG, "vr_nibbles_pack: bad argument");
  const int64_t need = (n + 1) / 2;
  int64_t len = 1;
  while (len < need) len <<= 1;  // MidRangeTree.cpp:1098-1106: buffer length rounded up to a power of two
  *out_len = len;
  if (!out) return VR_OK;
  if (cap < len) return fail(VR_ERR_ARG, "vr_nibbles_pack: buffer too small");
  std::memset(out, 0, (size_t)len);
  for (int64_t i = 0; i < n; ++i) {
    const int nib = (sym2(tree, i) << 2) | sym2(range, i);
    out[i >> 1] |= (uint8_t)((i & 1) ? nib : nib << 4);  // even node in the high nibble (:1124)
  }
  return VR_OK;
}

int vr_nibbles_unpack(const uint8_t* nib, int64_t nbytes, int64_t n, uint8_t* tree_out, uint8_t* range_out) {
  if (!nib || n < 0 || (n + 1) / 2 > nbytes || (!tree_out && !range_out)) return fail(VR_ERR_ARG, "vr_nibbles_unpack: bad argument");
  const size_t nb = (size_t)((n + 3) / 4);
  if (tree_out) std::memset(tree_out, 0, nb);
  if (range_out) std::memset(range_out, 0, nb);
  for (int64_t i = 0; i < n; ++i) {
    const int v = (i & 1) ? (nib[i >> 1] & 15) : (nib[i >> 1] >> 4);
    if (tree_out) tree_out[i >> 2] |= (uint8_t)((v >> 2) << (2 * (int)(i & 3)));
    if (range_out) range_out[i >> 2] |= (uint8_t)((v & 3) << (2 * (int)(i & 3)));
  }
  return VR_OK;
}

int vr_stream_from_nibbles(const vr_stream_header* hdr, const uint8_t* dm, const uint8_t* nib, int64_t nbytes, vr_stream** out,
                           uint8_t* range_out) {
  if (!hdr || !dm || !nib || !out) return fail(VR_ERR_ARG, "vr_stream_from_nibbles: null argument");
  const int64_t n = hdr->numActiveNodes;
  if (n <= 0 || (n + 1) / 2 > nbytes) return fail(VR_ERR_FORMAT, "4-bit stream shorter than numActiveNodes");
  std::vector<uint8_t> mid((size_t)((n + 3) / 4));
  int rc = vr_nibbles_unpack(nib, nbytes, n, mid.data(), range_out);
  if (rc != VR_OK) return rc;
  return vr_stream_from_memory(hdr, dm, mid.data(), (int64_t)mid.size(), out);
}

// ---- MidRangeTree::save / open (kd_mrt.cpp) -----------------------------------------------------------
int vr_mrt_save(const char* path, const vr_stream_header* hdr, const uint8_t* dm, const uint8_t* dm_range, const uint8_t* tree_bits,
                const uint8_t* range_bits, int64_t tree_nbytes) {
  if (!path || !hdr) return fail(VR_ERR_ARG, "vr_mrt_save: null argument");
  int rc = mrt_save(*reinterpret_cast<const vr_kd_header*>(hdr), dm, dm_range, tree_bits, range_bits, tree_nbytes, path);
  return rc == VR_OK ? VR_OK : fail(rc, "vr_mrt_save failed (no tree to save, bad depth, or I/O error)");
}
int vr_mrt_open(const char* path, vr_stream** mid_out, uint8_t* dm_range_out, int32_t dm_cap, uint8_t* range_bits_out, int64_t cap,
                int64_t* tree_nbytes) {
  if (!path || !mid_out || !tree_nbytes) return fail(VR_ERR_ARG, "vr_mrt_open: null argument");
  vr_stream* s = new vr_stream;
  std::vector<uint8_t> dmr, rb;
  int rc = mrt_open(path, &s->kd, &dmr, &rb);
  if (rc != VR_OK) { delete s; return fail(rc, rc == VR_ERR_IO ? "Tree File does not exist or is short" : "bad MidRangeTree file"); }
  *tree_nbytes = (int64_t)rb.size();
  if ((dm_range_out && dm_cap < (int32_t)dmr.size()) || (range_bits_out && cap < (int64_t)rb.size())) {
    delete s;
    return fail(VR_ERR_ARG, "vr_mrt_open: output buffer too small");
  }
  if (dm_range_out) std::memcpy(dm_range_out, dmr.data(), dmr.size());
  if (range_bits_out) std::memcpy(range_bits_out, rb.data(), rb.size());
  *mid_out = s;
  return VR_OK;
}

int vr_kd_compress(const uint8_t* brick, int64_t X, int64_t Y, int64_t Z, int tol, int epochs, vr_stream** out,
                   vr_build_stats* stats) {
  if (!brick || !out) return fail(VR_ERR_ARG, "null");
  vr_stream* s = new vr_stream;
  KdBuildStats st;
  int rc = kd_compress(brick, X, Y, Z, tol, epochs, &s->kd, &st);
  if (rc != VR_OK) { delete s; return fail(rc, "vr_kd_compress: extents must be powers of two"); }
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

}  // extern "C"
