// vr_capi_peer.cu -- peer-memory exchange (include/vr_b200.h, "peer-memory exchange"): per-rank windows mapped by all
// ranks of a node over CUDA IPC, data moved by one store kernel per step over NVLink, flags for completion.
// No counterpart in the reference (single GPU); replaces the NCCL calls of the sort-last frame.
#include <cuda_runtime.h>

#include <cstring>
#include <vector>

#include "../../include/vr_b200.h"
#include "render_march.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

namespace {

constexpr int MAX_PEERS = 32;
// Flag area behind every window (uint32 words): [slot][src] arrival counters written by the peers, then this rank's own
// bookkeeping: puts issued per (slot, dst), waits issued per (slot, src), error word, ticket.  All counters live on
// the device and are advanced by the kernels themselves, so a captured CUDA graph of a step can be replayed.
constexpr int FLAG_BYTES = 4096;
constexpr int SENT_WORD = VR_PEER_SLOTS * MAX_PEERS;
constexpr int SEEN_WORD = 2 * VR_PEER_SLOTS * MAX_PEERS;
constexpr int ERR_WORD = 3 * VR_PEER_SLOTS * MAX_PEERS;
constexpr int TICKET_WORD = ERR_WORD + 1;
static_assert((TICKET_WORD + 1) * 4 <= FLAG_BYTES, "flag area");

struct PutParams {
  uint8_t* dst[MAX_PEERS];
  const uint8_t* src[MAX_PEERS];
  int64_t bytes[MAX_PEERS];
  uint32_t* flag[MAX_PEERS];  // one per distinct destination
  uint32_t* sent[MAX_PEERS];  // local counter of puts to that destination on this slot
  int count, nFlags, ctasPerCopy;
  unsigned int* ticket;  // local
};

// grid = count * ctasPerCopy.  Stores go straight to the destination windows (peer memory over NVLink, or local).
__global__ void __launch_bounds__(256) peer_put_kernel(const __grid_constant__ PutParams p) {
  const int k = blockIdx.x / p.ctasPerCopy, part = blockIdx.x % p.ctasPerCopy;
  const int64_t n16 = p.bytes[k] >> 4;
  const uint4* s = reinterpret_cast<const uint4*>(p.src[k]);
  uint4* d = reinterpret_cast<uint4*>(p.dst[k]);
  for (int64_t i = (int64_t)part * blockDim.x + threadIdx.x; i < n16; i += (int64_t)p.ctasPerCopy * blockDim.x) d[i] = s[i];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(p.ticket, 1u);
    if (t == gridDim.x - 1) {  // every CTA has stored and fenced: publish
      *p.ticket = 0;
      __threadfence_system();
      for (int f = 0; f < p.nFlags; ++f) {
        const uint32_t v = *p.sent[f] + 1u;
        *p.sent[f] = v;
        *reinterpret_cast<volatile uint32_t*>(p.flag[f]) = v;
      }
    }
  }
}

struct WaitParams {
  const uint32_t* flag[MAX_PEERS];  // in this rank's own window
  uint32_t* seen[MAX_PEERS];        // local counter of waits on that source and slot
  int src[MAX_PEERS];
  int count, slot;
  unsigned int* err;
  long long limit;  // clock cycles
};
__global__ void peer_wait_kernel(const __grid_constant__ WaitParams p) {
  if ((int)threadIdx.x < p.count) {
    const volatile uint32_t* f = p.flag[threadIdx.x];
    const uint32_t want = *p.seen[threadIdx.x] + 1u;
    *p.seen[threadIdx.x] = want;
    const long long t0 = clock64();
    while ((int32_t)(*f - want) < 0) {
      if (clock64() - t0 > p.limit) { atomicExch(p.err, 0x10000u | ((unsigned)p.slot << 8) | (unsigned)p.src[threadIdx.x]); break; }  // which wait expired
      __nanosleep(200);
    }
  }
  __threadfence_system();
}

struct ReduceParams {
  const void* in[MAX_PEERS];
  void* out;
  int64_t count4;  // groups of four elements
  int n, op;
};
__global__ void __launch_bounds__(256) peer_reduce_kernel(const __grid_constant__ ReduceParams p) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < p.count4; i += (int64_t)gridDim.x * blockDim.x) {
    if (p.op == 0) {
      float4 a = reinterpret_cast<const float4*>(p.in[0])[i];
      for (int k = 1; k < p.n; ++k) {
        const float4 b = reinterpret_cast<const float4*>(p.in[k])[i];
        a.x = __fadd_rn(a.x, b.x); a.y = __fadd_rn(a.y, b.y); a.z = __fadd_rn(a.z, b.z); a.w = __fadd_rn(a.w, b.w);
      }
      reinterpret_cast<float4*>(p.out)[i] = a;
    } else {
      int4 a = reinterpret_cast<const int4*>(p.in[0])[i];
      for (int k = 1; k < p.n; ++k) {
        const int4 b = reinterpret_cast<const int4*>(p.in[k])[i];
        a.x = min(a.x, b.x); a.y = min(a.y, b.y); a.z = min(a.z, b.z); a.w = min(a.w, b.w);
      }
      reinterpret_cast<int4*>(p.out)[i] = a;
    }
  }
}

}  // namespace

struct vr_peer {
  vr_context* ctx = nullptr;
  int rank = 0, n = 0;
  int64_t windowBytes = 0;
  uint8_t* window = nullptr;  // local allocation: [windowBytes | FLAG_BYTES]
  uint8_t* mapped[MAX_PEERS] = {};
  bool ipc[MAX_PEERS] = {};
  bool opened = false;
  long long limit = 2000000000LL;
  uint32_t* flags_of(int r) const { return reinterpret_cast<uint32_t*>(mapped[r] + windowBytes); }
};

namespace vr {
int peer_targets(vr_peer* p, int slot, PeerTargets* t, int64_t* windowBytes) {
  if (!p || !t || slot < 0 || slot >= VR_PEER_SLOTS) return fail(VR_ERR_ARG, "peer_targets: bad argument");
  if (!p->opened) return fail(VR_ERR_STATE, "peer windows not opened");
  static_assert(VR_MAX_PEERS == MAX_PEERS, "peer table size");
  std::memset(t, 0, sizeof(*t));
  t->n = p->n; t->rank = p->rank;
  for (int r = 0; r < p->n; ++r) {
    t->base[r] = p->mapped[r];
    t->flag[r] = p->flags_of(r) + slot * MAX_PEERS + p->rank;
    t->sent[r] = p->flags_of(p->rank) + SENT_WORD + slot * MAX_PEERS + r;
  }
  t->ticket = reinterpret_cast<unsigned int*>(p->flags_of(p->rank) + TICKET_WORD);
  if (windowBytes) *windowBytes = p->windowBytes;
  return VR_OK;
}
}  // namespace vr

extern "C" {

int vr_peer_create(vr_context* ctx, int32_t rank, int32_t n, int64_t window_bytes, vr_peer** out) {
  if (!ctx || !out || n < 1 || n > MAX_PEERS || rank < 0 || rank >= n || window_bytes < 0 || (window_bytes & 15))
    return fail(VR_ERR_ARG, "vr_peer_create: bad argument (at most 32 ranks, window a multiple of 16 bytes)");
  CU(cudaSetDevice(ctx->device));
  vr_peer* p = new vr_peer;
  p->ctx = ctx; p->rank = rank; p->n = n; p->windowBytes = window_bytes;
  cudaError_t e = cudaMalloc((void**)&p->window, (size_t)window_bytes + FLAG_BYTES);
  if (e == cudaSuccess) e = cudaMemset(p->window + window_bytes, 0, FLAG_BYTES);
  if (e != cudaSuccess) { cudaFree(p->window); delete p; return fail_cuda(e, "vr_peer_create"); }
  int khz = 0;
  if (cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device) == cudaSuccess && khz > 0) p->limit = (long long)khz * 1000;  // 1 s
  p->mapped[rank] = p->window;
  *out = p;
  return VR_OK;
}

int vr_peer_destroy(vr_peer* p) {
  if (!p) return VR_OK;
  cudaSetDevice(p->ctx->device);
  cudaStreamSynchronize(p->ctx->stream);
  for (int r = 0; r < p->n; ++r)
    if (p->ipc[r] && p->mapped[r]) cudaIpcCloseMemHandle(p->mapped[r]);
  cudaFree(p->window);
  delete p;
  return VR_OK;
}

int vr_peer_handle(const vr_peer* p, uint8_t out[VR_PEER_HANDLE_BYTES]) {
  if (!p || !out) return fail(VR_ERR_ARG, "null");
  static_assert(sizeof(cudaIpcMemHandle_t) == VR_PEER_HANDLE_BYTES, "IPC handle size");
  CU(cudaSetDevice(p->ctx->device));
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, p->window));
  std::memcpy(out, &h, sizeof(h));
  return VR_OK;
}

int vr_peer_open(vr_peer* p, const uint8_t* handles) {
  if (!p || !handles) return fail(VR_ERR_ARG, "null");
  CU(cudaSetDevice(p->ctx->device));
  for (int r = 0; r < p->n; ++r) {
    if (r == p->rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handles + (size_t)r * VR_PEER_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    CU(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    p->mapped[r] = (uint8_t*)ptr;
    p->ipc[r] = true;
  }
  p->opened = true;
  return VR_OK;
}

int vr_peer_raw(const vr_peer* p, void** window) {
  if (!p || !window) return fail(VR_ERR_ARG, "null");
  *window = p->window;
  return VR_OK;
}

int vr_peer_open_raw(vr_peer* p, void* const* windows) {
  if (!p || !windows) return fail(VR_ERR_ARG, "null");
  for (int r = 0; r < p->n; ++r)
    if (r != p->rank) p->mapped[r] = (uint8_t*)windows[r];
  p->opened = true;
  return VR_OK;
}

int vr_peer_window(const vr_peer* p, int32_t rank, void** ptr) {
  if (!p || !ptr || rank < 0 || rank >= p->n || !p->mapped[rank]) return fail(VR_ERR_ARG, "vr_peer_window: bad rank or not opened");
  *ptr = p->mapped[rank];
  return VR_OK;
}

int vr_peer_put(vr_peer* p, int32_t slot, int32_t count, const int32_t* dst_rank, const int64_t* dst_off, const void* const* src_dev,
                const int64_t* bytes) {
  if (!p || slot < 0 || slot >= VR_PEER_SLOTS || count < 1 || count > MAX_PEERS || !dst_rank || !dst_off || !src_dev || !bytes)
    return fail(VR_ERR_ARG, "vr_peer_put: bad argument");
  if (!p->opened) return fail(VR_ERR_STATE, "vr_peer_put: windows not opened");
  CU(cudaSetDevice(p->ctx->device));
  PutParams q;
  std::memset(&q, 0, sizeof(q));
  q.count = count;
  int64_t most = 0;
  bool flagged[MAX_PEERS] = {};
  for (int k = 0; k < count; ++k) {
    const int d = dst_rank[k];
    if (d < 0 || d >= p->n || dst_off[k] < 0 || bytes[k] < 0 || dst_off[k] + bytes[k] > p->windowBytes || ((dst_off[k] | bytes[k]) & 15) ||
        ((uintptr_t)src_dev[k] & 15))
      return fail(VR_ERR_ARG, "vr_peer_put: copy outside the window or not 16-byte aligned");
    q.dst[k] = p->mapped[d] + dst_off[k];
    q.src[k] = (const uint8_t*)src_dev[k];
    q.bytes[k] = bytes[k];
    if (bytes[k] > most) most = bytes[k];
    if (!flagged[d]) {
      flagged[d] = true;
      q.flag[q.nFlags] = p->flags_of(d) + slot * MAX_PEERS + p->rank;
      q.sent[q.nFlags] = p->flags_of(p->rank) + SENT_WORD + slot * MAX_PEERS + d;
      ++q.nFlags;
    }
  }
  // enough CTAs to keep the links busy, not more than the data gives work to
  int per = (int)((most + 65535) / 65536);
  if (per < 1) per = 1;
  const int cap = 296 / count > 1 ? 296 / count : 1;
  if (per > cap) per = cap;
  q.ctasPerCopy = per;
  q.ticket = reinterpret_cast<unsigned int*>(p->flags_of(p->rank) + TICKET_WORD);
  peer_put_kernel<<<(unsigned)(count * per), 256, 0, p->ctx->stream>>>(q);
  CU(cudaGetLastError());
  return VR_OK;
}

int vr_peer_wait(vr_peer* p, int32_t slot, int32_t count, const int32_t* src_rank) {
  if (!p || slot < 0 || slot >= VR_PEER_SLOTS || count < 1 || count > MAX_PEERS || !src_rank) return fail(VR_ERR_ARG, "vr_peer_wait: bad argument");
  CU(cudaSetDevice(p->ctx->device));
  WaitParams q;
  std::memset(&q, 0, sizeof(q));
  q.count = count;
  for (int k = 0; k < count; ++k) {
    const int s = src_rank[k];
    if (s < 0 || s >= p->n) return fail(VR_ERR_ARG, "vr_peer_wait: bad rank");
    q.flag[k] = p->flags_of(p->rank) + slot * MAX_PEERS + s;
    q.seen[k] = p->flags_of(p->rank) + SEEN_WORD + slot * MAX_PEERS + s;
    q.src[k] = s;
  }
  q.slot = slot;
  q.err = reinterpret_cast<unsigned int*>(p->flags_of(p->rank) + ERR_WORD);
  q.limit = p->limit;
  peer_wait_kernel<<<1, 32, 0, p->ctx->stream>>>(q);
  CU(cudaGetLastError());
  return VR_OK;
}

int vr_peer_reduce(vr_peer* p, int32_t op, int32_t n, const void* const* in, void* out, int64_t count) {
  if (!p || !in || !out || n < 1 || n > MAX_PEERS || count < 0 || (count & 3) || op < 0 || op > 1) return fail(VR_ERR_ARG, "vr_peer_reduce: bad argument");
  CU(cudaSetDevice(p->ctx->device));
  if (count == 0) return VR_OK;
  ReduceParams q;
  std::memset(&q, 0, sizeof(q));
  for (int k = 0; k < n; ++k) q.in[k] = in[k];
  q.out = out; q.count4 = count >> 2; q.n = n; q.op = op;
  int64_t blocks = (q.count4 + 255) / 256;
  if (blocks > 1184) blocks = 1184;
  peer_reduce_kernel<<<(unsigned)blocks, 256, 0, p->ctx->stream>>>(q);
  CU(cudaGetLastError());
  return VR_OK;
}

int vr_peer_error(vr_peer* p, int32_t* timed_out) {
  if (!p || !timed_out) return fail(VR_ERR_ARG, "null");
  CU(cudaSetDevice(p->ctx->device));
  unsigned int e = 0;
  CU(cudaMemcpyAsync(&e, p->flags_of(p->rank) + ERR_WORD, sizeof(e), cudaMemcpyDeviceToHost, p->ctx->stream));
  CU(cudaStreamSynchronize(p->ctx->stream));
  *timed_out = (int32_t)e;
  return VR_OK;
}

}  // extern "C"
