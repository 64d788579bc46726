// kd_build.cu -- the stream producer on the GPU: VolumeKdtree::build
// (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:17-140: buildRecursive :143-201,
// compressGradientDescent :206-384, pruneTreeRecursive :596-629, convertToPreorder :631-724) as level-parallel
// kernels that emit byte-identical streams.  kd_compress.cpp is the same algorithm on the CPU and
// documents the reference's quirks; kd_build_core.h explains how the two sequential steps (running-mean estimate,
// order-dependent double accumulation) are made parallel without changing a bit.
//
// Trees are stored breadth-first: level d occupies [2^d - 1, 2^(d+1) - 1), children of i are 2i+1 and 2i+2.
// Host control (the gradient-descent loop of a level) reads one or two scalars per pass.
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>
#include <cub/device/device_scan.cuh>
#include <vector>

#include "kd_build.h"
#include "kd_build_core.h"

namespace vr {

#define KB_CHECK(x)                              \
  do {                                           \
    cudaError_t e_ = (x);                        \
    if (e_ != cudaSuccess) { err = e_; return VR_ERR_CUDA; } \
  } while (0)

namespace {

constexpr int TB = 256;  // threads per block of the level kernels (= VR_ACC_BLOCK)
static_assert(TB == VR_ACC_BLOCK, "block sums are per thread block");

__device__ __forceinline__ int parent_of(const uint8_t* parents, int depth, int64_t i) { return depth == 0 ? 0 : (int)parents[i >> 1]; }

// ---- pyramid ---------------------------------------------------------------------------------------
__global__ void leaves_kernel(const __grid_constant__ vr_kd_geom g, const uint8_t* vol, int64_t X, int64_t XY, uint8_t* temp, uint8_t* mn,
                              uint8_t* mx, int64_t numLeaves) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= numLeaves) return;
  const uint8_t s = vol[vr_leaf_voxel(g, (uint32_t)p, X, XY)];
  const int64_t idx = numLeaves - 1 + p;
  temp[idx] = s; mn[idx] = s; mx[idx] = s;
}
__global__ void pyramid_kernel(uint8_t* temp, uint8_t* mn, uint8_t* mx, int64_t start, int64_t n) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = start + k;
  const uint8_t a = min(mn[2 * i + 1], mn[2 * i + 2]), b = max(mx[2 * i + 1], mx[2 * i + 2]);
  mn[i] = a; mx[i] = b;
  temp[i] = (uint8_t)(((int)a + (int)b) / 2);  // (byte)((max+min)/2.0)  (:198)
}

// ---- estimate (running mean) -----------------------------------------------------------------------
// item = |parent - truth| << 32 | 1 when the node counts, else 0; the exclusive prefix sum of the items is the
// (sum, count) state in front of every node.
__global__ void est_init_kernel(const uint8_t* parents, const uint8_t* truth, int depth, int64_t n, uint64_t* items, uint8_t* dec) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int pd = abs(parent_of(parents, depth, i) - (int)truth[i]);
  dec[i] = pd > 0;
  items[i] = pd > 0 ? (((uint64_t)pd << 32) | 1u) : 0;
}
__global__ void est_iter_kernel(const uint8_t* parents, const uint8_t* truth, int depth, int64_t n, const uint64_t* prefix, uint64_t* items,
                                uint8_t* dec, unsigned int* changed) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int parent = parent_of(parents, depth, i), t = truth[i];
  const uint64_t st = prefix[i];
  const bool d = vr_estimate_counts(parent, t, (double)(st >> 32), (double)(uint32_t)st);
  if (d != (dec[i] != 0)) {
    dec[i] = d;
    items[i] = d ? (((uint64_t)abs(parent - t) << 32) | 1u) : 0;
    atomicAdd(changed, 1u);
    atomicMin(changed + 1, (unsigned int)i);  // everything in front of the first change is final
  }
}
// A pass that changes only a handful of decisions is usually a chain reaction creeping forward a few nodes per
// pass (runs of nodes that sit exactly on the threshold).  Everything in front of the first change is final and
// prefix[first] is exact, so one thread replays the next `len` nodes sequentially and leaves them final.
__global__ void est_window_kernel(const uint8_t* parents, const uint8_t* truth, int depth, int64_t n, const uint64_t* prefix, uint64_t* items,
                                  uint8_t* dec, int64_t first, int64_t len) {
  const uint64_t st = prefix[first];
  double sum = (double)(st >> 32), count = (double)(uint32_t)st;
  const int64_t end = first + len < n ? first + len : n;
  for (int64_t i = first; i < end; ++i) {
    const int parent = parent_of(parents, depth, i), t = truth[i];
    const bool d = vr_estimate_counts(parent, t, sum, count);
    dec[i] = d;
    items[i] = d ? (((uint64_t)abs(parent - t) << 32) | 1u) : 0;
    if (d) { sum += (double)abs(parent - t); count += 1.0; }
  }
}
__global__ void est_seq_kernel(const uint8_t* parents, const uint8_t* truth, int depth, int64_t n, double* out) {
  double sum = 0.0, count = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    const int parent = parent_of(parents, depth, i), t = truth[i];
    if (vr_estimate_counts(parent, t, sum, count)) { sum += (double)abs(parent - t); count += 1.0; }
  }
  out[0] = sum; out[1] = count;
}
__global__ void est_total_kernel(const uint64_t* prefix, const uint64_t* items, int64_t n, double* out) {
  const uint64_t st = prefix[n - 1] + items[n - 1];
  out[0] = (double)(st >> 32); out[1] = (double)(uint32_t)st;
}

// ---- encode passes ---------------------------------------------------------------------------------
template <bool WRITE>
__global__ void encode_kernel(const uint8_t* parents, const uint8_t* truth, int depth, int64_t n, int dist, uint8_t* code, uint8_t* recon,
                              uint32_t* blockSum, uint64_t* superSum, unsigned long long* total) {
  __shared__ uint32_t warpSum[TB / 32];
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  uint32_t e2 = 0;
  if (i < n) {
    const vr_enc e = vr_encode_node(parent_of(parents, depth, i), dist, truth[i]);
    if (WRITE) { code[i] = (uint8_t)e.code; recon[i] = (uint8_t)e.value; }
    e2 = (uint32_t)(e.err * e.err);
  }
  e2 = __reduce_add_sync(0xFFFFFFFFu, e2);
  if ((threadIdx.x & 31) == 0) warpSum[threadIdx.x >> 5] = e2;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = 0;
    for (int w = 0; w < TB / 32; ++w) s += warpSum[w];
    if (WRITE) {
      blockSum[blockIdx.x] = s;
      if (s) atomicAdd((unsigned long long*)&superSum[blockIdx.x / VR_ACC_SUPER], (unsigned long long)s);
    } else if (s) {
      atomicAdd(total, (unsigned long long)s);
    }
  }
}
// the order-exact accumulation of the level's squared errors (kd_build_core.h), one thread
__global__ void accumulate_kernel(const uint8_t* parents, const uint8_t* truth, int depth, int64_t n, int dist, const uint32_t* blockSum,
                                  const uint64_t* superSum, double x, double* out) {
  out[0] = vr_seq_accumulate(x, n, blockSum, superSum, [=](int64_t i) {
    const vr_enc e = vr_encode_node(parent_of(parents, depth, i), dist, truth[i]);
    return (uint32_t)(e.err * e.err);
  });
}

// ---- leaf statistics ------------------------------------------------------------------------------
__global__ void leaf_err_kernel(const uint8_t* truth, const uint8_t* recon, int64_t n, unsigned int* maxErr, unsigned long long* s1,
                                unsigned long long* s2) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  uint32_t e = 0;
  if (i < n) e = (uint32_t)abs((int)truth[i] - (int)recon[i]);
  const uint32_t m = __reduce_max_sync(0xFFFFFFFFu, e);
  const uint32_t a = __reduce_add_sync(0xFFFFFFFFu, e), b = __reduce_add_sync(0xFFFFFFFFu, e * e);
  if ((threadIdx.x & 31) == 0) {
    if (m) atomicMax(maxErr, m);
    if (a) { atomicAdd(s1, (unsigned long long)a); atomicAdd(s2, (unsigned long long)b); }
  }
}

// ---- prune ------------------------------------------------------------------------------------------
__global__ void prune_leaves_kernel(uint8_t* code, const uint8_t* recon, const uint8_t* truth, int64_t n, int tolerance) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (code[i] == 0 && abs((int)recon[i] - (int)truth[i]) < tolerance) code[i] = 3;
}
__global__ void prune_level_kernel(uint8_t* codes, int64_t start, int64_t n) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = start + k;
  if (codes[i] == 0 && codes[2 * i + 1] == 3 && codes[2 * i + 2] == 3) codes[i] = 3;
}

// ---- preorder emission ------------------------------------------------------------------------------
struct Dist7 { uint8_t d[VR_ADD_LEVELS]; };
// chain[i] = symbols (2 bits each) | count << 16; size = symbols this leaf emits
__global__ void chain_kernel(const uint8_t* code, uint8_t* recon, const uint8_t* truth, int64_t n, int tolerance, Dist7 dist, uint32_t* chain,
                             uint32_t* size, unsigned long long* rewrites) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t packed = 0, cnt = 0;
  if (code[i] != 3) {
    uint8_t sym[8];
    int r, rew;
    cnt = (uint32_t)vr_leaf_chain(recon[i], truth[i], tolerance, dist.d, sym, &r, &rew);
    for (uint32_t k = 0; k < cnt; ++k) packed |= (uint32_t)sym[k] << (2 * k);
    recon[i] = (uint8_t)r;
    if (rew) atomicAdd(rewrites, 1ull);
  }
  chain[i] = packed | (cnt << 16);
  size[i] = 1 + cnt;
}
__global__ void size_level_kernel(const uint8_t* codes, uint32_t* size, int64_t start, int64_t n) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = start + k;
  size[i] = codes[i] == 3 ? 1u : 1u + size[2 * i + 1] + size[2 * i + 2];
}
#define KB_NOT_EMITTED 0xFFFFFFFFu
// offsets of the children of level [start, start + n); symbols of the level itself
__global__ void offset_level_kernel(const uint8_t* codes, const uint32_t* size, uint32_t* off, uint8_t* sym, int64_t start, int64_t n, bool leaves,
                                    const uint32_t* chain) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = start + k;
  const uint32_t o = off[i];
  if (o == KB_NOT_EMITTED) {
    if (!leaves) { off[2 * i + 1] = KB_NOT_EMITTED; off[2 * i + 2] = KB_NOT_EMITTED; }
    return;
  }
  const uint8_t c = codes[i];
  sym[o] = c;
  if (leaves) {
    const uint32_t ch = chain[k], cnt = ch >> 16;
    for (uint32_t q = 0; q < cnt; ++q) sym[o + 1 + q] = (uint8_t)((ch >> (2 * q)) & 3u);
    return;
  }
  if (c == 3) { off[2 * i + 1] = KB_NOT_EMITTED; off[2 * i + 2] = KB_NOT_EMITTED; return; }
  off[2 * i + 1] = o + 1;
  off[2 * i + 2] = o + 1 + size[2 * i + 1];
}
__global__ void pack_kernel(const uint8_t* sym, int64_t nSym, uint8_t* bits, int64_t nBytes) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= nBytes) return;
  uint32_t b = 0;
  for (int q = 0; q < 4; ++q)
    if (4 * j + q < nSym) b |= (uint32_t)sym[4 * j + q] << (2 * q);
  bits[j] = (uint8_t)b;
}

// ---- decode index from the producer's own offsets ("the index for free") ---------------------
// The offsets of offset_level_kernel are exactly what the host walk of kd_stream.cpp (kd_prepare) recovers from the finished
// stream; the decoded value of every node follows from the codes top-down.  val[i] = value levelCut carries at node i
// (VolumeKdTree_recover.cpp:743,784-787); nodes under a pruned ancestor inherit the ancestor's value, which is what the
// decode writes into their boxes (:759-766).
__global__ void val_level_kernel(const uint8_t* codes, const uint32_t* off, uint8_t* val, int64_t start, int64_t n, int dist) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = start + k;
  const int parent = val[(i - 1) >> 1];
  val[i] = (uint8_t)(off[i] == KB_NOT_EMITTED ? parent : vr_apply(parent, codes[i], dist));
}
// final voxel values: the leaf's value with its refinement chain applied (:815-818 for depths > origTreeDepth)
__global__ void final_kernel(const uint8_t* leafCode, const uint32_t* leafOff, const uint8_t* leafVal, const uint32_t* chain, Dist7 dist, uint8_t* fin,
                             int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int v = leafVal[i];
  if (leafOff[i] != KB_NOT_EMITTED && leafCode[i] != 3) {
    const uint32_t ch = chain[i], cnt = ch >> 16;
    for (uint32_t q = 0; q < cnt; ++q) {
      const int c = (int)((ch >> (2 * q)) & 3u);
      if (c == 3) break;
      v = vr_apply(v, c, dist.d[q]);
    }
  }
  fin[i] = (uint8_t)v;
}
// one thread: tiles in path order, live tiles numbered as they come (kd_stream.cpp IndexVisitor::node)
__global__ void index_tiles_kernel(const uint8_t* codes, const uint32_t* off, const uint8_t* val, int64_t firstTile, int64_t nTiles,
                                   uint32_t* tileLeafBlock, uint32_t* tileBegin, uint8_t* tileVal, uint32_t* numLive) {
  uint32_t live = 0;
  for (int64_t t = 0; t < nTiles; ++t) {
    const int64_t i = firstTile + t;
    const bool isLive = off[i] != KB_NOT_EMITTED && codes[i] != 3;
    tileLeafBlock[t] = isLive ? live++ : VR_PRUNED;
    tileBegin[t] = isLive ? off[i] : 0u;
    tileVal[t] = isLive ? (uint8_t)0 : val[i];
  }
  *numLive = live;
}
__global__ void index_leaves_kernel(const uint8_t* codes, const uint32_t* off, const uint8_t* val, int64_t firstLeafNode, int64_t nLeafNodes,
                                    int lptLog2, const uint32_t* tileLeafBlock, uint32_t* leafOff, uint8_t* leafVal) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= nLeafNodes) return;
  const uint32_t blk = tileLeafBlock[k >> lptLog2];
  if (blk == VR_PRUNED) return;
  const int64_t j = firstLeafNode + k;
  const int64_t o = ((int64_t)blk << lptLog2) + (k & ((int64_t(1) << lptLog2) - 1));
  const bool em = off[j] != KB_NOT_EMITTED && codes[j] != 3;
  leafOff[o] = em ? off[j] : VR_PRUNED;
  leafVal[o] = val[j];
}
// exact bounds of the decoded voxels of every tile: one CTA per tile
__global__ void tile_bounds_kernel(const uint8_t* fin, int tileBits, uint8_t* tileMin, uint8_t* tileMax) {
  __shared__ uint32_t smn[TB / 32], smx[TB / 32];
  const int64_t base = (int64_t)blockIdx.x << tileBits, cnt = int64_t(1) << tileBits;
  uint32_t mn = 255, mx = 0;
  for (int64_t i = threadIdx.x; i < cnt; i += blockDim.x) { const uint32_t v = fin[base + i]; mn = min(mn, v); mx = max(mx, v); }
  mn = __reduce_min_sync(0xFFFFFFFFu, mn);
  mx = __reduce_max_sync(0xFFFFFFFFu, mx);
  if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < TB / 32; ++w) { mn = min(mn, smn[w]); mx = max(mx, smx[w]); }
    tileMin[blockIdx.x] = (uint8_t)mn; tileMax[blockIdx.x] = (uint8_t)mx;
  }
}

inline unsigned grid_for(int64_t n) { return (unsigned)((n + TB - 1) / TB); }

}  // namespace

struct KdBuilder {
  int64_t capLeaves = 0;   // leaves the buffers are sized for
  uint8_t *temp = nullptr, *mn = nullptr, *mx = nullptr, *codes = nullptr;
  uint8_t* lvl[3] = {nullptr, nullptr, nullptr};  // parents / recon / reconPrev (rotating)
  uint64_t *items = nullptr, *prefix = nullptr;   // estimate scan; later: size / off (uint32 views)
  uint8_t* dec = nullptr;
  uint32_t* blockSum = nullptr;
  uint64_t* superSum = nullptr;
  uint32_t* chain = nullptr;
  uint8_t *sym = nullptr, *bits = nullptr;
  int64_t capSym = 0;
  void* cubTemp = nullptr;
  size_t cubBytes = 0;
  uint8_t* idxTiles = nullptr;    // per tile: tileLeafBlock u32 | tileBegin u32 | tileVal | tileMin | tileMax
  int64_t capTiles = 0;
  double* scalarsDev = nullptr;   // 8 doubles / counters
  double* scalarsHost = nullptr;  // pinned mirror
  void release() {
    for (void* p : {(void*)temp, (void*)mn, (void*)mx, (void*)codes, (void*)lvl[0], (void*)lvl[1], (void*)lvl[2], (void*)items, (void*)prefix,
                    (void*)dec, (void*)blockSum, (void*)superSum, (void*)chain, (void*)sym, (void*)bits, cubTemp, (void*)scalarsDev,
                    (void*)idxTiles})
      if (p) cudaFree(p);
    if (scalarsHost) cudaFreeHost(scalarsHost);
    *this = KdBuilder();
  }
};

KdBuilder* kd_builder_create() { return new KdBuilder; }
void kd_builder_destroy(KdBuilder* b) {
  if (!b) return;
  b->release();
  delete b;
}

static cudaError_t builder_reserve(KdBuilder* b, int64_t numLeaves) {
  if (numLeaves <= b->capLeaves) return cudaSuccess;
  b->release();
  const size_t N = (size_t)numLeaves, T = 2 * N;  // T >= 2N - 1 tree nodes
  cudaError_t e;
#define KB_ALLOC(ptr, bytes) if ((e = cudaMalloc((void**)&(ptr), (bytes))) != cudaSuccess) { b->release(); return e; }
  KB_ALLOC(b->temp, T) KB_ALLOC(b->mn, T) KB_ALLOC(b->mx, T) KB_ALLOC(b->codes, T)
  for (auto& p : b->lvl) KB_ALLOC(p, N)
  KB_ALLOC(b->items, 8 * N) KB_ALLOC(b->prefix, 8 * N)  // as uint32: size[2N] and off[2N]
  KB_ALLOC(b->dec, N)
  KB_ALLOC(b->blockSum, 4 * (N / VR_ACC_BLOCK + 1))
  KB_ALLOC(b->superSum, 8 * (N / VR_ACC_BLOCK / VR_ACC_SUPER + 1))
  KB_ALLOC(b->chain, 4 * N)
  KB_ALLOC(b->scalarsDev, 8 * sizeof(double))
  if ((e = cudaHostAlloc((void**)&b->scalarsHost, 8 * sizeof(double), cudaHostAllocDefault)) != cudaSuccess) { b->release(); return e; }
  b->cubBytes = 0;
  if ((e = cub::DeviceScan::ExclusiveSum(nullptr, b->cubBytes, b->items, b->prefix, (int)N)) != cudaSuccess) { b->release(); return e; }
  KB_ALLOC(b->cubTemp, b->cubBytes + 16)
#undef KB_ALLOC
  b->capLeaves = numLeaves;
  return cudaSuccess;
}

namespace {
// reads k doubles (or 64-bit counters) of the device scalar block
cudaError_t fetch(KdBuilder* b, int k, cudaStream_t s) {
  cudaError_t e = cudaMemcpyAsync(b->scalarsHost, b->scalarsDev, (size_t)k * sizeof(double), cudaMemcpyDeviceToHost, s);
  return e != cudaSuccess ? e : cudaStreamSynchronize(s);
}
}  // namespace

int kd_build_device(KdBuilder* b, const uint8_t* volDev, int64_t X, int64_t Y, int64_t Z, int64_t rowPitch, int64_t slicePitch,
                    int tolerance, int maxEpochs, cudaStream_t stream, KdStream* out, KdBuildStats* stats, cudaError_t* errOut,
                    int64_t* iterStats, bool withIndex) {
  cudaError_t err = cudaSuccess;
  struct ErrGuard { cudaError_t& e; cudaError_t* o; ~ErrGuard() { if (o) *o = e; } } guard{err, errOut};
  if (!b || !volDev || !out) return VR_ERR_ARG;
  vr_kd_geom g;
  if (vr_kd_make_geom(X, Y, Z, &g) != 0) return VR_ERR_FORMAT;
  const int D0 = g.D0, Dmax = g.Dmax;
  const int64_t numLeaves = int64_t(1) << D0, firstLeaf = numLeaves - 1;
  KB_CHECK(builder_reserve(b, numLeaves));
  int64_t itTotal = 0, itWorst = 0;

  KB_CHECK(cudaMemsetAsync(b->codes, 0, (size_t)(2 * numLeaves - 1), stream));  // symbols start as 0 (maxEpochs < 1 writes none)
  // ---- BUILD: midrange pyramid ----
  leaves_kernel<<<grid_for(numLeaves), TB, 0, stream>>>(g, volDev, rowPitch, slicePitch, b->temp, b->mn, b->mx, numLeaves);
  for (int d = D0 - 1; d >= 0; --d) {
    const int64_t n = int64_t(1) << d;
    pyramid_kernel<<<grid_for(n), TB, 0, stream>>>(b->temp, b->mn, b->mx, n - 1, n);
  }
  KB_CHECK(cudaGetLastError());

  // ---- COMPRESS: per-level distance by gradient descent (same control flow as kd_compress.cpp) ----
  std::vector<uint8_t> distanceMap((size_t)Dmax + 1, 0);
  uint8_t *parents = b->lvl[0], *recon = b->lvl[1], *reconPrev = b->lvl[2];
  const double gamma = 1.25, hstep = 1.0, maxAbsStep = 4.0;
  double currentDistance = 0.0, currentError = 0.0, currentDF = 0.0, currentStepSize = 0.0;
  double previousStepSize = 0.0, previousDistance = 0.0, previousDF = 0.0, previousError = 0.0;
  unsigned int* changedDev = reinterpret_cast<unsigned int*>(b->scalarsDev + 2);
  unsigned long long* totalDev = reinterpret_cast<unsigned long long*>(b->scalarsDev + 3);

  for (int depth = 0; depth <= D0; ++depth) {
    const int64_t n = int64_t(1) << depth;
    const uint8_t* truth = b->temp + (n - 1);
    uint8_t* code = b->codes + (n - 1);
    int epoch = 0;
    double distanceSum = 0.0, distanceCount = 0.0;
    if (maxEpochs < 1) KB_CHECK(cudaMemsetAsync(recon, 0, (size_t)n, stream));  // the reference's resize() zero-fills
    // running-mean estimate
    if (n <= 2048) {
      est_seq_kernel<<<1, 1, 0, stream>>>(parents, truth, depth, n, b->scalarsDev);
      KB_CHECK(fetch(b, 2, stream));
    } else {
      est_init_kernel<<<grid_for(n), TB, 0, stream>>>(parents, truth, depth, n, b->items, b->dec);
      int64_t it = 0;
      for (;; ++it) {
        size_t tb = b->cubBytes;
        KB_CHECK(cub::DeviceScan::ExclusiveSum(b->cubTemp, tb, b->items, b->prefix, (int)n, stream));
        KB_CHECK(cudaMemsetAsync(changedDev, 0, sizeof(unsigned int), stream));
        KB_CHECK(cudaMemsetAsync(changedDev + 1, 0xFF, sizeof(unsigned int), stream));
        est_iter_kernel<<<grid_for(n), TB, 0, stream>>>(parents, truth, depth, n, b->prefix, b->items, b->dec, changedDev);
        KB_CHECK(fetch(b, 3, stream));
        unsigned int changed[2];
        std::memcpy(changed, b->scalarsHost + 2, sizeof(changed));
        if (changed[0] == 0) break;
        if (it >= 400) break;  // not converging: replay the level with one thread below
        if (changed[0] < 256)   // a creeping chain reaction: settle the next stretch sequentially
          est_window_kernel<<<1, 1, 0, stream>>>(parents, truth, depth, n, b->prefix, b->items, b->dec, (int64_t)changed[1], 4096);
      }
      itTotal += it + 1;
      if (it + 1 > itWorst) itWorst = it + 1;
      if (it >= 400) {
        est_seq_kernel<<<1, 1, 0, stream>>>(parents, truth, depth, n, b->scalarsDev);
      } else {
        est_total_kernel<<<1, 1, 0, stream>>>(b->prefix, b->items, n, b->scalarsDev);
      }
      KB_CHECK(fetch(b, 2, stream));
    }
    distanceSum = b->scalarsHost[0];
    distanceCount = b->scalarsHost[1];
    currentDistance = distanceCount > 0 ? std::round(distanceSum / distanceCount) : 0.0;

    previousDistance = 0.0;
    previousStepSize = 255.0;
    previousError = 65025.0;
    const unsigned nb = grid_for(n);
    while (epoch < maxEpochs && std::fabs(previousStepSize) >= 0.5) {
      if (epoch != 0) {
        previousDistance = currentDistance;
        previousError = currentError;
        previousDF = currentDF;
        previousStepSize = currentStepSize;
        currentDistance = std::round(std::fmin(255.0, std::fmax(0.0, previousDistance + previousStepSize)));
        if (currentDistance == previousDistance) break;
      }
      {  // error of the current distance; writes symbols and the reconstruction
        const int dist = (uint8_t)currentDistance;
        KB_CHECK(cudaMemsetAsync(b->superSum, 0, 8 * ((size_t)nb / VR_ACC_SUPER + 1), stream));
        encode_kernel<true><<<nb, TB, 0, stream>>>(parents, truth, depth, n, dist, code, recon, b->blockSum, b->superSum, nullptr);
        accumulate_kernel<<<1, 1, 0, stream>>>(parents, truth, depth, n, dist, b->blockSum, b->superSum, currentError, b->scalarsDev);
        KB_CHECK(fetch(b, 1, stream));
        currentError = b->scalarsHost[0] / (double)n;
      }
      if (currentError < 1.0) break;
      if (epoch != 0 && currentError > previousError) {
        currentError = previousError;
        currentDistance = previousDistance;
        currentDF = previousDF;
        currentStepSize = previousStepSize / 2.0;
        std::swap(recon, reconPrev);
        epoch++;
        continue;
      }
      // central difference of the mean squared error: integer sums, exact in any order
      const int est[2] = {(uint8_t)std::fmax(0.0, currentDistance - hstep), (uint8_t)std::fmin(255.0, currentDistance + hstep)};
      KB_CHECK(cudaMemsetAsync(totalDev, 0, 2 * sizeof(unsigned long long), stream));
      for (int k = 0; k < 2; ++k)
        encode_kernel<false><<<nb, TB, 0, stream>>>(parents, truth, depth, n, est[k], nullptr, nullptr, nullptr, nullptr, totalDev + k);
      KB_CHECK(fetch(b, 5, stream));
      unsigned long long acc[2];
      std::memcpy(acc, b->scalarsHost + 3, sizeof(acc));
      const double estErr[2] = {(double)acc[0] / (double)n, (double)acc[1] / (double)n};
      currentDF = (estErr[1] - estErr[0]) / (2.0 * hstep);
      currentStepSize = std::fmax(-maxAbsStep, std::fmin(maxAbsStep, -gamma * currentDF));
      KB_CHECK(cudaMemcpyAsync(reconPrev, recon, (size_t)n, cudaMemcpyDeviceToDevice, stream));
      epoch++;
    }
    distanceMap[(size_t)depth] = (uint8_t)currentDistance;
    if (depth < D0) std::swap(parents, recon);
  }

  // ---- leaf statistics before branch growth ----
  const uint8_t* leafTruth = b->temp + firstLeaf;
  uint8_t* leafCode = b->codes + firstLeaf;
  unsigned int* maxErrDev = reinterpret_cast<unsigned int*>(b->scalarsDev + 5);
  unsigned long long* sumsDev = reinterpret_cast<unsigned long long*>(b->scalarsDev + 6);  // s1, s2
  if (stats) {
    KB_CHECK(cudaMemsetAsync(b->scalarsDev + 5, 0, 3 * sizeof(double), stream));
    leaf_err_kernel<<<grid_for(numLeaves), TB, 0, stream>>>(leafTruth, recon, numLeaves, maxErrDev, sumsDev, sumsDev + 1);
    KB_CHECK(fetch(b, 8, stream));
    unsigned int m;
    std::memcpy(&m, b->scalarsHost + 5, sizeof(m));
    stats->maxErrorBeforeGrowth = (int)m;
  }

  // ---- PRUNE (:596-629) ----
  prune_leaves_kernel<<<grid_for(numLeaves), TB, 0, stream>>>(leafCode, recon, leafTruth, numLeaves, tolerance);
  for (int d = D0 - 1; d >= 0; --d) {
    const int64_t n = int64_t(1) << d;
    prune_level_kernel<<<grid_for(n), TB, 0, stream>>>(b->codes, n - 1, n);
  }

  // ---- fixed refinement distances (:92-97) ----
  Dist7 dist7;
  {
    const int add[VR_ADD_LEVELS] = {64, 32, 16, 8, 4, 2, 1};
    for (int k = 0; k < VR_ADD_LEVELS; ++k) { distanceMap[(size_t)(D0 + 1 + k)] = (uint8_t)add[k]; dist7.d[k] = (uint8_t)add[k]; }
  }

  // ---- CONVERT (:631-724): chains per leaf, subtree sizes bottom-up, offsets top-down, scatter, pack ----
  uint32_t* size = reinterpret_cast<uint32_t*>(b->items);   // [2N - 1]
  uint32_t* off = reinterpret_cast<uint32_t*>(b->prefix);   // [2N - 1]
  unsigned long long* rewritesDev = reinterpret_cast<unsigned long long*>(b->scalarsDev + 4);
  KB_CHECK(cudaMemsetAsync(rewritesDev, 0, sizeof(unsigned long long), stream));
  chain_kernel<<<grid_for(numLeaves), TB, 0, stream>>>(leafCode, recon, leafTruth, numLeaves, tolerance, dist7, b->chain, size + firstLeaf,
                                                       rewritesDev);
  for (int d = D0 - 1; d >= 0; --d) {
    const int64_t n = int64_t(1) << d;
    size_level_kernel<<<grid_for(n), TB, 0, stream>>>(b->codes, size, n - 1, n);
  }
  KB_CHECK(cudaMemcpyAsync(b->scalarsHost, size, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  KB_CHECK(cudaStreamSynchronize(stream));
  uint32_t numActive32;
  std::memcpy(&numActive32, b->scalarsHost, sizeof(numActive32));
  const int64_t numActive = (int64_t)numActive32, nBytes = (numActive + 3) / 4;
  if (numActive > b->capSym) {
    if (b->sym) cudaFree(b->sym);
    if (b->bits) cudaFree(b->bits);
    b->sym = b->bits = nullptr;
    b->capSym = 0;
    KB_CHECK(cudaMalloc((void**)&b->sym, (size_t)numActive + 16));
    KB_CHECK(cudaMalloc((void**)&b->bits, (size_t)nBytes + 16));
    b->capSym = numActive;
  }
  KB_CHECK(cudaMemsetAsync(off, 0, sizeof(uint32_t), stream));  // the root is emitted at symbol 0
  for (int d = 0; d <= D0; ++d) {
    const int64_t n = int64_t(1) << d;
    offset_level_kernel<<<grid_for(n), TB, 0, stream>>>(b->codes, size, off, b->sym, n - 1, n, d == D0, b->chain);
  }
  pack_kernel<<<grid_for(nBytes), TB, 0, stream>>>(b->sym, numActive, b->bits, nBytes);
  KB_CHECK(cudaGetLastError());

  // ---- decode index (kd_stream.h KdIndex) from the offsets above: no walk of the finished stream ----
  KdIndex ix;
  const bool emitIndex = withIndex && numActive < (int64_t)VR_PRUNED;
  if (emitIndex) {
    const int64_t nTiles = int64_t(1) << g.Dt, nLeafNodes = int64_t(1) << g.Dl;
    const int lptLog2 = g.tileBits - g.leafBits;
    if (nTiles > b->capTiles) {
      if (b->idxTiles) cudaFree(b->idxTiles);
      b->idxTiles = nullptr; b->capTiles = 0;
      KB_CHECK(cudaMalloc((void**)&b->idxTiles, (size_t)nTiles * 11 + 64));
      b->capTiles = nTiles;
    }
    uint32_t* tileLeafBlockDev = reinterpret_cast<uint32_t*>(b->idxTiles);
    uint32_t* tileBeginDev = tileLeafBlockDev + nTiles;
    uint8_t *tileValDev = b->idxTiles + 8 * nTiles, *tileMinDev = tileValDev + nTiles, *tileMaxDev = tileMinDev + nTiles;
    uint8_t *val = b->mn, *fin = b->mx;  // the pyramid's min / max arrays are free by now
    uint32_t* leafOffDev = b->chain;     // read by final_kernel first, then overwritten (stream order)
    uint8_t* leafValDev = b->dec;
    uint32_t* numLiveDev = reinterpret_cast<uint32_t*>(b->scalarsDev);
    KB_CHECK(cudaMemcpyAsync(val, distanceMap.data(), 1, cudaMemcpyHostToDevice, stream));  // root = distanceMap[0] (:743)
    for (int d = 1; d <= D0; ++d) {
      const int64_t n = int64_t(1) << d;
      val_level_kernel<<<grid_for(n), TB, 0, stream>>>(b->codes, off, val, n - 1, n, (int)distanceMap[(size_t)d]);
    }
    final_kernel<<<grid_for(numLeaves), TB, 0, stream>>>(leafCode, off + firstLeaf, val + firstLeaf, b->chain, dist7, fin, numLeaves);
    index_tiles_kernel<<<1, 1, 0, stream>>>(b->codes, off, val, nTiles - 1, nTiles, tileLeafBlockDev, tileBeginDev, tileValDev, numLiveDev);
    tile_bounds_kernel<<<(unsigned)nTiles, TB, 0, stream>>>(fin, g.tileBits, tileMinDev, tileMaxDev);
    index_leaves_kernel<<<grid_for(nLeafNodes), TB, 0, stream>>>(b->codes, off, val, nLeafNodes - 1, nLeafNodes, lptLog2, tileLeafBlockDev, leafOffDev,
                                                                 leafValDev);
    KB_CHECK(cudaGetLastError());
    std::vector<uint32_t> tileBegin((size_t)nTiles);
    ix.tileLeafBlock.resize((size_t)nTiles); ix.tileVal.resize((size_t)nTiles); ix.tileMin.resize((size_t)nTiles); ix.tileMax.resize((size_t)nTiles);
    KB_CHECK(cudaMemcpyAsync(ix.tileLeafBlock.data(), tileLeafBlockDev, (size_t)nTiles * 4, cudaMemcpyDeviceToHost, stream));
    KB_CHECK(cudaMemcpyAsync(tileBegin.data(), tileBeginDev, (size_t)nTiles * 4, cudaMemcpyDeviceToHost, stream));
    KB_CHECK(cudaMemcpyAsync(ix.tileVal.data(), tileValDev, (size_t)nTiles, cudaMemcpyDeviceToHost, stream));
    KB_CHECK(cudaMemcpyAsync(ix.tileMin.data(), tileMinDev, (size_t)nTiles, cudaMemcpyDeviceToHost, stream));
    KB_CHECK(cudaMemcpyAsync(ix.tileMax.data(), tileMaxDev, (size_t)nTiles, cudaMemcpyDeviceToHost, stream));
    KB_CHECK(fetch(b, 1, stream));
    uint32_t numLive;
    std::memcpy(&numLive, b->scalarsHost, sizeof(numLive));
    const size_t nLeafEntries = (size_t)numLive << lptLog2;
    ix.leafOff.resize(nLeafEntries); ix.leafVal.resize(nLeafEntries);
    if (nLeafEntries) {
      KB_CHECK(cudaMemcpyAsync(ix.leafOff.data(), leafOffDev, nLeafEntries * 4, cudaMemcpyDeviceToHost, stream));
      KB_CHECK(cudaMemcpyAsync(ix.leafVal.data(), leafValDev, nLeafEntries, cudaMemcpyDeviceToHost, stream));
      KB_CHECK(cudaStreamSynchronize(stream));
    }
    kd_index_finish(g, tileBegin, numActive, &ix);
  }

  if (stats) {
    KB_CHECK(cudaMemsetAsync(b->scalarsDev + 5, 0, 3 * sizeof(double), stream));
    leaf_err_kernel<<<grid_for(numLeaves), TB, 0, stream>>>(leafTruth, recon, numLeaves, maxErrDev, sumsDev, sumsDev + 1);
    KB_CHECK(fetch(b, 8, stream));
    unsigned int m;
    unsigned long long s[2], rw;
    std::memcpy(&m, b->scalarsHost + 5, sizeof(m));
    std::memcpy(s, b->scalarsHost + 6, sizeof(s));
    std::memcpy(&rw, b->scalarsHost + 4, sizeof(rw));
    stats->maxError = (int)m;
    stats->meanL1 = (double)s[0] / (double)numLeaves;  // integer sums: exact in any order
    stats->meanL2 = (double)s[1] / (double)numLeaves;
    stats->zeroRunRewrites = (int64_t)rw;
  }

  // ---- the stream object ----
  KdStream s;
  std::memset(&s.hdr, 0, sizeof(s.hdr));
  s.hdr.rootMax[0] = X; s.hdr.rootMax[1] = Y; s.hdr.rootMax[2] = Z;
  s.hdr.maxTreeDepth = Dmax;
  s.hdr.origTreeDepth = D0;
  s.hdr.X = X; s.hdr.Y = Y; s.hdr.Z = Z;
  s.hdr.numActiveNodes = numActive;
  s.geom = g;
  s.distanceMap = distanceMap;
  s.bits.resize((size_t)nBytes);
  KB_CHECK(cudaMemcpyAsync(s.bits.data(), b->bits, (size_t)nBytes, cudaMemcpyDeviceToHost, stream));
  KB_CHECK(cudaStreamSynchronize(stream));
  if (iterStats) { iterStats[0] = itTotal; iterStats[1] = itWorst; }
  if (emitIndex) s.index = std::move(ix);
  *out = std::move(s);
  return VR_OK;
}

}  // namespace vr
