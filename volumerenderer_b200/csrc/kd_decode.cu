// kd_decode.cu -- stage 1 of the hot path on the GPU: batched kd-tree stream decode into 8-bit
// bricks placed inside a device slab.  Replaces the reference's CPU
// VolumeKdtree::levelCut (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:726-835) and the
// brick placement of VolumeReader::LoadBricksToTexture (VolumeReader.h:184-205).
//
// Launch shape: one CTA per (brick, tile); a tile is the box of a depth-Dt tree node (32x32x16 for
// a 256x256x128 brick), at most 2^14 voxels.  256 parser threads, one per depth-Dl leaf subtree
// (4x4x4 voxels), walk their stream segments independently (kd_decode_core.h), writing path-ordered
// strips in shared memory; the CTA then gathers the strips into 16-byte row segments and stores
// them with 128-bit writes, so every 32-byte DRAM sector of the slab is written exactly once.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdlib>

#include "kd_decode.h"
#include "kd_decode_core.h"

namespace vr {

struct DecodeParams {
  vr_kd_geom g;
  const vr_kd_job* jobs;
  uint8_t* dst;
  int64_t rowPitch;    // slab X
  int64_t slicePitch;  // slab X*Y
  int vec16;
  const vr_kd_work* work;  // one entry per CTA (kd_worklist.h)
  const vr_fill_job* fills;  // constant groups attached to parse CTAs
  int variant;  // parser variant of the fast path: 0 = v4 with 3-word window, chain length by ALU (default), 1 = v3 (length by ALU), 2 = v2, 3 = v3, 4 = v4 (length by ALU), 5 = v4 3-word window, 6 = v4
  // occupancy (occupancy.h): when a decode tile is exactly one cell the kernel stores the tile's maximum
  // (8 writer slots per cell: one per warp, so no CTA-wide reduction is needed)
  uint8_t* occ;             // null = off
  const int32_t* tileCell;  // [2^Dt] cell of the tile relative to the cell of the brick's (0,0,0)
  // stream staging (VrWinStage): ranges[cta] = {first 16-byte chunk, chunks} of the tile's stream segment, copied to
  // shared memory at byte offset stripBytes by the whole CTA with 128-bit loads before the parsers start
  const vr_stage_range* ranges;      // null = off
  uint32_t stripBytes;
  // attached constant groups stored by the TMA unit (kd_decode.h): row images live at byte offset tmaOff of the dynamic
  // shared memory, VR_TMA_MAX_ATTACH images of (tx << 3) * VR_TMA_BOX_ROWS bytes
  int useTma;       // rows of a TMA box (0 = ordinary stores)
  int skipParse;    // timing experiments: parse entries only issue their attached groups
  uint32_t tmaOff;
  VrFillMaps maps;
  // persistent launch with SM roles (kd_decode_persistent_kernel)
  int nWork, nParse, fillEvery;
  uint32_t* queue;
  uint32_t lutOff;  // variant 13: byte offset of the two-level chain table in the dynamic shared memory
};

// occupancy cell of tile `tile` of a brick whose (0,0,0) is cell occBase: tiles are 32x32x16 = one cell
__device__ __forceinline__ int64_t tile_cell(const DecodeParams& p, const vr_kd_job& job, uint32_t tile) {
  return job.occBase + __ldg(p.tileCell + tile);
}

__device__ __forceinline__ void attached_fills(const DecodeParams& p, uint32_t first, uint32_t count) {
  for (uint32_t k = 0; k < count; ++k) {
    const vr_fill_job* fj = p.fills + first + k;
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(fj));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(fj) + 1);
    const uint64_t vals = (uint64_t)b.x | ((uint64_t)b.y << 32);
    const int64_t cell = (int64_t)((uint64_t)b.z | ((uint64_t)b.w << 32));
    if (p.occ && cell >= 0 && threadIdx.x < (1u << a.w)) {
      const uint32_t w = (uint32_t)((vals >> (8 * threadIdx.x)) & 0xFF) * 0x01010101u;
      *reinterpret_cast<uint2*>(p.occ + (cell + threadIdx.x) * 8) = make_uint2(w, w);
    }
    vr_group_fill(p.g, (int)a.w, vals, p.dst + (int64_t)(int)a.z * p.slicePitch + (int64_t)(int)a.y * p.rowPitch + (int)a.x, p.rowPitch,
                  p.slicePitch, threadIdx.x, blockDim.x);
  }
}

// TMA variant of attached_fills: the CTA writes one row image per group ((tx << L) x 8 bytes) into shared memory, warp 0
// issues the tensor stores (64 per group: 16 slices x 4 row blocks, two per lane) and everybody goes on to parse; the
// issuing lanes wait for their bulk groups right before the CTA exits.  Must be called by all threads of the CTA.
__device__ __forceinline__ void attached_fills_tma(const DecodeParams& p, uint32_t first, uint32_t count, uint32_t smemBase) {
  const uint32_t boxRows = (uint32_t)p.useTma;
  const uint32_t imgBytes = (uint32_t)(p.g.tx << 3) * boxRows;
  const uint32_t ltx = (uint32_t)(31 - __clz(p.g.tx));
  for (uint32_t k = 0; k < count; ++k) {
    const vr_fill_job* fj = p.fills + first + k;
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(fj));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(fj) + 1);
    const uint64_t vals = (uint64_t)b.x | ((uint64_t)b.y << 32);
    const int64_t cell = (int64_t)((uint64_t)b.z | ((uint64_t)b.w << 32));
    if (p.occ && cell >= 0 && threadIdx.x < (1u << a.w)) {
      const uint32_t w = (uint32_t)((vals >> (8 * threadIdx.x)) & 0xFF) * 0x01010101u;
      *reinterpret_cast<uint2*>(p.occ + (cell + threadIdx.x) * 8) = make_uint2(w, w);
    }
    const uint32_t lW = ltx + a.w;                                  // log2 row bytes of the group
    const uint32_t pieces = (boxRows << lW) >> 4;  // 16-byte pieces of the image
    const uint32_t img = smemBase + p.tmaOff + k * imgBytes;
    for (uint32_t q = threadIdx.x; q < pieces; q += blockDim.x) {
      const uint32_t xb = (q << 4) & ((1u << lW) - 1u);
      const uint32_t w = (uint32_t)((vals >> (8 * (xb >> ltx))) & 0xFF) * 0x01010101u;
      asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(img + 16u * q), "r"(w) : "memory");
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    const uint32_t rowBlocks = (uint32_t)p.g.ty / boxRows;
    const uint32_t per = (uint32_t)p.g.tz * rowBlocks;  // stores per group
    for (uint32_t k = 0; k < count; ++k) {
      const vr_fill_job* fj = p.fills + first + k;
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(fj));
      const void* tm = &p.maps.m[a.w][0];
      const uint32_t img = smemBase + p.tmaOff + k * imgBytes;
      for (uint32_t s = threadIdx.x; s < per; s += 32) {
        const int z = (int)a.z + (int)(s / rowBlocks), y = (int)a.y + (int)((s % rowBlocks) * boxRows);
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(tm), "r"((int)a.x), "r"(y), "r"(z),
                     "r"(img)
                     : "memory");
      }
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
}
__device__ __forceinline__ void attached_fills_tma_wait() {
  if (threadIdx.x < 32) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// One work-list entry, executed by a whole CTA (strips = the CTA's dynamic shared memory, sTileBase = one shared word).
__device__ __forceinline__ void decode_work_item(const DecodeParams& p, const uint32_t item, uint8_t* strips, int64_t& sTileBase) {
  const vr_kd_geom& g = p.g;
  const uint4 wk = __ldg(reinterpret_cast<const uint4*>(p.work) + item);  // {id, L, vals}
  const uint32_t brick = wk.x >> g.Dt;
  const uint32_t tile = wk.x & ((1u << g.Dt) - 1u);
  const vr_kd_job& job = p.jobs[brick];
  if (threadIdx.x == 0) {
    int ox, oy, oz;
    vr_tile_origin(g, tile, &ox, &oy, &oz);
    sTileBase = job.dstOffset + (int64_t)oz * p.slicePitch + (int64_t)oy * p.rowPitch + ox;
  }
  if (wk.y != VR_WORK_PARSE) {  // constant tiles: pure fill
    const uint64_t vals = (uint64_t)wk.z | ((uint64_t)wk.w << 32);
    const int L = (int)wk.y;
    if (p.occ && job.occBase >= 0 && threadIdx.x < (1u << L)) {  // one thread per member tile
      uint32_t t = tile;
      for (int j = 0; j < L; ++j) t |= ((threadIdx.x >> j) & 1u) << g.xrBit[j];
      const uint32_t w = (uint32_t)((vals >> (8 * threadIdx.x)) & 0xFF) * 0x01010101u;
      *reinterpret_cast<uint2*>(p.occ + tile_cell(p, job, t) * 8) = make_uint2(w, w);
    }
    __syncthreads();
    uint8_t* dst = p.dst + sTileBase;
    if (p.vec16) {
      vr_group_fill(g, L, vals, dst, p.rowPitch, p.slicePitch, threadIdx.x, blockDim.x);
    } else {
      const uint32_t items = vr_tile_items(g);
      for (uint32_t it = threadIdx.x; it < items; it += blockDim.x)
        vr_tile_fill_item(g, it, (int)(vals & 0xFF), dst, p.rowPitch, p.slicePitch, false);
    }
    return;
  }
  // constant groups that ride along with this parse CTA (wk.z = first, wk.w = count | late << 31): half of the CTAs
  // store them before parsing, the other half afterwards, so that at any time some CTAs of an SM write (HBM bound)
  // while the others parse (issue bound); with one order for all, the CTAs of a wave would fill and parse in step.
  const bool late = (wk.w >> 31) != 0;
  const bool tma = p.useTma && (wk.w & 0x7FFFFFFFu) != 0;  // CTA-uniform
  if (tma) attached_fills_tma(p, wk.z, wk.w & 0x7FFFFFFFu, (uint32_t)__cvta_generic_to_shared(strips));
  else if (!late) attached_fills(p, wk.z, wk.w & 0x7FFFFFFFu);
  if (p.skipParse) { if (tma) attached_fills_tma_wait(); else if (late) attached_fills(p, wk.z, wk.w & 0x7FFFFFFFu); return; }
  const uint32_t blk = __ldg(job.tileLeafBlock + tile);
  const uint32_t items = vr_tile_items(g);
  if (g.fastPath && p.vec16) {
    // fast path: 256 parsers, chains by table look-up, warp-local hand-off to the store phase (no CTA barrier)
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(strips);
    uint2 rg = make_uint2(0u, 0u);
    if (p.ranges) rg = __ldg(reinterpret_cast<const uint2*>(p.ranges) + item);
    if (p.variant == 13) {  // two-level chain table in shared memory, stream staged when the segment fits
      const uint32_t lutS = sbase + p.lutOff;
      for (uint32_t i = threadIdx.x; i < VR_LUT2_ENTRIES; i += blockDim.x) {
        const uint32_t v = __ldg(job.chainLut + VR_LUT2_OFF + i);
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(lutS + 4u * i), "r"(v) : "memory");
      }
      const uint32_t stage = sbase + p.stripBytes;
      if (rg.y) {
        const uint4* src = reinterpret_cast<const uint4*>(job.words) + rg.x;
        for (uint32_t j = threadIdx.x; j < rg.y; j += blockDim.x) {
          const uint4 v = __ldg(src + j);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stage + 16u * j), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
        }
      }
      __syncthreads();
      const uint32_t* lut2 = reinterpret_cast<const uint32_t*>((uintptr_t)lutS);
      if (rg.y) vr_tile_phase1_v5_l2<uint32_t, VrWinStageL>(job, g, blk, threadIdx.x, sbase, lut2, stage - 16u * rg.x);
      else vr_tile_phase1_v5_l2<uint32_t, VrWin32L>(job, g, blk, threadIdx.x, sbase, lut2, 0u);
    } else if (rg.y) {  // CTA-uniform: stage the tile's stream segment, then parse out of shared memory
      const uint4* src = reinterpret_cast<const uint4*>(job.words) + rg.x;
      const uint32_t stage = sbase + p.stripBytes;
      for (uint32_t j = threadIdx.x; j < rg.y; j += blockDim.x) {
        const uint4 v = __ldg(src + j);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stage + 16u * j), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
      }
      __syncthreads();
      vr_tile_phase1_v5_staged<true, uint32_t>(job, g, blk, threadIdx.x, sbase, stage - 16u * rg.x);
    } else if (p.variant == 0) vr_tile_phase1_v5<true, uint32_t, VrWin32>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 12) vr_tile_phase1_v5<true, uint32_t, VrWin128>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 10) vr_tile_phase1_v5<false, uint32_t, VrWin32>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 11) vr_tile_phase1_v4<true, uint32_t, VrWin96>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 5) vr_tile_phase1_v4<false, uint32_t, VrWin96>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 7) vr_tile_phase1_v4<true, uint32_t, VrWin96, 0>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 8) vr_tile_phase1_v4<true, uint32_t, VrWin96, 1>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 9) vr_tile_phase1_v4<true, uint32_t, VrWin96, 4>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 6) vr_tile_phase1_v4<false>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 1) vr_tile_phase1_v3<true>(job, g, blk, threadIdx.x, sbase);
    else if (p.variant == 2) vr_tile_phase1_v2(job, g, blk, threadIdx.x, strips);
    else if (p.variant == 3) vr_tile_phase1_v3<false>(job, g, blk, threadIdx.x, sbase);
    else vr_tile_phase1_v4<true>(job, g, blk, threadIdx.x, sbase);
    __syncwarp();
    int64_t tb = job.dstOffset;  // tile origin, recomputed per thread (no CTA barrier on this path)
    for (int d = 0; d < g.Dt; ++d)
      if ((tile >> (g.Dt - 1 - d)) & 1) {
        const int64_t st = g.splitDim[d] == 0 ? (int64_t)1 : (g.splitDim[d] == 1 ? p.rowPitch : p.slicePitch);
        tb += st * g.childOff[d];
      }
    uint8_t* dstw = p.dst + tb;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t perUnit = 1u << g.fbUnit;
    uint32_t m4 = 0;
    for (uint32_t q = lane; q < perUnit; q += 32) m4 = __vmaxu4(m4, vr_unit_phase2_item(g, warp, q, strips, dstw, p.rowPitch, p.slicePitch));
    if (p.occ && job.occBase >= 0) {  // this warp's maximum -> its slot of the tile's occupancy cell
      uint32_t m = max(max(m4 & 0xFF, (m4 >> 8) & 0xFF), max((m4 >> 16) & 0xFF, m4 >> 24));
      m = __reduce_max_sync(0xFFFFFFFFu, m);
      if (lane == 0) p.occ[tile_cell(p, job, tile) * 8 + warp] = (uint8_t)m;
    }
    if (tma) attached_fills_tma_wait();
    else if (late) attached_fills(p, wk.z, wk.w & 0x7FFFFFFFu);
    return;
  }
  const uint32_t lpt = 1u << (g.tileBits - g.leafBits);
  if (threadIdx.x < lpt) vr_tile_phase1(job, g, blk, threadIdx.x, strips);
  __syncthreads();
  uint8_t* dst = p.dst + sTileBase;
  for (uint32_t it = threadIdx.x; it < items; it += blockDim.x)
    vr_tile_phase2_item(g, it, strips, dst, p.rowPitch, p.slicePitch, p.vec16 != 0);
  if (tma) attached_fills_tma_wait();
  else if (late) attached_fills(p, wk.z, wk.w & 0x7FFFFFFFu);
}

size_t kd_decode_smem_bytes(const vr_kd_geom& g) {
  const size_t lpt = size_t(1) << (g.tileBits - g.leafBits);
  return lpt * ((size_t(1) << g.leafBits) + VR_LEAF_PITCH_PAD);
}

__global__ void __launch_bounds__(256, 6) kd_decode_tiles_kernel(const __grid_constant__ DecodeParams p) {
  extern __shared__ __align__(128) uint8_t strips[];
  __shared__ int64_t sTileBase;
  decode_work_item(p, blockIdx.x, strips, sTileBase);
}

// Persistent variant with SM roles: the work list holds the parse entries first ([0, nParse)) and the constant groups
// behind them; CTAs on every fillEvery-th SM (by %smid) drain the fill queue first, all others the parse queue first, and
// either kind helps the other once its own queue is empty.  While HBM is write-saturated a parse CTA's own 16 KB of voxel
// stores wait behind whatever its SM has queued; confining the fills to a subset of the SMs keeps the other SMs' store
// queues short (profiles/r1_decode_overlap.md).  queue[0] / queue[1] = next parse / fill entry, zeroed before the launch.
__global__ void __launch_bounds__(256, 6) kd_decode_persistent_kernel(const __grid_constant__ DecodeParams p) {
  extern __shared__ __align__(128) uint8_t strips[];
  __shared__ int64_t sTileBase;
  __shared__ uint32_t sItem;
  uint32_t smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  const uint32_t mine = (smid % (uint32_t)p.fillEvery) == 0 ? 1u : 0u;
  const uint32_t base[2] = {0u, (uint32_t)p.nParse}, count[2] = {(uint32_t)p.nParse, (uint32_t)(p.nWork - p.nParse)};
  for (;;) {
    if (threadIdx.x == 0) {
      uint32_t it = 0xFFFFFFFFu;
      for (uint32_t t = 0; t < 2 && it == 0xFFFFFFFFu; ++t) {
        const uint32_t q = mine ^ t;
        if (count[q] == 0) continue;
        const uint32_t k = atomicAdd(p.queue + q, 1u);
        if (k < count[q]) it = base[q] + k;
      }
      sItem = it;
    }
    __syncthreads();
    const uint32_t it = sItem;
    __syncthreads();
    if (it == 0xFFFFFFFFu) return;
    decode_work_item(p, it, strips, sTileBase);
  }
}

__global__ void __launch_bounds__(256) kd_prefetch_l2_kernel(const uint8_t* const* __restrict__ ptrs, const uint32_t* __restrict__ lines) {
  const uint8_t* base = ptrs[blockIdx.y];
  const uint32_t n = lines[blockIdx.y];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(base + (size_t)i * 128));
}

cudaError_t kd_prefetch_l2_launch(const uint8_t* const* ptrsDev, const uint32_t* linesDev, int n, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  kd_prefetch_l2_kernel<<<dim3(32, (unsigned)n), 256, 0, stream>>>(ptrsDev, linesDev);
  return cudaGetLastError();
}

typedef CUresult (*VrEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

bool kd_fill_maps_create(uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, int tx, int ty, int rows, VrFillMaps* out) {
  static_assert(sizeof(CUtensorMap) == 128, "CUtensorMap size");
  static VrEncodeFn enc = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      f = nullptr;
    }
    return (VrEncodeFn)f;
  }();
  if (!enc || tx < 16 || (tx << 3) > 256 || rows <= 0 || ty % rows) return false;
  if (SX % 16 || ((uintptr_t)vol) % 16 || SX >= (int64_t(1) << 32) || SY >= (int64_t(1) << 32) || SZ >= (int64_t(1) << 32)) return false;
  for (int L = 0; L <= 3; ++L) {
    cuuint64_t dims[3] = {(cuuint64_t)SX, (cuuint64_t)SY, (cuuint64_t)SZ};
    cuuint64_t strides[2] = {(cuuint64_t)SX, (cuuint64_t)SX * (cuuint64_t)SY};
    cuuint32_t box[3] = {(cuuint32_t)(tx << L), (cuuint32_t)rows, 1};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(reinterpret_cast<CUtensorMap*>(&out->m[L][0]), CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, vol, dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
  }
  return true;
}

cudaError_t kd_decode_launch(const vr_kd_geom& g, const vr_kd_job* jobsDev, const vr_kd_work* workDev, int nWork, uint8_t* dst,
                             int64_t rowPitch, int64_t slicePitch, bool vec16, cudaStream_t stream, uint8_t* occ,
                             const int32_t* tileCell, const vr_fill_job* fillsDev, const vr_stage_range* rangesDev, int stageBytes,
                             const VrFillMaps* tmaMaps, int nParse, int fillEvery, uint32_t* queueDev, int tmaRows) {
  if (nWork <= 0) return cudaSuccess;
  DecodeParams p;
  p.g = g;
  p.jobs = jobsDev;
  p.work = workDev;
  p.fills = fillsDev;
  p.dst = dst;
  p.rowPitch = rowPitch;
  p.slicePitch = slicePitch;
  p.vec16 = vec16 ? 1 : 0;
  p.occ = tileCell ? occ : nullptr; p.tileCell = tileCell;
  {
    static int variant = -1;
    if (variant < 0) { const char* e = getenv("VR_DECODE_VARIANT"); variant = e ? atoi(e) : 0; }
    p.variant = variant;
  }
  p.ranges = stageBytes > 0 ? rangesDev : nullptr;
  p.stripBytes = (uint32_t)kd_decode_smem_bytes(g);
  size_t smem = kd_decode_smem_bytes(g) + (p.ranges ? (size_t)stageBytes : 0);
  p.useTma = 0; p.tmaOff = 0;
  {
    static int skip = -1;
    if (skip < 0) { const char* e = getenv("VR_DECODE_ONLY"); skip = (e && atoi(e) == 3) ? 1 : 0; }
    p.skipParse = skip;
  }
  if (tmaMaps && tmaRows > 0 && fillsDev && vec16 && g.tx >= 16 && (g.tx << 3) <= 256 && g.ty % tmaRows == 0) {
    p.useTma = tmaRows;
    p.maps = *tmaMaps;
    p.tmaOff = (uint32_t)((smem + 127) & ~size_t(127));
    smem = p.tmaOff + (size_t)VR_TMA_MAX_ATTACH * (size_t)(g.tx << 3) * (size_t)tmaRows;
  }
  p.lutOff = 0;
  if (p.variant == 13) {
    p.lutOff = (uint32_t)((smem + 15) & ~size_t(15));
    smem = p.lutOff + 4u * VR_LUT2_ENTRIES + 64;
  }
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kd_decode_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  p.nWork = nWork; p.nParse = nParse; p.fillEvery = fillEvery; p.queue = queueDev;
  if (fillEvery > 0 && queueDev && nParse >= 0 && nParse <= nWork) {
    static int numSMs = 0;
    if (!numSMs) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&numSMs, cudaDevAttrMultiProcessorCount, dev); }
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(kd_decode_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    cudaError_t e = cudaMemsetAsync(queueDev, 0, 2 * sizeof(uint32_t), stream);
    if (e != cudaSuccess) return e;
    const int ctas = numSMs * 6 < nWork ? numSMs * 6 : nWork;
    kd_decode_persistent_kernel<<<(unsigned)ctas, 256, smem, stream>>>(p);
    return cudaGetLastError();
  }
  kd_decode_tiles_kernel<<<(unsigned)nWork, 256, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace vr
