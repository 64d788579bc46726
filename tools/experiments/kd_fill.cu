// kd_fill.cu -- see kd_fill.h.  Stage 1 of the hot path, the part of the reference's levelCut that writes
// pruned boxes (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:759-766), for boxes of tile size and up.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdlib>

#include "kd_fill.h"

namespace vr {

struct FillParams {
  FillMaps maps;
  const vr_fill_job* jobs;
  uint8_t* occ;
  int n;
  int ltx;  // log2 tile extent in x
  int ty, tz;
};

__global__ void __launch_bounds__(32) kd_fill_groups_kernel(const __grid_constant__ FillParams p) {
  extern __shared__ __align__(128) uint8_t img[];  // (tx << L) x ty bytes: one z slice of the group
  const uint32_t lane = threadIdx.x;
  const uint32_t simg = (uint32_t)__cvta_generic_to_shared(img);
  uint64_t have = 0;
  uint32_t haveL = 0xFFFFFFFFu;
  for (int i = blockIdx.x; i < p.n; i += gridDim.x) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(p.jobs + i));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(p.jobs + i) + 1);
    const uint64_t vals = (uint64_t)b.x | ((uint64_t)b.y << 32);
    const int64_t cell = (int64_t)((uint64_t)b.z | ((uint64_t)b.w << 32));
    const uint32_t L = a.w;
    if (L != haveL || vals != have) {  // new row image (most groups repeat the previous one: all 0 or all 230)
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      __syncwarp();
      const uint32_t lW = (uint32_t)p.ltx + L;
      const uint32_t pieces = ((uint32_t)p.ty << lW) >> 4;
      for (uint32_t q = lane; q < pieces; q += 32) {
        const uint32_t xb = (q << 4) & ((1u << lW) - 1u);
        const uint32_t w = (uint32_t)((vals >> (8 * (xb >> p.ltx))) & 0xFF) * 0x01010101u;
        *reinterpret_cast<uint4*>(img + (size_t)q * 16) = make_uint4(w, w, w, w);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      have = vals;
      haveL = L;
    }
    if (p.occ && cell >= 0 && lane < (1u << L)) {
      const uint32_t w = (uint32_t)((vals >> (8 * lane)) & 0xFF) * 0x01010101u;
      *reinterpret_cast<uint2*>(p.occ + (cell + lane) * 8) = make_uint2(w, w);
    }
    if (lane == 0) {
      const void* tm = &p.maps.m[L][0];
      for (int z = 0; z < p.tz; ++z)
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(tm), "r"(a.x), "r"(a.y),
                     "r"((int)a.z + z), "r"(simg)
                     : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// Plain-store variant: one CTA of 64 threads per group, no shared memory, so a CTA fits into whatever an SM has
// left next to six parse CTAs.
struct FillPlainParams {
  vr_kd_geom g;
  const vr_fill_job* jobs;
  uint8_t* dst;
  uint8_t* occ;
  int64_t rowPitch, slicePitch;
};

__global__ void __launch_bounds__(64) kd_fill_groups_plain_kernel(const __grid_constant__ FillPlainParams p) {
  const uint4 a = __ldg(reinterpret_cast<const uint4*>(p.jobs + blockIdx.x));
  const uint4 b = __ldg(reinterpret_cast<const uint4*>(p.jobs + blockIdx.x) + 1);
  const uint64_t vals = (uint64_t)b.x | ((uint64_t)b.y << 32);
  const int64_t cell = (int64_t)((uint64_t)b.z | ((uint64_t)b.w << 32));
  const uint32_t L = a.w;
  if (p.occ && cell >= 0 && threadIdx.x < (1u << L)) {
    const uint32_t w = (uint32_t)((vals >> (8 * threadIdx.x)) & 0xFF) * 0x01010101u;
    *reinterpret_cast<uint2*>(p.occ + (cell + threadIdx.x) * 8) = make_uint2(w, w);
  }
  uint8_t* dst = p.dst + (int64_t)(int)a.z * p.slicePitch + (int64_t)(int)a.y * p.rowPitch + (int)a.x;
  vr_group_fill(p.g, (int)L, vals, dst, p.rowPitch, p.slicePitch, threadIdx.x, 64);
}

cudaError_t kd_fill_plain_launch(const vr_kd_geom& g, const vr_fill_job* jobsDev, int n, uint8_t* dst, int64_t rowPitch, int64_t slicePitch,
                                 uint8_t* occ, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  FillPlainParams p;
  p.g = g;
  p.jobs = jobsDev;
  p.dst = dst;
  p.occ = occ;
  p.rowPitch = rowPitch;
  p.slicePitch = slicePitch;
  kd_fill_groups_plain_kernel<<<(unsigned)n, 64, 0, stream>>>(p);
  return cudaGetLastError();
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn encode_fn() {
  static EncodeFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      f = nullptr;
    }
    return (EncodeFn)f;
  }();
  return fn;
}

size_t kd_fill_smem_bytes(int tx, int ty, int maxL) { return (size_t)(tx << maxL) * (size_t)ty; }

bool kd_fill_maps_create(uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, int tx, int ty, int maxL, FillMaps* out) {
  static_assert(sizeof(CUtensorMap) == 128, "CUtensorMap size");
  EncodeFn enc = encode_fn();
  if (!enc || maxL < 0 || maxL > 3) return false;
  if ((tx << maxL) > 256 || ty > 256 || tx < 16 || kd_fill_smem_bytes(tx, ty, maxL) > 32768) return false;
  if (SX % 16 || ((uintptr_t)vol) % 16 || SX >= (int64_t(1) << 32) || SY >= (int64_t(1) << 32) || SZ >= (int64_t(1) << 32)) return false;
  for (int L = 0; L <= maxL; ++L) {
    cuuint64_t dims[3] = {(cuuint64_t)SX, (cuuint64_t)SY, (cuuint64_t)SZ};
    cuuint64_t strides[2] = {(cuuint64_t)SX, (cuuint64_t)SX * (cuuint64_t)SY};
    cuuint32_t box[3] = {(cuuint32_t)(tx << L), (cuuint32_t)ty, 1};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(reinterpret_cast<CUtensorMap*>(&out->m[L][0]), CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, vol, dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
  }
  return true;
}

cudaError_t kd_fill_launch(const FillMaps& maps, const vr_fill_job* jobsDev, int n, int tx, int ty, int tz, int maxL, uint8_t* occ, int ctas,
                           cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  FillParams p;
  p.maps = maps;
  p.jobs = jobsDev;
  p.occ = occ;
  p.n = n;
  p.ltx = 0;
  while ((1 << p.ltx) < tx) ++p.ltx;
  p.ty = ty;
  p.tz = tz;
  const size_t smem = kd_fill_smem_bytes(tx, ty, maxL);
  if (ctas > n) ctas = n;
  if (const char* e = getenv("VR_CARVEOUT")) cudaFuncSetAttribute(kd_fill_groups_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(e));
  kd_fill_groups_kernel<<<(unsigned)ctas, 32, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace vr
