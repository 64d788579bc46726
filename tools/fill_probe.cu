// fill_probe.cu -- what is the write-only HBM ceiling of this GPU, and does the store path matter?
//   (a) st.global.v4 grid-stride fill               (b) cudaMemsetAsync
//   (c) cp.async.bulk smem -> global, 16 KB chunks   (d) cp.async.bulk.tensor 3-D box stores (32x32x16 u8 boxes
//       of a 2048 x 256 x 1920 slab, the decode kernel's constant-tile shape)
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/fill_probe tools/fill_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("ERR %s line %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__global__ void fill_v4(uint4* dst, size_t n16, uint32_t w) {
  const uint4 v = make_uint4(w, w, w, w);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = v;
}

// each CTA: fill a 16 KB smem buffer once, then bulk-store it to successive 16 KB chunks
__global__ void fill_bulk(uint8_t* dst, size_t nChunks, uint32_t w, int chunk) {
  extern __shared__ __align__(128) uint8_t buf[];
  for (int i = threadIdx.x; i < chunk / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(buf)[i] = w;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(buf);
    for (size_t c = blockIdx.x; c < nChunks; c += gridDim.x) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + c * (size_t)chunk), "r"(s), "r"(chunk) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 8;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

// 3-D tensor-map box stores: box bx x by x bz from smem
__global__ void fill_tensor(const __grid_constant__ CUtensorMap tm, int nx, int ny, int nz, int bx, int by, int bz, uint32_t w) {
  extern __shared__ __align__(128) uint8_t buf[];
  const int bytes = bx * by * bz;
  for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(buf)[i] = w;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(buf);
    const size_t total = (size_t)nx * ny * nz;
    for (size_t t = blockIdx.x; t < total; t += gridDim.x) {
      const int ix = (int)(t % nx), iy = (int)((t / nx) % ny), iz = (int)(t / ((size_t)nx * ny));
      asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(&tm), "r"(ix * bx), "r"(iy * by),
                   "r"(iz * bz), "r"(s)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 8;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}


// (e) decode-shaped fills: one CTA of 256 threads per box of bx x 32 x 16 voxels of a 2048 x 256 x 1920 slab (bx = 32: one
// tile, 256: the 8 x-neighbouring tiles of a brick), 16-byte stores, CTA order = boxes in x, then y, then z
__global__ void __launch_bounds__(256) fill_boxes(uint8_t* dst, int bx, int lbx, uint32_t w, int work_every, const uint32_t* flag) {
  extern __shared__ uint8_t dummy[];
  if (flag && flag[blockIdx.x & 1023] == 12345u) return;  // one dependent load before any work
  if (work_every > 1 && (blockIdx.x % work_every) != 0) return;
  const uint32_t b = work_every > 1 ? blockIdx.x / work_every : blockIdx.x;
  const int nbx = 2048 >> lbx;
  const int ix = b % nbx, iy = (b / nbx) % 8, iz = b / (nbx * 8);
  uint8_t* base = dst + ((size_t)iz * 16 * 256 + (size_t)iy * 32) * 2048 + (size_t)ix * bx;
  const int lR = lbx - 4;
  const uint32_t piece = threadIdx.x & ((1u << lR) - 1u), rstep = 256u >> lR;
  const uint4 v = make_uint4(w, w, w, w);
  for (uint32_t row = threadIdx.x >> lR; row < 512; row += rstep) {
    const uint32_t y = row & 31, z = row >> 5;
    *reinterpret_cast<uint4*>(base + ((size_t)z * 256 + y) * 2048 + piece * 16) = v;
  }
}
__global__ void __launch_bounds__(256) empty_ctas(const uint32_t* flag, int load) {
  extern __shared__ uint8_t dummy[];
  if (load && flag[blockIdx.x & 1023] == 12345u) dummy[0] = 1;
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <typename F>
static float time_ms(F f, int reps) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) f();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  for (int i = 0; i < reps; ++i) f();
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  return ms / reps;
}

int main() {
  const int X = 2048, Y = 256, Z = 1920;
  const size_t bytes = (size_t)X * Y * Z;  // 1.007 GB > L2
  uint8_t* d;
  CK(cudaMalloc(&d, bytes));
  int sms;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const int reps = 20;
  for (int mult : {2, 4, 8, 16}) {
    float ms = time_ms([&] { fill_v4<<<sms * mult, 512>>>((uint4*)d, bytes / 16, 0x01010101u); }, reps);
    printf("st.v4 grid %2dx%d x512: %.4f ms  %.0f GB/s\n", mult, sms, ms, bytes / ms * 1e-6);
  }
  {
    float ms = time_ms([&] { CK(cudaMemsetAsync(d, 3, bytes)); }, reps);
    printf("cudaMemset: %.4f ms  %.0f GB/s\n", ms, bytes / ms * 1e-6);
  }
  for (int chunk : {4096, 16384, 65536}) {
    CK(cudaFuncSetAttribute(fill_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    for (int mult : {1, 2, 4}) {
      if ((size_t)chunk * mult > 200 * 1024) continue;
      float ms = time_ms([&] { fill_bulk<<<sms * mult, 128, chunk>>>(d, bytes / chunk, 0x02020202u, chunk); }, reps);
      printf("bulk chunk %6d grid %dx%d: %.4f ms  %.0f GB/s\n", chunk, mult, sms, ms, bytes / ms * 1e-6);
    }
  }

  {
    uint32_t* flag;
    CK(cudaMalloc(&flag, 4096));
    CK(cudaMemset(flag, 0, 4096));
    const int smem = 17408;
    for (int lbx : {5, 6, 7, 8}) {
      const int bx = 1 << lbx;
      const unsigned nb = (2048 / bx) * 8 * 120;
      for (int sm : {0, smem}) {
        float ms = time_ms([&] { fill_boxes<<<nb, 256, sm>>>(d, bx, lbx, 0x05050505u, 1, nullptr); }, reps);
        printf("boxes %3dx32x16 (%u CTAs, smem %5d): %.4f ms  %.0f GB/s\n", bx, nb, sm, ms, bytes / ms * 1e-6);
      }
    }
    {  // 61440 CTAs of which every 8th fills a 256-wide box (the grouped-fill launch shape)
      float ms = time_ms([&] { fill_boxes<<<61440, 256, smem>>>(d, 256, 8, 0x06060606u, 8, nullptr); }, reps);
      printf("boxes 256x32x16, 1 of 8 CTAs works: %.4f ms  %.0f GB/s\n", ms, bytes / ms * 1e-6);
      ms = time_ms([&] { fill_boxes<<<61440, 256, smem>>>(d, 256, 8, 0x06060606u, 8, flag); }, reps);
      printf("  same + one dependent load per CTA: %.4f ms  %.0f GB/s\n", ms, bytes / ms * 1e-6);
      ms = time_ms([&] { fill_boxes<<<61440, 256, smem>>>(d, 32, 5, 0x06060606u, 1, flag); }, reps);
      printf("boxes 32x32x16 + one dependent load per CTA: %.4f ms  %.0f GB/s\n", ms, bytes / ms * 1e-6);
    }
    for (int sm : {0, smem})
      for (int load : {0, 1}) {
        float ms = time_ms([&] { empty_ctas<<<61440, 256, sm>>>(flag, load); }, reps);
        printf("61440 empty CTAs (smem %5d, load %d): %.4f ms = %.1f ns per CTA per SM\n", sm, load, ms, ms * 1e6 / (61440.0 / sms));
      }
  }
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &qr));
  if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 0; }
  const int boxes[][3] = {{32, 32, 16}, {64, 32, 16}, {256, 32, 1}, {128, 16, 8}};
  for (auto& bx : boxes) {
    CUtensorMap tm;
    cuuint64_t dims[3] = {(cuuint64_t)X, (cuuint64_t)Y, (cuuint64_t)Z};
    cuuint64_t strides[2] = {(cuuint64_t)X, (cuuint64_t)X * Y};
    cuuint32_t box[3] = {(cuuint32_t)bx[0], (cuuint32_t)bx[1], (cuuint32_t)bx[2]};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); continue; }
    const int smem = bx[0] * bx[1] * bx[2];
    CK(cudaFuncSetAttribute(fill_tensor, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    for (int mult : {1, 2, 4}) {
      float ms = time_ms([&] { fill_tensor<<<sms * mult, 128, smem>>>(tm, X / bx[0], Y / bx[1], Z / bx[2], bx[0], bx[1], bx[2], 0x04040404u); }, reps);
      CK(cudaGetLastError());
      printf("tensor box %3dx%2dx%2d grid %dx%d: %.4f ms  %.0f GB/s\n", bx[0], bx[1], bx[2], mult, sms, ms, bytes / ms * 1e-6);
    }
  }
  CK(cudaDeviceSynchronize());
  // spot check of the last pattern
  uint32_t h[4];
  CK(cudaMemcpy(h, d + bytes - 16, 16, cudaMemcpyDeviceToHost));
  printf("tail word %08x\n", h[3]);
  return 0;
}
