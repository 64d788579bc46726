// occupancy.cu -- builds the per-cell maxima used for empty-space skipping (occupancy.h).
#include "occupancy.h"

namespace vr {
namespace {

// One CTA per cell that intersects the box: raw[cell] = max over (cell ∩ box).
__global__ void __launch_bounds__(256) occ_reduce_box_kernel(const uint8_t* __restrict__ vol, long long SX, long long SY,
                                                             int lox, int loy, int loz, int hix, int hiy, int hiz,
                                                             int c0x, int c0y, int c0z, uint8_t* __restrict__ raw, int nx, int ny,
                                                             int shx, int shy, int shz) {
  const int cx = c0x + blockIdx.x, cy = c0y + blockIdx.y, cz = c0z + blockIdx.z;
  // slab-local voxel range of the cell, clipped to the box
  const int x0 = max(lox, (cx << VR_OCC_LX) - shx), x1 = min(hix, ((cx + 1) << VR_OCC_LX) - shx);
  const int y0 = max(loy, (cy << VR_OCC_LY) - shy), y1 = min(hiy, ((cy + 1) << VR_OCC_LY) - shy);
  const int z0 = max(loz, (cz << VR_OCC_LZ) - shz), z1 = min(hiz, ((cz + 1) << VR_OCC_LZ) - shz);
  const int ex = x1 - x0, ey = y1 - y0, ez = z1 - z0;
  unsigned m = 0;
  if (ex > 0 && ey > 0 && ez > 0) {
    const bool v4 = ((ex & 3) == 0) && (((SX | (long long)x0) & 3) == 0) && ((reinterpret_cast<uintptr_t>(vol) & 3) == 0);
    if (v4) {
      const int wx = ex >> 2, items = wx * ey * ez;
      for (int it = threadIdx.x; it < items; it += 256) {
        const int xi = it % wx, y = (it / wx) % ey, z = it / (wx * ey);
        const unsigned w = __ldg(reinterpret_cast<const unsigned*>(vol + ((long long)(z0 + z) * SY + (y0 + y)) * SX + x0) + xi);
        m = __vmaxu4(m, w);
      }
      m = max(max(m & 0xFF, (m >> 8) & 0xFF), max((m >> 16) & 0xFF, m >> 24));
    } else {
      const int items = ex * ey * ez;
      for (int it = threadIdx.x; it < items; it += 256) {
        const int xi = it % ex, y = (it / ex) % ey, z = it / (ex * ey);
        m = max(m, (unsigned)__ldg(vol + ((long long)(z0 + z) * SY + (y0 + y)) * SX + x0 + xi));
      }
    }
  }
  __shared__ unsigned sm[8];
  m = __reduce_max_sync(0xFFFFFFFFu, m);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) m = max(m, sm[w]);
    raw[((long long)cz * ny + cy) * nx + cx] = (uint8_t)m;
  }
}

__global__ void occ_dilate_kernel(const uint8_t* __restrict__ raw, uint8_t* __restrict__ dil, int nx, int ny, int nz) {
  const long long n = (long long)nx * ny * nz;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int cx = (int)(i % nx), cy = (int)((i / nx) % ny), cz = (int)(i / ((long long)nx * ny));
  unsigned m = 0;
  for (int dz = 0; dz < 2; ++dz)
    for (int dy = 0; dy < 2; ++dy)
      for (int dx = 0; dx < 2; ++dx) {
        const int x = min(cx + dx, nx - 1), y = min(cy + dy, ny - 1), z = min(cz + dz, nz - 1);
        m = max(m, (unsigned)__ldg(raw + ((long long)z * ny + y) * nx + x));
      }
  dil[i] = (uint8_t)m;
}

// one warp per macro cell
__global__ void occ_macro_kernel(const uint8_t* __restrict__ dil, uint8_t* __restrict__ macro, int nx, int ny, int nz, int mx, int my,
                                 int mz) {
  const int warp = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (warp >= mx * my * mz) return;
  const int bx = warp % mx, by = (warp / mx) % my, bz = warp / (mx * my);
  const int W = 1 << VR_OCC_MACRO;
  unsigned m = 0;
  for (int it = lane; it < W * W * W; it += 32) {
    const int x = (bx << VR_OCC_MACRO) + (it & (W - 1)), y = (by << VR_OCC_MACRO) + ((it >> VR_OCC_MACRO) & (W - 1)),
              z = (bz << VR_OCC_MACRO) + (it >> (2 * VR_OCC_MACRO));
    if (x < nx && y < ny && z < nz) m = max(m, (unsigned)__ldg(dil + ((long long)z * ny + y) * nx + x));
  }
  m = __reduce_max_sync(0xFFFFFFFFu, m);
  if (lane == 0) macro[warp] = (uint8_t)m;
}

}  // namespace

cudaError_t occ_reduce_box_launch(const uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, const int64_t lo[3], const int64_t ext[3],
                                  uint8_t* raw, int nx, int ny, int nz, const int shift[3], cudaStream_t stream) {
  (void)SZ; (void)nz;
  const int hi[3] = {(int)(lo[0] + ext[0]), (int)(lo[1] + ext[1]), (int)(lo[2] + ext[2])};
  const int lg[3] = {VR_OCC_LX, VR_OCC_LY, VR_OCC_LZ};
  int c0[3], c1[3];
  for (int a = 0; a < 3; ++a) { c0[a] = ((int)lo[a] + shift[a]) >> lg[a]; c1[a] = (hi[a] - 1 + shift[a]) >> lg[a]; }
  dim3 grid((unsigned)(c1[0] - c0[0] + 1), (unsigned)(c1[1] - c0[1] + 1), (unsigned)(c1[2] - c0[2] + 1));
  occ_reduce_box_kernel<<<grid, 256, 0, stream>>>(vol, SX, SY, (int)lo[0], (int)lo[1], (int)lo[2], hi[0], hi[1], hi[2], c0[0], c0[1],
                                                  c0[2], raw, nx, ny, shift[0], shift[1], shift[2]);
  return cudaGetLastError();
}

cudaError_t occ_dilate_launch(const uint8_t* raw, uint8_t* dil, uint8_t* macro, int nx, int ny, int nz, int mx, int my, int mz,
                              cudaStream_t stream) {
  const long long n = (long long)nx * ny * nz;
  occ_dilate_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(raw, dil, nx, ny, nz);
  const long long threads = (long long)mx * my * mz * 32;
  occ_macro_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(dil, macro, nx, ny, nz, mx, my, mz);
  return cudaGetLastError();
}

}  // namespace vr
