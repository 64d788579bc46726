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
    const unsigned w = m * 0x01010101u;  // all writer slots
    *reinterpret_cast<uint2*>(raw + (((long long)cz * ny + cy) * nx + cx) * VR_OCC_SLOTS) = make_uint2(w, w);
  }
}

__device__ __forceinline__ unsigned raw_max(const uint8_t* __restrict__ raw, long long cell) {
  const uint2 w = __ldg(reinterpret_cast<const uint2*>(raw + cell * VR_OCC_SLOTS));
  const unsigned m = __vmaxu4(w.x, w.y);
  return max(max(m & 0xFF, (m >> 8) & 0xFF), max((m >> 16) & 0xFF, m >> 24));
}

// One CTA per level-3 cell = 8x8x8 level-0 cells: dilate raw into level 0, then reduce to levels 1..3.
__global__ void __launch_bounds__(512) occ_pyr_low_kernel(const uint8_t* __restrict__ raw, uint8_t* __restrict__ pyr,
                                                          const __grid_constant__ OccView L) {
  __shared__ uint8_t s0[512], s1[64], s2[8];
  const int t = threadIdx.x;
  const int cx = (blockIdx.x << 3) + (t & 7), cy = (blockIdx.y << 3) + ((t >> 3) & 7), cz = (blockIdx.z << 3) + (t >> 6);
  unsigned m = 0;
  if (cx < L.nx && cy < L.ny && cz < L.nz) {
    for (int dz = 0; dz < 2; ++dz)
      for (int dy = 0; dy < 2; ++dy)
        for (int dx = 0; dx < 2; ++dx) {
          const int x = min(cx + dx, L.nx - 1), y = min(cy + dy, L.ny - 1), z = min(cz + dz, L.nz - 1);
          m = max(m, raw_max(raw, ((long long)z * L.ny + y) * L.nx + x));
        }
    pyr[L.off[0] + ((long long)cz * L.ny + cy) * L.nx + cx] = (uint8_t)m;
  }
  s0[t] = (uint8_t)m;
  __syncthreads();
  for (int l = 1; l <= 3; ++l) {
    if (l >= L.nlev) break;
    const int w = 8 >> l, cells = w * w * w;  // cells of level l inside this CTA's region
    uint8_t* dst = l == 1 ? s1 : s2;
    const uint8_t* src = l == 1 ? s0 : (l == 2 ? s1 : s2);
    unsigned mm = 0;
    int x = 0, y = 0, z = 0;
    if (t < cells) {
      x = t % w; y = (t / w) % w; z = t / (w * w);
      const int ws = w * 2;
      for (int c = 0; c < 8; ++c) mm = max(mm, (unsigned)src[((2 * z + (c >> 2)) * ws + (2 * y + ((c >> 1) & 1))) * ws + 2 * x + (c & 1)]);
    }
    __syncthreads();  // src fully read before a level that aliases it is written (s2 serves levels 2 and 3)
    if (t < cells) {
      if (l < 3) dst[t] = (uint8_t)mm;
      const int nxl = ((L.nx - 1) >> l) + 1, nyl = ((L.ny - 1) >> l) + 1, nzl = ((L.nz - 1) >> l) + 1;
      const int gx = (blockIdx.x << (3 - l)) + x, gy = (blockIdx.y << (3 - l)) + y, gz = (blockIdx.z << (3 - l)) + z;
      if (gx < nxl && gy < nyl && gz < nzl) pyr[L.off[l] + ((long long)gz * nyl + gy) * nxl + gx] = (uint8_t)mm;
    }
    __syncthreads();
  }
}

// One CTA: levels 4.. from level 3, one level after the other.
__global__ void __launch_bounds__(256) occ_pyr_high_kernel(uint8_t* __restrict__ pyr, const __grid_constant__ OccView L) {
  for (int l = 4; l < L.nlev; ++l) {
    const int nxl = ((L.nx - 1) >> l) + 1, nyl = ((L.ny - 1) >> l) + 1, nzl = ((L.nz - 1) >> l) + 1;
    const int nxp = ((L.nx - 1) >> (l - 1)) + 1, nyp = ((L.ny - 1) >> (l - 1)) + 1, nzp = ((L.nz - 1) >> (l - 1)) + 1;
    for (int i = threadIdx.x; i < nxl * nyl * nzl; i += blockDim.x) {
      const int x = i % nxl, y = (i / nxl) % nyl, z = i / (nxl * nyl);
      unsigned m = 0;
      for (int c = 0; c < 8; ++c) {
        const int px = 2 * x + (c & 1), py = 2 * y + ((c >> 1) & 1), pz = 2 * z + (c >> 2);
        if (px < nxp && py < nyp && pz < nzp) m = max(m, (unsigned)pyr[L.off[l - 1] + ((long long)pz * nyp + py) * nxp + px]);
      }
      pyr[L.off[l] + i] = (uint8_t)m;
    }
    __syncthreads();
  }
}

// lvl[c] for the march (occupancy.h): VR_OCC_SOLID, or R = the largest radius such that every cell within Chebyshev distance R
// of c (in cells; cells outside the grid count as void -- the march clips its boxes to the stored voxels) is void for L.thr.
// Chebyshev distance to the nearest non-void cell, separably: along x, then min over y' of max(|dy|, .), then the same along z.
__global__ void occ_dist_x_kernel(const uint8_t* __restrict__ pyr0, uint8_t* __restrict__ gx, const __grid_constant__ OccView L) {
  const long long n = (long long)L.nx * L.ny * L.nz;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int cx = (int)(i % L.nx);
  const uint8_t* row = pyr0 + (i - cx);
  int g = VR_OCC_DIST_CAP;
  for (int d = 0; d < VR_OCC_DIST_CAP; ++d)
    if ((cx - d >= 0 && (int)row[cx - d] > L.thr) || (cx + d < L.nx && (int)row[cx + d] > L.thr)) { g = d; break; }
  gx[i] = (uint8_t)g;
}
// The hot path's first two passes in one launch, one CTA per z-slice of cells (nx * ny cells in shared memory): dilation of the
// raw maxima into level 0 of pyr (what occ_pyr_low_kernel writes there), distance along x, then along y.
__global__ void __launch_bounds__(256) occ_dilate_dist_xy_kernel(const uint8_t* __restrict__ raw, uint8_t* __restrict__ pyr0, uint8_t* __restrict__ gy,
                                                                 const __grid_constant__ OccView L) {
  extern __shared__ uint8_t sm_occ[];
  const int cells = L.nx * L.ny, cz = blockIdx.x;
  uint8_t *solid = sm_occ, *gx = sm_occ + cells;
  for (int i = threadIdx.x; i < cells; i += blockDim.x) {
    const int cx = i % L.nx, cy = i / L.nx;
    unsigned m = 0;
    for (int dz = 0; dz < 2; ++dz)
      for (int dy = 0; dy < 2; ++dy)
        for (int dx = 0; dx < 2; ++dx) {
          const int x = min(cx + dx, L.nx - 1), y = min(cy + dy, L.ny - 1), z = min(cz + dz, L.nz - 1);
          m = max(m, raw_max(raw, ((long long)z * L.ny + y) * L.nx + x));
        }
    pyr0[(long long)cz * cells + i] = (uint8_t)m;
    solid[i] = (int)m > L.thr;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cells; i += blockDim.x) {
    const int cx = i % L.nx;
    const uint8_t* row = solid + (i - cx);
    int g = VR_OCC_DIST_CAP;
    for (int d = 0; d < VR_OCC_DIST_CAP; ++d)
      if ((cx - d >= 0 && row[cx - d]) || (cx + d < L.nx && row[cx + d])) { g = d; break; }
    gx[i] = (uint8_t)g;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cells; i += blockDim.x) {
    const int cy = i / L.nx;
    int best = gx[i];
    for (int d = 1; d < best; ++d) {
      if (cy - d >= 0) best = min(best, max(d, (int)gx[i - d * L.nx]));
      if (cy + d < L.ny) best = min(best, max(d, (int)gx[i + d * L.nx]));
    }
    gy[(long long)cz * cells + i] = (uint8_t)best;
  }
}

// stride = distance between neighbours along the axis, c / nc = coordinate and extent along it; out = min over d of max(d, in[c +- d])
__global__ void occ_dist_axis_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, const __grid_constant__ OccView L, int axis, int finish) {
  const long long n = (long long)L.nx * L.ny * L.nz;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = axis == 1 ? (int)((i / L.nx) % L.ny) : (int)(i / ((long long)L.nx * L.ny));
  const int nc = axis == 1 ? L.ny : L.nz;
  const long long stride = axis == 1 ? L.nx : (long long)L.nx * L.ny;
  int best = in[i];
  for (int d = 1; d < best; ++d) {
    if (c - d >= 0) best = min(best, max(d, (int)in[i - d * stride]));
    if (c + d < nc) best = min(best, max(d, (int)in[i + d * stride]));
  }
  out[i] = finish ? (uint8_t)(best == 0 ? VR_OCC_SOLID : best - 1) : (uint8_t)best;
}

}  // namespace

namespace {
// pairs[2i] = cell, pairs[2i+1] = exact maximum of the cell's voxels: slot 0 takes the value, the other writer slots 0
__global__ void __launch_bounds__(256) occ_set_cells_kernel(uint8_t* __restrict__ raw, const long long* __restrict__ pairs, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const longlong2 e = __ldg(reinterpret_cast<const longlong2*>(pairs) + i);
  *reinterpret_cast<uint2*>(raw + e.x * VR_OCC_SLOTS) = make_uint2((unsigned)e.y & 0xFFu, 0u);
}
}  // namespace

namespace {
__global__ void __launch_bounds__(256) occ_merge_solid_kernel(uint8_t* __restrict__ lvl, const uint8_t* __restrict__ solid, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && __ldg(solid + i) && lvl[i] == VR_OCC_SOLID) lvl[i] = 0;  // skip this lattice position only (level 0)
}
}  // namespace

cudaError_t occ_merge_solid_launch(uint8_t* lvl, const uint8_t* solidDev, int n, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  occ_merge_solid_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(lvl, solidDev, n);
  return cudaGetLastError();
}

cudaError_t occ_set_cells_launch(uint8_t* raw, const int64_t* pairsDev, int n, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  occ_set_cells_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(raw, reinterpret_cast<const long long*>(pairsDev), n);
  return cudaGetLastError();
}

cudaError_t occ_levels_launch(const uint8_t* pyr, uint8_t* lvl, const OccView& L, cudaStream_t stream) {
  const long long n = (long long)L.nx * L.ny * L.nz;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  uint8_t *gx = lvl + n, *gy = lvl + 2 * n;  // scratch behind lvl (the array holds 3 n bytes)
  occ_dist_x_kernel<<<blocks, 256, 0, stream>>>(pyr + L.off[0], gx, L);
  occ_dist_axis_kernel<<<blocks, 256, 0, stream>>>(gx, gy, L, 1, 0);
  occ_dist_axis_kernel<<<blocks, 256, 0, stream>>>(gy, lvl, L, 2, 1);
  return cudaGetLastError();
}

// Level 0 of pyr from raw AND lvl for L.thr in two launches (what a freshly decoded timestep needs before its frame); false when
// a slice of cells does not fit in shared memory (the caller then takes occ_pyramid_launch + occ_levels_launch).
bool occ_refresh_launch(const uint8_t* raw, uint8_t* pyr, uint8_t* lvl, const OccView& L, cudaStream_t stream, cudaError_t* err) {
  const long long n = (long long)L.nx * L.ny * L.nz;
  const size_t smem = 2 * (size_t)L.nx * L.ny;
  if (smem > 40 * 1024) return false;
  uint8_t* gy = lvl + 2 * n;
  occ_dilate_dist_xy_kernel<<<(unsigned)L.nz, 256, smem, stream>>>(raw, pyr + L.off[0], gy, L);
  occ_dist_axis_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(gy, lvl, L, 2, 1);
  *err = cudaGetLastError();
  return true;
}

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

cudaError_t occ_pyramid_launch(const uint8_t* raw, uint8_t* pyr, const OccView& L, cudaStream_t stream) {
  dim3 grid((unsigned)(((L.nx - 1) >> 3) + 1), (unsigned)(((L.ny - 1) >> 3) + 1), (unsigned)(((L.nz - 1) >> 3) + 1));
  occ_pyr_low_kernel<<<grid, 512, 0, stream>>>(raw, pyr, L);
  if (L.nlev > 4) occ_pyr_high_kernel<<<1, 256, 0, stream>>>(pyr, L);
  return cudaGetLastError();
}

}  // namespace vr
