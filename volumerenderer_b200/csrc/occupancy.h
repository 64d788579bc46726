// occupancy.h -- min-max metadata for empty-space skipping (north_star: "empty-space skipping driven by
// the Octree / MidRangeTree min-max metadata").  The reference has no usable bound structure (Octree.cpp
// stores averages only, MidRangeTree's mid/half-range are lossy), so the bound kept here is the
// exact maximum of the decoded voxels per cell, produced by the decode kernel itself (one byte per decode
// tile) or, for raw uploads, by a reduction over the slab.
//
//   raw[c]   = max voxel of cell c (32 x 32 x 16 voxels, aligned to the global voxel grid)
//   dil[c]   = max of raw over c + {0,1}^3: bound on all 8 trilinear taps of any sample whose base texel
//              lies in c
//   macro[m] = max of dil over 8 x 8 x 8 cells (256 x 256 x 128 voxels = one reference brick)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vr {

#define VR_OCC_LX 5
#define VR_OCC_LY 5
#define VR_OCC_LZ 4
#define VR_OCC_MACRO 3  // log2 cells per macro cell and axis

struct OccView {        // what the march kernels read (null dil = no skipping)
  const uint8_t* dil;
  const uint8_t* macro;
  int nx, ny, nz;       // cells
  int mx, my, mz;       // macro cells
  int sx, sy, sz;       // cell of slab-local voxel l along x = (l + sx) >> VR_OCC_LX
  int thr;              // a cell / macro cell is skippable when its bound <= thr
};

cudaError_t occ_reduce_box_launch(const uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, const int64_t lo[3], const int64_t ext[3],
                                  uint8_t* raw, int nx, int ny, int nz, const int shift[3], cudaStream_t stream);
cudaError_t occ_dilate_launch(const uint8_t* raw, uint8_t* dil, uint8_t* macro, int nx, int ny, int nz, int mx, int my, int mz,
                              cudaStream_t stream);

}  // namespace vr
