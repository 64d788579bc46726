// occupancy.h -- min-max metadata for empty-space skipping (north_star: "empty-space skipping driven by
// the Octree / MidRangeTree min-max metadata").  The reference has no usable bound structure (Octree.cpp
// stores averages only, MidRangeTree's mid/half-range are lossy), so the bound kept here is the
// exact maximum of the decoded voxels per cell, produced by the decode kernel itself (one byte per warp
// and decode tile) or, for raw uploads, by a reduction over the slab.
//
//   raw[c][8] = maxima of cell c (32 x 32 x 16 voxels, aligned to the global voxel grid), one slot per
//               writer warp of the decode kernel; max over the 8 slots = max voxel of the cell
//   pyr, level 0 = max of raw over c + {0,1}^3: bound on all 8 trilinear taps of any sample whose base
//               texel lies in c ("dilated")
//   pyr, level l = max over 2^l x 2^l x 2^l level-0 cells (an octree of maxima up to one root cell)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vr {

#define VR_OCC_LX 5
#define VR_OCC_LY 5
#define VR_OCC_LZ 4
#define VR_OCC_SLOTS 8
#define VR_OCC_MAX_LEVELS 12

#define VR_OCC_SOLID 0xFF
#define VR_OCC_DIST_CAP 16  // lvl[] radii are capped below this many cells (step of the rm_yslab bench: 0.364 ms at 16, 0.366 at 32, 0.372 at 64)

struct OccView {        // what the march kernels read (null pyr = no skipping)
  const uint8_t* pyr;
  const uint8_t* lvl;   // per level-0 cell, for this thr: VR_OCC_SOLID, or the radius R (cells, Chebyshev) of the largest void cube of cells around it
  int nlev;
  int off[VR_OCC_MAX_LEVELS];  // byte offset of level l inside pyr
  int nx, ny, nz;       // level-0 cells; level l has ((n - 1) >> l) + 1 cells per axis
  int sx, sy, sz;       // cell of slab-local voxel l along x = (l + sx) >> VR_OCC_LX
  int thr;              // a cell is skippable when its bound <= thr
};

// raw[cell][0..7] = max over (cell ∩ box) for every cell the box touches
cudaError_t occ_reduce_box_launch(const uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, const int64_t lo[3], const int64_t ext[3],
                                  uint8_t* raw, int nx, int ny, int nz, const int shift[3], cudaStream_t stream);
// raw[cell] = value for n (cell, value) pairs of 2 x int64: cells whose voxels are known without being stored (lazy decode, kd_cull.h)
cudaError_t occ_set_cells_launch(uint8_t* raw, const int64_t* pairsDev, int n, cudaStream_t stream);
// lvl[c] = 0 (skip the fetch at this lattice position) where solid[c] != 0 and the cell is not void already: look-up cells
// of the isosurface march in which the crossing test cannot fire because every reachable voxel is >= isoValue (kd_cull.h)
cudaError_t occ_merge_solid_launch(uint8_t* lvl, const uint8_t* solidDev, int n, cudaStream_t stream);
// all pyramid levels from raw (two launches)
cudaError_t occ_pyramid_launch(const uint8_t* raw, uint8_t* pyr, const OccView& layout, cudaStream_t stream);
// lvl[c] for layout.thr: radius of the largest cube of void cells around c (three separable passes over level 0 of pyr; lvl holds
// 3 x cells bytes, the last two thirds are scratch) -- a ray leaves such a cube in one skip, where octree cells made it stop at
// every aligned boundary
cudaError_t occ_levels_launch(const uint8_t* pyr, uint8_t* lvl, const OccView& layout, cudaStream_t stream);
// level 0 of pyr (the dilated maxima) and lvl in two launches -- the per-timestep refresh; the octree levels above 0 are then stale
// (nothing on the render path reads them; vr_volume_occupancy_download rebuilds them on demand).  false = grid too wide, not launched
bool occ_refresh_launch(const uint8_t* raw, uint8_t* pyr, uint8_t* lvl, const OccView& layout, cudaStream_t stream, cudaError_t* err);

}  // namespace vr
