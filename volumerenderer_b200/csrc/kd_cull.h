// kd_cull.h -- host side: which tiles of a decode plan a frame can read (the intent behind the stub
// /root/reference/volume_renderer/isosurface_compressed.frag:23-29 `node{idx, boundMin, boundMax, scalarMin, scalarMax}`
// and the never-enabled SSBO path of main.cpp:203-237: render from the stream, touching only what the view needs).
//
// The march kernels (render_march.cu) fetch voxels only around lattice positions whose occupancy cell is not void
// for the mode's threshold `thr` (occupancy.h: level 0 of the pyramid = max of the raw cell maxima over c + {0,1}^3).
// The exact maximum of every tile is known from the stream index before anything is decoded (KdIndex::tileMax,
// kd_stream.cpp), so the set of voxels a frame can fetch is known too:
//   * a look-up at base texel b (cell c) is non-void only if some raw cell H in c + {0,1}^3 is "hot" (max > thr),
//     i.e. b >= lo(H) - cellsize and b <= hi(H);
//   * everything the kernel then reads lies within `reach` voxels of b per axis: composite reads the 8 taps of the
//     sample itself (b, b+1: reach 0); isosurface looks up at p + dirStep and reads the taps of p and p + dirStep
//     (isosurface.frag:120-121), of the bisection points between them (:26-41) and of the gradient probes at +-0.01
//     around the hit (:18,52-59): all base texels within ceil((|step| + 0.01) * G) of b, reach = that + 1.
// So a tile is needed iff its cell lies in [c - ceil(reach/cs), c + ceil((reach+1)/cs)] per axis for some look-up cell
// c that fetches (for max-only skipping: c in {H - 1, H} for a hot cell H).  Cells the plan does not cover (halo,
// bricks of another plan) count as hot.  Tiles that are not needed are never read by the frame: the lazy decode skips
// them and only records their maxima in the occupancy grid.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include "kd_stream.h"

namespace vr {

struct CullGrid {
  int nc[3];       // occupancy cells of the destination volume per axis
  int lg[3];       // log2 cell size per axis (VR_OCC_LX/LY/LZ)
};

// voxels per axis a frame reads beyond the base texel of a non-void look-up (see above); mode 0 = composite
inline void cull_reach(int mode, const float step[3], const int64_t G[3], int reach[3]) {
  for (int a = 0; a < 3; ++a) {
    if (mode == 0) { reach[a] = 0; continue; }
    const double r = ((double)std::fabs(step[a]) + 0.01) * (double)G[a];
    reach[a] = (int)std::ceil(r) + 1;
  }
}

// OR over a box window per cell: out[k] = any in[k - below[a] .. k + above[a]] along every axis (clipped to the grid)
inline void cull_dilate(const CullGrid& cg, std::vector<uint8_t>& grid, const int below[3], const int above[3]) {
  std::vector<uint8_t> tmp(grid.size());
  const int64_t stride[3] = {1, cg.nc[0], (int64_t)cg.nc[0] * cg.nc[1]};
  for (int a = 0; a < 3; ++a) {
    const int na = cg.nc[a], a1 = (a + 1) % 3, a2 = (a + 2) % 3;
    for (int j = 0; j < cg.nc[a2]; ++j)
      for (int i = 0; i < cg.nc[a1]; ++i) {
        const int64_t base = i * stride[a1] + j * stride[a2];
        int count = 0;  // sliding count over [k - below, k + above]
        for (int k = 0; k < std::min(na, above[a] + 1); ++k) count += grid[(size_t)(base + k * stride[a])];
        for (int k = 0; k < na; ++k) {
          tmp[(size_t)(base + k * stride[a])] = count > 0 ? 1 : 0;
          const int add = k + 1 + above[a], rem = k - below[a];
          if (add < na) count += grid[(size_t)(base + add * stride[a])];
          if (rem >= 0) count -= grid[(size_t)(base + rem * stride[a])];
        }
      }
    grid.swap(tmp);
  }
}

// The isosurface march needs a crossing s(p) < iso <= s(p + dirStep) (isosurface.frag:126).  Besides cells whose
// maximum proves s(p + dirStep) < iso ("void"), a look-up cell is "solid-high" when every voxel the taps of p can touch
// is >= thr + 2, i.e. s(p) >= (thr + 2)/255 - rounding > (thr + 1)/255 >= iso (one grey level of margin over the last-ulp
// behaviour of the fp32 lerps): the crossing test cannot fire and nothing else happens at that lattice position, so the
// fetch is skipped there too (the lazy path hands this mask to the march).  p's base texel lies within
// ceil(|step| * G) of the look-up texel; stepReach = that + 1.
inline void cull_step_reach(const float step[3], const int64_t G[3], int stepReach[3]) {
  for (int a = 0; a < 3; ++a) stepReach[a] = (int)std::ceil((double)std::fabs(step[a]) * (double)G[a]) + 1;
}

// cellOfTile(b, t) -> linear cell index of tile t of brick b (or < 0 if unknown).
// needed (n * nt bytes) receives 1 for tiles the frame can read; solidHigh (optional, one byte per cell of the grid;
// only filled for mode != 0) receives the mask described above.
template <typename CellOf>
inline void cull_tiles(const CullGrid& cg, int n, uint32_t nt, const KdIndex* const* idx, CellOf cellOfTile, int mode, int thr,
                       const int reach[3], const int stepReach[3], uint8_t* needed, std::vector<uint8_t>* solidHigh) {
  const size_t ncell = (size_t)cg.nc[0] * cg.nc[1] * cg.nc[2];
  std::vector<uint8_t> hot(ncell, 1), low(ncell, 1);  // unknown cells: hot, and not provably high
  for (int b = 0; b < n; ++b)
    for (uint32_t t = 0; t < nt; ++t) {
      const int64_t c = cellOfTile(b, t);
      if (c < 0 || (size_t)c >= ncell) continue;
      hot[(size_t)c] = idx[b]->tileMax[t] > thr ? 1 : 0;
      low[(size_t)c] = idx[b]->tileMin[t] <= thr + 1 ? 1 : 0;  // one grey level of margin, see above
    }
  int below[3], above[3];
  // look-up cell c fetches unless it is void: no hot cell in c + {0,1}^3 (level 0 of the occupancy pyramid) ...
  std::vector<uint8_t> fetch = hot;
  for (int a = 0; a < 3; ++a) { below[a] = 0; above[a] = 1; }
  cull_dilate(cg, fetch, below, above);
  // ... or solid-high: no voxel <= thr + 1 within stepReach of the cell (voxels [lo(c) - R, hi(c) + R + 1])
  if (mode != 0 && stepReach) {
    for (int a = 0; a < 3; ++a) {
      const int cs = 1 << cg.lg[a];
      below[a] = (stepReach[a] + cs - 1) / cs;
      above[a] = (stepReach[a] + 1 + cs - 1) / cs;
    }
    cull_dilate(cg, low, below, above);  // low[c] = some voxel <= thr + 1 in reach of c
    for (size_t c = 0; c < ncell; ++c) fetch[c] = fetch[c] && low[c];
    if (solidHigh) {
      solidHigh->resize(ncell);
      for (size_t c = 0; c < ncell; ++c) (*solidHigh)[c] = low[c] ? 0 : 1;
    }
  } else if (solidHigh) {
    solidHigh->clear();
  }
  // a fetching look-up cell c reads voxels [lo(c) - reach, hi(c) + reach + 1]
  for (int a = 0; a < 3; ++a) {
    const int cs = 1 << cg.lg[a];
    below[a] = (reach[a] + 1 + cs - 1) / cs;  // tile k is read from cells up to this far above it ...
    above[a] = (reach[a] + cs - 1) / cs;      // ... and this far below it
  }
  // needed[k] = any fetch cell c with k in [c - ceil(reach/cs), c + ceil((reach+1)/cs)]  <=>  c in [k - ceil((reach+1)/cs), k + ceil(reach/cs)]
  cull_dilate(cg, fetch, below, above);
  for (int b = 0; b < n; ++b)
    for (uint32_t t = 0; t < nt; ++t) {
      const int64_t c = cellOfTile(b, t);
      needed[(size_t)b * nt + t] = (c < 0 || (size_t)c >= ncell) ? 1 : fetch[(size_t)c];
    }
}

}  // namespace vr
