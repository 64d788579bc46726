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
// So a tile is needed iff its cell lies in [H - 1 - ceil(reach/cs), H + ceil((reach+1)/cs)] per axis for some hot
// cell H.  Cells the plan does not cover (halo, bricks of another plan) count as hot.  Tiles that are not needed are
// never read by the frame: the lazy decode skips them and only records their maxima in the occupancy grid.
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

// hot[c] = 1 for every cell of the grid that is not covered by a tile with max <= thr.
// cellOfTile(b, t) -> linear cell index of tile t of brick b (or < 0 if unknown).
// needed (n * nt bytes) receives 1 for tiles the frame can read.
template <typename CellOf>
inline void cull_tiles(const CullGrid& cg, int n, uint32_t nt, const KdIndex* const* idx, CellOf cellOfTile, int thr,
                       const int reach[3], uint8_t* needed) {
  const size_t ncell = (size_t)cg.nc[0] * cg.nc[1] * cg.nc[2];
  std::vector<uint8_t> hot(ncell, 1), tmp(ncell);
  for (int b = 0; b < n; ++b)
    for (uint32_t t = 0; t < nt; ++t) {
      const int64_t c = cellOfTile(b, t);
      if (c >= 0 && (size_t)c < ncell) hot[(size_t)c] = idx[b]->tileMax[t] > thr ? 1 : 0;
    }
  // separable dilation: out[i] = OR of in[i - up .. i + dn]  (a hot cell H marks [H - dn, H + up])
  const int64_t stride[3] = {1, cg.nc[0], (int64_t)cg.nc[0] * cg.nc[1]};
  for (int a = 0; a < 3; ++a) {
    const int cs = 1 << cg.lg[a];
    const int dn = 1 + (reach[a] + cs - 1) / cs;  // cells below the hot cell
    const int up = (reach[a] + 1 + cs - 1) / cs;  // cells above it
    const int na = cg.nc[a];
    const int a1 = (a + 1) % 3, a2 = (a + 2) % 3;
    for (int j = 0; j < cg.nc[a2]; ++j)
      for (int i = 0; i < cg.nc[a1]; ++i) {
        const int64_t base = i * stride[a1] + j * stride[a2];
        // tmp[k] = any hot cell in [k - up, k + dn]: sliding count
        int count = 0;
        for (int k = 0; k < std::min(na, dn + 1); ++k) count += hot[(size_t)(base + k * stride[a])];  // window of k = 0 is [0, dn]
        for (int k = 0; k < na; ++k) {
          tmp[(size_t)(base + k * stride[a])] = count > 0 ? 1 : 0;
          const int add = k + 1 + dn, rem = k - up;
          if (add < na) count += hot[(size_t)(base + add * stride[a])];
          if (rem >= 0) count -= hot[(size_t)(base + rem * stride[a])];
        }
      }
    hot.swap(tmp);
  }
  for (int b = 0; b < n; ++b)
    for (uint32_t t = 0; t < nt; ++t) {
      const int64_t c = cellOfTile(b, t);
      needed[(size_t)b * nt + t] = (c < 0 || (size_t)c >= ncell) ? 1 : hot[(size_t)c];
    }
}

}  // namespace vr
