// kd_worklist.h -- host side: the CTA work list of one decode launch (kd_decode.cu).
//
// One entry per CTA: a tile that has to be parsed, or the leader of a group of x-neighbouring constant tiles
// (vr_fill_group_of; the other members of a group get no CTA).  Entries are sorted by the address of the
// tile origin in the destination slab, so CTAs that run at the same time fill whole slab rows across brick
// borders: tools/fill_probe.cu measures 5.6-7.3 TB/s for fills issued in address order against 4.2 TB/s in
// kd-path order, where 32-byte pieces of one 128-byte line are written far apart in time.
#pragma once
#include <algorithm>
#include <utility>
#include <vector>

#include "kd_decode_core.h"
#include "kd_stream.h"

namespace vr {

// order: 0 = address order, 1 = parse tiles first, 2 = parse tiles last (each class in address order)
inline void kd_build_worklist(const vr_kd_geom& g, int n, const KdIndex* const* idx, const int64_t* dstOffset, int64_t rowPitch,
                              int64_t slicePitch, bool vec16, int order, std::vector<vr_kd_work>* out) {
  struct Item { int64_t key; vr_kd_work w; };
  std::vector<Item> items;
  const uint32_t nt = 1u << g.Dt;
  items.reserve((size_t)n * nt / 4 + 16);
  for (int b = 0; b < n; ++b) {
    vr_kd_job hj{};  // host view of the two tile tables, which is all vr_fill_group_of reads
    hj.tileLeafBlock = idx[b]->tileLeafBlock.data();
    hj.tileVal = idx[b]->tileVal.data();
    for (uint32_t t = 0; t < nt; ++t) {
      vr_kd_work w;
      w.id = ((uint32_t)b << g.Dt) | t;
      w.L = VR_WORK_PARSE;
      w.vals = 0;
      const bool parse = hj.tileLeafBlock[t] != VR_PRUNED;
      if (!parse) {
        if (vec16) {
          const int L = vr_fill_group_of(g, hj, t, &w.vals);
          if (L < 0) continue;
          w.L = (uint32_t)L;
        } else {
          w.L = 0;
          w.vals = hj.tileVal[t];
        }
      }
      int ox, oy, oz;
      vr_tile_origin(g, t, &ox, &oy, &oz);
      int64_t key = dstOffset[b] + (int64_t)oz * slicePitch + (int64_t)oy * rowPitch + ox;
      if (order == 1 && !parse) key += int64_t(1) << 60;
      if (order == 2 && parse) key += int64_t(1) << 60;
      items.push_back({key, w});
    }
  }
  std::sort(items.begin(), items.end(), [](const Item& a, const Item& b) { return a.key < b.key; });
  out->resize(items.size());
  for (size_t i = 0; i < items.size(); ++i) (*out)[i] = items[i].w;
}

}  // namespace vr
