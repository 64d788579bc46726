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

// order: 0 = address order, 1 = parse tiles first, 2 = parse tiles last (each class in address order).
// attach > 0 (needs fills != null): up to `attach` constant groups ride along with every parse CTA, which stores
// them before it starts parsing -- the fills (pure HBM writes) then overlap the parsing (issue bound) of the other
// CTAs on the same SM instead of forming a phase of their own.  Groups are dealt out evenly in address order;
// what does not fit stays a CTA of its own.  fills receives the attached groups; origins = voxel origin of every
// brick in the destination (3 per brick), occBase / tileCell as in the decode kernel (null = no occupancy).
// needed (n * 2^Dt bytes, kd_cull.h; needs occBase, tileCell and culled): tiles a frame cannot read get no work --
// a parse tile is dropped when it is not needed, a constant group when none of its members is -- and their exact
// maxima go to `culled` instead, to be written into the occupancy grid without decoding anything.
inline void kd_build_worklist(const vr_kd_geom& g, int n, const KdIndex* const* idx, const int64_t* dstOffset, int64_t rowPitch,
                              int64_t slicePitch, bool vec16, int order, std::vector<vr_kd_work>* out,
                              std::vector<vr_fill_job>* fills = nullptr, int attach = 0, const int64_t* origins = nullptr,
                              int lateMode = 0,
                              const int64_t* occBase = nullptr, const int32_t* tileCell = nullptr,
                              const uint8_t* needed = nullptr, std::vector<vr_occ_set>* culled = nullptr,
                              std::vector<vr_stage_range>* ranges = nullptr, int64_t stageBytes = 0,
                              int only = 0 /* timing experiments: 1 = parse tiles only, 2 = constant fills only */) {
  struct Item { int64_t key; vr_kd_work w; };
  struct FItem { int64_t key; vr_kd_work w; vr_fill_job f; };
  std::vector<Item> items;
  std::vector<FItem> fitems;
  const uint32_t nt = 1u << g.Dt;
  const bool split = fills && attach > 0 && vec16 && origins;
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
      if (needed && culled && occBase && tileCell && occBase[b] >= 0) {
        const uint32_t members = parse ? 1u : (1u << w.L);
        uint32_t mt[8];
        bool any = false;
        for (uint32_t i = 0; i < members; ++i) {  // member i of a group: x index of the leader + i, scattered to the x path bits
          uint32_t tm = t;
          for (uint32_t j = 0; j < w.L && !parse; ++j) tm |= ((i >> j) & 1u) << g.xrBit[j];
          mt[i] = tm;
          any = any || needed[(size_t)b * nt + tm];
        }
        if (!any) {
          for (uint32_t i = 0; i < members; ++i) culled->push_back({occBase[b] + tileCell[mt[i]], (int64_t)idx[b]->tileMax[mt[i]]});
          continue;
        }
      }
      if ((only == 1 && !parse) || (only == 2 && parse)) continue;
      int ox, oy, oz;
      vr_tile_origin(g, t, &ox, &oy, &oz);
      int64_t key = dstOffset[b] + (int64_t)oz * slicePitch + (int64_t)oy * rowPitch + ox;
      if (!parse && split) {
        vr_fill_job f;
        f.x = (int32_t)(origins[3 * b] + ox);
        f.y = (int32_t)(origins[3 * b + 1] + oy);
        f.z = (int32_t)(origins[3 * b + 2] + oz);
        f.L = w.L;
        f.vals = w.vals;
        f.occCell = (occBase && tileCell && occBase[b] >= 0) ? occBase[b] + tileCell[t] : -1;
        fitems.push_back({key, w, f});
        continue;
      }
      if (order == 1 && !parse) key += int64_t(1) << 60;
      if (order == 2 && parse) key += int64_t(1) << 60;
      items.push_back({key, w});
    }
  }
  std::sort(items.begin(), items.end(), [](const Item& a, const Item& b) { return a.key < b.key; });
  if (fills) fills->clear();
  if (split) {
    std::sort(fitems.begin(), fitems.end(), [](const FItem& a, const FItem& b) { return a.key < b.key; });
    const size_t P = items.size(), F = fitems.size();
    size_t next = 0;
    for (size_t i = 0; i < P && next < F; ++i) {  // even deal: parse CTA i gets groups [i*F/P, (i+1)*F/P), capped
      size_t want = std::min<size_t>((F * (i + 1)) / P - (F * i) / P, (size_t)attach);
      want = std::min(want, F - next);
      // bit 63 = store the groups after parsing instead of before (every other parse CTA, see kd_decode.cu)
      const uint64_t late = lateMode == 1 ? (i & 1) : (lateMode == 2 ? 1 : 0);
      items[i].w.vals = (uint64_t)fills->size() | ((uint64_t)want << 32) | (late << 63);
      for (size_t k = 0; k < want; ++k) fills->push_back(fitems[next++].f);
    }
    for (; next < F; ++next) items.push_back({fitems[next].key, fitems[next].w});  // leftovers: CTAs of their own, after the parse CTAs
  }
  out->resize(items.size());
  for (size_t i = 0; i < items.size(); ++i) (*out)[i] = items[i].w;
  // ranges[i] = {first 16-byte chunk, chunks} of the stream segment parse CTA i stages in shared memory (VrWinStage):
  // from the word of the tile's first symbol to 3 words past the word of its last one (the window's look-ahead);
  // {0, 0} = not staged (constant group, or a segment larger than the staging area: that CTA parses from global memory)
  if (ranges) {
    ranges->assign(items.size(), vr_stage_range{0u, 0u});
    const uint32_t tmask = (1u << g.Dt) - 1u;
    for (size_t i = 0; i < items.size() && stageBytes > 0; ++i) {
      const vr_kd_work& w = items[i].w;
      if (w.L != VR_WORK_PARSE) continue;
      const KdIndex& ix = *idx[w.id >> g.Dt];
      const uint32_t blk = ix.tileLeafBlock[w.id & tmask];
      const uint64_t begin = ix.blockSymBegin[blk], end = ix.blockSymEnd[blk];
      if (end <= begin) continue;
      const uint64_t c0 = (begin >> 4) >> 2, c1 = ((((end - 1) >> 4) + 3) >> 2) + 1;
      if ((int64_t)(c1 - c0) * 16 <= stageBytes) (*ranges)[i] = vr_stage_range{(uint32_t)c0, (uint32_t)(c1 - c0)};
    }
  }
}

}  // namespace vr

// ---- work lists of the persistent kernel (kd_decode.cu: kd_decode6_kernel) -------------------------------------------------
// Two queues instead of one CTA list: parse items (one tile each, self-contained so that the CTA needs one 48-byte read per
// tile) and constant groups (vr_fill_job, stored by the TMA unit).  Both in destination-address order.  Culling as above.
namespace vr {
inline void kd_build_worklist6(const vr_kd_geom& g, int n, const KdIndex* const* idx, const int64_t* dstOffset, int64_t rowPitch,
                               int64_t slicePitch, const int64_t* origins, const int64_t* occBase, const int32_t* tileCell,
                               const uint8_t* needed, std::vector<vr_occ_set>* culled, uint32_t stageChunksMax,
                               std::vector<vr_parse_item>* parse, std::vector<vr_fill_job>* fills) {
  struct PItem { int64_t key; vr_parse_item w; };
  struct FItem { int64_t key; vr_fill_job f; };
  std::vector<PItem> pitems;
  std::vector<FItem> fitems;
  const uint32_t nt = 1u << g.Dt;
  for (int b = 0; b < n; ++b) {
    vr_kd_job hj{};
    hj.tileLeafBlock = idx[b]->tileLeafBlock.data();
    hj.tileVal = idx[b]->tileVal.data();
    const bool occ = occBase && tileCell && occBase[b] >= 0;
    for (uint32_t t = 0; t < nt; ++t) {
      const bool isParse = hj.tileLeafBlock[t] != VR_PRUNED;
      uint64_t vals = 0;
      int L = 0;
      if (!isParse) {
        L = vr_fill_group_of(g, hj, t, &vals);
        if (L < 0) continue;
      }
      if (needed && culled && occ) {
        const uint32_t members = isParse ? 1u : (1u << L);
        uint32_t mt[8];
        bool any = false;
        for (uint32_t i = 0; i < members; ++i) {
          uint32_t tm = t;
          for (int j = 0; j < L && !isParse; ++j) tm |= ((i >> j) & 1u) << g.xrBit[j];
          mt[i] = tm;
          any = any || needed[(size_t)b * nt + tm];
        }
        if (!any) {
          for (uint32_t i = 0; i < members; ++i) culled->push_back({occBase[b] + tileCell[mt[i]], (int64_t)idx[b]->tileMax[mt[i]]});
          continue;
        }
      }
      int ox, oy, oz;
      vr_tile_origin(g, t, &ox, &oy, &oz);
      const int64_t key = dstOffset[b] + (int64_t)oz * slicePitch + (int64_t)oy * rowPitch + ox;
      if (isParse) {
        vr_parse_item w{};
        w.brick = (uint32_t)b;
        w.tile = t;
        w.blk = hj.tileLeafBlock[t];
        const uint64_t begin = idx[b]->blockSymBegin[w.blk], end = idx[b]->blockSymEnd[w.blk];
        // staged range: from the 16-byte chunk of the tile's first symbol to the chunk that holds the word after its last
        // symbol's word (the window keeps one word of look-ahead)
        const uint64_t c0 = (begin >> 4) >> 2, c1 = ((((end > begin ? end - 1 : begin) >> 4) + 2) >> 2) + 1;
        w.seg0 = (uint32_t)c0;
        w.segChunks = (c1 - c0) <= stageChunksMax ? (uint32_t)(c1 - c0) : 0u;
        w.dst = key;
        w.occCell = occ ? occBase[b] + tileCell[t] : -1;
        pitems.push_back({key, w});
      } else {
        vr_fill_job f;
        f.x = (int32_t)(origins[3 * b] + ox);
        f.y = (int32_t)(origins[3 * b + 1] + oy);
        f.z = (int32_t)(origins[3 * b + 2] + oz);
        f.L = (uint32_t)L;
        f.vals = vals;
        f.occCell = occ ? occBase[b] + tileCell[t] : -1;
        fitems.push_back({key, f});
      }
    }
  }
  std::sort(pitems.begin(), pitems.end(), [](const PItem& a, const PItem& b) { return a.key < b.key; });
  std::sort(fitems.begin(), fitems.end(), [](const FItem& a, const FItem& b) { return a.key < b.key; });
  parse->resize(pitems.size());
  for (size_t i = 0; i < pitems.size(); ++i) (*parse)[i] = pitems[i].w;
  fills->resize(fitems.size());
  for (size_t i = 0; i < fitems.size(); ++i) (*fills)[i] = fitems[i].f;
}
}  // namespace vr
