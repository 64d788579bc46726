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
#include "kd_parse6.h"
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
}

}  // namespace vr

// ---- work lists of the persistent kernel (kd_decode.cu: kd_decode6_kernel) -------------------------------------------------
// Two queues instead of one CTA list: parse items (one unit = an eighth of a tile = one warp's work each, self-contained
// 96-byte records) and constant groups (vr_fill_job, stored by the TMA unit).  Both in destination-address order.  Culling as above.
namespace vr {
// Returns false when a unit's segment exceeds stageChunksMax (cannot happen for a well-formed stream and the kernel's capacity;
// the caller then takes the classic kernel).
inline bool kd_build_worklist6(const vr_kd_geom& g, int n, const KdIndex* const* idx, const int64_t* dstOffset, int64_t rowPitch,
                               int64_t slicePitch, const int64_t* origins, const int64_t* occBase, const int32_t* tileCell,
                               const uint8_t* needed, std::vector<vr_occ_set>* culled, uint32_t stageChunksMax, const vr6::P6Geom& pg,
                               std::vector<vr_parse_item>* parse, std::vector<uint32_t>* parseBrick, std::vector<vr_fill_job>* fills,
                               bool heavyFirst = false) {
  struct PItem { int64_t key; vr_parse_item w; uint32_t brick; };
  struct FItem { int64_t key; vr_fill_job f; };
  std::vector<PItem> pitems;
  std::vector<FItem> fitems;
  const uint32_t nt = 1u << g.Dt;
  {
    size_t live = 0;
    for (int b = 0; b < n; ++b) live += (size_t)idx[b]->numLiveTiles;
    pitems.reserve(live * 8);
  }
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
        const uint32_t blk = hj.tileLeafBlock[t];
        const uint32_t* lo = idx[b]->leafOff.data() + (size_t)blk * 256;
        const uint64_t tileEnd = idx[b]->blockSymEnd[blk];
        // offset of the first live leaf at or behind every leaf index (the end of a run's symbols), in one backward pass
        uint64_t nextLive[257];
        nextLive[256] = tileEnd;
        for (int l = 255; l >= 0; --l) nextLive[l] = lo[l] != VR_PRUNED ? (uint64_t)lo[l] : nextLive[l + 1];
        const uint32_t runLen = 1u << pg.runLog, nRuns = 32u >> pg.runLog;
        for (uint32_t u = 0; u < 8; ++u) {
          vr_parse_item w{};
          w.tileUnit = (t << 3) | u;
          w.leaf0 = blk * 256u;
          uint32_t total = 0;
          for (uint32_t r = 0; r < nRuns; ++r) {
            const uint32_t base = vr6::p6_leaf_of(pg, u, r << pg.runLog);
            // the run's symbols lie between its first live leaf and the next live leaf behind it (or the end of the tile)
            uint64_t begin = ~0ull, last = 0;
            for (uint32_t l = base; l < base + runLen; ++l)
              if (lo[l] != VR_PRUNED) { if (begin == ~0ull) begin = lo[l]; last = lo[l]; }
            uint32_t chunks = 0, segOff = 0;
            if (begin != ~0ull) {
              uint64_t end = nextLive[base + runLen];
              // ... and a leaf subtree is at most 63 nodes + 64 voxel items of 8 symbols long
              if (last + 575u < end) end = last + 575u;
              // staged range: from the 16-byte chunk of the first symbol to the chunk that holds the word after the last
              // symbol's word (the window keeps one word of look-ahead)
              const uint64_t c0 = (begin >> 4) >> 2, c1 = ((((end > begin ? end - 1 : begin) >> 4) + 2) >> 2) + 1;
              chunks = (uint32_t)(c1 - c0);
              segOff = (uint32_t)(16u * c0);  // the packed stream sits at offset 0 of the arena
            }
            w.run[r].segOff = segOff;
            w.run[r].packed = VR_RUN_PACK(chunks, base, total);
            total += chunks;
          }
          for (uint32_t r = nRuns; r < 4u; ++r) w.run[r].packed = VR_RUN_PACK(0u, 0u, total);  // (stage_unit's one loop over all runs)
          if (total > stageChunksMax) return false;
          w.segChunks = total;
          w.dst = key;
          int32_t uo[3];
          vr6::p6_unit_origin(pg, u, uo);
          w.ux = (int32_t)(origins[3 * b] + ox + uo[0]);
          w.uy = (int32_t)(origins[3 * b + 1] + oy + uo[1]);
          w.uz = (int32_t)(origins[3 * b + 2] + oz + uo[2]);
          w.occCell = occ ? (int32_t)(occBase[b] + tileCell[t]) : -1;
          pitems.push_back({key + u, w, (uint32_t)b});
        }
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
  // Parse units in address order.  (heavyFirst = longest segments first, in classes of 1 KB: measured slower -- the two
  // 16-byte halves of a 32-byte sector belong to units four tickets apart, and in address order L2 merges them before they
  // reach DRAM: 0.70 ms against 0.87 ms on the config-3 layer.)
  // (sorted through an index array: the records are 112 bytes)
  std::vector<std::pair<int64_t, uint32_t>> ord(pitems.size());
  for (size_t i = 0; i < pitems.size(); ++i) {
    const int64_t cls = heavyFirst ? (int64_t)(0xFFFFu - (pitems[i].w.segChunks >> 6)) : 0;
    ord[i] = {(cls << 47) + pitems[i].key, (uint32_t)i};
  }
  std::sort(ord.begin(), ord.end());
  std::sort(fitems.begin(), fitems.end(), [](const FItem& a, const FItem& b) { return a.key < b.key; });
  parse->resize(pitems.size());
  if (parseBrick) parseBrick->resize(pitems.size());
  for (size_t i = 0; i < ord.size(); ++i) { (*parse)[i] = pitems[ord[i].second].w; if (parseBrick) (*parseBrick)[i] = pitems[ord[i].second].brick; }
  fills->resize(fitems.size());
  for (size_t i = 0; i < fitems.size(); ++i) (*fills)[i] = fitems[i].f;
  return true;
}
}  // namespace vr
