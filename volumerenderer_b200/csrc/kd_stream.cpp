// kd_stream.cpp -- container I/O, grammar walk and decode-index construction (host side).
// Reference behaviour mirrored here: VolumeKdtree::save / open
// (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:521-594) and the stream grammar
// produced by convertToPreorder (:631-724) and consumed by levelCut (:726-835).
#include "kd_stream.h"

#include <cstdio>
#include <cstring>

namespace vr {

int kd_check_header(const vr_kd_header& h, vr_kd_geom* g) {
  if (h.X <= 0 || h.Y <= 0 || h.Z <= 0 || h.numActiveNodes <= 0) return VR_ERR_FORMAT;
  if (vr_kd_make_geom(h.X, h.Y, h.Z, g) != 0) return VR_ERR_FORMAT;
  if (h.origTreeDepth != g->D0 || h.maxTreeDepth != g->Dmax) return VR_ERR_FORMAT;
  for (int i = 0; i < 3; ++i)
    if (h.rootMin[i] != 0) return VR_ERR_FORMAT;
  if (h.rootMax[0] != h.X || h.rootMax[1] != h.Y || h.rootMax[2] != h.Z) return VR_ERR_FORMAT;
  return VR_OK;
}

int kd_from_memory(const vr_kd_header* hdr, const uint8_t* distanceMap, const uint8_t* bits,
                   int64_t nbytes, KdStream* out) {
  if (!hdr || !distanceMap || !bits || !out) return VR_ERR_ARG;
  vr_kd_geom g;
  int rc = kd_check_header(*hdr, &g);
  if (rc != VR_OK) return rc;
  int64_t need = (hdr->numActiveNodes + 3) / 4;
  if (nbytes < need) return VR_ERR_FORMAT;  // longer is fine: open() of the reference over-sizes by 8
  out->hdr = *hdr;
  out->geom = g;
  out->distanceMap.assign(distanceMap, distanceMap + hdr->maxTreeDepth + 1);
  out->bits.assign(bits, bits + need);
  // unused high bits of the last byte are zero in a well-formed stream; normalise them
  int tail = (int)(hdr->numActiveNodes & 3);
  if (tail) out->bits[need - 1] &= (uint8_t)((1u << (2 * tail)) - 1);
  out->index = KdIndex();
  return VR_OK;
}

int kd_open(const char* path, KdStream* out) {
  if (!path || !out) return VR_ERR_ARG;
  FILE* f = std::fopen(path, "rb");
  if (!f) return VR_ERR_IO;  // the reference prints, waits for Enter and exit(-1)s here (:560-565)
  std::vector<uint8_t> buf;
  std::fseek(f, 0, SEEK_END);
  long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  if (sz < (long)sizeof(vr_kd_header)) { std::fclose(f); return VR_ERR_IO; }
  buf.resize((size_t)sz);
  size_t got = std::fread(buf.data(), 1, (size_t)sz, f);
  std::fclose(f);
  if (got != (size_t)sz) return VR_ERR_IO;
  vr_kd_header h;
  std::memcpy(&h, buf.data(), sizeof(h));
  if (h.maxTreeDepth < 0 || h.maxTreeDepth + 1 > VR_MAX_DEPTH) return VR_ERR_FORMAT;
  int64_t off = (int64_t)sizeof(h) + h.maxTreeDepth + 1;
  if (sz < off) return VR_ERR_IO;
  return kd_from_memory(&h, buf.data() + sizeof(h), buf.data() + off, sz - off, out);
}

int kd_save(const KdStream& s, const char* path) {
  if (!path) return VR_ERR_ARG;
  if (s.bits.empty()) return VR_ERR_STATE;  // "ERROR! No tree to save." (:527-531)
  if (!s.residentTiles.empty()) return VR_ERR_STATE;  // opened with a tile subset: not the whole stream
  FILE* f = std::fopen(path, "wb");
  if (!f) return VR_ERR_IO;
  size_t ok = std::fwrite(&s.hdr, sizeof(s.hdr), 1, f);
  ok &= std::fwrite(s.distanceMap.data(), 1, s.distanceMap.size(), f) == s.distanceMap.size();
  size_t n = (size_t)s.streamBytes();
  ok &= std::fwrite(s.bits.data(), 1, n, f) == n;
  std::fclose(f);
  return ok ? VR_OK : VR_ERR_IO;
}

// ---------------------------------------------------------------------------------------------
// Grammar walk.  A node at height h (= D0 - depth) starts at voxel path index v with v % 2^h == 0;
// after a subtree ends, the next node's height is ctz(v).  Visitor hooks:
//   node(pos, depth, h, v, code, value)  every tree node at depth <= D0 (not chain symbols)
//   leaf(v, h, value)                    every decoded box: 2^h voxels from path index v take `value` (what levelCut
//                                        writes at VolumeKdTree_recover.cpp:759-766), chains already applied
//   done_item(v, endPos)                 after a depth-D0 item (voxel + chain) was consumed
template <typename V>
static int walk(const KdStream& s, V& vis, int64_t* consumed) {
  const vr_kd_geom& g = s.geom;
  const int D0 = g.D0;
  const int64_t n = s.hdr.numActiveNodes;
  const uint8_t* bits = s.bits.data();
  const uint8_t* dm = s.distanceMap.data();
  const uint64_t total = uint64_t(1) << D0;
  int vals[VR_MAX_DEPTH + 2];
  uint64_t v = 0;
  int h = D0;
  int64_t pos = 0;
  for (;;) {
    if (pos >= n) return VR_ERR_FORMAT;
    const int depth = D0 - h;
    const int c = vr_sym(bits, pos);
    // root value is distanceMap[0] whatever its symbol says (VolumeKdTree_recover.cpp:743)
    int val = depth == 0 ? dm[0] : vr_apply(vals[h + 1], c, dm[depth]);
    vis.node(pos, depth, h, v, c, val);
    ++pos;
    if (c == 3) {
      vis.leaf(v, h, val);
      v += uint64_t(1) << h;
    } else if (h > 0) {
      vals[h] = val;
      --h;
      continue;
    } else {
      for (int k = 1; k <= VR_ADD_LEVELS; ++k) {
        if (pos >= n) return VR_ERR_FORMAT;
        int ck = vr_sym(bits, pos);
        ++pos;
        if (ck == 3) break;
        val = vr_apply(val, ck, dm[D0 + k]);
      }
      vis.leaf(v, 0, val);
      v += 1;
    }
    if (v == total) break;
    h = __builtin_ctzll(v);
  }
  if (consumed) *consumed = pos;
  return pos == n ? VR_OK : VR_ERR_FORMAT;
}

namespace {
struct CountVisitor {
  void node(int64_t, int, int, uint64_t, int, int) {}
  void leaf(uint64_t, int, int) {}
};

struct IndexVisitor {
  const vr_kd_geom& g;
  KdIndex& ix;
  uint32_t leavesPerTile;
  int64_t curBlock = -1;
  IndexVisitor(const vr_kd_geom& g_, KdIndex& ix_) : g(g_), ix(ix_) {
    leavesPerTile = 1u << (g.tileBits - g.leafBits);
  }
  void node(int64_t pos, int depth, int h, uint64_t v, int c, int val) {
    if (depth == g.Dt && c != 3) {  // live tile: open a block of leaf entries
      uint64_t t = v >> g.tileBits;
      if (curBlock >= 0) ix.blockSymEnd.push_back((uint32_t)pos);
      curBlock = (int64_t)ix.blockSymBegin.size();
      ix.tileLeafBlock[t] = (uint32_t)curBlock;
      ix.blockSymBegin.push_back((uint32_t)pos);
      ix.leafOff.resize(ix.leafOff.size() + leavesPerTile, VR_PRUNED);
      ix.leafVal.resize(ix.leafVal.size() + leavesPerTile, 0);
    }
    if (c == 3) {
      if (h >= g.tileBits) {  // constant box covering whole tiles
        uint64_t t0 = v >> g.tileBits, cnt = uint64_t(1) << (h - g.tileBits);
        for (uint64_t t = t0; t < t0 + cnt; ++t) { ix.tileLeafBlock[t] = VR_PRUNED; ix.tileVal[t] = (uint8_t)val; }
      } else if (h >= g.leafBits) {  // constant box covering whole leaf subtrees of the live tile
        uint64_t inTile = v & ((uint64_t(1) << g.tileBits) - 1);
        uint64_t l0 = inTile >> g.leafBits, cnt = uint64_t(1) << (h - g.leafBits);
        uint64_t base = (uint64_t)curBlock * leavesPerTile;
        for (uint64_t l = l0; l < l0 + cnt; ++l) { ix.leafOff[base + l] = VR_PRUNED; ix.leafVal[base + l] = (uint8_t)val; }
      }
    } else if (depth == g.Dl) {  // live leaf subtree
      uint64_t inTile = v & ((uint64_t(1) << g.tileBits) - 1);
      uint64_t base = (uint64_t)curBlock * leavesPerTile;
      ix.leafOff[base + (inTile >> g.leafBits)] = (uint32_t)pos;
      ix.leafVal[base + (inTile >> g.leafBits)] = (uint8_t)val;
    }
  }
  void leaf(uint64_t v, int h, int val) {  // exact bounds of the decoded voxels per tile (kd_cull.h)
    const uint64_t t0 = v >> g.tileBits, cnt = h > g.tileBits ? uint64_t(1) << (h - g.tileBits) : 1;
    for (uint64_t t = t0; t < t0 + cnt; ++t) {
      if (val > ix.tileMax[t]) ix.tileMax[t] = (uint8_t)val;
      if (val < ix.tileMin[t]) ix.tileMin[t] = (uint8_t)val;
    }
  }
};
}  // namespace

// longest parser segment: distance between successive live leaf offsets / block ends
static void finish_segments(const vr_kd_geom& g, KdIndex* ixp) {
  KdIndex& ix = *ixp;
  ix.numLiveTiles = (int64_t)ix.blockSymBegin.size();
  int64_t maxSeg = 0;
  const uint32_t lpt = 1u << (g.tileBits - g.leafBits);
  for (int64_t b = 0; b < ix.numLiveTiles; ++b) {
    int64_t prev = -1;
    for (uint32_t l = 0; l < lpt; ++l) {
      uint32_t o = ix.leafOff[b * lpt + l];
      if (o == VR_PRUNED) continue;
      if (prev >= 0 && (int64_t)o - prev > maxSeg) maxSeg = (int64_t)o - prev;
      prev = o;
    }
    if (prev >= 0 && (int64_t)ix.blockSymEnd[b] - prev > maxSeg) maxSeg = (int64_t)ix.blockSymEnd[b] - prev;
  }
  ix.maxSymbolsPerLeaf = maxSeg;
}

void kd_index_finish(const vr_kd_geom& g, const std::vector<uint32_t>& tileBegin, int64_t numActiveNodes, KdIndex* ix) {
  ix->blockSymBegin.clear();
  ix->blockSymEnd.clear();
  for (size_t t = 0; t < ix->tileLeafBlock.size(); ++t) {  // blocks are numbered in path order
    if (ix->tileLeafBlock[t] == VR_PRUNED) continue;
    if (!ix->blockSymBegin.empty()) ix->blockSymEnd.push_back(tileBegin[t]);
    ix->blockSymBegin.push_back(tileBegin[t]);
  }
  if (!ix->blockSymBegin.empty()) ix->blockSymEnd.push_back((uint32_t)numActiveNodes);
  finish_segments(g, ix);
  ix->built = true;
}

int kd_validate(const KdStream& s, int64_t* nodesConsumed) {
  if (s.bits.empty() || !s.residentTiles.empty()) return VR_ERR_STATE;
  CountVisitor cv;
  return walk(s, cv, nodesConsumed);
}

int kd_prepare(KdStream* s) {
  if (!s || s->bits.empty() || !s->residentTiles.empty()) return VR_ERR_STATE;
  if (s->hdr.numActiveNodes >= (int64_t)VR_PRUNED) return VR_ERR_UNSUPPORTED;
  const vr_kd_geom& g = s->geom;
  KdIndex ix;
  uint64_t nTiles = uint64_t(1) << g.Dt;
  ix.tileLeafBlock.assign(nTiles, VR_PRUNED);
  ix.tileVal.assign(nTiles, 0);
  ix.tileMax.assign(nTiles, 0);
  ix.tileMin.assign(nTiles, 255);
  IndexVisitor iv(g, ix);
  int64_t consumed = 0;
  int rc = walk(*s, iv, &consumed);
  if (rc != VR_OK) return rc;
  if (iv.curBlock >= 0) ix.blockSymEnd.push_back((uint32_t)consumed);
  finish_segments(g, &ix);
  ix.built = true;
  s->index = std::move(ix);
  return VR_OK;
}

}  // namespace vr
