// TEST HARNESS ONLY.  Runs the device decode logic of volumerenderer_b200/csrc/kd_decode_core.h on the
// CPU, CTA by CTA and thread by thread (phase 1 for every thread, then phase 2), so the parser and
// the strip->row gather can be checked against the oracle without a GPU.  Never linked into the
// product library.
#include <cstring>
#include <vector>
#include "../../volumerenderer_b200/csrc/kd_stream.h"
#include "../../volumerenderer_b200/csrc/kd_decode_core.h"
#include "../../volumerenderer_b200/csrc/kd_worklist.h"
#include "../../volumerenderer_b200/csrc/kd_parse6.h"

using namespace vr;

extern "C" {
// compress with the product's CPU compressor; returns a heap KdStream*
void* emu_compress(const uint8_t* vol, long long X, long long Y, long long Z, int tol, int epochs, long long* zeroRewrites) {
  KdStream* s = new KdStream;
  KdBuildStats st;
  if (kd_compress(vol, X, Y, Z, tol, epochs, s, &st) != VR_OK) { delete s; return nullptr; }
  if (zeroRewrites) *zeroRewrites = st.zeroRunRewrites;
  return s;
}
void* emu_from_memory(const vr_kd_header* h, const uint8_t* dm, const uint8_t* bits, long long n) {
  KdStream* s = new KdStream;
  if (kd_from_memory(h, dm, bits, n, s) != VR_OK) { delete s; return nullptr; }
  return s;
}
void emu_free(void* p) { delete static_cast<KdStream*>(p); }
void emu_meta(void* p, long long* m) {
  KdStream* s = static_cast<KdStream*>(p);
  m[0] = s->hdr.maxTreeDepth; m[1] = s->hdr.origTreeDepth; m[2] = s->hdr.X; m[3] = s->hdr.Y; m[4] = s->hdr.Z;
  m[5] = s->hdr.numActiveNodes; m[6] = (long long)s->bits.size();
}
void emu_get(void* p, uint8_t* dm, uint8_t* bits) {
  KdStream* s = static_cast<KdStream*>(p);
  std::memcpy(dm, s->distanceMap.data(), s->distanceMap.size());
  std::memcpy(bits, s->bits.data(), s->bits.size());
}
// brute force: the chain table equals the sequential saturating chain for every start value
long long emu_check_chain_lut(unsigned long long chainDist) {
  long long bad = 0;
  for (uint32_t codes = 0; codes < (1u << 14); ++codes) {
    const uint32_t e = vr_chain_lut_entry(codes, chainDist);
    for (int x = 0; x < 256; ++x) {
      int val = x, L = 7;
      for (int k = 0; k < 7; ++k) {
        int c = (codes >> (2 * k)) & 3;
        if (c == 3) { L = k + 1; break; }
        val = vr_apply(val, c, (int)((chainDist >> (8 * k)) & 0xFF));
      }
      int y = x + (int)(int8_t)(e & 0xFF);
      const int lo = (e >> 8) & 0xFF, hi = (e >> 16) & 0xFF;
      y = y < lo ? lo : y; y = y > hi ? hi : y;
      if (y != val || (int)(e >> 24) != L) ++bad;
    }
  }
  return bad;
}
int emu_validate(void* p, long long* nodes) {
  int64_t n = 0; int rc = kd_validate(*static_cast<KdStream*>(p), &n); *nodes = n; return rc;
}
// prepare + emulate the tile kernel into a slab of dims (SX,SY,SZ) at origin (ox,oy,oz)
// lazy (view-dependent) decode: the next emu_decode call drops the tiles whose byte in `needed` (2^Dt entries) is 0, as the
// device path does (kd_worklist.h with a kd_cull.h mask), and reports the (cell, max) pairs of the dropped tiles
static const uint8_t* g_needed = nullptr;
static long long* g_culled_out = nullptr;
static long long g_culled_cap = 0, g_culled_n = 0;
void emu_set_lazy(const uint8_t* needed, long long* culled_out, long long cap) { g_needed = needed; g_culled_out = culled_out; g_culled_cap = cap; g_culled_n = 0; }
long long emu_culled_count(void) { return g_culled_n; }
int emu_decode(void* p, uint8_t* slab, long long SX, long long SY, long long SZ, long long ox, long long oy, long long oz, int vec16) {
  KdStream* s = static_cast<KdStream*>(p);
  int rc = kd_prepare(s);
  if (rc != VR_OK) return rc;
  const vr_kd_geom& g = s->geom;
  std::vector<uint32_t> words((s->bits.size() + 3) / 4 + 16, 0);  // the windows read a few words past the end
  std::memcpy(words.data(), s->bits.data(), s->bits.size());
  vr_kd_job job;
  job.words = words.data();
  job.tileLeafBlock = s->index.tileLeafBlock.data();
  job.tileVal = s->index.tileVal.data();
  job.leafOff = s->index.leafOff.data();
  job.leafVal = s->index.leafVal.data();
  job.treeDist = 0; job.chainDist = 0;
  for (int h = 0; h <= g.leafBits && h < 8; ++h) job.treeDist |= (uint64_t)s->distanceMap[g.D0 - h] << (8 * h);
  for (int k = 1; k <= 7; ++k) job.chainDist |= (uint64_t)s->distanceMap[g.D0 + k] << (8 * (k - 1));
  job.dstOffset = (oz * SY + oy) * SX + ox;
  std::vector<uint32_t> lut(1u << 14);
  for (uint32_t i = 0; i < (1u << 14); ++i) lut[i] = vr_chain_lut_entry(i, job.chainDist);
  job.chainLut = lut.data();
  const uint32_t lpt = 1u << (g.tileBits - g.leafBits);
  std::vector<uint8_t> strips(lpt * ((1u << g.leafBits) + VR_LEAF_PITCH_PAD));
  const uint32_t items = vr_tile_items(g);
  // the launch's CTA list, exactly as the device path builds it (kd_worklist.h)
  std::vector<vr_kd_work> work;
  const KdIndex* idx = &s->index;
  std::vector<vr_fill_job> fills;
  const int64_t org[3] = {ox, oy, oz};
  if (g_needed) {  // cells of 32x32x16 voxels on the slab's grid, the brick on a cell boundary (as vr_decode_plan_create requires)
    if (g.tx != 32 || g.ty != 32 || g.tz != 16 || (ox & 31) || (oy & 31) || (oz & 15)) return VR_ERR_UNSUPPORTED;
    const int64_t ncx = (SX + 31) >> 5, ncy = (SY + 31) >> 5;
    const int64_t occBase = (ox >> 5) + ncx * ((oy >> 5) + ncy * (oz >> 4));
    std::vector<int32_t> tileCell(size_t(1) << g.Dt);
    for (uint32_t t = 0; t < tileCell.size(); ++t) {
      int tx0, ty0, tz0;
      vr_tile_origin(g, t, &tx0, &ty0, &tz0);
      tileCell[t] = (int32_t)((tx0 >> 5) + ncx * ((ty0 >> 5) + ncy * (tz0 >> 4)));
    }
    std::vector<vr_occ_set> culled;
    kd_build_worklist(g, 1, &idx, &job.dstOffset, SX, SX * SY, vec16 != 0, 0, &work, &fills, 2, org, 0, &occBase, tileCell.data(), g_needed,
                      &culled);
    g_culled_n = (long long)culled.size();
    for (size_t i = 0; i < culled.size() && (long long)i < g_culled_cap; ++i) { g_culled_out[2 * i] = culled[i].cell; g_culled_out[2 * i + 1] = culled[i].value; }
  } else {
    kd_build_worklist(g, 1, &idx, &job.dstOffset, SX, SX * SY, vec16 != 0, 0, &work, &fills, 4, org);
  }
  for (const vr_kd_work& wk : work) {
    const uint32_t tile = wk.id & ((1u << g.Dt) - 1u);
    int tox, toy, toz;
    vr_tile_origin(g, tile, &tox, &toy, &toz);
    uint8_t* dst = slab + job.dstOffset + ((int64_t)toz * SY + toy) * SX + tox;
    if (wk.L != VR_WORK_PARSE) {
      if (vec16) {
        for (uint32_t t = 0; t < 256; ++t) vr_group_fill(g, (int)wk.L, wk.vals, dst, SX, SX * SY, t, 256);
      } else {
        for (uint32_t it = 0; it < items; ++it) vr_tile_fill_item(g, it, (int)(wk.vals & 0xFF), dst, SX, SX * SY, false);
      }
      continue;
    }
    for (uint32_t k = 0; k < ((uint32_t)(wk.vals >> 32) & 0x7FFFFFFFu); ++k) {  // constant groups attached to this parse CTA
      const vr_fill_job& fj = fills[(size_t)(uint32_t)wk.vals + k];
      for (uint32_t t = 0; t < 256; ++t)
        vr_group_fill(g, (int)fj.L, fj.vals, slab + ((int64_t)fj.z * SY + fj.y) * SX + fj.x, SX, SX * SY, t, 256);
    }
    const uint32_t blk = job.tileLeafBlock[tile];
    std::fill(strips.begin(), strips.end(), 0xEE);
    if (g.fastPath && vec16) {  // mirrors the fast path of kd_decode_tiles_kernel
      for (uint32_t t = 0; t < 256; ++t) {
        vr_tile_phase1_v5<true, uint8_t*, VrWin32>(job, g, blk, t, strips.data());
      }
      for (uint32_t warp = 0; warp < 8; ++warp)
        for (uint32_t q = 0; q < (1u << g.fbUnit); ++q) vr_unit_phase2_item(g, warp, q, strips.data(), dst, SX, SX * SY);
      continue;
    }
    for (uint32_t t = 0; t < lpt; ++t) vr_tile_phase1(job, g, blk, t, strips.data());
    for (uint32_t it = 0; it < items; ++it) vr_tile_phase2_item(g, it, strips.data(), dst, SX, SX * SY, vec16 != 0);
  }
  return VR_OK;
}

// ---- the persistent kernel's parser (kd_parse6.h), thread by thread -------------------------------------------------------
// stageChunks: capacity of a warp's staging area in 16-byte chunks (304 on the device; a longer unit is a format error);
// needed/culled as in emu_set_lazy.  stats (4 values): parse units, longest staged segment in chunks, constant groups, 0.
}  // extern "C"
template <int ROT>
static void emu6_unit(const vr_kd_job& job, const vr_parse_item& it, const vr6::P6Geom& pg, const vr6::Dist& d, std::vector<uint8_t>& smem,
                      uint32_t lutOff, uint32_t stripOff, uint32_t stageOff, uint8_t* slab, int64_t SX, int64_t SY) {
  vr6::Mem::base() = smem.data();
  const uint32_t unit = it.tileUnit & 7u;
  const uint32_t runLen = 1u << pg.runLog, nRuns = 32u >> pg.runLog;
  // the runs' segments, one behind the other (as stage_unit of kd_decode.cu places them)
  for (uint32_t r = 0; r < nRuns; ++r) {
    const uint32_t pk = it.run[r].packed;
    if (VR_RUN_CHUNKS(pk))
      std::memcpy(smem.data() + stageOff + 16u * VR_RUN_SMEM(pk), reinterpret_cast<const uint8_t*>(job.words) + it.run[r].segOff, 16ull * VR_RUN_CHUNKS(pk));
  }
  std::fill(smem.begin() + stripOff, smem.begin() + stripOff + 32 * VR6_STRIP_PITCH, 0xEE);
  for (uint32_t lane = 0; lane < 32; ++lane) {
    const uint32_t r = lane >> pg.runLog, pk = it.run[r].packed;
    const uint32_t leaf = it.leaf0 + VR_RUN_LEAF(pk) + (lane & (runLen - 1u));
    if (VR_RUN_LEAF(pk) + (lane & (runLen - 1u)) != vr6::p6_leaf_of(pg, unit, lane)) { std::memset(slab, 0xCC, 64); return; }  // the work list and the geometry must agree
    const uint32_t off = job.leafOff[(size_t)leaf];
    const uint32_t val = job.leafVal[(size_t)leaf];
    const uint32_t strip = stripOff + lane * VR6_STRIP_PITCH;
    if (off == VR_PRUNED) { vr6::fill_leaf(strip, val); continue; }
    vr6::WinShared w;
    w.init(stageOff + 16u * VR_RUN_SMEM(pk) - it.run[r].segOff, off);  // shared address of virtual word 0 of the stream (wraps)
    vr6::parse_leaf<ROT>(w, val, d, lutOff, strip);
  }
  for (uint32_t m = 0; m < 128; ++m) {
    uint32_t sOff; int64_t gOff;
    vr6::p6_run_addr(pg, unit, m, SX, SX * SY, &sOff, &gOff);
    uint8_t* out = slab + it.dst + gOff;
    // the same run through the tensor-store path: dense box [z][y][32], placed at the unit's voxel coordinates
    const uint32_t bo = vr6::p6_box_off(pg, m);
    const int bx = (int)(bo & 31u), by = (int)((bo >> 5) & ((1u << pg.boxLy) - 1u)), bz = (int)(bo >> (5 + pg.boxLy));
    uint8_t* out2 = slab + ((int64_t)(it.uz + bz) * SY + (it.uy + by)) * SX + it.ux + bx;
    if (out2 != out) { std::memset(out, 0xDD, 16); continue; }  // the two address computations must agree
    for (uint32_t j = 0; j < 4; ++j) {
      const uint32_t a = stripOff + sOff + ((j & 1u) ? pg.xs[0] : 0u) + ((j & 2u) ? pg.xs[1] : 0u);
      std::memcpy(out + 4 * j, smem.data() + a, 4);
    }
  }
}
extern "C" {
static long long g6_stats[4];
void emu_decode6_stats(long long* out) { std::memcpy(out, g6_stats, sizeof(g6_stats)); }
int emu_decode6(void* p, uint8_t* slab, long long SX, long long SY, long long SZ, long long ox, long long oy, long long oz, int stageChunks) {
  KdStream* s = static_cast<KdStream*>(p);
  int rc = kd_prepare(s);
  if (rc != VR_OK) return rc;
  const vr_kd_geom& g = s->geom;
  vr6::P6Geom pg;
  if (!vr6::p6_make_geom(g, &pg) || (SX % 16) || (ox % 16)) return VR_ERR_UNSUPPORTED;
  std::vector<uint32_t> words((s->bits.size() + 3) / 4 + 16, 0);
  std::memcpy(words.data(), s->bits.data(), s->bits.size());
  vr_kd_job job{};
  job.words = words.data();
  job.tileLeafBlock = s->index.tileLeafBlock.data();
  job.tileVal = s->index.tileVal.data();
  job.leafOff = s->index.leafOff.data();
  job.leafVal = s->index.leafVal.data();
  uint64_t chainDist = 0;
  for (int k = 1; k <= 7; ++k) chainDist |= (uint64_t)s->distanceMap[g.D0 + k] << (8 * (k - 1));
  job.dstOffset = (oz * SY + oy) * SX + ox;
  vr6::Dist d{vr6::dist2(s->distanceMap[g.D0]), vr6::dist2(s->distanceMap[g.D0 - 1]), vr6::dist2(s->distanceMap[g.D0 - 2]),
              vr6::dist2(s->distanceMap[g.D0 - 3]), vr6::dist2(s->distanceMap[g.D0 - 4]), vr6::dist2(s->distanceMap[g.D0 - 5])};
  const uint32_t lutOff = 0, stripOff = VR6_LUT_BYTES, stageOff = VR6_LUT_BYTES + ((VR6_STRIPS_BYTES + 15u) & ~15u);
  std::vector<uint8_t> smem(stageOff + 16ull * (size_t)(stageChunks > 0 ? stageChunks : 1) + 64);
  for (uint32_t i = 0; i < VR6_LUT_ENTRIES; ++i) { const uint32_t e = vr6::chain_entry(i, chainDist); std::memcpy(smem.data() + lutOff + 4 * i, &e, 4); }
  const KdIndex* idx = &s->index;
  const int64_t org[3] = {ox, oy, oz};
  std::vector<vr_parse_item> parse;
  std::vector<vr_fill_job> fills;
  std::vector<vr_occ_set> culled;
  int64_t occBase = -1;
  std::vector<int32_t> tileCell;
  if (g_needed) {
    if (g.tx != 32 || g.ty != 32 || g.tz != 16 || (ox & 31) || (oy & 31) || (oz & 15)) return VR_ERR_UNSUPPORTED;
    const int64_t ncx = (SX + 31) >> 5, ncy = (SY + 31) >> 5;
    occBase = (ox >> 5) + ncx * ((oy >> 5) + ncy * (oz >> 4));
    tileCell.resize(size_t(1) << g.Dt);
    for (uint32_t t = 0; t < tileCell.size(); ++t) {
      int tx0, ty0, tz0;
      vr_tile_origin(g, t, &tx0, &ty0, &tz0);
      tileCell[t] = (int32_t)((tx0 >> 5) + ncx * ((ty0 >> 5) + ncy * (tz0 >> 4)));
    }
  }
  if (!kd_build_worklist6(g, 1, &idx, &job.dstOffset, SX, SX * SY, org, g_needed ? &occBase : nullptr, g_needed ? tileCell.data() : nullptr, g_needed,
                          g_needed ? &culled : nullptr, (uint32_t)(stageChunks > 0 ? stageChunks : 0), pg, &parse, nullptr, &fills))
    return VR_ERR_FORMAT;
  g_culled_n = (long long)culled.size();
  for (size_t i = 0; i < culled.size() && (long long)i < g_culled_cap; ++i) { g_culled_out[2 * i] = culled[i].cell; g_culled_out[2 * i + 1] = culled[i].value; }
  g6_stats[0] = (long long)parse.size(); g6_stats[1] = 0; g6_stats[2] = (long long)fills.size(); g6_stats[3] = 0;
  for (const vr_fill_job& f : fills)  // what the TMA unit stores: rows of (tx << L) bytes, tile j of the group takes byte j of vals
    for (int z = 0; z < g.tz; ++z)
      for (int y = 0; y < g.ty; ++y)
        for (uint32_t j = 0; j < (1u << f.L); ++j)
          std::memset(slab + ((int64_t)(f.z + z) * SY + (f.y + y)) * SX + f.x + (int64_t)j * g.tx, (int)((f.vals >> (8 * j)) & 0xFF), (size_t)g.tx);
  for (const vr_parse_item& it : parse) {
    if ((long long)it.segChunks > g6_stats[1]) g6_stats[1] = (long long)it.segChunks;
    if (pg.rot == 0) emu6_unit<0>(job, it, pg, d, smem, lutOff, stripOff, stageOff, slab, SX, SY);
    else emu6_unit<1>(job, it, pg, d, smem, lutOff, stripOff, stageOff, slab, SX, SY);
  }
  return VR_OK;
}
// brute force: the chain table of kd_parse6.h equals the sequential saturating chain for every start value
long long emu_check_chain_lut6(unsigned long long chainDist) {
  long long bad = 0;
  for (uint32_t codes = 0; codes < (1u << 14); ++codes) {
    const uint32_t e = vr6::chain_entry(codes, chainDist);
    for (int x = 0; x < 256; ++x) {
      int val = x, L = 7;
      for (int k = 0; k < 7; ++k) {
        int c = (codes >> (2 * k)) & 3;
        if (c == 3) { L = k + 1; break; }
        val = vr_apply(val, c, (int)((chainDist >> (8 * k)) & 0xFF));
      }
      int y = x + ((int)e >> 24);
      const int lo = (e >> 8) & 0xFF, hi = (e >> 16) & 0xFF;
      y = y < lo ? lo : y; y = y > hi ? hi : y;
      if (y != val || (int)(e & 0xFF) != 2 + 2 * L) ++bad;
    }
  }
  return bad;
}
}
