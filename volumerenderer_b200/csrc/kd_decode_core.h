// kd_decode_core.h -- the per-thread pieces of the GPU stream decoder, written as host/device
// code so the same source is exercised thread-by-thread on the CPU by tests/emu (test harness only)
// and compiled into kd_decode.cu for the device.
//
// What it decodes: the preorder 2-bit stream consumed by the reference's
// VolumeKdtree::levelCut (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:726-835):
//   node above origTreeDepth : 3 = constant box, else two children follow (left then right)
//   node at origTreeDepth    : 3 = voxel equals parent, else a refinement chain follows
//   chain (<= 7 symbols)     : 3 terminates, else saturating +/- distanceMap[depth]
// One parser thread owns one "leaf subtree" (root at depth Dl, 2^leafBits voxels) whose start
// offset and root value come from the side index (kd_stream.h).  It writes its voxels in PATH
// order into a private, padded shared-memory strip; a second phase gathers strips into x-fastest
// 16-byte row segments of the destination slab.
#pragma once
#include "kd_common.h"

#if defined(__CUDA_ARCH__)
#define VR_CTZ32(x) (__ffs((int)(x)) - 1)
#define VR_LDG32(p) __ldg(p)
#else
#define VR_CTZ32(x) __builtin_ctz((unsigned)(x))
#define VR_LDG32(p) (*(p))
#endif

#ifndef VR_PREFETCH_LINES
#define VR_PREFETCH_LINES 2  // 128-byte lines requested ahead per leaf subtree (kd_decode.cu, parse path)
#endif
#define VR_LEAF_PITCH_PAD 4  // strip pitch = 2^leafBits + 4 bytes (keeps strips on distinct banks)

// Per-brick decode job, resident in device memory.
typedef struct vr_kd_job {
  const uint32_t* words;         // packed stream, 4-byte aligned, padded by >= 16 readable bytes
  const uint32_t* tileLeafBlock; // [2^Dt]
  const uint8_t* tileVal;        // [2^Dt]
  const uint32_t* leafOff;       // [liveTiles * leavesPerTile]
  const uint8_t* leafVal;        // [liveTiles * leavesPerTile]
  uint64_t treeDist;             // byte h = distanceMap[D0 - h] for h = 0..leafBits (levels inside a leaf subtree)
  uint64_t chainDist;            // byte k-1 = distanceMap[D0 + k] for k = 1..7
  int64_t dstOffset;             // voxel offset of the brick's (0,0,0) inside the destination slab
  const uint32_t* chainLut;      // [16384] refinement-chain table for this stream's chainDist (vr_chain_lut_entry)
  int64_t occBase;               // occupancy cell of the brick's (0,0,0) in the destination's grid, < 0 = tiles are not cells
} vr_kd_job;

// One CTA of the decode launch (host: kd_worklist.h).  id = brick << Dt | tile.  L = VR_WORK_PARSE: the tile is
// parsed; otherwise the CTA fills the 2^L x-neighbouring constant tiles starting at `tile`, byte j of vals =
// value of the j-th of them.
#define VR_WORK_PARSE 0xFFFFFFFFu
typedef struct vr_kd_work {
  uint32_t id;
  uint32_t L;
  uint64_t vals;
} vr_kd_work;

// One group of 2^L x-neighbouring constant tiles stored by the TMA fill kernel (kd_fill.h); 32 bytes, read as two
// 16-byte words.
typedef struct vr_fill_job {
  int32_t x, y, z;   // voxel coordinates of the group's origin in the destination volume
  uint32_t L;
  uint64_t vals;     // byte j = value of the j-th tile
  int64_t occCell;   // occupancy cell of the first tile (members follow at +1), < 0 = none
} vr_fill_job;

// One unit of the persistent kernel's parse queue (kd_decode.cu: kd_decode6_kernel; host: kd_build_worklist6 fills the
// bookkeeping fields, the C ABI layer the device addresses).  A unit = 32 leaf boxes of a tile that together span the tile's
// whole x extent (32 x 8 x 8 voxels = 2 048, the work of one warp): its voxel box then leaves as 64 rows of 32 bytes --
// whole sectors.  (Units of 32 CONSECUTIVE leaf boxes are 16 voxels wide: 128 half-sector rows per box, which cost the
// TMA unit twice the time and wrote every sector in two halves: profiles/r2_step_final.md.)  In stream order the 32 leaf
// boxes are 2 or 4 runs of consecutive ones (vr6::P6Geom), each with its own staged stream segment.  96 bytes,
// self-contained: the warp needs nothing else to decode the unit.
typedef struct vr_parse_run {
  uint32_t segOff;  // byte offset inside the arena of the first 16-byte chunk of the run's stream segment
  uint32_t packed;  // bits 0-8 chunks to stage (0: no live leaf in the run), 9-16 first leaf of the run relative to the tile's
                    // first, 17-25 where the run's chunks go in the staging region (in chunks)
} vr_parse_run;
#define VR_RUN_PACK(chunks, leafRel, smemChunk) ((uint32_t)(chunks) | ((uint32_t)(leafRel) << 9) | ((uint32_t)(smemChunk) << 17))
#define VR_RUN_CHUNKS(p) ((p) & 0x1FFu)
#define VR_RUN_LEAF(p) (((p) >> 9) & 0xFFu)
#define VR_RUN_SMEM(p) (((p) >> 17) & 0x1FFu)
typedef struct vr_parse_item {
  // --- what the lane that requests the staging copies reads (32 bytes) ---
  const uint8_t* arena;  // device image of the prepared stream (the packed stream sits at offset 0)
  uint32_t segChunks;    // chunks to stage in shared memory, all runs together (0: no live leaf, nothing to parse)
  uint32_t pad0;
  uint32_t leafOffOff;   // byte offset inside the arena of the TILE's 256 leafOff entries ...
  uint32_t leafValOff;   // ... and of its 256 leafVal entries
  uint32_t leaf0;        // index of the tile's first leaf entry = tileLeafBlock[tile] * 256
  uint32_t tileUnit;     // tile << 3 | unit (unit = which eighth of the tile; also the warp slot of the occupancy cell)
  // --- what every lane reads ---
  uint64_t treeDist;     // byte h = distanceMap[D0 - h], h = 0..6
  int64_t dst;           // voxel offset of the TILE origin inside the destination slab
  int32_t ux, uy, uz;    // voxel coordinates of the UNIT's box in the destination (tensor-store coordinates)
  int32_t occCell;       // occupancy cell of the tile, < 0 = none
  // --- the runs (lanes 0..3 read one each) ---
  vr_parse_run run[4];
} vr_parse_item;

// A tile the lazy decode leaves out (kd_cull.h): its occupancy cell and the exact maximum of its voxels.
typedef struct vr_occ_set {
  int64_t cell;
  int64_t value;
} vr_occ_set;

// Sequential 2-bit reader over 32-bit words; always holds >= 32 valid bits after a refill.
struct VrBitReader {
  const uint32_t* w;
  uint64_t win;
  int have;
  uint32_t idx;
  VR_HD void init(const uint32_t* words, uint32_t sym) {
    w = words;
    idx = sym >> 4;
    int sh = (int)(sym & 15) * 2;
    win = (uint64_t)VR_LDG32(w + idx) >> sh;
    have = 32 - sh;
    ++idx;
    win |= (uint64_t)VR_LDG32(w + idx) << have;
    have += 32;
    ++idx;
  }
  VR_HD int take() {
    int c = (int)(win & 3);
    win >>= 2;
    have -= 2;
    if (have <= 32) {
      win |= (uint64_t)VR_LDG32(w + idx) << have;
      have += 32;
      ++idx;
    }
    return c;
  }
};

// Path-ordered byte sink with 4-byte packing.  Voxels arrive in increasing path index.
struct VrStripSink {
  uint8_t* buf;  // 4-byte aligned
  uint32_t acc;
  VR_HD void init(uint8_t* b) { buf = b; acc = 0; }
  VR_HD void put(uint32_t v, int val) {
    acc |= (uint32_t)val << (8 * (v & 3));
    if ((v & 3) == 3) {
      *reinterpret_cast<uint32_t*>(buf + (v & ~3u)) = acc;
      acc = 0;
    }
  }
  VR_HD void fill(uint32_t v, uint32_t cnt, int val) {
    if (cnt >= 4) {  // v is a multiple of cnt, hence of 4
      uint32_t word = (uint32_t)val * 0x01010101u;
      for (uint32_t i = 0; i < cnt; i += 4) *reinterpret_cast<uint32_t*>(buf + v + i) = word;
    } else {
      for (uint32_t i = 0; i < cnt; ++i) put(v + i, val);
    }
  }
  VR_HD void finish(uint32_t total) {  // strips shorter than one word
    if (total < 4)
      for (uint32_t i = 0; i < total; ++i) buf[i] = (uint8_t)(acc >> (8 * i));
  }
};

// Refinement chain below a depth-D0 node whose own value is `val` (VolumeKdTree_recover.cpp:806-833
// for depth >= origTreeDepth): up to 7 symbols, 3 terminates, no terminator after the 7th.
template <typename Reader>
VR_HD int vr_decode_chain(Reader& rd, int val, uint64_t chainDist) {
#pragma unroll 1
  for (int k = 0; k < VR_ADD_LEVELS; ++k) {
    int c = rd.take();
    if (c == 3) break;
    val = vr_apply(val, c, (int)((chainDist >> (8 * k)) & 0xFF));
  }
  return val;
}

// Decodes the leaf subtree whose root symbol the reader is positioned on.  The root is known to be
// an internal node (symbol != 3) with decoded value `ownVal`.
template <typename Reader, typename Sink>
VR_HD void vr_decode_leaf(Reader& rd, int leafBits, int ownVal, uint64_t treeDist, uint64_t chainDist,
                          Sink& sink) {
  const uint32_t total = 1u << leafBits;
  (void)rd.take();  // the root's own symbol; its effect is already in ownVal
  if (leafBits == 0) {
    sink.put(0, vr_decode_chain(rd, ownVal, chainDist));
    sink.finish(total);
    return;
  }
  uint64_t vals = (uint64_t)ownVal << (8 * leafBits);  // byte h = value of the path node at height h
  int h = leafBits - 1;
  uint32_t v = 0;
  for (;;) {
    const int c = rd.take();
    const int pv = (int)((vals >> (8 * (h + 1))) & 0xFF);
    if (c == 3) {
      sink.fill(v, 1u << h, pv);
      v += 1u << h;
    } else {
      const int val = vr_apply(pv, c, (int)((treeDist >> (8 * h)) & 0xFF));
      if (h > 0) {
        vals = (vals & ~((uint64_t)0xFF << (8 * h))) | ((uint64_t)val << (8 * h));
        --h;
        continue;
      }
      sink.put(v, vr_decode_chain(rd, val, chainDist));
      v += 1;
    }
    if (v == total) break;
    h = VR_CTZ32(v);
  }
  sink.finish(total);
}

// ---- phase 1: one parser thread ----------------------------------------------------------------
// strips: shared memory, leavesPerTile strips of pitch (2^leafBits + VR_LEAF_PITCH_PAD) bytes.
VR_HD void vr_tile_phase1(const vr_kd_job& job, const vr_kd_geom& g, uint32_t blk, uint32_t leaf,
                          uint8_t* strips) {
  const uint32_t lpt = 1u << (g.tileBits - g.leafBits);
  const uint32_t pitch = (1u << g.leafBits) + VR_LEAF_PITCH_PAD;
  const uint32_t off = job.leafOff[(size_t)blk * lpt + leaf];
  const int val = job.leafVal[(size_t)blk * lpt + leaf];
  VrStripSink sink;
  sink.init(strips + (size_t)leaf * pitch);
  if (off == VR_PRUNED) {
    sink.fill(0, 1u << g.leafBits, val);
    sink.finish(1u << g.leafBits);
    return;
  }
  VrBitReader rd;
  rd.init(job.words, off);
  vr_decode_leaf(rd, g.leafBits, val, job.treeDist, job.chainDist, sink);
}

// Tile-local path index of tile-local voxel (x,y,z).
VR_HD uint32_t vr_path_of(const vr_kd_geom& g, int x, int y, int z) {
  uint32_t p = 0;
  for (int j = 0; (1 << j) < g.tx; ++j) p |= (uint32_t)((x >> j) & 1) << g.pathBit[0][j];
  for (int j = 0; (1 << j) < g.ty; ++j) p |= (uint32_t)((y >> j) & 1) << g.pathBit[1][j];
  for (int j = 0; (1 << j) < g.tz; ++j) p |= (uint32_t)((z >> j) & 1) << g.pathBit[2][j];
  return p;
}

// Tile origin (brick-local voxel coordinates) of tile slot t (path prefix of Dt bits).
VR_HD void vr_tile_origin(const vr_kd_geom& g, uint32_t t, int* ox, int* oy, int* oz) {
  int o0 = 0, o1 = 0, o2 = 0;
  for (int d = 0; d < g.Dt; ++d)
    if ((t >> (g.Dt - 1 - d)) & 1) {
      const int s = g.splitDim[d], c = g.childOff[d];
      o0 += s == 0 ? c : 0; o1 += s == 1 ? c : 0; o2 += s == 2 ? c : 0;
    }
  *ox = o0; *oy = o1; *oz = o2;
}

VR_HD uint32_t vr_strip_addr(const vr_kd_geom& g, uint32_t path) {
  const uint32_t pitch = (1u << g.leafBits) + VR_LEAF_PITCH_PAD;
  return (path >> g.leafBits) * pitch + (path & ((1u << g.leafBits) - 1));
}

// Strip-address contribution of tile path bit b (strip address is linear over path bits).
VR_HD uint32_t vr_bit_addr(const vr_kd_geom& g, int b) {
  const uint32_t pitch = (1u << g.leafBits) + VR_LEAF_PITCH_PAD;
  return b < g.leafBits ? (1u << b) : (pitch << (b - g.leafBits));
}

// ---- phase 2: gather strips into x-fastest rows of the destination ------------------------------
// Work item `item` of a tile: a run of `seg` (= min(16, tx)) voxels along x at tile-local (x0,y,z).
// dst points at the tile origin inside the slab; rowPitch / slicePitch are the slab's y and z strides.
// When vec16 is set (seg == 16 and every address 16-byte aligned) one 128-bit store is issued.
VR_HD void vr_tile_phase2_item(const vr_kd_geom& g, uint32_t item, const uint8_t* strips, uint8_t* dst,
                               int64_t rowPitch, int64_t slicePitch, bool vec16) {
  // tile extents are powers of two: shifts and masks instead of integer division
  const int seg = g.tx < 16 ? g.tx : 16;
  const int lsr = g.tx < 16 ? 0 : VR_CTZ32((uint32_t)g.tx) - 4, lty = VR_CTZ32((uint32_t)g.ty);
  const int x0 = (int)(item & ((1u << lsr) - 1u)) * seg;
  const int y = (int)((item >> lsr) & (uint32_t)(g.ty - 1));
  const int z = (int)(item >> (lsr + lty));
  uint32_t base = 0;
  for (int j = 4; (1 << j) < g.tx; ++j) base += ((x0 >> j) & 1) ? vr_bit_addr(g, g.pathBit[0][j]) : 0u;
  for (int j = 0; (1 << j) < g.ty; ++j) base += ((y >> j) & 1) ? vr_bit_addr(g, g.pathBit[1][j]) : 0u;
  for (int j = 0; (1 << j) < g.tz; ++j) base += ((z >> j) & 1) ? vr_bit_addr(g, g.pathBit[2][j]) : 0u;
  uint32_t a[4] = {0, 0, 0, 0};
  for (int j = 0; j < 4 && (1 << j) < g.tx; ++j) a[j] = vr_bit_addr(g, g.pathBit[0][j]);
  uint8_t* out = dst + (int64_t)z * slicePitch + (int64_t)y * rowPitch + x0;
  if (vec16) {
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t r = 0;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int xi = q * 4 + i;
        const uint32_t ad = base + ((xi & 1) ? a[0] : 0u) + ((xi & 2) ? a[1] : 0u) + ((xi & 4) ? a[2] : 0u) +
                            ((xi & 8) ? a[3] : 0u);
        r |= (uint32_t)strips[ad] << (8 * i);
      }
      w[q] = r;
    }
#if defined(__CUDA_ARCH__)
    *reinterpret_cast<uint4*>(out) = make_uint4(w[0], w[1], w[2], w[3]);
#else
    for (int q = 0; q < 4; ++q)
      for (int i = 0; i < 4; ++i) out[q * 4 + i] = (uint8_t)(w[q] >> (8 * i));
#endif
  } else {
    for (int xi = 0; xi < seg; ++xi) {
      const uint32_t ad = base + ((xi & 1) ? a[0] : 0u) + ((xi & 2) ? a[1] : 0u) + ((xi & 4) ? a[2] : 0u) +
                          ((xi & 8) ? a[3] : 0u);
      out[xi] = strips[ad];
    }
  }
}

// Constant tile: same work items, value `val`.
VR_HD void vr_tile_fill_item(const vr_kd_geom& g, uint32_t item, int val, uint8_t* dst, int64_t rowPitch,
                             int64_t slicePitch, bool vec16) {
  // tile extents are powers of two: shifts and masks instead of integer division
  const int seg = g.tx < 16 ? g.tx : 16;
  const int lsr = g.tx < 16 ? 0 : VR_CTZ32((uint32_t)g.tx) - 4, lty = VR_CTZ32((uint32_t)g.ty);
  const int x0 = (int)(item & ((1u << lsr) - 1u)) * seg;
  const int y = (int)((item >> lsr) & (uint32_t)(g.ty - 1));
  const int z = (int)(item >> (lsr + lty));
  uint8_t* out = dst + (int64_t)z * slicePitch + (int64_t)y * rowPitch + x0;
  if (vec16) {
#if defined(__CUDA_ARCH__)
    const uint32_t w = (uint32_t)val * 0x01010101u;
    *reinterpret_cast<uint4*>(out) = make_uint4(w, w, w, w);
#else
    for (int i = 0; i < 16; ++i) out[i] = (uint8_t)val;
#endif
  } else {
    for (int xi = 0; xi < seg; ++xi) out[xi] = (uint8_t)val;
  }
}

// ---- grouped constant fills -----------------------------------------------------------------------
// A tile row is only tx (32) bytes wide, and a fill made of 32-byte pieces runs at ~4.2 TB/s on B200 where
// full 128-byte lines reach 6.3-7 TB/s (tools/fill_probe.cu: every warp store then touches 16 L2 lines
// instead of 4).  So constant tiles that are x-neighbours inside their brick are filled together: the tile
// with the lowest x of the largest aligned all-constant group (1, 2, 4 or 8 tiles) writes rows of
// n * tx bytes for the whole group, the other members write nothing.
#define VR_FILL_GROUP_MAX_ROW 4096  // bytes; rows are cut into 16-byte pieces, one per thread of a 256-thread CTA

// Returns log2 of the group size if `tile` (a constant tile) leads its group, -1 if another tile fills it.
// Byte j of *vals = value of the j-th tile of the group (in x order).
VR_HD int vr_fill_group_of(const vr_kd_geom& g, const vr_kd_job& job, uint32_t tile, uint64_t* vals) {
  int nb = g.nxr;
  while (nb > 0 && ((int64_t)g.tx << nb) > VR_FILL_GROUP_MAX_ROW) --nb;
  if (g.tx < 16) nb = 0;
  uint32_t xi = 0, rest = tile;
  for (int j = 0; j < nb; ++j) {
    xi |= ((tile >> g.xrBit[j]) & 1u) << j;
    rest &= ~(1u << g.xrBit[j]);
  }
  uint32_t m = 0;
  uint64_t v = 0;
#pragma unroll
  for (uint32_t i = 0; i < 8; ++i) {
    if (i < (1u << nb)) {
      uint32_t t = rest;
      for (int j = 0; j < nb; ++j) t |= ((i >> j) & 1u) << g.xrBit[j];
      if (job.tileLeafBlock[t] == VR_PRUNED) {
        m |= 1u << i;
        v |= (uint64_t)job.tileVal[t] << (8 * i);
      }
    }
  }
  int L = 0;
  for (int l = nb; l >= 1; --l) {
    const uint32_t lo = xi & ~((1u << l) - 1u);
    const uint32_t gm = ((1u << (1u << l)) - 1u) << lo;
    if ((m & gm) == gm) { L = l; break; }
  }
  if (xi & ((1u << L) - 1u)) return -1;
  *vals = v >> (8 * xi);
  return L;
}

VR_HD void vr_store16_const(uint8_t* out, int val) {
#if defined(__CUDA_ARCH__)
  const uint32_t w = (uint32_t)val * 0x01010101u;
#if !defined(VR_FILL_PLAIN_STORES)  // streaming stores (st.global.cs, evict-first in L2): measured 0.256 vs 0.262 ms per slab
  __stcs(reinterpret_cast<uint4*>(out), make_uint4(w, w, w, w));
#else
  *reinterpret_cast<uint4*>(out) = make_uint4(w, w, w, w);
#endif
#else
  for (int i = 0; i < 16; ++i) out[i] = (uint8_t)val;
#endif
}

// Thread `tid` of `nthreads` (a power of two >= pieces per row) fills its share of a group of 2^L tiles whose
// first tile has its origin at dst (16-byte aligned destination).  A thread keeps one 16-byte column of the
// group, so its value and x offset are fixed and a store costs one address update.
VR_HD void vr_group_fill(const vr_kd_geom& g, int L, uint64_t vals, uint8_t* dst, int64_t rowPitch, int64_t slicePitch,
                         uint32_t tid, uint32_t nthreads) {
  const int lpt = VR_CTZ32((uint32_t)g.tx) - 4;  // log2 pieces per tile row
  const int lR = lpt + L;                        // log2 pieces per group row
  const uint32_t piece = tid & ((1u << lR) - 1u);
  const uint32_t rstep = nthreads >> lR;
  const int val = (int)((vals >> (8 * (piece >> lpt))) & 0xFF);
  const int lty = VR_CTZ32((uint32_t)g.ty);
  const uint32_t row0 = tid >> lR;
  uint8_t* base = dst + (size_t)piece * 16;
  if (rstep <= (uint32_t)g.ty) {  // a pass covers rstep rows of one slice: walk y inside z with a fixed pointer step
    const int64_t ystep = (int64_t)rstep * rowPitch;
    for (int z = 0; z < g.tz; ++z) {
      uint8_t* q = base + (int64_t)z * slicePitch + (int64_t)row0 * rowPitch;
      for (uint32_t y = row0; y < (uint32_t)g.ty; y += rstep, q += ystep) vr_store16_const(q, val);
    }
  } else {  // a pass covers whole slices: this thread keeps its y
    const uint32_t zstep = rstep >> lty;
    uint8_t* q = base + (int64_t)(row0 >> lty) * slicePitch + (int64_t)(row0 & (uint32_t)(g.ty - 1)) * rowPitch;
    for (uint32_t z = row0 >> lty; z < (uint32_t)g.tz; z += zstep, q += (int64_t)zstep * slicePitch) vr_store16_const(q, val);
  }
}

VR_HD uint32_t vr_tile_items(const vr_kd_geom& g) {
  return (uint32_t)((g.tx < 16 ? 1 : (g.tx >> 4)) * g.ty * g.tz);
}

// =================================================================================================
// v2 fast path (full tiles of 256 leaf subtrees, 16-byte aligned destination)
// =================================================================================================

// Refinement-chain table.  A chain is <= 7 symbols (14 bits, LSB first); its effect on the voxel value
// is a composition of saturating +/- steps, which is always of the form x -> min(max(x + A, lo), hi).
// Entry = A (int8) | lo << 8 | hi << 16 | symbols consumed << 24.
VR_HD uint32_t vr_chain_lut_entry(uint32_t codes14, uint64_t chainDist) {
  int A = 0, lo = 0, hi = 255, L = VR_ADD_LEVELS;
  for (int k = 0; k < VR_ADD_LEVELS; ++k) {
    const int c = (int)((codes14 >> (2 * k)) & 3);
    if (c == 3) { L = k + 1; break; }
    const int d = (int)((chainDist >> (8 * k)) & 0xFF);
    const int a = c == 1 ? d : (c == 2 ? -d : 0);
    A += a;
    lo += a; lo = lo < 0 ? 0 : (lo > 255 ? 255 : lo);
    hi += a; hi = hi < 0 ? 0 : (hi > 255 ? 255 : hi);
  }
  return (uint32_t)(A & 0xFF) | ((uint32_t)lo << 8) | ((uint32_t)hi << 16) | ((uint32_t)L << 24);
}

// Phase 2 over a "unit" = the 2^unitBits voxels whose tile path index shares the top bits `unit`
// (a warp's 32 leaf subtrees when unitBits = leafBits + 5).  Item q = one 16-byte x-run; the other
// unitBits-4 free path bits are enumerated by the bits of q in increasing path-bit order.
// Requires the 4 low x bits to map to path bits < unitBits (checked on the host).
VR_HD uint32_t vr_bmax4(uint32_t a, uint32_t b) {  // per-byte unsigned max
#if defined(__CUDA_ARCH__)
  return __vmaxu4(a, b);
#else
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) { uint32_t x = (a >> (8 * i)) & 0xFF, y = (b >> (8 * i)) & 0xFF; r |= (x > y ? x : y) << (8 * i); }
  return r;
#endif
}
// Returns the per-byte maximum of the 16 voxels written (for the occupancy bound of the tile).
VR_HD uint32_t vr_unit_phase2_item(const vr_kd_geom& g, uint32_t unit, uint32_t q, const uint8_t* strips,
                                   uint8_t* dstTile, int64_t rowPitch, int64_t slicePitch) {
  const uint32_t m = q | (unit << g.fbUnit);  // one bit per free path bit, increasing path-bit order
  uint32_t sAddr = 0;
  int64_t gOff = 0;
#pragma unroll
  for (int k = 0; k < VR_TILE_BITS_MAX - 4; ++k) {
    if ((m >> k) & 1u) {
      sAddr += g.fbAddr[k];
      const int64_t st = g.fbAxis[k] == 0 ? (int64_t)1 : (g.fbAxis[k] == 1 ? rowPitch : slicePitch);
      gOff += st << g.fbCb[k];
    }
  }
  const uint32_t a[4] = {g.xa[0], g.xa[1], g.xa[2], g.xa[3]};
  uint32_t w[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int xi = k * 4 + i;
      const uint32_t ad = sAddr + ((xi & 1) ? a[0] : 0u) + ((xi & 2) ? a[1] : 0u) + ((xi & 4) ? a[2] : 0u) + ((xi & 8) ? a[3] : 0u);
      r |= (uint32_t)strips[ad] << (8 * i);
    }
    w[k] = r;
  }
  uint8_t* out = dstTile + gOff;
#if defined(__CUDA_ARCH__)
  *reinterpret_cast<uint4*>(out) = make_uint4(w[0], w[1], w[2], w[3]);
#else
  for (int k = 0; k < 4; ++k)
    for (int i = 0; i < 4; ++i) out[k * 4 + i] = (uint8_t)(w[k] >> (8 * i));
#endif
  return vr_bmax4(vr_bmax4(w[0], w[1]), vr_bmax4(w[2], w[3]));
}


// =================================================================================================
// v3 parser: 32-bit funnel-shift bit window, branch-free value update, shared-space strip stores,
// left/right voxel items handled as a pair.
// =================================================================================================
// Strip address type SA: uint32_t = shared-window byte address (device), uint8_t* = plain pointer (host emulation).
VR_HD void vr_st32(uint32_t addr, uint32_t val) {
#if defined(__CUDA_ARCH__)
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(val) : "memory");
#else
  (void)addr; (void)val;
#endif
}
VR_HD void vr_st32(uint8_t* addr, uint32_t val) { *reinterpret_cast<uint32_t*>(addr) = val; }
#define VR_ST32(addr, val) vr_st32((addr), (val))
#if defined(__CUDA_ARCH__)
#define VR_FUNNEL_R(lo, hi, sh) __funnelshift_r((lo), (hi), (sh))
#else
#define VR_FUNNEL_R(lo, hi, sh) ((uint32_t)(((((uint64_t)(hi)) << 32) | (uint64_t)(lo)) >> ((sh) & 31)))
#endif

struct VrWin32 {
  const uint32_t* w;  // next word to load
  uint32_t cur, nxt;
  int bo;             // bit offset into cur, 0..31
  VR_HD void init(const uint32_t* words, uint32_t sym) {
    w = words + (sym >> 4);
    cur = VR_LDG32(w);
    nxt = VR_LDG32(w + 1);
    w += 2;
    bo = (int)(sym & 15) * 2;
  }
  VR_HD uint32_t peek() const { return VR_FUNNEL_R(cur, nxt, bo); }  // 32 valid bits
  VR_HD void skip(int nbits) {
    bo += nbits;
    if (bo >= 32) {
      cur = nxt;
      nxt = VR_LDG32(w);
      ++w;
      bo -= 32;
    }
  }
};

// 1 = saturating add, 2 = saturating subtract, 0/3 = unchanged; no branches.
VR_HD int vr_apply_bf(int parent, int code, int dist) {
  const int s = (code & 1) - (code >> 1);
  int v = parent + s * dist;
  v = v < 0 ? 0 : v;
  return v > 255 ? 255 : v;
}

// v5: the three lowest levels of a leaf subtree (7 inner nodes, 8 voxels: an "octet") as straight-line code.
// v4 still pays ~31 instructions per inner node (dynamic height: 64-bit shifts for the value stack and the level
// distance, loop control) and ~12 per voxel pair for the pop.  Below height 3 the shape of the tree is fixed, so
// node values live in registers, level distances and strip byte positions are compile-time constants, and only
// the symbols decide: 3 at an inner node skips its subtree (the lanes that do not skip it keep the warp busy).
// Heights >= 3 keep v4's dynamic loop (at most 7 such nodes per leaf subtree).
template <int H, int POS, bool CHAIN_LEN_ALU, typename SA, typename WIN>
struct VrSub {
  static VR_HD void run(WIN& rd, int pv, uint64_t treeDist, const uint32_t* lut, SA base, uint32_t& acc) {
    const uint32_t bits = rd.peek();
    const int c = (int)(bits & 3);
    rd.skip(2);
    if (c == 3) {  // constant box of 2^H voxels, H = 1 or 2
      if (H == 2) {
        VR_ST32(base + POS, (uint32_t)pv * 0x01010101u);
      } else {
        acc |= ((uint32_t)pv * 0x0101u) << (8 * (POS & 3));
        if ((POS & 3) == 2) { VR_ST32(base + (POS & ~3), acc); acc = 0; }
      }
    } else {
      const int val = vr_apply_bf(pv, c, (int)((treeDist >> (8 * H)) & 0xFF));
      VrSub<H - 1, POS, CHAIN_LEN_ALU, SA, WIN>::run(rd, val, treeDist, lut, base, acc);
      VrSub<H - 1, POS + (1 << (H - 1)), CHAIN_LEN_ALU, SA, WIN>::run(rd, val, treeDist, lut, base, acc);
    }
  }
};
template <int POS, bool CHAIN_LEN_ALU, typename SA, typename WIN>
struct VrSub<0, POS, CHAIN_LEN_ALU, SA, WIN> {  // one voxel with its refinement chain
  static VR_HD void run(WIN& rd, int pv, uint64_t treeDist, const uint32_t* lut, SA base, uint32_t& acc) {
    const uint32_t bits = rd.peek();
    const int c = (int)(bits & 3);
    const uint32_t e = VR_LDG32(lut + ((bits >> 2) & 0x3FFFu));
    int L;
    if (CHAIN_LEN_ALU) {
      const uint32_t m = ((bits >> 2) & (bits >> 3) & 0x1555u) | 0x1000u;
      L = (VR_CTZ32(m) >> 1) + 1;
    } else {
      L = (int)(e >> 24);
    }
    const bool leaf = c == 3;
    rd.skip(leaf ? 2 : 2 + 2 * L);
    int val = vr_apply_bf(pv, c, (int)(treeDist & 0xFF)) + (int)(int8_t)(e & 0xFF);
    const int lo = (int)((e >> 8) & 0xFF), hi = (int)((e >> 16) & 0xFF);
    val = val < lo ? lo : val;
    val = val > hi ? hi : val;
    acc |= (uint32_t)(leaf ? pv : val) << (8 * (POS & 3));
    if ((POS & 3) == 3) { VR_ST32(base + (POS & ~3), acc); acc = 0; }
  }
};

template <bool CHAIN_LEN_ALU, typename SA, typename WIN>
VR_HD void vr_decode_leaf_v5(WIN& rd, int leafBits, int ownVal, uint64_t treeDist, const uint32_t* lut, SA base) {
  const uint32_t total = 1u << leafBits;  // leafBits >= 4 on this path
  rd.skip(2);  // the root's own symbol (known internal); its effect is already in ownVal
  uint64_t stack = (uint64_t)ownVal;
  int h = leafBits - 1;
  uint32_t v = 0;
  for (;;) {
    uint32_t bits = rd.peek();
    int c = (int)(bits & 3);
    while (h > 3 && c != 3) {  // inner nodes above the octets
      const int val = vr_apply_bf((int)(stack & 0xFF), c, (int)((treeDist >> (8 * h)) & 0xFF));
      stack = (stack << 8) | (uint64_t)val;
      --h;
      rd.skip(2);
      bits = rd.peek();
      c = (int)(bits & 3);
    }
    const int pv = (int)(stack & 0xFF);
    rd.skip(2);
    if (c == 3) {  // constant box of 2^h >= 8 voxels
      const uint32_t word = (uint32_t)pv * 0x01010101u;
      for (uint32_t i = 0; i < (1u << h); i += 4) VR_ST32(base + v + i, word);
      v += 1u << h;
    } else {  // h == 3: an octet
      const int val = vr_apply_bf(pv, c, (int)((treeDist >> 24) & 0xFF));
      uint32_t acc = 0;
      VrSub<2, 0, CHAIN_LEN_ALU, SA, WIN>::run(rd, val, treeDist, lut, base + v, acc);
      VrSub<2, 4, CHAIN_LEN_ALU, SA, WIN>::run(rd, val, treeDist, lut, base + v, acc);
      v += 8;
    }
    if (v == total) break;
    const int nh = VR_CTZ32(v);
    stack >>= 8 * (nh - h);  // pop back to the parent of the next node
    h = nh;
  }
}

template <bool CHAIN_LEN_ALU, typename SA, typename WIN = VrWin32>
VR_HD void vr_tile_phase1_v5(const vr_kd_job& job, const vr_kd_geom& g, uint32_t blk, uint32_t leaf, SA strips) {
  const uint32_t lpt = 1u << (g.tileBits - g.leafBits);
  const uint32_t pitch = (1u << g.leafBits) + VR_LEAF_PITCH_PAD;
  const uint32_t off = job.leafOff[(size_t)blk * lpt + leaf];
  const int val = job.leafVal[(size_t)blk * lpt + leaf];
  const SA base = strips + leaf * pitch;
  if (off == VR_PRUNED) {
    const uint32_t word = (uint32_t)val * 0x01010101u;
    for (uint32_t i = 0; i < (1u << g.leafBits); i += 4) VR_ST32(base + i, word);
    return;
  }
  WIN rd;
  rd.init(job.words, off);
  vr_decode_leaf_v5<CHAIN_LEN_ALU, SA, WIN>(rd, g.leafBits, val, job.treeDist, job.chainLut, base);
}
