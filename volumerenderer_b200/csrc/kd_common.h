// kd_common.h -- shared host/device definitions for the kd-tree stream path.
//
// Stream format and tree geometry follow the reference codec:
//   container  VolumeKdtree::save/open   /root/reference/volume_renderer/VolumeKdTree_recover.cpp:521-594
//   symbols    TwoBitArray               /root/reference/volume_renderer/TwoBitArray.h:30-49 (LSB-first, 4 per byte)
//   split rule buildRecursive/levelCut   VolumeKdTree_recover.cpp:151-163, 790-798
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define VR_HD __host__ __device__ __forceinline__
#else
#define VR_HD inline
#endif

#define VR_MAX_DEPTH 64       // distanceMap entries we are prepared to hold (maxTreeDepth+1 <= 64)
#define VR_ADD_LEVELS 7       // refinement-chain levels below origTreeDepth (VolumeKdTree_recover.cpp:22)
#define VR_TILE_BITS_MAX 14   // a decode tile holds at most 2^14 voxels (16 KiB of shared memory)
#define VR_LEAF_THREADS_LOG2 8  // 256 parser threads per tile
#define VR_PRUNED 0xFFFFFFFFu

// 88-byte container header, exactly as VolumeKdtree::save writes it (little endian).
#pragma pack(push, 1)
typedef struct vr_kd_header {
  int64_t rootMin[3];
  int64_t rootMax[3];
  int32_t maxTreeDepth;
  int32_t origTreeDepth;
  int64_t X, Y, Z;
  int64_t numActiveNodes;
} vr_kd_header;
#pragma pack(pop)

// Tree geometry of one power-of-two brick: which axis each depth splits and how far the right
// child is displaced.  All boxes at one depth have the same extents, so this is per depth only.
typedef struct vr_kd_geom {
  int32_t X, Y, Z;
  int32_t lx, ly, lz;          // log2 of the extents
  int32_t D0;                  // origTreeDepth = lx+ly+lz
  int32_t Dmax;                // maxTreeDepth = D0 + 7
  int32_t tileBits;            // log2 voxels per decode tile
  int32_t leafBits;            // log2 voxels per parser thread (height of a leaf subtree)
  int32_t Dt;                  // depth of tile roots   = D0 - tileBits
  int32_t Dl;                  // depth of leaf-subtree roots = D0 - leafBits
  int32_t tx, ty, tz;          // tile box extents
  uint8_t splitDim[VR_MAX_DEPTH];   // axis split by a node at depth d (d < D0)
  int32_t childOff[VR_MAX_DEPTH];   // displacement of the right child along that axis, in voxels
  // pathBit[a][j]: which bit of a tile-local path index (tileBits wide, MSB = first split below
  // the tile root) carries bit j of the tile-local coordinate along axis a
  uint8_t pathBit[3][16];
  // inverse map: tile path bit b carries coordinate bit pbCoordBit[b] of axis pbAxis[b]
  uint8_t pbAxis[16];
  uint8_t pbCoordBit[16];
  // gather tables of the fast decode path (valid when fastPath != 0): strip-address contribution of the
  // 4 low x bits, and of the other tileBits-4 path bits in increasing path-bit order with their axis and
  // coordinate bit; fbUnit = how many of those lie inside a warp's unit (path bits < leafBits+5)
  int32_t fastPath;
  int32_t fbUnit;
  // constant-tile fills: tile-index bits of the (up to 3) finest x splits above the tile roots, finest
  // first -- tiles that differ only in these bits are x-neighbours inside the brick
  int32_t nxr;
  uint8_t xrBit[4];
  uint16_t xa[4];
  uint16_t fbAddr[12];
  uint8_t fbAxis[12];
  uint8_t fbCb[12];
} vr_kd_geom;

static VR_HD int vr_ilog2_exact(int64_t v) {  // returns -1 when v is not a power of two
  if (v <= 0 || (v & (v - 1))) return -1;
  int l = 0;
  while ((int64_t(1) << l) < v) ++l;
  return l;
}

// Fills `g` for an X*Y*Z brick.  Returns 0, or -1 for extents the codec cannot represent
// (non powers of two are silently lossy in the reference; we reject them).
static inline int vr_kd_make_geom(int64_t X, int64_t Y, int64_t Z, vr_kd_geom* g) {
  int lx = vr_ilog2_exact(X), ly = vr_ilog2_exact(Y), lz = vr_ilog2_exact(Z);
  if (lx < 0 || ly < 0 || lz < 0) return -1;
  if (lx + ly + lz + VR_ADD_LEVELS + 1 > VR_MAX_DEPTH) return -1;
  g->X = (int32_t)X; g->Y = (int32_t)Y; g->Z = (int32_t)Z;
  g->lx = lx; g->ly = ly; g->lz = lz;
  g->D0 = lx + ly + lz;
  g->Dmax = g->D0 + VR_ADD_LEVELS;
  int64_t ext[3] = {X, Y, Z};
  for (int d = 0; d < VR_MAX_DEPTH; ++d) { g->splitDim[d] = 0; g->childOff[d] = 0; }
  for (int a = 0; a < 3; ++a) for (int j = 0; j < 16; ++j) g->pathBit[a][j] = 0;
  for (int j = 0; j < 16; ++j) { g->pbAxis[j] = 0; g->pbCoordBit[j] = 0; }
  g->tileBits = g->D0 < VR_TILE_BITS_MAX ? g->D0 : VR_TILE_BITS_MAX;
  g->leafBits = g->tileBits > VR_LEAF_THREADS_LOG2 ? g->tileBits - VR_LEAF_THREADS_LOG2 : 0;
  g->Dt = g->D0 - g->tileBits;
  g->Dl = g->D0 - g->leafBits;
  for (int d = 0; d < g->D0; ++d) {
    // axis = depth % 3, advanced cyclically while that extent is already 1
    int s = d % 3, i = 0;
    while (ext[s] == 1) s = (d + ++i) % 3;
    ext[s] >>= 1;
    g->splitDim[d] = (uint8_t)s;
    g->childOff[d] = (int32_t)ext[s];
    if (d + 1 == g->Dt) { g->tx = (int32_t)ext[0]; g->ty = (int32_t)ext[1]; g->tz = (int32_t)ext[2]; }
    if (d >= g->Dt) {
      const int b = g->D0 - 1 - d, cb = vr_ilog2_exact(ext[s]);
      g->pathBit[s][cb] = (uint8_t)b;
      g->pbAxis[b] = (uint8_t)s;
      g->pbCoordBit[b] = (uint8_t)cb;
    }
  }
  if (g->Dt == 0) { g->tx = (int32_t)X; g->ty = (int32_t)Y; g->tz = (int32_t)Z; }
  g->nxr = 0;
  for (int j = 0; j < 4; ++j) g->xrBit[j] = 0;
  for (int d = g->Dt - 1; d >= 0 && g->nxr < 3; --d)
    if (g->splitDim[d] == 0) g->xrBit[g->nxr++] = (uint8_t)(g->Dt - 1 - d);
  // fast-path gather tables
  g->fastPath = 0; g->fbUnit = 0;
  for (int j = 0; j < 4; ++j) g->xa[j] = 0;
  for (int j = 0; j < 12; ++j) { g->fbAddr[j] = 0; g->fbAxis[j] = 0; g->fbCb[j] = 0; }
  if (g->tileBits == VR_TILE_BITS_MAX && g->tx >= 16) {
    const int unitBits = g->leafBits + 5;
    const int pitch = (1 << g->leafBits) + 4;  // VR_LEAF_PITCH_PAD
    int ok = 1, n = 0;
    for (int j = 0; j < 4; ++j) if (g->pathBit[0][j] >= unitBits) ok = 0;
    for (int b = 0; b < g->tileBits && ok; ++b) {
      const int addr = b < g->leafBits ? (1 << b) : (pitch << (b - g->leafBits));
      if (g->pbAxis[b] == 0 && g->pbCoordBit[b] < 4) { g->xa[g->pbCoordBit[b]] = (uint16_t)addr; continue; }
      g->fbAddr[n] = (uint16_t)addr; g->fbAxis[n] = g->pbAxis[b]; g->fbCb[n] = g->pbCoordBit[b];
      if (b < unitBits) g->fbUnit = n + 1;
      ++n;
    }
    g->fastPath = ok;
  }
  return 0;
}

// 2-bit symbol i of a packed stream (TwoBitArray.h:30-32,47-49).
static VR_HD int vr_sym(const uint8_t* bits, int64_t i) {
  return (bits[i >> 2] >> (2 * (int)(i & 3))) & 3;
}

// Child value from parent value, symbol and level distance (VolumeKdTree_recover.cpp:784-787):
// 1 = saturating add, 2 = saturating subtract, 0/3 = unchanged.
static VR_HD int vr_apply(int parent, int code, int dist) {
  int up = parent + dist; up = up > 255 ? 255 : up;
  int dn = parent - dist; dn = dn < 0 ? 0 : dn;
  return code == 1 ? up : (code == 2 ? dn : parent);
}
