// kd_stream.h -- host-side kd-tree stream object: container I/O, grammar walk, decode index.
//
// Replaces the in-memory state of the reference's `class VolumeKdtree`
// (/root/reference/volume_renderer/VolumeKdtree_recover.h:57-79: tree, distanceMap, maxTreeDepth,
// origTreeDepth, numActiveNodes, X, Y, Z, rootMin, rootMax) plus a derived side index that lets the
// GPU start parsing the preorder stream at many places at once (the stream itself carries no
// subtree sizes, VolumeKdTree_recover.cpp:631-724).
#pragma once
#include <string>
#include <vector>
#include "kd_common.h"

namespace vr {

// Derived decode index ("prepared stream").  Never written to the .bin container.
//   tiles : one entry per depth-Dt node slot in path order (2^Dt of them).
//           tileLeafBlock[t] == VR_PRUNED  -> the whole tile box is the constant tileVal[t]
//           otherwise                      -> leaf entries at [tileLeafBlock[t]*leavesPerTile, +leavesPerTile)
//   leaves: one entry per depth-Dl node slot of a non-constant tile, in path order.
//           leafOff[i] == VR_PRUNED -> the leaf box is the constant leafVal[i]
//           otherwise               -> symbol offset of that node; leafVal[i] = the node's own decoded value
struct KdIndex {
  std::vector<uint32_t> tileLeafBlock;
  std::vector<uint8_t> tileVal;
  std::vector<uint32_t> leafOff;
  std::vector<uint8_t> leafVal;
  std::vector<uint8_t> tileMax, tileMin;  // exact bounds of the decoded voxels of every tile (host only; kd_cull.h)
  std::vector<uint32_t> blockSymBegin;  // per live tile: first symbol of the tile subtree
  std::vector<uint32_t> blockSymEnd;    // per live tile: one past its last symbol
  int64_t numLiveTiles = 0;
  int64_t maxSymbolsPerLeaf = 0;  // longest leaf-subtree segment, in symbols
  bool built = false;
};

struct KdStream {
  vr_kd_header hdr{};
  std::vector<uint8_t> distanceMap;  // maxTreeDepth+1 entries, index = depth
  std::vector<uint8_t> bits;         // ceil(numActiveNodes/4) bytes (trailing slack tolerated)
  vr_kd_geom geom{};
  KdIndex index;
  // Indexed container opened with a tile subset (kd_kdx.cpp): residentTiles[t] != 0 -> the symbols of tile t are in
  // `bits`; the rest of `bits` is zero.  Empty = the whole stream is resident (every other way of making a stream).
  std::vector<uint8_t> residentTiles;

  int64_t voxels() const { return hdr.X * hdr.Y * hdr.Z; }
  int64_t streamBytes() const { return (hdr.numActiveNodes + 3) / 4; }
};

#ifndef VR_OK  // same values as include/vr_b200.h
#define VR_OK 0
#define VR_ERR_ARG 1          // bad argument / null pointer
#define VR_ERR_IO 2           // file missing or short
#define VR_ERR_FORMAT 3       // header inconsistent, non power-of-two extents, grammar violation
#define VR_ERR_CUDA 4         // CUDA runtime error (see vr_last_error)
#define VR_ERR_UNSUPPORTED 5  // e.g. levelCut with cutDepth != maxTreeDepth (crashes in the reference)
#define VR_ERR_STATE 6        // call order (decode before prepare, render before volume, ...)
#endif

// Container -----------------------------------------------------------------------------------
int kd_from_memory(const vr_kd_header* hdr, const uint8_t* distanceMap, const uint8_t* bits,
                   int64_t nbytes, KdStream* out);
int kd_open(const char* path, KdStream* out);          // accepts exact and +8-byte-slack files
int kd_save(const KdStream& s, const char* path);      // byte-identical to VolumeKdtree::save

// Full grammar walk: returns VR_OK iff the stream parses to exactly numActiveNodes symbols and
// covers every voxel once.  It produces no voxels: decoding is the CUDA path only.
int kd_validate(const KdStream& s, int64_t* nodesConsumed);

// Builds s->index (sequential walk, O(numActiveNodes)).
int kd_prepare(KdStream* s);
// Derives blockSymBegin / blockSymEnd / numLiveTiles / maxSymbolsPerLeaf from tileLeafBlock + leafOff (what kd_prepare does
// after its walk); `tileBegin[t]` = symbol offset of live tile t.  Used by producers that know the offsets already.
void kd_index_finish(const vr_kd_geom& g, const std::vector<uint32_t>& tileBegin, int64_t numActiveNodes, KdIndex* ix);
int kd_check_header(const vr_kd_header& h, vr_kd_geom* g);

// Indexed ("random-access") container, SURVEY 8f-4 -- what HashedKdtree.cpp:96-214,406-507 was reaching for: the stream
// plus its decode index and exact per-tile bounds in one deterministic file, so that opening needs no walk of the stream
// and single tiles can be fetched from disk.  Layout in kd_kdx.cpp.
int kdx_serialize(KdStream* s, std::vector<uint8_t>* out);  // prepares the index if needed
int kdx_save(KdStream* s, const char* path);
// tileMask (optional, 2^Dt bytes): read only the stream bytes of these tiles.  verify: check the stream checksum (full opens only).
int kdx_from_memory(const uint8_t* buf, int64_t nbytes, const uint8_t* tileMask, bool verify, KdStream* out);
int kdx_open(const char* path, const uint8_t* tileMask, bool verify, KdStream* out, int64_t* bytesRead);
int kd_open_any(const char* path, KdStream* out);  // plain or indexed container, by magic

// MidRangeTree::save / open container (kd_mrt.cpp): midpoint stream + range stream in one file.
int mrt_save(const vr_kd_header& hdr, const uint8_t* dm, const uint8_t* dmRange, const uint8_t* treeBits, const uint8_t* rangeBits,
             int64_t treeSize, const char* path);
int mrt_open(const char* path, KdStream* mid, std::vector<uint8_t>* dmRange, std::vector<uint8_t>* rangeBits);

// CPU compressor: the reference's build() algorithm (VolumeKdTree_recover.cpp:17-140) restated;
// produces byte-identical streams.  `volume` is not modified (the reference empties its input).
struct KdBuildStats {
  int maxErrorBeforeGrowth = 0, maxError = 0;
  double meanL1 = 0, meanL2 = 0;
  int64_t zeroRunRewrites = 0;  // SURVEY C-13: must stay 0
};
int kd_compress(const uint8_t* volume, int64_t X, int64_t Y, int64_t Z, int tolerance, int maxEpochs,
                KdStream* out, KdBuildStats* stats);

}  // namespace vr
