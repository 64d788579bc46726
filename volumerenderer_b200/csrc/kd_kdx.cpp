// kd_kdx.cpp -- the indexed ("random-access") stream container.
//
// What the reference's HashedKdtree was reaching for (/root/reference/volume_renderer/HashedKdtree.cpp:96-214 build,
// :254-336 levelCutRecursive, :406-507 compress, :528-545 save): a layout from which any part of a brick can be decoded
// without parsing what lies in front of it.  The reference keys nodes by Morton code into an open hash with a collision
// overflow and seeds it from std::random_device (HashedKdtree.h:82-84); it crashes.  Here the preorder stream of
// VolumeKdtree::save (VolumeKdTree_recover.cpp:521-552) is kept bit for bit and the decode index that vr_stream_prepare
// derives from it (kd_stream.cpp) is stored beside it, with the exact bounds of every tile:
//   * opening needs no walk of the stream (the host walk costs 50-130 ms per dense 256x256x128 brick);
//   * a tile's symbols are one contiguous byte range [blockSymBegin/4, (blockSymEnd+3)/4), so a reader fetches single tiles
//     from disk (kdx_open with a tile mask) -- the view-dependent decode of kd_cull.h says which;
//   * the file is a pure function of the stream: no random seed, no collision handling, no growth policy.
//
// Layout (little endian, every section 16-byte aligned, gaps zero):
//   KdxHeader (352 bytes: magic, version, section count, file size, FNV-1a 64 of the stream section, tile/leaf geometry,
//              live-tile count, longest leaf segment, the reference's 88-byte header unchanged, distanceMap padded to 64,
//              section table {offset, bytes} x 9)
//   1 tileLeafBlock u32[2^Dt]   2 tileVal u8[2^Dt]   3 tileMin u8[2^Dt]   4 tileMax u8[2^Dt]
//   5 blockSymBegin u32[live]   6 blockSymEnd u32[live]   7 leafOff u32[live * leavesPerTile]   8 leafVal u8[live * leavesPerTile]
//   0 the 2-bit stream, ceil(numActiveNodes / 4) bytes -- last, so that header + index are one read
#include <cstdio>
#include <cstring>

#include "kd_stream.h"

namespace vr {

namespace {
constexpr int KDX_SECTIONS = 9;
constexpr char KDX_MAGIC[8] = {'V', 'R', 'K', 'D', 'X', 0, 0, 1};

#pragma pack(push, 1)
struct KdxHeader {
  char magic[8];
  uint32_t version;
  uint32_t nSections;
  uint64_t fileBytes;
  uint64_t bitsFnv1a;
  uint32_t tileBits, leafBits;
  uint64_t numLiveTiles;
  uint64_t maxSymbolsPerLeaf;
  vr_kd_header kd;
  uint8_t distanceMap[VR_MAX_DEPTH];
  struct { uint64_t offset, bytes; } sec[KDX_SECTIONS];
};
#pragma pack(pop)
static_assert(sizeof(KdxHeader) == 352, "header layout");

inline uint64_t align16(uint64_t v) { return (v + 15) & ~uint64_t(15); }

uint64_t fnv1a(const uint8_t* p, size_t n) {
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 1099511628211ull; }
  return h;
}

// one interface over a file and a memory buffer
struct Source {
  FILE* f = nullptr;
  const uint8_t* mem = nullptr;
  uint64_t size = 0;
  int64_t bytesRead = 0;
  bool read(uint64_t off, uint64_t n, void* dst) {
    if (off > size || n > size - off) return false;
    bytesRead += (int64_t)n;
    if (mem) { std::memcpy(dst, mem + off, (size_t)n); return true; }
    if (std::fseek(f, (long)off, SEEK_SET) != 0) return false;
    return std::fread(dst, 1, (size_t)n, f) == (size_t)n;
  }
};

template <typename T>
bool read_vec(Source& src, const KdxHeader& h, int sec, uint64_t count, std::vector<T>* out) {
  if (h.sec[sec].bytes != count * sizeof(T)) return false;
  out->resize((size_t)count);
  return count == 0 || src.read(h.sec[sec].offset, h.sec[sec].bytes, out->data());
}

int parse(Source& src, const uint8_t* tileMask, bool verify, KdStream* out) {
  KdxHeader h;
  if (!src.read(0, sizeof(h), &h)) return VR_ERR_IO;
  if (std::memcmp(h.magic, KDX_MAGIC, 8) != 0 || h.version != 1 || h.nSections != KDX_SECTIONS) return VR_ERR_FORMAT;
  if (h.fileBytes > src.size) return VR_ERR_IO;  // truncated
  vr_kd_geom g;
  if (h.kd.maxTreeDepth < 0 || h.kd.maxTreeDepth + 1 > VR_MAX_DEPTH) return VR_ERR_FORMAT;
  int rc = kd_check_header(h.kd, &g);
  if (rc != VR_OK) return rc;
  if (h.kd.numActiveNodes >= (int64_t)VR_PRUNED) return VR_ERR_UNSUPPORTED;
  if ((int)h.tileBits != g.tileBits || (int)h.leafBits != g.leafBits) return VR_ERR_FORMAT;  // index built for another tiling
  for (int k = 0; k < KDX_SECTIONS; ++k)
    if (h.sec[k].offset % 16 || h.sec[k].offset < sizeof(h) || h.sec[k].offset > h.fileBytes || h.sec[k].bytes > h.fileBytes - h.sec[k].offset)
      return VR_ERR_FORMAT;
  const uint64_t nTiles = uint64_t(1) << g.Dt, lpt = uint64_t(1) << (g.tileBits - g.leafBits), live = h.numLiveTiles;
  if (live > nTiles) return VR_ERR_FORMAT;
  const uint64_t nSym = (uint64_t)h.kd.numActiveNodes, nBytes = (nSym + 3) / 4;
  if (h.sec[0].bytes != nBytes) return VR_ERR_FORMAT;

  KdStream s;
  s.hdr = h.kd;
  s.geom = g;
  s.distanceMap.assign(h.distanceMap, h.distanceMap + h.kd.maxTreeDepth + 1);
  KdIndex& ix = s.index;
  if (!read_vec(src, h, 1, nTiles, &ix.tileLeafBlock) || !read_vec(src, h, 2, nTiles, &ix.tileVal) || !read_vec(src, h, 3, nTiles, &ix.tileMin) ||
      !read_vec(src, h, 4, nTiles, &ix.tileMax) || !read_vec(src, h, 5, live, &ix.blockSymBegin) || !read_vec(src, h, 6, live, &ix.blockSymEnd) ||
      !read_vec(src, h, 7, live * lpt, &ix.leafOff) || !read_vec(src, h, 8, live * lpt, &ix.leafVal))
    return VR_ERR_FORMAT;
  // structural checks that need no walk: blocks numbered in path order, symbol ranges ascending and inside the stream,
  // leaf offsets ascending inside their block's range
  uint64_t nextBlock = 0;
  for (uint64_t t = 0; t < nTiles; ++t) {
    const uint32_t b = ix.tileLeafBlock[t];
    if (b == VR_PRUNED) continue;
    if (b != nextBlock++) return VR_ERR_FORMAT;
  }
  if (nextBlock != live) return VR_ERR_FORMAT;
  for (uint64_t b = 0; b < live; ++b) {
    const uint64_t lo = ix.blockSymBegin[b], hi = ix.blockSymEnd[b];
    if (lo >= hi || hi > nSym || (b + 1 < live && hi != ix.blockSymBegin[b + 1])) return VR_ERR_FORMAT;
    int64_t prev = (int64_t)lo - 1;
    for (uint64_t l = 0; l < lpt; ++l) {
      const uint32_t o = ix.leafOff[b * lpt + l];
      if (o == VR_PRUNED) continue;
      if ((int64_t)o <= prev || o >= hi) return VR_ERR_FORMAT;
      prev = o;
    }
  }
  ix.numLiveTiles = (int64_t)live;
  ix.maxSymbolsPerLeaf = (int64_t)h.maxSymbolsPerLeaf;
  ix.built = true;

  s.bits.assign((size_t)nBytes, 0);
  if (!tileMask) {
    if (!src.read(h.sec[0].offset, nBytes, s.bits.data())) return VR_ERR_IO;
    if (verify && fnv1a(s.bits.data(), (size_t)nBytes) != h.bitsFnv1a) return VR_ERR_FORMAT;
  } else {
    // byte ranges of the selected live tiles, neighbouring ranges merged into one read
    s.residentTiles.assign((size_t)nTiles, 0);
    uint64_t runLo = 0, runHi = 0;
    bool open = false;
    auto flush = [&]() -> bool {
      if (!open) return true;
      open = false;
      return src.read(h.sec[0].offset + runLo, runHi - runLo, s.bits.data() + runLo);
    };
    for (uint64_t t = 0; t < nTiles; ++t) {
      if (!tileMask[t]) continue;
      s.residentTiles[(size_t)t] = 1;
      const uint32_t b = ix.tileLeafBlock[t];
      if (b == VR_PRUNED) continue;  // a constant tile has no symbols of its own
      const uint64_t lo = ix.blockSymBegin[b] / 4, hi = ((uint64_t)ix.blockSymEnd[b] + 3) / 4;
      if (open && lo <= runHi) { runHi = hi; continue; }
      if (!flush()) return VR_ERR_IO;
      runLo = lo; runHi = hi; open = true;
    }
    if (!flush()) return VR_ERR_IO;
  }
  *out = std::move(s);
  return VR_OK;
}
}  // namespace

int kdx_serialize(KdStream* s, std::vector<uint8_t>* out) {
  if (!s || !out || s->bits.empty()) return VR_ERR_STATE;
  if (!s->residentTiles.empty()) return VR_ERR_STATE;  // a tile subset is not the stream
  if (!s->index.built) {
    int rc = kd_prepare(s);
    if (rc != VR_OK) return rc;
  }
  const KdIndex& ix = s->index;
  KdxHeader h;
  std::memset(&h, 0, sizeof(h));
  std::memcpy(h.magic, KDX_MAGIC, 8);
  h.version = 1;
  h.nSections = KDX_SECTIONS;
  h.tileBits = (uint32_t)s->geom.tileBits;
  h.leafBits = (uint32_t)s->geom.leafBits;
  h.numLiveTiles = (uint64_t)ix.numLiveTiles;
  h.maxSymbolsPerLeaf = (uint64_t)ix.maxSymbolsPerLeaf;
  h.kd = s->hdr;
  std::memcpy(h.distanceMap, s->distanceMap.data(), s->distanceMap.size());
  const size_t nBytes = (size_t)s->streamBytes();
  h.bitsFnv1a = fnv1a(s->bits.data(), nBytes);
  const void* ptr[KDX_SECTIONS] = {s->bits.data(),          ix.tileLeafBlock.data(), ix.tileVal.data(), ix.tileMin.data(), ix.tileMax.data(),
                                   ix.blockSymBegin.data(), ix.blockSymEnd.data(),   ix.leafOff.data(), ix.leafVal.data()};
  const uint64_t bytes[KDX_SECTIONS] = {nBytes,
                                        ix.tileLeafBlock.size() * 4,
                                        ix.tileVal.size(),
                                        ix.tileMin.size(),
                                        ix.tileMax.size(),
                                        ix.blockSymBegin.size() * 4,
                                        ix.blockSymEnd.size() * 4,
                                        ix.leafOff.size() * 4,
                                        ix.leafVal.size()};
  uint64_t off = sizeof(h);
  const int order[KDX_SECTIONS] = {1, 2, 3, 4, 5, 6, 7, 8, 0};
  for (int k : order) { h.sec[k].offset = off; h.sec[k].bytes = bytes[k]; off = align16(off + bytes[k]); }
  h.fileBytes = off;
  out->assign((size_t)off, 0);
  std::memcpy(out->data(), &h, sizeof(h));
  for (int k = 0; k < KDX_SECTIONS; ++k)
    if (bytes[k]) std::memcpy(out->data() + h.sec[k].offset, ptr[k], (size_t)bytes[k]);
  return VR_OK;
}

int kdx_save(KdStream* s, const char* path) {
  if (!path) return VR_ERR_ARG;
  std::vector<uint8_t> buf;
  int rc = kdx_serialize(s, &buf);
  if (rc != VR_OK) return rc;
  FILE* f = std::fopen(path, "wb");
  if (!f) return VR_ERR_IO;
  const bool ok = std::fwrite(buf.data(), 1, buf.size(), f) == buf.size();
  return (std::fclose(f) == 0 && ok) ? VR_OK : VR_ERR_IO;
}

int kdx_from_memory(const uint8_t* buf, int64_t nbytes, const uint8_t* tileMask, bool verify, KdStream* out) {
  if (!buf || nbytes <= 0 || !out) return VR_ERR_ARG;
  Source src;
  src.mem = buf;
  src.size = (uint64_t)nbytes;
  return parse(src, tileMask, verify, out);
}

int kdx_open(const char* path, const uint8_t* tileMask, bool verify, KdStream* out, int64_t* bytesRead) {
  if (!path || !out) return VR_ERR_ARG;
  FILE* f = std::fopen(path, "rb");
  if (!f) return VR_ERR_IO;
  Source src;
  src.f = f;
  std::fseek(f, 0, SEEK_END);
  src.size = (uint64_t)std::ftell(f);
  int rc = parse(src, tileMask, verify, out);
  std::fclose(f);
  if (bytesRead) *bytesRead = src.bytesRead;
  return rc;
}

// VolumeKdtree::open on either container: the indexed one is recognised by its magic (a plain container starts with
// rootMin = 0,0,0: eight zero bytes) and opened without the walk.
int kd_open_any(const char* path, KdStream* out) {
  if (!path || !out) return VR_ERR_ARG;
  FILE* f = std::fopen(path, "rb");
  if (!f) return VR_ERR_IO;
  char magic[8] = {0};
  const size_t got = std::fread(magic, 1, 8, f);
  std::fclose(f);
  if (got == 8 && std::memcmp(magic, KDX_MAGIC, 8) == 0) return kdx_open(path, nullptr, false, out, nullptr);
  return kd_open(path, out);
}

}  // namespace vr
