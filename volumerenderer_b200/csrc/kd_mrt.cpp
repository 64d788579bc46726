// kd_mrt.cpp -- the container of the reference's MidRangeTree (/root/reference/volume_renderer/MidRangeTree.cpp:753-786 save,
// :788-833 open): the 88-byte header of VolumeKdtree::save, then distanceMap and distanceMap_range
// (maxTreeDepth + 1 bytes each), then tree.bits and tree_range.bits (treeSize bytes each, LSB-first 2-bit symbols).
// save() is reproduced byte for byte.  open() reads what save() wrote: the reference's own open() derives treeSize from
// the file size with one int64 too few in its header sum (:815, SURVEY C-3), so it over-sizes both arrays by 4 bytes and
// reads tree_range from 4 bytes too far in -- that misalignment is NOT reproduced.
#include <cstdio>
#include <cstring>

#include "kd_stream.h"

namespace vr {

int mrt_save(const vr_kd_header& hdr, const uint8_t* dm, const uint8_t* dmRange, const uint8_t* treeBits, const uint8_t* rangeBits,
             int64_t treeSize, const char* path) {
  if (!dm || !dmRange || !treeBits || !rangeBits || !path) return VR_ERR_ARG;
  if (treeSize <= 0) return VR_ERR_STATE;  // "ERROR! No tree to save." (:758-762)
  if (hdr.maxTreeDepth < 0 || hdr.maxTreeDepth + 1 > VR_MAX_DEPTH) return VR_ERR_FORMAT;
  FILE* f = std::fopen(path, "wb");
  if (!f) return VR_ERR_IO;
  const size_t nd = (size_t)hdr.maxTreeDepth + 1, nt = (size_t)treeSize;
  bool ok = std::fwrite(&hdr, sizeof(hdr), 1, f) == 1;
  ok = ok && std::fwrite(dm, 1, nd, f) == nd && std::fwrite(dmRange, 1, nd, f) == nd;
  ok = ok && std::fwrite(treeBits, 1, nt, f) == nt && std::fwrite(rangeBits, 1, nt, f) == nt;
  return (std::fclose(f) == 0 && ok) ? VR_OK : VR_ERR_IO;
}

int mrt_open(const char* path, KdStream* mid, std::vector<uint8_t>* dmRange, std::vector<uint8_t>* rangeBits) {
  if (!path || !mid || !dmRange || !rangeBits) return VR_ERR_ARG;
  FILE* f = std::fopen(path, "rb");
  if (!f) return VR_ERR_IO;  // the reference prints, waits for Enter and exit(-1)s (:794-799)
  std::fseek(f, 0, SEEK_END);
  const long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  std::vector<uint8_t> buf(sz > 0 ? (size_t)sz : 0);
  const size_t got = buf.empty() ? 0 : std::fread(buf.data(), 1, buf.size(), f);
  std::fclose(f);
  if (got != buf.size() || buf.size() < sizeof(vr_kd_header)) return VR_ERR_IO;
  vr_kd_header h;
  std::memcpy(&h, buf.data(), sizeof(h));
  if (h.maxTreeDepth < 0 || h.maxTreeDepth + 1 > VR_MAX_DEPTH) return VR_ERR_FORMAT;
  const int64_t nd = h.maxTreeDepth + 1;
  const int64_t rest = (int64_t)buf.size() - (int64_t)sizeof(h) - 2 * nd;
  if (rest <= 0 || (rest & 1)) return VR_ERR_FORMAT;
  const int64_t treeSize = rest / 2;
  const uint8_t* p = buf.data() + sizeof(h);
  int rc = kd_from_memory(&h, p, p + 2 * nd, treeSize, mid);  // checks the header and ceil(numActiveNodes/4) <= treeSize
  if (rc != VR_OK) return rc;
  dmRange->assign(p + nd, p + 2 * nd);
  rangeBits->assign(p + 2 * nd + treeSize, p + 2 * nd + 2 * treeSize);
  return VR_OK;
}

}  // namespace vr
