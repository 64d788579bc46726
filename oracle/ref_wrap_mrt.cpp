// TEST INFRASTRUCTURE ONLY -- not part of the product.
// C wrapper around the UNMODIFIED reference class MidRangeTree (MidRangeTree.h / MidRangeTree.cpp),
// separate translation unit because it re-declares `byte`, `MinMax`, `Point3i`.
#include "MidRangeTree.h"
#include <cstring>
#include <sstream>

namespace {
struct CoutMute {
  std::streambuf* old;
  std::ostringstream sink;
  CoutMute() : old(std::cout.rdbuf(sink.rdbuf())) {}
  ~CoutMute() { std::cout.rdbuf(old); }
};
struct RefMrt {
  std::vector<byte> input;
  MidRangeTree* t;
};
}  // namespace

extern "C" {
void* ref_mrt_build(const unsigned char* vol, long long X, long long Y, long long Z, int tolerance,
                    int maxEpochs, int useThreads) {
  CoutMute m;
  RefMrt* h = new RefMrt;
  h->input.assign(vol, vol + X * Y * Z);
  h->t = new MidRangeTree(h->input, X, Y, Z);
  h->t->setMaxEpochs(maxEpochs);
  h->t->setErrorTolerance(tolerance);
  h->t->build(useThreads != 0);
  return h;
}
void ref_mrt_free(void* hv) {
  RefMrt* h = static_cast<RefMrt*>(hv);
  delete h->t;
  delete h;
}
// meta[0..6] = maxTreeDepth, origTreeDepth, X, Y, Z, numActiveNodes, tree.bits.size()
void ref_mrt_meta(void* hv, long long* meta) {
  MidRangeTree* t = static_cast<RefMrt*>(hv)->t;
  meta[0] = t->maxTreeDepth; meta[1] = t->origTreeDepth;
  meta[2] = t->X; meta[3] = t->Y; meta[4] = t->Z;
  meta[5] = t->numActiveNodes; meta[6] = (long long)t->tree.bits.size();
}
void ref_mrt_get(void* hv, unsigned char* dmap, unsigned char* dmap_range, unsigned char* bits,
                 unsigned char* bits_range) {
  MidRangeTree* t = static_cast<RefMrt*>(hv)->t;
  std::memcpy(dmap, t->distanceMap.data(), t->distanceMap.size());
  std::memcpy(dmap_range, t->distanceMap_range.data(), t->distanceMap_range.size());
  std::memcpy(bits, t->tree.bits.data(), t->tree.bits.size());
  std::memcpy(bits_range, t->tree_range.bits.data(), t->tree_range.bits.size());
}
void ref_mrt_levelcut(void* hv, unsigned char* out) {
  MidRangeTree* t = static_cast<RefMrt*>(hv)->t;
  std::vector<byte> o;
  t->levelCut(t->maxTreeDepth, o);
  std::memcpy(out, o.data(), o.size());
}
// convertToByteArray (MidRangeTree.cpp:1095); returns byte length, copies when out != NULL
long long ref_mrt_bytearray(void* hv, unsigned char* out) {
  MidRangeTree* t = static_cast<RefMrt*>(hv)->t;
  std::vector<byte> o;
  t->convertToByteArray(o);
  if (out) std::memcpy(out, o.data(), o.size());
  return (long long)o.size();
}
}

extern "C" {
// MidRangeTree::save (MidRangeTree.cpp:753-786) of the unmodified reference
void ref_mrt_save(void* hv, const char* path) {
  CoutMute m;
  static_cast<RefMrt*>(hv)->t->save(path);
}
}
