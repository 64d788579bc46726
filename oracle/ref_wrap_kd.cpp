// TEST INFRASTRUCTURE ONLY -- not part of the product.
// C wrapper around the UNMODIFIED reference class VolumeKdtree, compiled from the sources where
// they lie under /root/reference (VolumeKdtree_recover.h / VolumeKdTree_recover.cpp) by
// oracle/Makefile into oracle/_ref/libvr_ref.so.  Used by tests/ and by bench.py's cpu_baseline /
// --impl reference leg only.
#include "VolumeKdtree_recover.h"
#include <cstring>
#include <sstream>
#include <thread>

namespace {
struct CoutMute {  // the reference prints progress on std::cout; keep test output clean
  std::streambuf* old;
  std::ostringstream sink;
  CoutMute() : old(std::cout.rdbuf(sink.rdbuf())) {}
  ~CoutMute() { std::cout.rdbuf(old); }
};
struct RefKd {
  std::vector<byte> input;  // build() empties this (VolumeKdTree_recover.cpp:51-52)
  VolumeKdtree* t;
};
}  // namespace

extern "C" {

// VolumeKdtree(in,x,y,z) + setMaxEpochs + setErrorTolerance + build  (main.cpp:251-258)
void* ref_kd_build(const unsigned char* vol, long long X, long long Y, long long Z, int tolerance,
                   int maxEpochs, int useThreads) {
  CoutMute m;
  RefKd* h = new RefKd;
  h->input.assign(vol, vol + X * Y * Z);
  h->t = new VolumeKdtree(h->input, X, Y, Z);
  h->t->setMaxEpochs(maxEpochs);
  h->t->setErrorTolerance(tolerance);
  h->t->build(useThreads != 0);
  return h;
}
// open(path)  (VolumeKdTree_recover.cpp:554)
void* ref_kd_open(const char* path) {
  CoutMute m;
  RefKd* h = new RefKd;
  h->t = new VolumeKdtree();
  h->t->open(path);
  return h;
}
void ref_kd_save(void* hv, const char* path) {
  CoutMute m;
  static_cast<RefKd*>(hv)->t->save(path);
}
void ref_kd_free(void* hv) {
  RefKd* h = static_cast<RefKd*>(hv);
  delete h->t;
  delete h;
}
// meta[0..6] = maxTreeDepth, origTreeDepth, X, Y, Z, numActiveNodes, bits.size()
void ref_kd_meta(void* hv, long long* meta) {
  VolumeKdtree* t = static_cast<RefKd*>(hv)->t;
  meta[0] = t->maxTreeDepth; meta[1] = t->origTreeDepth;
  meta[2] = t->X; meta[3] = t->Y; meta[4] = t->Z;
  meta[5] = t->numActiveNodes; meta[6] = (long long)t->tree.bits.size();
}
void ref_kd_get(void* hv, unsigned char* distanceMap, unsigned char* bits) {
  VolumeKdtree* t = static_cast<RefKd*>(hv)->t;
  std::memcpy(distanceMap, t->distanceMap.data(), t->distanceMap.size());
  std::memcpy(bits, t->tree.bits.data(), t->tree.bits.size());
}
// levelCut(maxTreeDepth, out)  (main.cpp:280-281, VolumeKdTree_recover.cpp:726)
void ref_kd_levelcut(void* hv, unsigned char* out) {
  VolumeKdtree* t = static_cast<RefKd*>(hv)->t;
  std::vector<byte> o;
  t->levelCut(t->maxTreeDepth, o);
  std::memcpy(out, o.data(), o.size());
}
// Timed CPU baseline: run levelCut `reps` times on each of `n` handles, bricks spread over
// `threads` host threads (one brick per thread at a time); returns seconds of wall time.
// levelCut only reads tree/distanceMap and writes the per-call vector + member pointer `output`,
// so distinct handles are independent.
double ref_kd_levelcut_bench(void** hv, int n, int reps, int threads) {
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (int w = 0; w < threads; ++w)
    pool.emplace_back([=]() {
      std::vector<byte> o;
      for (int r = 0; r < reps; ++r)
        for (int i = w; i < n; i += threads) {
          VolumeKdtree* t = static_cast<RefKd*>(hv[i])->t;
          t->levelCut(t->maxTreeDepth, o);
        }
    });
  for (auto& th : pool) th.join();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
}
