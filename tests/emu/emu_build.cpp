// emu_build.cpp -- TEST HARNESS ONLY: runs the GPU stream producer's algorithm (csrc/kd_build.cu) on the CPU,
// kernel by kernel as plain loops over the same host/device element functions (csrc/kd_build_core.h) and with the
// same parallel formulations (fixed-point running mean over prefix sums, hierarchical order-exact accumulation,
// subtree sizes -> offsets -> scatter), so that its exactness can be checked where there is no GPU.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../volumerenderer_b200/csrc/kd_build_core.h"
#include "../../volumerenderer_b200/csrc/kd_stream.h"

using namespace vr;

namespace {
inline int parent_of(const std::vector<uint8_t>& parents, int depth, int64_t i) { return depth == 0 ? 0 : (int)parents[(size_t)(i >> 1)]; }
}

extern "C" {

// naive sequential loop vs vr_seq_accumulate on e[0..n): returns 1 when bit-identical
int emu_accumulate_check(double x0, const uint32_t* e, long long n) {
  double ref = x0;
  for (long long i = 0; i < n; ++i) ref += (double)e[i];
  const long long nb = (n + VR_ACC_BLOCK - 1) / VR_ACC_BLOCK, ns = (nb + VR_ACC_SUPER - 1) / VR_ACC_SUPER;
  std::vector<uint32_t> bs((size_t)nb + 1, 0);
  std::vector<uint64_t> ss((size_t)ns + 1, 0);
  for (long long i = 0; i < n; ++i) { bs[(size_t)(i / VR_ACC_BLOCK)] += e[i]; ss[(size_t)(i / VR_ACC_BLOCK / VR_ACC_SUPER)] += e[i]; }
  const double got = vr_seq_accumulate(x0, (int64_t)n, bs.data(), ss.data(), [&](int64_t i) { return e[i]; });
  return std::memcmp(&ref, &got, sizeof(double)) == 0;
}

// Full producer.  Returns 0 and fills the stream handle; iters[0..1] = scan passes (total, worst level).
int emu_build(const uint8_t* volume, long long X, long long Y, long long Z, int tolerance, int maxEpochs, void** outStream, long long* iters) {
  vr_kd_geom g;
  if (vr_kd_make_geom(X, Y, Z, &g) != 0) return VR_ERR_FORMAT;
  const int D0 = g.D0, Dmax = g.Dmax;
  const int64_t N = int64_t(1) << D0, firstLeaf = N - 1, T = 2 * N - 1;
  std::vector<uint8_t> temp((size_t)T), mn((size_t)T), mx((size_t)T), codes((size_t)T, 0);
  for (int64_t p = 0; p < N; ++p) {
    const uint8_t s = volume[vr_leaf_voxel(g, (uint32_t)p, X, X * Y)];
    temp[(size_t)(firstLeaf + p)] = mn[(size_t)(firstLeaf + p)] = mx[(size_t)(firstLeaf + p)] = s;
  }
  for (int64_t i = firstLeaf - 1; i >= 0; --i) {
    const uint8_t a = std::min(mn[2 * i + 1], mn[2 * i + 2]), b = std::max(mx[2 * i + 1], mx[2 * i + 2]);
    mn[(size_t)i] = a; mx[(size_t)i] = b;
    temp[(size_t)i] = (uint8_t)(((int)a + (int)b) / 2);
  }
  std::vector<uint8_t> distanceMap((size_t)Dmax + 1, 0);
  std::vector<uint8_t> parents, recon, reconPrev;
  const double gamma = 1.25, hstep = 1.0, maxAbsStep = 4.0;
  double currentDistance = 0.0, currentError = 0.0, currentDF = 0.0, currentStepSize = 0.0;
  double previousStepSize = 0.0, previousDistance = 0.0, previousDF = 0.0, previousError = 0.0;
  long long itTotal = 0, itWorst = 0;
  for (int depth = 0; depth <= D0; ++depth) {
    const int64_t n = int64_t(1) << depth;
    const uint8_t* truth = temp.data() + (n - 1);
    uint8_t* code = codes.data() + (n - 1);
    int epoch = 0;
    recon.assign((size_t)n, 0);
    if (const char* dd = getenv("EMU_DUMP_LEVELS")) {  // level inputs of the running-mean estimate, for convergence studies
      char path[512];
      snprintf(path, sizeof(path), "%s/level_%02d.bin", dd, depth);
      if (FILE* f = fopen(path, "wb")) {
        long long nn = n;
        std::vector<uint8_t> par((size_t)n);
        for (int64_t i = 0; i < n; ++i) par[(size_t)i] = (uint8_t)parent_of(parents, depth, i);
        fwrite(&nn, 8, 1, f); fwrite(&depth, 4, 1, f); fwrite(par.data(), 1, (size_t)n, f); fwrite(truth, 1, (size_t)n, f);
        fclose(f);
      }
    }
    // running mean as a fixed point over prefix sums (est_init / scan / est_iter of kd_build.cu)
    double distanceSum = 0.0, distanceCount = 0.0;
    {
      std::vector<uint64_t> items((size_t)n), prefix((size_t)n);
      std::vector<uint8_t> dec((size_t)n);
      for (int64_t i = 0; i < n; ++i) {
        const int pd = std::abs(parent_of(parents, depth, i) - (int)truth[i]);
        dec[(size_t)i] = pd > 0;
        items[(size_t)i] = pd > 0 ? (((uint64_t)pd << 32) | 1u) : 0;
      }
      long long it = 0;
      for (;; ++it) {
        uint64_t run = 0;
        for (int64_t i = 0; i < n; ++i) { prefix[(size_t)i] = run; run += items[(size_t)i]; }
        unsigned changed = 0;
        int64_t first = -1;
        for (int64_t i = 0; i < n; ++i) {
          const int parent = parent_of(parents, depth, i), t = truth[i];
          const uint64_t st = prefix[(size_t)i];
          const bool d = vr_estimate_counts(parent, t, (double)(st >> 32), (double)(uint32_t)st);
          if (d != (dec[(size_t)i] != 0)) {
            dec[(size_t)i] = d;
            items[(size_t)i] = d ? (((uint64_t)std::abs(parent - t) << 32) | 1u) : 0;
            ++changed;
            if (first < 0) first = i;
          }
        }
        if (!changed) break;
        if (changed < 256) {  // est_window_kernel
          const uint64_t st = prefix[(size_t)first];
          double sum = (double)(st >> 32), count = (double)(uint32_t)st;
          const int64_t end = std::min<int64_t>(first + 4096, n);
          for (int64_t i = first; i < end; ++i) {
            const int parent = parent_of(parents, depth, i), t = truth[i];
            const bool d = vr_estimate_counts(parent, t, sum, count);
            dec[(size_t)i] = d;
            items[(size_t)i] = d ? (((uint64_t)std::abs(parent - t) << 32) | 1u) : 0;
            if (d) { sum += (double)std::abs(parent - t); count += 1.0; }
          }
        }
      }
      itTotal += it + 1;
      if (it + 1 > itWorst) itWorst = it + 1;
      const uint64_t st = prefix[(size_t)(n - 1)] + items[(size_t)(n - 1)];
      distanceSum = (double)(st >> 32);
      distanceCount = (double)(uint32_t)st;
    }
    currentDistance = distanceCount > 0 ? std::round(distanceSum / distanceCount) : 0.0;
    previousDistance = 0.0; previousStepSize = 255.0; previousError = 65025.0;
    while (epoch < maxEpochs && std::fabs(previousStepSize) >= 0.5) {
      if (epoch != 0) {
        previousDistance = currentDistance; previousError = currentError; previousDF = currentDF; previousStepSize = currentStepSize;
        currentDistance = std::round(std::fmin(255.0, std::fmax(0.0, previousDistance + previousStepSize)));
        if (currentDistance == previousDistance) break;
      }
      {
        const int dist = (uint8_t)currentDistance;
        const int64_t nb = (n + VR_ACC_BLOCK - 1) / VR_ACC_BLOCK;
        std::vector<uint32_t> bs((size_t)nb + 1, 0);
        std::vector<uint64_t> ss((size_t)(nb / VR_ACC_SUPER) + 2, 0);
        for (int64_t i = 0; i < n; ++i) {
          const vr_enc e = vr_encode_node(parent_of(parents, depth, i), dist, truth[i]);
          code[i] = (uint8_t)e.code; recon[(size_t)i] = (uint8_t)e.value;
          bs[(size_t)(i / VR_ACC_BLOCK)] += (uint32_t)(e.err * e.err);
          ss[(size_t)(i / VR_ACC_BLOCK / VR_ACC_SUPER)] += (uint64_t)(e.err * e.err);
        }
        currentError = vr_seq_accumulate(currentError, n, bs.data(), ss.data(), [&](int64_t i) {
          const vr_enc e = vr_encode_node(parent_of(parents, depth, i), dist, truth[i]);
          return (uint32_t)(e.err * e.err);
        });
        currentError /= (double)n;
      }
      if (currentError < 1.0) break;
      if (epoch != 0 && currentError > previousError) {
        currentError = previousError; currentDistance = previousDistance; currentDF = previousDF; currentStepSize = previousStepSize / 2.0;
        recon.swap(reconPrev);
        epoch++;
        continue;
      }
      const int est[2] = {(uint8_t)std::fmax(0.0, currentDistance - hstep), (uint8_t)std::fmin(255.0, currentDistance + hstep)};
      double estErr[2];
      for (int k = 0; k < 2; ++k) {
        uint64_t acc = 0;
        for (int64_t i = 0; i < n; ++i) {
          const vr_enc e = vr_encode_node(parent_of(parents, depth, i), est[k], truth[i]);
          acc += (uint64_t)(e.err * e.err);
        }
        estErr[k] = (double)acc / (double)n;
      }
      currentDF = (estErr[1] - estErr[0]) / (2.0 * hstep);
      currentStepSize = std::fmax(-maxAbsStep, std::fmin(maxAbsStep, -gamma * currentDF));
      reconPrev = recon;
      epoch++;
    }
    distanceMap[(size_t)depth] = (uint8_t)currentDistance;
    if (depth < D0) parents.swap(recon);
  }
  const uint8_t* leafTruth = temp.data() + firstLeaf;
  uint8_t* leafCode = codes.data() + firstLeaf;
  for (int64_t i = 0; i < N; ++i)
    if (leafCode[i] == 0 && std::abs((int)recon[(size_t)i] - (int)leafTruth[i]) < tolerance) leafCode[i] = 3;
  for (int64_t i = firstLeaf - 1; i >= 0; --i)
    if (codes[(size_t)i] == 0 && codes[(size_t)(2 * i + 1)] == 3 && codes[(size_t)(2 * i + 2)] == 3) codes[(size_t)i] = 3;
  uint8_t dist7[VR_ADD_LEVELS];
  {
    const int add[VR_ADD_LEVELS] = {64, 32, 16, 8, 4, 2, 1};
    for (int k = 0; k < VR_ADD_LEVELS; ++k) { distanceMap[(size_t)(D0 + 1 + k)] = (uint8_t)add[k]; dist7[k] = (uint8_t)add[k]; }
  }
  // chains, sizes, offsets, scatter
  std::vector<uint32_t> chain((size_t)N), size((size_t)T), off((size_t)T);
  for (int64_t i = 0; i < N; ++i) {
    uint32_t packed = 0, cnt = 0;
    if (leafCode[i] != 3) {
      uint8_t sym[8];
      int r, rew;
      cnt = (uint32_t)vr_leaf_chain(recon[(size_t)i], leafTruth[i], tolerance, dist7, sym, &r, &rew);
      for (uint32_t k = 0; k < cnt; ++k) packed |= (uint32_t)sym[k] << (2 * k);
      recon[(size_t)i] = (uint8_t)r;
    }
    chain[(size_t)i] = packed | (cnt << 16);
    size[(size_t)(firstLeaf + i)] = 1 + cnt;
  }
  for (int64_t i = firstLeaf - 1; i >= 0; --i) size[(size_t)i] = codes[(size_t)i] == 3 ? 1u : 1u + size[(size_t)(2 * i + 1)] + size[(size_t)(2 * i + 2)];
  const int64_t numActive = size[0];
  std::vector<uint8_t> sym((size_t)numActive, 0xEE);
  const uint32_t NOT = 0xFFFFFFFFu;
  off[0] = 0;
  for (int64_t i = 0; i < T; ++i) {
    const bool leaf = i >= firstLeaf;
    const uint32_t o = off[(size_t)i];
    if (o == NOT) { if (!leaf) { off[(size_t)(2 * i + 1)] = NOT; off[(size_t)(2 * i + 2)] = NOT; } continue; }
    const uint8_t c = codes[(size_t)i];
    sym[o] = c;
    if (leaf) {
      const uint32_t ch = chain[(size_t)(i - firstLeaf)], cnt = ch >> 16;
      for (uint32_t q = 0; q < cnt; ++q) sym[o + 1 + q] = (uint8_t)((ch >> (2 * q)) & 3u);
      continue;
    }
    if (c == 3) { off[(size_t)(2 * i + 1)] = NOT; off[(size_t)(2 * i + 2)] = NOT; continue; }
    off[(size_t)(2 * i + 1)] = o + 1;
    off[(size_t)(2 * i + 2)] = o + 1 + size[(size_t)(2 * i + 1)];
  }
  KdStream* s = new KdStream;
  std::memset(&s->hdr, 0, sizeof(s->hdr));
  s->hdr.rootMax[0] = X; s->hdr.rootMax[1] = Y; s->hdr.rootMax[2] = Z;
  s->hdr.maxTreeDepth = Dmax; s->hdr.origTreeDepth = D0;
  s->hdr.X = X; s->hdr.Y = Y; s->hdr.Z = Z;
  s->hdr.numActiveNodes = numActive;
  s->geom = g;
  s->distanceMap = distanceMap;
  s->bits.assign((size_t)((numActive + 3) / 4), 0);
  for (int64_t i = 0; i < numActive; ++i) s->bits[(size_t)(i >> 2)] |= (uint8_t)(sym[(size_t)i] << (2 * (i & 3)));
  *outStream = s;
  if (iters) { iters[0] = itTotal; iters[1] = itWorst; }
  return VR_OK;
}

// byte-wise comparison of an emu_build stream with the host compressor's on the same input; frees the emu stream
int emu_build_matches_host(void* emuStream, const uint8_t* volume, long long X, long long Y, long long Z, int tolerance, int maxEpochs) {
  KdStream* a = static_cast<KdStream*>(emuStream);
  KdStream b;
  KdBuildStats st;
  int rc = kd_compress(volume, X, Y, Z, tolerance, maxEpochs, &b, &st);
  int ok = rc == VR_OK && a->hdr.numActiveNodes == b.hdr.numActiveNodes && a->distanceMap == b.distanceMap && a->bits == b.bits;
  delete a;
  return ok;
}
}
