// kd_compress.cpp -- CPU stream producer: the reference's VolumeKdtree::build restated level-wise.
//
// Follows /root/reference/volume_renderer/VolumeKdTree_recover.cpp:
//   build :17-140, buildRecursive :143-201 (midrange pyramid), compressGradientDescent :206-384,
//   encodeNodeEstimate :415-455, encodeNode :457-502, pruneTreeRecursive :596-629,
//   convertToPreorder :631-724.
// The output must be byte-identical to the reference's stream (tests/test_host_codec.py: golden containers written by the unmodified reference, and live against oracle/_ref), so
// every data-dependent decision keeps the reference's order of evaluation, including its quirks:
//   * `currentError` is never reset: it carries over between epochs and levels (SURVEY C-1) and
//     is accumulated by sequential double additions, which we keep sequential;
//   * after a reverted epoch the symbols written by the rejected distance stay in the tree;
//   * pruning uses `error < tolerance` (strict) while branch growth uses `error > tolerance`.
// Not on the timed path; the recursion of the reference is replaced by loops over
// breadth-first levels (level d occupies [2^d-1, 2^(d+1)-1), children of i are 2i+1, 2i+2).
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "kd_stream.h"

namespace vr {

namespace {

struct Enc { int value; int code; int err; };

// encodeNode (:457-502): all quantities are integers in [0,255]; ties prefer 0, then 1, then 2.
static inline Enc encode_node(int parent, int dist, int truth) {
  int noneErr = std::abs(parent - truth);
  int addEst = parent + dist > 255 ? 255 : parent + dist;
  int addErr = std::abs(addEst - truth);
  int subEst = parent - dist < 0 ? 0 : parent - dist;
  int subErr = std::abs(subEst - truth);
  int minErr = subErr < (noneErr < addErr ? noneErr : addErr) ? subErr : (noneErr < addErr ? noneErr : addErr);
  if (minErr == noneErr) return {parent, 0, minErr};
  if (minErr == addErr) return {addEst, 1, minErr};
  return {subEst, 2, minErr};
}

// encodeNodeEstimate (:415-455): running mean of |parent-truth| over nodes that want a step.
static inline void estimate_node(int parent, int truth, double* sum, double* count) {
  double nodeTruth = (double)truth, parentEstimate = (double)parent;
  double parentDistance = std::fabs(parentEstimate - nodeTruth);
  double masterDistance = (*sum + parentDistance) / (*count + 1.0);
  double noneError = parentDistance;
  double addEstimate = std::fmin(255.0, parentEstimate + masterDistance);
  double addError = std::fabs(addEstimate - nodeTruth);
  double subEstimate = std::fmax(0.0, parentEstimate - masterDistance);
  double subError = std::fabs(subEstimate - nodeTruth);
  double minError = std::fmin(subError, std::fmin(noneError, addError));
  if (minError == noneError) return;
  *sum += parentDistance;
  *count += 1.0;
}

}  // namespace

int kd_compress(const uint8_t* volume, int64_t X, int64_t Y, int64_t Z, int tolerance, int maxEpochs,
                KdStream* out, KdBuildStats* stats) {
  if (!volume || !out) return VR_ERR_ARG;
  vr_kd_geom g;
  if (vr_kd_make_geom(X, Y, Z, &g) != 0) return VR_ERR_FORMAT;
  const int D0 = g.D0, Dmax = g.Dmax;
  const int64_t numLeaves = int64_t(1) << D0;
  const int64_t numOrig = 2 * numLeaves - 1;
  const int64_t firstLeaf = numLeaves - 1;

  // ---- BUILD: midrange pyramid --------------------------------------------------------------
  // Leaf with path index p (bit D0-1-d of p = right turn at depth d) samples the voxel at the
  // min corner of its box (:194-195).
  std::vector<uint8_t> mn((size_t)numOrig), mx((size_t)numOrig), temp((size_t)numOrig);
  {
    int64_t bitOff[VR_MAX_DEPTH];
    const int64_t stride[3] = {1, X, X * Y};
    for (int d = 0; d < D0; ++d) bitOff[D0 - 1 - d] = (int64_t)g.childOff[d] * stride[g.splitDim[d]];
    const int lowBits = D0 < 12 ? D0 : 12;
    std::vector<int64_t> lowLut((size_t)1 << lowBits);
    for (int64_t i = 0; i < (int64_t(1) << lowBits); ++i) {
      int64_t o = 0;
      for (int b = 0; b < lowBits; ++b) if (i >> b & 1) o += bitOff[b];
      lowLut[(size_t)i] = o;
    }
    for (int64_t hi = 0; hi < (numLeaves >> lowBits); ++hi) {
      int64_t o = 0;
      for (int b = lowBits; b < D0; ++b) if ((hi << lowBits) >> b & 1) o += bitOff[b];
      for (int64_t lo = 0; lo < (int64_t(1) << lowBits); ++lo) {
        uint8_t s = volume[o + lowLut[(size_t)lo]];
        int64_t idx = firstLeaf + (hi << lowBits) + lo;
        mn[(size_t)idx] = mx[(size_t)idx] = temp[(size_t)idx] = s;
      }
    }
    for (int64_t i = firstLeaf - 1; i >= 0; --i) {
      uint8_t a = mn[2 * i + 1] < mn[2 * i + 2] ? mn[2 * i + 1] : mn[2 * i + 2];
      uint8_t b = mx[2 * i + 1] > mx[2 * i + 2] ? mx[2 * i + 1] : mx[2 * i + 2];
      mn[(size_t)i] = a; mx[(size_t)i] = b;
      temp[(size_t)i] = (uint8_t)(((int)a + (int)b) / 2);  // (byte)((max+min)/2.0)  (:198)
    }
  }
  std::vector<uint8_t>().swap(mn);
  std::vector<uint8_t>().swap(mx);

  // ---- COMPRESS: per-level distance by gradient descent (:206-384) ---------------------------
  std::vector<uint8_t> codes((size_t)numOrig, 0);
  std::vector<uint8_t> distanceMap((size_t)Dmax + 1, 0);
  std::vector<uint8_t> parents, recon, reconPrev;
  const double gamma = 1.25, hstep = 1.0, maxAbsStep = 4.0;
  double currentDistance = 0.0, currentError = 0.0, currentDF = 0.0, currentStepSize = 0.0;
  double previousStepSize = 0.0, previousDistance = 0.0, previousDF = 0.0, previousError = 0.0;

  for (int depth = 0; depth <= D0; ++depth) {
    const int64_t numNodes = int64_t(1) << depth;
    const int64_t start = numNodes - 1;
    const uint8_t* truth = temp.data() + start;
    uint8_t* code = codes.data() + start;
    int epoch = 0;
    double distanceSum = 0.0, distanceCount = 0.0;
    recon.resize((size_t)numNodes);

    for (int64_t i = 0; i < numNodes; ++i) {
      int parent = depth == 0 ? 0 : parents[(size_t)(i >> 1)];
      estimate_node(parent, truth[i], &distanceSum, &distanceCount);
    }
    currentDistance = distanceCount > 0 ? std::round(distanceSum / distanceCount) : 0.0;

    previousDistance = 0.0;
    previousStepSize = 255.0;
    previousError = 65025.0;
    while (epoch < maxEpochs && std::fabs(previousStepSize) >= 0.5) {
      if (epoch != 0) {
        previousDistance = currentDistance;
        previousError = currentError;
        previousDF = currentDF;
        previousStepSize = currentStepSize;
        currentDistance = std::round(std::fmin(255.0, std::fmax(0.0, previousDistance + previousStepSize)));
        if (currentDistance == previousDistance) break;
      }
      {  // evaluate error for the current distance; writes symbols and the reconstruction
        const int dist = (uint8_t)currentDistance;
        for (int64_t i = 0; i < numNodes; ++i) {
          int parent = depth == 0 ? 0 : parents[(size_t)(i >> 1)];
          Enc e = encode_node(parent, dist, truth[i]);
          code[i] = (uint8_t)e.code;
          recon[(size_t)i] = (uint8_t)e.value;
          currentError += (double)(e.err * e.err);  // sequential: currentError is not an integer
        }
        currentError /= (double)numNodes;
      }
      if (currentError < 1.0) break;
      if (epoch != 0 && currentError > previousError) {
        currentError = previousError;
        currentDistance = previousDistance;
        currentDF = previousDF;
        currentStepSize = previousStepSize / 2.0;
        recon.swap(reconPrev);
        epoch++;
        continue;
      }
      // central difference of the mean squared error (:333-365); integer sums are exact in double
      const int est[2] = {(uint8_t)std::fmax(0.0, currentDistance - hstep),
                          (uint8_t)std::fmin(255.0, currentDistance + hstep)};
      double estErr[2];
      for (int k = 0; k < 2; ++k) {
        uint64_t acc = 0;
        for (int64_t i = 0; i < numNodes; ++i) {
          int parent = depth == 0 ? 0 : parents[(size_t)(i >> 1)];
          Enc e = encode_node(parent, est[k], truth[i]);
          acc += (uint64_t)(e.err * e.err);
        }
        estErr[k] = (double)acc / (double)numNodes;
      }
      currentDF = (estErr[1] - estErr[0]) / (2.0 * hstep);
      currentStepSize = std::fmax(-maxAbsStep, std::fmin(maxAbsStep, -gamma * currentDF));
      reconPrev = recon;
      epoch++;
    }
    distanceMap[(size_t)depth] = (uint8_t)currentDistance;
    if (depth < D0) {
      parents.swap(recon);
      recon.clear();
      reconPrev.clear();
      reconPrev.shrink_to_fit();
    }
  }
  std::vector<uint8_t>().swap(parents);
  std::vector<uint8_t>().swap(reconPrev);

  const uint8_t* leafTruth = temp.data() + firstLeaf;
  if (stats) {
    int m = 0;
    for (int64_t i = 0; i < numLeaves; ++i) {
      int e = std::abs((int)leafTruth[i] - (int)recon[(size_t)i]);
      if (e > m) m = e;
    }
    stats->maxErrorBeforeGrowth = m;
  }

  // ---- PRUNE (:596-629): bottom-up; a node becomes 3 iff both children did, its own symbol is 0
  // and (leaves only) its error is strictly below the tolerance.
  {
    uint8_t* leafCode = codes.data() + firstLeaf;
    for (int64_t i = 0; i < numLeaves; ++i)
      if (leafCode[i] == 0 && std::abs((int)recon[(size_t)i] - (int)leafTruth[i]) < tolerance) leafCode[i] = 3;
    for (int64_t i = firstLeaf - 1; i >= 0; --i)
      if (codes[(size_t)i] == 0 && codes[(size_t)(2 * i + 1)] == 3 && codes[(size_t)(2 * i + 2)] == 3) codes[(size_t)i] = 3;
  }
  // a leaf pruned to 3 had code 0, so a parent test of "child == 3" equals "child was pruned";
  // interior nodes can only be 3 by pruning, leaves likewise.

  // ---- fixed refinement distances (:92-97) --------------------------------------------------
  {
    const int add[VR_ADD_LEVELS] = {64, 32, 16, 8, 4, 2, 1};
    for (int k = 0; k < VR_ADD_LEVELS; ++k) distanceMap[(size_t)(D0 + 1 + k)] = (uint8_t)add[k];
  }

  // ---- CONVERT (:631-724): preorder emission with grown refinement chains ---------------------
  std::vector<uint8_t> sym;
  sym.reserve((size_t)numLeaves * 3);
  int64_t zeroRewrites = 0;
  {
    uint64_t v = 0;
    const uint64_t total = (uint64_t)numLeaves;
    int h = D0;
    for (;;) {
      const int depth = D0 - h;
      const int64_t idx = ((int64_t(1) << depth) - 1) + (int64_t)(v >> h);
      const int c = codes[(size_t)idx];
      sym.push_back((uint8_t)c);
      if (depth < D0) {
        if (c != 3) { --h; continue; }
        v += uint64_t(1) << h;
      } else {
        if (c != 3) {
          const int64_t li = (int64_t)v;
          const int t = leafTruth[li];
          int r = recon[(size_t)li];
          int64_t zeroStart = -1;
          for (int d = D0 + 1;; ++d) {
            if (std::abs(r - t) > tolerance) {  // grow the branch: eval = true
              Enc e = encode_node(r, distanceMap[(size_t)d], t);
              r = e.value;
              if (zeroStart != -1) { if (e.code != 0) zeroStart = -1; }
              else if (e.code == 0) zeroStart = (int64_t)sym.size();
              sym.push_back((uint8_t)e.code);
              if (d >= Dmax) {
                if (zeroStart != -1) { for (size_t q = (size_t)zeroStart; q < sym.size(); ++q) sym[q] = 3; ++zeroRewrites; }
                break;
              }
            } else {  // terminator
              sym.push_back(3);
              if (zeroStart != -1) { for (size_t q = (size_t)zeroStart; q < sym.size(); ++q) sym[q] = 3; ++zeroRewrites; }
              break;
            }
          }
          recon[(size_t)li] = (uint8_t)r;
        }
        v += 1;
      }
      if (v == total) break;
      h = __builtin_ctzll(v);
    }
  }

  if (stats) {
    int m = 0;
    double s1 = 0, s2 = 0;
    for (int64_t i = 0; i < numLeaves; ++i) {
      int e = std::abs((int)leafTruth[i] - (int)recon[(size_t)i]);
      if (e > m) m = e;
      s1 += e; s2 += (double)e * e;
    }
    stats->maxError = m;
    stats->meanL1 = s1 / (double)numLeaves;
    stats->meanL2 = s2 / (double)numLeaves;
    stats->zeroRunRewrites = zeroRewrites;
  }

  // ---- pack (TwoBitArray layout) and fill the stream object -----------------------------------
  KdStream s;
  std::memset(&s.hdr, 0, sizeof(s.hdr));
  s.hdr.rootMax[0] = X; s.hdr.rootMax[1] = Y; s.hdr.rootMax[2] = Z;
  s.hdr.maxTreeDepth = Dmax;
  s.hdr.origTreeDepth = D0;
  s.hdr.X = X; s.hdr.Y = Y; s.hdr.Z = Z;
  s.hdr.numActiveNodes = (int64_t)sym.size();
  s.geom = g;
  s.distanceMap = distanceMap;
  s.bits.assign((sym.size() + 3) / 4, 0);
  for (size_t i = 0; i < sym.size(); ++i) s.bits[i >> 2] |= (uint8_t)(sym[i] << (2 * (i & 3)));
  *out = std::move(s);
  return VR_OK;
}

}  // namespace vr
