// kd_build_core.h -- element functions of the GPU stream producer (kd_build.cu), host/device so that the
// pieces whose exactness is delicate are also exercised on the CPU (tests/emu/emu_build.cpp, test harness only).
//
// The producer must emit byte-identical streams to the reference's VolumeKdtree::build
// (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:17-140; restated level-wise for the CPU in
// kd_compress.cpp, which documents the quirks).  Two steps of that algorithm are sequential as written:
//   (1) encodeNodeEstimate (:415-455) keeps a running mean over the nodes of a level in visiting order, and each
//       node's decision depends on the mean so far;
//   (2) compressGradientDescent accumulates `currentError += err*err` in a double that starts from a
//       non-integer carried over from the previous level (SURVEY C-1), so the roundings depend on the order.
// (1) is solved as a fixed point: decisions d_i = f(state before i), the state is a prefix sum over the
// decisions, and the sequential result is the only assignment consistent with its own prefix sums; iterate
// "scan, re-decide" until nothing changes.  (2): all addends are non-negative integers, so partial sums only grow
// and an addition rounds only when the partial sum enters a new binade; between two such crossings the additions
// are exact and may be done in any grouping.  vr_seq_accumulate walks a two-level hierarchy of integer block sums
// and replays element by element only inside the (at most ~40) blocks where a crossing happens.
#pragma once
#include <stdint.h>

#include "kd_common.h"

#if defined(__CUDACC__)
#include <math.h>
#else
#include <cmath>
#endif

typedef struct vr_enc {
  int value, code, err;
} vr_enc;

// encodeNode (:457-502): integers in [0,255]; ties prefer 0, then 1, then 2.
static VR_HD vr_enc vr_encode_node(int parent, int dist, int truth) {
  const int noneErr = parent > truth ? parent - truth : truth - parent;
  const int addEst = parent + dist > 255 ? 255 : parent + dist;
  const int addErr = addEst > truth ? addEst - truth : truth - addEst;
  const int subEst = parent - dist < 0 ? 0 : parent - dist;
  const int subErr = subEst > truth ? subEst - truth : truth - subEst;
  const int m0 = noneErr < addErr ? noneErr : addErr;
  const int minErr = subErr < m0 ? subErr : m0;
  vr_enc e;
  if (minErr == noneErr) { e.value = parent; e.code = 0; e.err = minErr; }
  else if (minErr == addErr) { e.value = addEst; e.code = 1; e.err = minErr; }
  else { e.value = subEst; e.code = 2; e.err = minErr; }
  return e;
}

// encodeNodeEstimate (:415-455): does this node add |parent - truth| to the running mean, given the running sum
// and count BEFORE it (both integer-valued doubles)?  Same double operations, same order, as the reference.
static VR_HD bool vr_estimate_counts(int parent, int truth, double sum, double count) {
  const double nodeTruth = (double)truth, parentEstimate = (double)parent;
  const double parentDistance = fabs(parentEstimate - nodeTruth);
  const double masterDistance = (sum + parentDistance) / (count + 1.0);
  const double noneError = parentDistance;
  const double addEstimate = fmin(255.0, parentEstimate + masterDistance);
  const double addError = fabs(addEstimate - nodeTruth);
  const double subEstimate = fmax(0.0, parentEstimate - masterDistance);
  const double subError = fabs(subEstimate - nodeTruth);
  const double minError = fmin(subError, fmin(noneError, addError));
  return !(minError == noneError);
}

// ---- (2) order-exact accumulation ------------------------------------------------------------------
// x <- x + e[0] + e[1] + ... + e[n-1] with every addition rounded as the sequential double loop does.
// blockSum[b] = sum of e over block b (VR_ACC_BLOCK elements, the last block may be short), superSum[s] = sum over
// VR_ACC_SUPER consecutive blocks.  elem(i) returns e[i] (recomputed, not stored).
#define VR_ACC_BLOCK 256
#define VR_ACC_SUPER 64

// largest double that the sum may reach without leaving the binade of x (x > 0): additions of integers to x stay
// exact below it.  For x == 0 the first non-zero addend is added exactly anyway (integers below 2^53).
static VR_HD double vr_binade_top(double x) {
  int ex;
  (void)frexp(x, &ex);     // x = f * 2^ex, f in [0.5, 1)
  return ldexp(1.0, ex);   // first value of the next binade
}

template <typename ElemFn>
static VR_HD double vr_seq_accumulate(double x, int64_t n, const uint32_t* blockSum, const uint64_t* superSum, ElemFn elem) {
  const int64_t nBlocks = (n + VR_ACC_BLOCK - 1) / VR_ACC_BLOCK;
  const int64_t nSuper = (nBlocks + VR_ACC_SUPER - 1) / VR_ACC_SUPER;
  for (int64_t s = 0; s < nSuper; ++s) {
    if (superSum[s] == 0) continue;
    const double ss = (double)superSum[s];
    // exact iff every partial sum stays inside x's binade; x == 0 or an integer-valued x far below 2^53 is exact too,
    // but the carried-over value is not an integer in general, so test the binade
    if (x > 0.0 && x + ss < vr_binade_top(x) && (x + ss) - x == ss) { x += ss; continue; }
    const int64_t b1 = (s + 1) * VR_ACC_SUPER < nBlocks ? (s + 1) * VR_ACC_SUPER : nBlocks;
    for (int64_t b = s * VR_ACC_SUPER; b < b1; ++b) {
      if (blockSum[b] == 0) continue;
      const double bs = (double)blockSum[b];
      if (x > 0.0 && x + bs < vr_binade_top(x) && (x + bs) - x == bs) { x += bs; continue; }
      const int64_t i1 = (b + 1) * VR_ACC_BLOCK < n ? (b + 1) * VR_ACC_BLOCK : n;
      for (int64_t i = b * VR_ACC_BLOCK; i < i1; ++i) x += (double)elem(i);
    }
  }
  return x;
}

// ---- refinement chain of one leaf (convertToPreorder :655-704) ---------------------------------------
// Leaf with reconstruction r, truth t and a symbol != 3: grows the chain while |r - t| > tolerance, at most 7
// levels with the fixed distances 64..1; a 3 terminates unless the 7th level was reached.  A run of 0 symbols
// at the end of the chain is rewritten to 3s together with the terminator (kept from the reference, C-13).
// Returns the number of chain symbols (1..8), symbols in sym[], the final reconstruction in *rOut, *rewrote = 1
// when the zero-run rewrite fired.
static VR_HD int vr_leaf_chain(int r, int t, int tolerance, const uint8_t* dist7, uint8_t* sym, int* rOut, int* rewrote) {
  int n = 0, zeroStart = -1;
  *rewrote = 0;
  for (int k = 0;; ++k) {
    const int ad = r > t ? r - t : t - r;
    if (ad > tolerance) {
      const vr_enc e = vr_encode_node(r, dist7[k], t);
      r = e.value;
      if (zeroStart != -1) { if (e.code != 0) zeroStart = -1; }
      else if (e.code == 0) zeroStart = n;
      sym[n++] = (uint8_t)e.code;
      if (k + 1 >= VR_ADD_LEVELS) {
        if (zeroStart != -1) { for (int q = zeroStart; q < n; ++q) sym[q] = 3; *rewrote = 1; }
        break;
      }
    } else {
      sym[n++] = 3;
      if (zeroStart != -1) { for (int q = zeroStart; q < n; ++q) sym[q] = 3; *rewrote = 1; }
      break;
    }
  }
  *rOut = r;
  return n;
}

// Voxel offset (x fastest) of the leaf with path index p: bit D0-1-d of p = right turn at depth d.
static VR_HD int64_t vr_leaf_voxel(const vr_kd_geom& g, uint32_t p, int64_t X, int64_t XY) {
  int64_t o = 0;
  for (int d = 0; d < g.D0; ++d)
    if ((p >> (g.D0 - 1 - d)) & 1u) {
      const int64_t st = g.splitDim[d] == 0 ? (int64_t)1 : (g.splitDim[d] == 1 ? X : XY);
      o += st * g.childOff[d];
    }
  return o;
}
