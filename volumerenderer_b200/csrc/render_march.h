// render_march.h -- launch interface of the ray-marching kernels (render_march.cu, render_sortlast.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "occupancy.h"

namespace vr {
// Same field order as the oracle's Params (oracle/render_oracle.cpp): camera basis precomputed on the host.
struct RenderParams {
  float eye[3], s[3], u[3], f[3];
  float tw, th, znear, zfar;
  float step[3];
  float iso;
  int W, H, mode;  // mode 0 composite, 1 isosurface
  int skipCountMask;  // -1: skipped lattice positions count as samples (the reference's march steps); 0: only fetching steps
                      // count (vr_debug_set("count_fetched", 1); plain march only)
};
// occ.dil == null: march every sample; otherwise skip void cells (occupancy.h), same frame bit for bit
cudaError_t render_launch(const RenderParams& P, const OccView& occ, const uint8_t* vol, int X, int Y, int Z, float4* out,
                          int* hitIndex, unsigned long long* sampleCount, cudaStream_t stream);

// ---- sort-last: one device slab of a volume sharded over ranks -----------------------
// G = extent of the whole texture, o/S = origin and extent of the stored box (owned box + halo),
// [lo, hi) = owned box of base texels.
struct SlabGeom { int G[3], o[3], S[3], lo[3], hi[3]; };
#define VR_MAX_RANKS 64
#define VR_HIT_NONE 0x7FFFFFFF
// pass 1, composite: per pixel the segment partial (C_seg, T_seg) over the samples this slab owns
cudaError_t partial_composite_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, float* cseg,
                                     float* tseg, unsigned long long* sampleCount, cudaStream_t stream);
// pass 1, isosurface: first hit among the owned samples: colour (0 if none) and global sample index (VR_HIT_NONE)
cudaError_t partial_iso_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, float4* rgba, int* hit,
                               unsigned long long* sampleCount, cudaStream_t stream);
// pass 2, composite: tsegAll = n x (W*H) segment transmittances of all ranks; order = 8 x n visibility
// orders (device), one per direction octant.  contrib = per pixel (C, A) term of this rank; the sum of
// the terms over ranks is the frame's (C, A).
cudaError_t resolve_composite_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol,
                                     const float* tsegAll, int n, int rank, const int* order, const float* cseg, float2* contrib,
                                     unsigned long long* sampleCount, cudaStream_t stream);
// colour map of raycaster.frag:82-85 on the summed terms; white where the view ray misses the cube
cudaError_t finish_composite_launch(const RenderParams& P, const float2* sum, float4* out, cudaStream_t stream);
// pass 2, isosurface: keep this rank's colour only where its hit is the frame's first hit (hitMin =
// element-wise min over ranks); rank 0 supplies white where no rank hit.  Sum over ranks = the frame.
cudaError_t resolve_iso_launch(int W, int H, int rank, const int* hitMine, const int* hitMin, float4* rgba, cudaStream_t stream);

// ---- the same passes fused with their exchange step over peer-memory windows (vr_peer, include/vr_b200.h) --------
// A kernel stores its per-pixel result straight into the windows of the ranks that need it (NVLink stores from the
// compute kernel, no staging buffer, no copy kernel) and its last CTA raises the arrival flags.
#define VR_MAX_PEERS 32
struct PeerTargets {
  uint8_t* base[VR_MAX_PEERS];    // every rank's window as mapped here
  uint32_t* flag[VR_MAX_PEERS];   // arrival counter (slot, this rank) inside every rank's flag area
  uint32_t* sent[VR_MAX_PEERS];   // this rank's count of signals to that rank on this slot (device-side epoch)
  unsigned int* ticket;           // local CTA ticket
  int n, rank;
};
// pass 1: T_seg of this rank -> window[r] + gatherOff + rank * 4 * W*H for every r; C_seg stays local
cudaError_t partial_composite_peer_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, float* cseg,
                                          const PeerTargets& t, long long gatherOff, unsigned long long* sampleCount, cudaStream_t stream);
// pass 2: this rank's (C, A) term of pixel p -> stripe owner j = p / (W*H/n): window[j] + redOff + rank * 8 * W*H/n
cudaError_t resolve_composite_peer_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol,
                                          const float* tsegAll, const int* order, const float* cseg, const PeerTargets& t,
                                          long long redOff, unsigned long long* sampleCount, cudaStream_t stream);
// T_in exchange (replaces the all-gather of T_seg): the stripe owner walks its pixels' segments front to back and stores to every
// rank r the transmittance in front of r's segment (-1: terminated in front) at window[r] + tinOff; then the pass-2 kernel on T_in
cudaError_t tin_stripe_peer_launch(const RenderParams& P, const int* order, const PeerTargets& t, long long gatherOff, long long tinOff,
                                   cudaStream_t stream);
cudaError_t resolve_composite_tin_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, const float* tin,
                                         const float* tseg, int n, int rank, const int* order, const float* cseg, float2* contrib,
                                         unsigned long long* sampleCount, cudaStream_t stream);
// pass 3 (stripe owner): sum of the n terms in rank order + colour map -> RGBA of the stripe into every window + frameOff
cudaError_t finish_composite_peer_launch(const RenderParams& P, const PeerTargets& t, long long redOff, long long frameOff, int dstOnly,
                                         cudaStream_t stream);
}  // namespace vr
