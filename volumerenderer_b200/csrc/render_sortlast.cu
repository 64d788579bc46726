// render_sortlast.cu -- stage 2 on a volume sharded over GPUs: every rank marches the
// samples of the reference's ray lattice that fall into its own slab, and the per-slab segment
// partials are combined front to back ("sort-last").  The reference renders one texture on one GPU
// (main.cpp:404); the arithmetic per sample is unchanged (raycaster.frag:60-77, isosurface.frag:120-157).
//
// Lattice: every rank replays p_i = p_{i-1} + dirStep from the same entry point with the same fp32
// additions (raycaster.frag:39), so sample positions, the 300-sample budget and the exit test are
// identical on all ranks; a rank only fetches at the samples it owns (dev::owned).
//
// Composite: with T = 1 - A the recurrence a = s - s*A; C += a*s; A += 0.6*a is T *= (1 - 0.6 s),
// C += T_before * s^2, so a segment that starts with transmittance T_in contributes T_in * C_seg and
// hands on T_in * T_seg.  Early ray termination (A > 0.99, raycaster.frag:77) is a property of the whole
// ray: pass 1 stops only when the segment alone is opaque (always safe), pass 2 knows T_in from the
// other ranks' T_seg and re-marches the one segment in which the ray really terminates, with the
// reference recurrence started at A_in = 1 - T_in, so that the stop sample is the reference's.
// Isosurface: the first hit over the whole ray is the hit with the smallest sample index.
#include "render_common.cuh"

namespace vr {

using namespace dev;

namespace {

__device__ __forceinline__ bool pixel_of_thread(const RenderParams& P, int* px, int* py) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  *px = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
  *py = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
  return *px < P.W && *py < P.H;
}

__device__ __forceinline__ void count_samples(unsigned long long* sampleCount, unsigned n) {
  if (sampleCount) {
    unsigned tot = __reduce_add_sync(0xFFFFFFFFu, n);
    if ((threadIdx.x & 31) == 0 && tot) atomicAdd(sampleCount, (unsigned long long)tot);
  }
}

// Reference recurrence over the owned samples of one ray, started at (C, A).  A slab is convex, so the
// owned samples are one run of the lattice: stop at the first sample after it.
template <bool OCC>
__device__ __forceinline__ void march_segment(const Vol& v, const OccView& occ, V3 p, V3 dirStep, float* C, float* A, unsigned* n) {
  float c = *C, a = *A;
  const V3 inv = recip(dirStep);
  bool started = false;
  const int i0 = first_owned(v, p, dirStep);
  for (int j = 0; j < i0 && i0 < 300; ++j) p = add(p, dirStep);  // lattice replay up to the slab, additions only
  for (int i = i0; i < 300; ++i) {
    p = add(p, dirStep);
    if (!inside(p)) break;
    if (!owned(v, p)) {
      if (started) break;
      continue;
    }
    started = true;
    if (OCC) {  // void cell: exactly-zero samples change nothing (skipped positions need no ownership test either)
      const int k = occ_skip<true>(occ, v, p, dirStep, inv, 299 - i);
      if (k >= 0) {
        for (int j = 0; j < k; ++j) p = add(p, dirStep);
        i += k; *n += 1 + k;
        continue;
      }
    }
    float smp = tex<true>(v, p);
    ++*n;
    float prev_alpha = fsub(smp, fmul(smp, a));
    c = fadd(fmul(prev_alpha, smp), c);
    a = fadd(a, fmul(prev_alpha, 0.6f));
    if (a > 0.99f) break;
  }
  *C = c; *A = a;
}

template <bool OCC>
__device__ __forceinline__ void partial_pixel(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, int px, int py,
                                              float* C, float* T, unsigned* n) {
  float c = 0.0f, a = 0.0f;
  V3 vUV;
  if (ray_setup(P, px, py, &vUV)) {
    Vol v = make_vol(g, vol);
    V3 dir, dirStep;
    ray_dir(P, vUV, &dir, &dirStep);
    march_segment<OCC>(v, occ, vUV, dirStep, &c, &a, n);
  }
  *C = c;
  *T = fsub(1.0f, a);
}

template <bool OCC>
__global__ void __launch_bounds__(256) partial_composite_kernel(const __grid_constant__ RenderParams P,
                                                                const __grid_constant__ SlabGeom g,
                                                                const __grid_constant__ OccView occ,
                                                                const uint8_t* __restrict__ vol, float* __restrict__ cseg,
                                                                float* __restrict__ tseg,
                                                                unsigned long long* __restrict__ sampleCount) {
  int px, py;
  unsigned n = 0;
  if (pixel_of_thread(P, &px, &py)) {
    float C, T;
    partial_pixel<OCC>(P, g, occ, vol, px, py, &C, &T, &n);
    cseg[(long long)py * P.W + px] = C;
    tseg[(long long)py * P.W + px] = T;
  }
  count_samples(sampleCount, n);
}

// Last CTA of the grid publishes: every CTA has fenced its stores (system scope: they go to peer memory) before it
// takes a ticket, so whoever draws the last ticket knows all of them are visible before the flags are.
// only >= 0: raise the flag at that one rank (a result that a single rank consumes)
__device__ __forceinline__ void peer_signal(const PeerTargets& t, int only = -1) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned total = gridDim.x * gridDim.y;
    if (atomicAdd(t.ticket, 1u) == total - 1) {
      *t.ticket = 0;
      __threadfence_system();
      for (int r = 0; r < t.n; ++r) {
        if (only >= 0 && r != only) continue;
        const uint32_t v = *t.sent[r] + 1u;
        *t.sent[r] = v;
        *reinterpret_cast<volatile uint32_t*>(t.flag[r]) = v;
      }
    }
  }
}

template <bool OCC>
__global__ void __launch_bounds__(256) partial_composite_peer_kernel(const __grid_constant__ RenderParams P,
                                                                     const __grid_constant__ SlabGeom g,
                                                                     const __grid_constant__ OccView occ,
                                                                     const uint8_t* __restrict__ vol, float* __restrict__ cseg,
                                                                     const __grid_constant__ PeerTargets t, long long gatherOff,
                                                                     unsigned long long* __restrict__ sampleCount) {
  int px, py;
  unsigned n = 0;
  if (pixel_of_thread(P, &px, &py)) {
    float C, T;
    partial_pixel<OCC>(P, g, occ, vol, px, py, &C, &T, &n);
    const long long pix = (long long)py * P.W + px, npix = (long long)P.W * P.H;
    cseg[pix] = C;
    for (int r = 0; r < t.n; ++r) reinterpret_cast<float*>(t.base[r] + gatherOff)[(long long)t.rank * npix + pix] = T;
  }
  count_samples(sampleCount, n);
  peer_signal(t);
}

template <bool OCC>
__global__ void __launch_bounds__(256) partial_iso_kernel(const __grid_constant__ RenderParams P,
                                                          const __grid_constant__ SlabGeom g,
                                                          const __grid_constant__ OccView occ, const uint8_t* __restrict__ vol,
                                                          float4* __restrict__ rgbaOut, int* __restrict__ hitOut,
                                                          unsigned long long* __restrict__ sampleCount) {
  int px, py;
  unsigned n = 0;
  if (pixel_of_thread(P, &px, &py)) {
    float4 rgba = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    int hit = VR_HIT_NONE;
    V3 vUV;
    if (ray_setup(P, px, py, &vUV)) {
      Vol v = make_vol(g, vol);
      V3 dir, dirStep;
      ray_dir(P, vUV, &dir, &dirStep);
      const V3 inv = recip(dirStep);
      V3 p = vUV;
      bool started = false;
      const int i0 = first_owned(v, p, dirStep);
      for (int j = 0; j < i0 && i0 < 300; ++j) p = add(p, dirStep);
      for (int i = i0; i < 300; ++i) {
        p = add(p, dirStep);
        if (!inside(p)) break;
        if (!owned(v, p)) {
          if (started) break;
          continue;
        }
        started = true;
        const V3 pn = add(p, dirStep);
        if (OCC) {
          const int k = occ_skip<true>(occ, v, pn, dirStep, inv, 299 - i);
          if (k >= 0) {
            for (int j = 0; j < k; ++j) p = add(p, dirStep);
            i += k; n += 1 + k;
            continue;
          }
        }
        float smp = tex<true>(v, p);
        float smp2 = tex<true>(v, pn);
        ++n;
        if (fsub(smp, P.iso) < 0.0f && fsub(smp2, P.iso) >= 0.0f) {
          rgba = iso_shade<true>(v, P, p, dir, dirStep);
          hit = i;
          break;
        }
      }
    }
    rgbaOut[(long long)py * P.W + px] = rgba;
    hitOut[(long long)py * P.W + px] = hit;
  }
  count_samples(sampleCount, n);
}

template <bool OCC>
__device__ __forceinline__ float2 resolve_pixel(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* __restrict__ vol,
                                                const float* __restrict__ tsegAll, int nRanks, int rank, const int* __restrict__ order,
                                                const float* __restrict__ cseg, int px, int py, unsigned& n) {
    const long long pix = (long long)py * P.W + px, npix = (long long)P.W * P.H;
    float2 term = make_float2(0.0f, 0.0f);
    V3 vUV;
    if (ray_setup(P, px, py, &vUV)) {
      V3 dir, dirStep;
      ray_dir(P, vUV, &dir, &dirStep);
      // visibility order of the slabs for this ray's direction octant
      const int oct = (dirStep.x < 0.0f ? 1 : 0) | (dirStep.y < 0.0f ? 2 : 0) | (dirStep.z < 0.0f ? 4 : 0);
      const int* ord = order + oct * nRanks;
      // walk the segments front to back: every rank evaluates the same chain in the same order
      float T = 1.0f, Tin = 1.0f;
      int myPos = -1, crossPos = -1;
      for (int k = 0; k < nRanks; ++k) {
        const int r = __ldg(ord + k);
        if (r == rank) { myPos = k; Tin = T; }
        T = fmul(T, __ldg(tsegAll + (long long)r * npix + pix));
        if (crossPos < 0 && fsub(1.0f, T) > 0.99f) crossPos = k;
      }
      if (crossPos >= 0 && crossPos < myPos) {
        // the ray terminated in a slab in front of this one
      } else if (crossPos == myPos && Tin == 1.0f) {
        // nothing in front of this slab: pass 1 already ran the reference recurrence from A = 0
        term = make_float2(__ldg(cseg + pix), fsub(1.0f, __ldg(tsegAll + (long long)rank * npix + pix)));
      } else if (crossPos == myPos) {
        // the ray terminates inside this segment: replay it with the reference recurrence from A_in
        Vol v = make_vol(g, vol);
        float C = 0.0f, A = fsub(1.0f, Tin);
        march_segment<OCC>(v, occ, vUV, dirStep, &C, &A, &n);
        term = make_float2(C, A);
      } else {
        term.x = fmul(Tin, __ldg(cseg + pix));
        if (crossPos < 0 && myPos == nRanks - 1) term.y = fsub(1.0f, T);  // no termination anywhere: last slab closes alpha
      }
    }
    return term;
}

template <bool OCC>
__global__ void __launch_bounds__(256) resolve_composite_kernel(const __grid_constant__ RenderParams P,
                                                                const __grid_constant__ SlabGeom g,
                                                                const __grid_constant__ OccView occ,
                                                                const uint8_t* __restrict__ vol, const float* __restrict__ tsegAll,
                                                                int nRanks, int rank, const int* __restrict__ order,
                                                                const float* __restrict__ cseg, float2* __restrict__ contrib,
                                                                unsigned long long* __restrict__ sampleCount) {
  int px, py;
  unsigned n = 0;
  if (pixel_of_thread(P, &px, &py)) contrib[(long long)py * P.W + px] = resolve_pixel<OCC>(P, g, occ, vol, tsegAll, nRanks, rank, order, cseg, px, py, n);
  count_samples(sampleCount, n);
}

template <bool OCC>
__global__ void __launch_bounds__(256) resolve_composite_peer_kernel(const __grid_constant__ RenderParams P,
                                                                     const __grid_constant__ SlabGeom g,
                                                                     const __grid_constant__ OccView occ,
                                                                     const uint8_t* __restrict__ vol, const float* __restrict__ tsegAll,
                                                                     const int* __restrict__ order, const float* __restrict__ cseg,
                                                                     const __grid_constant__ PeerTargets t, long long redOff,
                                                                     unsigned long long* __restrict__ sampleCount) {
  int px, py;
  unsigned n = 0;
  if (pixel_of_thread(P, &px, &py)) {
    const float2 term = resolve_pixel<OCC>(P, g, occ, vol, tsegAll, t.n, t.rank, order, cseg, px, py, n);
    const long long pix = (long long)py * P.W + px, stripe = ((long long)P.W * P.H) / t.n;
    const long long j = pix / stripe;  // owner of the pixel's stripe
    reinterpret_cast<float2*>(t.base[j] + redOff)[(long long)t.rank * stripe + (pix - j * stripe)] = term;
  }
  count_samples(sampleCount, n);
  peer_signal(t);
}

// ---- the all-gather of T_seg replaced by two stripe exchanges ---------------------------------------------------------
// What a rank needs from the other ranks' T_seg is one number per pixel: the transmittance T_in in front of its own
// segment (or the fact that the ray has terminated in front of it).  So instead of every rank storing its whole T_seg into
// every window (n - 1 image-sized transfers per rank), stripe j of every rank's T_seg goes to rank j, which walks the
// segments of its pixels front to back exactly as resolve_pixel does (same order table, same fp32 products) and stores to
// every rank r the T_in of r's segment: 2 (n - 1) / n image-sized transfers per rank in total.
// Encoding: T_in >= 0 = the transmittance in front of the segment; -1 = the ray terminated in a slab in front.
__global__ void __launch_bounds__(256) tin_stripe_peer_kernel(const __grid_constant__ RenderParams P, const int* __restrict__ order,
                                                              const __grid_constant__ PeerTargets t, long long gatherOff, long long tinOff) {
  const long long stripe = ((long long)P.W * P.H) / t.n;
  const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < stripe) {
    const long long pix = (long long)t.rank * stripe + k;
    const float* in = reinterpret_cast<const float*>(t.base[t.rank] + gatherOff);  // n stripes, rank-major
    V3 vUV;
    if (ray_setup(P, (int)(pix % P.W), (int)(pix / P.W), &vUV)) {
      V3 dir, dirStep;
      ray_dir(P, vUV, &dir, &dirStep);
      const int oct = (dirStep.x < 0.0f ? 1 : 0) | (dirStep.y < 0.0f ? 2 : 0) | (dirStep.z < 0.0f ? 4 : 0);
      const int* ord = order + oct * t.n;
      float T = 1.0f;
      bool crossed = false;
      for (int q = 0; q < t.n; ++q) {
        const int r = __ldg(ord + q);
        reinterpret_cast<float*>(t.base[r] + tinOff)[pix] = crossed ? -1.0f : T;
        T = fmul(T, in[(long long)r * stripe + k]);
        if (!crossed && fsub(1.0f, T) > 0.99f) crossed = true;
      }
    } else {
      for (int r = 0; r < t.n; ++r) reinterpret_cast<float*>(t.base[r] + tinOff)[pix] = -1.0f;  // the ray misses the cube: no terms
    }
  }
  peer_signal(t);
}

// resolve_pixel with T_in given instead of all ranks' T_seg (same decisions, same arithmetic)
template <bool OCC>
__global__ void __launch_bounds__(256) resolve_composite_tin_kernel(const __grid_constant__ RenderParams P,
                                                                    const __grid_constant__ SlabGeom g,
                                                                    const __grid_constant__ OccView occ,
                                                                    const uint8_t* __restrict__ vol, const float* __restrict__ tin,
                                                                    const float* __restrict__ tseg, int nRanks, int rank,
                                                                    const int* __restrict__ order, const float* __restrict__ cseg,
                                                                    float2* __restrict__ contrib,
                                                                    unsigned long long* __restrict__ sampleCount) {
  int px, py;
  unsigned n = 0;
  if (pixel_of_thread(P, &px, &py)) {
    const long long pix = (long long)py * P.W + px;
    float2 term = make_float2(0.0f, 0.0f);
    const float Tin = __ldg(tin + pix);
    V3 vUV;
    if (Tin >= 0.0f && ray_setup(P, px, py, &vUV)) {
      V3 dir, dirStep;
      ray_dir(P, vUV, &dir, &dirStep);
      const float Tseg = __ldg(tseg + pix);
      const float Tout = fmul(Tin, Tseg);
      if (fsub(1.0f, Tout) > 0.99f) {           // the ray terminates inside this segment
        if (Tin == 1.0f) {                      // nothing in front: pass 1 already ran the reference recurrence from A = 0
          term = make_float2(__ldg(cseg + pix), fsub(1.0f, Tseg));
        } else {                                // replay it with the reference recurrence from A_in
          Vol v = make_vol(g, vol);
          float C = 0.0f, A = fsub(1.0f, Tin);
          march_segment<OCC>(v, occ, vUV, dirStep, &C, &A, &n);
          term = make_float2(C, A);
        }
      } else {
        term.x = fmul(Tin, __ldg(cseg + pix));
        const int oct = (dirStep.x < 0.0f ? 1 : 0) | (dirStep.y < 0.0f ? 2 : 0) | (dirStep.z < 0.0f ? 4 : 0);
        if (__ldg(order + oct * nRanks + nRanks - 1) == rank) term.y = fsub(1.0f, Tout);  // no termination anywhere: last slab closes alpha
      }
    }
    contrib[pix] = term;
  }
  count_samples(sampleCount, n);
}

// One thread per pixel of this rank's stripe: the ranks' terms in rank order, then the colour map (finish_composite_kernel).
__global__ void __launch_bounds__(256) finish_composite_peer_kernel(const __grid_constant__ RenderParams P,
                                                                    const __grid_constant__ PeerTargets t, long long redOff, long long frameOff,
                                                                    int dstOnly) {
  const long long stripe = ((long long)P.W * P.H) / t.n;
  const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < stripe) {
    const float2* in = reinterpret_cast<const float2*>(t.base[t.rank] + redOff);
    float2 s = in[k];
    for (int r = 1; r < t.n; ++r) {
      const float2 b = in[(long long)r * stripe + k];
      s.x = fadd(s.x, b.x);
      s.y = fadd(s.y, b.y);
    }
    const long long pix = (long long)t.rank * stripe + k;
    float4 rgba = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    V3 vUV;
    if (ray_setup(P, (int)(pix % P.W), (int)(pix / P.W), &vUV))
      rgba = make_float4(clamp01(fsub(1.0f, s.x)), clamp01(fsub(1.0f, s.x)), 1.0f, clamp01(s.y));
    if (dstOnly >= 0) reinterpret_cast<float4*>(t.base[dstOnly] + frameOff)[pix] = rgba;
    else for (int r = 0; r < t.n; ++r) reinterpret_cast<float4*>(t.base[r] + frameOff)[pix] = rgba;
  }
  peer_signal(t, dstOnly);
}

__global__ void __launch_bounds__(256) finish_composite_kernel(const __grid_constant__ RenderParams P,
                                                               const float2* __restrict__ sum, float4* __restrict__ out) {
  int px, py;
  if (pixel_of_thread(P, &px, &py)) {
    const long long pix = (long long)py * P.W + px;
    float4 rgba = make_float4(1.0f, 1.0f, 1.0f, 1.0f);  // clear colour (main.cpp:392)
    V3 vUV;
    if (ray_setup(P, px, py, &vUV)) {
      float2 s = sum[pix];
      rgba = make_float4(clamp01(fsub(1.0f, s.x)), clamp01(fsub(1.0f, s.x)), 1.0f, clamp01(s.y));  // raycaster.frag:82-85
    }
    out[pix] = rgba;
  }
}

__global__ void __launch_bounds__(256) resolve_iso_kernel(long long npix, int rank, const int* __restrict__ hitMine,
                                                          const int* __restrict__ hitMin, float4* __restrict__ rgba) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const int hm = hitMin[i], h = hitMine[i];
  if (hm == VR_HIT_NONE) {
    const float w = rank == 0 ? 1.0f : 0.0f;  // miss: vec4(255,255,255,1) clamped (isosurface.frag:79), supplied once
    rgba[i] = make_float4(w, w, w, w);
  } else if (h != hm) {
    rgba[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  }
}

dim3 pixel_grid(const RenderParams& P) { return dim3((P.W + 31) / 32, (P.H + 7) / 8); }

}  // namespace

cudaError_t partial_composite_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, float* cseg,
                                     float* tseg, unsigned long long* sampleCount, cudaStream_t stream) {
  if (occ.pyr) partial_composite_kernel<true><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, cseg, tseg, sampleCount);
  else partial_composite_kernel<false><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, cseg, tseg, sampleCount);
  return cudaGetLastError();
}
cudaError_t partial_iso_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, float4* rgba, int* hit,
                               unsigned long long* sampleCount, cudaStream_t stream) {
  if (occ.pyr) partial_iso_kernel<true><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, rgba, hit, sampleCount);
  else partial_iso_kernel<false><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, rgba, hit, sampleCount);
  return cudaGetLastError();
}
cudaError_t resolve_composite_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol,
                                     const float* tsegAll, int n, int rank, const int* order, const float* cseg, float2* contrib,
                                     unsigned long long* sampleCount, cudaStream_t stream) {
  if (occ.pyr) resolve_composite_kernel<true><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, tsegAll, n, rank, order, cseg, contrib, sampleCount);
  else resolve_composite_kernel<false><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, tsegAll, n, rank, order, cseg, contrib, sampleCount);
  return cudaGetLastError();
}
cudaError_t partial_composite_peer_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, float* cseg,
                                          const PeerTargets& t, long long gatherOff, unsigned long long* sampleCount, cudaStream_t stream) {
  if (occ.pyr) partial_composite_peer_kernel<true><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, cseg, t, gatherOff, sampleCount);
  else partial_composite_peer_kernel<false><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, cseg, t, gatherOff, sampleCount);
  return cudaGetLastError();
}
cudaError_t resolve_composite_peer_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol,
                                          const float* tsegAll, const int* order, const float* cseg, const PeerTargets& t,
                                          long long redOff, unsigned long long* sampleCount, cudaStream_t stream) {
  if (occ.pyr) resolve_composite_peer_kernel<true><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, tsegAll, order, cseg, t, redOff, sampleCount);
  else resolve_composite_peer_kernel<false><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, tsegAll, order, cseg, t, redOff, sampleCount);
  return cudaGetLastError();
}
cudaError_t finish_composite_peer_launch(const RenderParams& P, const PeerTargets& t, long long redOff, long long frameOff, int dstOnly,
                                         cudaStream_t stream) {
  const long long stripe = ((long long)P.W * P.H) / t.n;
  finish_composite_peer_kernel<<<dim3((unsigned)((stripe + 255) / 256), 1), 256, 0, stream>>>(P, t, redOff, frameOff, dstOnly);
  return cudaGetLastError();
}
cudaError_t tin_stripe_peer_launch(const RenderParams& P, const int* order, const PeerTargets& t, long long gatherOff, long long tinOff,
                                   cudaStream_t stream) {
  const long long stripe = ((long long)P.W * P.H) / t.n;
  tin_stripe_peer_kernel<<<dim3((unsigned)((stripe + 255) / 256), 1), 256, 0, stream>>>(P, order, t, gatherOff, tinOff);
  return cudaGetLastError();
}
cudaError_t resolve_composite_tin_launch(const RenderParams& P, const SlabGeom& g, const OccView& occ, const uint8_t* vol, const float* tin,
                                         const float* tseg, int n, int rank, const int* order, const float* cseg, float2* contrib,
                                         unsigned long long* sampleCount, cudaStream_t stream) {
  if (occ.pyr) resolve_composite_tin_kernel<true><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, tin, tseg, n, rank, order, cseg, contrib, sampleCount);
  else resolve_composite_tin_kernel<false><<<pixel_grid(P), 256, 0, stream>>>(P, g, occ, vol, tin, tseg, n, rank, order, cseg, contrib, sampleCount);
  return cudaGetLastError();
}
cudaError_t finish_composite_launch(const RenderParams& P, const float2* sum, float4* out, cudaStream_t stream) {
  finish_composite_kernel<<<pixel_grid(P), 256, 0, stream>>>(P, sum, out);
  return cudaGetLastError();
}
cudaError_t resolve_iso_launch(int W, int H, int rank, const int* hitMine, const int* hitMin, float4* rgba, cudaStream_t stream) {
  long long npix = (long long)W * H;
  resolve_iso_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, stream>>>(npix, rank, hitMine, hitMin, rgba);
  return cudaGetLastError();
}

}  // namespace vr
