// vr_capi_frame.cu -- the whole sort-last frame (and the halo exchange of a decoded timestep) as ONE call each, for hosts
// that are not Python: the renderer-mode surface of the reference is one Draw() per frame (main.cpp:404), and a C++ host
// of the sharded renderer should not have to re-implement the exchange protocol of volumerenderer_b200/sortlast.py (slot
// numbers, window layout, which wait follows which put).  Everything here is a sequence of the public entry points of
// include/vr_b200.h on the context's stream -- the same sequence sortlast.PeerExchange issues, so frames are bit-identical.
//
// Window layout (bytes; P = width * height, n ranks), identical on all ranks:
//   [0, 4Pn)            gather: this rank's stripe of the n ranks' T_seg in its first 4P bytes (composite) -- slot 0
//   [4Pn, 4Pn + 16P)    stripes-in: n stripes of this rank's stripe of the terms being combined -- slot 1; composite terms use
//                       8P of it, the other half holds T_in (slot 4)
//   [.., + 16P)         total: the combined buffer / the finished RGBA frame -- slot 2
//   [.., + B * halo)    B halo staging areas -- slot 3
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <vector>

#include "../../include/vr_b200.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)
#define RC(call) do { int rc_ = (call); if (rc_ != VR_OK) return rc_; } while (0)

namespace {
enum { SLOT_GATHER = 0, SLOT_RED_IN = 1, SLOT_RED_OUT = 2, SLOT_HALO = 3, SLOT_TIN = 4 };
inline int64_t pad16(int64_t v) { return (v + 15) / 16 * 16; }

struct HaloBox { int src, dst; int64_t lo[3], ext[3], bytes; };

// stored box of rank r for a halo of h voxels per side: owned box grown by h, clipped to the texture
void stored_box(const vr_slab_desc& d, const int64_t h[3], int64_t o[3], int64_t e[3]) {
  for (int a = 0; a < 3; ++a) {
    o[a] = std::max<int64_t>(0, d.own_lo[a] - h[a]);
    e[a] = std::min<int64_t>(d.global[a], d.own_hi[a] + h[a]) - o[a];
  }
}
// for every ordered pair (src, dst): the part of dst's stored box that src owns (sortlast.halo_transfers)
std::vector<HaloBox> halo_boxes(const std::vector<vr_slab_desc>& s, const int64_t h[3]) {
  std::vector<HaloBox> out;
  const int n = (int)s.size();
  for (int dst = 0; dst < n; ++dst) {
    int64_t o[3], e[3];
    stored_box(s[(size_t)dst], h, o, e);
    for (int src = 0; src < n; ++src) {
      if (src == dst) continue;
      HaloBox b{src, dst, {0, 0, 0}, {0, 0, 0}, 0};
      bool any = true;
      for (int a = 0; a < 3; ++a) {
        const int64_t lo = std::max(o[a], s[(size_t)src].own_lo[a]), hi = std::min(o[a] + e[a], s[(size_t)src].own_hi[a]);
        if (hi <= lo) any = false;
        b.lo[a] = lo; b.ext[a] = hi - lo;
      }
      if (!any) continue;
      b.bytes = pad16(b.ext[0] * b.ext[1] * b.ext[2]);
      out.push_back(b);
    }
  }
  return out;
}
}  // namespace

struct vr_sortlast_frame {
  vr_context* ctx = nullptr;
  vr_peer* peer = nullptr;
  vr_sortlast* sl = nullptr;
  int n = 0, rank = 0, W = 0, H = 0;
  int64_t P = 0, halo[3] = {0, 0, 0};
  int haloBuffers = 1;
  int64_t offGather = 0, offRedIn = 0, offTotal = 0, offHalo = 0, haloBytes = 0;
  std::vector<HaloBox> boxes;
  uint8_t* base = nullptr;  // this rank's window as mapped here
  float *cseg = nullptr, *tseg = nullptr, *contrib = nullptr, *rgba = nullptr, *red = nullptr;
  int32_t* hit = nullptr;
  uint8_t* pack = nullptr;  // packed halo boxes this rank sends, one area per staging buffer
  int64_t packBytes = 0;
};

namespace {
int all_ranks(vr_sortlast_frame* f, std::vector<int32_t>* v) {
  v->resize((size_t)f->n);
  for (int r = 0; r < f->n; ++r) (*v)[(size_t)r] = r;
  return VR_OK;
}
// all-reduce of nbytes (a multiple of 16 n) from src into the window's total area: reduce-scatter by stores, combine in rank
// order, stripes to every window (sortlast.PeerExchange.stages)
int all_reduce(vr_sortlast_frame* f, int op, const void* src, int64_t nbytes) {
  const int n = f->n;
  if (nbytes % (16 * n)) return fail(VR_ERR_ARG, "vr_sortlast_render: width * height must split into 16-byte aligned stripes per rank");
  const int64_t sb = nbytes / n;
  std::vector<int32_t> everyone;
  all_ranks(f, &everyone);
  std::vector<int64_t> off((size_t)n), bytes((size_t)n, sb);
  std::vector<const void*> ptr((size_t)n);
  for (int j = 0; j < n; ++j) { off[(size_t)j] = f->offRedIn + f->rank * sb; ptr[(size_t)j] = (const uint8_t*)src + j * sb; }
  RC(vr_peer_put(f->peer, SLOT_RED_IN, n, everyone.data(), off.data(), ptr.data(), bytes.data()));
  RC(vr_peer_wait(f->peer, SLOT_RED_IN, n, everyone.data()));
  for (int k = 0; k < n; ++k) ptr[(size_t)k] = f->base + f->offRedIn + k * sb;
  RC(vr_peer_reduce(f->peer, op, n, ptr.data(), f->red, sb / 4));
  for (int r = 0; r < n; ++r) { off[(size_t)r] = f->offTotal + f->rank * sb; ptr[(size_t)r] = f->red; }
  RC(vr_peer_put(f->peer, SLOT_RED_OUT, n, everyone.data(), off.data(), ptr.data(), bytes.data()));
  RC(vr_peer_wait(f->peer, SLOT_RED_OUT, n, everyone.data()));
  return VR_OK;
}
}  // namespace

extern "C" {

int vr_sortlast_frame_window_bytes(const vr_slab_desc* slabs, int32_t n, int32_t width, int32_t height, const int64_t halo[3],
                                   int32_t halo_buffers, int64_t* out) {
  if (!slabs || n < 1 || width <= 0 || height <= 0 || !halo || halo_buffers < 1 || !out) return fail(VR_ERR_ARG, "vr_sortlast_frame_window_bytes: bad argument");
  const std::vector<vr_slab_desc> s(slabs, slabs + n);
  std::vector<int64_t> need((size_t)n, 0);
  for (const HaloBox& b : halo_boxes(s, halo)) need[(size_t)b.dst] += b.bytes;
  const int64_t P = (int64_t)width * height, hb = *std::max_element(need.begin(), need.end());
  *out = 4 * P * n + 16 * P + 16 * P + pad16(hb) * halo_buffers;
  return VR_OK;
}

int vr_sortlast_frame_create(vr_context* ctx, vr_peer* peer, vr_sortlast* sl, int32_t width, int32_t height, const int64_t halo[3],
                             int32_t halo_buffers, vr_sortlast_frame** out) {
  if (!ctx || !peer || !sl || width <= 0 || height <= 0 || !halo || halo_buffers < 1 || !out) return fail(VR_ERR_ARG, "vr_sortlast_frame_create: bad argument");
  vr_sortlast_frame* f = new vr_sortlast_frame;
  f->ctx = ctx; f->peer = peer; f->sl = sl;
  f->n = sl->n; f->rank = sl->rank; f->W = width; f->H = height; f->P = (int64_t)width * height;
  f->haloBuffers = halo_buffers;
  for (int a = 0; a < 3; ++a) f->halo[a] = halo[a];
  f->boxes = halo_boxes(sl->slabs, halo);
  std::vector<int64_t> need((size_t)f->n, 0);
  int64_t mine = 0;
  for (const HaloBox& b : f->boxes) { need[(size_t)b.dst] += b.bytes; if (b.src == f->rank) mine += b.bytes; }
  f->haloBytes = pad16(*std::max_element(need.begin(), need.end()));
  f->offGather = 0;
  f->offRedIn = 4 * f->P * f->n;
  f->offTotal = f->offRedIn + 16 * f->P;
  f->offHalo = f->offTotal + 16 * f->P;
  void* w = nullptr;
  int rc = vr_peer_window(peer, f->rank, &w);
  if (rc != VR_OK) { delete f; return rc; }
  f->base = (uint8_t*)w;
  if ((4 * f->P) % (16 * f->n)) { delete f; return fail(VR_ERR_ARG, "vr_sortlast_frame_create: width * height must split into 16-byte aligned stripes per rank"); }
  cudaError_t e = cudaSetDevice(ctx->device);
  const size_t P = (size_t)f->P;
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->cseg, P * 4);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->tseg, P * 4);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->contrib, P * 8);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->rgba, P * 16);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->hit, P * 4);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->red, P * 16 / (size_t)f->n + 16);
  f->packBytes = mine;
  if (e == cudaSuccess && mine) e = cudaMalloc((void**)&f->pack, (size_t)mine * (size_t)halo_buffers);
  if (e != cudaSuccess) { vr_sortlast_frame_destroy(f); return fail_cuda(e, "vr_sortlast_frame_create: cudaMalloc"); }
  *out = f;
  return VR_OK;
}

int vr_sortlast_frame_destroy(vr_sortlast_frame* f) {
  if (!f) return VR_OK;
  if (f->ctx) cudaSetDevice(f->ctx->device);
  cudaFree(f->cseg); cudaFree(f->tseg); cudaFree(f->contrib); cudaFree(f->rgba); cudaFree(f->hit); cudaFree(f->red); cudaFree(f->pack);
  delete f;
  return VR_OK;
}

// Halo boxes of a freshly decoded slab: pack what the neighbours need, store it into their staging areas, wait for mine,
// unpack (which also refreshes the occupancy of the boxes).  part: 0 = both halves, 1 = send only, 2 = receive only -- a
// host that has other work for the GPU puts it between the halves (sortlast.InterleavedSteps).
int vr_sortlast_exchange_halos(vr_sortlast_frame* f, vr_volume* slab, int32_t buffer, int32_t part) {
  if (!f || !slab || buffer < 0 || buffer >= f->haloBuffers || part < 0 || part > 2) return fail(VR_ERR_ARG, "vr_sortlast_exchange_halos: bad argument");
  const vr_slab_desc& me = f->sl->slabs[(size_t)f->rank];
  int64_t so[3], se[3];
  stored_box(me, f->halo, so, se);
  if (so[0] != me.origin[0] || so[1] != me.origin[1] || so[2] != me.origin[2] || se[0] != slab->SX || se[1] != slab->SY || se[2] != slab->SZ)
    return fail(VR_ERR_ARG, "vr_sortlast_exchange_halos: the slab is not the owned box plus the halo given at creation");
  const int64_t base = f->offHalo + buffer * f->haloBytes;
  std::vector<int64_t> cursor((size_t)f->n, 0);
  std::vector<int32_t> dstRank, srcs;
  std::vector<int64_t> dstOff, bytes;
  std::vector<const void*> srcPtr;
  struct Recv { int64_t o[3], e[3], off; };
  std::vector<Recv> recvs;
  int64_t packed = 0;
  for (const HaloBox& b : f->boxes) {
    const int64_t o = cursor[(size_t)b.dst];
    cursor[(size_t)b.dst] += b.bytes;
    if (b.src == f->rank) {
      uint8_t* buf = f->pack + buffer * f->packBytes + packed;
      packed += b.bytes;
      if (part != 2) {
        const int64_t lo[3] = {b.lo[0] - so[0], b.lo[1] - so[1], b.lo[2] - so[2]};
        RC(vr_volume_pack_box(f->ctx, slab, lo, b.ext, buf));
        dstRank.push_back(b.dst); dstOff.push_back(base + o); srcPtr.push_back(buf); bytes.push_back(b.bytes);
      }
    } else if (b.dst == f->rank) {
      Recv r;
      for (int a = 0; a < 3; ++a) { r.o[a] = b.lo[a] - so[a]; r.e[a] = b.ext[a]; }
      r.off = base + o;
      recvs.push_back(r);
      if (std::find(srcs.begin(), srcs.end(), b.src) == srcs.end()) srcs.push_back(b.src);
    }
  }
  if (part != 2 && !dstRank.empty())
    RC(vr_peer_put(f->peer, SLOT_HALO, (int32_t)dstRank.size(), dstRank.data(), dstOff.data(), srcPtr.data(), bytes.data()));
  if (part != 1 && !recvs.empty()) {
    std::sort(srcs.begin(), srcs.end());
    RC(vr_peer_wait(f->peer, SLOT_HALO, (int32_t)srcs.size(), srcs.data()));
    for (const Recv& r : recvs) RC(vr_volume_unpack_box(f->ctx, slab, r.o, r.e, f->base + r.off));
  }
  return VR_OK;
}

// One frame of `params` over the sharded texture; *out_rgba_dev = W*H float4 in this rank's window (row 0 = bottom row),
// complete on every rank when the context's stream reaches this point, valid until the next frame's last exchange step.
int vr_sortlast_render(vr_sortlast_frame* f, const vr_volume* slab, const vr_render_params* params, const float** out_rgba_dev,
                       uint64_t* samples_dev) {
  if (!f || !slab || !params) return fail(VR_ERR_ARG, "vr_sortlast_render: null argument");
  if (params->width != f->W || params->height != f->H) return fail(VR_ERR_ARG, "vr_sortlast_render: frame size differs from vr_sortlast_frame_create");
  std::vector<int32_t> everyone;
  all_ranks(f, &everyone);
  const int n = f->n;
  if (params->mode == VR_MODE_COMPOSITE) {
    // pass 1; T_seg stripes to their owners, which hand every rank the transmittance in front of its segment
    // (vr_peer_tin_composite: two stripe exchanges instead of an all-gather of whole images); pass 2; terms to the stripe
    // owners; sum + colour map + stripes to every window
    RC(vr_partial_composite(f->ctx, f->sl, slab, params, f->cseg, f->tseg, samples_dev));
    std::vector<int64_t> off((size_t)n), bytes((size_t)n);
    std::vector<const void*> ptr((size_t)n);
    if (n > 4) {  // two stripe exchanges; up to 4 ranks the all-gather is as fast or faster and saves a round trip (measured)
      const int64_t s4 = 4 * f->P / n;
      for (int j = 0; j < n; ++j) { off[(size_t)j] = f->offGather + f->rank * s4; ptr[(size_t)j] = (const uint8_t*)f->tseg + j * s4; bytes[(size_t)j] = s4; }
      RC(vr_peer_put(f->peer, SLOT_GATHER, n, everyone.data(), off.data(), ptr.data(), bytes.data()));
      RC(vr_peer_wait(f->peer, SLOT_GATHER, n, everyone.data()));
      const int64_t offTin = f->offRedIn + 8 * f->P;  // the half of the stripes-in area the composite terms do not use
      RC(vr_peer_tin_composite(f->ctx, f->peer, SLOT_TIN, f->sl, params, f->offGather, offTin));
      RC(vr_peer_wait(f->peer, SLOT_TIN, n, everyone.data()));
      RC(vr_resolve_composite_tin(f->ctx, f->sl, slab, params, reinterpret_cast<const float*>(f->base + offTin), f->tseg, f->cseg, f->contrib,
                                  samples_dev));
    } else {
      const int64_t nb = 4 * f->P;
      for (int r = 0; r < n; ++r) { off[(size_t)r] = f->offGather + f->rank * nb; ptr[(size_t)r] = f->tseg; bytes[(size_t)r] = nb; }
      RC(vr_peer_put(f->peer, SLOT_GATHER, n, everyone.data(), off.data(), ptr.data(), bytes.data()));
      RC(vr_peer_wait(f->peer, SLOT_GATHER, n, everyone.data()));
      RC(vr_resolve_composite(f->ctx, f->sl, slab, params, reinterpret_cast<const float*>(f->base + f->offGather), f->cseg, f->contrib, samples_dev));
    }
    const int64_t sb = 8 * f->P / n;
    for (int j = 0; j < n; ++j) { off[(size_t)j] = f->offRedIn + f->rank * sb; ptr[(size_t)j] = (const uint8_t*)f->contrib + j * sb; bytes[(size_t)j] = sb; }
    RC(vr_peer_put(f->peer, SLOT_RED_IN, n, everyone.data(), off.data(), ptr.data(), bytes.data()));
    RC(vr_peer_wait(f->peer, SLOT_RED_IN, n, everyone.data()));
    RC(vr_peer_finish_composite(f->ctx, f->peer, SLOT_RED_OUT, params, f->offRedIn, f->offTotal));
    RC(vr_peer_wait(f->peer, SLOT_RED_OUT, n, everyone.data()));
  } else if (params->mode == VR_MODE_ISOSURFACE || params->mode == VR_MODE_ISOSURFACE_COMPRESSED) {
    // first hit per rank, min of the sample indices over ranks, mask, sum of the (single non-zero) colours
    RC(vr_partial_iso(f->ctx, f->sl, slab, params, f->rgba, f->hit, samples_dev));
    RC(all_reduce(f, 1, f->hit, 4 * f->P));
    RC(vr_resolve_iso(f->ctx, f->sl, params, f->hit, reinterpret_cast<const int32_t*>(f->base + f->offTotal), f->rgba));
    RC(all_reduce(f, 0, f->rgba, 16 * f->P));
  } else {
    return fail(VR_ERR_ARG, "vr_sortlast_render: unknown mode");
  }
  if (out_rgba_dev) *out_rgba_dev = reinterpret_cast<const float*>(f->base + f->offTotal);
  return VR_OK;
}

}  // extern "C"
