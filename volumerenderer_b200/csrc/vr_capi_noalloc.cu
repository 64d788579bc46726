// vr_capi_noalloc.cu -- vr_render_to_host_static: the frame of vr_render_to_host (vr_capi_render.cu) marched into scratch
// memory that is part of the library image, for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
// harnesses that forbid it).  No counterpart in the reference.
//
// vr_render_to_host keeps its device-side frame (RGBA float32, the optional hit indices and the sample counter) in buffers
// it allocates and frees per call.  Here they live in one file-scope __device__ array of VR_STATIC_SCRATCH_BYTES: every GPU
// that loads the library holds that much until the process exits.  All calls share the array and it keeps the last call's
// contents: the sample counter (the one thing the march accumulates into) is cleared at the start of every call, the frame
// and the hit indices are overwritten in full; two calls that overlap on the GPU (two contexts on one device, two threads)
// race on it.  The march runs without the occupancy grid -- building one allocates -- which gives the same frame bit for
// bit (render_march.h).
#include <cuda_runtime.h>

#include <mutex>

#include "../../include/vr_b200.h"
#include "render_march.h"
#include "vr_internal.h"

using namespace vr;

namespace {
constexpr size_t kMaxPixels = (size_t)VR_STATIC_MAX_PIXELS;
constexpr size_t kOffHit = kMaxPixels * 16;              // RGBA float32 first
constexpr size_t kOffCount = kOffHit + kMaxPixels * 4;   // then the hit indices, then the 64-bit sample counter
constexpr size_t kScratchBytes = kOffCount + 256;
static_assert(kScratchBytes == (size_t)VR_STATIC_SCRATCH_BYTES, "include/vr_b200.h states the size of the scratch array");

static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

// In host code g_static_scratch is a stand-in for the symbol, not a device address: the address the kernels get comes
// from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_addr[kMaxDevices];

int scratch_of(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxDevices) return fail(VR_ERR_UNSUPPORTED, "vr_render_to_host_static: device ordinal above 63");
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_addr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);  // the caller made `device` current
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(static scratch)");
    if (!p) return fail(VR_ERR_CUDA, "cudaGetSymbolAddress(static scratch) returned null");
    g_addr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_addr[device];
  return VR_OK;
}
}  // namespace

namespace vr {
int make_render_params(const vr_render_params* in, RenderParams* P);  // vr_capi_render.cu
}

extern "C" {

int64_t vr_static_scratch_max_pixels(void) { return (int64_t)kMaxPixels; }

int vr_render_to_host_static(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_host,
                             int32_t* hit_host, uint64_t* samples_host) {
  if (!ctx || !vol || !params || !out_rgba_host) return fail(VR_ERR_ARG, "vr_render_to_host_static: null argument");
  if (params->mode != VR_MODE_COMPOSITE && params->mode != VR_MODE_ISOSURFACE && params->mode != VR_MODE_ISOSURFACE_COMPRESSED)
    return fail(VR_ERR_ARG, "vr_render_to_host_static: unknown mode");
  if (params->width <= 0 || params->height <= 0) return fail(VR_ERR_ARG, "vr_render_to_host_static: empty frame");
  const size_t px = (size_t)params->width * (size_t)params->height;
  if (px > kMaxPixels)
    return fail(VR_ERR_UNSUPPORTED, "vr_render_to_host_static: width * height exceeds VR_STATIC_MAX_PIXELS (2 097 152); "
                                    "use vr_render_to_host, which allocates its frame");
  if (vol->is_partial())
    return fail(VR_ERR_STATE, "vr_render_to_host_static: the volume was decoded lazily for one frame; run its plan in full first");
  if (vol->device != ctx->device) return fail(VR_ERR_ARG, "vr_render_to_host_static: the volume lives on another device");
  if (vol->SX > 0x7FFFFFFF || vol->SY > 0x7FFFFFFF || vol->SZ > 0x7FFFFFFF) return fail(VR_ERR_UNSUPPORTED, "volume extent too large");
  RenderParams P;
  if (make_render_params(params, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) return fail_cuda(e, "cudaSetDevice");
  unsigned char* scratch = nullptr;
  { int rc = scratch_of(ctx->device, &scratch); if (rc != VR_OK) return rc; }
  float4* rgba = reinterpret_cast<float4*>(scratch);
  int* hit = hit_host ? reinterpret_cast<int*>(scratch + kOffHit) : nullptr;
  unsigned long long* count = samples_host ? reinterpret_cast<unsigned long long*>(scratch + kOffCount) : nullptr;
  // the array keeps the previous call's count: the march adds to it
  if (count) { e = cudaMemsetAsync(count, 0, 8, ctx->stream); if (e != cudaSuccess) return fail_cuda(e, "clear sample counter"); }
  const OccView none{};  // no occupancy grid: every sample is marched
  e = render_launch(P, none, vol->dev, (int)vol->SX, (int)vol->SY, (int)vol->SZ, rgba, hit, count, ctx->stream);
  if (e != cudaSuccess) return fail_cuda(e, "march launch");
  e = cudaMemcpyAsync(out_rgba_host, rgba, px * 16, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && hit_host) e = cudaMemcpyAsync(hit_host, hit, px * 4, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && samples_host) e = cudaMemcpyAsync(samples_host, count, 8, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) return fail_cuda(e, "vr_render_to_host_static: download");
  return VR_OK;
}

}  // extern "C"
