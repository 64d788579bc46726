// vr_capi_render.cu -- renderer half of the C ABI: camera set-up (GLM lookAt / perspectiveFov as used
// at main.cpp:396-397), shader-selectable mode (main.cpp:71-74), headless RGBA output.
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>

#include "../../include/vr_b200.h"
#include "occupancy.h"
#include "render_march.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

namespace vr {
int make_render_params(const vr_render_params* in, RenderParams* P) {
  if (in->width <= 0 || in->height <= 0) return VR_ERR_ARG;
  double f[3] = {in->front[0], in->front[1], in->front[2]}, up[3] = {in->up[0], in->up[1], in->up[2]};
  double fl = std::sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
  if (!(fl > 0)) return VR_ERR_ARG;
  for (double& c : f) c /= fl;
  double s[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
  double sl = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
  if (!(sl > 0)) return VR_ERR_ARG;
  for (double& c : s) c /= sl;
  double u[3] = {s[1] * f[2] - s[2] * f[1], s[2] * f[0] - s[0] * f[2], s[0] * f[1] - s[1] * f[0]};
  const double kPi = 3.14159265358979323846;
  double th = 1.0 / std::tan((double)in->vfov_deg * kPi / 180.0 / 2.0);
  double tw = th * (double)in->height / (double)in->width;
  for (int i = 0; i < 3; ++i) {
    P->eye[i] = in->eye[i]; P->s[i] = (float)s[i]; P->u[i] = (float)u[i]; P->f[i] = (float)f[i];
    P->step[i] = in->step_size[i];
  }
  P->tw = (float)tw; P->th = (float)th; P->znear = in->znear; P->zfar = in->zfar;
  P->iso = in->isoValue;
  P->W = in->width; P->H = in->height;
  P->mode = in->mode == VR_MODE_COMPOSITE ? 0 : 1;
  return VR_OK;
}
// Largest voxel value m whose sample can never reach the mode's threshold (occupancy.h): composite skips
// exactly-zero cells; isosurface skips cells where even the maximum tap stays below isoValue
// (smp2 <= m/255 in the kernel's own arithmetic, so `smp2 - iso >= 0` of isosurface.frag:126 cannot hold).
int skip_threshold(const RenderParams& P) {
  if (P.mode == 0) return 0;
  int thr = -1;
  for (int m = 0; m < 256; ++m) {
    volatile float s = (float)m * (1.0f / 255.0f);
    volatile float d = s - P.iso;
    if (d < 0.0f) thr = m; else break;
  }
  return thr;
}
}  // namespace vr

extern "C" {

int vr_render_defaults(vr_render_params* p) {
  if (!p) return fail(VR_ERR_ARG, "null");
  std::memset(p, 0, sizeof(*p));
  p->eye[2] = -0.75f;                 // camPos   main.cpp:33
  p->front[2] = 1.0f;                 // camFront main.cpp:34
  p->up[1] = 1.0f;                    // camUp    main.cpp:35
  p->vfov_deg = 50.0f;                // fov      main.cpp:40
  p->znear = 0.1f; p->zfar = 100.0f;  // main.cpp:397
  p->step_size[0] = 1.0f / 256.0f; p->step_size[1] = 1.0f / 256.0f; p->step_size[2] = 1.0f / 128.0f;  // main.cpp:330-331
  p->isoValue = 40.0f / 255.0f;       // main.cpp:52,334
  p->width = 1600; p->height = 1200;  // main.cpp:27
  p->mode = VR_MODE_COMPOSITE;
  return VR_OK;
}

int vr_camera_basis(const vr_render_params* in, float out[20]) {
  if (!in || !out) return fail(VR_ERR_ARG, "null");
  RenderParams P;
  if (make_render_params(in, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  std::memcpy(out, &P, 20 * sizeof(float));
  return VR_OK;
}

int vr_render(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_dev,
              int32_t* hit_dev, uint64_t* samples_dev) {
  if (!ctx || !vol || !params || !out_rgba_dev) return fail(VR_ERR_ARG, "vr_render: null argument");
  if (params->mode != VR_MODE_COMPOSITE && params->mode != VR_MODE_ISOSURFACE && params->mode != VR_MODE_ISOSURFACE_COMPRESSED)
    return fail(VR_ERR_ARG, "vr_render: unknown mode");
  RenderParams P;
  if (make_render_params(params, &P) != VR_OK) return fail(VR_ERR_ARG, "degenerate camera or image size");
  if (vol->SX > 0x7FFFFFFF || vol->SY > 0x7FFFFFFF || vol->SZ > 0x7FFFFFFF) return fail(VR_ERR_UNSUPPORTED, "volume extent too large");
  CU(cudaSetDevice(ctx->device));
  OccView occ;
  { int rc = volume_occ_view(ctx, vol, skip_threshold(P), &occ); if (rc != VR_OK) return rc; }
  CU(render_launch(P, occ, vol->dev, (int)vol->SX, (int)vol->SY, (int)vol->SZ, reinterpret_cast<float4*>(out_rgba_dev), hit_dev,
                   reinterpret_cast<unsigned long long*>(samples_dev), ctx->stream));
  return VR_OK;
}

int vr_render_to_host(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_host,
                      int32_t* hit_host, uint64_t* samples_host) {
  if (!ctx || !params || !out_rgba_host) return fail(VR_ERR_ARG, "vr_render_to_host: null argument");
  CU(cudaSetDevice(ctx->device));
  const size_t px = (size_t)params->width * (size_t)params->height;
  float* d = nullptr; int32_t* dh = nullptr; uint64_t* ds = nullptr;
  CU(cudaMalloc((void**)&d, px * 16));
  int rc = VR_OK;
  cudaError_t e = cudaSuccess;
  if (hit_host) e = cudaMalloc((void**)&dh, px * 4);
  if (e == cudaSuccess && samples_host) { e = cudaMalloc((void**)&ds, 8); if (e == cudaSuccess) e = cudaMemsetAsync(ds, 0, 8, ctx->stream); }
  if (e != cudaSuccess) rc = fail_cuda(e, "cudaMalloc(render outputs)");
  if (rc == VR_OK) rc = vr_render(ctx, vol, params, d, dh, ds);
  if (rc == VR_OK) {
    e = cudaMemcpyAsync(out_rgba_host, d, px * 16, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && hit_host) e = cudaMemcpyAsync(hit_host, dh, px * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && samples_host) e = cudaMemcpyAsync(samples_host, ds, 8, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = fail_cuda(e, "render download");
  }
  cudaFree(d); cudaFree(dh); cudaFree(ds);
  return rc;
}

int vr_render_compressed(vr_context* ctx, vr_decode_plan* plan, const vr_render_params* params, float* out_rgba_dev,
                         int32_t* hit_dev, uint64_t* samples_dev) {
  if (!plan) return fail(VR_ERR_ARG, "null plan");
  int rc = vr_decode_plan_run(plan);
  if (rc != VR_OK) return rc;
  return vr_render(ctx, plan->dst, params, out_rgba_dev, hit_dev, samples_dev);
}

}  // extern "C"
