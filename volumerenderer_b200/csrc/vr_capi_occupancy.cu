// vr_capi_occupancy.cu -- host side of the occupancy grids (occupancy.h): allocation, validity tracking
// across decode / upload / halo writes, lazy (re)build at render time.
#include <cuda_runtime.h>

#include <cstring>
#include <vector>

#include "../../include/vr_b200.h"
#include "occupancy.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

static const int kLg[3] = {VR_OCC_LX, VR_OCC_LY, VR_OCC_LZ};

void vr_volume::occ_free() {
  if (occRaw || occDil || occMacro) cudaSetDevice(device);
  cudaFree(occRaw); cudaFree(occDil); cudaFree(occMacro);
  occRaw = occDil = occMacro = nullptr;
  occValid = false; occDilDirty = true; ++occEpoch;
}

namespace vr {

int volume_occ_alloc(vr_volume* v) {
  if (v->occRaw) return VR_OK;
  const int64_t S[3] = {v->SX, v->SY, v->SZ};
  for (int a = 0; a < 3; ++a) {
    v->shift[a] = (int)(v->cellOrigin[a] & ((1 << kLg[a]) - 1));
    v->nc[a] = (int)((S[a] + v->shift[a] - 1) >> kLg[a]) + 1;
    v->nm[a] = ((v->nc[a] - 1) >> VR_OCC_MACRO) + 1;
  }
  const size_t n = (size_t)v->nc[0] * v->nc[1] * v->nc[2], m = (size_t)v->nm[0] * v->nm[1] * v->nm[2];
  CU(cudaSetDevice(v->device));
  CU(cudaMalloc((void**)&v->occRaw, n));
  CU(cudaMalloc((void**)&v->occDil, n));
  CU(cudaMalloc((void**)&v->occMacro, m));
  v->occValid = false; v->occDilDirty = true;
  return VR_OK;
}

int volume_occ_view(vr_context* ctx, const vr_volume* cv, int thr, OccView* out) {
  std::memset(out, 0, sizeof(*out));
  vr_volume* v = const_cast<vr_volume*>(cv);  // the grids are a cache of the voxel data
  if (!v->skipping || thr < 0) return VR_OK;
  if (v->SX > 0x7FFFFFFF || v->SY > 0x7FFFFFFF || v->SZ > 0x7FFFFFFF) return VR_OK;
  int rc = volume_occ_alloc(v);
  if (rc != VR_OK) return rc;
  if (!v->occValid) {
    const int64_t lo[3] = {0, 0, 0}, ext[3] = {v->SX, v->SY, v->SZ};
    CU(occ_reduce_box_launch(v->dev, v->SX, v->SY, v->SZ, lo, ext, v->occRaw, v->nc[0], v->nc[1], v->nc[2], v->shift, ctx->stream));
    v->occValid = true; v->occDilDirty = true;
  }
  if (v->occDilDirty) {
    CU(occ_dilate_launch(v->occRaw, v->occDil, v->occMacro, v->nc[0], v->nc[1], v->nc[2], v->nm[0], v->nm[1], v->nm[2], ctx->stream));
    v->occDilDirty = false;
  }
  out->dil = v->occDil; out->macro = v->occMacro;
  out->nx = v->nc[0]; out->ny = v->nc[1]; out->nz = v->nc[2];
  out->mx = v->nm[0]; out->my = v->nm[1]; out->mz = v->nm[2];
  out->sx = v->shift[0]; out->sy = v->shift[1]; out->sz = v->shift[2];
  out->thr = thr;
  return VR_OK;
}

// Voxels of the box [lo, lo+ext) were overwritten (upload, halo unpack).  When the box covers the stored
// part of every cell it touches, those cells are recomputed in place; otherwise the grid is rebuilt lazily.
void volume_occ_written(vr_context* ctx, vr_volume* v, const int64_t lo[3], const int64_t ext[3]) {
  if (!v->occRaw || !v->occValid) { v->occValid = false; return; }
  const int64_t S[3] = {v->SX, v->SY, v->SZ};
  bool covers = true;
  for (int a = 0; a < 3; ++a) {
    const int64_t cs = int64_t(1) << kLg[a], hi = lo[a] + ext[a];
    const bool loOk = lo[a] == 0 || ((lo[a] + v->shift[a]) % cs) == 0;
    const bool hiOk = hi == S[a] || ((hi + v->shift[a]) % cs) == 0;
    covers = covers && loOk && hiOk;
  }
  if (!covers) { v->occValid = false; return; }
  cudaSetDevice(v->device);
  if (occ_reduce_box_launch(v->dev, v->SX, v->SY, v->SZ, lo, ext, v->occRaw, v->nc[0], v->nc[1], v->nc[2], v->shift, ctx->stream) != cudaSuccess)
    v->occValid = false;
  v->occDilDirty = true;
}

}  // namespace vr

extern "C" {

int vr_volume_set_cell_origin(vr_volume* v, const int64_t origin[3]) {
  if (!v || !origin) return fail(VR_ERR_ARG, "null");
  bool same = true;
  for (int a = 0; a < 3; ++a) same = same && v->cellOrigin[a] == origin[a];
  if (same) return VR_OK;
  for (int a = 0; a < 3; ++a) { if (origin[a] < 0) return fail(VR_ERR_ARG, "negative origin"); v->cellOrigin[a] = origin[a]; }
  v->occ_free();
  return VR_OK;
}

int vr_volume_set_skipping(vr_volume* v, int enable) {
  if (!v) return fail(VR_ERR_ARG, "null");
  v->skipping = enable != 0;
  return VR_OK;
}

int vr_volume_invalidate(vr_volume* v) {
  if (!v) return fail(VR_ERR_ARG, "null");
  v->occValid = false;
  return VR_OK;
}

int vr_volume_occupancy_download(vr_context* ctx, const vr_volume* v, int32_t which, uint8_t* out_host, int64_t cap, int32_t dims[3]) {
  if (!ctx || !v || !dims) return fail(VR_ERR_ARG, "null");
  OccView ov;
  int rc = volume_occ_view(ctx, v, 0, &ov);
  if (rc != VR_OK) return rc;
  if (!ov.dil) return fail(VR_ERR_STATE, "skipping is disabled for this volume");
  const bool macro = which == 2;
  dims[0] = macro ? ov.mx : ov.nx; dims[1] = macro ? ov.my : ov.ny; dims[2] = macro ? ov.mz : ov.nz;
  const int64_t n = (int64_t)dims[0] * dims[1] * dims[2];
  if (!out_host) return VR_OK;
  if (cap < n) return fail(VR_ERR_ARG, "buffer too small");
  const uint8_t* src = which == 0 ? v->occRaw : (which == 1 ? v->occDil : v->occMacro);
  CU(cudaMemcpyAsync(out_host, src, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return VR_OK;
}

}  // extern "C"
