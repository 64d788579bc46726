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
  if (occRaw || occPyr || occLvl) cudaSetDevice(device);
  cudaFree(occRaw); cudaFree(occPyr); cudaFree(occLvl);
  occRaw = occPyr = occLvl = nullptr; occLvlThr = -2;
  occValid = false; occDilDirty = true; ++occEpoch;
}

namespace vr {

int volume_occ_alloc(vr_volume* v) {
  if (v->occRaw) return VR_OK;
  const int64_t S[3] = {v->SX, v->SY, v->SZ};
  for (int a = 0; a < 3; ++a) {
    v->shift[a] = (int)(v->cellOrigin[a] & ((1 << kLg[a]) - 1));
    v->nc[a] = (int)((S[a] + v->shift[a] - 1) >> kLg[a]) + 1;
  }
  const size_t n = (size_t)v->nc[0] * v->nc[1] * v->nc[2];
  size_t total = 0;
  v->occLevels = 0;
  for (int l = 0; l < VR_OCC_MAX_LEVELS; ++l) {
    const size_t nl = (size_t)(((v->nc[0] - 1) >> l) + 1) * (((v->nc[1] - 1) >> l) + 1) * (((v->nc[2] - 1) >> l) + 1);
    if (total + nl > 0x7FFFFFFF) break;
    v->occOff[l] = (int)total;
    total += (nl + 15) & ~size_t(15);
    v->occLevels = l + 1;
    if (nl == 1) break;
  }
  CU(cudaSetDevice(v->device));
  CU(cudaMalloc((void**)&v->occRaw, n * VR_OCC_SLOTS));
  CU(cudaMalloc((void**)&v->occPyr, total));
  CU(cudaMalloc((void**)&v->occLvl, 3 * n));  // radii + two scratch passes (occ_levels_launch)
  v->occValid = false; v->occDilDirty = true;
  return VR_OK;
}

int volume_occ_view(vr_context* ctx, const vr_volume* cv, int thr, OccView* out, bool wantPyramid) {
  std::memset(out, 0, sizeof(*out));
  vr_volume* v = const_cast<vr_volume*>(cv);  // the grids are a cache of the voxel data
  if (!v->skipping || thr < 0) return VR_OK;
  if (v->SX > 0x7FFFFFFF || v->SY > 0x7FFFFFFF || v->SZ > 0x7FFFFFFF) return VR_OK;
  int rc = volume_occ_alloc(v);
  if (rc != VR_OK) return rc;
  if (!v->occValid) {
    if (v->is_partial()) return fail(VR_ERR_STATE, "occupancy rebuild: the volume was decoded lazily for one frame (stale tiles); run its plan in full first");
    const int64_t lo[3] = {0, 0, 0}, ext[3] = {v->SX, v->SY, v->SZ};
    CU(occ_reduce_box_launch(v->dev, v->SX, v->SY, v->SZ, lo, ext, v->occRaw, v->nc[0], v->nc[1], v->nc[2], v->shift, ctx->stream));
    v->occValid = true; v->occDilDirty = true;
  }
  out->nlev = v->occLevels;
  for (int l = 0; l < v->occLevels; ++l) out->off[l] = v->occOff[l];
  out->nx = v->nc[0]; out->ny = v->nc[1]; out->nz = v->nc[2];
  out->sx = v->shift[0]; out->sy = v->shift[1]; out->sz = v->shift[2];
  out->thr = thr;
  if (v->occDilDirty || v->occLvlThr != thr) {
    // a freshly decoded timestep (or a new threshold): dilated level 0 + void radii in two launches; the octree levels above
    // level 0 are only built when somebody asks for them (wantPyramid: vr_volume_occupancy_download)
    cudaError_t e = cudaSuccess;
    if (occ_refresh_launch(v->occRaw, v->occPyr, v->occLvl, *out, ctx->stream, &e)) {
      CU(e);
      if (v->occDilDirty) v->occPyrHigh = false;
    } else {
      CU(occ_pyramid_launch(v->occRaw, v->occPyr, *out, ctx->stream));
      CU(occ_levels_launch(v->occPyr, v->occLvl, *out, ctx->stream));
      v->occPyrHigh = true;
    }
    v->occDilDirty = false;
    v->occLvlThr = thr;
  }
  if (wantPyramid && !v->occPyrHigh) {
    CU(occ_pyramid_launch(v->occRaw, v->occPyr, *out, ctx->stream));
    v->occPyrHigh = true;
  }
  out->pyr = v->occPyr;
  out->lvl = v->occLvl;
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
  v->partial.clear();  // the caller vouches for the stored voxels
  v->partialOrphan = false;
  return VR_OK;
}

int vr_volume_occupancy_download(vr_context* ctx, const vr_volume* v, int32_t which, uint8_t* out_host, int64_t cap, int32_t dims[3]) {
  if (!ctx || !v || !dims) return fail(VR_ERR_ARG, "null");
  OccView ov;
  int rc = volume_occ_view(ctx, v, 0, &ov, which >= 2);
  if (rc != VR_OK) return rc;
  if (!ov.pyr) return fail(VR_ERR_STATE, "skipping is disabled for this volume");
  if (which < 0 || which > ov.nlev) return fail(VR_ERR_ARG, "no such level");
  const int l = which == 0 ? 0 : which - 1;
  dims[0] = ((ov.nx - 1) >> l) + 1; dims[1] = ((ov.ny - 1) >> l) + 1; dims[2] = ((ov.nz - 1) >> l) + 1;
  const int64_t n = (int64_t)dims[0] * dims[1] * dims[2];
  if (!out_host) return VR_OK;
  if (cap < n) return fail(VR_ERR_ARG, "buffer too small");
  if (which == 0) {  // raw: max over the writer slots
    std::vector<uint8_t> slots((size_t)n * VR_OCC_SLOTS);
    CU(cudaMemcpyAsync(slots.data(), v->occRaw, slots.size(), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    for (int64_t i = 0; i < n; ++i) {
      uint8_t m = 0;
      for (int k = 0; k < VR_OCC_SLOTS; ++k) m = slots[(size_t)i * VR_OCC_SLOTS + k] > m ? slots[(size_t)i * VR_OCC_SLOTS + k] : m;
      out_host[i] = m;
    }
    return VR_OK;
  }
  CU(cudaMemcpyAsync(out_host, v->occPyr + ov.off[l], (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return VR_OK;
}

}  // extern "C"
