// kd_decode.h -- launch interface of the CUDA stream decoder (kd_decode.cu).
#pragma once
#include <cuda_runtime.h>
#include "kd_common.h"
#include "kd_decode_core.h"

namespace vr {
// Four CUtensorMap over the destination slab (uint8, 3-D), box = (tx << L) x rows x 1 for L = 0..3: the
// constant groups attached to a parse CTA are stored by the TMA unit (cp.async.bulk.tensor) from a small row image in
// shared memory while all eight warps of the CTA parse -- the fills then cost the parser neither issue slots nor CTA
// slots (it needs all six CTAs per SM to hide its latency).
#define VR_TMA_MAX_ATTACH 2
struct VrFillMaps { alignas(64) unsigned char m[4][128]; };
// false when the driver entry point is missing or the slab cannot be described (fills then use ordinary stores)
bool kd_fill_maps_create(uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, int tx, int ty, int rows, VrFillMaps* out);
size_t kd_decode_smem_bytes(const vr_kd_geom& g);
// Pulls n device arenas (ptrs[i], lines[i] x 128 bytes) into L2 with evict-last priority: the prepared streams of a
// plan (tens of MB, the L2 holds 126 MB) are then read by the parsers from L2 instead of DRAM.
cudaError_t kd_prefetch_l2_launch(const uint8_t* const* ptrsDev, const uint32_t* linesDev, int n, cudaStream_t stream);
// Decodes the prepared streams jobsDev[] into dst (slab, x fastest; rowPitch = slab X, slicePitch = slab X*Y),
// one CTA per entry of workDev (kd_worklist.h).  Asynchronous on `stream`.  occ (optional): raw occupancy grid
// of the slab (occupancy.h, 8 slots per cell); tileCell[tile] = cell of the tile relative to the brick's first
// cell; jobs with occBase >= 0 store their tile maxima into it.  fillsDev: the constant groups that parse entries
// refer to (vals = first | count << 32).
cudaError_t kd_decode_launch(const vr_kd_geom& g, const vr_kd_job* jobsDev, const vr_kd_work* workDev, int nWork, uint8_t* dst,
                             int64_t rowPitch, int64_t slicePitch, bool vec16, cudaStream_t stream, uint8_t* occ = nullptr,
                             const int32_t* tileCell = nullptr, const vr_fill_job* fillsDev = nullptr,
                             const vr_stage_range* rangesDev = nullptr, int stageBytes = 0,
                             const VrFillMaps* tmaMaps = nullptr, int nParse = -1, int fillEvery = 0, uint32_t* queueDev = nullptr,
                             int tmaRows = 0);
// fillEvery > 0 (needs queueDev = 2 counters and a work list with its nParse parse entries first): persistent launch in which
// every fillEvery-th SM works on the constant groups and the others on the parse tiles (kd_decode.cu).
// rangesDev (optional, one entry per CTA, kd_worklist.h): the stream segment a parse CTA copies into shared memory
// before parsing (stageBytes of dynamic shared memory behind the strips); null / stageBytes 0 = parse from global memory.
}  // namespace vr
