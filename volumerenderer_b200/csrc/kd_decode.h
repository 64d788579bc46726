// kd_decode.h -- launch interface of the CUDA stream decoder (kd_decode.cu).
#pragma once
#include <cuda_runtime.h>
#include "kd_common.h"
#include "kd_decode_core.h"
#include "kd_parse6.h"

namespace vr {
// Four CUtensorMap over the destination slab (uint8, 3-D), box = (tx << L) x rows x 1 for L = 0..3: constant groups are
// stored by the TMA unit (cp.async.bulk.tensor) from a small row image in shared memory, so the fills cost the parser
// neither issue slots nor load/store-unit entries.
struct VrFillMaps { alignas(64) unsigned char m[4][128]; };
// false when the driver entry point is missing or the slab cannot be described (the plan then uses the classic kernel)
bool kd_fill_maps_create(uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, int tx, int ty, int rows, VrFillMaps* out);
size_t kd_decode_smem_bytes(const vr_kd_geom& g);

// ---- persistent kernel (the product path for bricks whose leaf boxes are 4x4x4, e.g. 256x256x128) --------------------------
// One CTA per SM: 22 parse warps (one lane per leaf subtree, kd_parse6.h; a warp's work item is a "unit" = 32 leaf boxes = an
// eighth of a tile) + one warp that feeds the TMA unit with constant groups.  Work comes from two device-side queues (parse
// units / constant groups, both in destination-address order); a unit's stream segment and its 32 index entries are staged
// in shared memory by the warp's own 16-byte asynchronous copies (cp.async groups) -- no barrier between warps.
#define VR6_PARSE_WARPS 22
#define VR6_QUEUE_WORDS 64  // the two queue counters, 128 bytes apart
#define VR6_THREADS (VR6_PARSE_WARPS * 32 + 32)
#define VR6_IMG_SLOTS 3  // at most (rows = 8)
struct Decode6Plan {
  vr6::P6Geom pg;
  VrFillMaps maps;
  alignas(64) unsigned char unitMap[128];  // tensor map with a unit's box (32 x 8 x 8 for 256x256x128 bricks)
  int tmaRows;
  size_t smemBytes;
};
// false: this geometry / destination cannot take the persistent kernel
bool kd_decode6_plan(const vr_kd_geom& g, uint8_t* dst, int64_t SX, int64_t SY, int64_t SZ, Decode6Plan* out, int tmaRows = 8);
uint32_t kd_decode6_stage_chunks(void);  // staging capacity per unit in 16-byte chunks (host work list: kd_build_worklist6)
// Split mode (side != null): the constant groups run as a kernel of their own on side->stream, forked from and joined to
// `stream` by the two events (all three owned by the caller, one set per context).
struct Decode6Side { cudaStream_t stream; cudaEvent_t fork, join; };
// queueDev: VR6_QUEUE_WORDS words, zeroed by the launch.  lutDev: vr6::chain_entry table (16 384 entries).
cudaError_t kd_decode6_launch(const vr_kd_geom& g, const Decode6Plan& plan, const vr_kd_job* jobsDev, const vr_parse_item* parseDev, int nParse,
                              const vr_fill_job* fillsDev, int nFill, uint32_t* queueDev, const uint32_t* lutDev, uint8_t* dst,
                              int64_t rowPitch, int64_t slicePitch, uint8_t* occ, int only, cudaStream_t stream, int fillMode = 0,
                              int tmaOut = 0, const Decode6Side* side = nullptr);
// only (timing probes, vr_debug_set("decode_only")): 0 = everything, 1 = parse tiles, 2 = constant groups.

// ---- classic kernel: one CTA per work-list entry (kd_worklist.h) -----------------------------------------------------------
// Any power-of-two brick; also the challenger the persistent kernel is measured against (round 1's default).
// Decodes the prepared streams jobsDev[] into dst (slab, x fastest; rowPitch = slab X, slicePitch = slab X*Y).  occ
// (optional): raw occupancy grid of the slab (occupancy.h, 8 slots per cell); tileCell[tile] = cell of the tile relative
// to the brick's first cell; jobs with occBase >= 0 store their tile maxima into it.  fillsDev: the constant groups that
// parse entries refer to (vals = first | count << 32).
cudaError_t kd_decode_launch(const vr_kd_geom& g, const vr_kd_job* jobsDev, const vr_kd_work* workDev, int nWork, uint8_t* dst,
                             int64_t rowPitch, int64_t slicePitch, bool vec16, cudaStream_t stream, uint8_t* occ = nullptr,
                             const int32_t* tileCell = nullptr, const vr_fill_job* fillsDev = nullptr);
}  // namespace vr
