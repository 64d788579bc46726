// kd_fill.h -- constant-tile fills of a decode launch as TMA tensor stores (kd_fill.cu).
//
// The decode of a slab is two very different jobs: tiles that have to be parsed (issue/latency bound, almost no
// memory traffic) and constant tiles (pure HBM writes).  As CTAs of one kernel they compete for the same SM
// slots, and the parser needs all six CTAs per SM to hide its latency, so the two phases ran back to back.
// Here the fills run as a second, tiny kernel (one warp per CTA, two CTAs per SM, 8 KB of shared memory) next to
// the parse kernel: a CTA writes the row image of a group of x-neighbouring constant tiles into shared memory
// once and lets the TMA unit store it slice by slice (cp.async.bulk.tensor, one instruction per 8 KB), so the
// fills cost no issue slots and no registers worth mentioning while the parser keeps its occupancy.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "kd_decode_core.h"  // vr_fill_job

namespace vr {

struct FillMaps { alignas(64) unsigned char m[4][128]; };  // four CUtensorMap (box = (tx << L) x ty x 1)

// Encodes the four tensor maps for a volume.  Returns false when the driver entry point is missing or the
// shape cannot be expressed (caller then keeps the fills in the parse kernel's work list).
bool kd_fill_maps_create(uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, int tx, int ty, int maxL, FillMaps* out);
size_t kd_fill_smem_bytes(int tx, int ty, int maxL);
cudaError_t kd_fill_launch(const FillMaps& maps, const vr_fill_job* jobsDev, int n, int tx, int ty, int tz, int maxL, uint8_t* occ,
                           int ctas, cudaStream_t stream);

// Same jobs with ordinary 16-byte stores, one 64-thread CTA per group (no shared memory, no tensor map).
cudaError_t kd_fill_plain_launch(const vr_kd_geom& g, const vr_fill_job* jobsDev, int n, uint8_t* dst, int64_t rowPitch, int64_t slicePitch,
                                 uint8_t* occ, cudaStream_t stream);

}  // namespace vr
