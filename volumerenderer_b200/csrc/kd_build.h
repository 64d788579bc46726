// kd_build.h -- GPU stream producer (kd_build.cu): VolumeKdtree::build on the device, byte-identical streams.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "kd_stream.h"

namespace vr {

// Scratch of one producer (sized for the largest brick built so far; ~45 bytes per voxel); reuse it across bricks.
struct KdBuilder;
KdBuilder* kd_builder_create();
void kd_builder_destroy(KdBuilder* b);

// Builds the stream of one X*Y*Z brick.  `volDev` points at the brick's (0,0,0) inside a device array with row
// pitch `rowPitch` and slice pitch `slicePitch` (a tight brick: X and X*Y).  Synchronises `stream` several times
// per level (the gradient descent branches on the level's error).  *err receives the CUDA error on VR_ERR_CUDA.
// iterStats (optional, 2 entries): scan iterations of the running-mean fixed point (total, worst level).
// withIndex: also emit out->index (the decode index of kd_prepare, byte for byte) from the producer's own symbol offsets, so
// the finished stream is never walked on the host.
int kd_build_device(KdBuilder* b, const uint8_t* volDev, int64_t X, int64_t Y, int64_t Z, int64_t rowPitch, int64_t slicePitch,
                    int tolerance, int maxEpochs, cudaStream_t stream, KdStream* out, KdBuildStats* stats, cudaError_t* err,
                    int64_t* iterStats, bool withIndex = true);

}  // namespace vr
