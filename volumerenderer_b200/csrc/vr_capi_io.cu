// vr_capi_io.cu -- the file-facing side of the hot path: raw brick files -> device slab, and .bin tree
// containers -> uploaded streams, both pipelined so that the disk, the host cores and the PCIe link work at the same
// time.  Replaces (paths relative to /root/reference/volume_renderer/):
//   VolumeReader::LoadBricksToTexture  VolumeReader.h:151-223  (read every brick file, assemble one host array with
//                                      `unsigned int` offsets, then one glTexImage3D of the whole array :114-138)
//   VolumeReader::LoadVolumeFromBinaryFile  VolumeReader.h:244-289  (size check :258-260)
//   fillVolumeBrickMap                 main.cpp:599-619         (brick b -> grid (b % I, (b / I) % J, b / (I*J)))
//   VolumeKdtree::open                 VolumeKdTree_recover.cpp:554-594, called per brick at main.cpp:276-284
// Here a reader thread fills a ring of pinned staging buffers while the context's stream copies each brick straight
// to its place in the slab (3-D copy, 64-bit offsets); the host never holds the assembled array unless the caller
// asks for it (VolumeReader::data).
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/vr_b200.h"
#include "vr_internal.h"

using namespace vr;
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

namespace {

// 0 = ok, VR_ERR_IO = missing / short read, VR_ERR_FORMAT = size does not match the brick (the reference throws, :258-260)
int read_exact(const std::string& path, uint8_t* dst, int64_t want) {
  std::FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return VR_ERR_IO;
  std::fseek(f, 0, SEEK_END);
  const long long bytes = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  if (bytes != want) { std::fclose(f); return VR_ERR_FORMAT; }
  const size_t got = std::fread(dst, 1, (size_t)want, f);
  std::fclose(f);
  return got == (size_t)want ? VR_OK : VR_ERR_IO;
}

struct Ring {
  static const int kSlots = 4;
  uint8_t* buf[kSlots] = {};
  cudaEvent_t done[kSlots] = {};   // the copy out of the slot has finished
  bool used[kSlots] = {};
  ~Ring() {
    for (int i = 0; i < kSlots; ++i) { if (buf[i]) cudaFreeHost(buf[i]); if (done[i]) cudaEventDestroy(done[i]); }
  }
};

}  // namespace

extern "C" {

int vr_volume_load_bricks(vr_context* ctx, vr_volume* v, int64_t numBricks, int64_t I, int64_t J, int64_t K, int32_t timestep,
                          const int64_t brick_dims[3], vr_find_source_file find, void* user, const int64_t* brick_ijk,
                          uint8_t* host_data, int64_t stats[2]) {
  if (!ctx || !v || !brick_dims || !find || numBricks <= 0 || I <= 0 || J <= 0 || K <= 0)
    return fail(VR_ERR_ARG, "vr_volume_load_bricks: bad argument");
  const int64_t X = brick_dims[0], Y = brick_dims[1], Z = brick_dims[2];
  if (X <= 0 || Y <= 0 || Z <= 0) return fail(VR_ERR_ARG, "vr_volume_load_bricks: bad brick dimensions");
  if (I * X > v->SX || J * Y > v->SY || K * Z > v->SZ) return fail(VR_ERR_ARG, "vr_volume_load_bricks: the brick grid does not fit the volume");
  const int64_t brickBytes = X * Y * Z;
  // paths and grid positions are resolved on the calling thread (the callback need not be thread-safe)
  std::vector<std::string> paths((size_t)numBricks);
  std::vector<int64_t> pos((size_t)numBricks * 3);
  for (int64_t b = 0; b < numBricks; ++b) {
    const char* p = find((int32_t)b, timestep, user);
    if (!p) return fail(VR_ERR_IO, "findSourceFile returned no path");
    paths[(size_t)b] = p;
    int64_t* q = &pos[(size_t)b * 3];
    if (brick_ijk) { q[0] = brick_ijk[3 * b]; q[1] = brick_ijk[3 * b + 1]; q[2] = brick_ijk[3 * b + 2]; }
    else { q[0] = b % I; q[1] = (b / I) % J; q[2] = b / (I * J); }  // fillVolumeBrickMap, main.cpp:599-619
    if (q[0] < 0 || q[0] >= I || q[1] < 0 || q[1] >= J || q[2] < 0 || q[2] >= K) return fail(VR_ERR_ARG, "brick map entry outside the grid");
  }
  CU(cudaSetDevice(ctx->device));
  Ring ring;
  for (int i = 0; i < Ring::kSlots; ++i) {
    CU(cudaHostAlloc((void**)&ring.buf[i], (size_t)brickBytes, cudaHostAllocDefault));
    CU(cudaEventCreateWithFlags(&ring.done[i], cudaEventDisableTiming));
  }
  // reader thread: slot s is reusable once the main thread has seen its copy finish (freeSlots), filled slots are
  // handed over in brick order (ready)
  std::mutex mu;
  std::condition_variable cv;
  int freeSlots = Ring::kSlots;       // slots the reader may fill
  int64_t readyCount = 0;             // bricks read so far
  int readerRc = VR_OK;
  int64_t failedBrick = -1;
  bool stop = false;
  std::thread reader([&]() {
    for (int64_t b = 0; b < numBricks; ++b) {
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return freeSlots > 0 || stop; });
        if (stop) return;
        --freeSlots;
      }
      const int rc = read_exact(paths[(size_t)b], ring.buf[b % Ring::kSlots], brickBytes);
      std::lock_guard<std::mutex> lk(mu);
      if (rc != VR_OK) { readerRc = rc; failedBrick = b; cv.notify_all(); return; }
      ++readyCount;
      cv.notify_all();
    }
  });
  int rc = VR_OK;
  cudaError_t ce = cudaSuccess;
  int64_t copied = 0;
  for (int64_t b = 0; b < numBricks && rc == VR_OK; ++b) {
    const int slot = (int)(b % Ring::kSlots);
    {
      std::unique_lock<std::mutex> lk(mu);
      cv.wait(lk, [&] { return readyCount > b || readerRc != VR_OK; });
      if (readyCount <= b) { rc = readerRc; break; }
    }
    const int64_t* q = &pos[(size_t)b * 3];
    cudaMemcpy3DParms p;
    std::memset(&p, 0, sizeof(p));
    p.srcPtr = make_cudaPitchedPtr((void*)ring.buf[slot], (size_t)X, (size_t)X, (size_t)Y);
    p.dstPtr = make_cudaPitchedPtr((void*)v->dev, (size_t)v->SX, (size_t)v->SX, (size_t)v->SY);
    p.dstPos = make_cudaPos((size_t)(q[0] * X), (size_t)(q[1] * Y), (size_t)(q[2] * Z));
    p.extent = make_cudaExtent((size_t)X, (size_t)Y, (size_t)Z);
    p.kind = cudaMemcpyHostToDevice;
    ce = cudaMemcpy3DAsync(&p, ctx->stream);
    if (ce == cudaSuccess) ce = cudaEventRecord(ring.done[slot], ctx->stream);
    if (ce != cudaSuccess) { rc = VR_ERR_CUDA; break; }
    ring.used[slot] = true;
    if (host_data) {  // VolumeReader::data: the assembled array, same layout as the slab (VolumeReader.h:184-205, 64-bit)
      const int64_t GX = I * X, GY = J * Y;
      for (int64_t z = 0; z < Z; ++z)
        for (int64_t y = 0; y < Y; ++y)
          std::memcpy(host_data + ((q[2] * Z + z) * GY + (q[1] * Y + y)) * GX + q[0] * X, ring.buf[slot] + (z * Y + y) * X, (size_t)X);
    }
    ++copied;
    // hand the slot that the copy issued kSlots-1 bricks ago has left back to the reader
    const int64_t old = b - (Ring::kSlots - 1);
    if (old >= 0) {
      ce = cudaEventSynchronize(ring.done[old % Ring::kSlots]);
      if (ce != cudaSuccess) { rc = VR_ERR_CUDA; break; }
      std::lock_guard<std::mutex> lk(mu);
      ++freeSlots;
      cv.notify_all();
    }
  }
  {
    std::lock_guard<std::mutex> lk(mu);
    stop = true;
    cv.notify_all();
  }
  reader.join();
  cudaError_t se = cudaStreamSynchronize(ctx->stream);  // the staging ring is freed on return
  if (rc == VR_OK && se != cudaSuccess) { ce = se; rc = VR_ERR_CUDA; }
  v->occValid = false;  // reduced lazily at the next render (occupancy.h)
  if (copied * (int64_t)(brick_dims[0] * brick_dims[1] * brick_dims[2]) >= v->SX * v->SY * v->SZ) { v->partial.clear(); v->partialOrphan = false; }  // every voxel rewritten
  if (stats) { stats[0] = copied; stats[1] = copied * brickBytes; }
  if (rc == VR_ERR_CUDA) return fail_cuda(ce, "vr_volume_load_bricks: copy");
  if (rc == VR_ERR_IO) return fail(rc, ("cannot read brick file " + (failedBrick >= 0 ? paths[(size_t)failedBrick] : std::string("?"))).c_str());
  if (rc == VR_ERR_FORMAT)
    return fail(rc, ("size of " + (failedBrick >= 0 ? paths[(size_t)failedBrick] : std::string("?")) + " does not match the brick dimensions").c_str());
  return rc;
}

int vr_streams_load(vr_context* ctx, const char* const* paths, int32_t n, int32_t threads, vr_stream** out) {
  if (!paths || n <= 0 || !out) return fail(VR_ERR_ARG, "vr_streams_load: bad argument");
  for (int i = 0; i < n; ++i) { out[i] = nullptr; if (!paths[i]) return fail(VR_ERR_ARG, "vr_streams_load: null path"); }
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads <= 0) threads = 4;
  if (threads > n) threads = n;
  if (ctx) CU(cudaSetDevice(ctx->device));
  // Phase 1, workers: open (VolumeKdtree::open; an indexed container brings its index, a plain one is walked) and lay out the
  // device image.  Phase 2, workers again: write every image straight into its slice of the context's pinned staging buffer.
  // Then ONE device allocation and ONE H2D copy for the whole set (a timestep): the streams share the allocation.
  std::vector<int> err((size_t)n, VR_OK);
  auto run_pool = [&](auto&& fn) {
    std::atomic<int> next(0);
    std::vector<std::thread> pool;
    for (int w = 0; w < threads; ++w)
      pool.emplace_back([&]() {
        for (;;) {
          const int i = next.fetch_add(1);
          if (i >= n) return;
          fn(i);
        }
      });
    for (auto& t : pool) t.join();
  };
  run_pool([&](int i) {
    vr_stream* s = new vr_stream;
    int rc = kd_open_any(paths[i], &s->kd);
    if (rc == VR_OK) rc = ctx ? stream_image_layout(s) : stream_build_image(s);
    if (rc != VR_OK) { delete s; err[(size_t)i] = rc; return; }
    out[i] = s;
  });
  int rc = VR_OK, bad = -1;
  for (int i = 0; i < n && rc == VR_OK; ++i)
    if (err[(size_t)i] != VR_OK) { rc = err[(size_t)i]; bad = i; }
  std::string cudaMsg;
  if (rc == VR_OK && ctx) {
    std::vector<size_t> off((size_t)n);
    size_t total = 0;
    for (int i = 0; i < n; ++i) { off[(size_t)i] = total; total += (out[i]->imageBytes + 255) & ~size_t(255); }
    cudaError_t e = cudaSuccess;
    if (!ctx->stagingDone) e = cudaEventCreateWithFlags(&ctx->stagingDone, cudaEventDisableTiming);
    else e = cudaEventSynchronize(ctx->stagingDone);  // the previous load's copy has left the staging buffer
    if (e == cudaSuccess && total > ctx->stagingBytes) {
      if (ctx->stagingHost) cudaFreeHost(ctx->stagingHost);
      ctx->stagingHost = nullptr; ctx->stagingBytes = 0;
      const size_t want = total + total / 4;  // room to grow: pinning is the expensive part of a cold load
      e = cudaHostAlloc((void**)&ctx->stagingHost, want, cudaHostAllocDefault);
      if (e == cudaSuccess) ctx->stagingBytes = want;
    }
    auto mem = std::make_shared<vr::ArenaMem>();
    mem->device = ctx->device; mem->bytes = total;
    if (e == cudaSuccess) e = cudaMalloc((void**)&mem->dev, total);
    if (e == cudaSuccess) {
      uint8_t* stage = ctx->stagingHost;
      run_pool([&](int i) { stream_image_write(out[i], stage + off[(size_t)i]); });
      e = cudaMemcpyAsync(mem->dev, stage, total, cudaMemcpyHostToDevice, ctx->stream);
      if (e == cudaSuccess) e = cudaEventRecord(ctx->stagingDone, ctx->stream);
    }
    if (e != cudaSuccess) { rc = VR_ERR_CUDA; bad = -2; cudaMsg = std::string("vr_streams_load: ") + cudaGetErrorString(e); }
    else
      for (int i = 0; i < n; ++i) {
        vr_stream* s = out[i];
        s->devImage = mem->dev + off[(size_t)i]; s->device = ctx->device; s->devOwned = false; s->arena = mem;
      }
  }
  if (rc != VR_OK) {
    std::string msg = bad == -2 ? cudaMsg : "vr_streams_load: cannot open / parse " + std::string(bad >= 0 ? paths[bad] : "?");
    if (ctx) cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < n; ++i) { vr_stream_destroy(out[i]); out[i] = nullptr; }
    return fail(rc, msg.c_str());
  }
  return VR_OK;
}

}  // extern "C"
