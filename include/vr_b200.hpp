// vr_b200.hpp -- C++ drop-in classes on top of the C ABI (vr_b200.h), header only.
//
// Re-creates the three C++ surfaces main.cpp uses in the reference, with the reference's
// names and argument meaning, decoupled from GL / GLFW / Visual Studio:
//   class VolumeKdtree      /root/reference/volume_renderer/VolumeKdtree_recover.h:49-176
//   class VolumeReader<T>   /root/reference/volume_renderer/VolumeReader.h:36-291
//   renderer mode + uniforms + one Draw() per frame   main.cpp:71-74,319-341,396-404  ->  vr::Renderer
// Differences, all deliberate: build() leaves the caller's vector intact; errors are
// exceptions of type vr::Error instead of exit(-1) / GL error prints; levelCut runs on the GPU; the
// "texture" is a device slab (vr_volume) instead of a GL texture id.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "vr_b200.h"

namespace vr {

typedef unsigned char byte;  // VolumeKdtree_recover.h:22

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};
inline void check(int rc) {
  if (rc != VR_OK) throw Error(rc, std::string("vr_b200 error ") + std::to_string(rc) + ": " + vr_last_error());
}

// One device + CUDA stream, shared by the objects below (replaces the GL context of main.cpp:84-135).
class Context {
 public:
  explicit Context(int device = 0) { check(vr_context_create(device, &h_)); }
  ~Context() { vr_context_destroy(h_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  vr_context* handle() const { return h_; }
  void synchronize() { check(vr_context_synchronize(h_)); }
  // Process-wide context on device 0, created on first use: what reference-shaped calls that carry no context
  // (levelCut(cutDepth, out), VolumeReader's four-argument constructor) run on -- the role of the one GL context of main.cpp.
  static Context& shared() { static Context c(0); return c; }
 private:
  vr_context* h_ = nullptr;
};

// ---------------------------------------------------------------------------------------------------
class VolumeKdtree {
 public:
  // public fields of the reference class (VolumeKdtree_recover.h:57-79), refreshed by build() / open()
  std::vector<byte> tree;         // packed 2-bit symbols = TwoBitArray::bits
  std::vector<byte> distanceMap;  // per depth
  int maxTreeDepth = 0, origTreeDepth = 0;
  int64_t numActiveNodes = 0, X = 0, Y = 0, Z = 0;
  int tolerance = 6, maxEpochs = 5;  // defaults VolumeKdtree_recover.h:92-93
  vr_build_stats stats{};            // what build() prints in the reference (:76,120,129)

  VolumeKdtree() {}
  // :103.  ctx (optional) binds the device context levelCut / build(true-on-device) use; default = Context::shared().
  VolumeKdtree(std::vector<byte>& inData, int64_t x, int64_t y, int64_t z, Context* ctx = nullptr) : X(x), Y(y), Z(z), data_(&inData), ctx_(ctx) {}
  explicit VolumeKdtree(Context* ctx) : ctx_(ctx) {}
  void setContext(Context* ctx) { ctx_ = ctx; }
  ~VolumeKdtree() { vr_stream_destroy(s_); }
  VolumeKdtree(const VolumeKdtree&) = delete;
  VolumeKdtree& operator=(const VolumeKdtree&) = delete;

  void setErrorTolerance(int errorTolerance) { tolerance = errorTolerance; }  // :167
  void setMaxEpochs(int epochs) { maxEpochs = epochs; }                       // :175

  void build(bool /*useThreads*/ = true) {  // :126; CPU, reference algorithm, byte-identical stream
    if (!data_ || (int64_t)data_->size() != X * Y * Z) throw Error(VR_ERR_STATE, "build: no input volume of X*Y*Z bytes");
    reset();
    check(vr_kd_compress(data_->data(), X, Y, Z, tolerance, maxEpochs, &s_, &stats));
    refresh();
  }
  // The same build on the GPU (csrc/kd_build.cu): byte-identical stream, ~20 ms instead of ~0.7-2 s per brick.
  void build(Context& ctx) {
    if (!data_ || (int64_t)data_->size() != X * Y * Z) throw Error(VR_ERR_STATE, "build: no input volume of X*Y*Z bytes");
    reset();
    check(vr_kd_compress_device(ctx.handle(), data_->data(), X, Y, Z, tolerance, maxEpochs, &s_, &stats));
    refresh();
  }
  void save(const std::string& filename) { need(); check(vr_stream_save(s_, filename.c_str())); }  // :154
  void open(const std::string& filename) {                                                         // :160
    reset();
    check(vr_stream_open(filename.c_str(), &s_));
    refresh();
  }
  // Stage 1 on the GPU (:134, call site main.cpp:281), same two arguments as the reference; runs on the context bound at
  // construction (or the shared one).  cutDepth must be maxTreeDepth (any other depth crashes the reference).
  void levelCut(int cutDepth, std::vector<byte>& outData) { levelCut(cutDepth, outData, ctx_ ? *ctx_ : Context::shared()); }
  void levelCut(int cutDepth, std::vector<byte>& outData, Context& ctx) {
    need();
    outData.resize((size_t)(X * Y * Z));
    check(vr_kd_levelcut(ctx.handle(), s_, cutDepth, outData.data()));
    output_ = &outData;
  }
  // Error queries (:140-148): decoded output vs the input volume (usable here because build keeps the input).
  int measureMaxError() const {
    int m = 0;
    each_error([&](int e, size_t) { m = e > m ? e : m; });
    return m;
  }
  double measureMeanError() const {
    double s = 0;
    each_error([&](int e, size_t) { s += e; });
    return s / (double)(X * Y * Z);
  }
  void queryError(std::vector<byte>& out) const {
    out.resize((size_t)(X * Y * Z));
    each_error([&](int e, size_t i) { out[i] = (byte)e; });
  }
  vr_stream* stream() const { return s_; }

 private:
  template <class F> void each_error(F f) const {
    if (!data_ || !output_ || data_->size() != output_->size()) throw Error(VR_ERR_STATE, "error query needs the input volume and a levelCut output");
    for (size_t i = 0; i < data_->size(); ++i) f(std::abs((int)(*data_)[i] - (int)(*output_)[i]), i);
  }
  void need() const { if (!s_) throw Error(VR_ERR_STATE, "no tree: call build() or open() first"); }
  void reset() { vr_stream_destroy(s_); s_ = nullptr; }
  void refresh() {
    vr_stream_header h;
    check(vr_stream_get_header(s_, &h));
    maxTreeDepth = h.maxTreeDepth; origTreeDepth = h.origTreeDepth; numActiveNodes = h.numActiveNodes;
    X = h.X; Y = h.Y; Z = h.Z;
    distanceMap.resize((size_t)maxTreeDepth + 1);
    check(vr_stream_get_distance_map(s_, distanceMap.data(), (int32_t)distanceMap.size()));
    const uint8_t* b; int64_t n;
    check(vr_stream_get_bits(s_, &b, &n));
    tree.assign(b, b + n);
  }
  std::vector<byte>* data_ = nullptr;
  const std::vector<byte>* output_ = nullptr;
  vr_stream* s_ = nullptr;
  Context* ctx_ = nullptr;
};

// ---------------------------------------------------------------------------------------------------
typedef int64_t dim3D[3];  // (X, Y, Z)  VolumeReader.h:19

// Loads a volume stored as one raw binary file per brick into one x-fastest array and, optionally, into a
// device slab (the reference's GL_TEXTURE_3D).  Same constructor arguments and method names as the reference
// (VolumeReader.h:68, :91, :114, :151); 64-bit indexing throughout (the reference's `unsigned int` offsets
// overflow above 4 Gi voxels, SURVEY C-12).  A wrong file size throws (as :258-260), a missing file returns false.
template <typename T>
class VolumeReader {
 public:
  dim3D brickDims = {0, 0, 0};
  std::function<std::string(int, int)> findSourceFile;  // (brick, timestep) -> path
  std::map<int, dim3D>* brickMap = nullptr;  // brick -> (i, j, k) on the brick grid: the reference's type (VolumeReader.h:45)
  std::vector<T> data;
  dim3D dataDims = {0, 0, 0};
  vr_volume* volume = nullptr;  // replaces `GLuint textureId`

  VolumeReader() {
    findSourceFile = [](int, int) -> std::string { throw std::runtime_error("findSourceFile function not defined"); };
  }
  // VolumeReader.h:68-70, argument for argument; ctx (optional) = the device context of toGPU uploads, default Context::shared()
  VolumeReader(const dim3D brick, const dim3D /*grid*/, std::function<std::string(int, int)> findFileFunct, std::map<int, dim3D>* bMap,
               Context* ctx = nullptr)
      : findSourceFile(findFileFunct), brickMap(bMap), ctx_(ctx) {
    for (int a = 0; a < 3; ++a) brickDims[a] = brick[a];
  }
  // the same with a map of vectors (hosts that cannot hold arrays in a std::map)
  VolumeReader(const dim3D brick, const dim3D /*grid*/, std::function<std::string(int, int)> findFileFunct,
               std::map<int, std::vector<int64_t>>* bMap, Context* ctx = nullptr)
      : findSourceFile(findFileFunct), brickMapVec_(bMap), ctx_(ctx) {
    for (int a = 0; a < 3; ++a) brickDims[a] = brick[a];
  }
  ~VolumeReader() { vr_volume_destroy(volume); }
  void setContext(Context* ctx) { ctx_ = ctx; }

  bool LoadBrickToTexture(int brick, int timestep, bool dealloc, bool toGPU = true) {
    if (!read_brick(findSourceFile(brick, timestep))) return false;
    temp_.swap(data);
    for (int a = 0; a < 3; ++a) dataDims[a] = brickDims[a];
    if (toGPU) transferToGPU(dealloc);
    return true;
  }

  bool LoadBricksToTexture(int64_t numBricks, int64_t I, int64_t J, int64_t K, int timestep, bool dealloc, bool toGPU = true) {
    const int64_t X = brickDims[0], Y = brickDims[1], Z = brickDims[2];
    const int64_t GX = I * X, GY = J * Y;
    std::vector<int64_t> ijk((size_t)numBricks * 3);
    for (int b = 0; b < (int)numBricks; ++b) {
      if (brickMap && brickMap->count(b)) { for (int a = 0; a < 3; ++a) ijk[(size_t)b * 3 + a] = (*brickMap)[b][a]; }
      else if (brickMapVec_ && brickMapVec_->count(b)) { for (int a = 0; a < 3; ++a) ijk[(size_t)b * 3 + a] = (*brickMapVec_)[b][(size_t)a]; }
      else throw std::runtime_error("brick map has no entry for brick " + std::to_string(b));
    }
    dataDims[0] = GX; dataDims[1] = GY; dataDims[2] = K * Z;
    if (toGPU && !ctx_) ctx_ = &Context::shared();
    if (toGPU) {
      // pipelined path (vr_volume_load_bricks): file reads overlap the H2D copies, bricks go straight to their place in
      // the slab; the assembled host array (`data`) is only filled when the caller keeps it (dealloc == false)
      static_assert(sizeof(T) == 1, "the device slab is 8-bit (GL_RED / GL_UNSIGNED_BYTE in the reference)");
      vr_volume_destroy(volume); volume = nullptr;
      check(vr_volume_create(ctx_->handle(), dataDims[0], dataDims[1], dataDims[2], &volume));
      if (dealloc) std::vector<T>().swap(data); else data.assign((size_t)(GX * GY * K * Z), T(0));
      const int64_t bd[3] = {X, Y, Z};
      const int rc = vr_volume_load_bricks(ctx_->handle(), volume, numBricks, I, J, K, timestep, bd, &VolumeReader::find_trampoline, this,
                                           ijk.data(), dealloc ? nullptr : reinterpret_cast<uint8_t*>(data.data()), nullptr);
      if (rc == VR_ERR_IO) return false;                                  // missing file: the reference returns false
      if (rc == VR_ERR_FORMAT) throw std::runtime_error(vr_last_error());  // size mismatch: the reference throws (:258-260)
      check(rc);
      return true;
    }
    data.assign((size_t)(X * Y * Z * numBricks), T(0));
    for (int b = 0; b < (int)numBricks; ++b) {
      if (!read_brick(findSourceFile(b, timestep))) return false;
      for (int64_t z = 0; z < Z; ++z)
        for (int64_t y = 0; y < Y; ++y) {
          const T* src = temp_.data() + (z * Y + y) * X;
          T* dst = data.data() + ((ijk[(size_t)b * 3 + 2] * Z + z) * GY + (ijk[(size_t)b * 3 + 1] * Y + y)) * GX + ijk[(size_t)b * 3] * X;
          for (int64_t x = 0; x < X; ++x) dst[x] = src[x];
        }
    }
    if (toGPU) transferToGPU(dealloc);
    return true;
  }

  // glTexImage3D(GL_TEXTURE_3D, GL_RED, GL_UNSIGNED_BYTE) of VolumeReader.h:114-127 -> one device slab
  void transferToGPU(bool dealloc = true) {
    static_assert(sizeof(T) == 1, "the device slab is 8-bit (GL_RED / GL_UNSIGNED_BYTE in the reference)");
    if (!ctx_) ctx_ = &Context::shared();
    vr_volume_destroy(volume); volume = nullptr;
    check(vr_volume_create(ctx_->handle(), dataDims[0], dataDims[1], dataDims[2], &volume));
    check(vr_volume_upload_brick(ctx_->handle(), volume, reinterpret_cast<const uint8_t*>(data.data()), dataDims[0], dataDims[1],
                                 dataDims[2], 0, 0, 0));
    ctx_->synchronize();
    if (dealloc) std::vector<T>().swap(data);
  }

 private:
  bool read_brick(const std::string& path) {  // LoadVolumeFromBinaryFile, VolumeReader.h:244-289
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    const long long want = (long long)sizeof(T) * brickDims[0] * brickDims[1] * brickDims[2];
    if (bytes != want) { std::fclose(f); throw std::runtime_error("size of " + path + " does not match the brick dimensions"); }
    temp_.resize((size_t)(brickDims[0] * brickDims[1] * brickDims[2]));
    const size_t got = std::fread(temp_.data(), 1, (size_t)want, f);
    std::fclose(f);
    return got == (size_t)want;
  }
  static const char* find_trampoline(int32_t brick, int32_t timestep, void* self) {
    VolumeReader* r = static_cast<VolumeReader*>(self);
    r->path_ = r->findSourceFile(brick, timestep);
    return r->path_.c_str();
  }
  std::vector<T> temp_;
  std::string path_;
  std::map<int, std::vector<int64_t>>* brickMapVec_ = nullptr;
  Context* ctx_ = nullptr;
};

// ---------------------------------------------------------------------------------------------------
// The shader-selectable renderer of main.cpp: mode = which fragment shader is loaded (:71-74), the public
// fields are the uniforms and camera globals (:33-52, :319-341), Draw() is one brick.Draw() (:404) into a
// host RGBA float buffer (row 0 = bottom row) instead of the window's back buffer.
class Renderer {
 public:
  vr_render_params params;
  explicit Renderer(Context& ctx, int mode = VR_MODE_COMPOSITE, int width = 1600, int height = 1200) : ctx_(ctx) {
    check(vr_render_defaults(&params));
    params.mode = mode; params.width = width; params.height = height;
  }
  void setIsoValue(float v) { params.isoValue = v; }             // key_callback '0' / '1', main.cpp:489-498
  void setCamera(const float eye[3], const float front[3], const float up[3]) {
    for (int a = 0; a < 3; ++a) { params.eye[a] = eye[a]; params.front[a] = front[a]; params.up[a] = up[a]; }
  }
  uint64_t Draw(const vr_volume* volume, std::vector<float>& rgba) {
    rgba.resize((size_t)params.width * params.height * 4);
    uint64_t samples = 0;
    check(vr_render_to_host(ctx_.handle(), volume, &params, rgba.data(), nullptr, &samples));
    return samples;
  }
 private:
  Context& ctx_;
};

}  // namespace vr
