// dropin_main.cpp -- the reference's main.cpp pipeline (load bricks -> build -> save -> open -> levelCut ->
// upload -> draw, main.cpp:141-290,404) written against include/vr_b200.hpp.  Test program:
//   dropin_main cpu  <dir>   : everything that needs no GPU (brick files, build, save, open, stream equality)
//   dropin_main gpu  <dir>   : + levelCut on the GPU, error queries, one composite and one isosurface frame
// Prints "key=value" lines that tests/test_cpp_dropin.py checks against the Python mirror / the oracle.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "vr_b200.hpp"

static uint32_t fnv(const void* p, size_t n) {
  const unsigned char* b = (const unsigned char*)p;
  uint32_t h = 2166136261u;
  for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 16777619u; }
  return h;
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  const bool gpu = std::strcmp(argv[1], "gpu") == 0;
  const std::string dir = argv[2];
  try {
    vr::dim3D BRICK_DIM = {32, 32, 16}, VOLUME_GRID = {2, 2, 2};  // main.cpp:78-79 in miniature
    std::map<int, vr::dim3D> brickMap;                            // main.cpp:81, filled as fillVolumeBrickMap does (main.cpp:599-619)
    for (int b = 0; b < 8; ++b) { brickMap[b][0] = b % 2; brickMap[b][1] = (b / 2) % 2; brickMap[b][2] = b / 4; }
    auto findBrickBinaryFile = [&](int brick, int timestep) { return dir + "/brick_" + std::to_string(brick) + "_t" + std::to_string(timestep) + ".raw"; };
    vr::Context* ctx = gpu ? new vr::Context(0) : nullptr;
    vr::VolumeReader<vr::byte> volume(BRICK_DIM, VOLUME_GRID, findBrickBinaryFile, &brickMap, ctx);
    if (!volume.LoadBricksToTexture(8, 2, 2, 2, 7, /*dealloc*/ false, /*toGPU*/ false)) { std::puts("load=failed"); return 1; }
    std::printf("dims=%lld,%lld,%lld\n", (long long)volume.dataDims[0], (long long)volume.dataDims[1], (long long)volume.dataDims[2]);
    std::printf("assembled=%u\n", fnv(volume.data.data(), volume.data.size()));
    bool missing = !volume.LoadBrickToTexture(99, 7, false, false);
    std::printf("missing_brick_returns_false=%d\n", missing ? 1 : 0);
    volume.LoadBricksToTexture(8, 2, 2, 2, 7, false, false);

    vr::VolumeKdtree myTree(volume.data, volume.dataDims[0], volume.dataDims[1], volume.dataDims[2]);
    myTree.setMaxEpochs(2);       // main.cpp:253
    myTree.setErrorTolerance(1);  // main.cpp:254
    myTree.build(true);           // main.cpp:258
    std::printf("input_intact=%d\n", (int)(volume.data.size() == 64u * 64u * 32u));
    std::printf("nodes=%lld depth=%d,%d maxerr=%d\n", (long long)myTree.numActiveNodes, myTree.origTreeDepth, myTree.maxTreeDepth, myTree.stats.maxError);
    std::printf("stream=%u dmap=%u\n", fnv(myTree.tree.data(), myTree.tree.size()), fnv(myTree.distanceMap.data(), myTree.distanceMap.size()));
    myTree.save(dir + "/tree.bin");  // main.cpp:267
    vr::VolumeKdtree opened;
    opened.open(dir + "/tree.bin");  // main.cpp:269
    std::printf("reopened_equal=%d\n", (int)(opened.tree == myTree.tree && opened.distanceMap == myTree.distanceMap && opened.numActiveNodes == myTree.numActiveNodes));
    try { vr::VolumeKdtree t; t.open(dir + "/nope.bin"); std::puts("open_missing=nothrow"); }
    catch (const vr::Error& e) { std::printf("open_missing=error%d\n", e.code); }
    if (!gpu) return 0;

    {  // LoadBricksToTexture with toGPU: pipelined file reads + H2D copies straight into the slab (vr_volume_load_bricks)
      vr::VolumeReader<vr::byte> streamed(BRICK_DIM, VOLUME_GRID, findBrickBinaryFile, &brickMap, ctx);
      bool ok = streamed.LoadBricksToTexture(8, 2, 2, 2, 7, /*dealloc*/ false, /*toGPU*/ true);
      std::vector<vr::byte> dev(64u * 64u * 32u);
      if (ok) vr::check(vr_volume_download(ctx->handle(), streamed.volume, dev.data()));
      std::printf("streamed_ok=%d streamed_host=%u streamed_device=%u\n", ok ? 1 : 0, fnv(streamed.data.data(), streamed.data.size()),
                  fnv(dev.data(), dev.size()));
      ok = streamed.LoadBricksToTexture(8, 2, 2, 2, 7, /*dealloc*/ true, /*toGPU*/ true);
      if (ok) vr::check(vr_volume_download(ctx->handle(), streamed.volume, dev.data()));
      std::printf("streamed_dealloc=%d,%u,%u\n", ok ? 1 : 0, (unsigned)streamed.data.size(), fnv(dev.data(), dev.size()));
      bool gone = !streamed.LoadBricksToTexture(8, 2, 2, 2, /*timestep*/ 8, true, true);  // no files for timestep 8
      std::printf("streamed_missing_returns_false=%d\n", gone ? 1 : 0);
    }

    {  // the same build on the GPU gives the same stream
      vr::VolumeKdtree gpuTree(volume.data, volume.dataDims[0], volume.dataDims[1], volume.dataDims[2]);
      gpuTree.setMaxEpochs(2);
      gpuTree.setErrorTolerance(1);
      gpuTree.build(*ctx);
      std::printf("gpu_build_equal=%d\n", (int)(gpuTree.tree == myTree.tree && gpuTree.distanceMap == myTree.distanceMap &&
                                                 gpuTree.numActiveNodes == myTree.numActiveNodes && gpuTree.stats.maxError == myTree.stats.maxError));
    }
    std::vector<vr::byte> treeData;
    {
      std::vector<vr::byte> viaCtx;
      myTree.levelCut(myTree.maxTreeDepth, viaCtx, *ctx);  // explicit context
      myTree.levelCut(myTree.maxTreeDepth, treeData);      // main.cpp:280-281, the reference's two-argument call (shared device context)
      std::printf("levelcut_ctx_equal=%d\n", (int)(viaCtx == treeData));
    }
    std::printf("decoded=%u maxerr_measured=%d\n", fnv(treeData.data(), treeData.size()), myTree.measureMaxError());
    try { std::vector<vr::byte> o; myTree.levelCut(10, o, *ctx); std::puts("cut10=nothrow"); }
    catch (const vr::Error& e) { std::printf("cut10=error%d\n", e.code); }
    volume.data = treeData;  // main.cpp:290 replaces the raw texture by the decoded one
    volume.transferToGPU(false);
    vr::Renderer r(*ctx, VR_MODE_COMPOSITE, 96, 72);
    r.params.step_size[0] = r.params.step_size[1] = 1.0f / 64; r.params.step_size[2] = 1.0f / 32;
    std::vector<float> rgba;
    uint64_t S = r.Draw(volume.volume, rgba);
    std::printf("composite_samples=%llu composite=%u\n", (unsigned long long)S, fnv(rgba.data(), rgba.size() * 4));
    r.params.mode = VR_MODE_ISOSURFACE; r.setIsoValue(120.0f / 255.0f);
    S = r.Draw(volume.volume, rgba);
    std::printf("iso_samples=%llu\n", (unsigned long long)S);
    delete ctx;
  } catch (const std::exception& e) {
    std::printf("exception=%s\n", e.what());
    return 1;
  }
  return 0;
}
