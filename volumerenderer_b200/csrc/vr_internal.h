// vr_internal.h -- handle definitions behind the opaque C ABI types.
#pragma once
#include <cuda_runtime.h>
#include <map>
#include <memory>
#include <set>
#include <vector>
#include "../../include/vr_b200.h"
#include "kd_decode.h"
#include "kd_decode_core.h"
#include "kd_stream.h"

namespace vr { struct KdBuilder; void kd_builder_destroy(KdBuilder*); }

struct vr_peer;

namespace vr {
// One pinned host buffer and / or one device buffer that holds the images of several streams back to back (a timestep):
// vr_stream_arena_* (one H2D copy per timestep) and vr_streams_load (one allocation + one copy per load).  Streams keep it
// alive through a shared_ptr; their hostImage / devImage point inside it and are not theirs to free.
struct ArenaMem {
  uint8_t* host = nullptr;
  uint8_t* dev = nullptr;
  size_t bytes = 0;
  int device = -1;
  ~ArenaMem();
};
}  // namespace vr

struct vr_context {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool ownsStream = false;
  cudaEvent_t ev[16][2] = {};
  cudaEvent_t evOrder = nullptr;  // recorded by vr_context_wait_for(other = this)
  vr::KdBuilder* builder = nullptr;  // scratch of the GPU stream producer (kd_build.h), created on first use
  std::map<uint64_t, uint32_t*> chainLuts;   // chainDist -> device table of 16384 entries (classic kernel: vr_chain_lut_entry)
  std::map<uint64_t, uint32_t*> chainLuts6;  // same for the persistent kernel (vr6::chain_entry)
  // device-side frame of vr_render_to_host (RGBA, hit indices, sample counter): kept between calls, grown on demand, so
  // that one Draw() per frame (main.cpp:404) costs no cudaMalloc / cudaFree
  float* frameDev = nullptr; int32_t* frameHitDev = nullptr; uint64_t* frameCountDev = nullptr;
  size_t framePixels = 0, frameHitPixels = 0;
  // pinned staging of vr_streams_load (grow-only; stagingDone = the last copy out of it has finished)
  uint8_t* stagingHost = nullptr; size_t stagingBytes = 0; cudaEvent_t stagingDone = nullptr;
  vr::Decode6Side side{};  // side stream + fork / join events of the split decode launch (kd_decode.h), created with the first plan
};

// Device image of a prepared stream: one arena, sections 16-byte aligned:
//   [words (+16 B zero pad) | tileLeafBlock | leafOff | tileVal | leafVal]
struct vr_stream {
  vr::KdStream kd;
  uint8_t* hostImage = nullptr;   // pageable (new[]) or pinned (cudaHostAlloc)
  bool pinned = false;
  size_t imageBytes = 0;
  size_t offWords = 0, offTile = 0, offLeafOff = 0, offTileVal = 0, offLeafVal = 0;
  size_t wordsBytes = 0;          // stream section (algorithmic stream bytes, padded)
  uint8_t* devImage = nullptr;
  int device = -1;
  std::shared_ptr<vr::ArenaMem> arena;    // set when hostImage and / or devImage live inside a shared arena
  bool hostOwned = true, devOwned = true;  // false: the pointer belongs to `arena`
  void release();
  void drop_host();  // forget / free the host image
  void drop_dev();   // forget / free the device image
};

struct vr_volume {
  uint8_t* dev = nullptr;
  int64_t SX = 0, SY = 0, SZ = 0;
  bool owned = false;
  int device = 0;
  // occupancy for empty-space skipping (occupancy.h); grids are allocated on first use
  bool skipping = true;           // vr_volume_set_skipping
  int64_t cellOrigin[3] = {0, 0, 0};  // global voxel coordinate of voxel (0,0,0): cells align to the global grid
  int shift[3] = {0, 0, 0};       // cellOrigin mod cell size
  int nc[3] = {0, 0, 0};          // level-0 cells per axis
  int occLevels = 0;              // pyramid levels (level l has ((nc - 1) >> l) + 1 cells per axis)
  int occOff[12] = {0};           // byte offset of each level inside occPyr
  uint8_t *occRaw = nullptr, *occPyr = nullptr;  // raw: 8 writer slots per cell; pyr: dilated level 0 + octree of maxima
  bool occValid = false;          // occRaw bounds every stored voxel
  bool occDilDirty = true;        // occPyr needs recomputing from occRaw
  uint8_t* occLvl = nullptr;      // per level-0 cell: radius of the void cube of cells around it for threshold occLvlThr (occupancy.h)
  int occLvlThr = -2;             // -2 = occLvl is stale
  bool occPyrHigh = false;        // the octree levels above level 0 of occPyr are current (built on demand)
  int occEpoch = 0;               // bumped whenever the grids are freed (plans remember the epoch they were built for)
  // Plans whose last run was lazy (vr_render_compressed_lazy): the tiles their frame could not read hold stale voxels,
  // so every entry point that reads stored voxels (vr_render, vr_volume_download, vr_kd_compress_volume, vr_volume_pack_box,
  // vr_partial_*, occupancy rebuilds from the voxels) refuses the volume until each of these plans has been run in full
  // again.  partialOrphan: such a plan was destroyed while its tiles were stale (cleared by vr_volume_invalidate or by a
  // load that covers the whole volume).
  std::set<const void*> partial;
  bool partialOrphan = false;
  bool is_partial() const { return partialOrphan || !partial.empty(); }
  // Plans that decode into this volume.  vr_volume_destroy detaches them (their dst becomes null and they fail with
  // VR_ERR_STATE from then on), so a host that releases its handles in any order -- a garbage collector does -- is safe.
  std::set<struct vr_decode_plan*> plans;
  void occ_free();
};

struct vr_decode_plan {
  vr_context* ctx = nullptr;
  int device = 0;
  vr_kd_geom geom;
  vr_kd_job* jobsDev = nullptr;
  int n = 0;
  vr_volume* dst = nullptr;
  bool vec16 = false;
  bool occFused = false;  // every tile of every job is one occupancy cell: the kernel keeps occRaw current
  int occEpoch = 0;
  int32_t* tileCellDev = nullptr;  // [2^Dt] tile -> cell offset for dst's grid
  // persistent kernel (kd_decode.h): two device-side queues
  bool persistent6 = false;
  vr::Decode6Plan p6;
  vr_parse_item* parseDev = nullptr;
  int nParse = 0;
  int tmaOut = 0;  // the units' voxels leave through the TMA unit (launches dominated by constant groups)
  uint32_t* queueDev = nullptr;
  const uint32_t* lut6Dev = nullptr;
  // classic kernel: one CTA per entry (kd_worklist.h)
  vr_kd_work* workDev = nullptr;
  int nWork = 0;
  vr_fill_job* fillJobsDev = nullptr;  // constant groups (persistent: the fill queue; classic: attached to parse CTAs)
  int nFill = 0;
  int64_t bytesStream = 0, bytesVoxels = 0, bytesIndex = 0;
  // host copies kept for the lazy (view-dependent) work list, kd_cull.h
  std::vector<vr_stream*> streams;
  std::vector<int64_t> origins, dstOffset, occBase;
  std::vector<uint64_t> treeDist;
  std::vector<int32_t> tileCellHost;
  struct Lazy {
    bool built = false;
    int mode = 0, thr = 0, reach[3] = {0, 0, 0}, stepReach[3] = {0, 0, 0};   // cache key
    uint8_t* solidDev = nullptr;  // per occupancy cell of dst: 1 = "solid-high" look-up cell (kd_cull.h), isosurface only
    vr_parse_item* parseDev = nullptr; int nParse = 0; int tmaOut = 0;
    vr_kd_work* workDev = nullptr; int nWork = 0;
    vr_fill_job* fillJobsDev = nullptr; int nFill = 0;
    int64_t* occSetDev = nullptr; int nOccSet = 0;  // (cell, max) pairs of the tiles left out
    int64_t tilesTotal = 0, tilesWritten = 0, bytesVoxels = 0, bytesStream = 0;
  } lazy;
  Lazy subset;  // lists of the last vr_decode_plan_run_tiles
  void lazy_free(Lazy* which = nullptr);
};

// slab geometry of all ranks + the visibility orders (vr_capi_sortlast.cu)
struct vr_sortlast {
  int device = 0;
  int n = 0, rank = 0;
  std::vector<vr_slab_desc> slabs;
  int order[8][64 /* VR_MAX_RANKS, render_march.h */];
  int* orderDev = nullptr;  // 8 x n
};

namespace vr {
extern int g_count_fetched;  // vr_debug_set("count_fetched"), vr_capi_render.cu
int fail(int code, const char* msg);
int fail_cuda(cudaError_t e, const char* where);
int stream_build_image(vr_stream* s);  // prepare + pack host image
int stream_image_layout(vr_stream* s);                 // prepare + section offsets / imageBytes (no allocation)
void stream_image_write(const vr_stream* s, uint8_t* dst);  // the image (imageBytes) into dst
int context_chain_lut(vr_context* ctx, uint64_t chainDist, const uint32_t** out);
struct OccView;
struct PeerTargets;
// fills t for a fused kernel that signals `slot` at every rank (vr_capi_peer.cu); windowBytes = usable window size
int peer_targets(struct ::vr_peer* p, int slot, PeerTargets* t, int64_t* windowBytes);
int volume_occ_alloc(vr_volume* v);                       // allocate the grids for the current cellOrigin
int volume_occ_view(vr_context* ctx, const vr_volume* v, int thr, OccView* out, bool wantPyramid = false);  // build / refresh as needed; out->dil = null when off
// the march of vr_render without the partial-volume check (lazy path)
// solidMaskDev (optional, one byte per occupancy cell): look-up cells where the iso crossing test cannot fire because
// every reachable voxel is >= isoValue (kd_cull.h); the march skips the fetch there as it does in void cells
int render_volume(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_dev, int32_t* hit_dev,
                  uint64_t* samples_dev, bool allowPartial, const uint8_t* solidMaskDev = nullptr);
int skip_threshold_of(const vr_render_params* params, int* thr, int* mode);  // occupancy threshold of the frame's mode
// (re)builds plan->lazy for (thr, reach): the work list without the tiles the frame cannot read (kd_cull.h)
int plan_lazy_prepare(vr_decode_plan* plan, int mode, int thr, const int reach[3], const int stepReach[3]);
int plan_lazy_run(vr_decode_plan* plan, vr_decode_plan::Lazy* which = nullptr);  // decode launch over the lazy list + maxima of the tiles left out
void volume_occ_written(vr_context* ctx, vr_volume* v, const int64_t lo[3], const int64_t ext[3]);  // a box of voxels was overwritten
}
