/* vr_b200.h -- C ABI of the B200-native VolumeRenderer hot path (libvr_b200.so).
 *
 * Drop-in boundary for the reference's three C++ surfaces; every entry point cites
 * the reference interface it replaces (paths relative to /root/reference/volume_renderer/).
 * Conventions: plain pointers and sizes only; every function returns an int status (VR_OK = 0,
 * never throws or exits across the boundary -- the reference exit(-1)s on a missing tree file,
 * VolumeKdTree_recover.cpp:560-565, and throws on a brick size mismatch, VolumeReader.h:258-260);
 * host buffers are caller-owned; device buffers are library-owned unless wrapped; one CUDA stream
 * per context; handles are not thread-safe.  There is no CPU fallback: decode and render fail with
 * VR_ERR_CUDA when no device is usable.
 */
#ifndef VR_B200_H
#define VR_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VR_OK 0
#define VR_ERR_ARG 1
#define VR_ERR_IO 2
#define VR_ERR_FORMAT 3
#define VR_ERR_CUDA 4
#define VR_ERR_UNSUPPORTED 5
#define VR_ERR_STATE 6

typedef struct vr_context vr_context; /* device ordinal + CUDA stream                          */
typedef struct vr_stream vr_stream;   /* one kd-tree stream = state of a reference VolumeKdtree */
typedef struct vr_volume vr_volume;   /* device slab, x fastest = the reference's 3-D texture   */
typedef struct vr_decode_plan vr_decode_plan;

/* 88-byte container header exactly as VolumeKdtree::save writes it (VolumeKdTree_recover.cpp:535-542) */
#pragma pack(push, 1)
typedef struct vr_stream_header {
  int64_t rootMin[3];
  int64_t rootMax[3];
  int32_t maxTreeDepth;
  int32_t origTreeDepth;
  int64_t X, Y, Z;
  int64_t numActiveNodes;
} vr_stream_header;
#pragma pack(pop)

typedef struct vr_build_stats {
  int32_t maxErrorBeforeGrowth; /* "Before branch growth, Maximum Error" VolumeKdTree_recover.cpp:76 */
  int32_t maxError;             /* "Maximum Error" :120 */
  double meanL1, meanL2;        /* :129 */
  int64_t zeroRunRewrites;      /* firings of the zeroStartIdx rewrite (:684-689); expected 0 */
} vr_build_stats;

const char* vr_last_error(void); /* message of the last failing call on this thread */
const char* vr_version(void);

/* ---- context ------------------------------------------------------------------------------- */
/* Replaces the GL context + window set-up of main.cpp:84-135 with a headless CUDA device binding. */
int vr_context_create(int device, vr_context** out);
int vr_context_destroy(vr_context* ctx);
/* Use an externally created cudaStream_t (e.g. torch's current stream) for all later calls. */
int vr_context_set_stream(vr_context* ctx, void* cuda_stream);
int vr_context_synchronize(vr_context* ctx);
/* Cross-stream ordering for pipelines (e.g. a copy context uploading timestep t+1 while the render context
 * decodes and draws timestep t): work submitted to `ctx` after this call starts only after everything
 * submitted to `other` before this call has finished.  Device-side wait; the host does not block. */
int vr_context_wait_for(vr_context* ctx, vr_context* other);
/* Asynchronous device -> pinned/pageable host copy on the context stream (frame read-back in a pipeline). */
int vr_context_download(vr_context* ctx, void* dst_host, const void* src_dev, int64_t nbytes);

/* CUDA-event timing on the context stream (replaces DebugTimer::Begin/End, DebugTimer.cpp:6-35).
 * slot in [0,16): begin/end record events asynchronously; elapsed synchronises on the end event. */
int vr_timer_begin(vr_context* ctx, int slot);
int vr_timer_end(vr_context* ctx, int slot);
int vr_timer_elapsed_ms(vr_context* ctx, int slot, float* ms);

/* ---- stream container: VolumeKdtree::save / open (VolumeKdTree_recover.cpp:521-594) ---------- */
int vr_stream_from_memory(const vr_stream_header* hdr, const uint8_t* distanceMap, const uint8_t* bits,
                          int64_t nbytes, vr_stream** out);
int vr_stream_open(const char* path, vr_stream** out); /* accepts exact and +8-byte-slack files (SURVEY C-3) */
int vr_stream_save(const vr_stream* s, const char* path);
int vr_stream_destroy(vr_stream* s);
/* public fields of class VolumeKdtree (VolumeKdtree_recover.h:57-79) */
int vr_stream_get_header(const vr_stream* s, vr_stream_header* hdr);
int vr_stream_get_distance_map(const vr_stream* s, uint8_t* out, int32_t cap); /* maxTreeDepth+1 bytes */
int vr_stream_get_bits(const vr_stream* s, const uint8_t** bits, int64_t* nbytes); /* TwoBitArray::bits */
/* full grammar walk: VR_OK iff exactly numActiveNodes symbols parse */
int vr_stream_validate(const vr_stream* s, int64_t* nodesConsumed);
/* builds the derived decode index (side table of subtree offsets); idempotent */
int vr_stream_prepare(vr_stream* s);
/* bytes of the device image: [0] stream words, [1] index */
int vr_stream_device_bytes(const vr_stream* s, int64_t out[2]);
/* page-lock the host image so uploads run at full PCIe rate */
int vr_stream_pin(vr_stream* s);
/* H2D copy of stream + index (asynchronous on the context stream); implies prepare */
int vr_stream_upload(vr_context* ctx, vr_stream* s);

/* VolumeKdtree::open for a whole timestep (main.cpp:276-284 opens one tree per brick): `threads` host threads (0 = all
 * cores) open the n container files -- plain containers are walked to build their decode index, indexed containers
 * (vr_stream_save_indexed) bring it along -- and write the device images straight into the context's pinned staging
 * buffer (kept between calls); then ONE device allocation and ONE H2D copy carry the whole set (ctx may be NULL: open +
 * prepare only).  out receives n streams that share the allocation; on any failure every stream is destroyed and the
 * first error is returned (VR_ERR_IO missing / short file, VR_ERR_FORMAT). */
int vr_streams_load(vr_context* ctx, const char* const* paths, int32_t n, int32_t threads, vr_stream** out);
/* The prepared images of n streams (a timestep) moved into ONE pinned host buffer and ONE device buffer, so that a
 * timestep travels with one H2D copy instead of one per stream (VolumeReader::transferToGPU, VolumeReader.h:114-138, is
 * one glTexImage3D of the assembled array).  Create the arena BEFORE the decode plans of these streams (their device
 * addresses change).  The buffers live until the arena handle and every stream in it have been destroyed. */
typedef struct vr_stream_arena vr_stream_arena;
int vr_stream_arena_create(vr_context* ctx, vr_stream* const* streams, int32_t n, vr_stream_arena** out);
int vr_stream_arena_upload(vr_context* ctx, vr_stream_arena* a); /* asynchronous on the context stream */
int vr_stream_arena_bytes(const vr_stream_arena* a, int64_t* bytes);
int vr_stream_arena_destroy(vr_stream_arena* a);

/* ---- MidRangeTree's 4-bit/node shader stream (MidRangeTree::convertToByteArray, MidRangeTree.cpp:1095-1128) ----
 * byte k = tree[2k]<<6 | range[2k]<<4 | tree[2k+1]<<2 | range[2k+1]; length = next power of two >= ceil(n/2),
 * zero tail.  tree / range are TwoBitArray bit arrays (LSB first) of n symbols.  Returns the packed length
 * through *out_len; out may be NULL to query it. */
int vr_nibbles_pack(const uint8_t* tree, const uint8_t* range, int64_t n, uint8_t* out, int64_t cap, int64_t* out_len);
/* Inverse: split n nodes of a 4-bit stream into the two TwoBitArray bit arrays (ceil(n/4) bytes each). */
int vr_nibbles_unpack(const uint8_t* nibbles, int64_t nbytes, int64_t n, uint8_t* tree_out, uint8_t* range_out);
/* A decodable stream straight from the 4-bit shader stream: the midpoint codes are the codes levelCut walks
 * (MidRangeTree::levelCut decodes the midpoint tree only, MidRangeTree.cpp:984); range_out (optional,
 * ceil(numActiveNodes/4) bytes) receives the half-range codes. */
int vr_stream_from_nibbles(const vr_stream_header* hdr, const uint8_t* distanceMap, const uint8_t* nibbles, int64_t nbytes,
                           vr_stream** out, uint8_t* range_out);

/* ---- codec: VolumeKdtree(in,x,y,z) + setErrorTolerance + setMaxEpochs + build ----------------- */
/* (VolumeKdtree_recover.h:103,126,167,175; VolumeKdTree_recover.cpp:17-140).  CPU, reference
 * algorithm, byte-identical streams; `brick` is not modified (the reference empties its input). */
int vr_kd_compress(const uint8_t* brick, int64_t X, int64_t Y, int64_t Z, int tolerance, int maxEpochs,
                   vr_stream** out, vr_build_stats* stats);
/* The same producer on the GPU (csrc/kd_build.cu): level-parallel kernels, byte-identical streams
 * (the two sequential steps of build -- the running-mean estimate of encodeNodeEstimate :415-455 and the
 * order-dependent double accumulation of compressGradientDescent :206-384 -- are reproduced exactly, see
 * csrc/kd_build_core.h).  `brick` is host memory (uploaded); synchronous. */
int vr_kd_compress_device(vr_context* ctx, const uint8_t* brick, int64_t X, int64_t Y, int64_t Z, int tolerance, int maxEpochs,
                          vr_stream** out, vr_build_stats* stats);

/* ---- volume (device slab) --------------------------------------------------------------------- */
/* Replaces VolumeReader::data + textureId (VolumeReader.h:42-51): an SX*SY*SZ uint8 array. */
int vr_volume_create(vr_context* ctx, int64_t SX, int64_t SY, int64_t SZ, vr_volume** out);
int vr_volume_wrap(vr_context* ctx, void* device_ptr, int64_t SX, int64_t SY, int64_t SZ, vr_volume** out);
int vr_volume_destroy(vr_volume* v);
int vr_volume_device_ptr(const vr_volume* v, void** ptr);
int vr_volume_dims(const vr_volume* v, int64_t dims[3]);
int vr_volume_download(vr_context* ctx, const vr_volume* v, uint8_t* host); /* synchronous */
/* LoadBrickToTexture / LoadBricksToTexture raw path (VolumeReader.h:91-107,151-223): place one raw
 * brick (BX*BY*BZ, x fastest) at voxel origin (ox,oy,oz) of the slab; 64-bit indexing (SURVEY C-12). */
int vr_volume_upload_brick(vr_context* ctx, vr_volume* v, const uint8_t* brick, int64_t BX, int64_t BY,
                           int64_t BZ, int64_t ox, int64_t oy, int64_t oz);
/* LoadBricksToTexture (VolumeReader.h:151-223) as one call: brick b of `numBricks` raw files (one X*Y*Z uint8 file per
 * brick, path = find(b, timestep, user) = the reference's findSourceFile, main.cpp:580-597) goes to grid position
 * brick_ijk[3b..] (NULL = fillVolumeBrickMap's b -> (b % I, (b / I) % J, b / (I*J)), main.cpp:599-619) of the slab, 64-bit
 * offsets (SURVEY C-12).  Pipelined: a reader thread fills a ring of pinned staging buffers while the context's stream
 * copies each brick straight to its place (no assembled host array).  host_data (optional, I*X * J*Y * K*Z bytes)
 * receives the assembled array as well (= VolumeReader::data).  Synchronous.  A missing file gives VR_ERR_IO (the
 * reference returns false), a file whose size is not X*Y*Z gives VR_ERR_FORMAT (the reference throws, :258-260).
 * stats (optional): [0] bricks copied, [1] bytes copied. */
typedef const char* (*vr_find_source_file)(int32_t brick, int32_t timestep, void* user);
int vr_volume_load_bricks(vr_context* ctx, vr_volume* v, int64_t numBricks, int64_t I, int64_t J, int64_t K, int32_t timestep,
                          const int64_t brick_dims[3], vr_find_source_file find, void* user, const int64_t* brick_ijk,
                          uint8_t* host_data, int64_t stats[2]);
/* GPU producer on a brick that already lives in a device volume (e.g. one just decoded, simulated or uploaded):
 * compresses the X*Y*Z box at `origin` without a host round trip of the voxels.  fixed_point_iterations (optional,
 * 2 entries): scan passes of the running-mean fixed point, total and worst level. */
int vr_kd_compress_volume(vr_context* ctx, const vr_volume* vol, const int64_t origin[3], int64_t X, int64_t Y, int64_t Z,
                          int tolerance, int maxEpochs, vr_stream** out, vr_build_stats* stats, int64_t* fixed_point_iterations);

/* ---- empty-space skipping (north_star; no usable bound structure exists in the reference) --
 * Every volume carries a grid of exact per-cell maxima (cells of 32x32x16 voxels aligned to the global
 * voxel grid, plus a dilated copy and an octree of maxima above it).  The decode kernel maintains it for the
 * bricks it writes; raw uploads are reduced lazily at the next render.  The march skips the fetch where
 * the bound proves the sample is exactly 0 (composite: contributes nothing, raycaster.frag:69-72) or
 * below the iso value (isosurface.frag:126 cannot fire); sample positions are still produced by the
 * reference's repeated additions, so frames are bit-identical with skipping on or off. */
int vr_volume_set_cell_origin(vr_volume* v, const int64_t origin[3]); /* global coordinate of voxel (0,0,0); default 0 */
int vr_volume_set_skipping(vr_volume* v, int enable);                 /* default on */
int vr_volume_invalidate(vr_volume* v); /* the caller wrote voxels behind the library's back (wrapped memory) */
/* which: 0 raw cell maxima, 1 + l = level l of the octree of dilated maxima (VR_ERR_ARG past the root);
 * out_host may be NULL to query dims only */
int vr_volume_occupancy_download(vr_context* ctx, const vr_volume* v, int32_t which, uint8_t* out_host, int64_t cap, int32_t dims[3]);

/* ---- stage 1: decode = VolumeKdtree::levelCut (VolumeKdTree_recover.cpp:726-835) --------------- */
/* Plan a batched decode of n uploaded streams (all of one brick shape) into `dst`; origins holds
 * 3*n voxel coordinates (brick b of the reference grid sits at (i*X, j*Y, k*Z), main.cpp:599-619). */
int vr_decode_plan_create(vr_context* ctx, vr_stream* const* streams, int32_t n, const int64_t* origins,
                          vr_volume* dst, vr_decode_plan** out);
int vr_decode_plan_run(vr_decode_plan* plan); /* one kernel launch, asynchronous */
int vr_decode_plan_destroy(vr_decode_plan* plan);
/* Random-access decode (levelCutRecursive at an arbitrary node, HashedKdtree.cpp:254-336): decode only the
 * tiles with needed[b * tiles + t] != 0 (tiles = vr_stream_index_info out[0], kd path order; tile boxes from
 * vr_lazy_tile_mask).  Those tiles hold exactly the voxels levelCut produces there; every other tile is left alone, except
 * constant tiles in the x-aligned fill group of a needed tile, which receive their (correct) voxels too.  The occupancy
 * grid receives the exact maxima of the tiles left out.  Afterwards the destination is PARTIAL (see
 * vr_render_compressed_lazy).  Needs 32x32x16 tiles aligned with the destination's occupancy cells.  Streams opened with a
 * tile subset must hold every non-constant tile asked for (VR_ERR_STATE otherwise).  stats as vr_render_compressed_lazy. */
int vr_decode_plan_run_tiles(vr_decode_plan* plan, const uint8_t* needed, int64_t stats[4]);
/* algorithmic bytes of one run: [0] stream bytes read, [1] voxel bytes written, [2] index bytes read */
int vr_decode_plan_bytes(const vr_decode_plan* plan, int64_t out[3]);
/* which kernel the plan launches: 0 = persistent warp-specialised kernel (bricks with 4x4x4 leaf boxes such as the
 * reference's 256x256x128, 16-byte aligned placement), 1 = classic kernel, one CTA per tile (every other shape) */
int vr_decode_plan_kernel(const vr_decode_plan* plan, int32_t* which);
/* Debug knobs for A/B measurements; explicit, process-wide, read when a plan is created or run -- the library never
 * reads the environment.  "decode_kernel": 0 = choose (default), 1 = classic kernel always.  "decode_only": 0 = whole
 * launch, 1 = parse tiles only, 2 = constant groups only (timing probes; the destination is then incomplete).
 * "fill_attach": constant groups per parse CTA of the classic kernel (default 2).  Persistent kernel: "unit_store" 2 = a
 * unit's voxels as 16-byte stores of the lanes instead of one tensor store; "fill_mode" 0 / 2 = tensor stores of the constant
 * groups without / with (default) an evict-first L2 hint, + 256 * t = pace them by t * 100 ns per group while parse units are
 * queued; "tma_rows" 8 / 16 / 32 rows per constant-group tensor store; "parse_order" 1 = longest units first. */
int vr_debug_set(const char* key, int value);
int vr_debug_get(const char* key, int* value);
/* End-to-end levelCut: host stream -> host voxels (upload, decode, download; synchronous).
 * cutDepth must equal maxTreeDepth: any other depth crashes the reference and is
 * rejected with VR_ERR_UNSUPPORTED. */
int vr_kd_levelcut(vr_context* ctx, vr_stream* s, int32_t cutDepth, uint8_t* out_host);

/* ---- stage 2: ray marching = the reference's shader-selectable renderer -------------------------- */
/* Mode = which fragment shader main.cpp:71-74 would load. */
#define VR_MODE_COMPOSITE 0             /* raycaster.frag */
#define VR_MODE_ISOSURFACE 1            /* isosurface.frag */
#define VR_MODE_ISOSURFACE_COMPRESSED 2 /* isosurface_compressed.frag (a stub in the reference):
                                           defined here as isosurface.frag evaluated on the volume that
                                           levelCut produces from the streams, decoded on the device */

/* Uniforms + camera of main.cpp:33-40,52,319-341,396-397.  The whole volume maps to the unit cube. */
typedef struct vr_render_params {
  float eye[3], front[3], up[3]; /* camPos, camFront, camUp */
  float vfov_deg, znear, zfar;   /* perspectiveFov(radians(fov), w, h, 0.1, 100) */
  float step_size[3];            /* uniform step_size */
  float isoValue;                /* uniform isoValue */
  int32_t width, height;         /* framebuffer size */
  int32_t mode;
} vr_render_params;
int vr_render_defaults(vr_render_params* p); /* the reference's constants */
/* derived camera: eye[3], right[3], up[3], forward[3], tw, th, znear, zfar, step[3], iso (20 floats) */
int vr_camera_basis(const vr_render_params* p, float out[20]);
/* One frame = one UnitBrick::Draw() (main.cpp:404).  Output: width*height RGBA float32 in [0,1], row 0
 * = bottom row (GL convention), written to DEVICE memory; hit_dev (optional) = sample index of the
 * iso hit or -1; samples_dev (optional) += number of texture-fetching march steps.  Asynchronous. */
int vr_render(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_dev,
              int32_t* hit_dev, uint64_t* samples_dev);
/* Same, result copied to HOST buffers; synchronous. */
int vr_render_to_host(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_host,
                      int32_t* hit_host, uint64_t* samples_host);
/* The frame of vr_render_to_host without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  vr_render_to_host allocates and frees its device-side frame per call; this entry point marches into a
 * file-scope __device__ array of VR_STATIC_SCRATCH_BYTES that is part of the library image (every GPU that loads the
 * library holds it until the process exits) and calls no cudaMalloc / cudaFree.  The caller still passes the volume and
 * the host outputs; same arguments, same results (frame and hit indices bit for bit, same sample count; the march runs
 * without the occupancy grid, whose construction allocates).  width * height above VR_STATIC_MAX_PIXELS fails with
 * VR_ERR_UNSUPPORTED.  All calls share the array, which keeps the last call's contents (the sample counter is cleared
 * per call): calls that overlap on one GPU race on it.  Synchronous.  csrc/vr_capi_noalloc.cu. */
#define VR_STATIC_MAX_PIXELS 2097152 /* 2 Mi pixels: 1600 x 1200 (main.cpp:27) fits */
#define VR_STATIC_SCRATCH_BYTES (VR_STATIC_MAX_PIXELS * 20 + 256) /* RGBA float32 + hit indices + sample counter = 40 MiB */
int vr_render_to_host_static(vr_context* ctx, const vr_volume* vol, const vr_render_params* params, float* out_rgba_host,
                             int32_t* hit_host, uint64_t* samples_host);
int64_t vr_static_scratch_max_pixels(void);
/* ISOSURFACE_COMPRESSED / any mode straight from streams: run the decode plan, then march its
 * destination volume, all on the device with no host round trip. */
int vr_render_compressed(vr_context* ctx, vr_decode_plan* plan, const vr_render_params* params, float* out_rgba_dev,
                         int32_t* hit_dev, uint64_t* samples_dev);

/* Compressed-domain frame with a view-dependent ("lazy") decode -- the intent behind the reference's stub
 * isosurface_compressed.frag:23-29 (`node{idx, boundMin, boundMax, scalarMin, scalarMax}`) and the never-enabled
 * stream-as-SSBO path of main.cpp:203-237.  The stream index carries the exact maximum of every
 * 32x32x16 tile (vr_stream_tile_bounds), so before anything is decoded the library knows which tiles a frame of
 * `params` can read: the march skips every lattice position whose occupancy cell is void for the mode's threshold
 * (composite: maximum 0; isosurface: maximum below isoValue), and what it reads otherwise lies within a fixed reach
 * of a non-void cell (csrc/kd_cull.h).  Only those tiles are decoded; the others are not written at all and only their
 * maxima enter the occupancy grid.  The frame is bit-identical to vr_render_compressed.  Afterwards the destination
 * volume is PARTIAL (valid where this frame could read): vr_render on it fails with VR_ERR_STATE until
 * vr_decode_plan_run has decoded the plan in full again.  The culled work list is cached in the plan per
 * (threshold, step_size).  stats (optional): [0] tiles of the plan, [1] tiles written, [2] voxel bytes written,
 * [3] stream bytes read. */
int vr_render_compressed_lazy(vr_context* ctx, vr_decode_plan* plan, const vr_render_params* params, float* out_rgba_dev,
                              int32_t* hit_dev, uint64_t* samples_dev, int64_t stats[4]);
/* Host-only view of the same decision (no device needed): needed_out[b * tiles + t] = 1 iff a frame of `params` over a
 * volume of dims[] can read tile t of brick b placed at origins[3b..] (multiples of the 32x32x16 cell).
 * tile_boxes_out (optional) = voxel origin of every tile in the volume, 3 int32 per tile; tiles_per_brick (optional).
 * needed_out may be NULL to query tiles_per_brick only. */
int vr_lazy_tile_mask(vr_stream* const* streams, int32_t n, const int64_t* origins, const int64_t dims[3], const vr_render_params* params,
                      uint8_t* needed_out, int32_t* tile_boxes_out, int32_t* tiles_per_brick);
/* Exact minimum / maximum of the voxels levelCut would produce in every decode tile, read off the stream index
 * (implies prepare; nothing is decoded).  2^Dt entries each in tile (kd path) order; tile_dims (optional) = tile extent. */
int vr_stream_tile_bounds(vr_stream* s, uint8_t* mins, uint8_t* maxs, int64_t cap, int32_t tile_dims[3]);

/* ---- MidRangeTree::save / open (MidRangeTree.cpp:753-786, 788-833; layout SURVEY A.3) ---------------------------------
 * File = the 88-byte header, distanceMap, distanceMap_range (maxTreeDepth+1 bytes each), tree.bits, tree_range.bits
 * (tree_nbytes each).  vr_mrt_save writes what the reference's save() writes, byte for byte.  vr_mrt_open reads such a
 * file: the midpoint stream comes back as a decodable stream handle -- MidRangeTree::levelCut == VolumeKdtree::levelCut on it,
 * MidRangeTree.cpp:908-1017 -- and the range stream and its distance map as plain bytes (they feed convertToByteArray,
 * vr_nibbles_pack).  The reference's own open() mis-sizes both arrays by 4 bytes (one int64 missing from its header sum,
 * :815) and so reads tree_range misaligned (SURVEY C-3); that is not reproduced.  Call vr_mrt_open with NULL output
 * buffers first to learn tree_nbytes (the returned stream must still be destroyed). */
int vr_mrt_save(const char* path, const vr_stream_header* hdr, const uint8_t* distanceMap, const uint8_t* distanceMap_range,
                const uint8_t* tree_bits, const uint8_t* range_bits, int64_t tree_nbytes);
int vr_mrt_open(const char* path, vr_stream** mid_out, uint8_t* distanceMap_range_out, int32_t dm_cap, uint8_t* range_bits_out,
                int64_t cap, int64_t* tree_nbytes);

/* ---- indexed ("random-access") container -------------------------------------------------------------
 * What HashedKdtree.cpp:96-214 (build), :254-336 (levelCutRecursive), :406-507 (compress), :528-545 (save) was reaching
 * for: decode any part of a brick without parsing what lies in front of it.  The reference's Morton-keyed hash with
 * collision overflow and a std::random_device seed (HashedKdtree.h:82-84) crashes and is not reproduced;
 * the container keeps the stream of VolumeKdtree::save bit for bit and stores the decode index of vr_stream_prepare and
 * the exact per-tile bounds beside it (layout: csrc/kd_kdx.cpp).  Deterministic: a pure function of the stream.
 *   vr_stream_save_indexed    writes it (prepares the index first if the stream has none).
 *   vr_stream_open_indexed    opens it WITHOUT walking the stream.  tile_mask (optional, 2^Dt bytes in tile order, see
 *                             vr_stream_index_info) reads only the byte ranges of those tiles from disk; such a stream
 *                             decodes through vr_decode_plan_run_tiles / vr_render_compressed_lazy only, and only the
 *                             tiles it holds.  verify != 0 checks the stream checksum (full opens).
 *                             bytes_read (optional) = bytes fetched from the file.
 *   vr_stream_from_memory_indexed  the same from a buffer holding the file image.
 *   vr_stream_index_info      out[0] tiles per brick (2^Dt), [1] live (non-constant) tiles, [2] leaf subtrees per tile,
 *                             [3] 1 when only a tile subset is resident, [4] index bytes, [5] longest leaf segment (symbols).
 *   vr_stream_index_section   read-only view of one index array (implies prepare): 1 tileLeafBlock u32, 2 tileVal u8,
 *                             3 tileMin u8, 4 tileMax u8, 5 blockSymBegin u32, 6 blockSymEnd u32, 7 leafOff u32, 8 leafVal u8. */
int vr_stream_save_indexed(vr_stream* s, const char* path);
int vr_stream_open_indexed(const char* path, const uint8_t* tile_mask, int32_t verify, vr_stream** out, int64_t* bytes_read);
int vr_stream_from_memory_indexed(const uint8_t* image, int64_t nbytes, const uint8_t* tile_mask, int32_t verify, vr_stream** out);
int vr_stream_serialize_indexed(vr_stream* s, uint8_t* out, int64_t cap, int64_t* nbytes); /* out may be NULL to query the size */
int vr_stream_index_info(vr_stream* s, int64_t out[6]);
int vr_stream_index_section(vr_stream* s, int32_t which, const void** data, int64_t* nbytes);

/* ---- sort-last rendering of a volume sharded over GPUs (no counterpart in the reference, which
 * renders one texture on one GPU, main.cpp:290,404) ------------------------------------
 * One process per GPU; rank r holds a box of the whole texture (its bricks, decoded in place, plus a
 * halo) and renders the samples of the reference's ray lattice whose base texel lies in its owned
 * box.  The exchange between the passes (all-gather / all-reduce of image-sized buffers) is the
 * caller's: NCCL via torch.distributed in volumerenderer_b200/sortlast.py, or any other transport. */
typedef struct vr_slab_desc {
  int64_t global[3]; /* extent of the whole texture (what the unit cube maps to) */
  int64_t origin[3]; /* global coordinate of voxel (0,0,0) of the device slab (owned box minus halo) */
  int64_t own_lo[3]; /* owned box [own_lo, own_hi) in global coordinates; the boxes of all ranks   */
  int64_t own_hi[3]; /* partition the texture as a grid (slabs along one axis are the usual case)  */
} vr_slab_desc;
typedef struct vr_sortlast vr_sortlast;
/* ctx may be NULL (host-side queries only: vr_sortlast_order); the device tables are made on first use */
int vr_sortlast_create(vr_context* ctx, const vr_slab_desc* slabs, int32_t n, int32_t rank, vr_sortlast** out);
int vr_sortlast_destroy(vr_sortlast* sl);
/* front-to-back order of the n slabs for rays whose direction signs are octant (bit0: x<0, bit1: y<0, bit2: z<0) */
int vr_sortlast_order(const vr_sortlast* sl, int32_t octant, int32_t* out_ranks);
/* halo (voxels per axis and side) the device slab must hold around its owned box for `params`:
 * 1 for the trilinear taps of raycaster.frag:60; isosurface.frag additionally reaches one step ahead
 * (:121) and DELTA = 0.01 texture units around the hit (:18,52-59). */
int vr_sortlast_halo(const vr_render_params* params, const int64_t global[3], int64_t halo[3]);
/* pass 1 (asynchronous).  composite: cseg/tseg = W*H floats, the segment's premultiplied colour and
 * transmittance.  isosurface: rgba = W*H*4 floats (0 where no hit), hit = W*H sample indices
 * (VR_HIT_NONE where no hit). */
#define VR_HIT_NONE 0x7FFFFFFF
int vr_partial_composite(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                         float* cseg_dev, float* tseg_dev, uint64_t* samples_dev);
int vr_partial_iso(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                   float* rgba_dev, int32_t* hit_dev, uint64_t* samples_dev);
/* pass 2, composite: tseg_all_dev = the n ranks' tseg buffers, rank-major (all-gather); contrib_dev =
 * W*H*2 floats, this rank's (C, A) term.  The element-wise SUM of the terms over ranks goes to
 * vr_finish_composite, which applies the colour map of raycaster.frag:82-85. */
int vr_resolve_composite(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                         const float* tseg_all_dev, const float* cseg_dev, float* contrib_dev, uint64_t* samples_dev);
int vr_finish_composite(vr_context* ctx, const vr_render_params* params, const float* contrib_sum_dev, float* out_rgba_dev);
/* pass 2, isosurface: hit_min_dev = element-wise MIN of the ranks' hit buffers; rgba_dev is masked in
 * place so that the element-wise SUM over ranks is the finished frame. */
int vr_resolve_iso(vr_context* ctx, const vr_sortlast* sl, const vr_render_params* params, const int32_t* hit_mine_dev,
                   const int32_t* hit_min_dev, float* rgba_dev);
/* halo exchange helpers: copy a box of the slab (slab-local voxel coordinates) to / from a dense
 * x-fastest device buffer (asynchronous on the context stream). */
int vr_volume_pack_box(vr_context* ctx, const vr_volume* v, const int64_t origin[3], const int64_t extent[3], uint8_t* dst_dev);
int vr_volume_unpack_box(vr_context* ctx, vr_volume* v, const int64_t origin[3], const int64_t extent[3], const uint8_t* src_dev);

/* ---- peer-memory exchange for the sort-last frame (NVLink / NVSwitch, one process per GPU) ----------------------
 * The frame's exchange steps (halo planes, all-gather of segment transmittances, sum of the ranks' terms) move a few
 * MB between ranks that sit on one NVSwitch: as NCCL calls they cost 80-250 us each, mostly latency.  A vr_peer is a
 * window of device memory per rank that every other rank maps (CUDA IPC); data moves with ONE kernel per step that
 * stores straight into the destination windows over NVLink and then raises a flag there; the consumer waits on its own
 * flags with a one-warp kernel on its stream (bounded: a missing peer ends the wait after ~1 s and sets the error
 * flag instead of hanging the GPU).  No host synchronisation, no staging copies, no NCCL on this path.
 * Protocol: every put to (dst, slot) is matched by one wait on (slot, src) at dst, in the same order on both sides. */
typedef struct vr_peer vr_peer;
#define VR_PEER_HANDLE_BYTES 64
#define VR_PEER_SLOTS 8
int vr_peer_create(vr_context* ctx, int32_t rank, int32_t n, int64_t window_bytes, vr_peer** out);
int vr_peer_destroy(vr_peer* p);
int vr_peer_handle(const vr_peer* p, uint8_t out[VR_PEER_HANDLE_BYTES]);      /* to be all-gathered by the caller */
int vr_peer_open(vr_peer* p, const uint8_t* handles /* n * VR_PEER_HANDLE_BYTES, rank-major */);
/* same-process ranks (tests, one process driving several devices): raw pointers instead of IPC handles */
int vr_peer_raw(const vr_peer* p, void** window);
int vr_peer_open_raw(vr_peer* p, void* const* windows /* n, rank-major */);
int vr_peer_window(const vr_peer* p, int32_t rank, void** ptr);               /* rank's window as mapped in this process */
/* One kernel: for k < count copies bytes[k] from src_dev[k] (local memory) to window(dst_rank[k]) + dst_off[k], then
 * raises flag (slot, my rank) at every distinct destination.  16-byte aligned pointers, offsets and sizes. */
int vr_peer_put(vr_peer* p, int32_t slot, int32_t count, const int32_t* dst_rank, const int64_t* dst_off, const void* const* src_dev,
                const int64_t* bytes);
/* Waits (on the context's stream) until the ranks in src_rank[] have raised (slot, rank) once more. */
int vr_peer_wait(vr_peer* p, int32_t slot, int32_t count, const int32_t* src_rank);
/* out[j] = in[0][j] op in[1][j] op ... in[n-1][j] for j < count 4-byte elements, combined in index order
 * (deterministic); op 0 = float sum, 1 = int32 min.  in[] are device pointers (typically n stripes inside this rank's
 * window), count a multiple of 4. */
int vr_peer_reduce(vr_peer* p, int32_t op, int32_t n, const void* const* in, void* out, int64_t count);
int vr_peer_error(vr_peer* p, int32_t* timed_out); /* synchronises the stream; 0 = fine, else 0x10000 | slot << 8 | source rank of a wait that expired */

/* The composite passes of the sort-last frame FUSED with their exchange step: the march kernels store their per-pixel
 * results straight into the windows of the ranks that need them and raise `slot` there from their last CTA (no staging
 * buffers, no copy kernels).  Offsets are bytes inside the windows (16-byte aligned), identical on all ranks:
 *   pass 1: T_seg of this rank -> every window at gather_off + rank * 4*W*H      (then vr_peer_wait(slot, all ranks))
 *   pass 2: reads the n T_seg at gather_off of the own window; this rank's (C, A) term of pixel p goes to the owner of
 *           p's stripe, j = p / (W*H/n): window j at red_off + rank * 8*W*H/n     (then wait)
 *   pass 3: the stripe owner sums the n terms in rank order, applies the colour map of raycaster.frag:82-85 and stores
 *           the RGBA stripe into every window at frame_off                        (then wait: frame_off holds the frame) */
int vr_peer_partial_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_sortlast* sl, const vr_volume* slab,
                              const vr_render_params* params, float* cseg_dev, int64_t gather_off, uint64_t* samples_dev);
int vr_peer_resolve_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_sortlast* sl, const vr_volume* slab,
                              const vr_render_params* params, int64_t gather_off, const float* cseg_dev, int64_t red_off,
                              uint64_t* samples_dev);
int vr_peer_finish_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_render_params* params, int64_t red_off,
                             int64_t frame_off);
/* the same with the finished stripes stored into (and `slot` raised at) rank dst_rank only -- a frame that one rank displays
 * or reads back; dst_rank < 0 = every rank.  Only dst_rank waits for `slot` then. */
int vr_peer_finish_composite_to(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_render_params* params, int64_t red_off,
                                int64_t frame_off, int32_t dst_rank);

/* The all-gather of T_seg replaced by two stripe exchanges (what a rank needs of the others' T_seg is one number per
 * pixel, the transmittance T_in in front of its own segment): every rank puts stripe j of its T_seg into rank j's window at
 * gather_off + rank * 4*W*H/n; vr_peer_tin_composite (on rank j, after the wait) walks the segments of its stripe's pixels
 * front to back -- the same order table and fp32 products as vr_resolve_composite -- and stores into EVERY rank r's window,
 * at tin_off + 4 * pixel, the T_in of r's segment (-1 = the ray terminated in front of it), then raises `slot`;
 * vr_resolve_composite_tin is pass 2 on that buffer (tin_dev = own window + tin_off) and this rank's own T_seg.  Same
 * frame bit for bit, 2 (n-1)/n image-sized transfers per rank instead of n-1. */
int vr_peer_tin_composite(vr_context* ctx, vr_peer* peer, int32_t slot, const vr_sortlast* sl, const vr_render_params* params,
                          int64_t gather_off, int64_t tin_off);
int vr_resolve_composite_tin(vr_context* ctx, const vr_sortlast* sl, const vr_volume* slab, const vr_render_params* params,
                             const float* tin_dev, const float* tseg_dev, const float* cseg_dev, float* contrib_dev, uint64_t* samples_dev);

/* ---- the whole sort-last frame as ONE call (main.cpp:404: one Draw() per frame) ---------------------------------------
 * A C / C++ host of the sharded renderer needs no Python and no knowledge of the exchange protocol: it creates the peer
 * window (vr_peer_create with vr_sortlast_frame_window_bytes, handles exchanged by whatever transport it has), a
 * vr_sortlast and a vr_sortlast_frame, and then per decoded timestep calls vr_sortlast_exchange_halos and per frame
 * vr_sortlast_render.  Both are asynchronous on the context's stream and issue exactly the sequence of entry points above
 * that volumerenderer_b200/sortlast.py issues (frames are bit-identical).  Every rank makes the same calls in the same order.
 *   halo            voxels per axis and side around the owned box that every rank's device slab holds (vr_sortlast_halo);
 *                   slab origin / extent must be the owned box grown by it and clipped to the texture.
 *   halo_buffers    staging areas for halo planes (1; 2 when the host lets a neighbour run a timestep ahead).
 *   vr_sortlast_exchange_halos(part): 0 = send and receive, 1 = pack + send only, 2 = wait + unpack only (other GPU work
 *                   may be issued between the halves).
 *   vr_sortlast_render: COMPOSITE, ISOSURFACE or ISOSURFACE_COMPRESSED frame; *out_rgba_dev = width * height float4 inside
 *                   this rank's window, the finished frame on EVERY rank, valid until the next vr_sortlast_render. */
typedef struct vr_sortlast_frame vr_sortlast_frame;
int vr_sortlast_frame_window_bytes(const vr_slab_desc* slabs, int32_t n, int32_t width, int32_t height, const int64_t halo[3],
                                   int32_t halo_buffers, int64_t* out);
int vr_sortlast_frame_create(vr_context* ctx, vr_peer* peer, vr_sortlast* sl, int32_t width, int32_t height, const int64_t halo[3],
                             int32_t halo_buffers, vr_sortlast_frame** out);
int vr_sortlast_frame_destroy(vr_sortlast_frame* f);
int vr_sortlast_exchange_halos(vr_sortlast_frame* f, vr_volume* slab, int32_t buffer, int32_t part);
int vr_sortlast_render(vr_sortlast_frame* f, const vr_volume* slab, const vr_render_params* params, const float** out_rgba_dev,
                       uint64_t* samples_dev);

#ifdef __cplusplus
}
#endif
#endif /* VR_B200_H */
