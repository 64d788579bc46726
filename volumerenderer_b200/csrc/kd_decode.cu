// kd_decode.cu -- stage 1 of the hot path on the GPU: batched kd-tree stream decode into 8-bit bricks placed inside a
// device slab.  Replaces the reference's CPU VolumeKdtree::levelCut
// (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:726-835) and the brick placement of
// VolumeReader::LoadBricksToTexture (VolumeReader.h:184-205).
//
// Two kernels:
//  * kd_decode6_kernel -- the product path for bricks whose leaf boxes are 4x4x4 (256x256x128 and relatives).  Persistent,
//    one CTA per SM, warp-specialised: 22 parse warps, each a pipeline of its own over "units" (32 leaf subtrees = an
//    eighth of a 32x32x16 tile that spans the tile's whole x extent, 32 x 8 x 8 voxels, one lane per leaf subtree:
//    kd_parse6.h; the unit's stream segments (2 or 4 runs of consecutive leaf subtrees) and index entries arrive in shared
//    memory by 16-byte asynchronous copies (cp.async) of the warp's own lanes half a unit ahead, its voxels leave as one
//    tensor store of 64 whole-sector rows) and one warp that hands the constant groups to the TMA unit
//    (cp.async.bulk.tensor from a row image in shared memory).
//    Parsing is bound by instruction issue, the constant groups by HBM writes; in one CTA per work item (round 1) a CTA in
//    its fill phase parses nothing and holds its slot until its stores have drained, so the launch cost parse + fill.
//    Here the fills occupy one warp per SM and no issue slots to speak of, and the two overlap.
//  * kd_decode_tiles_kernel -- one CTA per work-list entry: any power-of-two brick shape; round 1's default, kept as the
//    challenger for A/B runs (vr_debug_set("decode_kernel", 1)).
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstring>

#include "kd_decode.h"
#include "kd_decode_core.h"
#include "kd_parse6.h"

namespace vr {

// =====================================================================================================================
// persistent kernel
// =====================================================================================================================
// Shared memory of the CTA: [chain table 64 KB | 22 staging areas of 4 864 B (two regions; a unit's dense voxel box for the
// tensor store is written over its own segment) | 22 x (leaf blocks 2 176 B, item 64 B, index entries 192 B) | 22 mbarriers |
// TMA row images 6 KB] = 227 KB to the byte.
#define VR6_OFF_LUT 0u
// The longest possible unit: 32 leaf subtrees of 63 nodes + 64 voxel items of 8 symbols = 575 symbols each.  A unit is 2 or 4
// runs of consecutive leaf subtrees; a run of 8 is at most 4 600 symbols = 1 150 bytes, staged as at most 74 16-byte chunks
// (alignment in front, one word of look-ahead behind: kd_build_worklist6): 296 chunks per unit, and real data gets there (a
// Richtmyer-Meshkov brick has units of 296).  Every unit is staged: a parser that reads its stream from global memory waits
// for HBM on its dependent chain, and next to the write-saturating constant groups that wait grows several-fold (measured:
// 15 % of the config-3 units took half of all long-scoreboard stalls).
#define VR6_STAGE_BYTES 4864u                                      // 304 chunks; a multiple of 128 (tensor stores read aligned boxes)
#define VR6_BOX_BYTES 2048u                                        // 32 x 8 x 8 voxels of a unit, x fastest
#define VR6_OFF_STAGE VR6_LUT_BYTES
#define VR6_OFF_MISC (VR6_OFF_STAGE + VR6_PARSE_WARPS * VR6_STAGE_BYTES)
#define VR6_WARP_STRIPS 0u                                         // 32 leaf blocks, pitch 68: 2 176 bytes
#define VR6_WARP_ITEM (VR6_WARP_STRIPS + 32u * VR6_STRIP_PITCH)    // 64 bytes
#define VR6_WARP_LEAFOFF (VR6_WARP_ITEM + 64u)                     // 128 bytes
#define VR6_WARP_LEAFVAL (VR6_WARP_LEAFOFF + 128u)                 // 64 bytes: one 16-byte chunk per run
#define VR6_MISC_BYTES (VR6_WARP_LEAFVAL + 64u)                    // 2 432, a multiple of 16
#define VR6_OFF_MBAR (VR6_OFF_MISC + VR6_PARSE_WARPS * VR6_MISC_BYTES)  // 8 bytes per warp: completion of the staging copies
#define VR6_QUEUE_FILL 32u  // the constant groups' counter lives on its own 128-byte line
#define VR6_BATCH_MAX 4u    // units per queue ticket: one atomic per unit on one address serialises in L2 (38 % of the stalls on
                            // 262 144 units); short queues take single units, whose costs differ by 10x, for balance
#define VR6_IMG_TOTAL 6144u  // row images of the constant groups: 6144 / (256 * rows) slots of (32 << 3) x rows bytes
#define VR6_OFF_IMG ((VR6_OFF_MBAR + VR6_PARSE_WARPS * 8u + 127u) & ~127u)  // tensor stores read 128-byte aligned images
static_assert(VR6_OFF_IMG + VR6_IMG_TOTAL <= 232448u, "shared memory of one CTA");

struct Decode6Params {
  const vr_parse_item* parse;
  const vr_fill_job* fills;
  uint32_t nParse, nFill;
  uint32_t* queue;  // [0] next parse unit, [VR6_QUEUE_FILL] next constant group
  uint8_t* dst;
  int64_t rowPitch, slicePitch;
  const uint32_t* lut;
  uint8_t* occ;
  vr6::P6Geom pg;
  int32_t tx, ty, tz;
  int32_t only;
  int32_t tmaRows;   // rows per tensor-store box (8, 16 or 32)
  int32_t tmaOut;    // 1: the units' voxels leave as one tensor store per unit, 0: as 16-byte stores of the lanes
  uint32_t batch;    // units per parse ticket
  int32_t fillMode;  // 0 = tensor stores (TMA), 1 = ordinary streaming stores by the fill warp, 2 = tensor stores with an evict-first L2 hint
  VrFillMaps maps;
  alignas(64) unsigned char unitMap[128];  // tensor map of the destination with a unit's box
};

__device__ __forceinline__ uint32_t lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds8(uint32_t a) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}

// Timing probe (only == 4): every unit reads the stream and index entries of one of the first 1 024 units, so its loads hit L2;
// what it decodes is garbage, where it stores is real.
__device__ __forceinline__ uint32_t src_of(const Decode6Params& p, uint32_t idx) { return p.only == 4 ? (idx & 1023u) : idx; }

// Stages unit `idx` (its 64-byte item, 32 index entries and stream segment) into the warp's shared-memory area with 16-byte
// asynchronous copies (LDGSTS), all lanes taking part.  Not the TMA unit: its queue is shared with the constant groups' tensor
// stores, and a unit's load stuck behind a few hundred KB of HBM-bound stores starves the parser (measured: the launch took
// longer than its two halves one after the other).  ia / ib = the first 32 bytes of the item, valid in lane 0.
// A queue ticket = n consecutive units; the counter counts tickets.  atom.inc on purpose: an atomicAdd() by one lane is
// rewritten (by ptxas, inline PTX included) into its warp-aggregated form, which reads the result back with a shuffle
// right behind the atomic -- the warp then waits out the whole round trip to L2 (several microseconds while the constant
// groups saturate HBM: 22 % of the parse warps' stall samples on the bench slab, profiles/r2_ticket_ab.md) although the
// ticket is not needed before the next unit.  A wrapping increment cannot be aggregated (with a bound of 2^32 - 1 ptxas turns it
// back into an add, hence 2^31 - 1: never reached).
__device__ __forceinline__ uint32_t take_ticket(uint32_t* q) {  // first unit = ticket * p.batch, multiplied where it is used
  uint32_t r;
  asm volatile("atom.global.inc.u32 %0, [%1], 0x7FFFFFFF;" : "=r"(r) : "l"(q) : "memory");
  return r;
}
__device__ __forceinline__ void cp16(uint32_t dstS, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dstS), "l"(src) : "memory");
}
// Completion of a unit's staging copies is tracked by the warp's mbarrier (one noinc arrival per lane once its copies have
// landed), not by cp.async groups: groups of asynchronous copies and bulk (TMA) groups count on ONE scoreboard in this part
// (cp.async.wait_group N and cp.async.bulk.wait_group[.read] N assemble to the same DEPBAR.LE SB0, N), so a wait for the
// staging would also wait for the previous unit's tensor store and the other way round.
// ia / ib = the first two 16-byte quarters of the unit's item (valid in lane 0), ir = its run descriptors (lane r holds run r).
__device__ __forceinline__ void stage_unit(const Decode6Params& p, uint32_t idx, const uint4& ia, const uint4& ib, const uint2& ir, uint32_t wbase,
                                           uint32_t segS, uint32_t barS, uint32_t lane) {
  const uint64_t ar = (uint64_t)__shfl_sync(0xFFFFFFFFu, ia.x, 0) | ((uint64_t)__shfl_sync(0xFFFFFFFFu, ia.y, 0) << 32);
  const uint8_t* arena = reinterpret_cast<const uint8_t*>(ar);
  const uint32_t leafOffOff = __shfl_sync(0xFFFFFFFFu, ib.x, 0), leafValOff = __shfl_sync(0xFFFFFFFFu, ib.y, 0);
  const uint32_t runLog = p.pg.runLog, nRuns = 32u >> runLog, cprLog = runLog - 2u;  // (1 << cprLog) 16-byte chunks of leafOff entries per run
  // lanes 0-3: the item's first 64 bytes; 4-11: the 32 leafOff entries (lane order = run after run); 12..: one 16-byte chunk
  // of leafVal entries per run
  const uint32_t myRun = lane < 12u ? ((lane - 4u) & 7u) >> cprLog : (lane - 12u) & 3u;
  const uint32_t leafRel = VR_RUN_LEAF(__shfl_sync(0xFFFFFFFFu, ir.y, myRun));
  // (timing probe only == 4: the stream-side half of the item comes from the same few units over and over, see src_of)
  if (lane < 4u) cp16(wbase + VR6_WARP_ITEM + 16u * lane, reinterpret_cast<const uint8_t*>(p.parse + (lane == 0u ? src_of(p, idx) : idx)) + 16u * lane);
  else if (lane < 12u) cp16(wbase + VR6_WARP_LEAFOFF + 16u * (lane - 4u), arena + leafOffOff + 4u * leafRel + 16u * ((lane - 4u) & ((1u << cprLog) - 1u)));
  else if (lane < 12u + nRuns) cp16(wbase + VR6_WARP_LEAFVAL + 16u * (lane - 12u), arena + ((leafValOff + leafRel) & ~15u));
  // the segments: the lanes of a run (8 or 16) copy that run's chunks, 128 or 256 contiguous bytes per step; the runs sit one
  // behind the other in the staging region
  {
    const uint32_t cr = lane >> runLog, j = lane & ((1u << runLog) - 1u);
    const uint32_t so = __shfl_sync(0xFFFFFFFFu, ir.x, cr), pk = __shfl_sync(0xFFFFFFFFu, ir.y, cr);
    const uint32_t sc = VR_RUN_CHUNKS(pk), sm = segS + 16u * VR_RUN_SMEM(pk);
    const uint8_t* src = arena + so;
#pragma unroll 1
    for (uint32_t c = j; c < sc; c += 1u << runLog) cp16(sm + 16u * c, src + 16u * c);
  }
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(barS) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t barS, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "W6_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@!p bra W6_%=;\n"
      "}\n" ::"r"(barS),
      "r"(parity)
      : "memory");
}
// Pulls everything unit `idx` will stage into L2 -- its stream segment (one 128-byte line per lane), its index entries and its
// item -- two units ahead of the staging copies: next to the constant groups HBM is write-saturated and a read that misses L2
// takes several microseconds, longer than a unit takes to parse.  ia / ib valid in lane 0.
__device__ __forceinline__ void prefetch_unit(const Decode6Params& p, uint32_t idx, const uint4& ia, const uint4& ib, const uint2& ir, uint32_t lane) {
  const uint64_t ar = (uint64_t)__shfl_sync(0xFFFFFFFFu, ia.x, 0) | ((uint64_t)__shfl_sync(0xFFFFFFFFu, ia.y, 0) << 32);
  const uint32_t leafOffOff = __shfl_sync(0xFFFFFFFFu, ib.x, 0), leafValOff = __shfl_sync(0xFFFFFFFFu, ib.y, 0);
  // eight lanes per run: one 128-byte line of its segment each, and the lines of its index entries
  const uint32_t r = lane >> 3, j = lane & 7u;
  const uint32_t so = __shfl_sync(0xFFFFFFFFu, ir.x, r), pk = __shfl_sync(0xFFFFFFFFu, ir.y, r);
  if (r < (32u >> p.pg.runLog)) {
    const uint32_t sc = VR_RUN_CHUNKS(pk);
#pragma unroll 1
    for (uint32_t l = j; l * 8u < sc + 7u && sc; l += 8u) asm volatile("prefetch.global.L2 [%0];" ::"l"(ar + so + 128ull * l));
    if (j == 6u) asm volatile("prefetch.global.L2 [%0];" ::"l"(ar + leafOffOff + 4u * VR_RUN_LEAF(pk)));
    if (j == 7u) asm volatile("prefetch.global.L2 [%0];" ::"l"(ar + leafValOff + VR_RUN_LEAF(pk)));
  }
  if (lane == 31u) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const uint8_t*>(p.parse + idx)));
  if (lane == 30u) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const uint8_t*>(p.parse + idx) + 95));
}

// One parse warp: its own queue tickets, staging area and leaf blocks -- no barrier with any other warp, so a warp whose unit
// is empty or heavily pruned moves on while its neighbours are still parsing (with one barrier per tile and group of 8 warps,
// the warps of a group waited 2.8 issue slots per instruction issued for the slowest of them: profiles/r2_decode.md).
// Software pipeline over units n: ticket n+3 (atomic) and item n+2 (two 16-byte loads by lane 0) are requested and unit
// n+1's segment is pulled into L2 while unit n is parsed; unit n+1 is staged while unit n's voxels are stored.
template <int ROT>
__device__ __forceinline__ void parse_warp(const Decode6Params& p, const uint32_t sb, const uint32_t warp, const uint32_t lane) {
  const uint32_t wbase = sb + VR6_OFF_MISC + warp * VR6_MISC_BYTES;
  const uint32_t strips = wbase + VR6_WARP_STRIPS, stageS = sb + VR6_OFF_STAGE + warp * VR6_STAGE_BYTES, barS = sb + VR6_OFF_MBAR + 8u * warp;
  const uint32_t lutS = sb + VR6_OFF_LUT;
  const uint32_t xs0 = p.pg.xs[0], xs1 = p.pg.xs[1];
  const uint32_t myStrip = strips + lane * VR6_STRIP_PITCH;
  // phase-2 constants of this lane for unit 0 of a tile (the unit index adds a per-unit offset): four 16-byte runs
  // (tensor-store mode: gRun = the run's place in the unit's dense box instead of its offset in the slab)
  uint32_t sRun[4], gRun[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint32_t s;
    int64_t go;
    vr6::p6_run_addr(p.pg, 0u, lane + 32u * (uint32_t)k, p.rowPitch, p.slicePitch, &s, &go);
    sRun[k] = strips + s;
    gRun[k] = p.tmaOut ? vr6::p6_box_off(p.pg, lane + 32u * (uint32_t)k) : (uint32_t)go;
  }
  uint32_t unitOff[3];  // destination offset of unit bits 0..2
#pragma unroll
  for (int j = 0; j < 3; ++j)
    unitOff[j] = (uint32_t)((p.pg.wAxis[j] == 0 ? (int64_t)1 : (p.pg.wAxis[j] == 1 ? p.rowPitch : p.slicePitch)) << p.pg.wCb[j]);

  // idxA = unit being parsed, idxB = next (item half in b0 / b1), idxC = the one after (item half requested below).  Tickets are
  // batches of p.batch consecutive units: (curBase, curPos) = where idxC sits in its batch, nextBase = the TICKET of the batch behind it,
  // requested a whole batch ahead (lane 0 only).
  const uint4 z4 = make_uint4(0, 0, 0, 0);
  uint32_t idxA = 0, idxB = 0, idxC = 0, curBase = 0, curPos = 0, nextBase = 0;
  uint4 a0 = z4, a1 = z4, b0 = z4, b1 = z4;
  // The first tickets of a warp are static (ticket k * W + gw, W = parse warps of the grid; 3 of them when a ticket is one
  // unit, else 2), the counter hands out the ones behind them: the first units start without a round trip to L2.
  const uint32_t W = gridDim.x * VR6_PARSE_WARPS, gw = blockIdx.x * VR6_PARSE_WARPS + warp;
  const uint32_t tOff = (p.batch == 1u ? 3u : 2u) * W;  // nextBase holds ticket - tOff (what the counter returns)
  if (lane == 0) {
    curBase = gw * p.batch;
    nextBase = W + gw - tOff;
    // the first three units: positions 0, 1, 2 of the ticket sequence (tickets may hold fewer than three units)
    idxA = curBase; curPos = 0u;
    if (curPos + 1u < p.batch) { ++curPos; idxB = curBase + curPos; } else { curBase = (nextBase + tOff) * p.batch; curPos = 0u; idxB = curBase; nextBase = 2u * W + gw - tOff; }
    if (curPos + 1u < p.batch) { ++curPos; idxC = curBase + curPos; } else { curBase = (nextBase + tOff) * p.batch; curPos = 0u; idxC = curBase; nextBase = take_ticket(p.queue); }
    if (idxA < p.nParse) { a0 = __ldg(reinterpret_cast<const uint4*>(p.parse + src_of(p, idxA))); a1 = __ldg(reinterpret_cast<const uint4*>(p.parse + src_of(p, idxA)) + 1); }
    if (idxB < p.nParse) { b0 = __ldg(reinterpret_cast<const uint4*>(p.parse + src_of(p, idxB))); b1 = __ldg(reinterpret_cast<const uint4*>(p.parse + src_of(p, idxB)) + 1); }
  }
  if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 32;" ::"r"(barS) : "memory");  // visible to the warp after the barrier below
  // the chain table, while those requests are in flight (the parse warps only: the fill warp does not wait for it)
  for (uint32_t i = threadIdx.x; i < VR6_LUT_BYTES / 16u; i += VR6_PARSE_WARPS * 32u) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p.lut) + i);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sb + 16u * i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
  asm volatile("bar.sync 1, %0;" ::"n"(VR6_PARSE_WARPS * 32) : "memory");
  idxA = __shfl_sync(0xFFFFFFFFu, idxA, 0);
  idxB = __shfl_sync(0xFFFFFFFFu, idxB, 0);
  idxC = __shfl_sync(0xFFFFFFFFu, idxC, 0);
  // run descriptors of the units (lane r: run r), loaded with the item halves: curR = the unit being parsed, bR / cR = the next two
  const uint2 z2 = make_uint2(0, 0);
  uint2 curR = z2, bR = z2;
  if (lane < 4u) {
    if (idxA < p.nParse) curR = __ldg(reinterpret_cast<const uint2*>(&p.parse[src_of(p, idxA)].run[lane]));
    if (idxB < p.nParse) bR = __ldg(reinterpret_cast<const uint2*>(&p.parse[src_of(p, idxB)].run[lane]));
  }
  // The staging area is two regions, R0 = [0, 2 048) and R1 = [2 048, 4 864).  A unit's segments sit in one of them and its
  // voxel box (2 048 B) is later written over it, so the OTHER region is free while the unit is parsed: the next unit is
  // staged there half-way through the parsing (its copies then have the rest of the unit to land), not after it.  A
  // segment longer than its region ("big": starts at 0 and reaches into R1; noise-like data) falls back to staging
  // after the parsing (if the next segment fits R1) or after the tensor store has read the box.
  uint32_t curOff = 0, parity = 0;
  bool curBig = false;
  if (idxA < p.nParse) {
    const uint32_t bytesA = 16u * __shfl_sync(0xFFFFFFFFu, a0.z, 0);
    curOff = bytesA <= VR6_STAGE_BYTES - VR6_BOX_BYTES ? VR6_BOX_BYTES : 0u;
    curBig = bytesA > VR6_STAGE_BYTES - VR6_BOX_BYTES;
    stage_unit(p, idxA, a0, a1, curR, wbase, stageS + curOff, barS, lane);
  }
  if (idxB < p.nParse) prefetch_unit(p, idxB, b0, b1, bR, lane);
  while (idxA < p.nParse) {
    // requests for later units first, so that they are in flight while this one is parsed
    uint4 c0 = z4, c1 = z4;
    uint2 cR = z2;
    uint32_t idxD = 0;
    if (lane < 4u && idxC < p.nParse) cR = __ldg(reinterpret_cast<const uint2*>(&p.parse[src_of(p, idxC)].run[lane]));
    if (lane == 0) {
      if (idxC < p.nParse) { c0 = __ldg(reinterpret_cast<const uint4*>(p.parse + src_of(p, idxC))); c1 = __ldg(reinterpret_cast<const uint4*>(p.parse + src_of(p, idxC)) + 1); }
      if (curPos + 1u < p.batch) { ++curPos; idxD = curBase + curPos; }
      else { curBase = (nextBase + tOff) * p.batch; curPos = 0u; idxD = curBase; nextBase = take_ticket(p.queue); }
    }
    mbar_wait(barS, parity);  // this unit's segment, index entries and item are in shared memory
    parity ^= 1u;
    const uint4 i2 = lds128(wbase + VR6_WARP_ITEM + 32u), i3 = lds128(wbase + VR6_WARP_ITEM + 48u);
    const uint64_t treeDist = (uint64_t)i2.x | ((uint64_t)i2.y << 32);
    const int64_t dstOff = (int64_t)((uint64_t)i2.z | ((uint64_t)i2.w << 32));
    const int32_t occCell = (int32_t)i3.w;
    const uint32_t unit = lds32(wbase + VR6_WARP_ITEM + 28u) & 7u;  // tileUnit
    // this lane's leaf box: run lane >> runLog of the unit, entry lane of the staged index entries (lane order = run after run)
    const uint32_t myRun = lane >> p.pg.runLog;
    const uint32_t runSeg = __shfl_sync(0xFFFFFFFFu, curR.x, myRun), runPk = __shfl_sync(0xFFFFFFFFu, curR.y, myRun);
    const uint32_t off = lds32(wbase + VR6_WARP_LEAFOFF + 4u * lane);
    const uint32_t val = lds8(wbase + VR6_WARP_LEAFVAL + 16u * myRun + (VR_RUN_LEAF(runPk) & 15u) + (lane & ((1u << p.pg.runLog) - 1u)));
    // where the next unit's segment can go
    const bool haveB = idxB < p.nParse;
    const uint32_t bytesB = 16u * __shfl_sync(0xFFFFFFFFu, b0.z, 0);
    const uint32_t otherOff = curOff ? 0u : VR6_BOX_BYTES, otherCap = curOff ? VR6_BOX_BYTES : VR6_STAGE_BYTES - VR6_BOX_BYTES;
    const bool fitsOther = haveB && bytesB <= otherCap;
    const bool early = fitsOther && !curBig;
    // ---- phase 1: one leaf subtree per lane, first half ----
    vr6::Dist d;
    d.d0 = vr6::dist2((uint32_t)treeDist & 0xFFu); d.d1 = vr6::dist2((uint32_t)(treeDist >> 8) & 0xFFu);
    d.d2 = vr6::dist2((uint32_t)(treeDist >> 16) & 0xFFu); d.d3 = vr6::dist2((uint32_t)(treeDist >> 24) & 0xFFu);
    d.d4 = vr6::dist2((uint32_t)(treeDist >> 32) & 0xFFu); d.d5 = vr6::dist2((uint32_t)(treeDist >> 40) & 0xFFu);
    vr6::WinShared w;
    uint32_t wo = 2u;  // bits consumed since the window's last re-alignment: the root's symbol
    const uint32_t word6 = vr6::mkword(val, d.d5);
    if (off == VR_PRUNED) {
      vr6::fill_leaf(myStrip, val);
    } else {
      w.init(stageS + curOff + 16u * VR_RUN_SMEM(runPk) - runSeg, off);
      vr6::parse_leaf_half<ROT, 0>(w, wo, word6, d, lutS, myStrip);
    }
    __syncwarp();  // every lane has read its index entry and the item
    if (early) {   // the other region held the previous unit's box: its tensor store (issued half a unit ago) must have read it
      if (p.tmaOut) {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
      }
      stage_unit(p, idxB, b0, b1, bR, wbase, stageS + otherOff, barS, lane);
    }
    // ---- second half ----
    if (off != VR_PRUNED) vr6::parse_leaf_half<ROT, 4>(w, wo, word6, d, lutS, myStrip);
    __syncwarp();  // every lane is done with the staged segment; the leaf blocks are complete
    bool staged = early;
    if (!staged && fitsOther) {  // this unit's segment was big; now that it is consumed the other region is free
      if (p.tmaOut) {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
      }
      stage_unit(p, idxB, b0, b1, bR, wbase, stageS + otherOff, barS, lane);
      staged = true;
    }
    // ---- phase 2: the warp's 32 leaf blocks -> 128 runs of 16 voxels along x, four per lane ----
    // (tensor-store mode: into the unit's box, which takes the place of its segment; no tensor store is reading that
    // region: every staging above waited for the stores before it, and the one store since then read the other region)
    uint8_t* dstUnit = p.dst + dstOff + ((unit & 1u) ? unitOff[0] : 0u) + ((unit & 2u) ? unitOff[1] : 0u) + ((unit & 4u) ? unitOff[2] : 0u);
    const uint32_t boxS = stageS + curOff;
    // byte maxima through 16-bit lanes: a 16-bit maximum is exact in its HIGH byte, so bytes 1 / 3 come from the words as they
    // are and bytes 0 / 2 from the words shifted left by 8 (the shift runs on the multiply-add pipe)
    uint32_t m13 = 0, m02 = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t a = sRun[k];
      const uint32_t w0 = lds32(a), w1 = lds32(a + xs0), w2 = lds32(a + xs1), w3 = lds32(a + xs0 + xs1);
      if (p.tmaOut) asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(boxS + gRun[k]), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
      else if (p.only != 3) *reinterpret_cast<uint4*>(dstUnit + gRun[k]) = make_uint4(w0, w1, w2, w3);  // (3: timing probe without the voxel stores)
      m13 = __vmaxu2(__vmaxu2(m13, w0), __vmaxu2(w1, __vmaxu2(w2, w3)));
      m02 = __vmaxu2(__vmaxu2(m02, w0 * 256u), __vmaxu2(w1 * 256u, __vmaxu2(w2 * 256u, w3 * 256u)));
    }
    if (p.tmaOut) {
      // The unit leaves as ONE tensor store (32 x 8 x 8 box: 64 rows of whole sectors).  Ordinary stores of the lanes wait behind the constant groups'
      // tensor stores whenever HBM is write-saturated, and since the load/store unit works in order, every shared-memory
      // access of the SM then waits with them: the launch took fill + parse (profiles/r2_decode.md).  In the TMA queue the
      // units simply take their turn.
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0 && p.only != 3) {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(&p.unitMap[0]), "r"((int)i3.x),
                     "r"((int)i3.y), "r"((int)i3.z), "r"(boxS)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (haveB && !staged) {  // the next segment needs more than the other region: from offset 0, once this unit's box has been read
      if (p.tmaOut) {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
      }
      stage_unit(p, idxB, b0, b1, bR, wbase, stageS, barS, lane);
    }
    curOff = staged ? otherOff : 0u;
    curBig = !staged && bytesB > VR6_BOX_BYTES;
    if (p.occ && occCell >= 0) {  // this unit's maximum -> its slot of the tile's occupancy cell
      uint32_t m = max(max((m13 >> 8) & 0xFFu, m13 >> 24), max((m02 >> 8) & 0xFFu, m02 >> 24));
      m = __reduce_max_sync(0xFFFFFFFFu, m);
      if (lane == 0) p.occ[(int64_t)occCell * 8 + unit] = (uint8_t)m;
    }
    __syncwarp();  // the leaf blocks are free for the next unit
    if (idxC < p.nParse) prefetch_unit(p, idxC, c0, c1, cR, lane);  // its item half was requested at the top of this iteration
    idxA = idxB; idxB = idxC; idxC = __shfl_sync(0xFFFFFFFFu, idxD, 0);
    b0 = c0; b1 = c1;
    curR = bR; bR = cR;
  }
  if (p.tmaOut && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // before the CTA's shared memory goes away
}

// The warp that feeds the TMA unit.  A constant group = 2^L x-neighbouring constant tiles (vr_fill_job): a row image of
// (tx << L) x VR6_TMA_ROWS bytes in shared memory (byte j * tx.. = value of tile j), stored tz * ty / rows times.  Images
// are cached by (L, values): a slab of mostly empty / mostly solid bricks rewrites them a handful of times per launch.
__device__ __forceinline__ void fill_warp(const Decode6Params& p, const uint32_t img0, const uint32_t lane) {
  const uint32_t ltx = 31u - (uint32_t)__clz(p.tx);
  const uint32_t rows = (uint32_t)p.tmaRows, imgBytes = 256u * rows, nSlots = VR6_IMG_TOTAL / imgBytes;
  const uint32_t rowBlocks = (uint32_t)p.ty / rows, per = (uint32_t)p.tz * rowBlocks;
  uint64_t keyVals[VR6_IMG_SLOTS];
  uint32_t keyL[VR6_IMG_SLOTS];
#pragma unroll
  for (int s = 0; s < VR6_IMG_SLOTS; ++s) { keyVals[s] = 0; keyL[s] = 0xFFFFFFFFu; }
  uint32_t victim = 0;
  uint64_t policy = 0;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  // tickets of 8 groups; the first one is static (the CTA's index), the next one is requested while this one is worked on
  uint32_t tk = blockIdx.x - gridDim.x;  // ticket - gridDim.x, as the counter returns it
  for (;;) {
    const uint32_t j0 = (__shfl_sync(0xFFFFFFFFu, tk, 0) + gridDim.x) * 8u;
    if (j0 >= p.nFill) break;
    if (lane == 0) tk = take_ticket(p.queue + VR6_QUEUE_FILL);
    const uint32_t n = min(8u, p.nFill - j0);
    uint4 a = make_uint4(0, 0, 0, 0), b = a;
    if (lane < n) {
      a = __ldg(reinterpret_cast<const uint4*>(p.fills + j0 + lane));
      b = __ldg(reinterpret_cast<const uint4*>(p.fills + j0 + lane) + 1);
    }
    for (uint32_t k = 0; k < n; ++k) {
      const int x = (int)__shfl_sync(0xFFFFFFFFu, a.x, k), y = (int)__shfl_sync(0xFFFFFFFFu, a.y, k), z = (int)__shfl_sync(0xFFFFFFFFu, a.z, k);
      const uint32_t L = __shfl_sync(0xFFFFFFFFu, a.w, k);
      const uint64_t vals = (uint64_t)__shfl_sync(0xFFFFFFFFu, b.x, k) | ((uint64_t)__shfl_sync(0xFFFFFFFFu, b.y, k) << 32);
      const int64_t cell = (int64_t)((uint64_t)__shfl_sync(0xFFFFFFFFu, b.z, k) | ((uint64_t)__shfl_sync(0xFFFFFFFFu, b.w, k) << 32));
      if (p.occ && cell >= 0 && lane < (1u << L)) {
        const uint32_t w = (uint32_t)((vals >> (8 * lane)) & 0xFF) * 0x01010101u;
        *reinterpret_cast<uint2*>(p.occ + (cell + lane) * 8) = make_uint2(w, w);
      }
      if (p.fillMode == 1) {  // A/B: the same group as 16-byte streaming stores of this one warp
        vr_kd_geom gg;
        gg.tx = p.tx; gg.ty = p.ty; gg.tz = p.tz;
        vr_group_fill(gg, (int)L, vals, p.dst + (int64_t)z * p.slicePitch + (int64_t)y * p.rowPitch + x, p.rowPitch, p.slicePitch, lane, 32u);
        continue;
      }
      int slotIdx = -1;
#pragma unroll
      for (int s = 0; s < VR6_IMG_SLOTS; ++s)
        if ((uint32_t)s < nSlots && keyL[s] == L && keyVals[s] == vals) slotIdx = s;
      if (slotIdx < 0) {  // warp-uniform: rewrite the least recently written image once the TMA unit has read it
        slotIdx = (int)victim;
        victim = (victim + 1u) % nSlots;
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
#pragma unroll
        for (int s = 0; s < VR6_IMG_SLOTS; ++s)
          if (s == slotIdx) { keyL[s] = L; keyVals[s] = vals; }
        const uint32_t lW = ltx + L;                             // log2 row bytes of the group
        const uint32_t pieces = (rows << lW) >> 4;               // 16-byte pieces of the image
        const uint32_t img = img0 + (uint32_t)slotIdx * imgBytes;
        for (uint32_t q = lane; q < pieces; q += 32u) {
          const uint32_t xb = (q << 4) & ((1u << lW) - 1u);
          const uint32_t w = (uint32_t)((vals >> (8 * (xb >> ltx))) & 0xFF) * 0x01010101u;
          asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(img + 16u * q), "r"(w) : "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
      }
      const uint32_t img = img0 + (uint32_t)slotIdx * imgBytes;
      const void* tm = &p.maps.m[L][0];
      for (uint32_t s = lane; s < per; s += 32u) {
        const int zz = z + (int)(s / rowBlocks), yy = y + (int)((s % rowBlocks) * rows);
        if (p.fillMode == 2)
          asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%1, %2, %3}], [%4], %5;" ::"l"(tm), "r"(x), "r"(yy),
                       "r"(zz), "r"(img), "l"(policy)
                       : "memory");
        else
          asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(tm), "r"(x), "r"(yy), "r"(zz), "r"(img)
                       : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the stores must have left before the CTA's shared memory goes away
}

template <int ROT>
__global__ void __launch_bounds__(VR6_THREADS, 1) kd_decode6_kernel(const __grid_constant__ Decode6Params p) {
  extern __shared__ __align__(128) uint8_t smem6[];
  const uint32_t sb = (uint32_t)__cvta_generic_to_shared(smem6);
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
  if (warp >= VR6_PARSE_WARPS) {
    if (p.only != 1 && p.nFill) fill_warp(p, sb + VR6_OFF_IMG, lane);
  } else if (p.only != 2 && p.nParse) {
    parse_warp<ROT>(p, sb, warp, lane);
  }
}

// The constant groups as a launch of their own (split mode): one warp per CTA, 8 KB of shared memory, next to the parse
// CTAs of the kernel above (launched without its fill warp and without the image area, so that both fit one SM).  A parse
// CTA holds 56 K registers until its last warp is done; as one kernel, those registers stay taken while the fill warp drains
// its HBM-bound queue, and a frame's march CTAs (a second stream) find no room.  As two kernels the parse CTAs leave when the
// parsing is done and the march runs beside the fills.
__global__ void __launch_bounds__(32, 1) kd_fill6_kernel(const __grid_constant__ Decode6Params p) {
  extern __shared__ __align__(128) uint8_t smem6[];
  fill_warp(p, (uint32_t)__cvta_generic_to_shared(smem6), threadIdx.x);
}

uint32_t kd_decode6_stage_chunks(void) { return VR6_STAGE_BYTES / 16u; }

typedef CUresult (*VrEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static VrEncodeFn tensor_map_encoder() {
  static VrEncodeFn enc = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      f = nullptr;
    }
    return (VrEncodeFn)f;
  }();
  return enc;
}

// 3-D uint8 tensor map over the slab with box bx x by x bz (dense in shared memory, x fastest)
static bool box_map_create(uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, int bx, int by, int bz, unsigned char out[128]) {
  static_assert(sizeof(CUtensorMap) == 128, "CUtensorMap size");
  VrEncodeFn enc = tensor_map_encoder();
  if (!enc || bx <= 0 || bx > 256 || by <= 0 || by > 256 || bz <= 0 || bz > 256) return false;
  if (SX % 16 || ((uintptr_t)vol) % 16 || SX >= (int64_t(1) << 32) || SY >= (int64_t(1) << 32) || SZ >= (int64_t(1) << 32)) return false;
  cuuint64_t dims[3] = {(cuuint64_t)SX, (cuuint64_t)SY, (cuuint64_t)SZ};
  cuuint64_t strides[2] = {(cuuint64_t)SX, (cuuint64_t)SX * (cuuint64_t)SY};
  cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, (cuuint32_t)bz};
  cuuint32_t es[3] = {1, 1, 1};
  return enc(reinterpret_cast<CUtensorMap*>(out), CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, vol, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool kd_fill_maps_create(uint8_t* vol, int64_t SX, int64_t SY, int64_t SZ, int tx, int ty, int rows, VrFillMaps* out) {
  VrEncodeFn enc = tensor_map_encoder();
  if (!enc || tx < 16 || (tx << 3) > 256 || rows <= 0 || ty % rows) return false;
  if (SX % 16 || ((uintptr_t)vol) % 16 || SX >= (int64_t(1) << 32) || SY >= (int64_t(1) << 32) || SZ >= (int64_t(1) << 32)) return false;
  for (int L = 0; L <= 3; ++L) {
    cuuint64_t dims[3] = {(cuuint64_t)SX, (cuuint64_t)SY, (cuuint64_t)SZ};
    cuuint64_t strides[2] = {(cuuint64_t)SX, (cuuint64_t)SX * (cuuint64_t)SY};
    cuuint32_t box[3] = {(cuuint32_t)(tx << L), (cuuint32_t)rows, 1};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(reinterpret_cast<CUtensorMap*>(&out->m[L][0]), CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, vol, dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
  }
  return true;
}

bool kd_decode6_plan(const vr_kd_geom& g, uint8_t* dst, int64_t SX, int64_t SY, int64_t SZ, Decode6Plan* out, int tmaRows) {
  if (tmaRows != 8 && tmaRows != 16) return false;  // (32 rows = 8 KB per image: no slot)
  out->tmaRows = tmaRows;
  if (!vr6::p6_make_geom(g, &out->pg)) return false;
  if (g.tx != 32 || g.ty % tmaRows) return false;                    // image slots are sized for 32-wide tiles
  if ((int64_t)g.tz * SX * SY + (int64_t)g.ty * SX >= (int64_t(1) << 31)) return false;  // 32-bit run offsets inside a tile
  if (!kd_fill_maps_create(dst, SX, SY, SZ, g.tx, g.ty, tmaRows, &out->maps)) return false;
  if (((uint32_t)VR6_UNIT_X << out->pg.boxLy << out->pg.boxLz) != VR6_BOX_BYTES) return false;
  if (!box_map_create(dst, SX, SY, SZ, VR6_UNIT_X, 1 << out->pg.boxLy, 1 << out->pg.boxLz, out->unitMap)) return false;
  out->smemBytes = (size_t)VR6_OFF_IMG + (size_t)VR6_IMG_TOTAL;
  return true;
}

template <int ROT>
static cudaError_t launch6(const Decode6Params& p, size_t smem, int ctas, int threads, cudaStream_t stream) {
  cudaError_t e = cudaFuncSetAttribute(kd_decode6_kernel<ROT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  kd_decode6_kernel<ROT><<<(unsigned)ctas, (unsigned)threads, smem, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t kd_decode6_launch(const vr_kd_geom& g, const Decode6Plan& plan, const vr_kd_job*, const vr_parse_item* parseDev, int nParse,
                              const vr_fill_job* fillsDev, int nFill, uint32_t* queueDev, const uint32_t* lutDev, uint8_t* dst,
                              int64_t rowPitch, int64_t slicePitch, uint8_t* occ, int only, cudaStream_t stream, int fillMode, int tmaOut,
                              const Decode6Side* side) {
  if (nParse <= 0 && nFill <= 0) return cudaSuccess;
  static int numSMs = 0;
  if (!numSMs) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&numSMs, cudaDevAttrMultiProcessorCount, dev); }
  Decode6Params p;
  p.parse = parseDev; p.fills = fillsDev;
  p.nParse = (uint32_t)(nParse > 0 ? nParse : 0); p.nFill = (uint32_t)(nFill > 0 ? nFill : 0);
  p.queue = queueDev;
  p.dst = dst; p.rowPitch = rowPitch; p.slicePitch = slicePitch;
  p.lut = lutDev; p.occ = occ;
  p.pg = plan.pg;
  p.tx = g.tx; p.ty = g.ty; p.tz = g.tz;
  p.only = only;
  p.fillMode = fillMode & 0xFF;
  p.tmaRows = plan.tmaRows;
  p.tmaOut = tmaOut ? 1 : 0;
  memcpy(p.unitMap, plan.unitMap, 128);
  p.maps = plan.maps;
  cudaError_t e = cudaMemsetAsync(queueDev, 0, VR6_QUEUE_WORDS * sizeof(uint32_t), stream);
  if (e != cudaSuccess) return e;
  // enough CTAs for the work: a parse warp takes units one at a time, a fill warp groups eight at a time
  const int wantP = (int)((p.nParse + VR6_PARSE_WARPS - 1u) / VR6_PARSE_WARPS), wantF = (int)((p.nFill + 7u) / 8u);
  const int want = wantP > wantF ? wantP : wantF;
  const int ctas = want < numSMs ? (want > 0 ? want : 1) : numSMs;
  const uint32_t perWarp = p.nParse / (uint32_t)(ctas * VR6_PARSE_WARPS);
  p.batch = perWarp >= 8u * VR6_BATCH_MAX ? VR6_BATCH_MAX : (perWarp >= 16u ? 2u : 1u);
  // rot == 2 (last split along z) puts tile x bit 3 outside a warp's 32 leaf boxes: p6_make_geom never accepts it
  if (plan.pg.rot != 0 && plan.pg.rot != 1) return cudaErrorInvalidValue;
  if (side && side->stream && (only == 0 || only == 1 || only == 2) && p.nParse && p.nFill && p.fillMode != 1) {
    // split mode: constant groups on the side stream (fork / join by events, capturable), parse CTAs without the fill warp
    // (timing probes only == 1 / 2: one of the two kernels alone)
    if (only != 1) {
      e = cudaEventRecord(side->fork, stream);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(side->stream, side->fork, 0);
      if (e != cudaSuccess) return e;
      Decode6Params pf = p;
      pf.only = 2;
      const int fillCtas = wantF < numSMs ? (wantF > 0 ? wantF : 1) : numSMs;
      kd_fill6_kernel<<<(unsigned)fillCtas, 32, VR6_IMG_TOTAL, side->stream>>>(pf);
      e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaEventRecord(side->join, side->stream);
      if (e != cudaSuccess) return e;
    }
    if (only != 2) {
      p.only = 1;
      const int parseCtas = wantP < numSMs ? (wantP > 0 ? wantP : 1) : numSMs;
      e = plan.pg.rot == 0 ? launch6<0>(p, VR6_OFF_IMG, parseCtas, VR6_PARSE_WARPS * 32, stream) : launch6<1>(p, VR6_OFF_IMG, parseCtas, VR6_PARSE_WARPS * 32, stream);
      if (e != cudaSuccess) return e;
    }
    return only != 1 ? cudaStreamWaitEvent(stream, side->join, 0) : cudaSuccess;
  }
  if (plan.pg.rot == 0) return launch6<0>(p, plan.smemBytes, ctas, VR6_THREADS, stream);
  return launch6<1>(p, plan.smemBytes, ctas, VR6_THREADS, stream);
}

// =====================================================================================================================
// classic kernel: one CTA per work-list entry
// =====================================================================================================================
struct DecodeParams {
  vr_kd_geom g;
  const vr_kd_job* jobs;
  uint8_t* dst;
  int64_t rowPitch;    // slab X
  int64_t slicePitch;  // slab X*Y
  int vec16;
  const vr_kd_work* work;    // one entry per CTA (kd_worklist.h)
  const vr_fill_job* fills;  // constant groups attached to parse CTAs
  // occupancy (occupancy.h): when a decode tile is exactly one cell the kernel stores the tile's maximum
  // (8 writer slots per cell: one per warp, so no CTA-wide reduction is needed)
  uint8_t* occ;             // null = off
  const int32_t* tileCell;  // [2^Dt] cell of the tile relative to the cell of the brick's (0,0,0)
};

// occupancy cell of tile `tile` of a brick whose (0,0,0) is cell occBase: tiles are 32x32x16 = one cell
__device__ __forceinline__ int64_t tile_cell(const DecodeParams& p, const vr_kd_job& job, uint32_t tile) {
  return job.occBase + __ldg(p.tileCell + tile);
}

__device__ __forceinline__ void attached_fills(const DecodeParams& p, uint32_t first, uint32_t count) {
  for (uint32_t k = 0; k < count; ++k) {
    const vr_fill_job* fj = p.fills + first + k;
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(fj));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(fj) + 1);
    const uint64_t vals = (uint64_t)b.x | ((uint64_t)b.y << 32);
    const int64_t cell = (int64_t)((uint64_t)b.z | ((uint64_t)b.w << 32));
    if (p.occ && cell >= 0 && threadIdx.x < (1u << a.w)) {
      const uint32_t w = (uint32_t)((vals >> (8 * threadIdx.x)) & 0xFF) * 0x01010101u;
      *reinterpret_cast<uint2*>(p.occ + (cell + threadIdx.x) * 8) = make_uint2(w, w);
    }
    vr_group_fill(p.g, (int)a.w, vals, p.dst + (int64_t)(int)a.z * p.slicePitch + (int64_t)(int)a.y * p.rowPitch + (int)a.x, p.rowPitch,
                  p.slicePitch, threadIdx.x, blockDim.x);
  }
}

size_t kd_decode_smem_bytes(const vr_kd_geom& g) {
  const size_t lpt = size_t(1) << (g.tileBits - g.leafBits);
  return lpt * ((size_t(1) << g.leafBits) + VR_LEAF_PITCH_PAD);
}

__global__ void __launch_bounds__(256, 6) kd_decode_tiles_kernel(const __grid_constant__ DecodeParams p) {
  extern __shared__ __align__(128) uint8_t strips[];
  __shared__ int64_t sTileBase;
  const vr_kd_geom& g = p.g;
  const uint4 wk = __ldg(reinterpret_cast<const uint4*>(p.work) + blockIdx.x);  // {id, L, vals}
  const uint32_t brick = wk.x >> g.Dt;
  const uint32_t tile = wk.x & ((1u << g.Dt) - 1u);
  const vr_kd_job& job = p.jobs[brick];
  if (threadIdx.x == 0) {
    int ox, oy, oz;
    vr_tile_origin(g, tile, &ox, &oy, &oz);
    sTileBase = job.dstOffset + (int64_t)oz * p.slicePitch + (int64_t)oy * p.rowPitch + ox;
  }
  if (wk.y != VR_WORK_PARSE) {  // constant tiles: pure fill
    const uint64_t vals = (uint64_t)wk.z | ((uint64_t)wk.w << 32);
    const int L = (int)wk.y;
    if (p.occ && job.occBase >= 0 && threadIdx.x < (1u << L)) {  // one thread per member tile
      uint32_t t = tile;
      for (int j = 0; j < L; ++j) t |= ((threadIdx.x >> j) & 1u) << g.xrBit[j];
      const uint32_t w = (uint32_t)((vals >> (8 * threadIdx.x)) & 0xFF) * 0x01010101u;
      *reinterpret_cast<uint2*>(p.occ + tile_cell(p, job, t) * 8) = make_uint2(w, w);
    }
    __syncthreads();
    uint8_t* dst = p.dst + sTileBase;
    if (p.vec16) {
      vr_group_fill(g, L, vals, dst, p.rowPitch, p.slicePitch, threadIdx.x, blockDim.x);
    } else {
      const uint32_t items = vr_tile_items(g);
      for (uint32_t it = threadIdx.x; it < items; it += blockDim.x)
        vr_tile_fill_item(g, it, (int)(vals & 0xFF), dst, p.rowPitch, p.slicePitch, false);
    }
    return;
  }
  // constant groups that ride along with this parse CTA (wk.z = first, wk.w = count): stored before parsing
  attached_fills(p, wk.z, wk.w & 0x7FFFFFFFu);
  const uint32_t blk = __ldg(job.tileLeafBlock + tile);
  const uint32_t items = vr_tile_items(g);
  if (g.fastPath && p.vec16) {
    // fast path: 256 parsers, chains by table look-up, warp-local hand-off to the store phase (no CTA barrier)
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(strips);
    vr_tile_phase1_v5<true, uint32_t, VrWin32>(job, g, blk, threadIdx.x, sbase);
    __syncwarp();
    int64_t tb = job.dstOffset;  // tile origin, recomputed per thread (no CTA barrier on this path)
    for (int d = 0; d < g.Dt; ++d)
      if ((tile >> (g.Dt - 1 - d)) & 1) {
        const int64_t st = g.splitDim[d] == 0 ? (int64_t)1 : (g.splitDim[d] == 1 ? p.rowPitch : p.slicePitch);
        tb += st * g.childOff[d];
      }
    uint8_t* dstw = p.dst + tb;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t perUnit = 1u << g.fbUnit;
    uint32_t m4 = 0;
    for (uint32_t q = lane; q < perUnit; q += 32) m4 = __vmaxu4(m4, vr_unit_phase2_item(g, warp, q, strips, dstw, p.rowPitch, p.slicePitch));
    if (p.occ && job.occBase >= 0) {  // this warp's maximum -> its slot of the tile's occupancy cell
      uint32_t m = max(max(m4 & 0xFF, (m4 >> 8) & 0xFF), max((m4 >> 16) & 0xFF, m4 >> 24));
      m = __reduce_max_sync(0xFFFFFFFFu, m);
      if (lane == 0) p.occ[tile_cell(p, job, tile) * 8 + warp] = (uint8_t)m;
    }
    return;
  }
  const uint32_t lpt = 1u << (g.tileBits - g.leafBits);
  if (threadIdx.x < lpt) vr_tile_phase1(job, g, blk, threadIdx.x, strips);
  __syncthreads();
  uint8_t* dst = p.dst + sTileBase;
  for (uint32_t it = threadIdx.x; it < items; it += blockDim.x)
    vr_tile_phase2_item(g, it, strips, dst, p.rowPitch, p.slicePitch, p.vec16 != 0);
}

cudaError_t kd_decode_launch(const vr_kd_geom& g, const vr_kd_job* jobsDev, const vr_kd_work* workDev, int nWork, uint8_t* dst,
                             int64_t rowPitch, int64_t slicePitch, bool vec16, cudaStream_t stream, uint8_t* occ,
                             const int32_t* tileCell, const vr_fill_job* fillsDev) {
  if (nWork <= 0) return cudaSuccess;
  DecodeParams p;
  p.g = g;
  p.jobs = jobsDev;
  p.work = workDev;
  p.fills = fillsDev;
  p.dst = dst;
  p.rowPitch = rowPitch;
  p.slicePitch = slicePitch;
  p.vec16 = vec16 ? 1 : 0;
  p.occ = tileCell ? occ : nullptr;
  p.tileCell = tileCell;
  const size_t smem = kd_decode_smem_bytes(g);
  kd_decode_tiles_kernel<<<(unsigned)nWork, 256, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace vr
