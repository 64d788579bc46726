// kd_parse6.h -- the leaf-subtree parser of the persistent decode kernel (kd_decode.cu), written as host/device code so
// that tests/emu runs exactly this source thread by thread on the CPU (test harness only).
//
// What it parses: one depth-(D0-6) subtree of the preorder 2-bit stream that the reference's VolumeKdtree::levelCut walks
// with a stack of boxes (/root/reference/volume_renderer/VolumeKdTree_recover.cpp:726-835): 63 inner nodes at most, 64
// voxels, every voxel followed by its refinement chain (<= 7 symbols, :806-833).  One thread owns one such subtree (its
// start offset and root value come from the side index, kd_stream.h) and writes the 4x4x4 voxels of its box, x fastest,
// into a 64-byte block of shared memory; the CTA then stores whole 16-byte x-runs (kd_decode.cu, phase 2).
//
// Why it looks the way it does (profiles/r2_step_final.md has the instruction counts and the A/B runs):
//  * values: a node's value is one of {parent, sat(parent + d), sat(parent - d), parent} selected by its symbol
//    (:784-787).  The parent builds that "select word" once (one fused add-min-relu in 16-bit lanes + one byte permute +
//    one multiply) and each child is ONE byte permute -- instead of 6 ALU operations per node.
//  * chains: one look-up in a 16 K-entry table (shared memory) gives the whole chain as x -> min(max(x + A, lo), hi) and
//    the number of bits the voxel item occupies.
//  * the two voxels of a height-1 node are finished together in 16-bit lanes (voxel_pair): 10 instead of 14 instructions
//    on the integer pipe.
//  * bit window: the three lowest levels of the tree (an octet: 7 nodes, 8 voxels) are straight-line code, so the
//    places where the 32-bit window has to be re-aligned are known at compile time: 6 per octet instead of one per
//    symbol, each branch-free (a refill is two predicated instructions).  Between them a token costs one shift.
//  * stream words come from shared memory: the unit's segments (at most 4.7 KB) are staged by asynchronous copies, so the
//    parser's dependent chain never waits on global memory; the window keeps a third word loaded one boundary ahead.
#pragma once
#include "kd_common.h"

#if defined(__CUDA_ARCH__)
#define VR6_DEV 1
#else
#define VR6_DEV 0
#endif
#if defined(__CUDACC__)
#define VR6_CE __host__ __device__ constexpr
#else
#define VR6_CE constexpr
#endif

namespace vr6 {

// ---- chain table ----------------------------------------------------------------------------------------------------
// Entry for the 14 bits that follow a voxel's own symbol (7 chain symbols, LSB first):
//   bits 0-7   number of stream bits of the whole voxel item (own symbol + chain symbols consumed) = 2 + 2 * L
//   bits 8-15  lo, bits 16-23 hi, bits 24-31 A (signed): the chain maps x to min(max(x + A, lo), hi)
// (a composition of saturating +-d steps is always of that form).  IDENT = the entry of "no chain" (own symbol 3).
#define VR6_IDENT 0x00FF0002u
VR_HD uint32_t chain_entry(uint32_t codes14, uint64_t chainDist) {
  int A = 0, lo = 0, hi = 255, L = VR_ADD_LEVELS;
  for (int k = 0; k < VR_ADD_LEVELS; ++k) {
    const int c = (int)((codes14 >> (2 * k)) & 3);
    if (c == 3) { L = k + 1; break; }
    const int d = (int)((chainDist >> (8 * k)) & 0xFF);
    const int a = c == 1 ? d : (c == 2 ? -d : 0);
    A += a;
    lo += a; lo = lo < 0 ? 0 : (lo > 255 ? 255 : lo);
    hi += a; hi = hi < 0 ? 0 : (hi > 255 ? 255 : hi);
  }
  return (uint32_t)(2 + 2 * L) | ((uint32_t)lo << 8) | ((uint32_t)hi << 16) | ((uint32_t)(A & 0xFF) << 24);
}
#define VR6_LUT_ENTRIES 16384u
#define VR6_LUT_BYTES 65536u

// ---- memory policy: "shared" addresses are 32-bit byte addresses (device: the shared window; host emulation: offsets
// into one array that stands for the CTA's shared memory) -----------------------------------------------------------------
#if VR6_DEV
struct Mem {
  static __device__ __forceinline__ uint32_t lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
  static __device__ __forceinline__ void st8(uint32_t a, uint32_t v) { asm volatile("st.shared.b8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
  static __device__ __forceinline__ void st16(uint32_t a, uint32_t v) { asm volatile("st.shared.b16 [%0], %1;" ::"r"(a), "h"((unsigned short)v) : "memory"); }
  static __device__ __forceinline__ void st32(uint32_t a, uint32_t v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
  static __device__ __forceinline__ uint32_t ldg32(const uint32_t* p) { return __ldg(p); }
  static __device__ __forceinline__ uint32_t funnel(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_r(lo, hi, s); }
  static __device__ __forceinline__ uint32_t perm(uint32_t a, uint32_t b, uint32_t s) {  // prmt as is (__byte_perm masks the selector first)
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(s));
    return r;
  }
};
#else
struct Mem {
  static inline uint8_t*& base() { static thread_local uint8_t* b = nullptr; return b; }
  static inline uint32_t lds32(uint32_t a) { uint32_t v; __builtin_memcpy(&v, base() + a, 4); return v; }
  static inline void st8(uint32_t a, uint32_t v) { base()[a] = (uint8_t)v; }
  static inline void st16(uint32_t a, uint32_t v) { uint16_t h = (uint16_t)v; __builtin_memcpy(base() + a, &h, 2); }
  static inline void st32(uint32_t a, uint32_t v) { __builtin_memcpy(base() + a, &v, 4); }
  static inline uint32_t ldg32(const uint32_t* p) { return *p; }
  static inline uint32_t funnel(uint32_t lo, uint32_t hi, uint32_t s) { return (uint32_t)(((((uint64_t)hi) << 32) | (uint64_t)lo) >> (s & 31)); }
  static inline uint32_t perm(uint32_t a, uint32_t b, uint32_t s) {
    const uint64_t src = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int k = 0; k < 4; ++k) {
      const uint32_t n = (s >> (4 * k)) & 0xF;
      uint32_t byte = (uint32_t)(src >> (8 * (n & 7))) & 0xFF;
      if (n & 8) byte = (byte & 0x80) ? 0xFF : 0x00;
      r |= byte << (8 * k);
    }
    return r;
  }
};
#endif

// ---- bit windows ------------------------------------------------------------------------------------------------------
// P = absolute bit position in the stream of the next unread symbol; (cur, nxt) = the stream words that hold bits
// [32 * (P >> 5), +64).  advance_peek(n) consumes (n & 0xFF) <= 32 bits and returns the next 32 bits; crossing a word boundary
// is the only conditional part (two moves and a load under one predicate, no branch).  Only bits 0..5 of P are ever looked at
// (the word address is tracked separately), so n may carry garbage above bit 7.
struct WinShared {  // stream segment staged in shared memory; base = shared address of (virtual) word 0 of the stream
  // nn = the word behind nxt, loaded one boundary ahead: the load of a refill is then not on the parser's dependent chain
  // (window -> table look-up -> shift -> window ...), where it cost a shared-memory round trip at every boundary
  uint32_t cur, nxt, nn, P, wi;
  VR_HD void init(uint32_t base, uint32_t sym) {
    P = sym * 2u;
    const uint32_t a = base + ((P >> 5) << 2);
    cur = Mem::lds32(a);
    nxt = Mem::lds32(a + 4u);
    nn = Mem::lds32(a + 8u);
    wi = a + 12u;
  }
  VR_HD uint32_t advance_peek(uint32_t n) {
    const uint32_t Pn = P + n;
    if ((P ^ Pn) & 32u) { cur = nxt; nxt = nn; nn = Mem::lds32(wi); wi += 4u; }
    P = Pn;
    return Mem::funnel(cur, nxt, P);
  }
};
// ---- geometry of a 4x4x4 leaf box -------------------------------------------------------------------------------------
// The six lowest tree levels of the bricks this path accepts split x, y, z cyclically (VolumeKdTree_recover.cpp:151-163 with
// all three extents still >= 4), so path bit b of a voxel inside its leaf box (bit 0 = last split) carries coordinate bit
// b / 3 of axis (ROT - b) mod 3, ROT = (D0 - 1) mod 3.  Block layout: byte offset = x + 4 y + 16 z.
template <int ROT>
struct Geo {
  static VR6_CE int axis(int b) { return ((ROT - b) % 3 + 3) % 3; }
  static VR6_CE int bitOff(int b) { return (axis(b) == 0 ? 1 : (axis(b) == 1 ? 4 : 16)) << (b / 3); }
  static VR6_CE int off(int path) {  // path: up to 6 bits
    int o = 0;
    for (int b = 0; b < 6; ++b) o += ((path >> b) & 1) ? bitOff(b) : 0;
    return o;
  }
  static constexpr int xbit = (ROT + 3) % 3;  // the path bit (0..2) that carries x0: axis(b) == 0  <=>  b == ROT (mod 3)
};

// All 2^H voxels of the box of a height-H node (H <= 3) at static path position POS take `val`.
template <int ROT, int H, int POS>
VR_HD void fill_small(uint32_t base, uint32_t val) {
  constexpr int xb = Geo<ROT>::xbit;
  if constexpr (xb < H) {  // the box spans both x0 values: pairs of x neighbours as 16-bit stores
    const uint32_t v2 = val * 0x0101u;
#pragma unroll
    for (int k = 0; k < (1 << H); ++k)
      if (!((k >> xb) & 1)) Mem::st16(base + (uint32_t)Geo<ROT>::off(POS + k), v2);
  } else {
#pragma unroll
    for (int k = 0; k < (1 << H); ++k) Mem::st8(base + (uint32_t)Geo<ROT>::off(POS + k), val);
  }
}
// Octets [i, i + n) of the leaf box take `val` (pruned nodes above the octets; n = 1, 2, 4).
template <int ROT>
VR_HD void fill_octets(uint32_t strip, uint32_t i, uint32_t n, uint32_t val) {
  for (uint32_t k = i; k < i + n; ++k) {
    const uint32_t base = strip + ((k & 1u) ? (uint32_t)Geo<ROT>::bitOff(3) : 0u) + ((k & 2u) ? (uint32_t)Geo<ROT>::bitOff(4) : 0u) +
                          ((k & 4u) ? (uint32_t)Geo<ROT>::bitOff(5) : 0u);
    fill_small<ROT, 3, 0>(base, val);
  }
}

// ---- values -----------------------------------------------------------------------------------------------------------
// Select word of a node with value n whose children sit at a level with distance d: bytes (n, sat(n+d), sat(n-d), n), so
// that a child with symbol c (0 same, 1 add, 2 subtract, 3 same-and-stop: VolumeKdTree_recover.cpp:784-787) is byte c.
VR_HD int imin(int a, int b) { return a < b ? a : b; }
VR_HD int imax(int a, int b) { return a > b ? a : b; }
// d2 = the level's distance packed for 16-bit lanes: d in the low half, -d in the high half (dist2()).
VR_HD uint32_t dist2(uint32_t d) { return d + ((0u - d) << 16); }
VR_HD uint32_t mkword(uint32_t n, uint32_t d2) {
  // (n, n) + (d, -d) in 16-bit lanes, capped at (255, 32767) and floored at 0 -- ONE fused add-min-relu gives sat(n+d) and
  // sat(n-d); one permute places them between two copies of n.  The copies come from a multiply (the multiply-add pipe is
  // nearly idle; the integer pipe that carries every logic operation, shift and min/max of this parser is its bottleneck).
  const uint32_t n2 = n * 0x00010001u;
#if VR6_DEV
  const uint32_t t = __viaddmin_s16x2_relu(n2, d2, 0x7FFF00FFu);
#else
  const int up = imin((int)n + (int)(d2 & 0xFFFFu), 255), dn = imax((int)n - (int)(d2 & 0xFFFFu), 0);
  const uint32_t t = (uint32_t)up | ((uint32_t)dn << 16);
#endif
  return Mem::perm(n2, t, 0x0640u);  // bytes (n, up, dn, n)
}
VR_HD uint32_t child(uint32_t word, uint32_t c) { return Mem::perm(word, 0u, c | 0x4440u); }
// the same from the window itself: selector = (bits & 3) | 0x4440 in one three-input logic operation
VR_HD uint32_t child_of_bits(uint32_t word, uint32_t bits) {
#if VR6_DEV
  uint32_t sel;
  asm("lop3.b32 %0, %1, 3, %2, 0xEA;" : "=r"(sel) : "r"(bits), "r"(0x4440u));
  return Mem::perm(word, 0u, sel);
#else
  return Mem::perm(word, 0u, (bits & 3u) | 0x4440u);
#endif
}

struct Dist { uint32_t d0, d1, d2, d3, d4, d5; };  // d_h = dist2(distanceMap[D0 - h]): applied to nodes at height h

// ---- the octet: straight-line code ------------------------------------------------------------------------------------
// State threaded through the octet: `bits` = the window shifted so that the next token is at bit 0, `o` = bits consumed
// since the last re-alignment.  REPEEK marks the static places where the window is re-aligned (see the header comment):
// before voxels 1, 2, 5, 6 and before the second height-2 node.
template <class WIN>
struct Cur {
  WIN& w;
  uint32_t bits, o;
  uint32_t lut;  // shared address of the chain table
  // o may carry garbage above bit 7 (voxel() adds whole table entries): the window only ever looks at bits 0..5 of its position
  VR_HD void repeek() { bits = w.advance_peek(o); o = 0; }
};

// Chain-table entry of the voxel item at the front of `bits`; the identity entry when the voxel's own symbol is 3.
VR_HD uint32_t item_entry(uint32_t lut, uint32_t bits) {
#if VR6_DEV
  uint32_t e;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b32 t;\n"
      "lop3.b32 t, %2, 3, 0, 0x0C;\n"   // ~bits & 3: zero iff the symbol is 3
      "setp.ne.u32 p, t, 0;\n"
      "mov.b32 %0, %3;\n"
      "@p ld.shared.b32 %0, [%1];\n"
      "}\n"
      : "=r"(e)
      : "r"(lut + (bits & 0xFFFCu)), "r"(bits), "n"(VR6_IDENT));
  return e;
#else
  return (bits & 3u) == 3u ? VR6_IDENT : Mem::lds32(lut + (bits & 0xFFFCu));
#endif
}
VR_HD bool is3(uint32_t bits) {  // front symbol == 3, as one logic operation with a predicate result
#if VR6_DEV
  uint32_t t;
  asm("lop3.b32 %0, %1, 3, 0, 0x0C;" : "=r"(t) : "r"(bits));  // ~bits & 3
  return t == 0u;
#else
  return (bits & 3u) == 3u;
#endif
}
VR_HD uint32_t shr_wrap(uint32_t v, uint32_t n) {  // v >> (n & 31): a funnel shift takes the count modulo 32, so no mask is needed
#if VR6_DEV
  return __funnelshift_r(v, 0u, n);
#else
  return v >> (n & 31u);
#endif
}

template <int ROT, int POS, class WIN>
VR_HD void voxel(Cur<WIN>& s, uint32_t word, uint32_t base) {
  if (POS == 1 || POS == 2 || POS == 5 || POS == 6) s.repeek();
  const uint32_t e = item_entry(s.lut, s.bits);
  const int x = (int)child_of_bits(word, s.bits);
  const int lo = (int)Mem::perm(e, 0u, 0x4441u), hi = (int)Mem::perm(e, 0u, 0x4442u);
  int v = x + ((int)e >> 24);
  v = imax(v, lo);
  v = imin(v, hi);
  Mem::st8(base + (uint32_t)Geo<ROT>::off(POS), (uint32_t)v);
  s.bits = shr_wrap(s.bits, e);  // low byte of e = bits of the item (<= 16)
  s.o += e;                      // only the low byte counts: repeek() masks
}

// Both voxels of a height-1 node at once: the two chains are looked up one after the other (the second item starts where the
// first one ends), the arithmetic runs in 16-bit lanes -- one permute each gathers (x0, x1), (A0, A1) sign-extended,
// (lo0, lo1), (hi0, hi1) from the select word and the two table entries, one fused add-max and one min finish both.
// 10 instructions instead of 14 on the pipe that bounds the parser (VR6_PAIR=0 keeps the one-by-one version for A/B).
#ifndef VR6_PAIR
#define VR6_PAIR 1
#endif
VR_HD uint32_t addmax_s16x2(uint32_t a, uint32_t b, uint32_t c) {  // per 16-bit lane: max(a + b, c), signed
#if VR6_DEV
  return __viaddmax_s16x2(a, b, c);
#else
  uint32_t r = 0;
  for (int k = 0; k < 2; ++k) {
    const int x = (int16_t)(a >> (16 * k)) + (int16_t)(b >> (16 * k)), y = (int16_t)(c >> (16 * k));
    r |= ((uint32_t)(uint16_t)(int16_t)(x > y ? x : y)) << (16 * k);
  }
  return r;
#endif
}
VR_HD uint32_t min_s16x2(uint32_t a, uint32_t b) {
#if VR6_DEV
  return __vmins2(a, b);
#else
  uint32_t r = 0;
  for (int k = 0; k < 2; ++k) {
    const int x = (int16_t)(a >> (16 * k)), y = (int16_t)(b >> (16 * k));
    r |= ((uint32_t)(uint16_t)(int16_t)(x < y ? x : y)) << (16 * k);
  }
  return r;
#endif
}
template <int ROT, int POS, class WIN>
VR_HD void voxel_pair(Cur<WIN>& s, uint32_t word, uint32_t base) {
  if (POS == 2 || POS == 6) s.repeek();
  const uint32_t bits0 = s.bits;
  const uint32_t e0 = item_entry(s.lut, bits0);
  s.bits = shr_wrap(bits0, e0);
  s.o += e0;
  if (POS == 0 || POS == 4) s.repeek();  // before voxels 1 and 5
  const uint32_t bits1 = s.bits;
  const uint32_t e1 = item_entry(s.lut, bits1);
  s.bits = shr_wrap(bits1, e1);
  s.o += e1;
  // selector of (x0, 0, x1, 0): the two symbols in nibbles 0 and 2, nibbles 1 and 3 = byte 4 (zero)
#if VR6_DEV
  uint32_t sel;
  asm("{\n"
      ".reg .b32 t;\n"
      "lop3.b32 t, %2, 0x300, 0x4040, 0xEA;\n"  // (a & b) | c
      "lop3.b32 %0, %1, 3, t, 0xEA;\n"
      "}\n"
      : "=r"(sel)
      : "r"(bits0), "r"(bits1 * 256u));
#else
  const uint32_t sel = (bits0 & 3u) | ((bits1 * 256u) & 0x300u) | 0x4040u;
#endif
  const uint32_t X = Mem::perm(word, 0u, sel);
  const uint32_t A = Mem::perm(e0, e1, 0xF7B3u);   // (A0, sign, A1, sign)
  const uint32_t LO = Mem::perm(e0, e1, 0x8581u);  // (lo0, 0, lo1, 0): nibble 8 = sign of the length byte = 0
  const uint32_t HI = Mem::perm(e0, e1, 0x8682u);
  const uint32_t V = min_s16x2(addmax_s16x2(X, A, LO), HI);
  if constexpr (Geo<ROT>::xbit == 0) {  // x neighbours: one 16-bit store
    Mem::st16(base + (uint32_t)Geo<ROT>::off(POS), Mem::perm(V, 0u, 0x4420u));
  } else {
    Mem::st8(base + (uint32_t)Geo<ROT>::off(POS), V);
    Mem::st8(base + (uint32_t)Geo<ROT>::off(POS + 1), V >> 16);
  }
}

template <int ROT, int H, int POS, class WIN>
struct Node {  // H = 1 or 2
  static VR_HD void run(Cur<WIN>& s, uint32_t word, const Dist& d, uint32_t base) {
    if (H == 2 && POS == 4) s.repeek();
    const uint32_t bits0 = s.bits;
    s.bits >>= 2;
    s.o += 2u;
    if (is3(bits0)) {
      fill_small<ROT, H, POS>(base, word >> 24);
    } else {
      const uint32_t cw = mkword(child_of_bits(word, bits0), H == 2 ? d.d1 : d.d0);
      if constexpr (H == 2) {
        Node<ROT, 1, POS, WIN>::run(s, cw, d, base);
        Node<ROT, 1, POS + 2, WIN>::run(s, cw, d, base);
      } else {
#if VR6_PAIR
        voxel_pair<ROT, POS, WIN>(s, cw, base);
#else
        voxel<ROT, POS, WIN>(s, cw, base);
        voxel<ROT, POS + 1, WIN>(s, cw, base);
#endif
      }
    }
  }
};

// One leaf subtree, in two halves (octets 0-3 = the first height-5 node, octets 4-7 = the second) so that the kernel can do
// warp-wide work between them (kd_decode.cu stages the next unit there).  The window stands on the root's own symbol
// (known to be an inner node; its value is rootVal) when the first half starts; `o` = bits consumed since the window's last
// re-alignment (2 = the root's symbol) is carried from one half to the next.
template <int ROT, int I0, class WIN>
VR_HD void parse_leaf_half(WIN& w, uint32_t& o, uint32_t word6, const Dist& d, uint32_t lut, uint32_t strip) {
  Cur<WIN> s{w, 0u, o, lut};
  uint32_t word5 = 0, word4 = 0;
  uint32_t i = (uint32_t)I0;  // octet index inside the leaf box (path bits 5..3)
  for (;;) {
    s.repeek();
    if ((i & 3u) == 0u) {  // a height-5 node starts here
      const uint32_t c = s.bits & 3u;
      s.bits >>= 2; s.o += 2u;
      if (c == 3u) {
        fill_octets<ROT>(strip, i, 4u, word6 >> 24);
        break;
      }
      word5 = mkword(child(word6, c), d.d4);
    }
    if ((i & 1u) == 0u) {  // a height-4 node
      const uint32_t c = s.bits & 3u;
      s.bits >>= 2; s.o += 2u;
      if (c == 3u) {
        fill_octets<ROT>(strip, i, 2u, word5 >> 24);
        i += 2u;
        if (i == (uint32_t)I0 + 4u) break;
        continue;
      }
      word4 = mkword(child(word5, c), d.d3);
    }
    {  // the octet's root (height 3)
      const uint32_t c = s.bits & 3u;
      s.bits >>= 2; s.o += 2u;
      const uint32_t base = strip + ((i & 1u) ? (uint32_t)Geo<ROT>::bitOff(3) : 0u) + ((i & 2u) ? (uint32_t)Geo<ROT>::bitOff(4) : 0u) +
                            ((i & 4u) ? (uint32_t)Geo<ROT>::bitOff(5) : 0u);
      if (c == 3u) {
        fill_small<ROT, 3, 0>(base, word4 >> 24);
      } else {
        const uint32_t word3 = mkword(child(word4, c), d.d2);
        Node<ROT, 2, 0, WIN>::run(s, word3, d, base);
        Node<ROT, 2, 4, WIN>::run(s, word3, d, base);
      }
    }
    i += 1u;
    if (i == (uint32_t)I0 + 4u) break;
  }
  o = s.o;
}
template <int ROT, class WIN>
VR_HD void parse_leaf(WIN& w, uint32_t rootVal, const Dist& d, uint32_t lut, uint32_t strip) {
  uint32_t o = 2u;  // the root's symbol
  const uint32_t word6 = mkword(rootVal, d.d5);
  parse_leaf_half<ROT, 0, WIN>(w, o, word6, d, lut, strip);
  parse_leaf_half<ROT, 4, WIN>(w, o, word6, d, lut, strip);
}

// Whole leaf box constant (index entry VR_PRUNED).
VR_HD void fill_leaf(uint32_t strip, uint32_t val) {
  const uint32_t w = val * 0x01010101u;
#pragma unroll
  for (int k = 0; k < 16; ++k) Mem::st32(strip + 4u * (uint32_t)k, w);
}

#define VR6_STRIP_PITCH 68u  // 64-byte leaf block + 4: the blocks of a warp's 32 lanes start in 32 different banks
#define VR6_STRIPS_BYTES (256u * VR6_STRIP_PITCH)

// ---- units and phase 2: leaf blocks -> 16-byte x-runs of the destination -------------------------------------------------
// A tile (14 path bits) has 256 leaf boxes, leaf-index bit k = path bit 6 + k.  A warp's unit = the 32 leaf boxes that share
// the three "unit" bits; the five "lane" bits are the three leaf-index bits that carry tile x bits 2, 3, 4 plus the two lowest
// other ones, so that the unit's box is 32 (all of x) x 8 x 8 voxels.  Lane l <-> the leaf box whose lane bits spell l.  The
// lane bits that form a block from path bit 6 upwards enumerate consecutive leaf boxes of the stream: a unit is 32 >> runLog
// runs of (1 << runLog) consecutive leaf boxes.
// The unit's 2 048 voxels = 128 runs of 16 voxels along x.  Run m (7 bits) = one value of the unit's free coordinate bits:
// the 4 row bits inside a leaf box (y0, y1, z0, z1) and the 3 lane bits that are not tile x bits 2 / 3 (one of them is x
// bit 4).  A run is four 4-byte rows of four x-neighbouring leaf blocks.
struct P6Geom {
  int32_t rot;           // (D0 - 1) mod 3
  uint16_t xs[2];        // strip offset of tile x bits 2, 3 (a lane bit each)
  uint16_t fs[7];        // strip offset of free bit k
  uint8_t fAxis[7];      // its axis (0 = x: tile x bit 4, 1 = y, 2 = z) ...
  uint8_t fCb[7];        // ... and tile-local coordinate bit
  uint8_t wAxis[3], wCb[3];  // the unit bits likewise
  uint8_t boxLy, boxLz;      // a unit's box is 32 x 2^boxLy x 2^boxLz voxels (32 x 8 x 8 for 256x256x128 bricks)
  uint8_t laneBit[5];        // leaf-index bit of lane bit k (ascending)
  uint8_t unitBit[3];        // leaf-index bit of unit bit j (ascending)
  uint8_t runLog;            // log2 of the run length: 3 (4 runs of 8) or 4 (2 runs of 16)
};
#define VR6_UNIT_X 32
// leaf index (0..255, relative to the tile's first) of lane l of unit u
VR_HD uint32_t p6_leaf_of(const P6Geom& p, uint32_t u, uint32_t l) {
  uint32_t r = 0;
#pragma unroll
  for (int k = 0; k < 5; ++k) r |= ((l >> k) & 1u) << p.laneBit[k];
#pragma unroll
  for (int j = 0; j < 3; ++j) r |= ((u >> j) & 1u) << p.unitBit[j];
  return r;
}
// false when the brick's tree does not have the shape this path is written for (kd_decode.cu then takes the generic kernel)
static inline bool p6_make_geom(const vr_kd_geom& g, P6Geom* out) {
  if (g.tileBits != 14 || g.leafBits != 6 || g.tx != VR6_UNIT_X) return false;
  const int rot = (g.D0 - 1) % 3;
  for (int b = 0; b < 6; ++b) {
    const int ax = ((rot - b) % 3 + 3) % 3;
    if (g.pbAxis[b] != ax || g.pbCoordBit[b] != b / 3) return false;
  }
  P6Geom p{};
  p.rot = rot;
  // lane bits: the x bits of the tile's leaf index and the two lowest others
  bool isLane[8] = {false, false, false, false, false, false, false, false};
  int nx = 0, nother = 0;
  for (int k = 0; k < 8; ++k)
    if (g.pbAxis[6 + k] == 0) { isLane[k] = true; ++nx; }
  if (nx != 3) return false;
  for (int k = 0; k < 8 && nother < 2; ++k)
    if (!isLane[k]) { isLane[k] = true; ++nother; }
  int nl = 0, nu = 0;
  for (int k = 0; k < 8; ++k) {
    if (isLane[k]) p.laneBit[nl++] = (uint8_t)k;
    else p.unitBit[nu++] = (uint8_t)k;
  }
  if (nl != 5 || nu != 3) return false;
  int run = 0;
  while (run < 5 && p.laneBit[run] == run) ++run;
  if (run < 3) return false;  // more than 4 runs per unit: not worth staging
  p.runLog = (uint8_t)(run > 4 ? 4 : run);  // (5 = one run of 32; handled as two of 16)
  auto lanePos = [&](int leafBit) { for (int k = 0; k < 5; ++k) if (p.laneBit[k] == leafBit) return k; return -1; };
  const int x2 = g.pathBit[0][2] - 6, x3 = g.pathBit[0][3] - 6, x4 = g.pathBit[0][4] - 6;
  if (lanePos(x2) < 0 || lanePos(x3) < 0 || lanePos(x4) < 0) return false;
  p.xs[0] = (uint16_t)(VR6_STRIP_PITCH << lanePos(x2));
  p.xs[1] = (uint16_t)(VR6_STRIP_PITCH << lanePos(x3));
  int n = 0;
  for (int b = 0; b < 6; ++b) {  // in-leaf row bits first (lowest) ...
    if (g.pbAxis[b] == 0) continue;  // x0, x1: inside a 4-byte row
    p.fs[n] = (uint16_t)((g.pbAxis[b] == 1 ? 4 : 16) << (b / 3));
    p.fAxis[n] = g.pbAxis[b];
    p.fCb[n] = g.pbCoordBit[b];
    ++n;
  }
  for (int k = 0; k < 5; ++k) {  // ... then the lane bits that are not x bits 2 / 3
    const int lb = p.laneBit[k];
    if (lb == x2 || lb == x3) continue;
    p.fs[n] = (uint16_t)(VR6_STRIP_PITCH << k);
    p.fAxis[n] = g.pbAxis[6 + lb];
    p.fCb[n] = g.pbCoordBit[6 + lb];
    ++n;
  }
  if (n != 7) return false;
  for (int k = 0; k < 7; ++k) {  // the free bits are x bit 4 and the low y and z bits of the unit's box, each axis without gaps
    if (p.fAxis[k] == 1) { if (p.fCb[k] != p.boxLy) return false; ++p.boxLy; }
    else if (p.fAxis[k] == 2) { if (p.fCb[k] != p.boxLz) return false; ++p.boxLz; }
    else if (p.fCb[k] != 4) return false;
  }
  for (int j = 0; j < 3; ++j) { p.wAxis[j] = g.pbAxis[6 + p.unitBit[j]]; p.wCb[j] = g.pbCoordBit[6 + p.unitBit[j]]; }
  *out = p;
  return true;
}
// offset of run m inside the unit's dense box [z][y][32 x] (what the tensor store of a unit reads)
VR_HD uint32_t p6_box_off(const P6Geom& p, uint32_t m) {
  uint32_t o = 0;
#pragma unroll
  for (int k = 0; k < 7; ++k)
    if ((m >> k) & 1u) o += p.fAxis[k] == 0 ? 16u : ((p.fAxis[k] == 1 ? (uint32_t)VR6_UNIT_X : ((uint32_t)VR6_UNIT_X << p.boxLy)) << p.fCb[k]);
  return o;
}
// voxel coordinates of unit u's box relative to the tile origin
VR_HD void p6_unit_origin(const P6Geom& p, uint32_t u, int32_t o[3]) {
  o[0] = o[1] = o[2] = 0;
  for (int j = 0; j < 3; ++j)
    if ((u >> j) & 1u) o[p.wAxis[j]] += 1 << p.wCb[j];
}
// strip offset (relative to the unit's 32 strips, lane order) and destination offset (relative to the tile origin) of run m of unit w
VR_HD void p6_run_addr(const P6Geom& p, uint32_t w, uint32_t m, int64_t rowPitch, int64_t slicePitch, uint32_t* sOff, int64_t* gOff) {
  uint32_t s = 0;
  int64_t go = 0;
#pragma unroll
  for (int k = 0; k < 7; ++k)
    if ((m >> k) & 1u) {
      s += p.fs[k];
      go += (p.fAxis[k] == 0 ? (int64_t)1 : (p.fAxis[k] == 1 ? rowPitch : slicePitch)) << p.fCb[k];
    }
#pragma unroll
  for (int j = 0; j < 3; ++j)
    if ((w >> j) & 1u) go += (p.wAxis[j] == 0 ? (int64_t)1 : (p.wAxis[j] == 1 ? rowPitch : slicePitch)) << p.wCb[j];
  *sOff = s;
  *gOff = go;
}

}  // namespace vr6
