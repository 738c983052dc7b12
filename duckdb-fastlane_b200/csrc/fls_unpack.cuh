// fls_unpack.cuh — per-thread FastLanes bit-unpack cores for sm_100a.
//
// A packed vector is W rows of 128 bytes.  One warp owns one vector and thread t owns the 32-bit column t
// of every packed row, so every row load is one fully-coalesced 128-byte request:
//   T=32 : thread t is lane t (32 values: rows r=0..31, position p = r*32 + t)
//   T=16 : thread t holds lanes 2t,2t+1 (2 x 16 values; a u32 store carries both -> p = r*64 + 2t + a)
//   T=8  : thread t holds lanes 4t..4t+3 (4 x 8 values; p = r*128 + 4t + a)
//   T=64 : thread t = (l = t&15, h = t>>4) holds half a lane: rows k = 32h + j, j=0..31 (p = k*16 + l).
//          The lane's stream is read as 32-bit words; thread h starts exactly at word h*W.
// W is a template parameter (all shifts/masks constant-folded); callers dispatch with FLS_SWITCH_W.
// Restates FastLanes `unffor` (missing upstream source; layout: SURVEY.md App. A.2).
//
// The same source compiles for the host when FLS_HOST_EMUL is defined (tests/emul): there a "warp" is a
// loop over t and every intrinsic below has a scalar twin.
#pragma once
#include <stdint.h>

#ifdef FLS_HOST_EMUL
#define FLS_DEV static inline
static inline uint32_t fls_shf_r(uint32_t lo, uint32_t hi, uint32_t s) {
  s &= 31;
  return s ? (lo >> s) | (hi << (32 - s)) : lo;
}
static inline uint32_t fls_prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint64_t v = ((uint64_t)b << 32) | a;
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xFF) << (8 * i);
  return r;
}
template <typename T> static inline T fls_ldg(const T *p) { return *p; }
#else
#define FLS_DEV __device__ __forceinline__
FLS_DEV uint32_t fls_shf_r(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_r(lo, hi, s); }
FLS_DEV uint32_t fls_prmt(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }
template <typename T> FLS_DEV T fls_ldg(const T *p) { return __ldg(p); }
#endif

// ---- T = 32 : v[r] = value at position r*32 + t
template <int W>
FLS_DEV void fls_unpack32(const uint32_t *__restrict__ in, uint32_t t, uint32_t base, uint32_t (&v)[32]) {
  if constexpr (W == 0) {
#pragma unroll
    for (int r = 0; r < 32; ++r) v[r] = base;
  } else {
    uint32_t w[W];
#pragma unroll
    for (int k = 0; k < W; ++k) w[k] = fls_ldg(in + k * 32 + t);
    constexpr uint32_t mask = (W == 32) ? 0xFFFFFFFFu : ((1u << (W & 31)) - 1u);
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      const int bit = r * W, k = bit >> 5, s = bit & 31;
      uint32_t x;
      if (s + W <= 32) x = w[k] >> s;
      else x = fls_shf_r(w[k], w[(k + 1 < W) ? k + 1 : k], s);
      v[r] = (x & mask) + base;
    }
  }
}

// ---- T = 64 : v[j] = value at row k = 32h + j of lane l  (t = h*16 + l); position p = k*16 + l
template <int W>
FLS_DEV void fls_unpack64(const uint32_t *__restrict__ in32, uint32_t t, uint64_t base, uint64_t (&v)[32]) {
  if constexpr (W == 0) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = base;
  } else {
    const uint32_t l = t & 15u, h = t >> 4;
    uint32_t x[W + 1];
#pragma unroll
    for (int i = 0; i < W; ++i) {
      const uint32_t m = h * W + i;  // 32-bit word index inside the lane stream
      x[i] = fls_ldg(in32 + ((m >> 1) * 16u + l) * 2u + (m & 1u));
    }
    x[W] = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int bit = j * W, a = bit >> 5, s = bit & 31;
      uint32_t lo, hi;
      // bits [s, s+W) of the word sequence x[a], x[a+1], x[a+2]
      const uint32_t x0 = x[a], x1 = x[(a + 1 <= W) ? a + 1 : W], x2 = x[(a + 2 <= W) ? a + 2 : W];
      lo = s ? fls_shf_r(x0, x1, s) : x0;
      if (W > 32) {
        hi = s ? fls_shf_r(x1, x2, s) : x1;
        if (W < 64) hi &= (1u << ((W - 32) & 31)) - 1u;
      } else {
        hi = 0;
        if (W < 32) lo &= (1u << (W & 31)) - 1u;
      }
      v[j] = (((uint64_t)hi << 32) | lo) + base;
    }
  }
}

// ---- T = 16 : v[r] packs lanes 2t (low half) and 2t+1 (high half) of row r; store as u32 at out32[r*32 + t]
template <int W>
FLS_DEV void fls_unpack16(const uint32_t *__restrict__ in, uint32_t t, uint16_t base, uint32_t (&v)[16]) {
  if constexpr (W == 0) {
#pragma unroll
    for (int r = 0; r < 16; ++r) v[r] = (uint32_t)base | ((uint32_t)base << 16);
  } else {
    uint32_t w[W];
#pragma unroll
    for (int k = 0; k < W; ++k) w[k] = fls_ldg(in + k * 32 + t);
    constexpr uint32_t mask = (1u << W) - 1u;  // W <= 16
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int bit = r * W, k = bit >> 4, s = bit & 15;
      const uint32_t w0 = w[k], w1 = w[(k + 1 < W) ? k + 1 : k];
      uint32_t a0 = fls_prmt(w0, w1, 0x5410), a1 = fls_prmt(w0, w1, 0x7632);  // {lane-half of w0, of w1}
      uint32_t lo = (((a0 >> s) & mask) + base) & 0xFFFFu;
      uint32_t hi = (((a1 >> s) & mask) + base) & 0xFFFFu;
      v[r] = lo | (hi << 16);
    }
  }
}

// ---- T = 8 : v[r] packs lanes 4t..4t+3 of row r (one byte each); store as u32 at out32[r*32 + t]
template <int W>
FLS_DEV void fls_unpack8(const uint32_t *__restrict__ in, uint32_t t, uint8_t base, uint32_t (&v)[8]) {
  if constexpr (W == 0) {
#pragma unroll
    for (int r = 0; r < 8; ++r) v[r] = (uint32_t)base * 0x01010101u;
  } else {
    uint32_t w[W];
#pragma unroll
    for (int k = 0; k < W; ++k) w[k] = fls_ldg(in + k * 32 + t);
    constexpr uint32_t mask = (1u << W) - 1u;  // W <= 8
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int bit = r * W, k = bit >> 3, s = bit & 7;
      const uint32_t w0 = w[k], w1 = w[(k + 1 < W) ? k + 1 : k];
      uint32_t out = 0;
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        uint32_t pair = fls_prmt(w0, w1, (uint32_t)(a | ((4 + a) << 4))) & 0xFFFFu;  // {byte a of w0, byte a of w1}
        out |= ((((pair >> s) & mask) + base) & 0xFFu) << (8 * a);
      }
      v[r] = out;
    }
  }
}

// Dispatch a statement over a runtime width: FLS_SWITCH_W(w, 32, { ... uses constexpr int W ... })
#define FLS_CASE_W(n, body) case n: { constexpr int W = n; body } break;
#define FLS_SWITCH_W8(w, body)                                                                      \
  switch (w) { FLS_CASE_W(0, body) FLS_CASE_W(1, body) FLS_CASE_W(2, body) FLS_CASE_W(3, body)      \
    FLS_CASE_W(4, body) FLS_CASE_W(5, body) FLS_CASE_W(6, body) FLS_CASE_W(7, body) FLS_CASE_W(8, body) default: break; }
#define FLS_CASES_9_16(body)                                                                        \
  FLS_CASE_W(9, body) FLS_CASE_W(10, body) FLS_CASE_W(11, body) FLS_CASE_W(12, body)                \
  FLS_CASE_W(13, body) FLS_CASE_W(14, body) FLS_CASE_W(15, body) FLS_CASE_W(16, body)
#define FLS_CASES_17_32(body)                                                                       \
  FLS_CASE_W(17, body) FLS_CASE_W(18, body) FLS_CASE_W(19, body) FLS_CASE_W(20, body)               \
  FLS_CASE_W(21, body) FLS_CASE_W(22, body) FLS_CASE_W(23, body) FLS_CASE_W(24, body)               \
  FLS_CASE_W(25, body) FLS_CASE_W(26, body) FLS_CASE_W(27, body) FLS_CASE_W(28, body)               \
  FLS_CASE_W(29, body) FLS_CASE_W(30, body) FLS_CASE_W(31, body) FLS_CASE_W(32, body)
#define FLS_CASES_33_64(body)                                                                       \
  FLS_CASE_W(33, body) FLS_CASE_W(34, body) FLS_CASE_W(35, body) FLS_CASE_W(36, body)               \
  FLS_CASE_W(37, body) FLS_CASE_W(38, body) FLS_CASE_W(39, body) FLS_CASE_W(40, body)               \
  FLS_CASE_W(41, body) FLS_CASE_W(42, body) FLS_CASE_W(43, body) FLS_CASE_W(44, body)               \
  FLS_CASE_W(45, body) FLS_CASE_W(46, body) FLS_CASE_W(47, body) FLS_CASE_W(48, body)               \
  FLS_CASE_W(49, body) FLS_CASE_W(50, body) FLS_CASE_W(51, body) FLS_CASE_W(52, body)               \
  FLS_CASE_W(53, body) FLS_CASE_W(54, body) FLS_CASE_W(55, body) FLS_CASE_W(56, body)               \
  FLS_CASE_W(57, body) FLS_CASE_W(58, body) FLS_CASE_W(59, body) FLS_CASE_W(60, body)               \
  FLS_CASE_W(61, body) FLS_CASE_W(62, body) FLS_CASE_W(63, body) FLS_CASE_W(64, body)
#define FLS_SWITCH_W16(w, body)                                                                     \
  switch (w) { FLS_CASE_W(0, body) FLS_CASE_W(1, body) FLS_CASE_W(2, body) FLS_CASE_W(3, body)      \
    FLS_CASE_W(4, body) FLS_CASE_W(5, body) FLS_CASE_W(6, body) FLS_CASE_W(7, body) FLS_CASE_W(8, body) \
    FLS_CASES_9_16(body) default: break; }
#define FLS_SWITCH_W32(w, body)                                                                     \
  switch (w) { FLS_CASE_W(0, body) FLS_CASE_W(1, body) FLS_CASE_W(2, body) FLS_CASE_W(3, body)      \
    FLS_CASE_W(4, body) FLS_CASE_W(5, body) FLS_CASE_W(6, body) FLS_CASE_W(7, body) FLS_CASE_W(8, body) \
    FLS_CASES_9_16(body) FLS_CASES_17_32(body) default: break; }
#define FLS_SWITCH_W64(w, body)                                                                     \
  switch (w) { FLS_CASE_W(0, body) FLS_CASE_W(1, body) FLS_CASE_W(2, body) FLS_CASE_W(3, body)      \
    FLS_CASE_W(4, body) FLS_CASE_W(5, body) FLS_CASE_W(6, body) FLS_CASE_W(7, body) FLS_CASE_W(8, body) \
    FLS_CASES_9_16(body) FLS_CASES_17_32(body) FLS_CASES_33_64(body) default: break; }
