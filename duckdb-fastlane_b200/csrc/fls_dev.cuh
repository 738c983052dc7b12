// fls_dev.cuh — warp-level building blocks of the sm_100a FLS decoders.
//
// Execution model: ONE WARP decodes ONE 1024-value FastLanes vector.
//   1. lane 0 issues a TMA 1-D bulk copy (cp.async.bulk, SASS UBLKCP) of the vector's packed payload from HBM
//      into the warp's shared-memory slot and the warp sleeps on its own mbarrier — no registers are tied up
//      by bytes in flight, so occupancy (and therefore memory-level parallelism) is set by shared memory only;
//   2. the lanes unpack from shared memory with the bit width as a RUNTIME value (two conflict-free LDS.128 +
//      funnel shifts per 4 values — there is no 0..64 switch and no per-width code bloat);
//   3. every thread owns 16 contiguous output bytes of a row, so 8 threads write one full 128-byte line with
//      STG.128 and a warp-wide store covers four lines.
//
// The same source compiles for the host when FLS_HOST_EMUL is defined (tests/emul): a "warp" is then a loop
// over t = 0..31 for each FLS_LANES region (regions are separated by a warp barrier on the device), the bulk
// copy is a memcpy and every intrinsic has a scalar twin.  That is TEST INFRASTRUCTURE for the GPU-less
// development container; the product library is built without it and has no CPU decode.
//
// Restated FastLanes primitives (upstream cwida/FastLanes source is absent from /root/reference; layout per
// SURVEY.md App. A.2/A.3): unffor = fls_unpack_*; transposed-layout walk = fls_chain_row.
#pragma once
#include <stdint.h>
#include <string.h>

#include "fls_format.h"

#ifdef FLS_HOST_EMUL
#define FLS_DEV static inline
#define FLS_MEM inline
#define FLS_LANES(t) for (uint32_t t = 0; t < 32u; ++t)
#define FLS_VAR(T, n) T n##_l[32]
#define FLS_AT(n) n##_l[t]
#define FLS_BCAST(n, lane) n##_l[lane]
#define FLS_VAR2(T, n, K) T n##_l[32][K]
#define FLS_AT2(n, j) n##_l[t][j]
struct fls_u4 { uint32_t x, y, z, w; };
struct fls_ul2 { uint64_t x, y; };
template <typename T> static inline T fls_ldg(const T *p) { return *p; }
template <typename T> static inline T fls_ldcg(const T *p) { return *p; }
static inline uint32_t fls_shf_r(uint32_t lo, uint32_t hi, uint32_t s) {
  s &= 31u;
  return s ? (lo >> s) | (hi << (32u - s)) : lo;
}
// funnel shift left (wrap): high word of (hi:lo) << (s & 31)
static inline uint32_t fls_shf_l(uint32_t lo, uint32_t hi, uint32_t s) {
  s &= 31u;
  return s ? (hi << s) | (lo >> (32u - s)) : hi;
}
// funnel shift right, shift amount clamped to 32: low word of (hi:lo) >> min(s, 32)
static inline uint32_t fls_shf_rc(uint32_t lo, uint32_t hi, uint32_t s) {
  return s >= 32u ? hi : (s ? (lo >> s) | (hi << (32u - s)) : lo);
}
// x << min(n, 32): 0 for n >= 32
static inline uint32_t fls_shl_c(uint32_t x, uint32_t n) { return n >= 32u ? 0u : x << n; }
static inline uint32_t fls_prmt(uint32_t a, uint32_t b, uint32_t sel) {
  const uint64_t v = ((uint64_t)b << 32) | a;
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
  return r;
}
template <typename V> static inline void fls_st128(void *p, const V &v) { memcpy(p, &v, 16); }
template <typename V> static inline V fls_lds128(const void *p) { V v; memcpy(&v, p, 16); return v; }
struct fls_u2 { uint32_t x, y; };
static inline fls_u2 fls_lds64(const void *p) { fls_u2 v; memcpy(&v, p, 8); return v; }
static inline void fls_sts64(void *p, uint32_t lo, uint32_t hi) { memcpy(p, &lo, 4); memcpy((uint8_t *)p + 4, &hi, 4); }
static inline void fls_st64(void *p, uint32_t lo, uint32_t hi) { memcpy(p, &lo, 4); memcpy((uint8_t *)p + 4, &hi, 4); }
static inline uint32_t fls_min(uint32_t a, uint32_t b) { return a < b ? a : b; }
static inline uint32_t fls_max(uint32_t a, uint32_t b) { return a > b ? a : b; }
static inline void fls_smem_max(uint32_t *p, uint32_t v) { if (v > *p) *p = v; }
#define FLS_BLOCK_SYNC() ((void)0)
#define FLS_GROUP_SYNC(nthreads) ((void)(nthreads))
extern int fls_emul_force_any; /* tests: make every vote succeed, i.e. every lane takes the most general variant */
#define FLS_ANY(x) ((x) || fls_emul_force_any) /* warp vote among the converged lanes: any variant it selects is correct for every lane */
// Exclusive prefix sum of a FLS_VAR(uint32_t, n) across the warp's lanes; `total` (warp-uniform) receives the sum.
// Use OUTSIDE FLS_LANES regions.
#define FLS_EXCL_SCAN(n, total)                                  \
  do {                                                           \
    uint32_t a_ = 0;                                             \
    for (uint32_t t_ = 0; t_ < 32u; ++t_) { const uint32_t x_ = n##_l[t_]; n##_l[t_] = a_; a_ += x_; } \
    (total) = a_;                                                \
  } while (0)
#else
#define FLS_DEV __device__ __forceinline__
#define FLS_MEM __device__ __forceinline__
// One pass on the device (t = lane id) followed by a warp barrier — the barrier is what a region boundary
// means: shared-memory (and same-address global) writes of the region are visible to the next region.
#define FLS_LANES(t) for (uint32_t t = threadIdx.x & 31u, fls_once_ = 1u; fls_once_; fls_once_ = 0u, __syncwarp())
#define FLS_VAR(T, n) T n##_l
#define FLS_AT(n) n##_l
#define FLS_BCAST(n, lane) __shfl_sync(0xFFFFFFFFu, n##_l, lane)
#define FLS_VAR2(T, n, K) T n##_l[K]
#define FLS_AT2(n, j) n##_l[j]
typedef uint4 fls_u4;
typedef ulonglong2 fls_ul2;
template <typename T> FLS_DEV T fls_ldg(const T *p) { return __ldg(p); }
// load through L2 only: data this warp has just written with streaming stores
template <typename T> FLS_DEV T fls_ldcg(const T *p) { return __ldcg(p); }
FLS_DEV uint32_t fls_shf_r(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_r(lo, hi, s); }
FLS_DEV uint32_t fls_shf_l(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_l(lo, hi, s); }
FLS_DEV uint32_t fls_shf_rc(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_rc(lo, hi, s); }
FLS_DEV uint32_t fls_shl_c(uint32_t x, uint32_t n) { return __funnelshift_lc(0u, x, n); }  // x << min(n, 32)
FLS_DEV uint32_t fls_prmt(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }
// streaming 128-bit store: decoded columns are written once and not re-read by this kernel
template <typename V> FLS_DEV void fls_st128(void *p, const V &v) { __stcs((V *)p, v); }
template <typename V> FLS_DEV V fls_lds128(const void *p) { return *(const V *)p; }
typedef uint2 fls_u2;
FLS_DEV fls_u2 fls_lds64(const void *p) { return *(const uint2 *)p; }
FLS_DEV void fls_sts64(void *p, uint32_t lo, uint32_t hi) { *(uint2 *)p = make_uint2(lo, hi); }
FLS_DEV void fls_st64(void *p, uint32_t lo, uint32_t hi) { __stcs((uint2 *)p, make_uint2(lo, hi)); }  // 8-byte aligned
FLS_DEV uint32_t fls_min(uint32_t a, uint32_t b) { return min(a, b); }
FLS_DEV uint32_t fls_max(uint32_t a, uint32_t b) { return max(a, b); }
FLS_DEV void fls_smem_max(uint32_t *p, uint32_t v) { atomicMax(p, v); }
#define FLS_BLOCK_SYNC() __syncthreads()
// barrier of a group of `nthreads` threads that all call it: one warp (32) or the whole block
#define FLS_GROUP_SYNC(nthreads) do { if ((nthreads) == 32u) __syncwarp(); else __syncthreads(); } while (0)
#define FLS_ANY(x) (__any_sync(__activemask(), (x)) != 0)
#define FLS_EXCL_SCAN(n, total)                                                   \
  do {                                                                            \
    const uint32_t x_ = n##_l;                                                    \
    uint32_t i_ = x_;                                                             \
    _Pragma("unroll") for (uint32_t d_ = 1; d_ < 32u; d_ <<= 1) {                 \
      const uint32_t y_ = __shfl_up_sync(0xFFFFFFFFu, i_, d_);                    \
      if ((threadIdx.x & 31u) >= d_) i_ += y_;                                    \
    }                                                                             \
    (total) = __shfl_sync(0xFFFFFFFFu, i_, 31);                                   \
    n##_l = i_ - x_;                                                              \
  } while (0)
#endif

#define FLS_ORDER_AT(i) ((0x73516240u >> (4u * (i))) & 0xFu) /* FL_ORDER 0,4,2,6,1,5,3,7 packed in nibbles */

// ---------------------------------------------------------------------------------------------------------
// The warp's staging slot + mbarrier.
// ---------------------------------------------------------------------------------------------------------
#define FLS_NO_VECTOR 0xFFFFFFFFu
struct FlsWarp {
  uint8_t *sm;  // this warp's shared-memory scratch, 128-byte aligned
  // pipelined kernels: vector whose payload has already been staged (or is in flight) and where
  uint32_t pf_v = FLS_NO_VECTOR, pf_off = 0;
#ifndef FLS_HOST_EMUL
  uint32_t bar;     // shared-space address of the warp's mbarrier
  uint32_t parity;  // phase parity the next wait() waits for

  __device__ __forceinline__ void init(uint8_t *slot, uint64_t *mbar) {
    sm = slot;
    bar = (uint32_t)__cvta_generic_to_shared(mbar);
    parity = 0;
    if ((threadIdx.x & 31u) == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
  }
  // Asynchronously copy `bytes` (multiple of 16; both addresses 16-byte aligned) from global memory to sm+off.
  // Warp-uniform arguments; call from all lanes, outside FLS_LANES regions.
  __device__ __forceinline__ void stage(uint32_t off, const uint8_t *src, uint32_t bytes) {
    if ((threadIdx.x & 31u) == 0 && bytes) {
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(sm + off);
      asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
      asm volatile(
          "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
          "l"(src), "r"(bytes), "r"(bar)
          : "memory");
    }
  }
  // Bulk copy OUT: `bytes` (multiple of 16) from sm+off to global memory at `dst` (both 16-byte aligned), issued by
  // lane 0 as one asynchronous copy that does not pass through the load/store pipe.  Call from all lanes after
  // the region that wrote the bytes; store_wait() before the bytes are overwritten (and before the block exits).
  __device__ __forceinline__ void store_out(uint8_t *dst, uint32_t off, uint32_t bytes) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the lanes' shared-memory writes -> async proxy
    __syncwarp();
    if ((threadIdx.x & 31u) == 0 && bytes) {
      const uint32_t src = (uint32_t)__cvta_generic_to_shared(sm + off);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  __device__ __forceinline__ void store_wait() {
    if ((threadIdx.x & 31u) == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
  }
  // Order earlier generic-proxy accesses to the slot before the next bulk copy (async proxy) overwrites it.
  __device__ __forceinline__ void fence_async() {
    __syncwarp();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  // Wait until every staged byte has landed.
  __device__ __forceinline__ void wait() {
    if ((threadIdx.x & 31u) == 0) {
      asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(bar) : "memory");
    }
    uint32_t done = 0;
    while (!done) {
      asm volatile(
          "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
          : "=r"(done)
          : "r"(bar), "r"(parity)
          : "memory");
    }
    parity ^= 1u;
  }
#else
  void init(uint8_t *slot, uint64_t *) { sm = slot; }
  void stage(uint32_t off, const uint8_t *src, uint32_t bytes) { memcpy(sm + off, src, bytes); }
  void wait() {}
  void fence_async() {}
  void store_out(uint8_t *dst, uint32_t off, uint32_t bytes) { memcpy(dst, sm + off, bytes); }
  void store_wait() {}
#endif
};

// ---------------------------------------------------------------------------------------------------------
// Generic single-value extract: value (row, lane) of a packed vector of T-bit lanes held in (shared) memory.
//   lane stream is LSB-first; word k of lane l at sm[k*L + l]; value r occupies stream bits [rW,(r+1)W).
// W in [1, 8*sizeof(T)].  Used by the schemes whose access order is not row-major (DELTA, RLE, ALP-RD left).
// ---------------------------------------------------------------------------------------------------------
template <typename T> FLS_DEV T fls_extract(const T *sm, uint32_t W, uint32_t row, uint32_t lane) {
  constexpr uint32_t TB = sizeof(T) * 8u, L = 1024u / TB;
  const uint32_t bit = row * W, k = bit / TB, s = bit % TB;
  if (sizeof(T) >= 4) {
    // branch-free: the word holding the value's first bit and the next one (clamped: it is only used when the value
    // straddles), joined by a funnel shift
    const T lo = sm[k * L + lane], hi = sm[fls_min(k + 1u, W - 1u) * L + lane];
    if (sizeof(T) == 4) return (T)(fls_shf_r((uint32_t)lo, (uint32_t)hi, s) & (W >= 32u ? 0xFFFFFFFFu : ((1u << W) - 1u)));
    const uint64_t x = ((uint64_t)lo >> s) | (((uint64_t)hi << 1) << (63u - s));  // hi << (64 - s), 0 for s == 0
    return (T)(W < 64u ? x & ((1ull << W) - 1ull) : x);
  }
  uint64_t x = (uint64_t)sm[k * L + lane] >> s;
  if (s + W > TB) x |= (uint64_t)sm[(k + 1u) * L + lane] << (TB - s);
  if (W < 64u) x &= (1ull << W) - 1ull;
  return (T)x;
}

// ---------------------------------------------------------------------------------------------------------
// Row-major vectorised unpack: thread (q = t&7, g = t>>3) produces the 16 output bytes [16q, 16q+16) of a row.
//   T=64: lanes 2q,2q+1          T=32: lanes 4q..4q+3        T=16: lanes 8q..8q+7       T=8: lanes 16q..16q+15
// `sm` is the staged payload (W rows of 128 bytes).  W in [1, T].  base is added with wrapping arithmetic.
// ---------------------------------------------------------------------------------------------------------
FLS_DEV fls_u4 fls_unpack32_row(const uint8_t *sm, uint32_t W, uint32_t row, uint32_t q, uint32_t base) {
  const uint32_t bit = row * W, k = bit >> 5, s = bit & 31u;
  const fls_u4 a = fls_lds128<fls_u4>(sm + k * 128u + q * 16u);
  const fls_u4 b = fls_lds128<fls_u4>(sm + fls_min(k + 1u, W - 1u) * 128u + q * 16u);
  const uint32_t mask = W >= 32u ? 0xFFFFFFFFu : ((1u << W) - 1u);
  fls_u4 o;
  o.x = (fls_shf_r(a.x, b.x, s) & mask) + base;
  o.y = (fls_shf_r(a.y, b.y, s) & mask) + base;
  o.z = (fls_shf_r(a.z, b.z, s) & mask) + base;
  o.w = (fls_shf_r(a.w, b.w, s) & mask) + base;
  return o;
}

FLS_DEV fls_ul2 fls_unpack64_row(const uint8_t *sm, uint32_t W, uint32_t row, uint32_t q, uint64_t base) {
  const uint32_t bit = row * W, k = bit >> 6, s = bit & 63u;
  const fls_ul2 a = fls_lds128<fls_ul2>(sm + k * 128u + q * 16u);
  const fls_ul2 b = fls_lds128<fls_ul2>(sm + fls_min(k + 1u, W - 1u) * 128u + q * 16u);
  const uint64_t mask = W >= 64u ? ~0ull : ((1ull << W) - 1ull);
  fls_ul2 o;  // (hi << 1) << (63 - s) == hi << (64 - s) for s in 1..63 and 0 for s == 0
  o.x = (((a.x >> s) | ((b.x << 1) << (63u - s))) & mask) + base;
  o.y = (((a.y >> s) | ((b.y << 1) << (63u - s))) & mask) + base;
  return o;
}

// the same with the straddle row clamped to `klast` (the last word row the caller staged) instead of W - 1
FLS_DEV fls_ul2 fls_unpack64_rows(const uint8_t *sm, uint32_t W, uint32_t row, uint32_t q, uint64_t base, uint32_t klast) {
  const uint32_t bit = row * W, k = bit >> 6, s = bit & 63u;
  const fls_ul2 a = fls_lds128<fls_ul2>(sm + k * 128u + q * 16u);
  const fls_ul2 b = fls_lds128<fls_ul2>(sm + fls_min(k + 1u, klast) * 128u + q * 16u);
  const uint64_t mask = W >= 64u ? ~0ull : ((1ull << W) - 1ull);
  fls_ul2 o;
  o.x = (((a.x >> s) | ((b.x << 1) << (63u - s))) & mask) + base;
  o.y = (((a.y >> s) | ((b.y << 1) << (63u - s))) & mask) + base;
  return o;
}

// two u16 lanes packed in each 32-bit word
FLS_DEV uint32_t fls_unpack16_word(uint32_t a, uint32_t b, uint32_t s, uint32_t mask, uint32_t base) {
  const uint32_t lo = ((fls_prmt(a, b, 0x5410u) >> s) & mask) + base;
  const uint32_t hi = ((fls_prmt(a, b, 0x7632u) >> s) & mask) + base;
  return (lo & 0xFFFFu) | (hi << 16);
}
FLS_DEV fls_u4 fls_unpack16_row(const uint8_t *sm, uint32_t W, uint32_t row, uint32_t q, uint32_t base) {
  const uint32_t bit = row * W, k = bit >> 4, s = bit & 15u;
  const fls_u4 a = fls_lds128<fls_u4>(sm + k * 128u + q * 16u);
  const fls_u4 b = fls_lds128<fls_u4>(sm + fls_min(k + 1u, W - 1u) * 128u + q * 16u);
  const uint32_t mask = (1u << W) - 1u;  // W <= 16
  fls_u4 o;
  o.x = fls_unpack16_word(a.x, b.x, s, mask, base);
  o.y = fls_unpack16_word(a.y, b.y, s, mask, base);
  o.z = fls_unpack16_word(a.z, b.z, s, mask, base);
  o.w = fls_unpack16_word(a.w, b.w, s, mask, base);
  return o;
}

// four u8 lanes packed in each 32-bit word
FLS_DEV uint32_t fls_unpack8_word(uint32_t a, uint32_t b, uint32_t s, uint32_t mask, uint32_t base) {
  uint32_t o = 0;
#pragma unroll
  for (uint32_t i = 0; i < 4u; ++i) {
    const uint32_t pair = fls_prmt(a, b, i | ((4u + i) << 4)) & 0xFFFFu;  // {byte i of a, byte i of b}
    o |= ((((pair >> s) & mask) + base) & 0xFFu) << (8u * i);
  }
  return o;
}
FLS_DEV fls_u4 fls_unpack8_row(const uint8_t *sm, uint32_t W, uint32_t row, uint32_t q, uint32_t base) {
  const uint32_t bit = row * W, k = bit >> 3, s = bit & 7u;
  const fls_u4 a = fls_lds128<fls_u4>(sm + k * 128u + q * 16u);
  const fls_u4 b = fls_lds128<fls_u4>(sm + fls_min(k + 1u, W - 1u) * 128u + q * 16u);
  const uint32_t mask = (1u << W) - 1u;  // W <= 8
  fls_u4 o;
  o.x = fls_unpack8_word(a.x, b.x, s, mask, base);
  o.y = fls_unpack8_word(a.y, b.y, s, mask, base);
  o.z = fls_unpack8_word(a.z, b.z, s, mask, base);
  o.w = fls_unpack8_word(a.w, b.w, s, mask, base);
  return o;
}

// ---------------------------------------------------------------------------------------------------------
// Unified transposed layout (App. A.3).  Lane l of a T-bit vector walks chain steps r = 0..T-1; step r lives at
// packed position p = ORDER[r/8]*16 + (r%8)*128 + l, i.e. packed row p / L of the SAME lane, and reconstructs
// original index lane_start(l) + r.
// ---------------------------------------------------------------------------------------------------------
template <uint32_t TB> FLS_DEV uint32_t fls_chain_row(uint32_t r) {
  constexpr uint32_t L = 1024u / TB;
  return (FLS_ORDER_AT(r >> 3) * 16u + (r & 7u) * 128u) / L;  // the lane offset never carries into the row
}
FLS_DEV uint32_t fls_lane_start(uint32_t l) { return 64u * (l & 15u) + 8u * FLS_ORDER_AT(l >> 4); }

// Padded shared tile in ORIGINAL order (one extra element per 32 x 4 bytes / 64 x 8 bytes) so that the
// lane-strided writes of a chain and the row-major reads of the copy-out are both bank-conflict free.
template <typename T> FLS_DEV uint32_t fls_tile_slot(uint32_t e) {
  if (sizeof(T) == 4) return e + (e >> 5);
  if (sizeof(T) == 8) return e + (e >> 6);
  return e;
}
#define FLS_TILE_BYTES(T) ((uint32_t)(sizeof(T) == 8 ? 8320 : sizeof(T) == 4 ? 4224 : 1024 * sizeof(T)))

// copy a tile out to global memory, 4 (or 8) bytes per thread, 128-byte rows
template <typename T> FLS_DEV void fls_tile_copy_out(const uint8_t *tile_bytes, T *out, uint32_t t, uint32_t add = 0) {
  if (sizeof(T) == 8) {
    const uint64_t *tile = (const uint64_t *)tile_bytes;
    uint64_t *o = (uint64_t *)out;
#pragma unroll 8
    for (uint32_t j = 0; j < 32u; ++j) { const uint32_t e = j * 32u + t; o[e] = tile[e + (e >> 6)]; }
  } else if (sizeof(T) == 4) {
    const uint32_t *tile = (const uint32_t *)tile_bytes;
    uint32_t *o = (uint32_t *)out;
#pragma unroll 8
    for (uint32_t j = 0; j < 32u; ++j) o[j * 32u + t] = tile[j * 33u + t] + add;
  } else {
    const uint32_t *tile = (const uint32_t *)tile_bytes;
    uint32_t *o = (uint32_t *)out;
    for (uint32_t j = 0; j < 8u * sizeof(T); ++j) o[j * 32u + t] = tile[j * 32u + t];
  }
}
