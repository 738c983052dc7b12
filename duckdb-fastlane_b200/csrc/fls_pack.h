// fls_pack.h — host-side FastLanes layout primitives used by the ENCODER (fls_writer.cpp).
// Interleaved bit-packing and the unified transposed layout (App. A.3).
// Decoding on the product path is done only by the CUDA kernels (flsgpu_kernels.cuh); the CPU
// decode restatement lives in oracle/ and does not include this file.
#pragma once
#include <stdint.h>
#include <string.h>
#include "fls_format.h"

namespace flsw {

// transposed position of (row r, lane l)           — App. A.3
static inline uint32_t t_index(uint32_t r, uint32_t l) {
  return FLS_ORDER[r >> 3] * 16u + (r & 7u) * 128u + l;
}
// original index held at transposed position i      — App. A.3
static inline uint32_t t_orig(uint32_t i) {
  return (i & 15u) * 64u + FLS_ORDER[(i >> 4) & 7u] * 8u + (i >> 7);
}

template <typename T>
static inline uint32_t bits_needed(T v) {
  uint32_t n = 0;
  while (v) { ++n; v >>= 1; }
  return n;
}

// Interleaved pack of 1024 values (position p = k*L + l holds row k of lane l).
// out: W rows x L words of T; word k of lane l at out[k*L + l]; lane stream is LSB-first.
template <typename T>
static void pack_vec(const T *in, uint32_t W, T *out) {
  constexpr uint32_t TB = sizeof(T) * 8, L = 1024 / TB;
  if (W == 0) return;
  memset(out, 0, 128 * W);
  const T mask = (W == TB) ? (T)~(T)0 : (T)(((T)1 << W) - 1);
  for (uint32_t l = 0; l < L; ++l) {
    for (uint32_t r = 0; r < TB; ++r) {
      T v = in[r * L + l] & mask;
      uint32_t bit = r * W, k = bit / TB, s = bit % TB;
      out[k * L + l] |= (T)(v << s);
      if (s + W > TB) out[(k + 1) * L + l] |= (T)(v >> (TB - s));
    }
  }
}

// Delta-encode a vector in the unified transposed layout.
//   v[1024]      original order
//   prev_of_lane: for lane l the chain starts from bases[l] := value preceding the lane's first original
//                 (callers choose what "preceding" means; see encoders)
//   d[1024]      deltas at packed positions p = t_index(r,l)   (so unffor(d) feeds the lane chains)
// Lane l visits originals start(l)+r, r = 0..T-1, start(l) = 64*(l%16) + 8*FLS_ORDER[l/16].
template <typename T>
static inline uint32_t lane_start(uint32_t l) {
  return 64u * (l & 15u) + 8u * FLS_ORDER[l >> 4];
}

template <typename T>
static void delta_vec(const T *v, const T *bases, T *d) {
  constexpr uint32_t TB = sizeof(T) * 8, L = 1024 / TB;
  for (uint32_t l = 0; l < L; ++l) {
    T prev = bases[l];
    uint32_t s = lane_start<T>(l);
    for (uint32_t r = 0; r < TB; ++r) {
      T cur = v[s + r];
      d[t_index(r, l)] = (T)(cur - prev);
      prev = cur;
    }
  }
}

static inline uint64_t align_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

}  // namespace flsw
