// fls_filter.cuh — late materialisation on the device: evaluate a conjunction of `column <op> constant` predicates
// on decoded columns that are still in HBM, build an ordered selection, and compact the projected columns so that
// only surviving rows cross PCIe (the reference leaves filter pushdown switched off,
// /root/reference/src/scanner/scan_fastlanes.cpp:154-155, so there is no reference twin: the semantics are those of
// DuckDB's comparison operators, which the engine re-applies on top).
//
// One warp owns one 1024-row vector everywhere: row 32 j + lane is bit `lane` of mask word j, so a vector's 32 mask
// words are one coalesced 128-byte line and ranks inside a vector are popcounts — no atomics, output order = row order.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define FLS_MAX_PREDS 8

struct FlsPredDev {
  const void *data;         // fixed width: decoded values (padded rows)
  const uint32_t *offsets;  // strings
  const uint8_t *payload;
  const uint32_t *validity; // bit per padded row (null: no NULLs); a NULL value fails every comparison
  const uint8_t *text;      // strings: the constant's bytes (device)
  uint32_t text_len;
  int32_t kind;             // 0 signed int, 1 unsigned int, 2 f32, 3 f64, 4 string
  uint32_t width;           // bytes per value
  uint32_t op;              // enum flsgpu_cmp
  int64_t i64;
  double f64;
};
struct FlsPredSet {
  uint32_t n;
  FlsPredDev p[FLS_MAX_PREDS];
};

namespace fls_filter {

template <typename T> __device__ __forceinline__ bool cmp_ord(T a, T b, uint32_t op) {
  switch (op) {
    case 0: return a == b;
    case 1: return a != b;
    case 2: return a < b;
    case 3: return a <= b;
    case 4: return a > b;
    default: return a >= b;
  }
}
// DuckDB's total order on floating point: NaN equals NaN and is greater than every other value
template <typename F> __device__ __forceinline__ bool cmp_fp(F a, F b, uint32_t op) {
  const bool an = a != a, bn = b != b;
  const bool eq = (an && bn) || a == b;
  const bool lt = !an && (bn || a < b);
  switch (op) {
    case 0: return eq;
    case 1: return !eq;
    case 2: return lt;
    case 3: return lt || eq;
    case 4: return !(lt || eq);
    default: return !lt;
  }
}
__device__ __forceinline__ bool eval(const FlsPredDev &p, uint64_t row) {
  if (p.validity && !((p.validity[row >> 5] >> (row & 31u)) & 1u)) return false;
  switch (p.kind) {
    case 0: {
      int64_t v;
      if (p.width == 1) v = ((const int8_t *)p.data)[row];
      else if (p.width == 2) v = ((const int16_t *)p.data)[row];
      else if (p.width == 4) v = ((const int32_t *)p.data)[row];
      else v = ((const int64_t *)p.data)[row];
      return cmp_ord<int64_t>(v, p.i64, p.op);
    }
    case 1: {
      uint64_t v;
      if (p.width == 1) v = ((const uint8_t *)p.data)[row];
      else if (p.width == 2) v = ((const uint16_t *)p.data)[row];
      else if (p.width == 4) v = ((const uint32_t *)p.data)[row];
      else v = ((const uint64_t *)p.data)[row];
      return cmp_ord<uint64_t>(v, (uint64_t)p.i64, p.op);
    }
    case 2: return cmp_fp<float>(((const float *)p.data)[row], (float)p.f64, p.op);
    case 3: return cmp_fp<double>(((const double *)p.data)[row], p.f64, p.op);
    default: {
      const uint32_t a = p.offsets[row], b = p.offsets[row + 1];
      bool eq = b - a == p.text_len;
      for (uint32_t i = 0; eq && i < p.text_len; ++i) eq = p.payload[a + i] == p.text[i];
      return p.op == 0 ? eq : !eq;
    }
  }
}

// mask[v * 32 + j] bit l = row v * 1024 + 32 j + l passes every predicate (and is a real row: < valid[v]);
// vcnt[v] = rows of vector v that pass
__global__ void __launch_bounds__(256) mask_kernel(FlsPredSet ps, const uint32_t *__restrict__ valid, uint32_t n_vec,
                                                   uint32_t *__restrict__ mask, uint32_t *__restrict__ vcnt) {
  const uint32_t v = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (v >= n_vec) return;
  const uint32_t nv = valid[v];
  uint32_t mine = 0, cnt = 0;
  for (uint32_t j = 0; j < 32u; ++j) {
    const uint32_t r = 32u * j + lane;
    bool ok = r < nv;
    const uint64_t row = (uint64_t)v * 1024u + r;
    for (uint32_t k = 0; ok && k < ps.n; ++k) ok = eval(ps.p[k], row);
    const uint32_t w = __ballot_sync(0xFFFFFFFFu, ok);
    if (lane == j) mine = w;
    cnt += __popc(w);
  }
  mask[(size_t)v * 32u + lane] = mine;
  if (lane == 0) vcnt[v] = cnt;
}

// exclusive scan of cnt[0..n) into off[0..n) (one block); total[0] += sum, off[] starts at the old total[0]
__global__ void __launch_bounds__(1024) scan_kernel(const uint32_t *__restrict__ cnt, uint32_t n, uint64_t *__restrict__ off,
                                                    uint64_t *total) {
  __shared__ uint64_t part[1024];
  const uint32_t tid = threadIdx.x, per = (n + 1023u) / 1024u;
  const uint32_t a = min(n, tid * per), b = min(n, a + per);
  uint64_t s = 0;
  for (uint32_t i = a; i < b; ++i) s += cnt[i];
  part[tid] = s;
  __syncthreads();
  for (uint32_t d = 1; d < 1024u; d <<= 1) {
    const uint64_t x = tid >= d ? part[tid - d] : 0;
    __syncthreads();
    part[tid] += x;
    __syncthreads();
  }
  uint64_t run = total[0] + (tid ? part[tid - 1] : 0);
  for (uint32_t i = a; i < b; ++i) { off[i] = run; run += cnt[i]; }
  __syncthreads();
  if (tid == 1023u) total[0] += part[1023];
}

template <typename T>
__global__ void __launch_bounds__(256) compact_fixed_kernel(const T *__restrict__ in, const uint32_t *__restrict__ mask,
                                                            const uint64_t *__restrict__ voff, uint32_t n_vec, T *__restrict__ out) {
  const uint32_t v = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (v >= n_vec) return;
  const uint32_t mine = mask[(size_t)v * 32u + lane];
  uint64_t dst = voff[v];
  for (uint32_t j = 0; j < 32u; ++j) {
    const uint32_t w = __shfl_sync(0xFFFFFFFFu, mine, j);
    if (!w) continue;  // warp-uniform
    if ((w >> lane) & 1u) out[dst + __popc(w & ((1u << lane) - 1u))] = in[(size_t)v * 1024u + 32u * j + lane];
    dst += __popc(w);
  }
}

// validity bits of the selected rows (out zeroed by the caller)
__global__ void __launch_bounds__(256) compact_validity_kernel(const uint32_t *__restrict__ valid, const uint32_t *__restrict__ mask,
                                                               const uint64_t *__restrict__ voff, uint32_t n_vec,
                                                               unsigned long long *__restrict__ out) {
  const uint32_t v = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (v >= n_vec) return;
  const uint32_t mine = mask[(size_t)v * 32u + lane], vmine = valid[(size_t)v * 32u + lane];
  uint64_t dst = voff[v];
  for (uint32_t j = 0; j < 32u; ++j) {
    const uint32_t w = __shfl_sync(0xFFFFFFFFu, mine, j), vw = __shfl_sync(0xFFFFFFFFu, vmine, j);
    if (!w) continue;
    if ((w >> lane) & (vw >> lane) & 1u) {
      const uint64_t at = dst + __popc(w & ((1u << lane) - 1u));
      atomicOr(out + (at >> 6), 1ull << (at & 63u));
    }
    dst += __popc(w);
  }
}

// logical row id (position in the unpadded range) of every selected row
__global__ void __launch_bounds__(256) compact_rowid_kernel(const uint64_t *__restrict__ row0, const uint32_t *__restrict__ mask,
                                                            const uint64_t *__restrict__ voff, uint32_t n_vec, uint64_t *__restrict__ out) {
  const uint32_t v = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (v >= n_vec) return;
  const uint32_t mine = mask[(size_t)v * 32u + lane];
  uint64_t dst = voff[v];
  const uint64_t r0 = row0[v];
  for (uint32_t j = 0; j < 32u; ++j) {
    const uint32_t w = __shfl_sync(0xFFFFFFFFu, mine, j);
    if (!w) continue;
    if ((w >> lane) & 1u) out[dst + __popc(w & ((1u << lane) - 1u))] = r0 + 32u * j + lane;
    dst += __popc(w);
  }
}

// strings, pass 1: bytes of the vector's selected strings
__global__ void __launch_bounds__(256) str_bytes_kernel(const uint32_t *__restrict__ offs, const uint32_t *__restrict__ mask,
                                                        uint32_t n_vec, uint32_t *__restrict__ vbytes) {
  const uint32_t v = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (v >= n_vec) return;
  const uint32_t mine = mask[(size_t)v * 32u + lane];
  uint32_t sum = 0;
  for (uint32_t j = 0; j < 32u; ++j) {
    const uint32_t w = __shfl_sync(0xFFFFFFFFu, mine, j);
    if (!w) continue;
    const size_t r = (size_t)v * 1024u + 32u * j + lane;
    if ((w >> lane) & 1u) sum += offs[r + 1] - offs[r];
  }
  for (uint32_t d = 16; d; d >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, d);
  if (lane == 0) vbytes[v] = sum;
}

// strings, pass 2: out_offs[k + 1] = end of the k-th selected string, its bytes copied behind the previous one's
__global__ void __launch_bounds__(256) compact_str_kernel(const uint32_t *__restrict__ offs, const uint8_t *__restrict__ pay,
                                                          const uint32_t *__restrict__ mask, const uint64_t *__restrict__ voff,
                                                          const uint64_t *__restrict__ vboff, uint32_t n_vec,
                                                          uint32_t *__restrict__ out_offs, uint8_t *__restrict__ out_pay) {
  const uint32_t v = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (v >= n_vec) return;
  const uint32_t mine = mask[(size_t)v * 32u + lane];
  uint64_t dst = voff[v], bdst = vboff[v];
  for (uint32_t j = 0; j < 32u; ++j) {
    const uint32_t w = __shfl_sync(0xFFFFFFFFu, mine, j);
    if (!w) continue;
    const size_t r = (size_t)v * 1024u + 32u * j + lane;
    const bool on = (w >> lane) & 1u;
    const uint32_t a = on ? offs[r] : 0u, len = on ? offs[r + 1] - a : 0u;
    uint32_t incl = len;  // inclusive scan of the lengths
    for (uint32_t d = 1; d < 32u; d <<= 1) {
      const uint32_t x = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= d) incl += x;
    }
    if (on) {
      const uint64_t at = bdst + incl - len;
      out_offs[dst + __popc(w & ((1u << lane) - 1u)) + 1] = (uint32_t)(at + len);
      for (uint32_t i = 0; i < len; ++i) out_pay[at + i] = pay[a + i];
    }
    dst += __popc(w);
    bdst += __shfl_sync(0xFFFFFFFFu, incl, 31);
  }
}

}  // namespace fls_filter
