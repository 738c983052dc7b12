// fls_kernels.cuh — chunk descriptors and the per-vector entry points the kernels (and tests/emul) call.
//
// One launch decodes every chunk of one (scheme, width) family in a batch of row groups: grid = 32 * n_chunks
// blocks of 2 warps; warp w of block b decodes vector 2*(b%32)+w of chunk b/32.  A chunk = one (row group,
// column).
#pragma once
#include "fls_decode.cuh"

// Device-visible descriptor of one column chunk; built on the host by fls_plan.h.
struct FlsChunkDev {
  const uint8_t *seg[FLS_MAX_SEGS];  // device addresses of the chunk's segments
  void *out;      // fixed width: first output value of this row group; strings: &offsets[first row of rg]
  uint8_t *pay;   // strings: batch payload base
  uint32_t n_vec;
  uint32_t aux32;
  uint32_t pay_base;  // strings: this row group's first payload byte, relative to `pay`
  uint32_t scheme;
  uint64_t aux64[2];
};

FLS_DEV fls_vechdr fls_load_hdr(const FlsChunkDev &k, uint32_t seg, uint32_t v) {
  const uint64_t *p = (const uint64_t *)(k.seg[seg]) + 2ull * v;
  const uint64_t a = fls_ldg(p), b = fls_ldg(p + 1);
  fls_vechdr h;
  h.ep = (uint32_t)a; h.meta = (uint32_t)(a >> 32); h.base = b;
  return h;
}

template <typename T> struct VecFFOR {
  static constexpr uint32_t SMEM = DecFFOR<T>::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    const uint32_t bw = FLS_META_BW(h.meta);
    DecFFOR<T>::run(w, k.seg[1] + 16ull * h.ep, bw > sizeof(T) * 8u ? (uint32_t)sizeof(T) * 8u : bw, h.base,
                    (T *)k.out + (size_t)v * FLS_VEC);
  }
};
#ifndef FLS_DELTA64_MIN_BLOCKS
#define FLS_DELTA64_MIN_BLOCKS 16 /* 64 registers: 16 blocks = 32 warps per SM (0.105 -> 0.091 ms on l_orderkey; 18 / 20 blocks spill: 0.096 / 0.105) */
#endif
template <typename T> struct VecDELTA {
  static constexpr uint32_t SMEM = DecDELTA<T>::SMEM, BLOCK_SMEM = 0;
  static constexpr int MIN_BLOCKS = sizeof(T) == 8 ? FLS_DELTA64_MIN_BLOCKS : 1;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    const uint32_t bw = FLS_META_BW(h.meta);
    DecDELTA<T>::run(w, k.seg[1] + 16ull * h.ep, bw > sizeof(T) * 8u ? (uint32_t)sizeof(T) * 8u : bw, h.base,
                     (T *)k.out + (size_t)v * FLS_VEC);
  }
};
template <typename T> struct VecRLE {
  static constexpr uint32_t SMEM = DecRLE<T>::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    DecRLE<T>::run(w, k.seg[1] + 16ull * h.ep, FLS_META_BW(h.meta) ? 1u : 0u, FLS_META_N(h.meta),
                   (T *)k.out + (size_t)v * FLS_VEC);
  }
};
template <typename T> struct VecDICT {
  static constexpr uint32_t SMEM = DecDICT<T>::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    const uint32_t bw = FLS_META_BW(h.meta);
    DecDICT<T>::run(w, k.seg[1] + 16ull * h.ep, bw > 32u ? 32u : bw, h.base, (const T *)k.seg[2], k.aux32,
                    (T *)k.out + (size_t)v * FLS_VEC);
  }
};
// FFOR + exception list (SLPATCH; FREQUENCY when bw = 0): FFOR, then the listed positions take their own values
template <typename T> struct VecFFORPATCH {
  static constexpr uint32_t SMEM = DecFFOR<T>::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    uint32_t bw = FLS_META_BW(h.meta);
    if (bw > sizeof(T) * 8u) bw = (uint32_t)sizeof(T) * 8u;
    const uint32_t n_exc = FLS_META_N(h.meta);
    const uint8_t *in = k.seg[1] + 16ull * h.ep;
    T *out = (T *)k.out + (size_t)v * FLS_VEC;
    DecFFOR<T>::run(w, in, bw, h.base, out);
    if (n_exc) {  // patches land after the region barrier of the decode, so they win over the unpacked values
      const T *xv = (const T *)(in + 128ull * bw);
      const uint16_t *xp = (const uint16_t *)(in + 128ull * bw + (((uint64_t)n_exc * sizeof(T) + 15u) & ~15ull));
      FLS_LANES(t) {
        for (uint32_t i = t; i < n_exc; i += 32u) out[fls_ldg(xp + i) & 1023u] = fls_ldg(xv + i);
      }
    }
  }
};
// CROSS_RLE: seg[0] = run values (aux32 x T), seg[1] = cumulative run ends u32 (padded rows of the row group); the
// run of position p is the first one whose end is > p (binary search; the ends are L1 / L2 resident)
template <typename T> struct VecCROSSRLE {
  static constexpr uint32_t SMEM = 16, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &, const uint8_t *) {
    const T *vals = (const T *)k.seg[0];
    const uint32_t *ends = (const uint32_t *)k.seg[1];
    const uint32_t n = k.aux32;
    T *out = (T *)k.out + (size_t)v * FLS_VEC;
    FLS_LANES(t) {
      // lane t owns the 32 consecutive positions 32 t .. 32 t + 31: one search, then a walk along the runs
      const uint32_t p0 = v * FLS_VEC + 32u * t;
      uint32_t lo = 0, hi = n ? n - 1u : 0u;
      while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (fls_ldg(ends + mid) > p0) hi = mid; else lo = mid + 1u;
      }
      uint32_t r = lo, end = n ? fls_ldg(ends + r) : 0u;
      T val = n ? fls_ldg(vals + r) : (T)0;
      for (uint32_t j = 0; j < 32u; ++j) {
        while (p0 + j >= end && r + 1u < n) { ++r; end = fls_ldg(ends + r); val = fls_ldg(vals + r); }
        out[32u * t + j] = val;
      }
    }
  }
};
struct VecALP {
  static constexpr uint32_t SMEM = DecALP::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    const uint32_t bw = FLS_META_BW(h.meta);
    DecALP::run(w, k.seg[1] + 16ull * h.ep, bw > 64u ? 64u : bw, FLS_META_E(h.meta), FLS_META_F(h.meta),
                FLS_META_N(h.meta), h.base, (double *)k.out + (size_t)v * FLS_VEC);
  }
};
struct VecALP32 {
  static constexpr uint32_t SMEM = DecALP32::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    const uint32_t bw = FLS_META_BW(h.meta);
    DecALP32::run(w, k.seg[1] + 16ull * h.ep, bw > 32u ? 32u : bw, FLS_META_E(h.meta), FLS_META_F(h.meta),
                  FLS_META_N(h.meta), (int64_t)h.base, (float *)k.out + (size_t)v * FLS_VEC);
  }
};
struct VecALPRD {
  static constexpr uint32_t SMEM = DecALPRD::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    uint32_t rbw = k.aux32 & 0xFFu, lbw = (k.aux32 >> 8) & 0xFFu;
    if (rbw > 63u) rbw = 63u;
    if (lbw > 16u) lbw = 16u;
    DecALPRD::run(w, k.seg[1] + 16ull * h.ep, rbw, lbw, FLS_META_N(h.meta), k.aux64[0], k.aux64[1],
                  (uint64_t *)k.out + (size_t)v * FLS_VEC);
  }
};

// 2 warps per block: measured 6 % faster than 4 on ALP and 2 % on FFOR u64 (more, smaller blocks fill the SMs'
// shared memory more evenly); 1 is no better and costs FFOR u32 its occupancy (32 resident blocks per SM)
#ifndef FLS_WARPS_PER_BLOCK
#define FLS_WARPS_PER_BLOCK 2
#endif
#define FLS_BLOCKS_PER_CHUNK (FLS_RG_VECS / FLS_WARPS_PER_BLOCK)

#ifndef FLS_HOST_EMUL
// Shared memory layout of a block: [warp slots | block-shared (V::BLOCK_SMEM) | one mbarrier per warp]
template <class V> constexpr uint32_t fls_block_smem() {
  return FLS_WARPS_PER_BLOCK * V::SMEM + V::BLOCK_SMEM + FLS_WARPS_PER_BLOCK * 8u;
}

// V::MIN_BLOCKS (optional): resident blocks per SM the register allocation must allow
template <class V, class = void> struct FlsMinBlocks { static constexpr int value = 1; };
template <class V> struct FlsMinBlocks<V, decltype((void)V::MIN_BLOCKS)> { static constexpr int value = V::MIN_BLOCKS; };
template <class V>
__global__ void __launch_bounds__(FLS_WARPS_PER_BLOCK * 32, FlsMinBlocks<V>::value) fls_vec_kernel(const FlsChunkDev *__restrict__ chunks) {
  extern __shared__ __align__(128) uint8_t fls_smem[];
  const FlsChunkDev &k = chunks[blockIdx.x / FLS_BLOCKS_PER_CHUNK];
  const uint32_t warp = threadIdx.x >> 5, bx = blockIdx.x % FLS_BLOCKS_PER_CHUNK;
  const uint32_t v = bx * FLS_WARPS_PER_BLOCK + warp;
  uint8_t *bsm = fls_smem + FLS_WARPS_PER_BLOCK * V::SMEM;
  if constexpr (V::BLOCK_SMEM != 0) {
    if (bx * FLS_WARPS_PER_BLOCK >= k.n_vec) return;
    V::block_init(k, bsm, threadIdx.x, FLS_WARPS_PER_BLOCK * 32);
    __syncthreads();
  }
  if (v >= k.n_vec) return;
  FlsWarp w;
  w.init(fls_smem + warp * V::SMEM, (uint64_t *)(bsm + V::BLOCK_SMEM) + warp);
  V::run(k, v, w, bsm);
}

// String columns with a per-chunk lookup table (DICT, FSST): a PERSISTENT grid of (at most) one block per SM.
// The launch's vectors are numbered chunk-major (FLS_RG_VECS slots per chunk) and block b owns the contiguous
// slot range [S*b/G, S*(b+1)/G): every block gets the same number of vectors (+-1) whatever the number of
// chunks, so there is no wave quantisation.  A block's range is a sequence of SEGMENTS (chunk, vector range);
// a segment needs its chunk's lookup table (V::block_init).  The block keeps V::NBUF table buffers; segment s
// uses buffer s % NBUF.  Warps take vectors from the segment's shared counter and never meet at a block-wide
// barrier after the prologue: the warp that retires LAST from segment s (an atomic count) is by construction
// the only one still touching that buffer, so it builds the table of segment s + NBUF in place and publishes
// it (ready[buf] = s + NBUF + 1); warps arriving at a segment spin on its ready word.  With NBUF = 2 the table
// of the next segment is always there already and nobody waits; with NBUF = 1 (a table too large to double)
// the others wait for the last warp's vector + one table build per segment.
// Each table buffer is followed by a copy of its chunk's descriptor: the decoders read descriptor fields (segment
// addresses, output addresses) for every vector, and a shared-memory load is an order of magnitude closer than the
// descriptor array in global memory.
// Shared memory: [warp slots | constant area (V::CONST_SMEM) | tables (V::TABLES_SMEM; buffer b at V::tab_off(b)) |
//                 NBUF descriptors of 128 B | mbarriers | control]
#define FLS_DESC_SMEM 128u
static_assert(sizeof(FlsChunkDev) <= FLS_DESC_SMEM && sizeof(FlsChunkDev) % 4u == 0, "descriptor copy");
template <class V> constexpr uint32_t fls_chunk_smem() {
  return V::WARPS * V::SMEM + V::CONST_SMEM + V::TABLES_SMEM + V::NBUF * FLS_DESC_SMEM + V::WARPS * 8u + 64u;
}
template <class V>
__global__ void __launch_bounds__(V::WARPS * 32, 1) fls_chunk_kernel(const FlsChunkDev *__restrict__ chunks, uint32_t n_chunks) {
  extern __shared__ __align__(128) uint8_t fls_smem[];
  constexpr uint32_t NB = V::NBUF;
  const uint32_t S = n_chunks * FLS_RG_VECS;
  const uint32_t G0 = (uint32_t)((uint64_t)S * blockIdx.x / gridDim.x), G1 = (uint32_t)((uint64_t)S * (blockIdx.x + 1u) / gridDim.x);
  if (G0 >= G1) return;
  const uint32_t c0 = G0 / FLS_RG_VECS, nseg = (G1 - 1u) / FLS_RG_VECS - c0 + 1u;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  uint8_t *csm = fls_smem + V::WARPS * V::SMEM;
  uint8_t *tabs = csm + V::CONST_SMEM;
  uint8_t *descs = tabs + V::TABLES_SMEM;
  uint64_t *mbar = (uint64_t *)(descs + NB * FLS_DESC_SMEM);
  auto copy_desc = [&](uint32_t buf, uint32_t chunk, uint32_t tid) {  // tid: 0 .. 31 at least
    if (tid < sizeof(FlsChunkDev) / 4u)
      ((uint32_t *)(descs + buf * FLS_DESC_SMEM))[tid] = __ldg((const uint32_t *)(chunks + chunk) + tid);
  };
  uint32_t *next = (uint32_t *)(mbar + V::WARPS);  // next[NB]: vector counter of the segment in each buffer
  uint32_t *done = next + NB;                      // done[NB]: warps that have retired from it
  volatile uint32_t *ready = done + NB;            // ready[NB]: 1 + the segment whose table the buffer holds
  auto seg_lo = [&](uint32_t s) { const uint32_t b0 = (c0 + s) * FLS_RG_VECS; return max(G0, b0) - b0; };
  auto seg_hi = [&](uint32_t s) { const uint32_t b0 = (c0 + s) * FLS_RG_VECS; return min(min(G1, b0 + FLS_RG_VECS) - b0, chunks[c0 + s].n_vec); };
  // prologue: the first NB tables (one table: by the whole block; several: one warp each, so that their
  // global-memory round trips overlap)
  V::const_init(csm, threadIdx.x, V::WARPS * 32u);
  if (NB == 1u) V::block_init(chunks[c0], tabs + V::tab_off(0), threadIdx.x, V::WARPS * 32u);
  else if (warp < NB && warp < nseg) V::block_init(chunks[c0 + warp], tabs + V::tab_off(warp), lane, 32u);  // side by side
  if (warp < NB && warp < nseg) copy_desc(warp, c0 + warp, lane);
  if (threadIdx.x < NB) {
    next[threadIdx.x] = threadIdx.x < nseg ? seg_lo(threadIdx.x) : 0u;
    done[threadIdx.x] = 0u;
    ready[threadIdx.x] = threadIdx.x + 1u;
  }
  __syncthreads();
  FlsWarp w;
  w.init(fls_smem + warp * V::SMEM, mbar + warp);
  for (uint32_t s = 0; s < nseg; ++s) {
    const uint32_t b = s % NB;
    const uint8_t *bsm = tabs + V::tab_off(b);
    const FlsChunkDev &k = *(const FlsChunkDev *)(descs + b * FLS_DESC_SMEM);  // the shared-memory copy
    const uint32_t v1 = seg_hi(s);
    if (s >= NB) {
      while (ready[b] != s + 1u) __nanosleep(32);
      __threadfence_block();
      __syncwarp();
    }
    auto grab = [&]() {
      uint32_t x = 0;
      if (lane == 0) x = atomicAdd(next + b, 1u);
      return __shfl_sync(0xFFFFFFFFu, x, 0);
    };
    for (uint32_t v = grab(); v < v1;) {  // one vector ahead, so that run() can prefetch the next one's payload
      const uint32_t vn = grab();
      V::run(k, v, vn < v1 ? vn : FLS_NO_VECTOR, w, bsm, csm);
      v = vn;
    }
    if (s + NB < nseg) {  // block-uniform: somebody has to build the table of segment s + NB in this buffer
      __syncwarp();
      uint32_t last = 0;
      if (lane == 0) {
        __threadfence_block();
        last = atomicAdd(done + b, 1u) == V::WARPS - 1u ? 1u : 0u;
      }
      if (__shfl_sync(0xFFFFFFFFu, last, 0)) {
        __threadfence_block();
        V::block_init(chunks[c0 + s + NB], tabs + V::tab_off(b), lane, 32u);
        copy_desc(b, c0 + s + NB, lane);
        if (lane == 0) { next[b] = seg_lo(s + NB); done[b] = 0u; }
        __threadfence_block();
        __syncwarp();
        if (lane == 0) ready[b] = s + NB + 1u;
      }
    }
  }
  w.store_wait();  // a bulk copy out of the warp's slot may still be reading it
}

// UNCOMPRESSED / blob copy: chunk by = blockIdx.y copies aux64[1] bytes seg[0] -> out (both 16-byte aligned).
__global__ void __launch_bounds__(256) fls_copy_kernel(const FlsChunkDev *__restrict__ chunks, uint32_t bpc) {
  const FlsChunkDev &k = chunks[blockIdx.x / bpc];
  const uint32_t bx = blockIdx.x % bpc;
  const uint4 *src = (const uint4 *)k.seg[0];
  uint4 *dst = (uint4 *)k.out;
  const uint64_t n16 = k.aux64[1] >> 4;
  for (uint64_t i = bx * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)bpc * blockDim.x)
    __stcs(dst + i, __ldg(src + i));
}
// validity bitmaps: 128 bytes per vector, copied from the chunk's bitmap segment or set to all ones
__global__ void __launch_bounds__(256) fls_validity_kernel(const FlsChunkDev *__restrict__ chunks) {
  const FlsChunkDev &k = chunks[blockIdx.x];
  const uint4 *src = (const uint4 *)k.seg[0];
  uint4 *dst = (uint4 *)k.out;
  for (uint32_t i = threadIdx.x; i < k.n_vec * 8u; i += blockDim.x)
    dst[i] = src ? __ldg(src + i) : make_uint4(~0u, ~0u, ~0u, ~0u);
}
// CONSTANT: fill n_vec*1024 values of width aux32 bytes with aux64[0].
__global__ void __launch_bounds__(256) fls_fill_kernel(const FlsChunkDev *__restrict__ chunks, uint32_t bpc) {
  const FlsChunkDev &k = chunks[blockIdx.x / bpc];
  const uint32_t bx = blockIdx.x % bpc;
  const uint32_t wb = k.aux32;
  uint64_t pat = k.aux64[0];
  if (wb == 1) pat = (pat & 0xFF) * 0x0101010101010101ull;
  else if (wb == 2) pat = (pat & 0xFFFF) * 0x0001000100010001ull;
  else if (wb == 4) pat = (pat & 0xFFFFFFFFull) * 0x0000000100000001ull;
  uint64_t *dst = (uint64_t *)k.out;
  const uint64_t n8 = (uint64_t)k.n_vec * FLS_VEC * wb / 8;
  for (uint64_t i = bx * (uint64_t)blockDim.x + threadIdx.x; i < n8; i += (uint64_t)bpc * blockDim.x)
    dst[i] = pat;
}
#endif
