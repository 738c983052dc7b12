// fls_decode.cuh — per-vector decoders for fixed-width columns (one warp = one 1024-value vector).
//
// Each decoder is a struct with
//   SMEM            bytes of per-warp shared scratch it needs (staging slot + tiles)
//   run(w, ...)     called by ALL lanes of the warp with warp-uniform arguments; per-lane work sits inside
//                   FLS_LANES(t) { ... } regions (see fls_dev.cuh).
//
// Restated FastLanes primitives (upstream source absent):
//   DecFFOR  = unffor                      DecDELTA = unffor + rsum + untranspose (App. A.3)
//   DecRLE   = rle_map                     DecDICT  = dict gather
//   DecALP   = alp decode + patch (twin: duckdb .../alp/algorithm/alp.hpp:129-134,376-404)
//   DecALPRD = alp_rd decode + patch (twin: duckdb .../alprd/algorithm/alprd.hpp:206-233)
#pragma once
#include "fls_dev.cuh"

template <typename T> struct FlsRowGeom {            // thread (q,g): rows g*RPT .. g*RPT+RPT-1, bytes [16q,16q+16)
  static constexpr uint32_t TB = sizeof(T) * 8u;     // bits per value = rows per vector
  static constexpr uint32_t RPT = TB / 4u;           // rows per thread
};

// broadcast a T-sized pattern into 16 bytes
template <typename T> FLS_DEV fls_u4 fls_splat(uint64_t v) {
  if (sizeof(T) == 1) v = (v & 0xFFull) * 0x0101010101010101ull;
  else if (sizeof(T) == 2) v = (v & 0xFFFFull) * 0x0001000100010001ull;
  else if (sizeof(T) == 4) v = (v & 0xFFFFFFFFull) * 0x0000000100000001ull;
  fls_u4 o;
  o.x = (uint32_t)v; o.y = (uint32_t)(v >> 32); o.z = o.x; o.w = o.y;
  return o;
}

// one row (16 bytes per thread) of a packed T-vector
template <typename T> FLS_DEV fls_u4 fls_unpack_row(const uint8_t *sm, uint32_t W, uint32_t row, uint32_t q, uint64_t base) {
  if (sizeof(T) == 8) {
    const fls_ul2 v = fls_unpack64_row(sm, W, row, q, base);
    fls_u4 o;
    o.x = (uint32_t)v.x; o.y = (uint32_t)(v.x >> 32); o.z = (uint32_t)v.y; o.w = (uint32_t)(v.y >> 32);
    return o;
  } else if (sizeof(T) == 4) {
    return fls_unpack32_row(sm, W, row, q, (uint32_t)base);
  } else if (sizeof(T) == 2) {
    return fls_unpack16_row(sm, W, row, q, (uint32_t)base & 0xFFFFu);
  } else {
    return fls_unpack8_row(sm, W, row, q, (uint32_t)base & 0xFFu);
  }
}

// ===================================================================== FFOR
// payload: [packed 128*bw]
template <typename T> struct DecFFOR {
  static constexpr uint32_t SMEM = 128u * sizeof(T) * 8u;
  FLS_MEM static void run(FlsWarp &w, const uint8_t *in, uint32_t bw, uint64_t base, T *out) {
    constexpr uint32_t RPT = FlsRowGeom<T>::RPT;
    if (bw == 0) {
      const fls_u4 pat = fls_splat<T>(base);
      FLS_LANES(t) {
        const uint32_t q = t & 7u, g = t >> 3;
#pragma unroll 4
        for (uint32_t j = 0; j < RPT; ++j) fls_st128((uint8_t *)out + (g * RPT + j) * 128u + q * 16u, pat);
      }
      return;
    }
    w.stage(0, in, 128u * bw);
    w.wait();
    FLS_LANES(t) {
      const uint32_t q = t & 7u, g = t >> 3;
#pragma unroll 4
      for (uint32_t j = 0; j < RPT; ++j) {
        const uint32_t row = g * RPT + j;
        fls_st128((uint8_t *)out + row * 128u + q * 16u, fls_unpack_row<T>(w.sm, bw, row, q, base));
      }
    }
  }
};

// ===================================================================== DELTA
// payload: [lane bases 128 B][packed deltas 128*bw]; deltas are FFOR'd (wrapping + base).
template <typename T> struct DecDELTA {
  static constexpr uint32_t TB = sizeof(T) * 8u, L = 1024u / TB;
  static constexpr uint32_t STAGE = 128u + 128u * TB;
  // T=64: see run64 (half a vector per slot: 4224 bytes + the 16 half-lane totals)
  static constexpr uint32_t SMEM = sizeof(T) == 8 ? 4224u + 128u : STAGE + FLS_TILE_BYTES(T);

  // Reconstruct into the warp's padded tile (original order).  `add` is added to every value (string offsets).
  FLS_MEM static void to_tile(FlsWarp &w, const uint8_t *in, uint32_t bw, uint64_t base) {
    w.stage(0, in, 128u + 128u * bw);
    w.wait();
    const T *bases = (const T *)w.sm;
    const T *pk = (const T *)(w.sm + 128);
    T *tile = (T *)(w.sm + STAGE);
    {
      constexpr uint32_t LPT = L / 32u;  // lanes per thread: 1 (T=32), 2 (T=16), 4 (T=8)
      FLS_LANES(t) {
        for (uint32_t a = 0; a < LPT; ++a) {
          const uint32_t l = t * LPT + a;
          T acc = bases[l];
          const uint32_t e0 = fls_lane_start(l);
          for (uint32_t r = 0; r < TB; ++r) {
            acc = (T)(acc + (bw ? fls_extract<T>(pk, bw, fls_chain_row<TB>(r), l) : (T)0) + (T)base);
            tile[fls_tile_slot<T>(e0 + r)] = acc;
          }
        }
      }
    }
  }
  // ---- T=64: 16 lanes x 64 steps; thread (l = t&15, h = t>>4) owns steps [32h, 32h+32) = originals 64l+32h .. +31 and
  // holds its 32 deltas in registers; only the 16 half totals cross lanes (xch).  The slot is HALF a vector wide: the
  // staged payload (deltas of up to 16 bits: 128 + 128*16 bytes) and, once every thread holds its deltas, one half of
  // the output at a time — the thread's values 16p .. 16p+15 (p = 0, 1) are the 128-byte output row 4l + 2h + p, so
  // a half is every other output row; its 32 rows sit at 16 m + m/2 (m = 2l + h; the m/2 keeps the sixteen threads
  // of a 64-bit store phase in sixteen different banks).  4.3 KB per warp instead of 8.4 KB: twice the resident warps.
  // Deltas wider than 16 bits (rare: DELTA is picked for narrow deltas) are read from global memory, twice.
  static constexpr uint32_t HALF_TILE = 4224u;  // >= 128 + 128*16 (payload) and >= 8 * (31*16 + 15 + 16) (half tile)
  static constexpr uint32_t XCH = HALF_TILE;    // 16 x u64
  FLS_MEM static uint32_t half_slot(uint32_t m, uint32_t jj) { return 16u * m + (m >> 1) + jj; }

  // Thread (l, h) needs packed rows ORDER[4h + j/8] + 8 (j%8) = {0,4,2,6}[j/8] + 8 (j%8) + h of lane l: the even rows
  // (h = 0) or the odd rows (h = 1) of the lane's bit stream.  W <= FLS_DELTA_FASTW (sorted keys: l_orderkey packs into
  // 5-6 bits): the lane's W 64-bit words are loaded once into registers, shifted right by h W bits (so that both halves
  // extract at the SAME, compile-time bit positions), and every delta is one funnel shift and one mask — against two
  // shared loads, two 64-bit variable shifts and a 64-bit mask in fls_extract.
  template <uint32_t W> FLS_MEM static void narrow_deltas(const uint8_t *lane_words, uint32_t h, uint32_t (&d)[32]) {
    uint32_t wd[2u * W + 1u];
#pragma unroll
    for (uint32_t k = 0; k < W; ++k) {
      const fls_u2 v = fls_lds64(lane_words + 128u * k);
      wd[2u * k] = v.x; wd[2u * k + 1u] = v.y;
    }
    wd[2u * W] = 0u;
    const uint32_t s0 = h * W;  // <= 16: a funnel shift's count
#pragma unroll
    for (uint32_t i = 0; i < 2u * W; ++i) wd[i] = fls_shf_r(wd[i], wd[i + 1u], s0);
    constexpr uint32_t mask = (1u << W) - 1u;
    // ascending rows, so that the words die as the deltas appear (a thread never holds much more than 2W + 1 or 32
    // registers of either); even row 2e is chain step j = 8 * {0,2,1,3}[e % 4] + e / 4
#pragma unroll
    for (uint32_t e = 0; e < 32u; ++e) {
      const uint32_t j = 8u * ((0x3120u >> (4u * (e & 3u))) & 0xFu) + (e >> 2);
      const uint32_t bit = 2u * e * W, i = bit >> 5, s = bit & 31u;
      d[j] = (s + W <= 32u ? wd[i] >> s : fls_shf_r(wd[i], wd[i + 1u], s)) & mask;
    }
  }
  FLS_MEM static void run64(FlsWarp &w, const uint8_t *in, uint32_t bw, uint64_t base, T *out) {
    const bool wide = bw > 16u;  // no register-resident deltas: two generic passes over the payload in global memory
    w.stage(0, in, wide ? 128u : 128u + 128u * bw);
    w.wait();
    const T *bases = (const T *)w.sm;
    const T *pk = (const T *)(in + 128);  // wide only (the half tiles below overwrite the slot while deltas are still needed)
    uint64_t *xch = (uint64_t *)(w.sm + XCH);
    T *tile = (T *)w.sm;  // overwrites the staged payload once every thread holds its deltas
    FLS_VAR2(uint32_t, dl, 32);
    FLS_VAR(uint64_t, accv);
    FLS_LANES(t) {
      const uint32_t l = t & 15u, h = t >> 4;
      FLS_AT(accv) = (uint64_t)bases[l];
      uint64_t tot = 0;
      if (wide) {
#pragma unroll 2
        for (uint32_t j = 0; j < 32u; ++j) tot += (uint64_t)fls_extract<T>(pk, bw, fls_chain_row<TB>(32u * h + j), l);
      } else {
        uint32_t d[32];
        const uint8_t *lw = w.sm + 128u + 8u * l;
        switch (bw) {  // warp-uniform
          case 0:
#pragma unroll
            for (uint32_t j = 0; j < 32u; ++j) d[j] = 0u;
            break;
#define FLS_DELTA_CASE(W) case W: narrow_deltas<W>(lw, h, d); break;
          FLS_DELTA_CASE(1) FLS_DELTA_CASE(2) FLS_DELTA_CASE(3) FLS_DELTA_CASE(4) FLS_DELTA_CASE(5) FLS_DELTA_CASE(6)
          FLS_DELTA_CASE(7) FLS_DELTA_CASE(8) FLS_DELTA_CASE(9) FLS_DELTA_CASE(10) FLS_DELTA_CASE(11) FLS_DELTA_CASE(12)
          FLS_DELTA_CASE(13) FLS_DELTA_CASE(14) FLS_DELTA_CASE(15)
#undef FLS_DELTA_CASE
          default: narrow_deltas<16>(lw, h, d); break;
        }
#pragma unroll
        for (uint32_t j = 0; j < 32u; ++j) { FLS_AT2(dl, j) = d[j]; tot += d[j]; }
      }
      if (h == 0u) xch[l] = tot + 32ull * base;  // the upper half of a lane starts where the lower half ends
    }
    FLS_LANES(t) {
      const uint32_t l = t & 15u, h = t >> 4;
      if (h) FLS_AT(accv) += xch[l];
    }
#pragma unroll
    for (uint32_t p = 0; p < 2u; ++p) {
      FLS_LANES(t) {
        const uint32_t l = t & 15u, h = t >> 4, m = 2u * l + h;
        uint64_t acc = FLS_AT(accv);
        if (wide) {
#pragma unroll 2
          for (uint32_t jj = 0; jj < 16u; ++jj) {
            acc += (uint64_t)fls_extract<T>(pk, bw, fls_chain_row<TB>(32u * h + 16u * p + jj), l) + base;
            tile[half_slot(m, jj)] = (T)acc;
          }
        } else {
#pragma unroll
          for (uint32_t jj = 0; jj < 16u; ++jj) {
            acc += (uint64_t)FLS_AT2(dl, 16u * p + jj) + base;
            tile[half_slot(m, jj)] = (T)acc;
          }
        }
        FLS_AT(accv) = acc;
      }
      FLS_LANES(t) {  // the half's 32 rows leave two at a time: threads 0..15 one 128-byte row, threads 16..31 the next
        const uint32_t jj = t & 15u;
#pragma unroll 8
        for (uint32_t i = 0; i < 16u; ++i) {
          const uint32_t m = 2u * i + (t >> 4);
          out[16u * (2u * m + p) + jj] = tile[half_slot(m, jj)];
        }
      }
    }
  }
  FLS_MEM static void run(FlsWarp &w, const uint8_t *in, uint32_t bw, uint64_t base, T *out) {
    if constexpr (sizeof(T) == 8) {
      run64(w, in, bw, base, out);
    } else {
      to_tile(w, in, bw, base);
      FLS_LANES(t) { fls_tile_copy_out<T>(w.sm + STAGE, out, t); }
    }
  }
};

// ===================================================================== RLE
// payload: [run values n x T, pad16][idx lane bases 64 x u16][idx packed 128*bw, bw in {0,1}]
// The run-index stream is a u16 vector in the transposed layout with 0/1 deltas.  Rare scheme: the index
// stream is staged, run values are gathered straight from global memory (L1-resident, <= 8 KB).
template <typename T> struct DecRLE {
  static constexpr uint32_t STAGE = 256u;
  static constexpr uint32_t SMEM = STAGE + FLS_TILE_BYTES(T);
  FLS_MEM static void run(FlsWarp &w, const uint8_t *in, uint32_t bw, uint32_t n_runs, T *out) {
    const T *runs = (const T *)in;
    const uint8_t *q = in + (((uint64_t)n_runs * sizeof(T) + 15u) & ~(uint64_t)15u);
    w.stage(0, q, 128u + 128u * bw);
    w.wait();
    const uint16_t *bases = (const uint16_t *)w.sm;
    const uint16_t *pk = (const uint16_t *)(w.sm + 128);
    T *tile = (T *)(w.sm + STAGE);
    const uint32_t last_run = n_runs ? n_runs - 1u : 0u;  // a run index past the run values (corrupt stream) reads the last one
    FLS_LANES(t) {
      for (uint32_t a = 0; a < 2u; ++a) {
        const uint32_t l = 2u * t + a;
        uint32_t idx = bases[l];
        const uint32_t e0 = fls_lane_start(l);
        for (uint32_t r = 0; r < 16u; ++r) {
          if (bw) idx += fls_extract<uint16_t>(pk, bw, fls_chain_row<16>(r), l);
          tile[fls_tile_slot<T>(e0 + r)] = fls_ldg(runs + fls_min(idx & 0xFFFFu, last_run));
        }
      }
    }
    FLS_LANES(t) { fls_tile_copy_out<T>(w.sm + STAGE, out, t); }
  }
};

// ===================================================================== DICT (u32 FFOR indices -> gather)
// payload: [packed 128*bw] of u32 lanes; dictionary = raw T values, one per row group (L1/L2 resident).
template <typename T> struct DecDICT {
  static constexpr uint32_t SMEM = 4096u;
  FLS_MEM static void run(FlsWarp &w, const uint8_t *in, uint32_t bw, uint64_t base, const T *dict, uint32_t n_dict,
                          T *out) {
    if (bw) { w.stage(0, in, 128u * bw); w.wait(); }
    FLS_LANES(t) {
      const uint32_t q = t & 7u, g = t >> 3;
      for (uint32_t j = 0; j < 8u; ++j) {
        const uint32_t row = g * 8u + j;
        fls_u4 ix;
        if (bw) ix = fls_unpack32_row(w.sm, bw, row, q, (uint32_t)base);
        else { ix.x = ix.y = ix.z = ix.w = (uint32_t)base; }
        T v[4];
        v[0] = fls_ldg(dict + fls_min(ix.x, n_dict - 1u));
        v[1] = fls_ldg(dict + fls_min(ix.y, n_dict - 1u));
        v[2] = fls_ldg(dict + fls_min(ix.z, n_dict - 1u));
        v[3] = fls_ldg(dict + fls_min(ix.w, n_dict - 1u));
        T *o = out + row * 32u + q * 4u;
        if (sizeof(T) == 8) {
          fls_ul2 a, b;
          a.x = (uint64_t)v[0]; a.y = (uint64_t)v[1]; b.x = (uint64_t)v[2]; b.y = (uint64_t)v[3];
          fls_st128(o, a); fls_st128(o + 2, b);
        } else if (sizeof(T) == 4) {
          fls_u4 a;
          a.x = (uint32_t)v[0]; a.y = (uint32_t)v[1]; a.z = (uint32_t)v[2]; a.w = (uint32_t)v[3];
          fls_st128(o, a);
        } else {
          o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; o[3] = v[3];
        }
      }
    }
  }
};

// ===================================================================== ALP (f64)
// payload: [packed 128*bw][exc values n x f64, pad16][exc positions n x u16, pad16]
// value = (double)(int64)(bits + for) * (double)FACT[f] * FRAC[e]  — int->fp convert, then two separately
// rounded IEEE multiplies in this order (alp.hpp:129-134); never fused, never reassociated.
#ifdef FLS_HOST_EMUL
#define FLS_CONST_TABLE static const
#else
#define FLS_CONST_TABLE __device__ const
#endif
FLS_CONST_TABLE double fls_alp_frac[21] = {  // alp_constants.hpp:112-132 (FRAC_ARR), literals verbatim
    1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001, 0.00000001, 0.000000001, 0.0000000001,
    0.00000000001, 0.000000000001, 0.0000000000001, 0.00000000000001, 0.000000000000001, 0.0000000000000001,
    0.00000000000000001, 0.000000000000000001, 0.0000000000000000001, 0.00000000000000000001};
FLS_CONST_TABLE double fls_alp_fact[19] = {  // (double)FACT_ARR[f] (alp_constants.hpp:43-61); 10^f, f<=18, is exact
    1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18};

FLS_DEV uint64_t fls_alp_value(uint64_t bits, double fact, double frac) {
#ifdef FLS_HOST_EMUL
  volatile double a = (double)(int64_t)bits * fact;
  volatile double b = a * frac;
  double r = b;
  uint64_t u;
  memcpy(&u, &r, 8);
  return u;
#else
  return (uint64_t)__double_as_longlong(__dmul_rn(__dmul_rn(__ll2double_rn((long long)bits), fact), frac));
#endif
}

struct DecALP {
  // The slot holds 32 packed word rows.  ALP integers are narrow (a vector that needs more than 32 bits is rare), and
  // an 8 KB slot — what a 64-bit-wide vector would need — halves the resident warps of every launch.  Wider vectors
  // are decoded in two passes of 32 output rows: the packed words of rows [32 p, 32 p + 32) are word rows
  // [32 p W / 64, ((32 p + 32) W - 1) / 64], at most 32 of them.
#ifndef FLS_ALP_SLOT
#define FLS_ALP_SLOT 4096u
#endif
  static constexpr uint32_t SMEM = FLS_ALP_SLOT;
  FLS_MEM static void run(FlsWarp &w, const uint8_t *in, uint32_t bw, uint32_t e, uint32_t f, uint32_t n_exc,
                          uint64_t base, double *out) {
    const double fact = fls_alp_fact[f < 19u ? f : 18u], frac = fls_alp_frac[e < 21u ? e : 20u];
    const uint32_t passes = 128u * bw > SMEM ? 2u : 1u, rows = 64u / passes;
    for (uint32_t ps = 0; ps < passes; ++ps) {
      const uint32_t r0 = ps * rows;
      uint32_t k0 = 0, k1 = 0;  // staged word rows [k0, k1]
      if (bw) {
        k0 = (r0 * bw) >> 6; k1 = ((r0 + rows) * bw - 1u) >> 6;
        if (ps) w.fence_async();  // pass 0's reads of the slot before the next bulk copy lands in it
        w.stage(0, in + 128u * k0, 128u * (k1 - k0 + 1u));
        w.wait();
      }
      const uint8_t *sm0 = w.sm - 128u * k0;  // fls_unpack64_rows addresses word rows absolutely; it reads [k0, k1] only
      FLS_LANES(t) {
        const uint32_t q = t & 7u, g = t >> 3, per = rows / 4u;
#pragma unroll 4
        for (uint32_t j = 0; j < per; ++j) {
          const uint32_t row = r0 + g * per + j;
          fls_ul2 v;
          if (bw) v = fls_unpack64_rows(sm0, bw, row, q, base, k1);
          else { v.x = base; v.y = base; }
          v.x = fls_alp_value(v.x, fact, frac);
          v.y = fls_alp_value(v.y, fact, frac);
          fls_st128((uint8_t *)out + row * 128u + q * 16u, v);
        }
      }
    }
    if (n_exc) {  // patches land after the region barrier above, so they win over the decoded values
      const uint64_t *xv = (const uint64_t *)(in + 128ull * bw);
      const uint16_t *xp = (const uint16_t *)(in + 128ull * bw + (((uint64_t)n_exc * 8u + 15u) & ~15ull));
      FLS_LANES(t) {
        for (uint32_t i = t; i < n_exc; i += 32u) ((uint64_t *)out)[fls_ldg(xp + i) & 1023u] = fls_ldg(xv + i);
      }
    }
  }
};

// ===================================================================== ALP (f32)
// payload: [packed 128*bw (u32 lanes)][exc values n x f32, pad16][exc positions n x u16, pad16]
// value = (float)(int64)(bits + for) * (float)FACT[f] * FRAC[e] in float arithmetic (AlpTypedConstants<float>,
// alp_constants.hpp:66-83; DecodeValue alp.hpp:129-134): one int64 -> float conversion, two separately rounded multiplies.
FLS_CONST_TABLE float fls_alpf_frac[11] = {1.0F, 0.1F, 0.01F, 0.001F, 0.0001F, 0.00001F, 0.000001F, 0.0000001F, 0.00000001F,
                                           0.000000001F, 0.0000000001F};
FLS_DEV uint32_t fls_alpf_value(int64_t n, float fact, float frac) {
#ifdef FLS_HOST_EMUL
  volatile float a = (float)n * fact;
  volatile float b = a * frac;
  float r = b;
  uint32_t u;
  memcpy(&u, &r, 4);
  return u;
#else
  return (uint32_t)__float_as_int(__fmul_rn(__fmul_rn(__ll2float_rn((long long)n), fact), frac));
#endif
}
struct DecALP32 {
  static constexpr uint32_t SMEM = 4096u;
  FLS_MEM static void run(FlsWarp &w, const uint8_t *in, uint32_t bw, uint32_t e, uint32_t f, uint32_t n_exc,
                          int64_t base, float *out) {
    const float fact = (float)fls_alp_fact[f < 19u ? f : 18u], frac = fls_alpf_frac[e < 11u ? e : 10u];
    if (bw) { w.stage(0, in, 128u * bw); w.wait(); }
    FLS_LANES(t) {
      const uint32_t q = t & 7u, g = t >> 3;
#pragma unroll 4
      for (uint32_t j = 0; j < 8u; ++j) {
        const uint32_t row = g * 8u + j;
        fls_u4 v;
        if (bw) v = fls_unpack32_row(w.sm, bw, row, q, 0u);
        else { v.x = 0; v.y = 0; v.z = 0; v.w = 0; }
        v.x = fls_alpf_value(base + (int64_t)v.x, fact, frac);
        v.y = fls_alpf_value(base + (int64_t)v.y, fact, frac);
        v.z = fls_alpf_value(base + (int64_t)v.z, fact, frac);
        v.w = fls_alpf_value(base + (int64_t)v.w, fact, frac);
        fls_st128((uint8_t *)out + row * 128u + q * 16u, v);
      }
    }
    if (n_exc) {  // patches land after the region barrier above, so they win over the decoded values
      const uint32_t *xv = (const uint32_t *)(in + 128ull * bw);
      const uint16_t *xp = (const uint16_t *)(in + 128ull * bw + (((uint64_t)n_exc * 4u + 15u) & ~15ull));
      FLS_LANES(t) {
        for (uint32_t i = t; i < n_exc; i += 32u) ((uint32_t *)out)[fls_ldg(xp + i) & 1023u] = fls_ldg(xv + i);
      }
    }
  }
};

// ===================================================================== ALP-RD (f64 bit patterns)
// payload: [right packed 128*rbw (u64 lanes)][left packed 128*lbw (u16 lanes)][exc left n x u16, pad16]
//          [exc pos n x u16, pad16];  out = (dict[left] << rbw) | right; exceptions replace the left part.
struct DecALPRD {
  static constexpr uint32_t SMEM = 8192u + 128u * 16u;
  FLS_MEM static uint64_t dict_at(uint64_t lo, uint64_t hi, uint32_t i) {
    const uint64_t d = (i & 4u) ? hi : lo;
    return (d >> (16u * (i & 3u))) & 0xFFFFull;
  }
  FLS_MEM static void run(FlsWarp &w, const uint8_t *in, uint32_t rbw, uint32_t lbw, uint32_t n_exc, uint64_t dict_lo,
                          uint64_t dict_hi, uint64_t *out) {
    w.stage(0, in, 128u * (rbw + lbw));  // right and left parts are adjacent in the payload
    w.wait();
    const uint8_t *right = w.sm;
    const uint16_t *left = (const uint16_t *)(w.sm + 128u * rbw);
    const uint32_t sh = rbw & 63u;  // rbw in [48,63]
    FLS_LANES(t) {
      const uint32_t q = t & 7u, g = t >> 3;
#pragma unroll 2
      for (uint32_t j = 0; j < 16u; ++j) {
        const uint32_t row = g * 16u + j, p = row * 16u + q * 2u;  // positions p, p+1
        fls_ul2 v;
        if (rbw) v = fls_unpack64_row(right, rbw, row, q, 0);
        else { v.x = 0; v.y = 0; }
        const uint32_t l0 = lbw ? fls_extract<uint16_t>(left, lbw, p >> 6, p & 63u) : 0u;
        const uint32_t l1 = lbw ? fls_extract<uint16_t>(left, lbw, (p + 1u) >> 6, (p + 1u) & 63u) : 0u;
        v.x |= dict_at(dict_lo, dict_hi, l0 & 7u) << sh;
        v.y |= dict_at(dict_lo, dict_hi, l1 & 7u) << sh;
        fls_st128((uint8_t *)out + row * 128u + q * 16u, v);
      }
    }
    if (n_exc) {
      const uint16_t *xl = (const uint16_t *)(in + 128ull * (rbw + lbw));
      const uint16_t *xp = (const uint16_t *)((const uint8_t *)xl + (((uint64_t)n_exc * 2u + 15u) & ~15ull));
      FLS_LANES(t) {
        for (uint32_t i = t; i < n_exc; i += 32u) {
          const uint32_t p = fls_ldg(xp + i) & 1023u;
          const uint64_t r = rbw ? fls_extract<uint64_t>((const uint64_t *)right, rbw, p >> 4, p & 15u) : 0ull;
          out[p] = ((uint64_t)fls_ldg(xl + i) << sh) | r;
        }
      }
    }
  }
};
