// fls_writer.cpp — FLS encoder + file writer (host C++17).
//
// Role: produces the .fls images the read path consumes.  It is the write half the reference only
// declares (/root/reference/src/include/fastlanes_facade.hpp:40-44 createFile/writeChunk/finalizeFile;
// COPY ... TO 'x.fls' at /root/reference/src/writer/write_fastlane_stream.cpp:125-177) and is used by
// the tests and bench.py to mint synthetic workloads.  NOT on the timed decode path.
// Layout constants: fls_format.h.  Arithmetic twins: ALP follows
// duckdb/src/include/duckdb/storage/compression/alp/algorithm/alp.hpp:100-134,300-364; ALP-RD follows
// .../alprd/algorithm/alprd.hpp:70-199; FSST stream semantics follow duckdb/third_party/fsst/fsst.h:93-104.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "fls_format.h"
#include "fls_pack.h"
#include "fls_reader.h"

namespace flsw {

typedef std::vector<uint8_t> Bytes;

static void pad_to(Bytes &b, size_t a) { b.resize(align_up(b.size(), a), 0); }
template <typename T>
static void put(Bytes &b, const T &v) {
  const uint8_t *p = (const uint8_t *)&v;
  b.insert(b.end(), p, p + sizeof(T));
}
static void put_bytes(Bytes &b, const void *p, size_t n) {
  b.insert(b.end(), (const uint8_t *)p, (const uint8_t *)p + n);
}

// One encoded column chunk: segments in scheme order.
struct Chunk {
  uint8_t scheme = 0;
  uint32_t aux32 = 0;
  uint64_t aux64[2] = {0, 0};
  std::vector<Bytes> segs;
  size_t bytes() const {
    size_t s = 0;
    for (auto &x : segs) s += align_up(x.size(), FLS_SEG_ALIGN);
    return s;
  }
};

// A stream under construction: header table + data segment.
struct Stream {
  Bytes hdr, data;
  void begin_vec(uint32_t meta, uint64_t base) {
    pad_to(data, 16);
    fls_vechdr h;
    h.ep = (uint32_t)(data.size() / 16);
    h.meta = meta;
    h.base = base;
    put(hdr, h);
  }
  void move_into(Chunk &c) {
    c.segs.push_back(std::move(hdr));
    c.segs.push_back(std::move(data));
  }
};

template <typename T>
static void append_packed(Bytes &data, const T *vals, uint32_t W) {
  if (!W) return;
  size_t o = data.size();
  data.resize(o + 128 * W);
  pack_vec<T>(vals, W, (T *)(data.data() + o));
}

// ------------------------------------------------------------------ integer schemes
// `v` always holds n_vec*1024 values (tail padded by the caller).

template <typename T>
static void ffor_vec(Stream &s, const T *v, uint32_t e = 0, uint32_t f = 0, uint32_t n = 0, int force_bw = -1) {
  // Frame of reference = min in whichever of the unsigned / signed orders gives the narrower range
  // (decode is a wrapping add, so both are valid for either signedness).
  typedef typename std::make_signed<T>::type ST;
  T umn = v[0], umx = v[0];
  ST smn = (ST)v[0], smx = (ST)v[0];
  for (uint32_t i = 1; i < FLS_VEC; ++i) {
    umn = std::min(umn, v[i]); umx = std::max(umx, v[i]);
    smn = std::min(smn, (ST)v[i]); smx = std::max(smx, (ST)v[i]);
  }
  T urange = (T)(umx - umn), srange = (T)((T)smx - (T)smn);
  T mn = urange <= srange ? umn : (T)smn;
  T range = std::min(urange, srange);
  T tmp[FLS_VEC];
  for (uint32_t i = 0; i < FLS_VEC; ++i) tmp[i] = (T)(v[i] - mn);
  uint32_t W = bits_needed<T>(range);
  if (force_bw >= 0 && (uint32_t)force_bw > W) W = force_bw;
  s.begin_vec(FLS_META(W, e, f, n), (uint64_t)mn);
  append_packed<T>(s.data, tmp, W);
}

template <typename T>
static Chunk enc_ffor(const T *v, uint32_t n_vec, int force_bw) {
  Chunk c; c.scheme = FLS_S_FFOR;
  Stream s;
  for (uint32_t i = 0; i < n_vec; ++i) ffor_vec<T>(s, v + i * FLS_VEC, 0, 0, 0, force_bw);
  s.move_into(c);
  return c;
}

// DELTA: lane bases = the value preceding the lane's first original in the vector; for the lane that
// starts at original 0 the base is `first_prev` (chosen by the caller).  Deltas are then FFOR'd as unsigned T
// with wrapping arithmetic (base = min delta in the signed sense so ascending data packs tightly).
template <typename T>
static void delta_stream_vec(Stream &s, const T *v, T first_prev) {
  constexpr uint32_t TB = sizeof(T) * 8, L = 1024 / TB;
  typedef typename std::make_signed<T>::type ST;
  T bases[L], d[FLS_VEC];
  for (uint32_t l = 0; l < L; ++l) {
    uint32_t st = lane_start<T>(l);
    bases[l] = st ? v[st - 1] : first_prev;
  }
  delta_vec<T>(v, bases, d);
  ST mn = (ST)d[0], mx = (ST)d[0];
  for (uint32_t i = 1; i < FLS_VEC; ++i) { mn = std::min(mn, (ST)d[i]); mx = std::max(mx, (ST)d[i]); }
  T range = (T)((T)mx - (T)mn);
  uint32_t W = bits_needed<T>(range);
  for (uint32_t i = 0; i < FLS_VEC; ++i) d[i] = (T)(d[i] - (T)mn);
  s.begin_vec(FLS_META(W, 0, 0, 0), (uint64_t)(T)mn);
  put_bytes(s.data, bases, sizeof(bases));  // 128 bytes for every T
  append_packed<T>(s.data, d, W);
}

template <typename T>
static Chunk enc_delta(const T *v, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_DELTA;
  Stream s;
  for (uint32_t i = 0; i < n_vec; ++i) {
    // value "before" the vector: the previous vector's last value, or a linear extrapolation for the first
    T prev = i ? v[i * FLS_VEC - 1] : (T)(v[0] - (T)(v[1] - v[0]));
    delta_stream_vec<T>(s, v + i * FLS_VEC, prev);
  }
  s.move_into(c);
  return c;
}

// RLE: run values raw + u16 run-index stream, delta-coded in the transposed layout with bw in {0,1}.
template <typename T>
static Chunk enc_rle(const T *v, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_RLE;
  Stream s;
  for (uint32_t i = 0; i < n_vec; ++i) {
    const T *x = v + i * FLS_VEC;
    std::vector<T> runs;
    uint16_t idx[FLS_VEC];
    runs.push_back(x[0]); idx[0] = 0;
    for (uint32_t j = 1; j < FLS_VEC; ++j) {
      if (x[j] != x[j - 1]) runs.push_back(x[j]);
      idx[j] = (uint16_t)(runs.size() - 1);
    }
    uint16_t bases[64], d[FLS_VEC];
    for (uint32_t l = 0; l < 64; ++l) {
      uint32_t st = lane_start<uint16_t>(l);
      bases[l] = st ? idx[st - 1] : 0;
    }
    delta_vec<uint16_t>(idx, bases, d);
    uint32_t W = runs.size() > 1 ? 1 : 0;
    s.begin_vec(FLS_META(W, 0, 0, runs.size()), 0);
    put_bytes(s.data, runs.data(), runs.size() * sizeof(T));
    pad_to(s.data, 16);
    put_bytes(s.data, bases, sizeof(bases));
    append_packed<uint16_t>(s.data, d, W);
  }
  s.move_into(c);
  return c;
}

// DICT: per-row-group dictionary (sorted distinct values, raw) + u32 FFOR indices (base 0).
template <typename T>
static bool enc_dict(const T *v, uint32_t n_vec, uint32_t max_dict, Chunk &c) {
  std::vector<T> dict(v, v + (size_t)n_vec * FLS_VEC);
  std::sort(dict.begin(), dict.end());
  dict.erase(std::unique(dict.begin(), dict.end()), dict.end());
  if (dict.size() > max_dict) return false;
  c = Chunk(); c.scheme = FLS_S_DICT; c.aux32 = (uint32_t)dict.size();
  Stream s;
  std::vector<uint32_t> idx(FLS_VEC);
  for (uint32_t i = 0; i < n_vec; ++i) {
    for (uint32_t j = 0; j < FLS_VEC; ++j)
      idx[j] = (uint32_t)(std::lower_bound(dict.begin(), dict.end(), v[i * FLS_VEC + j]) - dict.begin());
    ffor_vec<uint32_t>(s, idx.data());
  }
  s.move_into(c);
  Bytes d; put_bytes(d, dict.data(), dict.size() * sizeof(T));
  c.segs.push_back(std::move(d));
  return true;
}

// FFOR_PATCH (SLPATCH / FREQUENCY): per vector the frame [lo, lo + 2^W) that makes 128 W + exceptions x (sizeof T + 2)
// smallest; values outside the frame go to the exception list and pack as 0.
template <typename T>
static Chunk enc_ffor_patch(const T *v, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_FFOR_PATCH;
  Stream s;
  constexpr uint32_t TB = sizeof(T) * 8;
  for (uint32_t i = 0; i < n_vec; ++i) {
    const T *x = v + i * FLS_VEC;
    std::vector<T> srt(x, x + FLS_VEC);
    std::sort(srt.begin(), srt.end());
    uint64_t best_cost = ~0ull;
    uint32_t best_w = TB;
    T best_lo = srt[0];
    for (uint32_t W = 0; W <= TB; ++W) {
      // widest coverage of a window of 2^W consecutive values over the sorted vector (two pointers)
      uint32_t cover = 0, at = 0;
      for (uint32_t a = 0, b = 0; a < FLS_VEC; ++a) {
        if (b < a) b = a;
        while (b < FLS_VEC && (W >= TB || (T)(srt[b] - srt[a]) <= (T)(((T)1 << W) - 1))) ++b;
        if (b - a > cover) { cover = b - a; at = a; }
        if (b == FLS_VEC) break;
      }
      const uint64_t cost = 128ull * W + (uint64_t)(FLS_VEC - cover) * (sizeof(T) + 2);
      if (cost < best_cost) { best_cost = cost; best_w = W; best_lo = srt[at]; }
    }
    T tmp[FLS_VEC];
    std::vector<T> xv;
    std::vector<uint16_t> xp;
    for (uint32_t j = 0; j < FLS_VEC; ++j) {
      const T d = (T)(x[j] - best_lo);
      const bool in = best_w >= TB || d <= (T)(((T)1 << best_w) - 1);
      tmp[j] = in ? d : (T)0;
      if (!in) { xv.push_back(x[j]); xp.push_back((uint16_t)j); }
    }
    s.begin_vec(FLS_META(best_w, 0, 0, xv.size()), (uint64_t)best_lo);
    append_packed<T>(s.data, tmp, best_w);
    put_bytes(s.data, xv.data(), xv.size() * sizeof(T)); pad_to(s.data, 16);
    put_bytes(s.data, xp.data(), xp.size() * 2); pad_to(s.data, 16);
  }
  s.move_into(c);
  return c;
}

// CROSS_RLE: runs over the whole row group (they span vectors): values + cumulative ends.
template <typename T>
static Chunk enc_cross_rle(const T *v, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_CROSS_RLE;
  const size_t n = (size_t)n_vec * FLS_VEC;
  std::vector<T> vals;
  std::vector<uint32_t> ends;
  for (size_t i = 0; i < n; ++i) {
    if (i == 0 || v[i] != v[i - 1]) { if (i) ends.push_back((uint32_t)i); vals.push_back(v[i]); }
  }
  ends.push_back((uint32_t)n);
  c.aux32 = (uint32_t)vals.size();
  Bytes a, b;
  put_bytes(a, vals.data(), vals.size() * sizeof(T));
  put_bytes(b, ends.data(), ends.size() * 4);
  c.segs.push_back(std::move(a));
  c.segs.push_back(std::move(b));
  return c;
}

template <typename T>
static Chunk enc_uncompressed(const T *v, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_UNCOMPRESSED;
  Bytes d; put_bytes(d, v, (size_t)n_vec * FLS_VEC * sizeof(T));
  c.segs.push_back(std::move(d));
  return c;
}

template <typename T>
static bool enc_constant(const T *v, uint32_t n_vec, Chunk &c) {
  for (size_t i = 1; i < (size_t)n_vec * FLS_VEC; ++i)
    if (v[i] != v[0]) return false;
  c = Chunk(); c.scheme = FLS_S_CONSTANT; c.aux64[0] = (uint64_t)v[0];
  return true;
}

// Pick the smallest of the integer schemes (or the forced one).
template <typename T>
static Chunk enc_int_auto(const T *v, uint32_t n_vec, int force, int force_bw) {
  Chunk c;
  if (force == FLS_S_UNCOMPRESSED) return enc_uncompressed<T>(v, n_vec);
  if (force == FLS_S_FFOR) return enc_ffor<T>(v, n_vec, force_bw);
  if (force == FLS_S_DELTA) return enc_delta<T>(v, n_vec);
  if (force == FLS_S_RLE) return enc_rle<T>(v, n_vec);
  if (force == FLS_S_DICT) { if (enc_dict<T>(v, n_vec, 1u << 20, c)) return c; return enc_ffor<T>(v, n_vec, -1); }
  if (force == FLS_S_CONSTANT) { if (enc_constant<T>(v, n_vec, c)) return c; return enc_ffor<T>(v, n_vec, -1); }
  if (force == FLS_S_FFOR_PATCH) return enc_ffor_patch<T>(v, n_vec);
  if (force == FLS_S_CROSS_RLE) return enc_cross_rle<T>(v, n_vec);
  if (enc_constant<T>(v, n_vec, c)) return c;
  Chunk best = enc_ffor<T>(v, n_vec, -1);
  // (the exception list and row-group-wide runs are taken only when they pay clearly: their decoders do more work per value)
  { Chunk p = enc_ffor_patch<T>(v, n_vec); if (p.bytes() + p.bytes() / 4 < best.bytes()) best = std::move(p); }
  { Chunk r = enc_cross_rle<T>(v, n_vec); if (2 * r.bytes() < best.bytes()) best = std::move(r); }
  { Chunk d = enc_delta<T>(v, n_vec); if (d.bytes() < best.bytes()) best = std::move(d); }
  { Chunk r = enc_rle<T>(v, n_vec); if (r.bytes() < best.bytes()) best = std::move(r); }
  { Chunk d; if (enc_dict<T>(v, n_vec, 4096, d) && d.bytes() < best.bytes()) best = std::move(d); }
  return best;
}

// ------------------------------------------------------------------ ALP (doubles)
// Arithmetic follows DuckDB's in-tree ALP (same authors as FastLanes):
//   encode  alp.hpp:118-123   n = v * EXP[e] * FRAC[f]; round by +MAGIC-MAGIC (alp.hpp:108-115)
//   decode  alp.hpp:129-134   (double)n * (double)FACT[f] * FRAC[e]      — two IEEE multiplies, in order
//   special alp.hpp:100-103   NaN/inf/out-of-range/-0.0 cannot be encoded -> exception
// This translation unit must be compiled with -ffp-contract=off.
static const double ALP_EXP[] = {1.0, 10.0, 100.0, 1000.0, 10000.0, 100000.0, 1000000.0, 10000000.0, 100000000.0,
                                 1000000000.0, 10000000000.0, 100000000000.0, 1000000000000.0, 10000000000000.0,
                                 100000000000000.0, 1000000000000000.0, 10000000000000000.0, 100000000000000000.0,
                                 1000000000000000000.0};
static const double ALP_FRAC[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001, 0.00000001,
                                  0.000000001, 0.0000000001, 0.00000000001, 0.000000000001, 0.0000000000001,
                                  0.00000000000001, 0.000000000000001, 0.0000000000000001, 0.00000000000000001,
                                  0.000000000000000001};
static const int64_t ALP_FACT[] = {1LL, 10LL, 100LL, 1000LL, 10000LL, 100000LL, 1000000LL, 10000000LL, 100000000LL,
                                   1000000000LL, 10000000000LL, 100000000000LL, 1000000000000LL, 10000000000000LL,
                                   100000000000000LL, 1000000000000000LL, 10000000000000000LL, 100000000000000000LL,
                                   1000000000000000000LL};
static const double ALP_MAGIC = 6755399441055744.0;  // 2^51 + 2^52
static const double ALP_LIMIT = 9223372036854774784.0;
static const int ALP_MAX_E = 18;

static inline bool alp_impossible(double n) {
  return !std::isfinite(n) || std::isnan(n) || n > ALP_LIMIT || n < -ALP_LIMIT || (n == 0.0 && std::signbit(n));
}
static inline int64_t alp_encode(double v, int e, int f) {
  double t = v * ALP_EXP[e] * ALP_FRAC[f];
  if (alp_impossible(t)) return (int64_t)ALP_LIMIT;
  t = t + ALP_MAGIC - ALP_MAGIC;
  return (int64_t)t;
}
static inline double alp_decode(int64_t n, int e, int f) {
  return (double)n * (double)ALP_FACT[f] * ALP_FRAC[e];
}
static inline bool same_bits(double a, double b) { return memcmp(&a, &b, 8) == 0; }

// estimated bits for a sample under (e,f); exceptions cost 64+16 bits (alp.hpp:160-193)
static uint64_t alp_estimate(const std::vector<double> &smp, int e, int f, bool penalize) {
  uint64_t exc = 0, ok = 0;
  int64_t mn = INT64_MAX, mx = INT64_MIN;
  for (double v : smp) {
    int64_t n = alp_encode(v, e, f);
    if (alp_decode(n, e, f) == v) { ++ok; mn = std::min(mn, n); mx = std::max(mx, n); }
    else ++exc;
  }
  if (penalize && ok < 2) return UINT64_MAX;
  uint64_t delta = ok ? (uint64_t)mx - (uint64_t)mn : 0;
  uint32_t bits = bits_needed<uint64_t>(delta);
  return smp.size() * bits + exc * (64 + 16);
}

struct AlpCombo { int e, f; uint64_t n, size; };
static bool alp_better(const AlpCombo &a, const AlpCombo &b) {  // alp.hpp:143-154
  if (a.n != b.n) return a.n > b.n;
  if (a.size != b.size) return a.size < b.size;
  if (a.e != b.e) return a.e > b.e;
  return a.f > b.f;
}

// first-level sampling: up to 8 equidistant vectors x 32 equidistant values; top 5 combinations
static std::vector<AlpCombo> alp_topk(const double *v, uint32_t n_vec) {
  std::map<std::pair<int, int>, uint64_t> votes;
  uint32_t step = std::max(1u, n_vec / 8);
  for (uint32_t vi = 0; vi < n_vec; vi += step) {
    std::vector<double> smp;
    for (uint32_t j = 0; j < FLS_VEC; j += 32) smp.push_back(v[vi * FLS_VEC + j]);
    AlpCombo best = {ALP_MAX_E, ALP_MAX_E, 0, UINT64_MAX};
    bool have = false;
    for (int e = ALP_MAX_E; e >= 0; --e)
      for (int f = e; f >= 0; --f) {
        AlpCombo c = {e, f, 0, alp_estimate(smp, e, f, true)};
        if (!have || alp_better(c, best)) { best = c; have = true; }
      }
    votes[{best.e, best.f}]++;
  }
  std::vector<AlpCombo> out;
  for (auto &kv : votes) out.push_back({kv.first.first, kv.first.second, kv.second, 0});
  std::sort(out.begin(), out.end(), alp_better);
  if (out.size() > 5) out.resize(5);
  return out;
}

static void alp_vec(Stream &s, const double *v, const std::vector<AlpCombo> &topk) {
  int e = topk[0].e, f = topk[0].f;
  if (topk.size() > 1) {  // second-level sampling (alp.hpp:250-285)
    std::vector<double> smp;
    for (uint32_t j = 0; j < FLS_VEC; j += 32) smp.push_back(v[j]);
    uint64_t best = UINT64_MAX; int worse = 0;
    for (auto &c : topk) {
      uint64_t sz = alp_estimate(smp, c.e, c.f, false);
      if (sz >= best) { if (++worse == 2) break; continue; }
      best = sz; e = c.e; f = c.f; worse = 0;
    }
  }
  int64_t enc[FLS_VEC];
  std::vector<uint16_t> pos; std::vector<double> exc;
  for (uint32_t i = 0; i < FLS_VEC; ++i) {
    enc[i] = alp_encode(v[i], e, f);
    double d = alp_decode(enc[i], e, f);
    if (!same_bits(d, v[i])) { pos.push_back((uint16_t)i); exc.push_back(v[i]); }
  }
  // exceptions take the first non-exception encoded value so they do not widen the range (alp.hpp:314-328)
  int64_t filler = 0;
  { size_t k = 0;
    for (uint32_t i = 0; i < FLS_VEC; ++i) { if (k < pos.size() && pos[k] == i) { ++k; continue; } filler = enc[i]; break; } }
  for (uint16_t p : pos) enc[p] = filler;
  int64_t mn = enc[0], mx = enc[0];
  for (uint32_t i = 1; i < FLS_VEC; ++i) { mn = std::min(mn, enc[i]); mx = std::max(mx, enc[i]); }
  uint64_t tmp[FLS_VEC];
  for (uint32_t i = 0; i < FLS_VEC; ++i) tmp[i] = (uint64_t)enc[i] - (uint64_t)mn;
  uint32_t W = bits_needed<uint64_t>((uint64_t)mx - (uint64_t)mn);
  s.begin_vec(FLS_META(W, e, f, pos.size()), (uint64_t)mn);
  append_packed<uint64_t>(s.data, tmp, W);
  put_bytes(s.data, exc.data(), exc.size() * 8); pad_to(s.data, 16);
  put_bytes(s.data, pos.data(), pos.size() * 2); pad_to(s.data, 16);
}

static thread_local int tl_force_alp_ef = -1;  // test knob: e * 32 + f, or -1 (flsw_force_alp_ef)
static Chunk enc_alp(const double *v, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_ALP;
  Stream s;
  std::vector<AlpCombo> topk = alp_topk(v, n_vec);
  if (tl_force_alp_ef >= 0) topk.assign(1, AlpCombo{tl_force_alp_ef / 32, tl_force_alp_ef % 32, 1, 0});
  for (uint32_t i = 0; i < n_vec; ++i) alp_vec(s, v + i * FLS_VEC, topk);
  s.move_into(c);
  return c;
}

// ------------------------------------------------------------------ ALP (floats)
// DuckDB's AlpTypedConstants<float> (alp_constants.hpp:66-83): float arithmetic throughout, magic number 2^22 + 2^23,
// exponents 0..10; decode = (float)n * (float)FACT[f] * FRAC[e], two float multiplies (alp.hpp:129-134).  The encoded
// integers of a vector's non-exceptions differ by less than 2^32 (a float holds 24 bits), so the stream is u32 FFOR.
static const float ALPF_EXP[] = {1.0F, 10.0F, 100.0F, 1000.0F, 10000.0F, 100000.0F, 1000000.0F, 10000000.0F, 100000000.0F,
                                 1000000000.0F, 10000000000.0F};
static const float ALPF_FRAC[] = {1.0F, 0.1F, 0.01F, 0.001F, 0.0001F, 0.00001F, 0.000001F, 0.0000001F, 0.00000001F,
                                  0.000000001F, 0.0000000001F};
static const float ALPF_MAGIC = 12582912.0F;
static const int ALPF_MAX_E = 10;
static inline int64_t alpf_encode(float v, int e, int f) {
  float t = v * ALPF_EXP[e] * ALPF_FRAC[f];
  if (!std::isfinite(t) || std::isnan(t) || t > ALP_LIMIT || t < -ALP_LIMIT || (t == 0.0f && std::signbit(t))) return (int64_t)ALP_LIMIT;
  t = t + ALPF_MAGIC - ALPF_MAGIC;
  return (int64_t)t;
}
static inline float alpf_decode(int64_t n, int e, int f) { return (float)n * (float)ALP_FACT[f] * ALPF_FRAC[e]; }
static inline bool same_bits32(float a, float b) { return memcmp(&a, &b, 4) == 0; }
static uint64_t alpf_estimate(const std::vector<float> &smp, int e, int f, bool penalize) {
  uint64_t exc = 0, ok = 0;
  int64_t mn = INT64_MAX, mx = INT64_MIN;
  for (float v : smp) {
    int64_t n = alpf_encode(v, e, f);
    if (alpf_decode(n, e, f) == v) { ++ok; mn = std::min(mn, n); mx = std::max(mx, n); }
    else ++exc;
  }
  if (penalize && ok < 2) return UINT64_MAX;
  uint64_t delta = ok ? (uint64_t)mx - (uint64_t)mn : 0;
  return smp.size() * bits_needed<uint64_t>(delta) + exc * (32 + 16);
}
static std::vector<AlpCombo> alpf_topk(const float *v, uint32_t n_vec) {
  std::map<std::pair<int, int>, uint64_t> votes;
  uint32_t step = std::max(1u, n_vec / 8);
  for (uint32_t vi = 0; vi < n_vec; vi += step) {
    std::vector<float> smp;
    for (uint32_t j = 0; j < FLS_VEC; j += 32) smp.push_back(v[vi * FLS_VEC + j]);
    AlpCombo best = {ALPF_MAX_E, ALPF_MAX_E, 0, UINT64_MAX};
    bool have = false;
    for (int e = ALPF_MAX_E; e >= 0; --e)
      for (int f = e; f >= 0; --f) {
        AlpCombo c = {e, f, 0, alpf_estimate(smp, e, f, true)};
        if (!have || alp_better(c, best)) { best = c; have = true; }
      }
    votes[{best.e, best.f}]++;
  }
  std::vector<AlpCombo> out;
  for (auto &kv : votes) out.push_back({kv.first.first, kv.first.second, kv.second, 0});
  std::sort(out.begin(), out.end(), alp_better);
  if (out.size() > 5) out.resize(5);
  return out;
}
// false: some vector's encoded range does not fit 32 bits (then the column takes another scheme)
static bool enc_alp_f32(const float *v, uint32_t n_vec, Chunk &c) {
  c = Chunk(); c.scheme = FLS_S_ALP;
  Stream s;
  std::vector<AlpCombo> topk = alpf_topk(v, n_vec);
  if (tl_force_alp_ef >= 0) topk.assign(1, AlpCombo{std::min(tl_force_alp_ef / 32, ALPF_MAX_E), std::min(tl_force_alp_ef % 32, ALPF_MAX_E), 1, 0});
  for (uint32_t vi = 0; vi < n_vec; ++vi) {
    const float *x = v + vi * FLS_VEC;
    int e = topk[0].e, f = topk[0].f;
    if (topk.size() > 1) {
      std::vector<float> smp;
      for (uint32_t j = 0; j < FLS_VEC; j += 32) smp.push_back(x[j]);
      uint64_t best = UINT64_MAX; int worse = 0;
      for (auto &cb : topk) {
        uint64_t sz = alpf_estimate(smp, cb.e, cb.f, false);
        if (sz >= best) { if (++worse == 2) break; continue; }
        best = sz; e = cb.e; f = cb.f; worse = 0;
      }
    }
    int64_t enc[FLS_VEC];
    std::vector<uint16_t> pos; std::vector<float> exc;
    for (uint32_t i = 0; i < FLS_VEC; ++i) {
      enc[i] = alpf_encode(x[i], e, f);
      if (!same_bits32(alpf_decode(enc[i], e, f), x[i])) { pos.push_back((uint16_t)i); exc.push_back(x[i]); }
    }
    int64_t filler = 0;
    { size_t k = 0;
      for (uint32_t i = 0; i < FLS_VEC; ++i) { if (k < pos.size() && pos[k] == i) { ++k; continue; } filler = enc[i]; break; } }
    for (uint16_t p : pos) enc[p] = filler;
    int64_t mn = enc[0], mx = enc[0];
    for (uint32_t i = 1; i < FLS_VEC; ++i) { mn = std::min(mn, enc[i]); mx = std::max(mx, enc[i]); }
    if ((uint64_t)mx - (uint64_t)mn > 0xFFFFFFFFull) return false;
    uint32_t tmp[FLS_VEC];
    for (uint32_t i = 0; i < FLS_VEC; ++i) tmp[i] = (uint32_t)((uint64_t)enc[i] - (uint64_t)mn);
    uint32_t W = bits_needed<uint32_t>((uint32_t)((uint64_t)mx - (uint64_t)mn));
    s.begin_vec(FLS_META(W, e, f, pos.size()), (uint64_t)(int64_t)mn);  // base: the frame of reference, sign-extended
    append_packed<uint32_t>(s.data, tmp, W);
    put_bytes(s.data, exc.data(), exc.size() * 4); pad_to(s.data, 16);
    put_bytes(s.data, pos.data(), pos.size() * 2); pad_to(s.data, 16);
  }
  s.move_into(c);
  return true;
}
static Chunk enc_float_auto(const float *v, uint32_t n_vec, int force, int force_bw) {
  const uint32_t *bits = (const uint32_t *)v;
  Chunk a;
  if (force == FLS_S_ALP) { if (enc_alp_f32(v, n_vec, a)) return a; return enc_int_auto<uint32_t>(bits, n_vec, -1, -1); }
  if (force >= 0) return enc_int_auto<uint32_t>(bits, n_vec, force, force_bw);
  Chunk best = enc_int_auto<uint32_t>(bits, n_vec, -1, force_bw);
  if (best.scheme != FLS_S_CONSTANT && enc_alp_f32(v, n_vec, a) && a.bytes() < best.bytes()) best = std::move(a);
  return best;
}

// ------------------------------------------------------------------ ALP-RD (doubles, "real doubles")
// Split each 64-bit pattern at rbw: right part bit-packed as u64, left part (<=16 bits) dictionary-coded
// with <=8 entries / <=3 bits, misses stored as (u16 left, u16 pos) exceptions.  alprd.hpp:70-199.
static double alprd_estimate(const uint64_t *bits, size_t n, uint32_t rbw, std::vector<uint16_t> *dict_out) {
  std::unordered_map<uint64_t, uint32_t> cnt;
  for (size_t i = 0; i < n; ++i) cnt[bits[i] >> rbw]++;
  std::vector<std::pair<uint32_t, uint64_t>> srt;
  for (auto &kv : cnt) srt.push_back({kv.second, kv.first});
  std::sort(srt.begin(), srt.end(), [](auto &a, auto &b) { return a.first != b.first ? a.first > b.first : a.second < b.second; });
  uint32_t exc = 0;
  for (size_t i = 8; i < srt.size(); ++i) exc += srt[i].first;
  size_t nd = std::min<size_t>(8, srt.size());
  uint32_t lbw = std::max(1u, bits_needed<uint32_t>((uint32_t)nd - 1));
  if (dict_out) { dict_out->clear(); for (size_t i = 0; i < nd; ++i) dict_out->push_back((uint16_t)srt[i].second); }
  return rbw + lbw + (double)exc * 32.0 / (double)n;
}

// Test knob (flsw_inject_alprd): the NEXT ALP-RD chunk is assembled from streams another encoder produced — the
// reference tree's own AlpRD compressor (tests/golden/alprd.npz) — instead of this file's: cut position, dictionary,
// dictionary indices, right parts and exceptions go into the container verbatim, so a decoder under test consumes
// bytes the reference wrote.
struct AlprdInject {
  uint32_t rbw, lbw, ndict, n_vec;
  const uint16_t *dict, *left;      // dict[8]; left[n_vec * 1024] = dictionary indices
  const uint64_t *right;            // right[n_vec * 1024]
  const uint32_t *nexc;             // per vector
  const uint16_t *xpos, *xleft;     // per vector: 1024 slots each, the first nexc[v] used
};
static thread_local const AlprdInject *tl_alprd_inject = nullptr;

static Chunk enc_alprd_injected(const AlprdInject &in, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_ALP_RD;
  c.aux32 = in.rbw | (in.lbw << 8) | (in.ndict << 16);
  uint16_t d8[8] = {0};
  for (uint32_t i = 0; i < in.ndict && i < 8; ++i) d8[i] = in.dict[i];
  memcpy(c.aux64, d8, 16);
  Stream s;
  for (uint32_t vi = 0; vi < n_vec; ++vi) {
    const uint32_t src = vi < in.n_vec ? vi : in.n_vec - 1;  // padding vectors repeat the last one
    s.begin_vec(FLS_META(0, 0, 0, in.nexc[src]), 0);
    append_packed<uint64_t>(s.data, in.right + (size_t)src * FLS_VEC, in.rbw);
    append_packed<uint16_t>(s.data, in.left + (size_t)src * FLS_VEC, in.lbw);
    put_bytes(s.data, in.xleft + (size_t)src * FLS_VEC, in.nexc[src] * 2); pad_to(s.data, 16);
    put_bytes(s.data, in.xpos + (size_t)src * FLS_VEC, in.nexc[src] * 2); pad_to(s.data, 16);
  }
  s.move_into(c);
  return c;
}

static Chunk enc_alprd(const double *v, uint32_t n_vec) {
  if (tl_alprd_inject) { const AlprdInject *in = tl_alprd_inject; tl_alprd_inject = nullptr; return enc_alprd_injected(*in, n_vec); }
  Chunk c; c.scheme = FLS_S_ALP_RD;
  const uint64_t *bits = (const uint64_t *)v;
  std::vector<uint64_t> smp;
  for (size_t i = 0; i < (size_t)n_vec * FLS_VEC; i += 16) smp.push_back(bits[i]);
  uint32_t rbw = 48; double best = 1e300;
  for (uint32_t cut = 1; cut <= 16; ++cut) {
    double est = alprd_estimate(smp.data(), smp.size(), 64 - cut, nullptr);
    if (est <= best) { best = est; rbw = 64 - cut; }
  }
  std::vector<uint16_t> dict;
  alprd_estimate(smp.data(), smp.size(), rbw, &dict);
  uint32_t nd = (uint32_t)dict.size();
  uint32_t lbw = std::max(1u, bits_needed<uint32_t>(nd - 1));
  c.aux32 = rbw | (lbw << 8) | (nd << 16);
  uint16_t d8[8] = {0};
  for (uint32_t i = 0; i < nd; ++i) d8[i] = dict[i];
  memcpy(c.aux64, d8, 16);
  Stream s;
  const uint64_t rmask = (rbw == 64) ? ~0ULL : ((1ULL << rbw) - 1);
  for (uint32_t vi = 0; vi < n_vec; ++vi) {
    uint64_t right[FLS_VEC]; uint16_t left[FLS_VEC];
    std::vector<uint16_t> epos, eleft;
    for (uint32_t i = 0; i < FLS_VEC; ++i) {
      uint64_t b = bits[vi * FLS_VEC + i];
      right[i] = b & rmask;
      uint16_t lp = (uint16_t)(b >> rbw);
      uint32_t k = 0;
      while (k < nd && d8[k] != lp) ++k;
      if (k == nd) { epos.push_back((uint16_t)i); eleft.push_back(lp); k = 0; }
      left[i] = (uint16_t)k;
    }
    s.begin_vec(FLS_META(0, 0, 0, epos.size()), 0);
    append_packed<uint64_t>(s.data, right, rbw);
    append_packed<uint16_t>(s.data, left, lbw);
    put_bytes(s.data, eleft.data(), eleft.size() * 2); pad_to(s.data, 16);
    put_bytes(s.data, epos.data(), epos.size() * 2); pad_to(s.data, 16);
  }
  s.move_into(c);
  return c;
}

static Chunk enc_double_auto(const double *v, uint32_t n_vec, int force) {
  const uint64_t *bits = (const uint64_t *)v;
  Chunk c;
  if (force == FLS_S_ALP) return enc_alp(v, n_vec);
  if (force == FLS_S_ALP_RD) return enc_alprd(v, n_vec);
  if (force >= 0) return enc_int_auto<uint64_t>(bits, n_vec, force, -1);
  if (enc_constant<uint64_t>(bits, n_vec, c)) return c;
  Chunk best = enc_alp(v, n_vec);
  { Chunk r = enc_alprd(v, n_vec); if (r.bytes() < best.bytes()) best = std::move(r); }
  { Chunk r = enc_rle<uint64_t>(bits, n_vec); if (r.bytes() < best.bytes()) best = std::move(r); }
  { Chunk d; if (enc_dict<uint64_t>(bits, n_vec, 4096, d) && d.bytes() < best.bytes()) best = std::move(d); }
  { Chunk u = enc_uncompressed<uint64_t>(bits, n_vec); if (u.bytes() < best.bytes()) best = std::move(u); }
  return best;
}

// ------------------------------------------------------------------ FSST symbol table (own construction)
// Stream semantics follow duckdb/third_party/fsst/fsst.h:93-104: a compressed string is a sequence of
// 1-byte codes; code c < 255 expands to symbol[c] (1..8 bytes, little-endian in a u64); code 255 is an
// escape followed by one literal byte.  Table construction here is a plain greedy bottom-up builder
// (frequency x length gain over 5 rounds) — simpler than libfsst's, any valid table decodes the same way.
struct Fsst {
  uint32_t n = 0;
  uint64_t sym[255];
  uint8_t len[255];
  std::vector<uint16_t> by_first[256];  // codes sorted by descending length
  void index() {
    for (auto &b : by_first) b.clear();
    for (uint32_t c = 0; c < n; ++c) by_first[sym[c] & 0xFF].push_back((uint16_t)c);
    for (auto &b : by_first) std::sort(b.begin(), b.end(), [&](uint16_t x, uint16_t y) { return len[x] > len[y]; });
  }
  // longest match at p (rem bytes remain); returns code or -1
  inline int match(const uint8_t *p, size_t rem) const {
    uint64_t w = 0;
    memcpy(&w, p, rem < 8 ? rem : 8);
    for (uint16_t c : by_first[p[0]]) {
      uint32_t L = len[c];
      if (L > rem) continue;
      uint64_t m = L == 8 ? ~0ULL : ((1ULL << (8 * L)) - 1);
      if ((w & m) == sym[c]) return c;
    }
    return -1;
  }
  void blob(Bytes &out) const {  // u64 symbol[256] | u8 len[256]
    uint64_t s[256] = {0}; uint8_t l[256] = {0};
    for (uint32_t c = 0; c < n; ++c) { s[c] = sym[c]; l[c] = len[c]; }
    l[255] = 1;  // escape emits one literal byte
    put_bytes(out, s, sizeof(s)); put_bytes(out, l, sizeof(l));
  }
  void compress(const uint8_t *p, size_t n_in, Bytes &out) const {
    size_t i = 0;
    while (i < n_in) {
      int c = match(p + i, n_in - i);
      if (c < 0) { out.push_back(255); out.push_back(p[i]); ++i; }
      else { out.push_back((uint8_t)c); i += len[c]; }
    }
  }
};

static Fsst fsst_train(const uint8_t *bytes, const uint64_t *off, size_t n_str) {
  // sample ~32 KB of strings, evenly spread
  std::vector<std::pair<const uint8_t *, size_t>> smp;
  size_t total = off[n_str] - off[0], budget = 32768, step = 1;
  if (total > budget && n_str) step = std::max<size_t>(1, total / budget);
  for (size_t i = 0; i < n_str; i += step) smp.push_back({bytes + off[i], (size_t)(off[i + 1] - off[i])});
  Fsst t; t.index();
  struct Cand { uint64_t sym; uint8_t len; uint64_t gain; };
  for (int round = 0; round < 5; ++round) {
    // pseudo-symbol ids: 0..254 table codes, 256+b literal byte b
    std::unordered_map<uint64_t, uint64_t> c1, c2;  // key = id, key = id1<<16|id2
    auto id_sym = [&](int id, uint64_t &s, uint8_t &l) {
      if (id >= 256) { s = (uint64_t)(id - 256); l = 1; } else { s = t.sym[id]; l = t.len[id]; }
    };
    for (auto &sp : smp) {
      int prev = -1; size_t i = 0;
      while (i < sp.second) {
        int c = t.match(sp.first + i, sp.second - i);
        int id = c < 0 ? 256 + sp.first[i] : c;
        i += c < 0 ? 1 : t.len[c];
        c1[id]++;
        if (prev >= 0) c2[((uint64_t)prev << 16) | (uint64_t)id]++;
        prev = id;
      }
    }
    std::map<std::pair<uint64_t, uint8_t>, uint64_t> gains;
    for (auto &kv : c1) { uint64_t s; uint8_t l; id_sym((int)kv.first, s, l); gains[{s, l}] += kv.second * l; }
    for (auto &kv : c2) {
      uint64_t s1, s2; uint8_t l1, l2;
      id_sym((int)(kv.first >> 16), s1, l1); id_sym((int)(kv.first & 0xFFFF), s2, l2);
      if (l1 + l2 > 8) continue;
      gains[{s1 | (s2 << (8 * l1)), (uint8_t)(l1 + l2)}] += kv.second * (l1 + l2);
    }
    std::vector<Cand> cands;
    for (auto &kv : gains) cands.push_back({kv.first.first, kv.first.second, kv.second});
    std::sort(cands.begin(), cands.end(), [](const Cand &a, const Cand &b) {
      if (a.gain != b.gain) return a.gain > b.gain;
      if (a.len != b.len) return a.len > b.len;
      return a.sym < b.sym; });
    t.n = (uint32_t)std::min<size_t>(255, cands.size());
    for (uint32_t c = 0; c < t.n; ++c) { t.sym[c] = cands[c].sym; t.len[c] = cands[c].len; }
    t.index();
  }
  return t;
}

// ------------------------------------------------------------------ string schemes
// Input: bytes + u64 offsets[n+1] (n = n_vec*1024, tail padded with empty strings by the caller).
// All offsets stored are relative to the row group's decoded payload start and fit u32.

static void offsets_stream(Stream &s, const std::vector<uint32_t> &ends, uint32_t n_vec) {
  for (uint32_t i = 0; i < n_vec; ++i) {
    uint32_t first_prev = i ? ends[i * FLS_VEC - 1] : 0;
    delta_stream_vec<uint32_t>(s, ends.data() + i * FLS_VEC, first_prev);
  }
}

static Chunk enc_str_uncompressed(const uint8_t *bytes, const uint64_t *off, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_STR_UNCOMPRESSED;
  size_t n = (size_t)n_vec * FLS_VEC;
  std::vector<uint32_t> ends(n);
  for (size_t i = 0; i < n; ++i) ends[i] = (uint32_t)(off[i + 1] - off[0]);
  Stream s; offsets_stream(s, ends, n_vec); s.move_into(c);
  Bytes b; put_bytes(b, bytes + off[0], off[n] - off[0]);
  c.aux64[0] = off[n] - off[0];
  c.segs.push_back(std::move(b));
  return c;
}

// Test knob (flsw_inject_fsst): the NEXT FSST chunk takes its symbol table and compressed bytes from another
// encoder — libfsst itself (tests/golden/fsst.npz: fsst_create / fsst_compress of the reference tree's
// duckdb/third_party/fsst) — verbatim; only the offset streams around them are this file's.
struct FsstInject {
  const uint8_t *symtab;     // u64 symbol[256] | u8 len[256]
  const uint8_t *comp;       // compressed bytes of the row group's strings, back to back
  const uint64_t *comp_off;  // [n_strings + 1]
  uint32_t n_strings;
};
static thread_local const FsstInject *tl_fsst_inject = nullptr;

static Chunk enc_str_fsst(const uint8_t *bytes, const uint64_t *off, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_STR_FSST;
  size_t n = (size_t)n_vec * FLS_VEC;
  const FsstInject *inj = tl_fsst_inject;
  tl_fsst_inject = nullptr;
  Fsst t;
  if (!inj) t = fsst_train(bytes, off, n);
  std::vector<uint32_t> dec_ends(n), cmp_ends(n);
  Bytes comp;
  for (size_t i = 0; i < n; ++i) {
    if (!inj) t.compress(bytes + off[i], off[i + 1] - off[i], comp);
    else if (i < inj->n_strings) put_bytes(comp, inj->comp + inj->comp_off[i], inj->comp_off[i + 1] - inj->comp_off[i]);
    dec_ends[i] = (uint32_t)(off[i + 1] - off[0]);
    cmp_ends[i] = (uint32_t)comp.size();
  }
  Stream sd, sc;
  offsets_stream(sd, dec_ends, n_vec); sd.move_into(c);
  offsets_stream(sc, cmp_ends, n_vec); sc.move_into(c);
  Bytes tb;
  if (inj) put_bytes(tb, inj->symtab, 2304); else t.blob(tb);
  c.segs.push_back(std::move(tb));
  c.aux64[0] = off[n] - off[0];
  c.aux64[1] = comp.size();
  pad_to(comp, 16);
  comp.resize(comp.size() + 16, 0);  // slack so word-wise readers may over-read
  c.segs.push_back(std::move(comp));
  return c;
}

// FSST12: up to 4096 symbols, 12-bit codes, codes 0..255 = the bytes themselves (no escape).  The table here is the
// 3840 substrings of 2..8 bytes with the largest count x (length - 1) in a sample; compression is greedy longest match.
static Chunk enc_str_fsst12(const uint8_t *bytes, const uint64_t *off, uint32_t n_vec) {
  Chunk c; c.scheme = FLS_S_STR_FSST12;
  const size_t n = (size_t)n_vec * FLS_VEC;
  struct Key { uint64_t s; uint32_t l; bool operator==(const Key &o) const { return s == o.s && l == o.l; } };
  struct KeyHash { size_t operator()(const Key &k) const { return (size_t)((k.s * 0x9E3779B97F4A7C15ull) ^ (k.l * 0xC2B2AE3D27D4EB4Full)); } };
  auto key_at = [](const uint8_t *p, uint32_t l) { uint64_t s = 0; memcpy(&s, p, l); return Key{s, l}; };
  std::unordered_map<Key, uint64_t, KeyHash> cnt;
  const size_t total = off[n] - off[0], step = std::max<size_t>(1, total / 65536);
  for (size_t i = 0; i < n; i += step) {
    const uint8_t *p = bytes + off[i];
    const size_t len = off[i + 1] - off[i];
    for (size_t a = 0; a < len; ++a)
      for (uint32_t l = 2; l <= 8 && a + l <= len; ++l) cnt[key_at(p + a, l)]++;
  }
  std::vector<std::pair<uint64_t, Key>> cand;
  for (auto &kv : cnt) if (kv.second > 1) cand.push_back({kv.second * (kv.first.l - 1), kv.first});
  std::sort(cand.begin(), cand.end(), [](const std::pair<uint64_t, Key> &a, const std::pair<uint64_t, Key> &b) {
    return a.first != b.first ? a.first > b.first : (a.second.l != b.second.l ? a.second.l > b.second.l : a.second.s < b.second.s); });
  if (cand.size() > 4096 - 256) cand.resize(4096 - 256);
  std::vector<uint64_t> sym(256 + cand.size());
  std::vector<uint8_t> slen(256 + cand.size());
  std::unordered_map<Key, uint32_t, KeyHash> code;
  for (uint32_t b = 0; b < 256; ++b) { sym[b] = b; slen[b] = 1; }
  for (size_t k = 0; k < cand.size(); ++k) { sym[256 + k] = cand[k].second.s; slen[256 + k] = (uint8_t)cand[k].second.l; code[cand[k].second] = (uint32_t)(256 + k); }
  std::vector<uint32_t> dec_ends(n), cmp_ends(n);
  Bytes comp;
  for (size_t i = 0; i < n; ++i) {
    const uint8_t *p = bytes + off[i];
    const size_t len = off[i + 1] - off[i];
    std::vector<uint32_t> codes;
    for (size_t a = 0; a < len;) {
      uint32_t cd = p[a], adv = 1;
      for (uint32_t l = (uint32_t)std::min<size_t>(8, len - a); l >= 2; --l) {
        auto it = code.find(key_at(p + a, l));
        if (it != code.end()) { cd = it->second; adv = l; break; }
      }
      codes.push_back(cd);
      a += adv;
    }
    for (size_t k = 0; k + 1 < codes.size(); k += 2) {  // two codes in three bytes, LSB first
      const uint32_t w = codes[k] | (codes[k + 1] << 12);
      comp.push_back((uint8_t)w); comp.push_back((uint8_t)(w >> 8)); comp.push_back((uint8_t)(w >> 16));
    }
    if (codes.size() & 1) { const uint32_t w = codes.back(); comp.push_back((uint8_t)w); comp.push_back((uint8_t)(w >> 8)); }
    dec_ends[i] = (uint32_t)(off[i + 1] - off[0]);
    cmp_ends[i] = (uint32_t)comp.size();
  }
  Stream sd, sc;
  offsets_stream(sd, dec_ends, n_vec); sd.move_into(c);
  offsets_stream(sc, cmp_ends, n_vec); sc.move_into(c);
  Bytes tb;
  put_bytes(tb, sym.data(), sym.size() * 8);
  put_bytes(tb, slen.data(), slen.size());
  c.segs.push_back(std::move(tb));
  c.aux32 = (uint32_t)sym.size();
  c.aux64[0] = off[n] - off[0];
  c.aux64[1] = comp.size();
  pad_to(comp, 16);
  comp.resize(comp.size() + 16, 0);
  c.segs.push_back(std::move(comp));
  return c;
}

static bool enc_str_dict(const uint8_t *bytes, const uint64_t *off, uint32_t n_vec, uint32_t max_dict, bool fsst,
                         Chunk &c, int force_bw = -1) {
  size_t n = (size_t)n_vec * FLS_VEC;
  std::map<std::string, uint32_t> m;
  for (size_t i = 0; i < n; ++i) {
    m.emplace(std::string((const char *)bytes + off[i], off[i + 1] - off[i]), 0);
    if (m.size() > max_dict) return false;
  }
  uint32_t k = 0;
  for (auto &kv : m) kv.second = k++;
  c = Chunk(); c.scheme = fsst ? FLS_S_STR_FSST_DICT : FLS_S_STR_DICT; c.aux32 = (uint32_t)m.size();
  Stream s;
  std::vector<uint32_t> idx(FLS_VEC), vec_ep(n_vec + 1);
  uint64_t run = 0;
  for (uint32_t v = 0; v < n_vec; ++v) {
    vec_ep[v] = (uint32_t)run;
    for (uint32_t j = 0; j < FLS_VEC; ++j) {
      size_t i = (size_t)v * FLS_VEC + j;
      idx[j] = m[std::string((const char *)bytes + off[i], off[i + 1] - off[i])];
      run += off[i + 1] - off[i];
    }
    ffor_vec<uint32_t>(s, idx.data(), 0, 0, 0, force_bw);
  }
  vec_ep[n_vec] = (uint32_t)run;
  c.aux64[0] = run;
  s.move_into(c);
  // dictionary payload
  Bytes dbytes; std::vector<uint64_t> doff64(1, 0); std::vector<uint32_t> doff(1, 0);
  for (auto &kv : m) { put_bytes(dbytes, kv.first.data(), kv.first.size()); doff64.push_back(dbytes.size()); doff.push_back((uint32_t)dbytes.size()); }
  Bytes o, ve; put_bytes(ve, vec_ep.data(), vec_ep.size() * 4);
  if (!fsst) {
    put_bytes(o, doff.data(), doff.size() * 4);
    c.segs.push_back(std::move(o));
    dbytes.resize(dbytes.size() + 16, 0);
    c.segs.push_back(std::move(dbytes));
    c.segs.push_back(std::move(ve));
  } else {
    Fsst t = fsst_train(dbytes.data(), doff64.data(), m.size());
    Bytes comp; std::vector<uint32_t> coff(1, 0);
    for (size_t i = 0; i < m.size(); ++i) { t.compress(dbytes.data() + doff64[i], doff64[i + 1] - doff64[i], comp); coff.push_back((uint32_t)comp.size()); }
    put_bytes(o, coff.data(), coff.size() * 4);
    c.segs.push_back(std::move(o));
    comp.resize(comp.size() + 16, 0);
    c.segs.push_back(std::move(comp));
    c.segs.push_back(std::move(ve));
    Bytes tb; t.blob(tb); c.segs.push_back(std::move(tb));
    Bytes od; put_bytes(od, doff.data(), doff.size() * 4); c.segs.push_back(std::move(od));
  }
  return true;
}

static Chunk enc_str_auto(const uint8_t *bytes, const uint64_t *off, uint32_t n_vec, int force, int force_bw = -1) {
  size_t n = (size_t)n_vec * FLS_VEC;
  Chunk c;
  if (force == FLS_S_STR_UNCOMPRESSED) return enc_str_uncompressed(bytes, off, n_vec);
  if (force == FLS_S_STR_FSST) return enc_str_fsst(bytes, off, n_vec);
  if (force == FLS_S_STR_FSST12) return enc_str_fsst12(bytes, off, n_vec);
  if (force == FLS_S_STR_DICT || force == FLS_S_STR_FSST_DICT) {
    if (enc_str_dict(bytes, off, n_vec, 1u << 20, force == FLS_S_STR_FSST_DICT, c, force_bw)) return c;
    return enc_str_fsst(bytes, off, n_vec);
  }
  bool all_same = true;
  size_t l0 = off[1] - off[0];
  for (size_t i = 1; i < n && all_same; ++i)
    all_same = (off[i + 1] - off[i] == l0) && memcmp(bytes + off[i], bytes + off[0], l0) == 0;
  if (all_same || force == FLS_S_STR_CONSTANT) {
    c.scheme = FLS_S_STR_CONSTANT; c.aux32 = (uint32_t)l0; c.aux64[0] = l0 * n;
    Bytes b; put_bytes(b, bytes + off[0], l0); b.resize(b.size() + 16, 0);
    c.segs.push_back(std::move(b));
    return c;
  }
  Chunk best;
  if (enc_str_dict(bytes, off, n_vec, 4096, false, best)) {
    Chunk fd;
    if (best.aux32 >= 256 && enc_str_dict(bytes, off, n_vec, 4096, true, fd) && fd.bytes() < best.bytes()) best = std::move(fd);
    return best;
  }
  best = enc_str_fsst(bytes, off, n_vec);
  // (FSST12 is written only on request — flsw_force_scheme — so that the lineitem-shaped benchmark keeps its FSST column)
  { Chunk u = enc_str_uncompressed(bytes, off, n_vec); if (u.bytes() < best.bytes()) best = std::move(u); }
  return best;
}

// ------------------------------------------------------------------ file assembly
struct Writer {
  std::vector<std::string> names;
  std::vector<int> types;
  std::vector<int> force;     // per column: forced scheme or -1
  std::vector<int> force_bw;  // per column: minimum FFOR width or -1
  std::vector<int> force_ef;  // per column: forced ALP exponent * 32 + factor, or -1
  Bytes file;
  Bytes rg_footer;  // serialized row-group descriptors
  uint32_t n_rg = 0;
  uint64_t n_rows = 0;
  bool last_partial = false, finished = false;
  std::string err;
  std::vector<int> rg_schemes;                   // schemes chosen for the columns of the row group being appended
  std::map<uint32_t, FsstInject> inject_fsst;    // column -> streams for its next row group (test knobs)
  std::map<uint32_t, AlprdInject> inject_alprd;
};

static inline bool bit_at(const uint64_t *bits, size_t r) { return (bits[r >> 6] >> (r & 63u)) & 1u; }

template <typename T>
static Chunk enc_fixed(const void *src, uint32_t n_rows, uint32_t n_vec, int force, int force_bw, bool is_double,
                       const uint64_t *valid = nullptr) {
  std::vector<T> v((size_t)n_vec * FLS_VEC);
  memcpy(v.data(), src, (size_t)n_rows * sizeof(T));
  if (valid) {  // NULL rows take the value before them (the first non-NULL value for leading NULLs): cheap to encode
    size_t first = 0;
    while (first < n_rows && !bit_at(valid, first)) ++first;
    T prev = first < n_rows ? v[first] : T();
    for (size_t i = 0; i < n_rows; ++i) { if (bit_at(valid, i)) prev = v[i]; else v[i] = prev; }
  }
  for (size_t i = n_rows; i < v.size(); ++i) v[i] = v[n_rows - 1];  // pad the tail vector
  if (is_double) return sizeof(T) == 8 ? enc_double_auto((const double *)v.data(), n_vec, force)
                                       : enc_float_auto((const float *)v.data(), n_vec, force, force_bw);  // (is_double: floating point)
  return enc_int_auto<T>(v.data(), n_vec, force, force_bw);
}

static int append_rowgroup(Writer &w, uint32_t n_rows, const void *const *col_data, const uint64_t *const *str_off,
                           const uint64_t *const *validity = nullptr) {
  if (w.finished) { w.err = "writer already finished"; return 1; }
  if (w.last_partial) { w.err = "only the last row group may end in a partial vector (rows not a multiple of 1024)"; return 1; }
  if (n_rows == 0 || n_rows > FLS_RG_VECS * FLS_VEC) { w.err = "row group must hold 1..65536 rows"; return 1; }
  if (n_rows % FLS_VEC) w.last_partial = true;
  uint32_t n_vec = (n_rows + FLS_VEC - 1) / FLS_VEC;
  pad_to(w.file, FLS_RG_ALIGN);
  uint64_t rg_off = w.file.size();
  Bytes cc;  // column-chunk records
  w.rg_schemes.clear();
  for (size_t c = 0; c < w.types.size(); ++c) {
    Chunk ch;
    int t = w.types[c], f = w.force[c];
    const uint64_t *vb = validity ? validity[c] : nullptr;
    if (vb) {  // a bitmap without a cleared bit is not stored
      bool all = true;
      for (uint32_t r = 0; all && r < n_rows; ++r) all = bit_at(vb, r);
      if (all) vb = nullptr;
    }
    tl_force_alp_ef = c < w.force_ef.size() ? w.force_ef[c] : -1;
    tl_fsst_inject = nullptr; tl_alprd_inject = nullptr;
    FsstInject fi; AlprdInject ai;
    if (w.inject_fsst.count((uint32_t)c)) { fi = w.inject_fsst[(uint32_t)c]; w.inject_fsst.erase((uint32_t)c); tl_fsst_inject = &fi; f = FLS_S_STR_FSST; }
    if (w.inject_alprd.count((uint32_t)c)) { ai = w.inject_alprd[(uint32_t)c]; w.inject_alprd.erase((uint32_t)c); tl_alprd_inject = &ai; f = FLS_S_ALP_RD; }
    // EQUAL: a fixed-width column whose values (and NULLs) repeat an earlier column of the same type in this row group
    int equal_to = -1;
    if (fls_type_bits(t) && (f < 0 || f == FLS_S_EQUAL) && !vb)
      for (size_t e = 0; e < c && equal_to < 0; ++e)
        if (w.types[e] == t && !(validity && validity[e]) &&
            memcmp(col_data[e], col_data[c], (size_t)n_rows * fls_type_bits(t) / 8) == 0)
          equal_to = (int)e;
    if (equal_to >= 0 && w.rg_schemes.size() > (size_t)equal_to && w.rg_schemes[equal_to] != FLS_S_EQUAL) {
      ch.scheme = FLS_S_EQUAL;
      ch.aux32 = (uint32_t)equal_to;
    } else
    switch (fls_type_bits(t)) {
      case 8: ch = enc_fixed<uint8_t>(col_data[c], n_rows, n_vec, f == FLS_S_EQUAL ? -1 : f, w.force_bw[c], false, vb); break;
      case 16: ch = enc_fixed<uint16_t>(col_data[c], n_rows, n_vec, f == FLS_S_EQUAL ? -1 : f, w.force_bw[c], false, vb); break;
      case 32: ch = enc_fixed<uint32_t>(col_data[c], n_rows, n_vec, f == FLS_S_EQUAL ? -1 : f, w.force_bw[c], t == FLS_T_FLOAT, vb); break;
      case 64: ch = enc_fixed<uint64_t>(col_data[c], n_rows, n_vec, f == FLS_S_EQUAL ? -1 : f, w.force_bw[c], t == FLS_T_DOUBLE, vb); break;
      default: {
        if (t != FLS_T_STR && t != FLS_T_BYTE_ARRAY) { w.err = "unsupported column type"; return 1; }
        std::vector<uint64_t> off((size_t)n_vec * FLS_VEC + 1);
        memcpy(off.data(), str_off[c], (size_t)(n_rows + 1) * 8);
        for (size_t i = n_rows + 1; i < off.size(); ++i) off[i] = off[n_rows];
        if (off[n_rows] - off[0] >= (1ull << 31)) { w.err = "row group string payload exceeds 2 GiB"; return 1; }
        ch = enc_str_auto((const uint8_t *)col_data[c], off.data(), n_vec, f, w.force_bw[c]);
      }
    }
    if (vb) {  // validity bitmap: the chunk's last segment, 128 bytes per vector, padding rows cleared
      Bytes bm((size_t)n_vec * 128, 0);
      for (uint32_t r = 0; r < n_rows; ++r)
        if (bit_at(vb, r)) bm[r >> 3] |= (uint8_t)(1u << (r & 7u));
      ch.segs.push_back(std::move(bm));
    }
    if (ch.segs.size() > FLS_MAX_SEGS) { w.err = "internal: too many segments"; return 1; }
    w.rg_schemes.push_back(ch.scheme);
    put<uint8_t>(cc, ch.scheme); put<uint8_t>(cc, (uint8_t)ch.segs.size()); put<uint16_t>(cc, vb ? FLS_CHUNK_VALIDITY : 0);
    put<uint32_t>(cc, ch.aux32); put<uint64_t>(cc, ch.aux64[0]); put<uint64_t>(cc, ch.aux64[1]);
    for (auto &sg : ch.segs) {
      pad_to(w.file, FLS_SEG_ALIGN);
      put<uint64_t>(cc, w.file.size() - rg_off); put<uint64_t>(cc, sg.size());
      put_bytes(w.file, sg.data(), sg.size());
    }
  }
  pad_to(w.file, FLS_SEG_ALIGN);
  fls_rowgroup rg = {rg_off, w.file.size() - rg_off, n_vec, n_rows};
  put(w.rg_footer, rg);
  put_bytes(w.rg_footer, cc.data(), cc.size());
  w.n_rg++; w.n_rows += n_rows;
  return 0;
}

static int finish(Writer &w) {
  if (w.finished) return 0;
  pad_to(w.file, FLS_RG_ALIGN);
  uint64_t foot_off = w.file.size();
  Bytes f;
  put<uint32_t>(f, FLS_FOOTER_VERSION); put<uint32_t>(f, (uint32_t)w.types.size());
  put<uint32_t>(f, w.n_rg); put<uint32_t>(f, FLS_RG_VECS); put<uint64_t>(f, w.n_rows);
  for (size_t c = 0; c < w.types.size(); ++c) {
    put<uint8_t>(f, (uint8_t)w.types[c]); put<uint8_t>(f, 0); put<uint16_t>(f, (uint16_t)w.names[c].size());
    put_bytes(f, w.names[c].data(), w.names[c].size());
  }
  pad_to(f, 8);
  put_bytes(f, w.rg_footer.data(), w.rg_footer.size());
  put_bytes(w.file, f.data(), f.size());
  put<uint64_t>(w.file, foot_off); put<uint64_t>(w.file, (uint64_t)f.size()); put<uint64_t>(w.file, FLS_MAGIC);
  w.finished = true;
  return 0;
}

}  // namespace flsw

// ------------------------------------------------------------------ C API (see include/flsgpu.h, "writer" section)
extern "C" {
typedef struct flsw_writer flsw_writer;

flsw_writer *flsw_create(uint32_t n_cols, const char *const *names, const int32_t *types) {
  auto *w = new flsw::Writer();
  for (uint32_t i = 0; i < n_cols; ++i) {
    w->names.push_back(names[i]); w->types.push_back(types[i]); w->force.push_back(-1); w->force_bw.push_back(-1);
  }
  flsw::put<uint64_t>(w->file, FLS_MAGIC);
  flsw::put_bytes(w->file, FLS_VERSION_BYTES, 8);
  return (flsw_writer *)w;
}
int flsw_force_scheme(flsw_writer *h, uint32_t col, int32_t scheme, int32_t min_bw) {
  auto *w = (flsw::Writer *)h;
  if (col >= w->types.size()) return 1;
  w->force[col] = scheme; w->force_bw[col] = min_bw;
  return 0;
}
int flsw_append_rowgroup(flsw_writer *h, uint32_t n_rows, const void *const *col_data, const uint64_t *const *str_off) {
  return flsw::append_rowgroup(*(flsw::Writer *)h, n_rows, col_data, str_off);
}
int flsw_append_rowgroup_v(flsw_writer *h, uint32_t n_rows, const void *const *col_data, const uint64_t *const *str_off,
                           const uint64_t *const *validity) {
  return flsw::append_rowgroup(*(flsw::Writer *)h, n_rows, col_data, str_off, validity);
}
int flsw_finish(flsw_writer *h) { return flsw::finish(*(flsw::Writer *)h); }
const void *flsw_data(flsw_writer *h) { return ((flsw::Writer *)h)->file.data(); }
uint64_t flsw_size(flsw_writer *h) { return ((flsw::Writer *)h)->file.size(); }
const char *flsw_error(flsw_writer *h) { return ((flsw::Writer *)h)->err.c_str(); }
int flsw_write_file(flsw_writer *h, const char *path) {
  auto *w = (flsw::Writer *)h;
  if (!w->finished) flsw::finish(*w);
  FILE *f = fopen(path, "wb");
  if (!f) { w->err = std::string("cannot open for writing: ") + path; return 1; }
  size_t n = fwrite(w->file.data(), 1, w->file.size(), f);
  fclose(f);
  if (n != w->file.size()) { w->err = "short write"; return 1; }
  return 0;
}
void flsw_destroy(flsw_writer *h) { delete (flsw::Writer *)h; }
// Test knobs: pin a column's ALP exponent/factor; expose the scalar ALP encode used by the writer.
int flsw_force_alp_ef(flsw_writer *h, uint32_t col, int32_t e, int32_t f) {
  flsw::Writer *w = (flsw::Writer *)h;
  if (col >= w->types.size() || e < 0 || e > 18 || f < 0 || f > e) return 1;
  w->force_ef.resize(w->types.size(), -1);
  w->force_ef[col] = e * 32 + f;
  return 0;
}
int64_t flsw_alp_encode_value(double v, int32_t e, int32_t f) { return flsw::alp_encode(v, e, f); }
int64_t flsw_alpf_encode_value(float v, int32_t e, int32_t f) { return flsw::alpf_encode(v, e, f); }
// Test knobs: the next row group appended for `col` uses these externally produced streams verbatim (the pointers
// must stay valid until that flsw_append_rowgroup returns).
int flsw_inject_fsst(flsw_writer *h, uint32_t col, const uint8_t *symtab2304, const uint8_t *comp, const uint64_t *comp_off,
                     uint32_t n_strings) {
  flsw::Writer *w = (flsw::Writer *)h;
  if (col >= w->types.size() || fls_type_bits(w->types[col]) != 0 || !symtab2304 || !comp_off || n_strings > FLS_RG_VECS * FLS_VEC) return 1;
  w->inject_fsst[col] = flsw::FsstInject{symtab2304, comp, comp_off, n_strings};
  return 0;
}
int flsw_inject_alprd(flsw_writer *h, uint32_t col, uint32_t rbw, uint32_t lbw, uint32_t ndict, const uint16_t *dict8,
                      uint32_t n_vec, const uint16_t *left, const uint64_t *right, const uint32_t *nexc, const uint16_t *xpos,
                      const uint16_t *xleft) {
  flsw::Writer *w = (flsw::Writer *)h;
  if (col >= w->types.size() || w->types[col] != FLS_T_DOUBLE || rbw == 0 || rbw > 63 || lbw > 16 || ndict > 8 || n_vec == 0 || n_vec > FLS_RG_VECS)
    return 1;
  for (uint32_t v = 0; v < n_vec; ++v) if (nexc[v] > FLS_VEC) return 1;
  w->inject_alprd[col] = flsw::AlprdInject{rbw, lbw, ndict, n_vec, dict8, left, right, nexc, xpos, xleft};
  return 0;
}

// Build a new image holding `n_total` row groups by repeating the (full) row groups of `image` round-robin.
// Every repeat is a separate copy of the bytes.  Returns a finished writer (flsw_data/flsw_size) or NULL.
flsw_writer *flsw_tile(const void *image, uint64_t size, uint32_t n_total) {
  FlsTable t;
  if (!fls_parse((const uint8_t *)image, size, t).empty() || t.n_rg == 0) return nullptr;
  for (auto &r : t.rgs)
    if (r.n_tuples != FLS_RG_VECS * FLS_VEC) return nullptr;  // only full row groups can be repeated
  flsw::Writer *w = new flsw::Writer();
  w->names = t.names;
  w->types = t.types;
  w->force.assign(t.n_cols, -1);
  w->force_bw.assign(t.n_cols, -1);
  uint64_t total = t.rgs[0].off;
  for (uint32_t i = 0; i < n_total; ++i) total += flsw::align_up(t.rgs[i % t.n_rg].size, FLS_RG_ALIGN);
  w->file.reserve(total + 4096 + (size - t.data_end));
  w->file.assign((const uint8_t *)image, (const uint8_t *)image + t.rgs[0].off);
  for (uint32_t i = 0; i < n_total; ++i) {
    const uint32_t g = i % t.n_rg;
    const fls_rowgroup &r = t.rgs[g];
    flsw::pad_to(w->file, FLS_RG_ALIGN);
    fls_rowgroup nr = {w->file.size(), r.size, r.n_vec, r.n_tuples};
    flsw::put_bytes(w->file, (const uint8_t *)image + r.off, r.size);
    flsw::put(w->rg_footer, nr);
    for (uint32_t c = 0; c < t.n_cols; ++c) {
      const fls_colchunk &k = t.chunk(g, c);
      flsw::put<uint8_t>(w->rg_footer, k.scheme); flsw::put<uint8_t>(w->rg_footer, k.n_seg);
      flsw::put<uint16_t>(w->rg_footer, k.flags); flsw::put<uint32_t>(w->rg_footer, k.aux32);
      flsw::put<uint64_t>(w->rg_footer, k.aux64[0]); flsw::put<uint64_t>(w->rg_footer, k.aux64[1]);
      for (uint32_t sg = 0; sg < k.n_seg; ++sg) {
        flsw::put<uint64_t>(w->rg_footer, k.seg[sg].off); flsw::put<uint64_t>(w->rg_footer, k.seg[sg].size);
      }
    }
    w->n_rg++;
    w->n_rows += r.n_tuples;
  }
  flsw::finish(*w);
  return (flsw_writer *)w;
}
}
