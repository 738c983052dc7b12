/* fls_oracle.c — CPU ORACLE for the FLS read/decode path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * The product (libflsgpu.so) never links or calls it.
 *
 * What it restates: what FastLanes `RowgroupReader::materialize()` does for the reference at
 * /root/reference/src/fastlanes_facade.cpp:33-48 (connect -> read_fls -> get_rowgroup_reader -> materialize),
 * i.e. footer parse + per-vector decode of every column into plain arrays, plus the value-type contract
 * the facade copies out at /root/reference/src/fastlanes_facade.cpp:125-172 (i32/i64/float/double/strings).
 *
 * PARITY STATUS: the FastLanes library itself (cwida/FastLanes @ 928fdb82, /root/reference/.SUBMODULES.json)
 * is absent from the reference checkout, so the container + FFOR/DELTA/RLE/DICT layouts are restated from
 * public knowledge and are **parity unpinned** against upstream bytes.  The ALP, ALP-RD
 * and FSST arithmetic IS pinned: tests/golden/ holds vectors produced by DuckDB's in-tree twins of those
 * algorithms (oracle/ref_twin.cpp compiled against /root/reference/duckdb), and tests/test_oracle_golden.py
 * checks this file against them.
 *
 * Plain C99, scalar loops, written independently of duckdb-fastlane_b200/csrc (constants restated here).
 * Build: see oracle/Makefile (-O3 -fno-fast-math -ffp-contract=off).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define O_MAGIC 0x656E614C74736146ULL
#define O_VEC 1024u

enum { T_BOOL = 1, T_I8, T_I16, T_I32, T_I64, T_U8, T_U16, T_U32, T_U64, T_FLT, T_DBL, T_STR, T_DATE, T_BYTES };
enum { S_UNCOMP = 0, S_CONST, S_FFOR, S_DELTA, S_RLE, S_DICT, S_ALP, S_ALPRD,
       S_STR_UNCOMP, S_STR_CONST, S_STR_DICT, S_STR_FSST, S_STR_FSST_DICT,
       S_FFOR_PATCH, S_CROSS_RLE, S_EQUAL, S_STR_FSST12 };

static const uint8_t ORDER[8] = {0, 4, 2, 6, 1, 5, 3, 7};

typedef struct { uint64_t off, size; } oseg;
typedef struct { uint8_t scheme, n_seg; uint16_t flags; uint32_t aux32; uint64_t aux64[2]; oseg seg[8]; } ochunk;
typedef struct { uint64_t off, size; uint32_t n_vec, n_tuples; ochunk *cols; } org;
typedef struct {
  const uint8_t *img; uint64_t size;
  uint32_t n_cols, n_rg; uint64_t n_rows;
  uint8_t *types; char **names;
  org *rg;
} flso_table;

static unsigned type_bits(int t) {
  switch (t) {
    case T_BOOL: case T_I8: case T_U8: return 8;
    case T_I16: case T_U16: return 16;
    case T_I32: case T_U32: case T_FLT: case T_DATE: return 32;
    case T_I64: case T_U64: case T_DBL: return 64;
    default: return 0;
  }
}

#define RD(T, p) (*(const T *)(p))

/* ---- footer parse (restates Connection::read_fls, call site fastlanes_facade.cpp:34) */
int flso_open(const void *image, uint64_t size, flso_table **out) {
  const uint8_t *b = (const uint8_t *)image;
  if (size < 40 || RD(uint64_t, b) != O_MAGIC || RD(uint64_t, b + size - 8) != O_MAGIC) return 1;
  uint64_t foff = RD(uint64_t, b + size - 24), fsz = RD(uint64_t, b + size - 16);
  if (foff + fsz + 24 != size) return 2;
  const uint8_t *p = b + foff;
  if (RD(uint32_t, p) != 1) return 3;
  flso_table *t = (flso_table *)calloc(1, sizeof(*t));
  t->img = b; t->size = size;
  t->n_cols = RD(uint32_t, p + 4); t->n_rg = RD(uint32_t, p + 8); t->n_rows = RD(uint64_t, p + 16);
  p += 24;
  t->types = (uint8_t *)calloc(t->n_cols, 1);
  t->names = (char **)calloc(t->n_cols, sizeof(char *));
  for (uint32_t c = 0; c < t->n_cols; ++c) {
    t->types[c] = p[0];
    uint16_t nl = RD(uint16_t, p + 2);
    t->names[c] = (char *)calloc(nl + 1, 1);
    memcpy(t->names[c], p + 4, nl);
    p += 4 + nl;
  }
  p = b + foff + ((uint64_t)(p - (b + foff)) + 7) / 8 * 8;
  t->rg = (org *)calloc(t->n_rg, sizeof(org));
  for (uint32_t g = 0; g < t->n_rg; ++g) {
    org *r = &t->rg[g];
    r->off = RD(uint64_t, p); r->size = RD(uint64_t, p + 8);
    r->n_vec = RD(uint32_t, p + 16); r->n_tuples = RD(uint32_t, p + 20);
    p += 24;
    r->cols = (ochunk *)calloc(t->n_cols, sizeof(ochunk));
    for (uint32_t c = 0; c < t->n_cols; ++c) {
      ochunk *k = &r->cols[c];
      k->scheme = p[0]; k->n_seg = p[1]; k->flags = RD(uint16_t, p + 2);
      k->aux32 = RD(uint32_t, p + 4); k->aux64[0] = RD(uint64_t, p + 8); k->aux64[1] = RD(uint64_t, p + 16);
      p += 24;
      for (uint32_t s = 0; s < k->n_seg; ++s) { k->seg[s].off = RD(uint64_t, p); k->seg[s].size = RD(uint64_t, p + 8); p += 16; }
    }
  }
  *out = t;
  return 0;
}

void flso_close(flso_table *t) {
  if (!t) return;
  for (uint32_t c = 0; c < t->n_cols; ++c) free(t->names[c]);
  for (uint32_t g = 0; g < t->n_rg; ++g) free(t->rg[g].cols);
  free(t->names); free(t->types); free(t->rg); free(t);
}
uint32_t flso_num_columns(const flso_table *t) { return t->n_cols; }
uint32_t flso_num_rowgroups(const flso_table *t) { return t->n_rg; }
uint64_t flso_num_rows(const flso_table *t) { return t->n_rows; }
int flso_column_type(const flso_table *t, uint32_t c) { return t->types[c]; }
const char *flso_column_name(const flso_table *t, uint32_t c) { return t->names[c]; }
uint32_t flso_rowgroup_vectors(const flso_table *t, uint32_t g) { return t->rg[g].n_vec; }
uint32_t flso_rowgroup_rows(const flso_table *t, uint32_t g) { return t->rg[g].n_tuples; }
int flso_chunk_scheme(const flso_table *t, uint32_t g, uint32_t c) { return t->rg[g].cols[c].scheme; }
uint64_t flso_chunk_bytes(const flso_table *t, uint32_t g, uint32_t c) {
  uint64_t s = 0;
  for (uint32_t i = 0; i < t->rg[g].cols[c].n_seg; ++i) s += t->rg[g].cols[c].seg[i].size;
  return s;
}
uint64_t flso_chunk_payload(const flso_table *t, uint32_t g, uint32_t c) { return t->rg[g].cols[c].aux64[0]; }
/* file extent of segment s of a chunk (tests that damage files on purpose); 0 = no such segment */
int flso_chunk_segment(const flso_table *t, uint32_t g, uint32_t c, uint32_t s, uint64_t *off, uint64_t *size) {
  const ochunk *k = &t->rg[g].cols[c];
  if (s >= k->n_seg) return 0;
  *off = t->rg[g].off + k->seg[s].off; *size = k->seg[s].size;
  return 1;
}

/* ---- per-vector primitives, instantiated for T in {u8,u16,u32,u64} ------------------------------------
 * unffor : interleaved bit-unpack + frame of reference: value r of lane l occupies bits
 *          [rW,(r+1)W) of the lane's LSB-first stream; word k of lane l is in[k*L+l]; out[r*L+l] = bits + base.
 * undelta: unified transposed layout (App. A.3): lane l walks packed positions
 *          p(r,l) = ORDER[r/8]*16 + (r%8)*128 + l, accumulating from bases[l], and writes originals
 *          start(l)+r with start(l) = 64*(l%16) + 8*ORDER[l/16].
 */
#define DEF_PRIMS(T, TB)                                                                              \
  static void unffor_##TB(const T *in, unsigned W, T base, T *out) {                                  \
    enum { L = 1024 / TB };                                                                           \
    if (W == 0) { for (unsigned i = 0; i < O_VEC; ++i) out[i] = base; return; }                       \
    const T mask = (W == TB) ? (T) ~(T)0 : (T)((((T)1) << W) - 1);                                    \
    for (unsigned r = 0; r < TB; ++r) {                                                               \
      unsigned bit = r * W, k = bit / TB, s = bit % TB;                                               \
      const T *w0 = in + k * L, *w1 = in + (k + 1) * L;                                               \
      T *o = out + r * L;                                                                             \
      if (s + W <= TB) { for (unsigned l = 0; l < L; ++l) o[l] = (T)(((w0[l] >> s) & mask) + base); } \
      else { for (unsigned l = 0; l < L; ++l) o[l] = (T)((((w0[l] >> s) | (T)(w1[l] << (TB - s))) & mask) + base); } \
    }                                                                                                 \
  }                                                                                                   \
  static void undelta_##TB(const T *d, const T *bases, T *out) {                                      \
    enum { L = 1024 / TB };                                                                           \
    for (unsigned l = 0; l < L; ++l) {                                                                \
      T acc = bases[l];                                                                               \
      unsigned st = 64u * (l % 16u) + 8u * ORDER[l / 16u];                                            \
      for (unsigned r = 0; r < TB; ++r) {                                                             \
        acc = (T)(acc + d[ORDER[r >> 3] * 16u + (r & 7u) * 128u + l]);                                \
        out[st + r] = acc;                                                                            \
      }                                                                                               \
    }                                                                                                 \
  }
DEF_PRIMS(uint8_t, 8)
DEF_PRIMS(uint16_t, 16)
DEF_PRIMS(uint32_t, 32)
DEF_PRIMS(uint64_t, 64)

typedef struct { uint32_t ep, meta; uint64_t base; } ovhdr;
#define M_BW(m) ((m) & 0x7Fu)
#define M_E(m) (((m) >> 8) & 0x1Fu)
#define M_F(m) (((m) >> 13) & 0x1Fu)
#define M_N(m) (((m) >> 18) & 0x7FFu)
#define PAD16(x) (((x) + 15u) & ~(uint64_t)15u)

/* ---- integer column chunk -> out[n_vec*1024]  (restates materialize()'s per-vector loop, call site
 *      fastlanes_facade.cpp:48) */
#define DEF_INT_CHUNK(T, TB)                                                                              \
  static int chunk_int_##TB(const uint8_t *rgp, const ochunk *k, uint32_t n_vec, T *out) {                \
    const ovhdr *h = (const ovhdr *)(rgp + k->seg[0].off);                                                \
    const uint8_t *data = rgp + k->seg[1].off;                                                            \
    T tmp[O_VEC];                                                                                         \
    switch (k->scheme) {                                                                                  \
      case S_UNCOMP: memcpy(out, rgp + k->seg[0].off, (size_t)n_vec * O_VEC * sizeof(T)); return 0;       \
      case S_CONST: for (size_t i = 0; i < (size_t)n_vec * O_VEC; ++i) out[i] = (T)k->aux64[0]; return 0; \
      case S_FFOR:                                                                                        \
        for (uint32_t v = 0; v < n_vec; ++v)                                                              \
          unffor_##TB((const T *)(data + 16ull * h[v].ep), M_BW(h[v].meta), (T)h[v].base, out + v * O_VEC); \
        return 0;                                                                                         \
      case S_DELTA:                                                                                       \
        for (uint32_t v = 0; v < n_vec; ++v) {                                                            \
          const uint8_t *p = data + 16ull * h[v].ep;                                                      \
          unffor_##TB((const T *)(p + 128), M_BW(h[v].meta), (T)h[v].base, tmp);                          \
          undelta_##TB(tmp, (const T *)p, out + v * O_VEC);                                               \
        }                                                                                                 \
        return 0;                                                                                         \
      case S_RLE:                                                                                         \
        for (uint32_t v = 0; v < n_vec; ++v) {                                                            \
          const uint8_t *p = data + 16ull * h[v].ep;                                                      \
          uint32_t n_runs = M_N(h[v].meta);                                                               \
          const T *runs = (const T *)p;                                                                   \
          const uint8_t *q = p + PAD16((uint64_t)n_runs * sizeof(T));                                     \
          uint16_t d[O_VEC], idx[O_VEC];                                                                  \
          unffor_16((const uint16_t *)(q + 128), M_BW(h[v].meta), 0, d);                                  \
          undelta_16(d, (const uint16_t *)q, idx);                                                        \
          for (unsigned i = 0; i < O_VEC; ++i) out[v * O_VEC + i] = runs[idx[i]];                         \
        }                                                                                                 \
        return 0;                                                                                         \
      case S_FFOR_PATCH: /* SLPATCH / FREQUENCY: unffor, then the exception list overwrites */ \
        for (uint32_t v = 0; v < n_vec; ++v) {                                                            \
          const uint8_t *p = data + 16ull * h[v].ep;                                                      \
          const uint32_t bw = M_BW(h[v].meta), nx = M_N(h[v].meta);                                       \
          unffor_##TB((const T *)p, bw, (T)h[v].base, out + v * O_VEC);                                   \
          const T *xv = (const T *)(p + 128ull * bw);                                                     \
          const uint16_t *xp = (const uint16_t *)(p + 128ull * bw + PAD16((uint64_t)nx * sizeof(T)));     \
          for (uint32_t i = 0; i < nx; ++i) out[v * O_VEC + (xp[i] & 1023u)] = xv[i];                     \
        }                                                                                                 \
        return 0;                                                                                         \
      case S_CROSS_RLE: { /* runs over the whole row group: values + cumulative ends */                   \
        const T *vals = (const T *)(rgp + k->seg[0].off);                                                 \
        const uint32_t *ends = (const uint32_t *)(rgp + k->seg[1].off);                                   \
        size_t at = 0, n = (size_t)n_vec * O_VEC;                                                         \
        for (uint32_t r = 0; r < k->aux32 && at < n; ++r)                                                 \
          for (; at < ends[r] && at < n; ++at) out[at] = vals[r];                                         \
        for (; at < n; ++at) out[at] = k->aux32 ? vals[k->aux32 - 1] : (T)0;                              \
        return 0;                                                                                         \
      }                                                                                                   \
      case S_DICT: {                                                                                      \
        const T *dict = (const T *)(rgp + k->seg[2].off);                                                 \
        uint32_t idx[O_VEC];                                                                              \
        for (uint32_t v = 0; v < n_vec; ++v) {                                                            \
          unffor_32((const uint32_t *)(data + 16ull * h[v].ep), M_BW(h[v].meta), (uint32_t)h[v].base, idx); \
          for (unsigned i = 0; i < O_VEC; ++i) out[v * O_VEC + i] = dict[idx[i]];                         \
        }                                                                                                 \
        return 0;                                                                                         \
      }                                                                                                   \
    }                                                                                                     \
    return 10;                                                                                            \
  }
DEF_INT_CHUNK(uint8_t, 8)
DEF_INT_CHUNK(uint16_t, 16)
DEF_INT_CHUNK(uint32_t, 32)
DEF_INT_CHUNK(uint64_t, 64)

/* ---- ALP: decode = (double)(int64)(bits + for) * FACT[f] * FRAC[e], then patch exceptions.
 *      duckdb/src/include/duckdb/storage/compression/alp/algorithm/alp.hpp:129-134 (DecodeValue) and
 *      :376-404 (Decompress: unFOR, decode, patch); constants alp_constants.hpp:43-61,112-132.
 *      FACT is an int64 converted to double at use; two separate IEEE multiplies, left to right. */
static const int64_t A_FACT[] = {1LL, 10LL, 100LL, 1000LL, 10000LL, 100000LL, 1000000LL, 10000000LL, 100000000LL,
                                 1000000000LL, 10000000000LL, 100000000000LL, 1000000000000LL, 10000000000000LL,
                                 100000000000000LL, 1000000000000000LL, 10000000000000000LL, 100000000000000000LL,
                                 1000000000000000000LL};
static const double A_FRAC[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001, 0.00000001,
                                0.000000001, 0.0000000001, 0.00000000001, 0.000000000001, 0.0000000000001,
                                0.00000000000001, 0.000000000000001, 0.0000000000000001, 0.00000000000000001,
                                0.000000000000000001, 0.0000000000000000001, 0.00000000000000000001};

double flso_alp_decode_value(int64_t n, unsigned e, unsigned f) {
  double t = (double)n * (double)A_FACT[f];
  return t * A_FRAC[e];
}

static int chunk_alp(const uint8_t *rgp, const ochunk *k, uint32_t n_vec, double *out) {
  const ovhdr *h = (const ovhdr *)(rgp + k->seg[0].off);
  const uint8_t *data = rgp + k->seg[1].off;
  uint64_t tmp[O_VEC];
  for (uint32_t v = 0; v < n_vec; ++v) {
    uint32_t m = h[v].meta, W = M_BW(m), e = M_E(m), f = M_F(m), nx = M_N(m);
    const uint8_t *p = data + 16ull * h[v].ep;
    unffor_64((const uint64_t *)p, W, h[v].base, tmp);
    double *o = out + (size_t)v * O_VEC;
    for (unsigned i = 0; i < O_VEC; ++i) o[i] = flso_alp_decode_value((int64_t)tmp[i], e, f);
    const double *xv = (const double *)(p + 128ull * W);
    const uint16_t *xp = (const uint16_t *)(p + 128ull * W + PAD16((uint64_t)nx * 8));
    for (uint32_t i = 0; i < nx; ++i) o[xp[i]] = xv[i];
  }
  return 0;
}

/* ---- ALP on FLOAT columns: AlpTypedConstants<float> (alp_constants.hpp:66-83) — float arithmetic, exponents 0..10;
 *      decode = (float)(int64)n * (float)FACT[f] * FRAC[e] (alp.hpp:129-134 with T = float).  The stream is u32 FFOR,
 *      the vector header's base the (sign-extended) frame of reference. */
static const float AF_FRAC[] = {1.0F, 0.1F, 0.01F, 0.001F, 0.0001F, 0.00001F, 0.000001F, 0.0000001F, 0.00000001F,
                                0.000000001F, 0.0000000001F};
float flso_alpf_decode_value(int64_t n, unsigned e, unsigned f) {
  volatile float a = (float)n * (float)A_FACT[f > 18 ? 18 : f];
  volatile float b = a * AF_FRAC[e > 10 ? 10 : e];
  return b;
}
static int chunk_alp32(const uint8_t *rgp, const ochunk *k, uint32_t n_vec, float *out) {
  const ovhdr *h = (const ovhdr *)(rgp + k->seg[0].off);
  const uint8_t *data = rgp + k->seg[1].off;
  uint32_t tmp[O_VEC];
  for (uint32_t v = 0; v < n_vec; ++v) {
    uint32_t m = h[v].meta, W = M_BW(m), e = M_E(m), f = M_F(m), nx = M_N(m);
    const uint8_t *p = data + 16ull * h[v].ep;
    unffor_32((const uint32_t *)p, W, 0, tmp);
    float *o = out + (size_t)v * O_VEC;
    for (unsigned i = 0; i < O_VEC; ++i) o[i] = flso_alpf_decode_value((int64_t)h[v].base + (int64_t)tmp[i], e, f);
    const float *xv = (const float *)(p + 128ull * W);
    const uint16_t *xp = (const uint16_t *)(p + 128ull * W + PAD16((uint64_t)nx * 4));
    for (uint32_t i = 0; i < nx; ++i) o[xp[i] & 1023u] = xv[i];
  }
  return 0;
}

/* ---- ALP-RD: out = (dict[left] << rbw) | right, then patch left-part exceptions.
 *      duckdb/.../alprd/algorithm/alprd.hpp:206-233. */
void flso_alprd_decode(const uint16_t *dict, const uint16_t *left, const uint64_t *right, uint32_t rbw, uint32_t n,
                       const uint16_t *xpos, const uint16_t *xleft, uint32_t nx, uint64_t *o) {
  for (uint32_t i = 0; i < n; ++i) o[i] = ((uint64_t)dict[left[i] & 7] << rbw) | right[i];
  for (uint32_t i = 0; i < nx; ++i) o[xpos[i]] = ((uint64_t)xleft[i] << rbw) | right[xpos[i]];
}

static int chunk_alprd(const uint8_t *rgp, const ochunk *k, uint32_t n_vec, uint64_t *out) {
  const ovhdr *h = (const ovhdr *)(rgp + k->seg[0].off);
  const uint8_t *data = rgp + k->seg[1].off;
  uint32_t rbw = k->aux32 & 0xFF, lbw = (k->aux32 >> 8) & 0xFF;
  uint16_t dict[8];
  memcpy(dict, k->aux64, 16);
  uint64_t right[O_VEC]; uint16_t left[O_VEC];
  for (uint32_t v = 0; v < n_vec; ++v) {
    uint32_t nx = M_N(h[v].meta);
    const uint8_t *p = data + 16ull * h[v].ep;
    unffor_64((const uint64_t *)p, rbw, 0, right);
    unffor_16((const uint16_t *)(p + 128ull * rbw), lbw, 0, left);
    uint64_t *o = out + (size_t)v * O_VEC;
    const uint16_t *xl = (const uint16_t *)(p + 128ull * rbw + 128ull * lbw);
    const uint16_t *xp = (const uint16_t *)((const uint8_t *)xl + PAD16((uint64_t)nx * 2));
    flso_alprd_decode(dict, left, right, rbw, O_VEC, xp, xl, nx, o);
  }
  return 0;
}

/* ---- FSST: each code byte c<255 expands to symbol[c] (len[c] bytes, little-endian in a u64); 255 escapes
 *      one literal byte.  duckdb/third_party/fsst/fsst.h:93-104 (decoder struct), :162-241 (decompress).
 *      Symbol-table blob in the file: u64 symbol[256] | u8 len[256]. */
uint64_t flso_fsst_decode(const uint8_t *symtab, const uint8_t *in, uint64_t n_in, uint8_t *out) {
  const uint64_t *sym = (const uint64_t *)symtab;
  const uint8_t *len = symtab + 2048;
  uint64_t o = 0;
  for (uint64_t i = 0; i < n_in;) {
    uint8_t c = in[i++];
    if (c == 255) { out[o++] = in[i++]; }
    else { uint64_t s = sym[c]; for (unsigned b = 0; b < len[c]; ++b) out[o++] = (uint8_t)(s >> (8 * b)); }
  }
  return o;
}

/* ---- FSST12 (upstream libfsst's fsst12.h is not in the reference tree: restated, UNVERIFIED): 12-bit
 *      codes packed LSB first, two codes per three bytes; L compressed bytes hold floor(2 L / 3) codes; code c expands to
 *      symbol[c] (len[c] bytes); codes 0..255 are the bytes themselves, there is no escape.  Table blob: u64 symbol[n] | u8
 *      len[n]. */
uint64_t flso_fsst12_decode(const uint8_t *symtab, uint32_t n_sym, const uint8_t *in, uint64_t n_in, uint8_t *out) {
  const uint64_t *sym = (const uint64_t *)symtab;
  const uint8_t *len = symtab + 8ull * n_sym;
  uint64_t o = 0, n_codes = 2 * n_in / 3;
  for (uint64_t k = 0; k < n_codes; ++k) {
    const uint8_t *p = in + 3 * (k / 2);
    uint32_t c = (k & 1) ? ((uint32_t)p[1] >> 4) | ((uint32_t)p[2] << 4) : (uint32_t)p[0] | (((uint32_t)p[1] & 15u) << 8);
    if (c >= n_sym) c = 0;
    uint64_t s = sym[c];
    for (unsigned b = 0; b < len[c] && b < 8; ++b) out[o++] = (uint8_t)(s >> (8 * b));
  }
  return o;
}

/* offsets stream: DELTA u32 over END offsets -> ends[n_vec*1024] */
static void ends_stream(const uint8_t *rgp, const ochunk *k, unsigned s0, uint32_t n_vec, uint32_t *ends) {
  const ovhdr *h = (const ovhdr *)(rgp + k->seg[s0].off);
  const uint8_t *data = rgp + k->seg[s0 + 1].off;
  uint32_t tmp[O_VEC];
  for (uint32_t v = 0; v < n_vec; ++v) {
    const uint8_t *p = data + 16ull * h[v].ep;
    unffor_32((const uint32_t *)(p + 128), M_BW(h[v].meta), (uint32_t)h[v].base, tmp);
    undelta_32(tmp, (const uint32_t *)p, ends + (size_t)v * O_VEC);
  }
}

/* string column chunk -> offs[n+1] (relative to `pay`, offs[0] = base) and payload bytes.
 * Restates FLSStrColumn / str_col_t materialisation consumed at fastlanes_facade.cpp:153-171. */
static int chunk_str(const uint8_t *rgp, const ochunk *k, uint32_t n_vec, uint32_t base, uint32_t *offs, uint8_t *pay) {
  size_t n = (size_t)n_vec * O_VEC;
  offs[0] = base;
  switch (k->scheme) {
    case S_STR_CONST:
      for (size_t i = 0; i < n; ++i) {
        memcpy(pay + (size_t)i * k->aux32, rgp + k->seg[0].off, k->aux32);
        offs[i + 1] = base + (uint32_t)((i + 1) * k->aux32);
      }
      return 0;
    case S_STR_UNCOMP: {
      uint32_t *ends = (uint32_t *)malloc(n * 4);
      ends_stream(rgp, k, 0, n_vec, ends);
      for (size_t i = 0; i < n; ++i) offs[i + 1] = base + ends[i];
      memcpy(pay, rgp + k->seg[2].off, k->aux64[0]);
      free(ends);
      return 0;
    }
    case S_STR_FSST: {
      uint32_t *de = (uint32_t *)malloc(n * 4), *ce = (uint32_t *)malloc(n * 4);
      ends_stream(rgp, k, 0, n_vec, de);
      ends_stream(rgp, k, 2, n_vec, ce);
      const uint8_t *symtab = rgp + k->seg[4].off, *comp = rgp + k->seg[5].off;
      for (size_t i = 0; i < n; ++i) {
        uint32_t cs = i ? ce[i - 1] : 0, ds = i ? de[i - 1] : 0;
        uint64_t got = flso_fsst_decode(symtab, comp + cs, ce[i] - cs, pay + ds);
        if (got != de[i] - ds) { free(de); free(ce); return 21; }
        offs[i + 1] = base + de[i];
      }
      free(de); free(ce);
      return 0;
    }
    case S_STR_FSST12: {
      uint32_t *de = (uint32_t *)malloc(n * 4), *ce = (uint32_t *)malloc(n * 4);
      ends_stream(rgp, k, 0, n_vec, de);
      ends_stream(rgp, k, 2, n_vec, ce);
      const uint8_t *symtab = rgp + k->seg[4].off, *comp = rgp + k->seg[5].off;
      for (size_t i = 0; i < n; ++i) {
        uint32_t cs = i ? ce[i - 1] : 0, ds = i ? de[i - 1] : 0;
        uint64_t got = flso_fsst12_decode(symtab, k->aux32, comp + cs, ce[i] - cs, pay + ds);
        if (got != de[i] - ds) { free(de); free(ce); return 23; }
        offs[i + 1] = base + de[i];
      }
      free(de); free(ce);
      return 0;
    }
    case S_STR_DICT:
    case S_STR_FSST_DICT: {
      const ovhdr *h = (const ovhdr *)(rgp + k->seg[0].off);
      const uint8_t *data = rgp + k->seg[1].off;
      const uint32_t *doff = (const uint32_t *)(rgp + k->seg[2].off);
      const uint8_t *dbytes = rgp + k->seg[3].off;
      uint8_t *dec = NULL;
      uint32_t *dec_off = NULL;
      if (k->scheme == S_STR_FSST_DICT) { /* decode the dictionary once per row group */
        const uint8_t *symtab = rgp + k->seg[5].off;
        const uint32_t *want = (const uint32_t *)(rgp + k->seg[6].off);
        dec = (uint8_t *)malloc((size_t)want[k->aux32] + 8);
        dec_off = (uint32_t *)malloc(((size_t)k->aux32 + 1) * 4);
        dec_off[0] = 0;
        for (uint32_t d = 0; d < k->aux32; ++d) {
          uint64_t got = flso_fsst_decode(symtab, dbytes + doff[d], doff[d + 1] - doff[d], dec + dec_off[d]);
          dec_off[d + 1] = dec_off[d] + (uint32_t)got;
          if (dec_off[d + 1] != want[d + 1]) { free(dec); free(dec_off); return 22; }
        }
        doff = dec_off; dbytes = dec;
      }
      uint32_t idx[O_VEC], run = 0;
      for (uint32_t v = 0; v < n_vec; ++v) {
        unffor_32((const uint32_t *)(data + 16ull * h[v].ep), M_BW(h[v].meta), (uint32_t)h[v].base, idx);
        for (unsigned i = 0; i < O_VEC; ++i) {
          uint32_t l = doff[idx[i] + 1] - doff[idx[i]];
          memcpy(pay + run, dbytes + doff[idx[i]], l);
          run += l;
          offs[(size_t)v * O_VEC + i + 1] = base + run;
        }
      }
      free(dec); free(dec_off);
      return 0;
    }
  }
  return 20;
}


/* ---- tiny pthread parallel-for over row groups (no OpenMP runtime in this image) */
typedef int (*rg_fn)(void *ctx, uint32_t g);
typedef struct { rg_fn fn; void *ctx; uint32_t g1; volatile uint32_t next; volatile int rc; } par_job;
static void *par_worker(void *a) {
  par_job *j = (par_job *)a;
  for (;;) {
    uint32_t g = __sync_fetch_and_add(&j->next, 1u);
    if (g >= j->g1) break;
    int r = j->fn(j->ctx, g);
    if (r) j->rc = r;
  }
  return NULL;
}
static int par_for(uint32_t g0, uint32_t g1, int threads, rg_fn fn, void *ctx) {
  par_job j = {fn, ctx, g1, g0, 0};
  if (threads <= 1 || g1 - g0 <= 1) { par_worker(&j); return j.rc; }
  if (threads > 256) threads = 256;
  pthread_t th[256];
  int n = 0;
  for (; n < threads - 1; ++n) if (pthread_create(&th[n], NULL, par_worker, &j)) break;
  par_worker(&j);
  for (int i = 0; i < n; ++i) pthread_join(th[i], NULL);
  return j.rc;
}

/* ---- public decode entry points.  Output convention (identical to libflsgpu's):
 *   fixed width : out[(sum of n_vec over the range) * 1024] values, tail vectors padded
 *   strings     : offsets[rows_padded + 1] (u32, relative to payload[0]) + payload bytes */
/* Validity (FORMAT.md, chunk flag 1: the chunk's last segment is a bitmap, 128 bytes per vector, bit r = row r holds a
 * value).  out = padded rows / 8 bytes; chunks without a bitmap contribute all ones.  Returns 1 when some chunk of the
 * range has a bitmap, else 0.  The reference has no twin: it never reads FastLanes' null maps
 * (fastlanes_facade.cpp:113-118,173-181). */
int flso_decode_validity(const flso_table *t, uint32_t c, uint32_t g0, uint32_t g1, uint8_t *out) {
  int any = 0;
  uint64_t at = 0;
  for (uint32_t g = g0; g < g1; ++g) {
    const ochunk *k = &t->rg[g].cols[c];
    const uint64_t nb = 128ull * t->rg[g].n_vec;
    if ((k->flags & 1) && k->n_seg) { memcpy(out + at, t->img + t->rg[g].off + k->seg[k->n_seg - 1].off, nb); any = 1; }
    else memset(out + at, 0xFF, nb);
    at += nb;
  }
  return any;
}
uint64_t flso_range_padded_rows(const flso_table *t, uint32_t g0, uint32_t g1) {
  uint64_t n = 0;
  for (uint32_t g = g0; g < g1; ++g) n += (uint64_t)t->rg[g].n_vec * O_VEC;
  return n;
}
uint64_t flso_range_payload(const flso_table *t, uint32_t c, uint32_t g0, uint32_t g1) {
  uint64_t n = 0;
  for (uint32_t g = g0; g < g1; ++g) n += t->rg[g].cols[c].aux64[0];
  return n;
}

static int decode_fixed_rg(const flso_table *t, uint32_t c, uint32_t g, void *dst) {
  const org *r = &t->rg[g];
  const ochunk *k = &r->cols[c];
  const uint8_t *rgp = t->img + r->off;
  if (k->scheme == S_EQUAL) { /* the chunk equals another column's chunk of the same row group */
    if (k->aux32 >= t->n_cols || r->cols[k->aux32].scheme == S_EQUAL || type_bits(t->types[k->aux32]) != type_bits(t->types[c])) return 13;
    k = &r->cols[k->aux32];
  }
  if (k->scheme == S_ALP && type_bits(t->types[c]) == 32) return chunk_alp32(rgp, k, r->n_vec, (float *)dst);
  if (k->scheme == S_ALP) return chunk_alp(rgp, k, r->n_vec, (double *)dst);
  if (k->scheme == S_ALPRD) return chunk_alprd(rgp, k, r->n_vec, (uint64_t *)dst);
  switch (type_bits(t->types[c])) {
    case 8: return chunk_int_8(rgp, k, r->n_vec, (uint8_t *)dst);
    case 16: return chunk_int_16(rgp, k, r->n_vec, (uint16_t *)dst);
    case 32: return chunk_int_32(rgp, k, r->n_vec, (uint32_t *)dst);
    case 64: return chunk_int_64(rgp, k, r->n_vec, (uint64_t *)dst);
  }
  return 11;
}

typedef struct { const flso_table *t; uint32_t c, g0; uint64_t *rs, *ps; void *out; uint32_t *offsets; uint8_t *payload; } dec_ctx;
static int fixed_job(void *a, uint32_t g) {
  dec_ctx *x = (dec_ctx *)a;
  return decode_fixed_rg(x->t, x->c, g, (uint8_t *)x->out + x->rs[g - x->g0]);
}
int flso_decode_fixed(const flso_table *t, uint32_t c, uint32_t g0, uint32_t g1, void *out, int threads) {
  unsigned tb = type_bits(t->types[c]);
  if (!tb) return 12;
  uint64_t *start = (uint64_t *)malloc((size_t)(g1 - g0 + 1) * 8);
  start[0] = 0;
  for (uint32_t g = g0; g < g1; ++g) start[g - g0 + 1] = start[g - g0] + (uint64_t)t->rg[g].n_vec * O_VEC * (tb / 8);
  dec_ctx x = {t, c, g0, start, NULL, out, NULL, NULL};
  int rc = par_for(g0, g1, threads, fixed_job, &x);
  free(start);
  return rc;
}

static int str_job(void *a, uint32_t g) {
  dec_ctx *x = (dec_ctx *)a;
  const org *r = &x->t->rg[g];
  return chunk_str(x->t->img + r->off, &r->cols[x->c], r->n_vec, (uint32_t)x->ps[g - x->g0],
                   x->offsets + x->rs[g - x->g0], x->payload + x->ps[g - x->g0]);
}
int flso_decode_str(const flso_table *t, uint32_t c, uint32_t g0, uint32_t g1, uint32_t *offsets, uint8_t *payload,
                    int threads) {
  if (type_bits(t->types[c])) return 13;
  uint64_t *rs = (uint64_t *)malloc((size_t)(g1 - g0 + 1) * 8), *ps = (uint64_t *)malloc((size_t)(g1 - g0 + 1) * 8);
  rs[0] = ps[0] = 0;
  for (uint32_t g = g0; g < g1; ++g) {
    rs[g - g0 + 1] = rs[g - g0] + (uint64_t)t->rg[g].n_vec * O_VEC;
    ps[g - g0 + 1] = ps[g - g0] + t->rg[g].cols[c].aux64[0];
  }
  dec_ctx x = {t, c, g0, rs, ps, NULL, offsets, payload};
  offsets[0] = 0;
  int rc = par_for(g0, g1, threads, str_job, &x);
  free(rs); free(ps);
  return rc;
}
