// fls_format.h — single source of truth for the FLS container as this repo reads and writes it.
//
// The upstream FastLanes library (cwida/FastLanes @ 928fdb82, the code behind
// `connection->read_fls(...)` at /root/reference/src/fastlanes_facade.cpp:34) is NOT present in the
// reference checkout, so the container below is a restatement from public knowledge of FastLanes
//  and is UNVERIFIED against upstream bytes.  See FORMAT.md for the prose spec.
//
// Shared by: fls_writer.cpp (encoder), fls_reader.cpp (footer parser), flsgpu.cu (kernels).
// The oracle (oracle/fls_oracle.c) deliberately restates these constants instead of including this file.
#pragma once
#include <stdint.h>

#define FLS_MAGIC 0x656E614C74736146ULL /* "FastLane" little-endian; echoed by
                                           /root/reference/src/writer/write_fastlane.cpp:117 */
#define FLS_VERSION_BYTES "v0.1.4\0"    /* /root/reference/src/writer/write_fastlane.cpp:122 */
#define FLS_FOOTER_VERSION 1u

#define FLS_VEC 1024u           /* values per FastLanes vector */
/* dictionary-string kernel, short configuration (fls_strings.cuh): its staging buffer, and the longest entry of which any
 * 256 fit it — what the planner routes by (fls_plan.h) */
#ifndef FLS_DICT_SHORT_STG
#define FLS_DICT_SHORT_STG 6656u
#define FLS_DICT_SHORT_WARPS 20u
#endif
#define FLS_DICT_SHORT_MAXLEN ((FLS_DICT_SHORT_STG - 15u) / 256u) /* 25 */
#define FLS_RG_VECS 64u         /* vectors per row group (65,536 rows;
                                   /root/reference/src/writer/write_fastlane_stream.cpp:21-24) */
#define FLS_RG_ALIGN 256u       /* row groups start on 256-byte file offsets */
#define FLS_SEG_ALIGN 16u       /* segments start on 16-byte offsets inside a row group */
#define FLS_MAX_SEGS 8u

// ---- data types: order mirrors the reference's type table
//      (/root/reference/src/type_mapping.cpp:11-105).
enum fls_type {
  FLS_T_INVALID = 0,
  FLS_T_BOOLEAN = 1,
  FLS_T_INT8 = 2,
  FLS_T_INT16 = 3,
  FLS_T_INT32 = 4,
  FLS_T_INT64 = 5,
  FLS_T_UINT8 = 6,
  FLS_T_UINT16 = 7,
  FLS_T_UINT32 = 8,
  FLS_T_UINT64 = 9,
  FLS_T_FLOAT = 10,
  FLS_T_DOUBLE = 11,
  FLS_T_STR = 12,
  FLS_T_DATE = 13,
  FLS_T_BYTE_ARRAY = 14,
  FLS_T_N
};

// ---- column-chunk schemes (the top-level operator of the column's decode expression)
enum fls_scheme {
  FLS_S_UNCOMPRESSED = 0,  // B raw
  FLS_S_CONSTANT = 1,      // aux64[0] = value bits
  FLS_S_FFOR = 2,          // S
  FLS_S_DELTA = 3,         // S   (lane bases inline)
  FLS_S_RLE = 4,           // S   (run values + u16 index delta stream inline)
  FLS_S_DICT = 5,          // S idx(u32 FFOR) + B dict
  FLS_S_ALP = 6,           // S   (f64; exceptions inline)
  FLS_S_ALP_RD = 7,        // S   (f64; aux32 = rbw | lbw<<8 | ndict<<16, aux64[0..1] = 8 x u16 dict)
  FLS_S_STR_UNCOMPRESSED = 8,  // S end-offsets(DELTA u32) + B bytes
  FLS_S_STR_CONSTANT = 9,      // B bytes (the single value)
  FLS_S_STR_DICT = 10,         // S idx(u32 FFOR) + B dict_offsets(u32 x (n+1)) + B dict_bytes + B vec_ep(u32 x (n_vec+1))
  FLS_S_STR_FSST = 11,         // S dec end-offsets(DELTA u32) + S cmp end-offsets(DELTA u32) + B symtab + B bytes
  FLS_S_STR_FSST_DICT = 12,    // S idx + B dict_cmp_offsets(u32 x (n+1)) + B dict_cmp_bytes + B vec_ep + B symtab
                               //   + B dict_dec_offsets(u32 x (n+1))
  FLS_S_FFOR_PATCH = 13,   // S   FFOR + exception list (FastLanes SLPATCH; with bw = 0 it is FREQUENCY: the base is the
                           //     vector's most frequent value and every other value an exception)
  FLS_S_CROSS_RLE = 14,    // B run values (aux32 x T) + B run ends (aux32 x u32, cumulative, in padded rows of the row
                           //     group): runs span vectors
  FLS_S_EQUAL = 15,        // the chunk equals column aux32's chunk of the same row group (no segments)
  FLS_S_STR_FSST12 = 16,   // FSST with 12-bit codes: S dec end-offsets + S cmp end-offsets (both DELTA u32) + B symbol table
                           //   (aux32 = n symbols <= 4096: u64 symbol[n] | u8 len[n]; codes 0..255 are the bytes themselves,
                           //   there is no escape) + B compressed bytes (+16 slack).  A string's codes are packed LSB first,
                           //   12 bits each, padded to whole bytes: L compressed bytes hold floor(2 L / 3) codes
  FLS_S_N
};

// Segment roles inside a column chunk (index into fls_colchunk.seg[]).
//   stream k  = { seg[2k] = vector-header table, seg[2k+1] = data }   (S above)
//   blobs follow the streams in the order listed per scheme.
enum { FLS_SEG_HDR0 = 0, FLS_SEG_DATA0 = 1, FLS_SEG_HDR1 = 2, FLS_SEG_DATA1 = 3 };

// ---- per-vector header: 16 bytes, one per vector, table = seg[2k]
//   ep   : offset of this vector's payload inside the stream's data segment, in 16-byte units
//   meta : bw[0:7) | e[8:13) | f[13:18) | n[18:29)      (e,f: ALP exponent/factor; n: n_exc or n_runs)
//   base : FFOR base (wrapping add after unpack); ALP: frame of reference of the encoded int64s
typedef struct {
  uint32_t ep;
  uint32_t meta;
  uint64_t base;
} fls_vechdr;

#define FLS_META(bw, e, f, n) \
  ((uint32_t)(bw) | ((uint32_t)(e) << 8) | ((uint32_t)(f) << 13) | ((uint32_t)(n) << 18))
#define FLS_META_BW(m) ((m) & 0x7Fu)
#define FLS_META_E(m) (((m) >> 8) & 0x1Fu)
#define FLS_META_F(m) (((m) >> 13) & 0x1Fu)
#define FLS_META_N(m) (((m) >> 18) & 0x7FFu)

// Payload layouts at `ep` (all pieces padded to 16 bytes):
//   FFOR    : [packed 128*bw]
//   DELTA   : [lane bases 128 B = L x T/8][packed 128*bw]          (deltas are FFOR'd: + base)
//   RLE     : [run values n x T/8, pad16][idx lane bases 64 x u16][idx packed 128*bw, bw in {0,1}]
//   ALP     : [packed 128*bw][exc values n x f64, pad16][exc positions n x u16, pad16]
//   FFOR_PATCH : [packed 128*bw][exc values n x T, pad16][exc positions n x u16, pad16]   (exception slots pack 0)
//   ALP_RD  : [right packed 128*rbw][left packed 128*lbw][exc left n x u16, pad16][exc pos n x u16, pad16]

// Unified transposed layout.
static const uint8_t FLS_ORDER[8] = {0, 4, 2, 6, 1, 5, 3, 7};

// ---- in-memory descriptors produced by the footer parser
typedef struct {
  uint64_t off;   // relative to the row group's first byte
  uint64_t size;
} fls_seg;

// fls_colchunk.flags
#define FLS_CHUNK_VALIDITY 1u  // the chunk's LAST segment is a validity bitmap: 128 bytes per vector, bit r of the vector
                               // (LSB first) = row r holds a value; NULL rows keep a filler in the value streams

typedef struct {
  uint8_t scheme;   // enum fls_scheme
  uint8_t n_seg;
  uint16_t flags;
  uint32_t aux32;   // DICT/STR_DICT: dictionary entries; ALP_RD: rbw|lbw<<8|ndict<<16
  uint64_t aux64[2];  // CONSTANT: value; ALP_RD: dict; STR_*: [0]=decoded payload bytes of the row group
  fls_seg seg[FLS_MAX_SEGS];
} fls_colchunk;

typedef struct {
  uint64_t off;     // file offset
  uint64_t size;
  uint32_t n_vec;
  uint32_t n_tuples;
} fls_rowgroup;

static inline uint32_t fls_type_bits(int t) {
  switch (t) {
    case FLS_T_BOOLEAN: case FLS_T_INT8: case FLS_T_UINT8: return 8;
    case FLS_T_INT16: case FLS_T_UINT16: return 16;
    case FLS_T_INT32: case FLS_T_UINT32: case FLS_T_FLOAT: case FLS_T_DATE: return 32;
    case FLS_T_INT64: case FLS_T_UINT64: case FLS_T_DOUBLE: return 64;
    default: return 0;  // variable width
  }
}
