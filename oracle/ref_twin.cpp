// ref_twin.cpp — thin C shim over the REFERENCE TREE's own ALP / ALP-RD / FSST code.  TEST INFRASTRUCTURE ONLY.
//
// FastLanes itself is absent from /root/reference, but DuckDB (vendored there, same CWI authors) carries
// in-tree twins of three of the arithmetic kernels the read path needs.  This file
// contains NO algorithm: it includes the reference headers / compiles libfsst.cpp where they lie
// (oracle/Makefile `ref`, __graft_entry__.build) and exposes them over a C ABI so that
// tests/golden/make_golden.py can mint golden vectors in the dev container.  The built library goes to
// oracle/_ref/ (git-ignored); nothing on the GPU box needs it — only the committed vectors travel.
//
//   ALP     duckdb/src/include/duckdb/storage/compression/alp/algorithm/alp.hpp:108-134 (EncodeValue/DecodeValue),
//           :289-366 (Compress), :376-404 (Decompress)
//   ALP-RD  duckdb/src/include/duckdb/storage/compression/alprd/algorithm/alprd.hpp:84-199 (dictionary + Compress),
//           :206-233 (Decompress)
//   FSST    duckdb/third_party/fsst/fsst.h:93-104,162-241 ; libfsst.cpp (create / compress / export / import)
#include <cmath>
#include <cstring>
#include <vector>

#include "duckdb/common/types/value.hpp"
#include "duckdb/storage/compression/alp/algorithm/alp.hpp"
#include "duckdb/storage/compression/alprd/algorithm/alprd.hpp"
#include "fsst.h"

// Glue only: the two Value:: predicates alp.hpp calls live in duckdb's value.cpp, which drags the whole engine
// in at link time.  Their meaning is std::isfinite / std::isnan (duckdb/src/common/types/value.cpp:362-386).
namespace duckdb {
template <> bool Value::IsFinite(double v) { return std::isfinite(v); }
template <> bool Value::IsNan(double v) { return std::isnan(v); }
template <> bool Value::IsFinite(float v) { return std::isfinite(v); }
template <> bool Value::IsNan(float v) { return std::isnan(v); }
// Error-path glue: the three exception helpers the headers reference (bounds checks of duckdb::vector, numeric
// casts).  They only run when an internal invariant breaks; here they abort instead of formatting a message.
Exception::Exception(ExceptionType, const std::string &msg) : std::runtime_error(msg) { abort(); }
InternalException::InternalException(const std::string &msg) : Exception(ExceptionType::INTERNAL, msg) { }
ExceptionFormatValue::ExceptionFormatValue(int64_t) { }
std::string Exception::ConstructMessageRecursive(const std::string &msg, std::vector<ExceptionFormatValue> &) { return msg; }
}  // namespace duckdb

using namespace duckdb;

extern "C" {

int64_t twin_alp_encode(double v, uint8_t e, uint8_t f) {
  alp::AlpEncodingIndices idx(e, f);
  return alp::AlpCompression<double, true>::EncodeValue(v, idx);
}
double twin_alp_decode(int64_t n, uint8_t e, uint8_t f) {
  alp::AlpEncodingIndices idx(e, f);
  return alp::AlpCompression<double, true>::DecodeValue(n, idx);
}

// The same for FLOAT columns (AlpTypedConstants<float>: float arithmetic, exponents 0..10).
int64_t twin_alpf_encode(float v, uint8_t e, uint8_t f) {
  alp::AlpEncodingIndices idx(e, f);
  return alp::AlpCompression<float, true>::EncodeValue(v, idx);
}
float twin_alpf_decode(int64_t n, uint8_t e, uint8_t f) {
  alp::AlpEncodingIndices idx(e, f);
  return alp::AlpCompression<float, true>::DecodeValue(n, idx);
}
// DuckDB's sampler + Compress + Decompress on one float vector: decoded floats, chosen (e, f), exception count
int twin_alpf_roundtrip(const float *in, uint32_t n, float *out, uint8_t *e, uint8_t *f) {
  alp::AlpCompressionState<float, false> st;
  duckdb::vector<duckdb::vector<float>> samples; samples.emplace_back(in, in + n);
  alp::AlpCompression<float, false>::FindTopKCombinations(samples, st);
  alp::AlpCompression<float, false>::Compress(in, n, st);
  alp::AlpDecompression<float>::Decompress(st.values_encoded, out, n, st.vector_encoding_indices.factor,
                                           st.vector_encoding_indices.exponent, st.exceptions_count, st.exceptions,
                                           st.exceptions_positions, st.frame_of_reference, (uint8_t)st.bit_width);
  *e = st.vector_encoding_indices.exponent;
  *f = st.vector_encoding_indices.factor;
  return st.exceptions_count;
}

// One ALP vector through DuckDB's sampler + compressor.  Outputs: chosen (e,f), FOR base, bit width,
// FOR-subtracted encoded integers, exceptions.  Returns the exception count.
int twin_alp_compress(const double *in, uint32_t n, uint8_t *e, uint8_t *f, uint64_t *frame, uint8_t *bw,
                      uint64_t *encoded, uint16_t *exc_pos, double *exc_val) {
  alp::AlpCompressionState<double, false> st;
  duckdb::vector<duckdb::vector<double>> samples; samples.emplace_back(in, in + n);
  alp::AlpCompression<double, false>::FindTopKCombinations(samples, st);
  alp::AlpCompression<double, false>::Compress(in, n, st);
  *e = st.vector_encoding_indices.exponent;
  *f = st.vector_encoding_indices.factor;
  *frame = st.frame_of_reference;
  *bw = (uint8_t)st.bit_width;
  memcpy(encoded, st.encoded_integers, n * 8);
  for (uint32_t i = 0; i < st.exceptions_count; ++i) { exc_pos[i] = st.exceptions_positions[i]; exc_val[i] = st.exceptions[i]; }
  return st.exceptions_count;
}

// DuckDB's decompressor on (bit-packed by DuckDB) data produced by its compressor: returns decoded doubles.
void twin_alp_roundtrip(const double *in, uint32_t n, double *out) {
  alp::AlpCompressionState<double, false> st;
  duckdb::vector<duckdb::vector<double>> samples; samples.emplace_back(in, in + n);
  alp::AlpCompression<double, false>::FindTopKCombinations(samples, st);
  alp::AlpCompression<double, false>::Compress(in, n, st);
  alp::AlpDecompression<double>::Decompress(st.values_encoded, out, n, st.vector_encoding_indices.factor,
                                            st.vector_encoding_indices.exponent, st.exceptions_count, st.exceptions,
                                            st.exceptions_positions, st.frame_of_reference, (uint8_t)st.bit_width);
}

// One ALP-RD vector: dictionary from the vector itself, then Compress.  Outputs right/left widths, dictionary,
// per-value dictionary index and right part, exceptions (left parts).  Returns the exception count.
int twin_alprd_compress(const uint64_t *in, uint32_t n, uint8_t *rbw, uint8_t *lbw, uint8_t *ndict, uint16_t *dict,
                        uint16_t *left_idx, uint64_t *right, uint16_t *exc_pos, uint16_t *exc_left, uint64_t *roundtrip) {
  alp::AlpRDCompressionState<double, false> st;
  duckdb::vector<uint64_t> sample(in, in + n);
  alp::AlpRDCompression<double, false>::FindBestDictionary(sample, st);
  alp::AlpRDCompression<double, false>::Compress(in, n, st);
  *rbw = st.right_bit_width;
  *lbw = st.left_bit_width;
  *ndict = st.actual_dictionary_size;
  for (int i = 0; i < 8; ++i) dict[i] = st.left_parts_dict[i];
  for (uint32_t i = 0; i < n; ++i) {
    right[i] = in[i] & ((1ULL << st.right_bit_width) - 1);
    uint16_t key = (uint16_t)(in[i] >> st.right_bit_width);
    auto it = st.left_parts_dict_map.find(key);
    uint16_t di = it == st.left_parts_dict_map.end() ? st.actual_dictionary_size : (uint16_t)it->second;
    left_idx[i] = di >= st.actual_dictionary_size ? st.actual_dictionary_size : di;
  }
  for (uint32_t i = 0; i < st.exceptions_count; ++i) { exc_pos[i] = st.exceptions_positions[i]; exc_left[i] = st.exceptions[i]; }
  alp::AlpRDDecompression<double>::Decompress(st.left_parts_encoded, st.right_parts_encoded, st.left_parts_dict, roundtrip, n,
                                              st.exceptions_count, st.exceptions, st.exceptions_positions,
                                              st.left_bit_width, st.right_bit_width);
  return st.exceptions_count;
}

// FSST: train libfsst on the strings, compress each, export the decoder table in this repo's container layout
// (u64 symbol[256] little-endian + u8 len[256]; code 255 = escape).  comp_off[n+1] are offsets into comp.
// Returns compressed bytes, or 0 on failure.  Also decompresses with libfsst itself into `roundtrip`.
uint64_t twin_fsst_compress(const uint8_t *bytes, const uint64_t *off, uint32_t n, uint8_t *symtab2304, uint8_t *comp,
                            uint64_t comp_cap, uint64_t *comp_off, uint8_t *roundtrip) {
  std::vector<size_t> lens(n), clens(n);
  std::vector<unsigned char *> ptrs(n), cptrs(n);
  for (uint32_t i = 0; i < n; ++i) { lens[i] = off[i + 1] - off[i]; ptrs[i] = (unsigned char *)bytes + off[i]; }
  duckdb_fsst_encoder_t *enc = duckdb_fsst_create(n, lens.data(), ptrs.data(), 0);
  if (!enc) return 0;
  size_t done = duckdb_fsst_compress(enc, n, lens.data(), ptrs.data(), comp_cap, comp, clens.data(), cptrs.data());
  if (done != n) { duckdb_fsst_destroy(enc); return 0; }
  duckdb_fsst_decoder_t dec = duckdb_fsst_decoder(enc);
  memset(symtab2304, 0, 2304);
  for (int c = 0; c < 255; ++c) {
    uint64_t s = dec.symbol[c];
    uint32_t L = dec.len[c];
    if (L < 8) s &= (1ULL << (8 * L)) - 1;
    memcpy(symtab2304 + 8 * c, &s, 8);
    symtab2304[2048 + c] = (uint8_t)L;
  }
  comp_off[0] = 0;
  uint64_t pos = 0, rt = 0;
  std::vector<uint8_t> packed;
  for (uint32_t i = 0; i < n; ++i) {
    packed.insert(packed.end(), cptrs[i], cptrs[i] + clens[i]);
    pos += clens[i];
    comp_off[i + 1] = pos;
    rt += duckdb_fsst_decompress(&dec, clens[i], cptrs[i], lens[i] + 8, roundtrip + rt);
  }
  memcpy(comp, packed.data(), packed.size());
  duckdb_fsst_destroy(enc);
  return pos;
}
}
