// fls_reader.h — host-side FLS footer parser (no CUDA).
//
// Replaces what `fastlanes::Connection::read_fls(path)` + `TableReader` give the reference
// (/root/reference/src/fastlanes_facade.cpp:33-41): the table descriptor — schema, row groups, and for every
// (row group, column) the scheme and segment extents.  No data is decoded here (the reference's bind decodes
// row group 0 as a side effect, /root/reference/src/scan_fastlanes.cpp:40-43; this build does not).
#pragma once
#include <stdint.h>
#include <string>
#include <vector>

#include "fls_format.h"

struct FlsTable {
  uint32_t n_cols = 0, n_rg = 0, rg_vecs = 0;
  uint64_t n_rows = 0;
  std::vector<int> types;
  std::vector<std::string> names;
  std::vector<fls_rowgroup> rgs;
  std::vector<fls_colchunk> chunks;  // [rg * n_cols + col]
  uint64_t data_begin = 0, data_end = 0;  // file extent of the row-group area

  const fls_colchunk &chunk(uint32_t rg, uint32_t col) const { return chunks[(size_t)rg * n_cols + col]; }
  uint64_t padded_rows(uint32_t g0, uint32_t g1) const {
    uint64_t n = 0;
    for (uint32_t g = g0; g < g1; ++g) n += (uint64_t)rgs[g].n_vec * FLS_VEC;
    return n;
  }
  uint64_t rows(uint32_t g0, uint32_t g1) const {
    uint64_t n = 0;
    for (uint32_t g = g0; g < g1; ++g) n += rgs[g].n_tuples;
    return n;
  }
};

// Parse the footer of an in-memory image.  Returns "" on success, else an error message.
std::string fls_parse(const uint8_t *image, uint64_t size, FlsTable &out);
// Parse the footer of a file without reading the row-group area.
std::string fls_parse_file(const char *path, FlsTable &out, uint64_t *file_size);
