/* flsw.h — C ABI of the FLS encoder/writer (libflsw.so, host code; csrc/fls_writer.cpp).
 *
 * This is the write half the reference declares and never implements:
 *   FastLanesFacade::createFile / writeChunk / finalizeFile   /root/reference/src/include/fastlanes_facade.hpp:40-44
 * and the COPY ... TO 'x.fls' (FORMAT FLS) sink built on it   /root/reference/src/writer/write_fastlane_stream.cpp:65-177
 * (SURVEY.md section 8f rank 2).  It is not on the timed decode path: the tests and bench.py use it to mint
 * workloads, and ext/write_fastlane.cpp binds it as the DuckDB copy function.
 *
 * A writer takes whole row groups (<= 65,536 rows) as column arrays, picks a scheme per column chunk (FFOR, DELTA,
 * RLE, DICT, ALP, ALP-RD, constant, uncompressed; strings: dictionary, FSST, FSST-compressed dictionary, constant,
 * uncompressed) and produces the container described in FORMAT.md. */
#ifndef FLSW_H
#define FLSW_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct flsw_writer flsw_writer;

/* createFile: column names and types (enum flsgpu_type of flsgpu.h).  Never NULL. */
flsw_writer *flsw_create(uint32_t n_cols, const char *const *names, const int32_t *types);
/* writeChunk: append one row group of n_rows <= 65,536 rows.  col_data[c] = the column's values (fixed width: n_rows
 * elements of the type's width; strings: the payload bytes), str_off[c] = u64 offsets[n_rows + 1] for string columns
 * (NULL for fixed width).  Returns 0, or 1 with flsw_error(). */
int flsw_append_rowgroup(flsw_writer *w, uint32_t n_rows, const void *const *col_data, const uint64_t *const *str_off);
/* The same with NULLs: validity[c] = bitmap of column c (bit r of word r / 64, LSB first = row r holds a value; DuckDB's
 * ValidityMask layout) or NULL for a column without NULLs; `validity` itself may be NULL.  The values / strings at NULL
 * rows are ignored (strings: the offsets must still ascend). */
int flsw_append_rowgroup_v(flsw_writer *w, uint32_t n_rows, const void *const *col_data, const uint64_t *const *str_off,
                           const uint64_t *const *validity);
/* finalizeFile: write the footer (flsw_finish) / the whole image to `path` (finishes first if needed). */
int flsw_finish(flsw_writer *w);
int flsw_write_file(flsw_writer *w, const char *path);
const void *flsw_data(flsw_writer *w);
uint64_t flsw_size(flsw_writer *w);
const char *flsw_error(flsw_writer *w);
void flsw_destroy(flsw_writer *w);

/* Test knobs: force a column's scheme (enum fls_scheme, -1 = choose) and a minimum FFOR width (-1 = none); pin a
 * double column's ALP exponent/factor; the scalar ALP encode the writer uses. */
int flsw_force_scheme(flsw_writer *w, uint32_t col, int32_t scheme, int32_t min_bw);
int flsw_force_alp_ef(flsw_writer *w, uint32_t col, int32_t e, int32_t f);
int64_t flsw_alp_encode_value(double v, int32_t e, int32_t f);
int64_t flsw_alpf_encode_value(float v, int32_t e, int32_t f); /* FLOAT columns: float arithmetic, e, f <= 10 */
/* Test knobs for parity against the reference tree's own encoders: the NEXT row group appended for `col` is built
 * from externally produced streams, verbatim.  FSST: libfsst's symbol table (u64 symbol[256] | u8 len[256]) and
 * compressed bytes (comp_off[n_strings + 1]); the row group still receives the raw strings (decoded offsets).
 * ALP-RD: DuckDB's cut position, dictionary, per-value dictionary indices / right parts and exceptions
 * (per vector: 1024 slots of xpos / xleft, the first nexc[v] used).  Pointers must stay valid until that
 * flsw_append_rowgroup returns. */
int flsw_inject_fsst(flsw_writer *w, uint32_t col, const uint8_t *symtab2304, const uint8_t *comp, const uint64_t *comp_off,
                     uint32_t n_strings);
int flsw_inject_alprd(flsw_writer *w, uint32_t col, uint32_t rbw, uint32_t lbw, uint32_t ndict, const uint16_t *dict8,
                      uint32_t n_vec, const uint16_t *left, const uint64_t *right, const uint32_t *nexc, const uint16_t *xpos,
                      const uint16_t *xleft);
/* A new finished writer holding n_total row groups: the (full) row groups of `image` repeated round-robin. */
flsw_writer *flsw_tile(const void *image, uint64_t size, uint32_t n_total);

#ifdef __cplusplus
}
#endif
#endif
