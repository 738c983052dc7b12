/* flsgpu.h — C ABI of libflsgpu.so, the B200 (sm_100a) FLS read/decode library.
 *
 * This is the drop-in boundary of the read_fls / scan_fastlanes path: it is what the reference's
 * FastLanesFacade would bind instead of the FastLanes C++ reader objects.  Reference interface each entry
 * point replaces (paths relative to /root/reference):
 *
 *   flsgpu_open / flsgpu_open_image   src/fastlanes_facade.cpp:26-67  FastLanesFacade::openFile
 *                                     (fastlanes::connect :33, Connection::read_fls :34,
 *                                      TableReader::get_rowgroup_reader :41) — WITHOUT the eager
 *                                     materialize() of :48: open only parses the footer.
 *   flsgpu_schema                     src/include/fastlanes_facade.hpp:34-35 getColumnTypes/getColumnNames
 *                                     (declared, never implemented) and src/type_mapping.cpp:64-105.
 *   flsgpu_num_rows / _num_rowgroups /
 *   _rowgroup_rows                    src/fastlanes_facade.cpp:52-56 Rowgroup::ColCount/RowCount.
 *   flsgpu_decode                     src/fastlanes_facade.cpp:48 RowgroupReader::materialize() (the decode)
 *                                     plus :69-200 readNextChunk's copy-out: decoded columns come back as
 *                                     flat typed arrays (strings: u32 offsets + payload bytes), host or device.
 *   flsgpu_make_resident              no reference twin: stages compressed row groups into HBM once so that
 *                                     repeated decodes time the kernels only (BASELINE "HBM-resident").
 *   flsgpu_batch_release / flsgpu_close   src/fastlanes_facade.cpp:202-210 closeFile.
 *   flsgpu_last_error                 the facade's bool returns + BinderException text
 *                                     (src/scan_fastlanes.cpp:42 "Failed to open FastLanes file: <path>").
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or an FLSGPU_E_* code;
 * no exception crosses this boundary.  One caller at a time per handle, from any thread.  There is NO CPU
 * decode behind this ABI: without a CUDA device every decode call fails with FLSGPU_E_CUDA.
 */
#ifndef FLSGPU_H
#define FLSGPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLSGPU_ABI_VERSION 2

enum flsgpu_status {
  FLSGPU_OK = 0,
  FLSGPU_E_IO = 1,      /* cannot open / read the file */
  FLSGPU_E_FORMAT = 2,  /* not an FLS file, or corrupt footer */
  FLSGPU_E_RANGE = 3,   /* bad row-group range / column id / argument */
  FLSGPU_E_CUDA = 4,    /* CUDA runtime failure (message in flsgpu_last_error) */
  FLSGPU_E_LIMIT = 5,   /* batch exceeds a limit (a filtered decode over a string column with >= 4 GiB of payload) */
  FLSGPU_E_NOMEM = 6
};

/* Column types: values of enum fls_type, the order of the reference's type table
 * (src/type_mapping.cpp:11-62). */
enum flsgpu_type {
  FLSGPU_T_BOOLEAN = 1, FLSGPU_T_INT8, FLSGPU_T_INT16, FLSGPU_T_INT32, FLSGPU_T_INT64, FLSGPU_T_UINT8,
  FLSGPU_T_UINT16, FLSGPU_T_UINT32, FLSGPU_T_UINT64, FLSGPU_T_FLOAT, FLSGPU_T_DOUBLE, FLSGPU_T_STR,
  FLSGPU_T_DATE, FLSGPU_T_BYTE_ARRAY
};

/* open flags */
#define FLSGPU_OPEN_PINNED_IMAGE 1u /* open_image: the caller's image is already page-locked */
#define FLSGPU_OPEN_NO_PIN 2u       /* do not page-lock the image (H2D then goes through the driver's bounce buffer) */
#define FLSGPU_OPEN_TIME_KERNELS 4u /* record a CUDA-event pair around every kernel launch (flsgpu_kernel_stats) */
#define FLSGPU_OPEN_LOAD_FILE 16u /* flsgpu_open: read the whole file into page-locked memory at the first decode (the
                                     round-1 behaviour) instead of mapping it and staging each pipeline slice through a
                                     page-locked bounce buffer */
#define FLSGPU_OPEN_GDS 32u       /* flsgpu_open: read each slice's compressed bytes straight from the file into HBM with
                                     cuFile (GPUDirect Storage; SURVEY.md section 8f rank 4, no reference twin).  Fails
                                     with FLSGPU_E_IO and the cuFile error text when libcufile or the driver is not
                                     available — never silently falls back */
#define FLSGPU_OPEN_NO_PLAN_CACHE 8u /* rebuild and re-upload the chunk descriptors on every decode (what the first
                                        scan of a range pays; bench.py reports it beside the cached figure) */

/* decode flags */
#define FLSGPU_OUT_DEVICE 0u /* leave decoded columns in HBM; column pointers are device pointers */
#define FLSGPU_OUT_HOST 1u   /* also copy decoded columns to pinned host memory; column pointers are host pointers */

typedef struct flsgpu flsgpu_t;
typedef struct flsgpu_batch flsgpu_batch_t;

typedef struct {
  int32_t device;        /* CUDA device ordinal */
  uint32_t flags;        /* FLSGPU_OPEN_* */
  uint32_t slice_rowgroups; /* row groups per H2D/decode/D2H pipeline slice when the range is not resident; 0 = default */
  uint32_t reserved;
} flsgpu_opts;

typedef struct {
  int32_t type;           /* enum flsgpu_type */
  uint32_t width;         /* bytes per value; 0 for strings */
  uint64_t rows;          /* valid rows in [g0,g1) */
  uint64_t padded_rows;   /* rows the buffers hold (every row group padded to whole 1024-value vectors) */
  void *data;             /* fixed width: padded_rows * width bytes */
  uint32_t *offsets;      /* strings: padded_rows + 1 Arrow-style offsets into payload, offsets[0] = 0 */
  uint8_t *payload;       /* strings: payload bytes */
  uint64_t payload_bytes; /* strings: bytes of payload in use */
  uint64_t compressed_bytes; /* algorithmic bytes read: segment bytes of this column in [g0,g1) */
  uint64_t decoded_bytes;    /* algorithmic bytes written: values, or offsets + payload */
  uint64_t *offsets64;    /* strings whose payload in this batch reaches 4 GiB: `offsets` is NULL and this holds
                             padded_rows + 1 64-bit offsets instead (NULL otherwise) */
  uint64_t *validity;     /* NULL: every row holds a value.  Else a bitmap over the padded rows, bit r (LSB first, 64
                             rows per word, DuckDB's ValidityMask layout) = row r is not NULL; a NULL row's value /
                             string is a filler.  The reference ignores FastLanes' null maps
                             (src/fastlanes_facade.cpp:113-118,173-181 emit NULL only for unknown columns). */
} flsgpu_column;

typedef struct {
  float ms_total;    /* first enqueue -> last completion, on the device (CUDA events) */
  float ms_h2d;      /* sum over slices of the compressed-bytes H2D copies */
  float ms_kernels;  /* sum over slices of the decode kernels */
  float ms_d2h;      /* sum over slices of the decoded-bytes D2H copies */
  uint64_t h2d_bytes, d2h_bytes;
  uint64_t compressed_bytes, decoded_bytes; /* algorithmic, over the selected columns */
  uint32_t kernel_launches, slices;
  float ms_host_plan; /* host time spent building chunk descriptors (inside ms_total) */
  float ms_host_total; /* host wall time of the whole call */
} flsgpu_timing;

typedef struct {
  const char *name;     /* kernel family, e.g. "ffor_u32" */
  uint64_t launches;
  double ms;            /* sum of per-launch CUDA-event durations */
  uint64_t bytes_read;  /* algorithmic */
  uint64_t bytes_written;
  uint64_t vectors;
} flsgpu_kernel_stat;

/* ---- open / close ------------------------------------------------------------------------------------ */
int flsgpu_open(const char *path, const flsgpu_opts *opts, flsgpu_t **out);
/* The image must stay valid until flsgpu_close. */
int flsgpu_open_image(const void *image, uint64_t size, const flsgpu_opts *opts, flsgpu_t **out);
void flsgpu_close(flsgpu_t *h);
/* Closed handles leave their pooled device and page-locked host buffers in a bounded process-wide cache for the next
 * handle (a query = open + scan + close; FLSGPU_CACHE_HOST_MB / FLSGPU_CACHE_DEV_MB set the bounds, 0 = keep nothing).
 * This frees the cache. */
void flsgpu_trim_cache(void);
/* Error text of the last failing call on `h`; with h == NULL, of the last failing open on this thread. */
const char *flsgpu_last_error(const flsgpu_t *h);
int flsgpu_abi_version(void);
/* CUDA devices visible to the process (0 when there is none or no driver). */
int flsgpu_device_count(void);

/* ---- schema ------------------------------------------------------------------------------------------- */
uint32_t flsgpu_num_columns(const flsgpu_t *h);
uint32_t flsgpu_num_rowgroups(const flsgpu_t *h);
uint64_t flsgpu_num_rows(const flsgpu_t *h);
uint32_t flsgpu_rowgroup_rows(const flsgpu_t *h, uint32_t rg);
/* names[i] / types[i] stay valid until flsgpu_close. */
int flsgpu_schema(const flsgpu_t *h, uint32_t *n_cols, const char *const **names, const int32_t **types);
/* scheme (enum fls_scheme) and segment bytes of one column chunk */
int flsgpu_chunk_info(const flsgpu_t *h, uint32_t rg, uint32_t col, int32_t *scheme, uint64_t *bytes);
/* Bounds of an integer / DATE column inside one row group, read off the compressed form: an FFOR vector holds
 * values in [base, base + 2^W), a CONSTANT chunk one value — the per-vector headers are a zone map that costs no
 * decode (row-group skipping for the filters the reference leaves switched off,
 * /root/reference/src/scanner/scan_fastlanes.cpp:154-155).  Needs no CUDA device.  *known = 1 and min <= every
 * value <= max (signed order for signed types and DATE; unsigned types are reported when they fit int64), or
 * *known = 0 when the chunk's scheme gives no bounds. */
int flsgpu_rowgroup_minmax(flsgpu_t *h, uint32_t rg, uint32_t col, int64_t *min, int64_t *max, int32_t *known);
/* Dictionary predicate on a string column (same row of SURVEY.md section 8f): does the row group's dictionary hold
 * the string bytes[0..len)?  A STR_DICT chunk answers from its dictionary segment and a STR_CONSTANT chunk from its
 * one value — read on the host, nothing is decoded; *known = 1 and *found = 0 means no row of the row group equals
 * the string, so `column = 'x'` can skip it.  *known = 0 for every other scheme (FSST-compressed dictionaries
 * included: the product has no CPU decoder to open them with).  Needs no CUDA device. */
int flsgpu_rowgroup_dict_contains(flsgpu_t *h, uint32_t rg, uint32_t col, const uint8_t *bytes, uint64_t len,
                                  int32_t *found, int32_t *known);

/* ---- decode ------------------------------------------------------------------------------------------- */
/* Stage the compressed bytes of row groups [g0,g1) into HBM and keep them until close / the next call. */
int flsgpu_make_resident(flsgpu_t *h, uint32_t g0, uint32_t g1);
/* Decode columns cols[0..n_cols) (NULL = all) of row groups [g0,g1).  Returns when the batch is complete.
 * Unless the range is resident, only the byte ranges of the listed columns' chunks are copied to the device
 * (flsgpu_timing.h2d_bytes says how many). */
int flsgpu_decode(flsgpu_t *h, uint32_t g0, uint32_t g1, const uint32_t *cols, uint32_t n_cols, uint32_t flags,
                  flsgpu_batch_t **out);
uint32_t flsgpu_batch_num_columns(const flsgpu_batch_t *b);
int flsgpu_batch_column(const flsgpu_batch_t *b, uint32_t i, flsgpu_column *out);
int flsgpu_batch_timing(const flsgpu_batch_t *b, flsgpu_timing *out);
/* Copy column i of a FLSGPU_OUT_DEVICE batch to caller memory: `values` receives padded_rows * width bytes
 * (fixed width) or payload_bytes bytes (strings); `offsets` (strings only) receives padded_rows + 1 entries. */
int flsgpu_batch_fetch(const flsgpu_batch_t *b, uint32_t i, void *values, uint32_t *offsets);
/* The same for a column that carries offsets64: padded_rows + 1 entries. */
int flsgpu_batch_fetch_offsets64(const flsgpu_batch_t *b, uint32_t i, uint64_t *offsets64);
/* The same for column i's validity bitmap: (padded_rows + 63) / 64 words; FLSGPU_E_RANGE when the column has none. */
int flsgpu_batch_fetch_validity(const flsgpu_batch_t *b, uint32_t i, uint64_t *bits);
/* Optional result gather (the reference merges its threads' DataChunks on the host,
 * src/scanner/scan_fastlanes.cpp:89-120): copy column i of a batch that holds device buffers (FLSGPU_OUT_DEVICE) into
 * buffers the caller owns on device `dst_device` — NVLink peer copies when the two GPUs can reach each other,
 * staged through the host by the driver otherwise.  Sizes are those of flsgpu_batch_fetch; `dst_offsets` receives
 * 32-bit entries, or 64-bit ones when the column carries offsets64.  NULL destinations are skipped.  Returns when
 * the copies are complete; *ms (may be NULL) = their device time. */
int flsgpu_batch_copy_to_device(const flsgpu_batch_t *b, uint32_t i, int32_t dst_device, void *dst_values,
                                void *dst_offsets, float *ms);
/* Returns the batch's buffers to the handle's pool. */
void flsgpu_batch_release(flsgpu_batch_t *b);

/* ---- filtered decode (late materialisation) -------------------------------------------------------------
 * SURVEY.md section 8f rank 3, second half.  The reference declares filter pushdown and leaves it switched off
 * (src/scanner/scan_fastlanes.cpp:154-155); DuckDB's hooks are table_function.hpp:377-379,414-422.
 * flsgpu_decode_filtered decodes the predicate columns and the projected columns of [g0,g1) into HBM, evaluates the
 * conjunction preds[0..n_preds) there, and hands back only the rows that pass, in row order: every column of the
 * batch has rows == padded_rows == the number of surviving rows, strings are re-based (offsets[0] = 0).  With
 * FLSGPU_OUT_HOST only the survivors cross PCIe.  Comparison semantics are DuckDB's: integers and DATE by value
 * (signed or unsigned as the column type says; the constant is i64, reinterpreted for UINT64), FLOAT / DOUBLE in
 * DuckDB's total order (NaN = NaN, NaN greater than everything else; the constant is f64), strings byte-wise for
 * EQ / NE only.  A predicate column need not be projected.  At most FLSGPU_MAX_PREDICATES predicates. */
enum flsgpu_cmp { FLSGPU_CMP_EQ = 0, FLSGPU_CMP_NE, FLSGPU_CMP_LT, FLSGPU_CMP_LE, FLSGPU_CMP_GT, FLSGPU_CMP_GE };
#define FLSGPU_MAX_PREDICATES 8
typedef struct {
  uint32_t col;         /* file column id */
  uint32_t op;          /* enum flsgpu_cmp */
  int64_t i64;          /* constant for integer / BOOLEAN / DATE columns */
  double f64;           /* constant for FLOAT / DOUBLE columns */
  const uint8_t *text;  /* constant for STR / BYTE_ARRAY columns (EQ / NE) */
  uint64_t text_len;
} flsgpu_predicate;
int flsgpu_decode_filtered(flsgpu_t *h, uint32_t g0, uint32_t g1, const uint32_t *cols, uint32_t n_cols,
                           const flsgpu_predicate *preds, uint32_t n_preds, uint32_t flags, flsgpu_batch_t **out);
/* Row ids (positions in the unpadded row range of [g0,g1)) of a filtered batch's rows: *n entries, host memory for
 * FLSGPU_OUT_HOST batches and device memory otherwise; valid until the batch is released.  *ids = NULL, *n = the
 * batch's row count for an unfiltered batch. */
int flsgpu_batch_selection(const flsgpu_batch_t *b, const uint64_t **ids, uint64_t *n);

/* ---- selection without allocation ----------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * flsgpu_select_fixed_noalloc runs the selection kernels of flsgpu_decode_filtered (mask, scan, compaction) for ONE
 * predicate `values[row] <op> constant` over a fixed-width column that is already decoded in HBM, with every scratch
 * buffer (mask words, per-vector counts and offsets, the total) in a file-scope `static __device__` array of
 * FLSGPU_NOALLOC_SCRATCH_BYTES that each GPU holds from the moment the library is loaded on it until exit.  It calls
 * no cudaMalloc / cudaFree / allocating library routine; the caller passes every input and output:
 *   device      CUDA device the pointers live on
 *   values      device, n_rows values of `type` (enum fls_type: a fixed-width type, FLS_T_* of fls_format.h)
 *   n_rows      at most FLSGPU_NOALLOC_MAX_ROWS, else FLSGPU_E_LIMIT and nothing is launched
 *   op, i64, f64  the comparison, as in flsgpu_predicate (i64 for integer / BOOLEAN / DATE, f64 for FLOAT / DOUBLE)
 *   out_values  device, room for n_rows values: receives the surviving values in row order (may be NULL)
 *   out_row_ids device, room for n_rows uint64: receives the surviving rows' indices in row order (may be NULL)
 *   out_count   device, one uint64: receives the number of surviving rows
 *   stream      a cudaStream_t (NULL = the default stream); the kernels are launched on it and NOT waited for
 * Every call on a device shares that device's scratch array, which keeps the previous call's contents (the entry
 * point clears the one word that must start at zero): calls that overlap on the GPU — two streams, two threads —
 * race; serialise them.  The first call on a device resolves the array's address (cudaGetSymbolAddress); call once
 * with n_rows = 0 before a graph capture to get that out of the way.  Errors: flsgpu_last_error(NULL). */
#define FLSGPU_NOALLOC_MAX_ROWS (4096u * 1024u)
#define FLSGPU_NOALLOC_SCRATCH_BYTES (256u + 4096u * (4u + 4u + 8u + 8u + 128u))
int flsgpu_select_fixed_noalloc(int32_t device, const void *values, int32_t type, uint64_t n_rows, uint32_t op,
                                int64_t i64, double f64, void *out_values, uint64_t *out_row_ids,
                                uint64_t *out_count, void *stream);

/* ---- instrumentation ---------------------------------------------------------------------------------- */
/* Per-kernel-family accumulators (needs FLSGPU_OPEN_TIME_KERNELS).  i in [0, flsgpu_num_kernel_stats). */
uint32_t flsgpu_num_kernel_stats(const flsgpu_t *h);
int flsgpu_kernel_stats(const flsgpu_t *h, uint32_t i, flsgpu_kernel_stat *out);
void flsgpu_reset_kernel_stats(flsgpu_t *h);
/* Write `bytes` bytes of a device scratch buffer (L2 flush between timed iterations). */
int flsgpu_flush_l2(flsgpu_t *h, uint64_t bytes);

#ifdef __cplusplus
}
#endif
#endif
