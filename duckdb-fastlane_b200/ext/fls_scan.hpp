// fls_scan.hpp — host-side scan cursor over the C ABI (include/flsgpu.h), free of DuckDB types.
//
// This is the state the reference keeps in FastLanesFacade::Impl (/root/reference/src/fastlanes_facade.cpp:12-20:
// connection, table reader, row-group reader, materialised row group, current_row/total_rows) rebuilt on top of
// libflsgpu: a row cursor over batches of row groups decoded on the GPU into pinned host memory.  While the
// caller consumes batch i in <= 2048-row slices, batch i+1 is decoded by a worker thread, so the H2D copy, the
// kernels and the D2H copy of the next batch overlap the engine's processing of the current one.
//
// The DuckDB-facing classes (fastlanes_facade.hpp, scan_fastlanes.cpp) sit on top of this; tests drive it
// directly through ext/fls_scan_check.cpp.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "flsgpu.h"

namespace flsgpu_host {

// A run of rows of one column, valid until the next call to FlsScanner::next()/close().
struct ColumnSlice {
  int32_t type = 0;        // enum flsgpu_type
  uint32_t width = 0;      // bytes per value; 0 for strings
  const void *data = nullptr;          // fixed width: `rows` values
  const uint32_t *offsets = nullptr;   // strings: rows + 1 offsets into payload (offsets[0] need not be 0)
  const uint8_t *payload = nullptr;
  // NULLs: bit (validity_bit0 + r) of `validity` (64 rows per word, LSB first) = row r of the run holds a value;
  // null pointer = the column has no NULLs in this batch
  const uint64_t *validity = nullptr;
  uint64_t validity_bit0 = 0;
};

// A comparison `column <op> constant` on an integer / DATE column, used to skip row groups whose bounds
// (flsgpu_rowgroup_minmax: read off the FFOR / DELTA vector headers, no decode) cannot satisfy it.
enum ZoneOp { ZONE_EQ, ZONE_NE, ZONE_LT, ZONE_LE, ZONE_GT, ZONE_GE };
struct ZonePredicate {
  uint32_t col = 0;  // file column id
  ZoneOp op = ZONE_EQ;
  int64_t value = 0;
  // string columns: `column = text` against the row group's dictionary (flsgpu_rowgroup_dict_contains); pruning
  // uses ZONE_EQ only, the device filter ZONE_EQ and ZONE_NE
  bool is_text = false;
  std::string text;
  // FLOAT / DOUBLE columns: device filter only (their chunks carry no bounds)
  bool is_double = false;
  double dvalue = 0;
};

class FlsScanner {
 public:
  FlsScanner();
  ~FlsScanner();
  FlsScanner(const FlsScanner &) = delete;
  FlsScanner &operator=(const FlsScanner &) = delete;

  // Parses the footer only (no decode, unlike the reference's openFile which materialises row group 0).
  // On failure returns false and error() holds the text the reference's bind raises
  // ("Failed to open FastLanes file: <path>", /root/reference/src/scan_fastlanes.cpp:42).
  bool open(const std::string &path, int device = 0, uint32_t batch_rowgroups = 0);
  // Several GPUs behind one cursor (`gpus := [0, 1, ..]` on the table function): one handle and
  // one decode in flight per device, batches of row groups dealt to the devices round-robin and served in file order.
  // A device may be listed twice (two decodes in flight on it).
  bool open(const std::string &path, const std::vector<int> &devices, uint32_t batch_rowgroups = 0);
  uint32_t deviceCount() const;
  void close();
  bool isOpen() const;
  const std::string &error() const;

  uint32_t columnCount() const;
  uint64_t rowCount() const;
  uint32_t rowGroupCount() const;
  const std::vector<std::string> &columnNames() const;
  const std::vector<int32_t> &columnTypes() const;  // enum flsgpu_type per column

  // Projection pushdown: decode only these columns (file column ids, any order, duplicates allowed).
  // Must be called before the first next(); an empty list means "all columns".
  bool setProjection(const std::vector<uint32_t> &cols);
  // Restrict the scan to row groups [g0, g1) — one shard of a multi-GPU / multi-thread scan.
  bool setRowGroupRange(uint32_t g0, uint32_t g1);
  // After a scan has ended (or before it began): back to the state right after open(), handles and their buffer pools
  // kept, so that setRowGroupRange / pruneRowGroups can aim the cursor at another shard.  Projection and device filter stay.
  void rewind();

  // Row-group skipping: a row group is dropped from the scan when one of the predicates
  // (a conjunction) is false for every value its bounds allow.  The caller still applies the predicates to the rows
  // it gets — this only saves the H2D copy and the decode of row groups that cannot contribute.  Call after
  // setRowGroupRange / before the first next().  Returns the number of row groups skipped.
  uint32_t pruneRowGroups(const std::vector<ZonePredicate> &preds);
  uint32_t rowGroupsSkipped() const;
  // Late materialisation (same SURVEY row, second half): the conjunction is evaluated on the GPU after the decode
  // (flsgpu_decode_filtered) and only the rows that pass are copied to the host and served by next() — the runs it
  // returns then hold no row that fails a predicate, and may span row groups.  At most FLSGPU_MAX_PREDICATES
  // predicates are taken (the first ones); the caller may re-apply all of them.  Before the first next().
  bool setDeviceFilter(const std::vector<ZonePredicate> &preds);
  uint64_t rowsFilteredOut() const;

  // Next run of at most max_rows rows (never crossing a row-group boundary).  Returns false at the end of the
  // scan or on error (error() non-empty).  out[i] describes projected column i.
  bool next(uint64_t max_rows, uint64_t &rows, std::vector<ColumnSlice> &out);
  // Rows served so far / rows in the scan range: table_scan_progress.
  uint64_t rowsServed() const;
  uint64_t rowsInRange() const;

 private:
  struct Impl;
  std::unique_ptr<Impl> p;
};

}  // namespace flsgpu_host
