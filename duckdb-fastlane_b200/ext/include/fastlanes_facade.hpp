//===----------------------------------------------------------------------===//
// fastlanes_facade.hpp — FastLanesFacade over libflsgpu (B200 decode) instead of the FastLanes CPU reader.
//
// Same class, namespace and reading methods as the reference's facade
// (/root/reference/src/include/fastlanes_facade.hpp:22-47): openFile / getColumnTypes / getColumnNames /
// readNextChunk / closeFile / isValid, PIMPL so no decoder header leaks into DuckDB translation units.
// Differences, all on purpose:
//   * openFile parses the footer only — the reference materialises row group 0 on the CPU at bind AND at init
//     (/root/reference/src/fastlanes_facade.cpp:41-48, called from src/scan_fastlanes.cpp:41 and :57);
//   * readNextChunk writes typed flat vectors straight from the decoded column arrays instead of one boxed
//     Value per cell cast to VARCHAR (/root/reference/src/fastlanes_facade.cpp:125-172), and serves every row
//     group of the file, not only the first;
//   * getColumnTypes / getColumnNames are implemented (declared but never defined in the reference);
//   * setProjection: projection pushdown (the reference's dead scanner sets the flag,
//     /root/reference/src/scanner/scan_fastlanes.cpp:153, but never uses column_ids).
// The writing half of the reference class (createFile / writeChunk / finalizeFile) is declared there and never
// implemented; COPY ... TO binds the writer's C ABI directly (ext/write_fastlane.cpp).
//===----------------------------------------------------------------------===//
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "duckdb/common/types.hpp"
#include "duckdb/common/types/data_chunk.hpp"
#include "fls_scan.hpp"

namespace duckdb {

class FastLanesFacade {
private:
	class Impl;
	std::unique_ptr<Impl> pImpl;

public:
	FastLanesFacade();
	~FastLanesFacade();

	// Reading operations
	bool openFile(const std::string &file_path);
	//! The same on several GPUs (CUDA device ordinals): batches of row groups are decoded on the devices in turn and
	//! served in file order.  An empty list means device 0.
	bool openFile(const std::string &file_path, const std::vector<int> &devices);
	std::vector<LogicalType> getColumnTypes();
	std::vector<std::string> getColumnNames();
	//! Projected file column ids, in output order; call between openFile and the first readNextChunk.
	bool setProjection(const std::vector<idx_t> &column_ids);
	//! Skip the row groups whose FFOR / DELTA bounds rule out `column <op> constant` (a conjunction); returns how
	//! many were skipped.  The engine keeps the filters: this only avoids decoding what cannot match.
	idx_t pruneRowGroups(const std::vector<flsgpu_host::ZonePredicate> &predicates);
	//! Evaluate the conjunction on the GPU and serve only the rows that pass (flsgpu_decode_filtered); the engine
	//! keeps the filters.  Call between openFile and the first readNextChunk.
	bool setDeviceFilter(const std::vector<flsgpu_host::ZonePredicate> &predicates);
	//! One shard of a parallel scan: aim the (possibly finished) reader at row groups [g0, g1) of the open file.
	//! Projection and device filter stay; row-group pruning must be asked for again.
	bool setRowGroupRange(idx_t g0, idx_t g1);
	idx_t rowGroupCount() const;
	//! A scan that needs no column (count(*)): readNextChunk serves the row counts of the current row-group range
	//! from the footer and decodes nothing.
	void setCountOnly(bool on);
	//! Fills `result` with the next <= STANDARD_VECTOR_SIZE rows; false at the end of the file or on error.
	bool readNextChunk(DataChunk &result);
	void closeFile();

	// Utility
	bool isValid() const;
	idx_t rowCount() const;
	idx_t rowsRead() const;
	//! Text of the last failure ("Failed to open FastLanes file: <path>" for a bad path).
	const std::string &lastError() const;
};

//! FastLanes data type (enum flsgpu_type = the reference's type table, src/type_mapping.cpp:64-105) -> LogicalType
LogicalType FastLanesTypeToDuckDB(int32_t fls_type);

} // namespace duckdb
