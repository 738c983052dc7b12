// fastlane_extension.cpp — entry points of the `fastlane` DuckDB extension, GPU read path only.
// Keeps the two C symbols DuckDB's loader resolves (`fastlane_init`, `fastlane_version`,
// /root/reference/src/fastlane_extension.cpp:113-124) and the Load() -> ScanFastLanes::Register chain (:98-101).
// Of the reference's dead LoadInternal (:44-94) the stream COPY writer is registered too (ext/write_fastlane.cpp);
// the csv/json converters are out of scope.
#define DUCKDB_EXTENSION_MAIN

#include "fastlane_extension.hpp"

#include "duckdb.hpp"
#include "scan_fastlanes.hpp"
#include "write_fastlane.hpp"

namespace duckdb {

void FastlaneExtension::Load(DuckDB &db) {
	ScanFastLanes::Register(*db.instance);
	ext_fastlane::RegisterFastlaneStreamCopyFunction(*db.instance); // COPY ... TO 'x.fls' (FORMAT FLS)
}

std::string FastlaneExtension::Name() {
	return "fastlane";
}

std::string FastlaneExtension::Version() const {
#ifdef EXT_VERSION_FASTLANE
	return EXT_VERSION_FASTLANE;
#else
	return "";
#endif
}

} // namespace duckdb

extern "C" {

DUCKDB_EXTENSION_API void fastlane_init(duckdb::DatabaseInstance &db) {
	duckdb::FastlaneExtension extension;
	duckdb::DuckDB db_wrapper(db);
	extension.Load(db_wrapper);
}

DUCKDB_EXTENSION_API const char *fastlane_version() {
	return duckdb::DuckDB::LibraryVersion();
}
}

#ifndef DUCKDB_EXTENSION_MAIN
#error DUCKDB_EXTENSION_MAIN not defined
#endif
