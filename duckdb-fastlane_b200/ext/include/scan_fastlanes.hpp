// scan_fastlanes.hpp — registration hook of the read_fls / scan_fastlanes table function.
// Same class and method as the reference (/root/reference/src/scan_fastlanes.hpp; used at
// /root/reference/src/fastlane_extension.cpp:98-101).
#pragma once

#include "duckdb.hpp"

namespace duckdb {

class ScanFastLanes {
public:
	static void Register(DatabaseInstance &db);
};

} // namespace duckdb
