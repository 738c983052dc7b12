// write_fastlane.hpp — registration hook of COPY ... TO 'x.fls' (FORMAT FLS).
// Same entry point name as the reference's (dead) stream writer:
// RegisterFastlaneStreamCopyFunction, /root/reference/src/writer/write_fastlane_stream.cpp:292-314.
#pragma once

#include "duckdb.hpp"

namespace duckdb {
namespace ext_fastlane {

void RegisterFastlaneStreamCopyFunction(DatabaseInstance &db);

} // namespace ext_fastlane
} // namespace duckdb
