// duckdb_sql_check — a DuckDB v1.3.2 instance with the GPU read path registered exactly as the extension's Load()
// registers it (ScanFastLanes::Register, ext/fastlane_extension.cpp), driven with SQL.  One JSON line per
// statement on stdout: {"sql": ..., "error": "..."} or {"sql": ..., "columns": [...], "rows": [[...], ...]}
// (values rendered with Value::ToString).  tests/test_duckdb_sql.py runs the reference's own sqllogictest
// statements (/root/reference/test/sql/fastlane.test) and typed multi-column scans through it.
//
// Test infrastructure for the drop-in boundary: it links libduckdb built from the reference tree's DuckDB sources
// (dev container only; the binary and the library travel to the GPU box with the snapshot) and libflsgpu.so.
//
//   duckdb_sql_check "SELECT count(*) FROM read_fls('x.fls')" ["SELECT ..."]...
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "duckdb.hpp"
#include "scan_fastlanes.hpp"
#include "write_fastlane.hpp"

static std::string esc(const std::string &s) {
	std::string o;
	for (unsigned char c : s) {
		if (c == '"' || c == '\\') {
			o += '\\';
			o += (char)c;
		} else if (c == '\n') {
			o += "\\n";
		} else if (c == '\t') {
			o += "\\t";
		} else if (c < 0x20) {
			char b[8];
			snprintf(b, sizeof b, "\\u%04x", c);
			o += b;
		} else {
			o += (char)c;
		}
	}
	return o;
}

int main(int argc, char **argv) {
	// FLS_SQL_CHECK_LOAD=<path of fastlane.duckdb_extension>: nothing is registered statically, the functions come from
	// LOAD '<path>' (an unsigned extension), i.e. through DuckDB's own extension loader
	const char *load = getenv("FLS_SQL_CHECK_LOAD");
	duckdb::DBConfig config;
	if (load) {
		config.options.allow_unsigned_extensions = true;
	}
	duckdb::DuckDB db(nullptr, &config);
	duckdb::Connection con(db);
	if (load) {
		auto res = con.Query(std::string("LOAD '") + load + "'");
		if (res->HasError()) {
			printf("{\"sql\": \"LOAD\", \"error\": \"%s\"}\n", esc(res->GetError()).c_str());
			return 1;
		}
	} else {
		duckdb::ScanFastLanes::Register(*db.instance);
		duckdb::ext_fastlane::RegisterFastlaneStreamCopyFunction(*db.instance);
	}
	int rc = 0;
	for (int i = 1; i < argc; i++) {
		fprintf(stderr, "sql_check: statement %d\n", i - 1);
		const auto t0 = std::chrono::steady_clock::now();
		auto res = con.Query(argv[i]);
		const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
		printf("{\"sql\": \"%s\", \"ms\": %.3f, ", esc(argv[i]).c_str(), ms);
		if (res->HasError()) {
			printf("\"error\": \"%s\"}\n", esc(res->GetError()).c_str());
			rc = 1;
			continue;
		}
		printf("\"columns\": [");
		for (duckdb::idx_t c = 0; c < res->ColumnCount(); c++) {
			printf("%s\"%s\"", c ? ", " : "", esc(res->ColumnName(c)).c_str());
		}
		printf("], \"types\": [");
		for (duckdb::idx_t c = 0; c < res->ColumnCount(); c++) {
			printf("%s\"%s\"", c ? ", " : "", esc(res->types[c].ToString()).c_str());
		}
		printf("], \"rows\": [");
		for (duckdb::idx_t r = 0; r < res->RowCount(); r++) {
			printf("%s[", r ? ", " : "");
			for (duckdb::idx_t c = 0; c < res->ColumnCount(); c++) {
				auto v = res->GetValue(c, r);
				printf("%s\"%s\"", c ? ", " : "", esc(v.IsNull() ? "NULL" : v.ToString()).c_str());
			}
			printf("]");
		}
		printf("]}\n");
	}
	return rc;
}
