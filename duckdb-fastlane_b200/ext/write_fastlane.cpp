// write_fastlane.cpp — COPY <table|query> TO 'x.fls' (FORMAT FLS [, ROW_GROUP_SIZE n]) over libflsw's C ABI
// (include/flsw.h).  SURVEY.md section 8f rank 2: the reference declares this sink and leaves it a stub
// (/root/reference/src/writer/write_fastlane_stream.cpp: options :77-104, sink :125-143, registration :292-314;
// examples/basic_usage.sql:27-36 is the statement it is meant to serve).  Option names, their validation and the
// error texts follow the reference; the encoding itself is csrc/fls_writer.cpp (host code, not a GPU path).
//
// Types follow the reference's table (/root/reference/src/type_mapping.cpp:11-62): BOOLEAN, the eight integer types,
// FLOAT, DOUBLE, VARCHAR, DATE, BLOB map one to one; everything else is FALLBACK -> written as VARCHAR (:58-60).
// FastLanes null maps are ignored by the reference's reader (src/fastlanes_facade.cpp:125-172), so a NULL has no
// representation: the sink refuses it instead of writing a made-up value.
#include "write_fastlane.hpp"

#include <mutex>

#include "duckdb/common/exception.hpp"
#include "duckdb/common/string_util.hpp"
#include "duckdb/common/vector_operations/vector_operations.hpp"
#include "duckdb/function/copy_function.hpp"
#include "duckdb/main/extension_util.hpp"
#include "flsgpu.h"
#include "flsw.h"

namespace duckdb {
namespace ext_fastlane {
namespace {

constexpr idx_t kMaxRowGroup = 65536; // 64 FastLanes vectors of 1024 values (FORMAT.md)

int32_t FlsTypeOf(const LogicalType &t) {
	switch (t.id()) {
	case LogicalTypeId::BOOLEAN: return FLSGPU_T_BOOLEAN;
	case LogicalTypeId::TINYINT: return FLSGPU_T_INT8;
	case LogicalTypeId::SMALLINT: return FLSGPU_T_INT16;
	case LogicalTypeId::INTEGER: return FLSGPU_T_INT32;
	case LogicalTypeId::BIGINT: return FLSGPU_T_INT64;
	case LogicalTypeId::UTINYINT: return FLSGPU_T_UINT8;
	case LogicalTypeId::USMALLINT: return FLSGPU_T_UINT16;
	case LogicalTypeId::UINTEGER: return FLSGPU_T_UINT32;
	case LogicalTypeId::UBIGINT: return FLSGPU_T_UINT64;
	case LogicalTypeId::FLOAT: return FLSGPU_T_FLOAT;
	case LogicalTypeId::DOUBLE: return FLSGPU_T_DOUBLE;
	case LogicalTypeId::VARCHAR: return FLSGPU_T_STR;
	case LogicalTypeId::DATE: return FLSGPU_T_DATE;
	case LogicalTypeId::BLOB: return FLSGPU_T_BYTE_ARRAY;
	default: return 0; // FALLBACK: cast to VARCHAR
	}
}

idx_t WidthOf(int32_t fls_type) {
	switch (fls_type) {
	case FLSGPU_T_BOOLEAN:
	case FLSGPU_T_INT8:
	case FLSGPU_T_UINT8: return 1;
	case FLSGPU_T_INT16:
	case FLSGPU_T_UINT16: return 2;
	case FLSGPU_T_INT32:
	case FLSGPU_T_UINT32:
	case FLSGPU_T_FLOAT:
	case FLSGPU_T_DATE: return 4;
	case FLSGPU_T_INT64:
	case FLSGPU_T_UINT64:
	case FLSGPU_T_DOUBLE: return 8;
	default: return 0; // strings / blobs
	}
}

struct FastlaneWriteBindData : public TableFunctionData {
	vector<LogicalType> sql_types;
	vector<string> column_names;
	vector<int32_t> fls_types; // per column; FALLBACK columns are FLSGPU_T_STR here
	vector<bool> fallback;
	idx_t row_group_size = kMaxRowGroup;
	bool row_group_size_set = false;
	optional_idx row_group_size_bytes; // accepted for compatibility; row groups are cut by rows
	optional_idx row_groups_per_file;  // accepted for compatibility; one file is written
};

struct FastlaneWriteGlobalState : public GlobalFunctionData {
	std::mutex lock;
	flsw_writer *writer = nullptr;
	string file_path;
	idx_t rows = 0; // rows buffered for the current row group
	vector<vector<uint8_t>> data;   // per column: values / string payload
	vector<vector<uint64_t>> offs;  // per string column: offsets[rows + 1]
	vector<vector<uint64_t>> valid; // per column: validity bitmap of the buffered rows (bit = holds a value)
	vector<bool> has_null;          // per column: a NULL was buffered for the current row group
	~FastlaneWriteGlobalState() override {
		if (writer) {
			flsw_destroy(writer);
		}
	}
};

struct FastlaneWriteLocalState : public LocalFunctionData {};

unique_ptr<FunctionData> FastlaneWriteBind(ClientContext &, CopyFunctionBindInput &input, const vector<string> &names,
                                           const vector<LogicalType> &sql_types) {
	D_ASSERT(names.size() == sql_types.size());
	auto bind_data = make_uniq<FastlaneWriteBindData>();
	for (auto &option : input.info.options) {
		const auto loption = StringUtil::Lower(option.first);
		if (option.second.size() != 1) {
			throw BinderException("%s requires exactly one argument", StringUtil::Upper(loption));
		}
		if (loption == "row_group_size" || loption == "chunk_size") {
			if (bind_data->row_group_size_bytes.IsValid()) {
				throw BinderException("ROW_GROUP_SIZE and ROW_GROUP_SIZE_BYTES are mutually exclusive");
			}
			const auto n = option.second[0].GetValue<uint64_t>();
			if (n == 0) {
				throw BinderException("ROW_GROUP_SIZE must be at least 1");
			}
			// a row group is 1..64 whole FastLanes vectors (only the file's last vector may be partial): round up
			bind_data->row_group_size = MinValue<idx_t>((n + 1023) / 1024 * 1024, kMaxRowGroup);
			bind_data->row_group_size_set = true;
		} else if (loption == "row_group_size_bytes") {
			if (bind_data->row_group_size_set) {
				throw BinderException("ROW_GROUP_SIZE and ROW_GROUP_SIZE_BYTES are mutually exclusive");
			}
			auto roption = option.second[0];
			bind_data->row_group_size_bytes = roption.type().id() == LogicalTypeId::VARCHAR
			                                      ? DBConfig::ParseMemoryLimit(roption.ToString())
			                                      : roption.GetValue<uint64_t>();
		} else if (loption == "row_groups_per_file") {
			bind_data->row_groups_per_file = option.second[0].GetValue<uint64_t>();
		} else {
			throw BinderException("Unknown option for FastLanes: %s", StringUtil::Upper(loption));
		}
	}
	bind_data->sql_types = sql_types;
	bind_data->column_names = names;
	for (auto &t : sql_types) {
		const int32_t ft = FlsTypeOf(t);
		bind_data->fls_types.push_back(ft ? ft : FLSGPU_T_STR);
		bind_data->fallback.push_back(ft == 0);
	}
	return std::move(bind_data);
}

unique_ptr<GlobalFunctionData> FastlaneWriteInitializeGlobal(ClientContext &, FunctionData &bind_data_p,
                                                             const string &file_path) {
	auto &bind_data = bind_data_p.Cast<FastlaneWriteBindData>();
	auto result = make_uniq<FastlaneWriteGlobalState>();
	vector<const char *> names;
	for (auto &n : bind_data.column_names) {
		names.push_back(n.c_str());
	}
	result->writer = flsw_create((uint32_t)names.size(), names.data(), bind_data.fls_types.data());
	if (!result->writer) {
		throw IOException("Failed to create FastLanes file: " + file_path);
	}
	result->file_path = file_path;
	result->data.resize(names.size());
	result->offs.resize(names.size());
	result->valid.resize(names.size());
	result->has_null.assign(names.size(), false);
	for (idx_t c = 0; c < names.size(); c++) {
		if (WidthOf(bind_data.fls_types[c]) == 0) {
			result->offs[c].push_back(0);
		}
	}
	return std::move(result);
}

unique_ptr<LocalFunctionData> FastlaneWriteInitializeLocal(ExecutionContext &, FunctionData &) {
	return make_uniq<FastlaneWriteLocalState>();
}

void FlushRowGroup(FastlaneWriteGlobalState &g, const FastlaneWriteBindData &bind_data) {
	if (g.rows == 0) {
		return;
	}
	vector<const void *> cols;
	vector<const uint64_t *> offs, valid;
	for (idx_t c = 0; c < g.data.size(); c++) {
		valid.push_back(g.has_null[c] ? g.valid[c].data() : nullptr);
		const bool str = WidthOf(bind_data.fls_types[c]) == 0;
		static const uint8_t empty = 0;
		cols.push_back(g.data[c].empty() ? (const void *)&empty : (const void *)g.data[c].data());
		offs.push_back(str ? g.offs[c].data() : nullptr);
	}
	if (flsw_append_rowgroup_v(g.writer, (uint32_t)g.rows, cols.data(), offs.data(), valid.data())) {
		throw IOException("Failed to write FastLanes rowgroup: %s", flsw_error(g.writer));
	}
	g.rows = 0;
	for (idx_t c = 0; c < g.data.size(); c++) {
		g.data[c].clear();
		g.valid[c].clear();
		g.has_null[c] = false;
		if (!g.offs[c].empty()) {
			g.offs[c].assign(1, 0);
		}
	}
}

// rows [from, from + n) of the chunk -> the row group buffers
void AppendRows(FastlaneWriteGlobalState &g, const FastlaneWriteBindData &bind_data, DataChunk &input,
                vector<Vector *> &cols, idx_t from, idx_t n) {
	for (idx_t c = 0; c < cols.size(); c++) {
		UnifiedVectorFormat fmt;
		cols[c]->ToUnifiedFormat(input.size(), fmt);
		const idx_t wb = WidthOf(bind_data.fls_types[c]);
		g.valid[c].resize((g.rows + n + 63) / 64, 0);
		for (idx_t i = from; i < from + n; i++) {
			const idx_t idx = fmt.sel->get_index(i);
			const idx_t row = g.rows + (i - from);
			if (!fmt.validity.RowIsValid(idx)) { // a NULL: a filler value / an empty string, the validity bit stays 0
				g.has_null[c] = true;
				if (wb) {
					g.data[c].insert(g.data[c].end(), wb, 0);
				} else {
					g.offs[c].push_back(g.data[c].size());
				}
				continue;
			}
			g.valid[c][row >> 6] |= uint64_t(1) << (row & 63);
			if (wb) {
				const auto *src = fmt.data + idx * wb;
				g.data[c].insert(g.data[c].end(), src, src + wb);
			} else {
				const auto &s = UnifiedVectorFormat::GetData<string_t>(fmt)[idx];
				const auto *p = (const uint8_t *)s.GetData();
				g.data[c].insert(g.data[c].end(), p, p + s.GetSize());
				g.offs[c].push_back(g.data[c].size());
			}
		}
	}
	g.rows += n;
}

void FastlaneWriteSink(ExecutionContext &context, FunctionData &bind_data_p, GlobalFunctionData &gstate,
                       LocalFunctionData &, DataChunk &input) {
	auto &bind_data = bind_data_p.Cast<FastlaneWriteBindData>();
	auto &g = gstate.Cast<FastlaneWriteGlobalState>();
	// FALLBACK columns are written as their VARCHAR rendering (/root/reference/src/type_mapping.cpp:58-60)
	vector<unique_ptr<Vector>> casted;
	vector<Vector *> cols;
	for (idx_t c = 0; c < input.ColumnCount(); c++) {
		if (bind_data.fallback[c]) {
			casted.push_back(make_uniq<Vector>(LogicalType::VARCHAR, input.size()));
			VectorOperations::DefaultCast(input.data[c], *casted.back(), input.size());
			cols.push_back(casted.back().get());
		} else {
			cols.push_back(&input.data[c]);
		}
	}
	std::lock_guard<std::mutex> guard(g.lock);
	idx_t from = 0;
	while (from < input.size()) {
		const idx_t n = MinValue<idx_t>(input.size() - from, bind_data.row_group_size - g.rows);
		AppendRows(g, bind_data, input, cols, from, n);
		from += n;
		if (g.rows >= bind_data.row_group_size) {
			FlushRowGroup(g, bind_data);
		}
	}
}

void FastlaneWriteCombine(ExecutionContext &, FunctionData &, GlobalFunctionData &, LocalFunctionData &) {
}

void FastlaneWriteFinalize(ClientContext &, FunctionData &bind_data_p, GlobalFunctionData &gstate) {
	auto &bind_data = bind_data_p.Cast<FastlaneWriteBindData>();
	auto &g = gstate.Cast<FastlaneWriteGlobalState>();
	std::lock_guard<std::mutex> guard(g.lock);
	FlushRowGroup(g, bind_data);
	if (flsw_write_file(g.writer, g.file_path.c_str())) {
		throw IOException("Failed to finalize FastLanes file: %s", flsw_error(g.writer));
	}
}

CopyFunctionExecutionMode FastlaneWriteExecutionMode(bool, bool) {
	return CopyFunctionExecutionMode::REGULAR_COPY_TO_FILE; // one sink at a time: rows keep their order
}

} // namespace

void RegisterFastlaneStreamCopyFunction(DatabaseInstance &db) {
	CopyFunction function("fls");
	function.copy_to_bind = FastlaneWriteBind;
	function.copy_to_initialize_global = FastlaneWriteInitializeGlobal;
	function.copy_to_initialize_local = FastlaneWriteInitializeLocal;
	function.copy_to_sink = FastlaneWriteSink;
	function.copy_to_combine = FastlaneWriteCombine;
	function.copy_to_finalize = FastlaneWriteFinalize;
	function.execution_mode = FastlaneWriteExecutionMode;
	function.extension = "fls";
	ExtensionUtil::RegisterFunction(db, function);

	function.name = "fastlane";
	function.extension = "fastlane";
	ExtensionUtil::RegisterFunction(db, function);
}

} // namespace ext_fastlane
} // namespace duckdb
