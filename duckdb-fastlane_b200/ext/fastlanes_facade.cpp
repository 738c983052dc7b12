// fastlanes_facade.cpp — see include/fastlanes_facade.hpp.  Mirrors /root/reference/src/fastlanes_facade.cpp with
// the FastLanes objects (Connection / TableReader / RowgroupReader / Rowgroup, :12-20) replaced by one
// flsgpu_host::FlsScanner, i.e. by the C ABI of include/flsgpu.h.
#include "fastlanes_facade.hpp"

#include <cstring>

#include "duckdb/common/types/vector.hpp"
#include "duckdb/common/types/vector_buffer.hpp"
#include "utf8proc_wrapper.hpp"
#include "fls_scan.hpp"

namespace duckdb {

class FastLanesFacade::Impl {
public:
	flsgpu_host::FlsScanner scanner;
	std::vector<flsgpu_host::ColumnSlice> slices;
	std::vector<idx_t> projection; // file column id per output column; empty = identity
	std::string error;
	bool initialized = false;
	bool count_only = false; // the scan needs no column: serve row counts from the footer, decode nothing
	uint64_t count_left = 0;
};

LogicalType FastLanesTypeToDuckDB(int32_t t) {
	switch (t) {
	case FLSGPU_T_BOOLEAN: return LogicalType::BOOLEAN;
	case FLSGPU_T_INT8: return LogicalType::TINYINT;
	case FLSGPU_T_INT16: return LogicalType::SMALLINT;
	case FLSGPU_T_INT32: return LogicalType::INTEGER;
	case FLSGPU_T_INT64: return LogicalType::BIGINT;
	case FLSGPU_T_UINT8: return LogicalType::UTINYINT;
	case FLSGPU_T_UINT16: return LogicalType::USMALLINT;
	case FLSGPU_T_UINT32: return LogicalType::UINTEGER;
	case FLSGPU_T_UINT64: return LogicalType::UBIGINT;
	case FLSGPU_T_FLOAT: return LogicalType::FLOAT;
	case FLSGPU_T_DOUBLE: return LogicalType::DOUBLE;
	case FLSGPU_T_STR: return LogicalType::VARCHAR;
	case FLSGPU_T_DATE: return LogicalType::DATE;
	case FLSGPU_T_BYTE_ARRAY: return LogicalType::BLOB;
	default: return LogicalType::VARCHAR; // the reference's FALLBACK (src/type_mapping.cpp:100-103)
	}
}

FastLanesFacade::FastLanesFacade() : pImpl(std::make_unique<Impl>()) {
}
FastLanesFacade::~FastLanesFacade() {
	closeFile();
}

bool FastLanesFacade::openFile(const std::string &file_path) {
	return openFile(file_path, std::vector<int>());
}

bool FastLanesFacade::openFile(const std::string &file_path, const std::vector<int> &devices) {
	closeFile();
	if (!pImpl->scanner.open(file_path, devices)) {
		pImpl->error = pImpl->scanner.error();
		return false;
	}
	pImpl->error.clear();
	pImpl->initialized = true;
	return true;
}

std::vector<LogicalType> FastLanesFacade::getColumnTypes() {
	std::vector<LogicalType> r;
	for (auto t : pImpl->scanner.columnTypes()) {
		r.push_back(FastLanesTypeToDuckDB(t));
	}
	return r;
}

std::vector<std::string> FastLanesFacade::getColumnNames() {
	return pImpl->scanner.columnNames();
}

bool FastLanesFacade::setProjection(const std::vector<idx_t> &column_ids) {
	std::vector<uint32_t> cols(column_ids.begin(), column_ids.end());
	if (!pImpl->scanner.setProjection(cols)) {
		pImpl->error = pImpl->scanner.error();
		return false;
	}
	pImpl->projection = column_ids;
	return true;
}

bool FastLanesFacade::setRowGroupRange(idx_t g0, idx_t g1) {
	pImpl->scanner.rewind();
	if (!pImpl->scanner.setRowGroupRange((uint32_t)g0, (uint32_t)g1)) {
		pImpl->error = pImpl->scanner.error();
		return false;
	}
	pImpl->error.clear();
	pImpl->count_left = pImpl->scanner.rowsInRange();
	return true;
}
void FastLanesFacade::setCountOnly(bool on) {
	pImpl->count_only = on;
	pImpl->count_left = pImpl->scanner.rowsInRange();
}
idx_t FastLanesFacade::rowGroupCount() const {
	return pImpl->scanner.rowGroupCount();
}

idx_t FastLanesFacade::pruneRowGroups(const std::vector<flsgpu_host::ZonePredicate> &predicates) {
	return pImpl->scanner.pruneRowGroups(predicates);
}

bool FastLanesFacade::setDeviceFilter(const std::vector<flsgpu_host::ZonePredicate> &predicates) {
	if (!pImpl->scanner.setDeviceFilter(predicates)) {
		pImpl->error = pImpl->scanner.error();
		return false;
	}
	return true;
}

// no byte of [p, p + n) has its top bit set: ASCII is valid UTF-8, and eight bytes at a time is an order of magnitude
// faster than decoding code points
static bool AllAscii(const uint8_t *p, uint64_t n) {
	uint64_t acc = 0, i = 0;
	for (; i + 32 <= n; i += 32) {
		uint64_t w[4];
		memcpy(w, p + i, 32);
		acc |= w[0] | w[1] | w[2] | w[3];
	}
	for (; i < n; i++) {
		acc |= (uint64_t)p[i] << 56;
	}
	return (acc & 0x8080808080808080ull) == 0;
}

template <class T>
static void CopyFixed(Vector &out, const flsgpu_host::ColumnSlice &s, idx_t rows) {
	memcpy(FlatVector::GetData<T>(out), s.data, rows * sizeof(T));
}

bool FastLanesFacade::readNextChunk(DataChunk &result) {
	if (!pImpl->initialized) {
		return false;
	}
	if (pImpl->count_only) {
		// count(*): DuckDB asks for the row-id pseudo column alone and reads none of its values; the row counts are in
		// the footer, so nothing is staged, decoded or copied (the reference boxes every cell of every column here,
		// /root/reference/src/fastlanes_facade.cpp:125-172)
		if (pImpl->count_left == 0) {
			return false;
		}
		const uint64_t n = MinValue<uint64_t>(STANDARD_VECTOR_SIZE, pImpl->count_left);
		pImpl->count_left -= n;
		for (idx_t c = 0; c < result.ColumnCount(); c++) {
			result.data[c].SetVectorType(VectorType::CONSTANT_VECTOR);
			ConstantVector::SetNull(result.data[c], true);
		}
		result.SetCardinality(n);
		return true;
	}
	uint64_t rows = 0;
	if (!pImpl->scanner.next(STANDARD_VECTOR_SIZE, rows, pImpl->slices)) {
		pImpl->error = pImpl->scanner.error();
		return false;
	}
	// like the reference (src/fastlanes_facade.cpp:102-104) never touch more columns than the chunk has
	const idx_t ncol = MinValue<idx_t>(result.ColumnCount(), pImpl->slices.size());
	result.SetCardinality(rows);
	for (idx_t c = 0; c < ncol; c++) {
		auto &s = pImpl->slices[c];
		auto &out = result.data[c];
		out.SetVectorType(VectorType::FLAT_VECTOR);
		switch (s.type) {
		case FLSGPU_T_BOOLEAN: { // bytes come from a file: anything non-zero is true
			auto dst = FlatVector::GetData<bool>(out);
			auto src = (const uint8_t *)s.data;
			for (idx_t r = 0; r < rows; r++) {
				dst[r] = src[r] != 0;
			}
			break;
		}
		case FLSGPU_T_INT8:
		case FLSGPU_T_UINT8: CopyFixed<uint8_t>(out, s, rows); break;
		case FLSGPU_T_INT16:
		case FLSGPU_T_UINT16: CopyFixed<uint16_t>(out, s, rows); break;
		case FLSGPU_T_INT32:
		case FLSGPU_T_UINT32:
		case FLSGPU_T_FLOAT:
		case FLSGPU_T_DATE: CopyFixed<uint32_t>(out, s, rows); break;
		case FLSGPU_T_INT64:
		case FLSGPU_T_UINT64:
		case FLSGPU_T_DOUBLE: CopyFixed<uint64_t>(out, s, rows); break;
		case FLSGPU_T_STR:
		case FLSGPU_T_BYTE_ARRAY: {
			// strings <= 12 bytes are inlined in string_t; longer ones are copied into the vector's heap in ONE
			// allocation per chunk (the chunk stays valid after the scanner recycles its pinned batch buffers).
			// The offsets come from a file: they are checked before they address anything, and VARCHAR bytes are
			// validated as UTF-8 like the reference's Value(std::string) does (src/fastlanes_facade.cpp:157-172).
			auto dst = FlatVector::GetData<string_t>(out);
			const uint32_t first = s.offsets[0], last = s.offsets[rows];
			bool ok = first <= last;
			idx_t long_bytes = 0;
			for (idx_t r = 0; r < rows; r++) { // one pass: ascending offsets, bytes of the strings that are not inlined
				const uint32_t a = s.offsets[r], b = s.offsets[r + 1];
				ok = ok && a <= b;
				const uint32_t len = b - a;
				long_bytes += len > string_t::INLINE_LENGTH ? len : 0;
			}
			if (!ok) {
				pImpl->error = "FastLanes file is damaged: string offsets of a chunk are not ascending";
				return false;
			}
			if (s.type == FLSGPU_T_STR && last > first && !AllAscii(s.payload + first, last - first) &&
			    Utf8Proc::Analyze(const_char_ptr_cast(s.payload + first), last - first) == UnicodeType::INVALID) {
				pImpl->error = "FastLanes file holds a VARCHAR value that is not valid UTF-8";
				return false;
			}
			// the chunk's strings are one contiguous run of payload bytes: ONE copy into the vector's heap, then the
			// long strings point into it (short ones are inlined by string_t's constructor)
			const char *heap = nullptr;
			if (long_bytes) {
				auto buf = StringVector::GetStringBuffer(out).AllocateShrinkableBuffer(last - first);
				memcpy(buf, s.payload + first, last - first);
				heap = const_char_ptr_cast(buf);
			}
			for (idx_t r = 0; r < rows; r++) {
				const uint32_t a = s.offsets[r], len = s.offsets[r + 1] - a;
				dst[r] = len <= string_t::INLINE_LENGTH ? string_t(const_char_ptr_cast(s.payload + a), len) : string_t(heap + (a - first), len);
			}
			break;
		}
		default: // unknown column type -> NULLs, as the reference does (src/fastlanes_facade.cpp:173-181)
			for (idx_t r = 0; r < rows; r++) {
				FlatVector::SetNull(out, r, true);
			}
		}
		if (s.validity) { // NULLs: the batch's bitmap, from bit validity_bit0 on, into the vector's validity mask
			auto &mask = FlatVector::Validity(out);
			for (idx_t r = 0; r < rows; r++) {
				const uint64_t bit = s.validity_bit0 + r;
				if (!((s.validity[bit >> 6] >> (bit & 63)) & 1)) {
					mask.SetInvalid(r);
				}
			}
		}
	}
	return true;
}

void FastLanesFacade::closeFile() {
	pImpl->scanner.close();
	pImpl->slices.clear();
	pImpl->projection.clear();
	pImpl->initialized = false;
}

bool FastLanesFacade::isValid() const {
	return pImpl->initialized;
}
idx_t FastLanesFacade::rowCount() const {
	return pImpl->scanner.rowsInRange();
}
idx_t FastLanesFacade::rowsRead() const {
	return pImpl->scanner.rowsServed();
}
const std::string &FastLanesFacade::lastError() const {
	return pImpl->error;
}

} // namespace duckdb
