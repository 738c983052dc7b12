// fls_scan_check — drives FlsScanner the way FastLanesScan drives the facade (one <= 2048-row chunk per call,
// /root/reference/src/scan_fastlanes.cpp:65-77) and prints one JSON line: row count and, per projected column,
// a digest "A:B" of the served values: A = sum of byte_k * (k + 1) over the column's value bytes in row order
// (fixed width: the values; strings: the payload bytes), B = the same over the strings' lengths (0 for fixed
// width), both mod 2^64.  tests/test_ext_scan.py compares them with digests of the CPU oracle's decode.
//
//   fls_scan_check <file.fls> [--cols 3,0,7] [--chunk 2048] [--batch 16] [--range g0:g1] [--device 0]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "fls_scan.hpp"

struct Digest {
  uint64_t a = 0, ka = 0, b = 0, kb = 0;
  void bytes(const void *p, size_t n) {
    const uint8_t *x = (const uint8_t *)p;
    for (size_t i = 0; i < n; ++i) a += (uint64_t)x[i] * ++ka;
  }
  void length(uint32_t len) { b += (uint64_t)len * ++kb; }
};

int main(int argc, char **argv) {
  if (argc < 2) { fprintf(stderr, "usage: fls_scan_check <file.fls> [--cols a,b] [--chunk n] [--batch n] [--range g0:g1]\n"); return 2; }
  std::vector<uint32_t> cols;
  uint64_t chunk = 2048;
  uint32_t batch = 0, g0 = 0, g1 = 0;
  bool have_range = false;
  int device = 0;
  for (int i = 2; i + 1 < argc; i += 2) {
    const std::string k = argv[i];
    const char *v = argv[i + 1];
    if (k == "--cols") { for (const char *s = v; *s;) { cols.push_back((uint32_t)strtoul(s, (char **)&s, 10)); if (*s == ',') ++s; } }
    else if (k == "--chunk") chunk = strtoull(v, nullptr, 10);
    else if (k == "--batch") batch = (uint32_t)strtoul(v, nullptr, 10);
    else if (k == "--device") device = atoi(v);
    else if (k == "--range") { char *e; g0 = (uint32_t)strtoul(v, &e, 10); g1 = (uint32_t)strtoul(e + 1, nullptr, 10); have_range = true; }
  }
  flsgpu_host::FlsScanner sc;
  if (!sc.open(argv[1], device, batch)) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
  if (!cols.empty() && !sc.setProjection(cols)) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
  if (have_range && !sc.setRowGroupRange(g0, g1)) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
  const size_t nc = cols.empty() ? sc.columnCount() : cols.size();
  std::vector<Digest> dig(nc);
  std::vector<flsgpu_host::ColumnSlice> out;
  uint64_t rows = 0, total = 0, chunks = 0;
  while (sc.next(chunk, rows, out)) {
    for (size_t c = 0; c < nc; ++c) {
      const auto &s = out[c];
      if (s.width) dig[c].bytes(s.data, rows * s.width);
      else
        for (uint64_t r = 0; r < rows; ++r) {
          const uint32_t len = s.offsets[r + 1] - s.offsets[r];
          dig[c].length(len);
          dig[c].bytes(s.payload + s.offsets[r], len);
        }
    }
    total += rows;
    ++chunks;
  }
  if (!sc.error().empty()) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
  printf("{\"rows\": %llu, \"chunks\": %llu, \"columns\": %zu, \"names\": [", (unsigned long long)total, (unsigned long long)chunks, nc);
  for (size_t c = 0; c < nc; ++c) printf("%s\"%s\"", c ? ", " : "", sc.columnNames()[cols.empty() ? c : cols[c]].c_str());
  printf("], \"digests\": [");
  for (size_t c = 0; c < nc; ++c) printf("%s\"%016llx:%016llx\"", c ? ", " : "", (unsigned long long)dig[c].a, (unsigned long long)dig[c].b);
  printf("]}\n");
  return 0;
}
