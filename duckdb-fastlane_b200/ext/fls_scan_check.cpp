// fls_scan_check — drives FlsScanner the way FastLanesScan drives the facade (one <= 2048-row chunk per call,
// /root/reference/src/scan_fastlanes.cpp:65-77) and prints one JSON line: row count and, per projected column,
// a digest "A:B" of the served values: A = sum of byte_k * (k + 1) over the column's value bytes in row order
// (fixed width: the values; strings: the payload bytes), B = the same over the strings' lengths (0 for fixed
// width), both mod 2^64.  tests/test_ext_scan.py compares them with digests of the CPU oracle's decode.
//
//   fls_scan_check <file.fls> [--cols 3,0,7] [--chunk 2048] [--batch 16] [--range g0:g1] [--device 0] [--gpus 0,1]
//                  [--bench passes]   (--bench: no digests; every served run is touched (one byte per 4 KB page) and
//                  the line carries wall-clock seconds per pass and decoded bytes: the in-process multi-GPU timing)
#include <chrono>
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
  int device = 0, bench = 0, lanes = 1;  // --lanes n: every device listed n times = n decodes in flight on it
  std::vector<int> gpus;
  for (int i = 2; i + 1 < argc; i += 2) {
    const std::string k = argv[i];
    const char *v = argv[i + 1];
    if (k == "--cols") { for (const char *s = v; *s;) { cols.push_back((uint32_t)strtoul(s, (char **)&s, 10)); if (*s == ',') ++s; } }
    else if (k == "--chunk") chunk = strtoull(v, nullptr, 10);
    else if (k == "--batch") batch = (uint32_t)strtoul(v, nullptr, 10);
    else if (k == "--device") device = atoi(v);
    else if (k == "--bench") bench = atoi(v);
    else if (k == "--lanes") lanes = atoi(v);
    else if (k == "--gpus") { for (const char *s = v; *s;) { gpus.push_back((int)strtol(s, (char **)&s, 10)); if (*s == ',') ++s; } }
    else if (k == "--range") { char *e; g0 = (uint32_t)strtoul(v, &e, 10); g1 = (uint32_t)strtoul(e + 1, nullptr, 10); have_range = true; }
  }
  if (gpus.empty()) gpus.push_back(device);
  if (lanes > 1) {
    std::vector<int> rep;
    for (int l = 0; l < lanes; ++l) rep.insert(rep.end(), gpus.begin(), gpus.end());
    gpus.swap(rep);
  }
  if (bench > 0) {
    double best = 1e30, sum = 0;
    uint64_t bytes = 0, total = 0;
    for (int pass = 0; pass <= bench; ++pass) {  // pass 0 warms up (pinning the image, buffer pools)
      flsgpu_host::FlsScanner sc;
      if (!sc.open(argv[1], gpus, batch)) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
      if (!cols.empty() && !sc.setProjection(cols)) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
      const auto t0 = std::chrono::steady_clock::now();
      std::vector<flsgpu_host::ColumnSlice> out;
      uint64_t rows = 0, touched = 0;
      bytes = 0; total = 0;
      while (sc.next(chunk, rows, out)) {
        for (const auto &s : out) {
          if (s.width) { const uint8_t *p = (const uint8_t *)s.data; const size_t n = rows * s.width; for (size_t i = 0; i < n; i += 4096) touched += p[i]; bytes += n; }
          else { const size_t n = s.offsets[rows] - s.offsets[0]; const uint8_t *p = s.payload + s.offsets[0]; for (size_t i = 0; i < n; i += 4096) touched += p[i]; bytes += n + 4 * rows; }
        }
        total += rows;
      }
      if (!sc.error().empty()) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
      const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      if (pass) { best = sec < best ? sec : best; sum += sec; }
      if (touched == 0x7fffffffffffffffull) printf(" ");
    }
    printf("{\"rows\": %llu, \"decoded_bytes\": %llu, \"gpus\": %zu, \"passes\": %d, \"sec_best\": %.6f, \"sec_mean\": %.6f}\n",
           (unsigned long long)total, (unsigned long long)bytes, gpus.size(), bench, best, sum / bench);
    return 0;
  }
  flsgpu_host::FlsScanner sc;
  if (!sc.open(argv[1], gpus, batch)) { printf("{\"error\": \"%s\"}\n", sc.error().c_str()); return 1; }
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
