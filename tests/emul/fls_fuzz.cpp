// fls_fuzz.cpp — decode-path fuzzer over the HOST EMULATION of the CUDA decoders.  TEST INFRASTRUCTURE ONLY.
//
// Built with -fsanitize=address (tests/test_footer_fuzz.py): the exact device source (fls_decode.cuh /
// fls_strings.cuh under FLS_HOST_EMUL) and the exact host gate (fls_validate.h) run on damaged images whose every
// buffer — image, outputs, scratch, "shared memory" — is a heap block of exactly the size the library would hand
// a kernel.  A damaged file must end in a validation error or in a decode that stays inside those blocks; an
// out-of-bounds read or write aborts the process (ASAN), which fails the test.
//
//   fls_fuzz <image> <mode> <seed> <iterations>
//     mode  hdr     random bytes inside the per-vector header tables (entry points, widths, counts, bases)
//           footer  random bytes inside the footer (schemes, aux fields, segment tables)
//           data    random bytes anywhere in the row-group area (string offsets, indices, exception positions)
//           wide    8 aligned footer bytes replaced by a large value
//           segs    random bytes inside one randomly chosen segment (uniform over segments, so the small ones —
//                   dictionary offsets, entry points, symbol tables, lane bases — are hit as often as the payloads)
#include <cstdio>
#include <cstdlib>
#include <random>

#include "fls_emul.cpp"

static std::vector<uint8_t> read_file(const char *p) {
  std::vector<uint8_t> v;
  FILE *f = fopen(p, "rb");
  if (!f) return v;
  fseek(f, 0, SEEK_END);
  v.resize((size_t)ftell(f));
  fseek(f, 0, SEEK_SET);
  if (fread(v.data(), 1, v.size(), f) != v.size()) v.clear();
  fclose(f);
  return v;
}

int main(int argc, char **argv) {
  if (argc < 5) return 2;
  const std::vector<uint8_t> good = read_file(argv[1]);
  const std::string mode = argv[2];
  std::mt19937_64 rng((uint64_t)atoll(argv[3]));
  const int iters = atoi(argv[4]);
  FlsTable t0;
  if (good.empty() || !fls_parse(good.data(), good.size(), t0).empty()) return 3;
  uint64_t tail[3];
  memcpy(tail, good.data() + good.size() - 24, 24);
  const uint64_t foot_off = tail[0], foot_end = good.size() - 24;
  // header tables of the pristine file: (file offset, size)
  std::vector<std::pair<uint64_t, uint64_t>> hdrs, segs;
  for (uint32_t g = 0; g < t0.n_rg; ++g)
    for (uint32_t c = 0; c < t0.n_cols; ++c) {
      const fls_colchunk &k = t0.chunk(g, c);
      const uint32_t n_streams = k.scheme == FLS_S_STR_FSST ? 2u : (k.scheme == FLS_S_UNCOMPRESSED || k.scheme == FLS_S_CONSTANT || k.scheme == FLS_S_STR_CONSTANT) ? 0u : 1u;
      for (uint32_t s = 0; s < n_streams; ++s)
        if (k.seg[2 * s].size) hdrs.push_back({t0.rgs[g].off + k.seg[2 * s].off, k.seg[2 * s].size});
      for (uint32_t s = 0; s < k.n_seg; ++s)
        if (k.seg[s].size) segs.push_back({t0.rgs[g].off + k.seg[s].off, k.seg[s].size});
    }
  auto rnd = [&](uint64_t lo, uint64_t hi) { return lo + rng() % (hi - lo); };  // [lo, hi)
  long ok = 0, refused = 0, unparsed = 0;
  for (int it = 0; it < iters; ++it) {
    // the image is an exact-size heap block: a read one byte past it is an ASAN report
    uint8_t *img = (uint8_t *)malloc(good.size());
    memcpy(img, good.data(), good.size());
    if (mode == "hdr" && !hdrs.empty()) {
      for (int k = 0, n = 1 + (int)rnd(0, 4); k < n; ++k) {
        const auto &hs = hdrs[rnd(0, hdrs.size())];
        const uint64_t at = hs.first + rnd(0, hs.second);
        // half of the hits go for the fields that form addresses: ep (bytes 0-3) and meta (bytes 4-7)
        const uint64_t a2 = (rng() & 1) ? (at & ~15ull) + rnd(0, 8) : at;
        img[a2 < hs.first + hs.second ? a2 : at] = (uint8_t)rng();
      }
    } else if (mode == "segs" && !segs.empty()) {
      const auto &sg = segs[rnd(0, segs.size())];
      for (int k = 0, n = 1 + (int)rnd(0, 8); k < n; ++k) {
        // small segments anywhere; large ones near their start (lane bases, first offsets) half of the time
        const uint64_t span = (rng() & 1) ? (sg.second < 256 ? sg.second : 256) : sg.second;
        img[sg.first + rnd(0, span)] = (rng() & 3) ? (uint8_t)rng() : 0xFF;
      }
    } else if (mode == "footer") {
      for (int k = 0, n = 1 + (int)rnd(0, 3); k < n; ++k) img[rnd(foot_off, foot_end)] = (uint8_t)rng();
    } else if (mode == "wide") {
      const uint64_t at = foot_off + 8 * rnd(0, (foot_end - foot_off) / 8);
      const uint64_t v = rng() >> (rng() % 40);
      memcpy(img + at, &v, 8);
    } else {
      for (int k = 0; k < 16; ++k) img[rnd(16, foot_off)] = (uint8_t)rng();
    }
    FlsTable t;
    if (!fls_parse(img, good.size(), t).empty()) { ++unparsed; free(img); continue; }
    bool any_refused = false;
    for (uint32_t c = 0; c < t.n_cols; ++c) {
      uint64_t need[4];
      const uint64_t rc = emul_need(img, good.size(), c, 0, t.n_rg, need);
      if (rc) { any_refused = true; continue; }
      if (need[0] > (1ull << 31) || need[1] > (1ull << 31) || need[2] > (1ull << 31)) { any_refused = true; continue; }  // the product sizes device buffers by these too; here they must stay mallocable
      void *data = need[0] ? malloc(need[0]) : nullptr;
      uint32_t *offs = need[1] ? (uint32_t *)malloc(need[1]) : nullptr;
      uint8_t *pay = need[2] ? (uint8_t *)malloc(need[2]) : nullptr;
      if (emul_decode(img, good.size(), c, 0, t.n_rg, data, offs, pay)) any_refused = true;
      free(data); free(offs); free(pay);
    }
    (any_refused ? refused : ok)++;
    free(img);
  }
  printf("done %ld %ld %ld\n", ok, refused, unparsed);
  return 0;
}
