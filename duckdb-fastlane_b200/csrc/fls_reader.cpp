// fls_reader.cpp — FLS footer parser.  See fls_reader.h; layout in fls_format.h / FORMAT.md.
#include "fls_reader.h"

#include <cstdio>
#include <cstring>

namespace {
struct Cur {
  const uint8_t *p, *end;
  bool ok = true;
  template <typename T> T get() {
    T v{};
    if (p + sizeof(T) > end) { ok = false; return v; }
    memcpy(&v, p, sizeof(T));
    p += sizeof(T);
    return v;
  }
};

std::string parse_footer(const uint8_t *foot, uint64_t fsz, uint64_t foot_off, FlsTable &t) {
  Cur c{foot, foot + fsz};
  if (c.get<uint32_t>() != FLS_FOOTER_VERSION) return "unsupported FLS footer version";
  t.n_cols = c.get<uint32_t>();
  t.n_rg = c.get<uint32_t>();
  t.rg_vecs = c.get<uint32_t>();
  t.n_rows = c.get<uint64_t>();
  if (!c.ok || t.n_cols == 0 || t.n_cols > 65535 || t.rg_vecs != FLS_RG_VECS) return "corrupt FLS footer header";
  for (uint32_t i = 0; i < t.n_cols; ++i) {
    int ty = c.get<uint8_t>();
    c.get<uint8_t>();
    uint16_t nl = c.get<uint16_t>();
    if (!c.ok || c.p + nl > c.end || ty <= FLS_T_INVALID || ty >= FLS_T_N) return "corrupt FLS column descriptor";
    t.types.push_back(ty);
    t.names.emplace_back((const char *)c.p, nl);
    c.p += nl;
  }
  c.p = foot + ((uint64_t)(c.p - foot) + 7) / 8 * 8;
  t.rgs.resize(t.n_rg);
  t.chunks.resize((size_t)t.n_rg * t.n_cols);
  uint64_t rows = 0;
  for (uint32_t g = 0; g < t.n_rg; ++g) {
    fls_rowgroup &r = t.rgs[g];
    r = c.get<fls_rowgroup>();
    if (!c.ok || r.off % FLS_RG_ALIGN || r.off + r.size > foot_off || r.n_vec == 0 || r.n_vec > FLS_RG_VECS ||
        r.n_tuples > r.n_vec * FLS_VEC || r.n_tuples + FLS_VEC <= r.n_vec * FLS_VEC)
      return "corrupt FLS row-group descriptor";
    // only the last row group may be partial (decoded rows stay contiguous across row groups)
    if (g + 1 < t.n_rg && r.n_tuples != r.n_vec * FLS_VEC) return "unsupported: partial vector before the last row group";
    if (g && r.off < t.rgs[g - 1].off + t.rgs[g - 1].size) return "corrupt FLS row-group order";
    rows += r.n_tuples;
    for (uint32_t col = 0; col < t.n_cols; ++col) {
      fls_colchunk &k = t.chunks[(size_t)g * t.n_cols + col];
      memset(&k, 0, sizeof(k));
      k.scheme = c.get<uint8_t>();
      k.n_seg = c.get<uint8_t>();
      k.flags = c.get<uint16_t>();
      k.aux32 = c.get<uint32_t>();
      k.aux64[0] = c.get<uint64_t>();
      k.aux64[1] = c.get<uint64_t>();
      if (!c.ok || k.scheme >= FLS_S_N || k.n_seg > FLS_MAX_SEGS) return "corrupt FLS column-chunk descriptor";
      for (uint32_t s = 0; s < k.n_seg; ++s) {
        k.seg[s].off = c.get<uint64_t>();
        k.seg[s].size = c.get<uint64_t>();
        if (!c.ok || k.seg[s].off % FLS_SEG_ALIGN || k.seg[s].off + k.seg[s].size > r.size)
          return "corrupt FLS segment descriptor";
      }
    }
  }
  if (rows != t.n_rows) return "FLS footer row count mismatch";
  t.data_begin = t.n_rg ? t.rgs[0].off : 16;
  t.data_end = t.n_rg ? t.rgs[t.n_rg - 1].off + t.rgs[t.n_rg - 1].size : 16;
  return "";
}
}  // namespace

std::string fls_parse(const uint8_t *img, uint64_t size, FlsTable &t) {
  uint64_t magic0 = 0, tail[3] = {0, 0, 0};
  if (size < 40) return "file too small to be an FLS file";
  memcpy(&magic0, img, 8);
  memcpy(tail, img + size - 24, 24);
  if (magic0 != FLS_MAGIC || tail[2] != FLS_MAGIC) return "bad FLS magic";
  if (tail[0] + tail[1] + 24 != size || tail[0] < 16) return "corrupt FLS trailer";
  return parse_footer(img + tail[0], tail[1], tail[0], t);
}

std::string fls_parse_file(const char *path, FlsTable &t, uint64_t *file_size) {
  FILE *f = fopen(path, "rb");
  if (!f) return std::string("cannot open ") + path;
  std::string err;
  uint64_t head = 0, tail[3];
  fseek(f, 0, SEEK_END);
  uint64_t size = (uint64_t)ftell(f);
  if (file_size) *file_size = size;
  if (size < 40) { fclose(f); return "file too small to be an FLS file"; }
  fseek(f, 0, SEEK_SET);
  if (fread(&head, 8, 1, f) != 1) err = "short read";
  fseek(f, (long)(size - 24), SEEK_SET);
  if (err.empty() && fread(tail, 24, 1, f) != 1) err = "short read";
  if (err.empty() && (head != FLS_MAGIC || tail[2] != FLS_MAGIC)) err = "bad FLS magic";
  if (err.empty() && (tail[0] + tail[1] + 24 != size || tail[0] < 16)) err = "corrupt FLS trailer";
  if (err.empty()) {
    std::vector<uint8_t> foot(tail[1]);
    fseek(f, (long)tail[0], SEEK_SET);
    if (tail[1] && fread(foot.data(), 1, tail[1], f) != tail[1]) err = "short read";
    else err = parse_footer(foot.data(), tail[1], tail[0], t);
  }
  fclose(f);
  return err;
}
