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

// well-formed UTF-8 (no overlong forms, no surrogates, <= U+10FFFF): column names end up in the engine's catalog
bool valid_utf8(const uint8_t *s, size_t n) {
  for (size_t i = 0; i < n;) {
    const uint8_t c = s[i];
    size_t k;
    uint32_t cp;
    if (c < 0x80) { ++i; continue; }
    else if ((c & 0xE0) == 0xC0) { k = 1; cp = c & 0x1Fu; }
    else if ((c & 0xF0) == 0xE0) { k = 2; cp = c & 0x0Fu; }
    else if ((c & 0xF8) == 0xF0) { k = 3; cp = c & 0x07u; }
    else return false;
    if (i + k >= n) return false;  // the sequence runs past the end
    for (size_t j = 1; j <= k; ++j) {
      if ((s[i + j] & 0xC0) != 0x80) return false;
      cp = (cp << 6) | (s[i + j] & 0x3Fu);
    }
    if ((k == 1 && cp < 0x80) || (k == 2 && cp < 0x800) || (k == 3 && cp < 0x10000) || cp > 0x10FFFF ||
        (cp >= 0xD800 && cp <= 0xDFFF))
      return false;
    i += k + 1;
  }
  return true;
}

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
    if (!valid_utf8(c.p, nl) || memchr(c.p, 0, nl)) return "corrupt FLS column name";
    t.types.push_back(ty);
    t.names.emplace_back((const char *)c.p, nl);
    c.p += nl;
  }
  c.p = foot + ((uint64_t)(c.p - foot) + 7) / 8 * 8;
  // every row group costs its descriptor and, per column, a chunk descriptor of at least 24 bytes: a count the
  // footer cannot hold is refused before anything is sized by it
  if (c.p > c.end || (uint64_t)t.n_rg * (sizeof(fls_rowgroup) + 24ull * t.n_cols) > (uint64_t)(c.end - c.p))
    return "corrupt FLS footer header";
  t.rgs.resize(t.n_rg);
  t.chunks.resize((size_t)t.n_rg * t.n_cols);
  uint64_t rows = 0;
  for (uint32_t g = 0; g < t.n_rg; ++g) {
    fls_rowgroup &r = t.rgs[g];
    r = c.get<fls_rowgroup>();
    if (!c.ok || r.off % FLS_RG_ALIGN || r.size > foot_off || r.off > foot_off - r.size || r.n_vec == 0 || r.n_vec > FLS_RG_VECS ||
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
        if (!c.ok || k.seg[s].off % FLS_SEG_ALIGN || k.seg[s].size > r.size || k.seg[s].off > r.size - k.seg[s].size)
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
  if (tail[0] > size || tail[1] > size || tail[0] + tail[1] + 24 != size || tail[0] < 16) return "corrupt FLS trailer";
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
  if (err.empty() && (tail[0] > size || tail[1] > size || tail[0] + tail[1] + 24 != size || tail[0] < 16)) err = "corrupt FLS trailer";
  if (err.empty()) {
    std::vector<uint8_t> foot(tail[1]);
    fseek(f, (long)tail[0], SEEK_SET);
    if (tail[1] && fread(foot.data(), 1, tail[1], f) != tail[1]) err = "short read";
    else err = parse_footer(foot.data(), tail[1], tail[0], t);
  }
  fclose(f);
  return err;
}
