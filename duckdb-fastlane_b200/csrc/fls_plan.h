// fls_plan.h — turns (table descriptor, row-group range, column, staged base address, output buffers) into
// launch groups of chunk descriptors.  Host-only, header-only; shared by flsgpu.cu (device addresses) and the
// host emulation of the kernels in tests/emul (host addresses).
//
// This is the per-column "expression interpreter" of the missing FastLanes reader reduced to a table:
// scheme -> kernel kind.
#pragma once
#include <algorithm>
#include <string>
#include <vector>

#include "fls_kernels.cuh"
#include "fls_reader.h"

enum FlsKernelKind {
  FLS_K_FFOR = 0, FLS_K_DELTA, FLS_K_RLE, FLS_K_DICT, FLS_K_ALP, FLS_K_ALPRD, FLS_K_COPY, FLS_K_FILL,
  FLS_K_STR_OFFSETS,    // DELTA u32 end-offset stream -> offsets[i+1]
  FLS_K_STR_COPY,       // payload blob copy (unaligned)
  FLS_K_STR_CONST,
  FLS_K_STR_DICT,
  FLS_K_STR_FSST,
  FLS_K_STR_FSST_DICTPREP,  // decode an FSST-compressed dictionary into scratch
  FLS_K_VALIDITY,           // validity bitmap of a chunk (copied, or all ones) -> the column's bitmap
  FLS_K_FFOR_PATCH, FLS_K_CROSS_RLE, FLS_K_STR_FSST12,
  FLS_K_STR_FLAGS,           // STR_DICT chunk with up to 8 one-byte entries, every vector full (flag columns)
  FLS_K_STR_DICT_LONG,       // STR_DICT chunk with an entry longer than FLS_DICT_SHORT_MAXLEN bytes (or unknown lengths)
  FLS_K_N
};

struct FlsLaunch {
  int kind;
  int width;  // bytes per value (fixed width) or 0
  std::vector<FlsChunkDev> chunks;
  // algorithmic traffic of the launch: segment bytes read, decoded bytes written
  uint64_t bytes_read = 0, bytes_written = 0, vectors = 0;
  // persistent chunk kernels hand each block a contiguous range of `chunks`: sort_by_key() interleaves the
  // columns (row-group-major order) so that every range holds the same mix of cheap and expensive columns
  std::vector<uint32_t> keys;
  void add(const FlsChunkDev &d, uint64_t rd, uint64_t wr, uint32_t key = 0) {
    chunks.push_back(d);
    keys.push_back(key);
    bytes_read += rd; bytes_written += wr; vectors += d.n_vec;
  }
};

inline void fls_sort_by_key(FlsLaunch &l) {
  std::vector<uint32_t> idx(l.chunks.size());
  for (uint32_t i = 0; i < idx.size(); ++i) idx[i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return l.keys[a] < l.keys[b]; });
  std::vector<FlsChunkDev> c(l.chunks.size());
  for (uint32_t i = 0; i < idx.size(); ++i) c[i] = l.chunks[idx[i]];
  l.chunks.swap(c);
  std::sort(l.keys.begin(), l.keys.end());
}

struct FlsColOut {
  void *data = nullptr;        // fixed width: padded_rows * width bytes
  uint32_t *offsets = nullptr; // strings: padded_rows + 1
  uint8_t *payload = nullptr;  // strings: sum of row-group payload bytes (+16 slack)
  uint8_t *scratch = nullptr;  // strings (FSST_DICT): decoded dictionaries
  uint8_t *validity = nullptr; // 128 bytes per vector of the range, or null when no chunk of the range has NULLs
  uint64_t pay_base0 = 0;      // strings: payload bytes that precede row group g0 in `payload` (pipeline slices)
};

struct FlsColNeed {  // sizes the caller must provide for a column over a row-group range
  uint64_t data_bytes = 0, offsets_bytes = 0, payload_bytes = 0, scratch_bytes = 0;
  uint64_t validity_bytes = 0;    // 128 per vector when some chunk of the range carries a validity bitmap, else 0
  uint64_t compressed_bytes = 0;  // segment bytes read
  uint64_t decoded_bytes = 0;     // algorithmic bytes written (values, or offsets + payload)
};

// One launch per (kernel kind, value width): chunks of different columns share it.
static inline FlsLaunch &fls_launch_for(std::vector<FlsLaunch> &ls, int kind, int width) {
  for (auto &l : ls)
    if (l.kind == kind && l.width == width) return l;
  ls.emplace_back();
  ls.back().kind = kind;
  ls.back().width = width;
  return ls.back();
}

static inline FlsColNeed fls_column_need(const FlsTable &t, uint32_t col, uint32_t g0, uint32_t g1,
                                         const uint8_t *host_image /* may be null */) {
  FlsColNeed n;
  const uint32_t wb = fls_type_bits(t.types[col]) / 8;
  const uint64_t rows = t.padded_rows(g0, g1);
  for (uint32_t g = g0; g < g1; ++g) {
    const fls_colchunk &k = t.chunk(g, col);
    for (uint32_t s = 0; s < k.n_seg; ++s) n.compressed_bytes += k.seg[s].size;
    if (k.flags & FLS_CHUNK_VALIDITY) n.validity_bytes = rows / 8;
    if (!wb) {
      n.payload_bytes += k.aux64[0];
      if (k.scheme == FLS_S_STR_FSST_DICT && host_image) {
        // decoded dictionary size = last entry of dict_dec_offsets (seg[6])
        const uint8_t *p = host_image + t.rgs[g].off + k.seg[6].off + 4ull * k.aux32;
        uint32_t sz;
        memcpy(&sz, p, 4);
        n.scratch_bytes += ((uint64_t)sz + 31) & ~15ull;
      }
    }
  }
  if (wb) { n.data_bytes = rows * wb; n.decoded_bytes = n.data_bytes; }
  else {
    n.offsets_bytes = (rows + 1) * 4;
    n.decoded_bytes = n.offsets_bytes + n.payload_bytes;
    n.payload_bytes += 16;
  }
  return n;
}

// Dictionary offsets u32[n + 1] (in the host image): is every entry at most FLS_DICT_SHORT_MAXLEN bytes long, so that any
// 256 of them fit the staging buffer of the 20-warp configuration of the DICT kernel?
static inline bool fls_dict_is_short(const uint8_t *offsets, uint64_t seg_size, uint32_t n) {
  if (seg_size < 4ull * ((uint64_t)n + 1u)) return false;
  uint32_t prev;
  memcpy(&prev, offsets, 4);
  for (uint32_t e = 0; e < n; ++e) {
    uint32_t next;
    memcpy(&next, offsets + 4ull * (e + 1u), 4);
    if (next < prev || next - prev > FLS_DICT_SHORT_MAXLEN) return false;
    prev = next;
  }
  return true;
}

// STR_DICT chunk `k` of a row group that starts at `rgp` in the host image: are its entries 1 .. 8 strings of one byte,
// and does its payload hold exactly one byte per padded row?  *flags8 = the entries' bytes, entry e in byte e.
static inline bool fls_dict_is_flags(const uint8_t *rgp, const fls_colchunk &k, uint32_t n_vec, uint64_t *flags8) {
  const uint32_t n = k.aux32;
  if (k.n_seg < 5 || n < 1u || n > 8u || k.aux64[0] != (uint64_t)n_vec * FLS_VEC) return false;
  if (k.seg[2].size < 4ull * (n + 1u)) return false;
  uint32_t off[9];
  memcpy(off, rgp + k.seg[2].off, 4u * (n + 1u));
  uint64_t v = 0;
  for (uint32_t e = 0; e < n; ++e) {
    if (off[e + 1u] != off[e] + 1u || (uint64_t)off[e] + 1u > k.seg[3].size) return false;
    v |= (uint64_t)rgp[k.seg[3].off + off[e]] << (8u * e);
  }
  *flags8 = v;
  return true;
}

// stage_base = address (device or host) of the byte at file offset t.rgs[g0].off
static inline std::string fls_plan_column(const FlsTable &t, uint32_t col, uint32_t g0, uint32_t g1,
                                          const uint8_t *stage_base, const uint8_t *host_image,
                                          const FlsColOut &out, std::vector<FlsLaunch> &ls) {
  const uint32_t wb = fls_type_bits(t.types[col]) / 8;
  uint64_t row_base = 0, pay_base = out.pay_base0, scratch_off = 0;
  for (uint32_t g = g0; g < g1; ++g) {
    const fls_rowgroup &rg = t.rgs[g];
    const fls_colchunk &k0 = t.chunk(g, col);
    // EQUAL: the chunk is column aux32's chunk of this row group, decoded into this column's buffers
    const bool eq = k0.scheme == FLS_S_EQUAL;
    if (eq && (k0.aux32 >= t.n_cols || t.chunk(g, k0.aux32).scheme == FLS_S_EQUAL ||
               fls_type_bits(t.types[k0.aux32]) != fls_type_bits(t.types[col]) || !wb))
      return "EQUAL chunk does not name a fixed-width column of the same width";
    const fls_colchunk &k = eq ? t.chunk(g, k0.aux32) : k0;
    FlsChunkDev d;
    memset(&d, 0, sizeof(d));
    const uint8_t *rgp = stage_base + (rg.off - t.rgs[g0].off);
    for (uint32_t s = 0; s < k.n_seg; ++s) d.seg[s] = rgp + k.seg[s].off;
    d.n_vec = rg.n_vec;
    d.aux32 = k.aux32;
    d.aux64[0] = k.aux64[0];
    d.aux64[1] = k.aux64[1];
    d.scheme = k.scheme;
    uint64_t rd = 0;
    for (uint32_t s = 0; s < k.n_seg; ++s) rd += k.seg[s].size;
    const uint64_t vals = (uint64_t)rg.n_vec * FLS_VEC;
    if (out.validity) {  // seg[0] = the chunk's bitmap, or null: all rows hold values
      FlsChunkDev vd;
      memset(&vd, 0, sizeof(vd));
      vd.n_vec = rg.n_vec;
      const bool hv = (k.flags & FLS_CHUNK_VALIDITY) && k.n_seg;
      vd.seg[0] = hv ? d.seg[k.n_seg - 1u] : nullptr;
      vd.out = out.validity + row_base / 8;
      fls_launch_for(ls, FLS_K_VALIDITY, 0).add(vd, hv ? 128ull * rg.n_vec : 0, 128ull * rg.n_vec);
    }
    if (wb) {
      const uint64_t wr = vals * wb;
      d.out = (uint8_t *)out.data + row_base * wb;
      switch (k.scheme) {
        case FLS_S_FFOR: fls_launch_for(ls, FLS_K_FFOR, wb).add(d, rd, wr); break;
        case FLS_S_DELTA: fls_launch_for(ls, FLS_K_DELTA, wb).add(d, rd, wr); break;
        case FLS_S_RLE: fls_launch_for(ls, FLS_K_RLE, wb).add(d, rd, wr); break;
        case FLS_S_FFOR_PATCH: fls_launch_for(ls, FLS_K_FFOR_PATCH, wb).add(d, rd, wr); break;
        case FLS_S_CROSS_RLE: fls_launch_for(ls, FLS_K_CROSS_RLE, wb).add(d, rd, wr); break;
        case FLS_S_DICT: fls_launch_for(ls, FLS_K_DICT, wb).add(d, rd, wr); break;
        case FLS_S_ALP:
          if (t.types[eq ? k0.aux32 : col] != FLS_T_DOUBLE && t.types[eq ? k0.aux32 : col] != FLS_T_FLOAT) return "ALP on a column that is neither DOUBLE nor FLOAT";
          fls_launch_for(ls, FLS_K_ALP, wb).add(d, rd, wr); break;
        case FLS_S_ALP_RD:
          if (!(t.types[eq ? k0.aux32 : col] == FLS_T_DOUBLE)) return "ALP_RD on a non-double column";
          fls_launch_for(ls, FLS_K_ALPRD, wb).add(d, rd, wr); break;
        case FLS_S_UNCOMPRESSED:
          d.aux64[1] = (uint64_t)rg.n_vec * FLS_VEC * wb;
          fls_launch_for(ls, FLS_K_COPY, wb).add(d, rd, wr); break;
        case FLS_S_CONSTANT:
          d.aux32 = wb;
          fls_launch_for(ls, FLS_K_FILL, wb).add(d, rd, wr); break;
        default: return "scheme not valid for a fixed-width column";
      }
    } else {
      d.out = out.offsets + row_base;
      d.pay = out.payload;
      d.pay_base = (uint32_t)pay_base;
      const uint64_t wr = vals * 4 + k.aux64[0];  // offsets + payload
      switch (k.scheme) {
        case FLS_S_STR_UNCOMPRESSED: {
          fls_launch_for(ls, FLS_K_STR_OFFSETS, 0).add(d, k.seg[0].size + k.seg[1].size, vals * 4);
          FlsChunkDev c2 = d;
          c2.seg[0] = d.seg[2];
          fls_launch_for(ls, FLS_K_STR_COPY, 0).add(c2, k.seg[2].size, k.aux64[0]);
          break;
        }
        case FLS_S_STR_CONSTANT: fls_launch_for(ls, FLS_K_STR_CONST, 0).add(d, rd, wr); break;
        case FLS_S_STR_DICT: {
          // Flag columns (l_returnflag, l_linestatus): a dictionary of up to 8 one-byte strings and no short vector.
          // Nothing about them needs the table / staging machinery of the persistent DICT kernel — the offsets are
          // arithmetic and 8 strings are two byte permutes — so they go to a plain warp-per-vector kernel that
          // runs with the bandwidth-bound launches; the dictionary's bytes travel in the descriptor.
          uint64_t flags8 = 0;
          if (host_image && fls_dict_is_flags(host_image + rg.off, k, rg.n_vec, &flags8)) {
            d.aux64[1] = flags8;
            fls_launch_for(ls, FLS_K_STR_FLAGS, 0).add(d, rd, wr);
          } else {
            const bool sh = host_image && fls_dict_is_short(host_image + rg.off + k.seg[2].off, k.seg[2].size, k.aux32);
            fls_launch_for(ls, sh ? FLS_K_STR_DICT : FLS_K_STR_DICT_LONG, 0).add(d, rd, wr, g);
          }
          break;
        }
        case FLS_S_STR_FSST: fls_launch_for(ls, FLS_K_STR_FSST, 0).add(d, rd, wr, g); break;
        case FLS_S_STR_FSST12: fls_launch_for(ls, FLS_K_STR_FSST12, 0).add(d, rd, wr); break;
        case FLS_S_STR_FSST_DICT: {
          if (!host_image) return "FSST_DICT planning needs the host image";
          uint32_t sz;
          memcpy(&sz, host_image + rg.off + k.seg[6].off + 4ull * k.aux32, 4);
          FlsChunkDev p = d;
          p.out = out.scratch + scratch_off;  // decoded dictionary bytes
          fls_launch_for(ls, FLS_K_STR_FSST_DICTPREP, 0).add(p, k.seg[2].size + k.seg[3].size + k.seg[5].size, sz);
          // gather stage reads the decoded dictionary: offsets = dict_dec_offsets (seg[6]), bytes = scratch
          d.seg[2] = d.seg[6];
          d.seg[3] = out.scratch + scratch_off;
          scratch_off += ((uint64_t)sz + 31) & ~15ull;
          const bool sh = fls_dict_is_short(host_image + rg.off + k.seg[6].off, k.seg[6].size, k.aux32);
          fls_launch_for(ls, sh ? FLS_K_STR_DICT : FLS_K_STR_DICT_LONG, 0)
              .add(d, k.seg[0].size + k.seg[1].size + k.seg[4].size + k.seg[6].size, wr, g);
          break;
        }
        default: return "scheme not valid for a string column";
      }
      pay_base += k.aux64[0];
    }
    row_base += (uint64_t)rg.n_vec * FLS_VEC;
  }
  return "";
}
