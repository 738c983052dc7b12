// fls_strings.cuh — string column decoders (Arrow layout out: u32 offsets[n+1] + payload bytes).
//
// Thread t of the warp that owns a vector is responsible for the 32 strings e = 32t .. 32t+31 (DICT/CONST)
// or for lane t's 32 contiguous originals (FSST / offset streams, unified transposed layout) — in both cases
// a CONTIGUOUS run of output bytes, written through FlsByteSink as aligned 8-byte stores.
//
// Restated FastLanes primitives (upstream source absent); FSST stream semantics
// follow duckdb/third_party/fsst/fsst.h:93-104,162-241 (code<255 -> symbol, 255 -> escaped literal byte).
#pragma once
#include "fls_kernels.cuh"

// ---- contiguous byte writer: head/tail bytes individually, everything between as aligned u64 stores
struct FlsByteSink {
  uint64_t *word;  // aligned address of the word being assembled
  uint64_t acc;
  uint32_t n;      // bytes valid in acc (including the `head` foreign bytes of the first word)
  uint32_t head;   // low bytes of the first word that belong to someone else
  FLS_MEM void init(uint8_t *p) {
    const uint32_t a = (uint32_t)((uintptr_t)p & 7u);
    word = (uint64_t *)(p - a);
    acc = 0; n = a; head = a;
  }
  FLS_MEM void flush_partial(uint32_t lo, uint32_t hi) {
    uint8_t *b = (uint8_t *)word;
    for (uint32_t i = lo; i < hi; ++i) b[i] = (uint8_t)(acc >> (8u * i));
  }
  // append the low L bytes of s (bytes above L must be zero), 1 <= L <= 8
  FLS_MEM void push(uint64_t s, uint32_t L) {
    acc |= s << (8u * n);
    n += L;
    if (n >= 8u) {
      if (head) { flush_partial(head, 8u); head = 0; }
      else *word = acc;
      ++word;
      n -= 8u;
      acc = n ? (s >> (8u * (L - n))) : 0ull;
    }
  }
  FLS_MEM void finish() {
    if (n > head) flush_partial(head, n);
  }
};

// ---- contiguous byte writer, lean variant for the decode loops: 64-bit accumulator, one aligned 8-byte store per
// completed word, no byte loops on the hot path.  The first (possibly shared) word is parked in a register and
// written byte-wise in finish(), like the tail, so neighbouring lanes never touch each other's bytes.
struct FlsWordSink {
  uint64_t *word;   // aligned address of the word being assembled
  uint8_t *base;    // aligned address of the first word
  uint64_t acc, headw;
  uint32_t n, head;
  bool first;
  FLS_MEM void init(uint8_t *p) {
    const uint32_t a = (uint32_t)((uintptr_t)p & 7u);
    base = p - a;
    word = (uint64_t *)base;
    acc = 0; headw = 0; n = a; head = a; first = true;
  }
  // append the low L bytes of s (bytes above L must be zero), 0 <= L <= 8
  FLS_MEM void push(uint64_t s, uint32_t L) {
    const uint32_t sh = 8u * n;
    acc |= s << sh;
    const uint64_t carry = (s >> 1) >> (63u - sh);  // the bytes that did not fit (none when n == 0)
    n += L;
    const bool full = n >= 8u;
    if (full && !first) *word = acc;   // predicated store
    headw = (full && first) ? acc : headw;
    first = first && !full;
    word += full ? 1 : 0;
    acc = full ? carry : acc;
    n &= 7u;
  }
  FLS_MEM void finish() {
    if (!first) {
      for (uint32_t i = head; i < 8u; ++i) base[i] = (uint8_t)(headw >> (8u * i));
      uint8_t *b = (uint8_t *)word;
      for (uint32_t i = 0; i < n; ++i) b[i] = (uint8_t)(acc >> (8u * i));
    } else {
      for (uint32_t i = head; i < n; ++i) base[i] = (uint8_t)(acc >> (8u * i));
    }
  }
};

// ---- sequential byte reader over global memory with three 8-byte words of lookahead.
// The next words are requested long before their bytes are consumed, so the serial decode loop of a lane never
// waits for DRAM.  Loads are clamped to [.., limit): nothing beyond the lane's own run (rounded up to 8) is read.
struct FlsByteStream {
  const uint64_t *next;   // address of the word that follows n2
  const uint64_t *limit;  // first word that must not be read
  uint64_t cur, n0, n1, n2;
  uint32_t bi;            // bytes of `cur` already consumed (0..7)
  FLS_MEM uint64_t fetch() {
    const uint64_t v = next < limit ? fls_ldg(next) : 0ull;
    ++next;
    return v;
  }
  FLS_MEM void init(const uint8_t *p, const uint8_t *end) {
    const uint32_t a = (uint32_t)((uintptr_t)p & 7u);
    next = (const uint64_t *)(p - a);
    limit = (const uint64_t *)(((uintptr_t)end + 7u) & ~(uintptr_t)7u);
    cur = fetch(); n0 = fetch(); n1 = fetch(); n2 = fetch();
    bi = a;
  }
  // the next two bytes (low 16 bits), without consuming them
  FLS_MEM uint32_t peek2() const {
    const uint32_t sh = 8u * bi;
    return (uint32_t)((cur >> sh) | ((n0 << 1) << (63u - sh)));
  }
  FLS_MEM void skip(uint32_t k) {  // k in {1,2}
    bi += k;
    if (bi >= 8u) { bi -= 8u; cur = n0; n0 = n1; n1 = n2; n2 = fetch(); }
  }
};

// ===================================================================== end-offset stream -> offsets[i+1]
// chunk: seg[0]=hdr, seg[1]=data ([lane bases 128 B][packed deltas]); out = &offsets[first row of rg]
struct VecSTROFFS {
  static constexpr uint32_t SMEM = DecDELTA<uint32_t>::SMEM, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    const uint32_t bw = FLS_META_BW(h.meta);
    DecDELTA<uint32_t>::to_tile(w, k.seg[1] + 16ull * h.ep, bw > 32u ? 32u : bw, h.base);
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC + 1;
    FLS_LANES(t) { fls_tile_copy_out<uint32_t>(w.sm + DecDELTA<uint32_t>::STAGE, offs, t, k.pay_base); }
  }
};

// ===================================================================== constant string
// chunk: seg[0] = the value's bytes, aux32 = its length
struct VecSTRCONST {
  static constexpr uint32_t SMEM = 16, BLOCK_SMEM = 0;
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &, const uint8_t *) {
    const uint32_t len = k.aux32, add = k.pay_base + v * FLS_VEC * len;
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC;
    FLS_LANES(t) {
      for (uint32_t j = 0; j < 32u; ++j) offs[j * 32u + t + 1u] = add + (j * 32u + t + 1u) * len;
      if (len) {
        FlsByteSink sink;
        sink.init(k.pay + add + (uint64_t)32u * t * len);
        for (uint32_t i = 0; i < 32u; ++i)
          for (uint32_t b = 0; b < len; ++b) sink.push(fls_ldg(k.seg[0] + b), 1);
        sink.finish();
      }
    }
  }
};

// ===================================================================== dictionary strings
// chunk: seg[0]/[1] = idx stream (u32 FFOR), seg[2] = dict_offsets u32[n+1], seg[3] = dict bytes,
//        seg[4] = vec_ep u32[n_vec+1] (payload offset of each vector inside the row group)
struct VecSTRDICT {
  static constexpr uint32_t TILE = 4096u, XCH = TILE + 4224u;
  static constexpr uint32_t SMEM = XCH + 128u;
  // Block-shared dictionary cache (the block's 4 warps decode vectors of the same chunk):
  //   [hdr: u32 maxlen, u32 stride][lens u16 x 1024][tab: n_dict x stride u64, entries zero padded]
  // Built when the dictionary is small (<= 1024 entries and <= 8 KB padded); stride = 0 means "not cached".
  static constexpr uint32_t CACHE_ENTRIES = 1024u, CACHE_BYTES = 8192u;
  static constexpr uint32_t LENS = 16u, TAB = LENS + 2u * CACHE_ENTRIES;
  static constexpr uint32_t BLOCK_SMEM = TAB + CACHE_BYTES;
  FLS_MEM static void block_init(const FlsChunkDev &k, uint8_t *bsm, uint32_t tid, uint32_t nthreads) {
    uint32_t *hdr = (uint32_t *)bsm;
    uint16_t *lens = (uint16_t *)(bsm + LENS);
    uint64_t *tab = (uint64_t *)(bsm + TAB);
    const uint32_t *doff = (const uint32_t *)k.seg[2];
    const uint8_t *dbytes = k.seg[3];
    const uint32_t n_dict = k.aux32;
    if (tid == 0) { hdr[0] = 0; hdr[1] = 0; hdr[2] = 0; }
    FLS_BLOCK_SYNC();
    if (n_dict > CACHE_ENTRIES) return;  // uniform over the block
    uint32_t m = 0, mn = 0xFFFFFFFFu;
    for (uint32_t i = tid; i < n_dict; i += nthreads) {
      const uint32_t L = fls_ldg(doff + i + 1) - fls_ldg(doff + i);
      m = fls_max(m, L);
      mn = fls_min(mn, L);
    }
    fls_smem_max(hdr, m);
    fls_smem_max(hdr + 2, ~mn);  // max of the complement = complement of the min
    FLS_BLOCK_SYNC();
    const uint32_t stride = (hdr[0] + 7u) >> 3;
    if (stride == 0 || n_dict * stride * 8u > CACHE_BYTES) return;
    for (uint32_t i = tid; i < n_dict; i += nthreads) {
      const uint32_t a = fls_ldg(doff + i), L = fls_ldg(doff + i + 1) - a;
      lens[i] = (uint16_t)L;
      for (uint32_t wd = 0; wd < stride; ++wd) {
        uint64_t x = 0;
        for (uint32_t b = 0; b < 8u && 8u * wd + b < L; ++b) x |= (uint64_t)fls_ldg(dbytes + a + 8u * wd + b) << (8u * b);
        tab[i * stride + wd] = x;
      }
    }
    if (tid == 0) hdr[1] = stride;
  }
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *bsm) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    uint32_t bw = FLS_META_BW(h.meta);
    if (bw > 32u) bw = 32u;
    const uint32_t *doff = (const uint32_t *)k.seg[2];
    const uint8_t *dbytes = k.seg[3];
    const uint32_t n_dict = k.aux32;
    const uint32_t vec_ep = fls_ldg((const uint32_t *)k.seg[4] + v);
    const uint32_t stride = ((const uint32_t *)bsm)[1];
    const uint32_t maxlen = ((const uint32_t *)bsm)[0], minlen = ~((const uint32_t *)bsm)[2];
    // every entry has the same power-of-two length (flags, codes): strings pack into whole words
    const uint32_t ulen = (stride == 1u && maxlen == minlen && (maxlen & (maxlen - 1u)) == 0u) ? maxlen : 0u;
    const uint16_t *lens = (const uint16_t *)(bsm + LENS);
    const uint64_t *tab = (const uint64_t *)(bsm + TAB);
    uint32_t *tile = (uint32_t *)(w.sm + TILE), *xch = (uint32_t *)(w.sm + XCH);
    if (bw) { w.stage(0, k.seg[1] + 16ull * h.ep, 128u * bw); w.wait(); }
    // 1. indices -> tile (original order: FFOR positions are original positions)
    FLS_LANES(t) {
      const uint32_t q = t & 7u, g = t >> 3;
      for (uint32_t j = 0; j < 8u; ++j) {
        const uint32_t row = g * 8u + j, p = row * 32u + q * 4u;
        fls_u4 ix;
        if (bw) ix = fls_unpack32_row(w.sm, bw, row, q, (uint32_t)h.base);
        else { ix.x = ix.y = ix.z = ix.w = (uint32_t)h.base; }
        tile[fls_tile_slot<uint32_t>(p)] = fls_min(ix.x, n_dict - 1u);
        tile[fls_tile_slot<uint32_t>(p + 1u)] = fls_min(ix.y, n_dict - 1u);
        tile[fls_tile_slot<uint32_t>(p + 2u)] = fls_min(ix.z, n_dict - 1u);
        tile[fls_tile_slot<uint32_t>(p + 3u)] = fls_min(ix.w, n_dict - 1u);
      }
    }
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC + 1;
    if (ulen) {
      // fixed-length fast path: offsets are arithmetic, 8/ulen strings fill one 8-byte word
      const uint32_t per = 8u / ulen, sh = 8u * ulen;
      FLS_LANES(t) {
        FlsWordSink sink;
        sink.init(k.pay + k.pay_base + vec_ep + 32u * t * ulen);
        for (uint32_t i = 0; i < 32u; i += per) {
          uint64_t x = 0;
          for (uint32_t j = 0; j < per; ++j) x |= tab[tile[33u * t + i + j]] << (sh * j);
          sink.push(x, 8u);
        }
        sink.finish();
#pragma unroll 8
        for (uint32_t j = 0; j < 32u; ++j) offs[j * 32u + t] = k.pay_base + vec_ep + (j * 32u + t + 1u) * ulen;
      }
      return;
    }
    // 2. bytes produced by each thread's 32 strings
    FLS_LANES(t) {
      uint32_t tot = 0;
      if (stride) {
#pragma unroll 8
        for (uint32_t i = 0; i < 32u; ++i) tot += lens[tile[33u * t + i]];
      } else {
        for (uint32_t i = 0; i < 32u; ++i) {
          const uint32_t d = tile[33u * t + i];
          tot += fls_ldg(doff + d + 1) - fls_ldg(doff + d);
        }
      }
      xch[t] = tot;
    }
    // 3. copy the strings, replace indices by end offsets
    FLS_LANES(t) {
      uint32_t pos = vec_ep;
      for (uint32_t u = 0; u < t; ++u) pos += xch[u];
      FlsWordSink sink;
      sink.init(k.pay + k.pay_base + pos);
      if (stride == 1u) {  // every entry fits one word: one append per string
#pragma unroll 4
        for (uint32_t i = 0; i < 32u; ++i) {
          const uint32_t d = tile[33u * t + i], L = lens[d];
          sink.push(tab[d], L);
          pos += L;
          tile[33u * t + i] = pos;
        }
      } else if (stride) {
        for (uint32_t i = 0; i < 32u; ++i) {
          const uint32_t d = tile[33u * t + i], L = lens[d];
          for (uint32_t o = 0; o < L; o += 8u) sink.push(tab[d * stride + (o >> 3)], fls_min(8u, L - o));
          pos += L;
          tile[33u * t + i] = pos;
        }
      } else {
        for (uint32_t i = 0; i < 32u; ++i) {
          const uint32_t d = tile[33u * t + i];
          const uint32_t a = fls_ldg(doff + d), b = fls_ldg(doff + d + 1);
          for (uint32_t c = a; c < b; ++c) sink.push(fls_ldg(dbytes + c), 1);
          pos += b - a;
          tile[33u * t + i] = pos;
        }
      }
      sink.finish();
    }
    FLS_LANES(t) { fls_tile_copy_out<uint32_t>(w.sm + TILE, offs, t, k.pay_base); }
  }
};

// ===================================================================== FSST strings
// chunk: seg[0]/[1] = decoded end-offset stream, seg[2]/[3] = compressed end-offset stream (both DELTA u32),
//        seg[4] = symbol table (u64 symbol[256] | u8 len[256]), seg[5] = compressed bytes (+16 slack).
// Lane t's 32 strings are contiguous in both the compressed and the decoded space, so the thread decodes ONE
// contiguous compressed run [c0,c1) into ONE contiguous output run starting at d0.
struct VecSTRFSST {
  // Phase 1 uses the slot as [decoded-ends DELTA stage + tile | compressed-ends DELTA stage]; phase 2 reuses the
  // whole slot as the window the vector's compressed bytes are bulk-copied into.
  static constexpr uint32_t CSTAGE = DecDELTA<uint32_t>::SMEM;  // compressed-ends payload staged after the tile
  static constexpr uint32_t WINDOW = 18432u;                    // 3 blocks of 4 warps per SM
  static constexpr uint32_t SMEM = WINDOW;
  static constexpr uint32_t BLOCK_SMEM = 2304;  // symbol table, shared by the block's warps (same chunk)
  static_assert(CSTAGE + 128u + 4096u <= WINDOW, "phase-1 buffers must fit the slot");
  FLS_MEM static void block_init(const FlsChunkDev &k, uint8_t *bsmem, uint32_t tid, uint32_t nthreads) {
    const uint64_t *src = (const uint64_t *)k.seg[4];
    uint64_t *dst = (uint64_t *)bsmem;
    for (uint32_t i = tid; i < 2304u / 8u; i += nthreads) {
      uint64_t x = fls_ldg(src + i);
      if (i == 255u) x = 0;                                                      // sym[255]
      if (i == 256u + 31u) x = (x & 0x00FFFFFFFFFFFFFFull) | (1ull << 56);       // len[255] = 1
      dst[i] = x;
    }
  }
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *symtab) {
    const fls_vechdr hd = fls_load_hdr(k, 0, v), hc = fls_load_hdr(k, 2, v);
    uint32_t bwd = FLS_META_BW(hd.meta), bwc = FLS_META_BW(hc.meta);
    if (bwd > 32u) bwd = 32u;
    if (bwc > 32u) bwc = 32u;
    // both offset streams in flight at once
    w.stage(CSTAGE, k.seg[3] + 16ull * hc.ep, 128u + 128u * bwc);
    DecDELTA<uint32_t>::to_tile(w, k.seg[1] + 16ull * hd.ep, bwd, hd.base);  // stages, waits for both, fills tile
    const uint32_t *dbases = (const uint32_t *)w.sm;
    const uint32_t *cbases = (const uint32_t *)(w.sm + CSTAGE);
    const uint32_t *cpk = (const uint32_t *)(w.sm + CSTAGE + 128u);
    const uint64_t *sym = (const uint64_t *)symtab;
    const uint8_t *len = symtab + 2048;
    const uint8_t *comp = k.seg[5];
    uint8_t *pay = k.pay + k.pay_base;
    // lane t's 32 strings are contiguous originals: compressed run [c0,c1) -> decoded run starting at d0
    FLS_VAR(uint32_t, c0); FLS_VAR(uint32_t, c1); FLS_VAR(uint32_t, d0);
    FLS_LANES(t) {
      FLS_AT(c0) = cbases[t];
      FLS_AT(d0) = dbases[t];
      uint32_t e = FLS_AT(c0) + 32u * (uint32_t)hc.base;
      if (bwc)
        for (uint32_t r = 0; r < 32u; ++r) e += fls_extract<uint32_t>(cpk, bwc, r, t);
      FLS_AT(c1) = e;
    }
    // offsets leave first: afterwards nothing in the slot is live and it becomes the compressed-byte window
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC + 1;
    FLS_LANES(t) { fls_tile_copy_out<uint32_t>(w.sm + DecDELTA<uint32_t>::STAGE, offs, t, k.pay_base); }
    // lane 0 owns originals [0,32), lane 31 owns [992,1024): the vector's compressed bytes are [cv0, cv1)
    const uint32_t cv0 = FLS_BCAST(c0, 0), cv1 = FLS_BCAST(c1, 31);
    const uint32_t wbase = cv0 & ~15u;
    const uint32_t span = cv1 > wbase ? cv1 - wbase : 0u;
    if (span != 0u && span + 16u <= WINDOW) {
      w.fence_async();
      w.stage(0, comp + wbase, (span + 15u) & ~15u);
      w.wait();
      const uint8_t *cb = w.sm;
      FLS_LANES(t) {
        uint32_t p = FLS_AT(c0) - wbase;
        const uint32_t pend = FLS_AT(c1) - wbase;
        if (pend > p && pend <= span) {
          FlsWordSink sink;
          sink.init(pay + FLS_AT(d0));
          // len[255] is 1 and sym[255] is 0 in the block's copy of the table (block_init), so an escape is
          // just "take the literal instead of the symbol": no divergent path in the loop.
          uint32_t code = cb[p], lit = cb[p + 1u];
          for (;;) {
            const uint32_t esc = code == 255u ? 1u : 0u;
            const uint32_t np = p + 1u + esc;
            if (np > pend) break;  // end of the run (or a dangling escape at its end)
            const uint32_t L = len[code];
            uint64_t sv = sym[code];
            if (esc) sv = lit;
            p = np;
            code = cb[p];
            lit = cb[p + 1u];
            sink.push(sv, L);
          }
          sink.finish();
        }
      }
      return;
    }
    // vectors whose compressed bytes do not fit the window: per-lane streaming from global memory
    FLS_LANES(t) {
      const uint32_t a0 = FLS_AT(c0), a1 = FLS_AT(c1);
      if (a1 > a0) {
        FlsByteStream src;
        src.init(comp + a0, comp + a1);
        FlsWordSink sink;
        sink.init(pay + FLS_AT(d0));
        uint32_t left = a1 - a0;
        while (left) {
          const uint32_t pk = src.peek2(), code = pk & 0xFFu;
          const uint32_t kk = code == 255u ? 2u : 1u;
          if (kk > left) break;
          const uint32_t L = code == 255u ? 1u : (uint32_t)len[code];
          const uint64_t sv = code == 255u ? (uint64_t)((pk >> 8) & 0xFFu) : sym[code];
          left -= kk;
          src.skip(kk);
          sink.push(sv, L);
        }
        sink.finish();
      }
    }
  }
};

// ===================================================================== chunk-level helpers
// FSST-compressed dictionary -> decoded dictionary bytes in scratch (k.out); one thread per entry.
//   seg[2] = compressed offsets u32[n+1], seg[3] = compressed bytes, seg[5] = symtab, seg[6] = decoded offsets
FLS_DEV void fls_fsst_dict_entry(const FlsChunkDev &k, uint32_t d) {
  const uint32_t *coff = (const uint32_t *)k.seg[2], *doff = (const uint32_t *)k.seg[6];
  const uint64_t *sym = (const uint64_t *)k.seg[5];
  const uint8_t *len = k.seg[5] + 2048, *in = k.seg[3];
  uint8_t *o = (uint8_t *)k.out + fls_ldg(doff + d);
  for (uint32_t i = fls_ldg(coff + d), e = fls_ldg(coff + d + 1); i < e;) {
    const uint32_t code = fls_ldg(in + i++);
    if (code == 255u) { *o++ = fls_ldg(in + i++); }
    else { uint64_t s = fls_ldg(sym + code); for (uint32_t b = fls_ldg(len + code); b; --b) { *o++ = (uint8_t)s; s >>= 8; } }
  }
}

#ifndef FLS_HOST_EMUL
__global__ void __launch_bounds__(128) fls_fsst_dictprep_kernel(const FlsChunkDev *__restrict__ chunks, uint32_t bpc) {
  const FlsChunkDev &k = chunks[blockIdx.x / bpc];
  for (uint32_t d = (blockIdx.x % bpc) * blockDim.x + threadIdx.x; d < k.aux32; d += bpc * blockDim.x)
    fls_fsst_dict_entry(k, d);
}
// payload blob copy seg[0] -> pay + pay_base, aux64[0] bytes, arbitrary destination alignment
__global__ void __launch_bounds__(256) fls_str_copy_kernel(const FlsChunkDev *__restrict__ chunks, uint32_t bpc) {
  const FlsChunkDev &k = chunks[blockIdx.x / bpc];
  const uint8_t *src = k.seg[0];
  uint8_t *dst = k.pay + k.pay_base;
  const uint64_t n = k.aux64[0];
  const uint32_t head = (uint32_t)((16u - ((uintptr_t)dst & 15u)) & 15u);  // bytes until dst is 16-byte aligned
  const uint64_t tid = (blockIdx.x % bpc) * (uint64_t)blockDim.x + threadIdx.x, nthr = (uint64_t)bpc * blockDim.x;
  if (n <= head + 16u) {
    for (uint64_t i = tid; i < n; i += nthr) dst[i] = __ldg(src + i);
    return;
  }
  const uint64_t n16 = (n - head) >> 4;
  for (uint64_t i = tid; i < n16; i += nthr) {  // aligned stores, byte-granular (L1-cached) loads
    const uint8_t *s = src + head + 16u * i;
    uint32_t wv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      wv[j] = (uint32_t)__ldg(s + 4 * j) | ((uint32_t)__ldg(s + 4 * j + 1) << 8) | ((uint32_t)__ldg(s + 4 * j + 2) << 16) |
              ((uint32_t)__ldg(s + 4 * j + 3) << 24);
    *(uint4 *)(dst + head + 16u * i) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
  }
  for (uint64_t i = tid; i < head; i += nthr) dst[i] = __ldg(src + i);
  for (uint64_t i = head + 16u * n16 + tid; i < n; i += nthr) dst[i] = __ldg(src + i);
}
#else
static void fls_str_copy_chunk_host(const FlsChunkDev &k) { memcpy(k.pay + k.pay_base, k.seg[0], k.aux64[0]); }
static void fls_fsst_dictprep_chunk_host(const FlsChunkDev &k) {
  for (uint32_t d = 0; d < k.aux32; ++d) fls_fsst_dict_entry(k, d);
}
#endif
