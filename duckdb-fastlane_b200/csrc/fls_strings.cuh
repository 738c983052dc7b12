// fls_strings.cuh — string column decoders (Arrow layout out: u32 offsets[n+1] + payload bytes).
//
// DICT and FSST columns: a block of WARPS warps owns (a slice of) one column chunk and keeps the chunk's lookup
// table in shared memory; each warp decodes whole vectors, a quarter (256 strings) at a time, lane t owning 8
// consecutive strings = one CONTIGUOUS run of output bytes that it appends to the warp's staging buffer
// (FlsPush); the quarter then leaves for HBM as coalesced 128-bit stores.
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

// ---- staged word writer.  A lane appends pieces of <= 8 bytes to ONE contiguous run inside the warp's shared
// staging buffer.  State: `o` = 4-aligned address of the word being filled, `n8` = 8 * (bytes pending in that
// word, 0..3), `prev` = a word whose TOP n8/8 bytes are the pending bytes.  A piece comes as a 16-byte table
// entry {lo, hi, tail, L8}: lo/hi = the bytes (zero above the length), L8 = 8 * length, tail = the piece's last
// four bytes (= lo when the length is < 4).  One push is two funnel shifts, two predicated word stores and one
// clamped funnel shift:
//   w0 = (prev:lo) << n8 >> 32      completes the pending word             stored when n8 + L8 >= 32
//   w1 = (lo:hi)   << n8 >> 32      the next four bytes                    stored when n8 + L8 >= 64
//   prev' = bytes [L, L+4) of (prev, lo, hi) = L < 4 ? (prev:lo) >> L8 : tail      (SHF.R.CLAMP does the select)
// A push with L8 = 0 changes nothing.  The first word of a run may be shared with the previous lane's run: it is
// stored with zeros in the foreign low bytes, and every lane writes its own trailing partial word BYTE-wise in
// finish(), which runs in a later warp phase — so shared words always end up complete, whatever the order.
#ifdef FLS_HOST_EMUL
typedef uint8_t *fls_sptr;  // address inside the warp's shared-memory slot
FLS_DEV fls_sptr fls_sptr_of(uint8_t *p) { return p; }
template <uint32_t THR> FLS_DEV void fls_sts32_ge(uint32_t m, fls_sptr a, uint32_t v) { if (m >= THR) *(uint32_t *)a = v; }
FLS_DEV void fls_sts8(fls_sptr a, uint32_t v) { *a = (uint8_t)v; }
#else
typedef uint32_t fls_sptr;  // shared-space address: STS with a 32-bit address register, no generic-address checks
FLS_DEV fls_sptr fls_sptr_of(uint8_t *p) { return (uint32_t)__cvta_generic_to_shared(p); }
// predicated store (no branch): if (m >= THR) *(u32 *)a = v
template <uint32_t THR> FLS_DEV void fls_sts32_ge(uint32_t m, fls_sptr a, uint32_t v) {
  asm volatile("{\n .reg .pred p;\n setp.ge.u32 p, %0, %1;\n @p st.shared.u32 [%2], %3;\n}" ::"r"(m), "n"(THR), "r"(a), "r"(v)
               : "memory");
}
FLS_DEV void fls_sts8(fls_sptr a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
#endif

struct FlsPush {
  fls_sptr o0;   // address of the run's first (possibly shared) word
  uint32_t T;    // bits appended so far + 8 * (bytes of the first word that precede the run)
  uint32_t T0;   // the initial T
  uint32_t prev; // a word whose top (T & 31) / 8 bytes are the pending bytes
  FLS_MEM void init(uint8_t *stg, uint32_t s) {  // s = byte offset of the run's first byte inside stg
    o0 = fls_sptr_of(stg) + (s & ~3u);
    T = T0 = 8u * (s & 3u);
    prev = 0;
  }
  // everything a piece needs is a function of (T before, T after, prev before): the pieces of a batch only
  // chain through one add (T) and one funnel shift (prev) each, the rest is independent work
  FLS_MEM static void emit(fls_sptr o0, uint32_t Tb, uint32_t Ta, uint32_t pv, uint32_t lo, uint32_t hi) {
    const uint32_t a = Tb >> 5, d = (Ta >> 5) - a;
    const fls_sptr addr = o0 + 4u * a;
    fls_sts32_ge<1u>(d, addr, fls_shf_l(pv, lo, Tb));       // funnel shifts use T mod 32
    fls_sts32_ge<2u>(d, addr + 4u, fls_shf_l(lo, hi, Tb));
  }
  FLS_MEM void push(uint32_t lo, uint32_t hi, uint32_t tail, uint32_t L8) {
    const uint32_t Ta = T + L8;
    emit(o0, T, Ta, prev, lo, hi);
    prev = fls_shf_rc(prev, tail, L8);
    T = Ta;
  }
  FLS_MEM void push(const fls_u4 &e) { push(e.x, e.y, e.z, e.w); }
  template <uint32_t N> FLS_MEM void push_all(const fls_u4 (&e)[N]) {
    uint32_t Tj[N + 1], pj[N + 1];
    Tj[0] = T; pj[0] = prev;
#pragma unroll
    for (uint32_t j = 0; j < N; ++j) { Tj[j + 1] = Tj[j] + e[j].w; pj[j + 1] = fls_shf_rc(pj[j], e[j].z, e[j].w); }
#pragma unroll
    for (uint32_t j = 0; j < N; ++j) emit(o0, Tj[j], Tj[j + 1], pj[j], e[j].x, e[j].y);
    T = Tj[N]; prev = pj[N];
  }
  FLS_MEM void push_byte(uint32_t b) { push(b, 0u, b, 8u); }
  FLS_MEM void finish() {
    const uint32_t n = (T & 31u) >> 3, first = (T >> 5) == 0u ? (T0 >> 3) : 0u;
    const fls_sptr o = o0 + 4u * (T >> 5);
    for (uint32_t i = first; i < n; ++i) fls_sts8(o + i, (prev >> (8u * (4u - n + i))) & 0xFFu);
  }
};

// 16-byte table entry of a piece of L <= 8 bytes held in the low bytes of `v`
FLS_DEV fls_u4 fls_piece_entry(uint64_t v, uint32_t L) {
  if (L < 8u) v &= (1ull << (8u * L)) - 1ull;
  fls_u4 e;
  e.x = (uint32_t)v;
  e.y = (uint32_t)(v >> 32);
  e.z = L < 4u ? e.x : (uint32_t)(v >> (8u * (L - 4u)));
  e.w = 8u * L;
  return e;
}

// Copy `len` staged bytes stg[h .. h+len) to global memory at `g` (h = g & 15, so 16-byte pieces line up):
// whole pieces as coalesced 128-bit loads/stores, the partial first/last piece byte-wise (those bytes' neighbours
// belong to another quarter / vector / row group and are written by their own warp).  Call inside FLS_LANES.
FLS_DEV void fls_stage_copy_out(const uint8_t *stg, uint8_t *g, uint32_t h, uint32_t len, uint32_t t) {
  const uint32_t end = h + len;
  const uint32_t p0 = h ? 1u : 0u, p1 = end >> 4;
  const uint32_t head_end = fls_min(end, 16u * p0);
  const uint32_t tail_start = fls_max(16u * p1, head_end);
  uint8_t *gb = g - h;  // 16-byte aligned
  for (uint32_t p = p0 + t; p < p1; p += 32u) fls_st128(gb + 16u * p, fls_lds128<fls_u4>(stg + 16u * p));
  for (uint32_t i = h + t; i < head_end; i += 32u) gb[i] = stg[i];
  for (uint32_t i = tail_start + t; i < end; i += 32u) gb[i] = stg[i];
}

// The same, for a whole warp (call OUTSIDE FLS_LANES): the whole 16-byte pieces leave as ONE bulk copy shared ->
// global (no load/store-pipe work; w.store_wait() before the staging buffer is written again), the partial first /
// last piece byte-wise.  `stg` is the warp's slot (w.sm).
FLS_DEV void fls_stage_store(FlsWarp &w, uint8_t *g, uint32_t h, uint32_t len) {
  const uint8_t *stg = w.sm;
  const uint32_t end = h + len;
  const uint32_t p0 = h ? 1u : 0u, p1 = end >> 4;
  const uint32_t head_end = fls_min(end, 16u * p0);
  const uint32_t tail_start = fls_max(16u * p1, head_end);
  uint8_t *gb = g - h;  // 16-byte aligned
  w.store_out(gb + 16u * p0, 16u * p0, p1 > p0 ? 16u * (p1 - p0) : 0u);
  FLS_LANES(t) {
    for (uint32_t i = h + t; i < head_end; i += 32u) gb[i] = stg[i];
    for (uint32_t i = tail_start + t; i < end; i += 32u) gb[i] = stg[i];
  }
}

// The same with the staging buffer at byte `off` (a multiple of 16) of the warp's slot.
FLS_DEV void fls_stage_store_at(FlsWarp &w, uint32_t off, uint8_t *g, uint32_t h, uint32_t len) {
  const uint8_t *stg = w.sm + off;
  const uint32_t end = h + len;
  const uint32_t p0 = h ? 1u : 0u, p1 = end >> 4;
  const uint32_t head_end = fls_min(end, 16u * p0);
  const uint32_t tail_start = fls_max(16u * p1, head_end);
  uint8_t *gb = g - h;  // 16-byte aligned
  w.store_out(gb + 16u * p0, off + 16u * p0, p1 > p0 ? 16u * (p1 - p0) : 0u);
  FLS_LANES(t) {
    for (uint32_t i = h + t; i < head_end; i += 32u) gb[i] = stg[i];
    for (uint32_t i = tail_start + t; i < end; i += 32u) gb[i] = stg[i];
  }
}

// DELTA u32 chain walk from a staged payload ([lane bases 128 B][packed deltas]) into a padded tile (original order)
FLS_DEV void fls_walk32(const uint8_t *stage, uint32_t bw, uint32_t base, uint32_t *tile, uint32_t t) {
  const uint32_t *bases = (const uint32_t *)stage, *pk = (const uint32_t *)(stage + 128);
  uint32_t acc = bases[t];
  const uint32_t e0 = fls_lane_start(t);
#pragma unroll
  for (uint32_t r = 0; r < 32u; ++r) {
    acc += (bw ? fls_extract<uint32_t>(pk, bw, fls_chain_row<32>(r), t) : 0u) + base;
    tile[fls_tile_slot<uint32_t>(e0 + r)] = acc;
  }
}

// ===================================================================== dictionary strings
// chunk: seg[0]/[1] = idx stream (u32 FFOR), seg[2] = dict_offsets u32[n+1], seg[3] = dict bytes,
//        seg[4] = vec_ep u32[n_vec+1] (payload offset of each vector inside the row group)
// A vector is decoded in four quarters of 256 strings; in a quarter lane t owns the 8 consecutive strings
// 256q + 8t .. +7: their indices sit in ONE packed row (row 8q + t/4, packed lanes 8(t%4) .. +7), so the lane
// unpacks them straight from the staged payload with 128-bit loads; its strings form one contiguous run of
// output bytes, appended piece by piece (FlsPush) to the warp's staging buffer, and its 8 end offsets leave as
// two aligned 128-bit stores {run start, end 0 .. end 6} (end 7 is the next lane's run start).
// Two configurations of the same kernel: short entries (the planner routes a chunk here when 256 of its longest entry
// fit the staging buffer) run 20 warps with a 6.5 KB staging buffer; every other dictionary the round-1 configuration,
// 16 warps and 8.25 KB (quarters that still do not fit go byte-wise).  More warps keep paying on this kernel: 12 / 14 /
// 16 / 18 / 20 warps = 0.508 / 0.468 / 0.437 / 0.420 / 0.403 ms on the two dictionary columns of the SF10 launch
// (profiles/r02_dict_warps.log); at 20 the compiler is held to 96 registers.
template <uint32_t WARPS_, uint32_t STG_> struct VecSTRDICTT {
  static constexpr uint32_t WARPS = WARPS_;
  static constexpr uint32_t STG = STG_;      // staging of one quarter's payload
  // staged index payload: two halves of HALF bytes (this vector's and the prefetched next one's, widths <= 10 bits),
  // or one vector's payload up to 20 bits wide, or — wider still — the packed rows of one quarter at a time
  static constexpr uint32_t IDX = STG, HALF = 1280u;
  static constexpr uint32_t SMEM = STG + 2u * HALF;
  // block-shared: [hdr 16 B: u32 pieces, u32 maxlen, u32 ~minlen][lens u32 x 1024][pieces: n_dict x P x 16 B]
  static constexpr uint32_t CACHE_ENTRIES = 1024u, TAB_BYTES = 16384u;
  static constexpr uint32_t LENS = 16u, TAB = LENS + 4u * CACHE_ENTRIES;
  // ... [vector headers 64 x 16 B][vec_ep u32 x 65]: read once per block instead of once per vector from HBM
  static constexpr uint32_t HDRS = TAB + TAB_BYTES, VEP = HDRS + 64u * 16u;
  static constexpr uint32_t BLOCK_SMEM = VEP + 68u * 4u;
  static constexpr uint32_t NBUF = 2, CONST_SMEM = 0;  // two tables per block: the next chunk's is built ahead
  static constexpr uint32_t TABLES_SMEM = NBUF * BLOCK_SMEM;
  FLS_MEM static constexpr uint32_t tab_off(uint32_t b) { return b * BLOCK_SMEM; }
  static constexpr bool PIPELINED = true;
  FLS_MEM static void const_init(uint8_t *, uint32_t, uint32_t) {}
  // called by all `nthreads` threads of a group (a whole block, or ONE warp: nthreads == 32)
  // Global loads are issued in batches of independent loads held in registers (B per thread) and stored
  // afterwards: a loop of load -> store iterations would expose one full memory latency per iteration, and this
  // runs on ONE warp between two segments.  Three dependent rounds in all: (headers, entry points, entry lengths),
  // entry offsets, entry bytes.
  FLS_MEM static void block_init(const FlsChunkDev &k, uint8_t *bsm, uint32_t tid, uint32_t nthreads) {
    constexpr uint32_t B = 4;
    uint32_t *hdr = (uint32_t *)bsm;
    uint32_t *lens = (uint32_t *)(bsm + LENS);
    const uint32_t *doff = (const uint32_t *)k.seg[2];
    const uint8_t *dbytes = k.seg[3];
    const uint32_t n_dict = k.aux32;
    const bool cached = n_dict <= CACHE_ENTRIES;
    if (tid == 0) { hdr[0] = 0; hdr[1] = 0; hdr[2] = 0; }
    FLS_GROUP_SYNC(nthreads);
    uint32_t m = 0, mn = 0xFFFFFFFFu;
    {
      const uint32_t nh = k.n_vec <= 64u ? 2u * k.n_vec : 0u, ne = k.n_vec <= 64u ? k.n_vec + 1u : 0u, nl = cached ? n_dict : 0u;
      for (uint32_t i0 = 0; i0 < fls_max(fls_max(nh, ne), nl); i0 += B * nthreads) {
        uint64_t rh[B];
        uint32_t re[B], ra[B], rb[B];
#pragma unroll
        for (uint32_t j = 0; j < B; ++j) {
          const uint32_t i = i0 + j * nthreads + tid;
          rh[j] = i < nh ? fls_ldg((const uint64_t *)k.seg[0] + i) : 0ull;
          re[j] = i < ne ? fls_ldg((const uint32_t *)k.seg[4] + i) : 0u;
          ra[j] = i < nl ? fls_ldg(doff + i) : 0u;
          rb[j] = i < nl ? fls_ldg(doff + i + 1) : 0u;
        }
#pragma unroll
        for (uint32_t j = 0; j < B; ++j) {
          const uint32_t i = i0 + j * nthreads + tid;
          if (i < nh) ((uint64_t *)(bsm + HDRS))[i] = rh[j];
          if (i < ne) ((uint32_t *)(bsm + VEP))[i] = re[j];
          if (i < nl) { const uint32_t L = rb[j] - ra[j]; lens[i] = L; m = fls_max(m, L); mn = fls_min(mn, L); }
        }
      }
    }
    if (!cached) return;  // uniform over the group: P stays 0 = "not cached"
    fls_smem_max(hdr + 1, m);
    fls_smem_max(hdr + 2, ~mn);  // max of the complement = complement of the min
    FLS_GROUP_SYNC(nthreads);
    const uint32_t P = (hdr[1] + 7u) >> 3;
    if (P == 0 || n_dict * P * 16u > TAB_BYTES) return;
    // small tables keep only the pieces' 8 bytes, 16 times (piece i, copy r at i * 128 + r * 8; lane t reads copy
    // t & 15 with one 64-bit load): the sixteen lanes of a 64-bit load phase never share a bank, whatever pieces they
    // look up, and a lookup costs 2 shared-memory wavefronts instead of 4 — a piece's length and last four bytes
    // follow from the string's length, which the lane already holds
    const uint32_t sh = n_dict * P * 128u <= TAB_BYTES ? 7u : 4u;
    for (uint32_t i0 = 0; i0 < n_dict * P; i0 += B * nthreads) {
      uint32_t a[B], pos[B], Lp[B];
      uint64_t lo[B], hi[B];
#pragma unroll
      for (uint32_t j = 0; j < B; ++j) {
        const uint32_t i = i0 + j * nthreads + tid, d = i / P;
        a[j] = i < n_dict * P ? fls_ldg(doff + d) : 0u;
      }
#pragma unroll
      for (uint32_t j = 0; j < B; ++j) {
        const uint32_t i = i0 + j * nthreads + tid, d = fls_min(i / P, n_dict - 1u), pc = i % P;
        const uint32_t L = lens[d];
        Lp[j] = (i < n_dict * P && L > 8u * pc) ? fls_min(8u, L - 8u * pc) : 0u;
        pos[j] = a[j] + 8u * pc;
        // the piece's <= 8 bytes at an arbitrary byte address = two aligned 8-byte words (the dictionary's
        // segment is 16-byte aligned and followed by 16 bytes of slack)
        const uint64_t *q = (const uint64_t *)(dbytes + (pos[j] & ~7u));
        lo[j] = Lp[j] ? fls_ldg(q) : 0ull;
        hi[j] = Lp[j] ? fls_ldg(q + 1) : 0ull;
      }
#pragma unroll
      for (uint32_t j = 0; j < B; ++j) {
        const uint32_t i = i0 + j * nthreads + tid, bs = 8u * (pos[j] & 7u);
        const uint64_t x = bs ? (lo[j] >> bs) | (hi[j] << (64u - bs)) : lo[j];
        if (i < n_dict * P) {
          const fls_u4 en = fls_piece_entry(x, Lp[j]);
          if (sh == 7u) { for (uint32_t r = 0; r < 16u; ++r) fls_sts64(bsm + TAB + 128u * i + 8u * r, en.x, en.y); }
          else *(fls_u4 *)(bsm + TAB + 16u * i) = en;
        }
      }
    }
    if (tid == 0) { hdr[0] = P; hdr[3] = sh; }
  }
  // the 8 strings of a lane, P pieces each: table loads are issued in batches of up to 8 entries BEFORE the
  // pushes that consume them (the compiler cannot hoist shared loads over the pushes' shared stores itself)
  // piece pc of a string of length L8 / 8, from the replicated 8-byte table
  FLS_MEM static fls_u4 piece_rep(const uint8_t *tabr, uint32_t idx, uint32_t pc, uint32_t L8) {
    const fls_u2 v = fls_lds64(tabr + 128u * idx);
    const uint32_t Lp8 = L8 > 64u * pc ? fls_min(64u, L8 - 64u * pc) : 0u;
    fls_u4 e;
    e.x = v.x; e.y = v.y; e.w = Lp8;
    e.z = fls_shf_rc(v.x, v.y, fls_max(Lp8, 32u) - 32u);  // last four bytes: lo for <= 4 bytes, else (lo:hi) >> 8(L - 4)
    return e;
  }
  template <uint32_t P> FLS_MEM static void push_strings_rep(FlsPush &p, const uint8_t *tabr, const uint32_t (&d)[8], const uint32_t (&len)[8]) {
    constexpr uint32_t SPB = P == 1 ? 8u : P == 2 ? 4u : 2u;  // strings per batch
#pragma unroll
    for (uint32_t s0 = 0; s0 < 8u; s0 += SPB) {
      fls_u4 e[SPB * P];
#pragma unroll
      for (uint32_t i = 0; i < SPB; ++i)
#pragma unroll
        for (uint32_t pc = 0; pc < P; ++pc) e[i * P + pc] = piece_rep(tabr, d[s0 + i] * P + pc, pc, 8u * len[s0 + i]);
      p.push_all(e);
    }
  }
  template <uint32_t P> FLS_MEM static void push_strings(FlsPush &p, const uint8_t *tab, uint32_t sh, const uint32_t (&d)[8]) {
    constexpr uint32_t SPB = P == 1 ? 8u : P == 2 ? 4u : 2u;  // strings per batch
#pragma unroll
    for (uint32_t s0 = 0; s0 < 8u; s0 += SPB) {
      fls_u4 e[SPB * P];
#pragma unroll
      for (uint32_t i = 0; i < SPB; ++i)
#pragma unroll
        for (uint32_t pc = 0; pc < P; ++pc) e[i * P + pc] = fls_lds128<fls_u4>(tab + ((d[s0 + i] * P + pc) << sh));
      p.push_all(e);
    }
  }
  // every entry has the same length U in {1, 2, 4}: the lane's 8 strings are 8U bytes, merged in registers into
  // U eight-byte pieces — U pushes instead of 8
  template <uint32_t U> FLS_MEM static void push_uniform(FlsPush &p, const uint8_t *tab, uint32_t sh, const uint32_t (&d)[8]) {
    uint32_t x[8], wd[2u * U];
#pragma unroll
    for (uint32_t i = 0; i < 8u; ++i) x[i] = *(const uint32_t *)(tab + (d[i] << sh));  // the entry's bytes, zero above U
    if (U == 1u) {
      wd[0] = x[0] | (x[1] << 8) | (x[2] << 16) | (x[3] << 24);
      wd[1] = x[4] | (x[5] << 8) | (x[6] << 16) | (x[7] << 24);
    } else if (U == 2u) {
#pragma unroll
      for (uint32_t j = 0; j < 4u; ++j) wd[j] = x[2u * j] | (x[2u * j + 1u] << 16);
    } else {
#pragma unroll
      for (uint32_t j = 0; j < 8u; ++j) wd[j % (2u * U)] = x[j];
    }
#pragma unroll
    for (uint32_t j = 0; j < U; ++j) p.push(wd[2u * j], wd[2u * j + 1u], wd[2u * j + 1u], 64u);
  }
  // the lane's 8 indices: packed row `row`, packed lanes 8(t%4) .. +7 of the staged FFOR payload
  FLS_MEM static void unpack8(const uint8_t *sm, uint32_t bw, uint32_t row, uint32_t t, uint32_t base, uint32_t lim,
                              uint32_t (&d)[8]) {
    if (!bw) {
#pragma unroll
      for (uint32_t i = 0; i < 8u; ++i) d[i] = fls_min(base, lim);
      return;
    }
    const uint32_t bit = row * bw, kk = bit >> 5, s = bit & 31u;
    const uint8_t *pa = sm + kk * 128u + 32u * (t & 3u), *pb = sm + fls_min(kk + 1u, bw - 1u) * 128u + 32u * (t & 3u);
    const fls_u4 a0 = fls_lds128<fls_u4>(pa), a1 = fls_lds128<fls_u4>(pa + 16);
    // the next word row only where the row's values straddle it (never for widths that divide 32): a quarter-warp
    // shares its row, so the skipped loads cost no shared-memory wavefronts
    fls_u4 b0, b1;
    b0.x = b0.y = b0.z = b0.w = 0u;
    b1 = b0;
    if (s + bw > 32u) { b0 = fls_lds128<fls_u4>(pb); b1 = fls_lds128<fls_u4>(pb + 16); }
    const uint32_t mask = bw >= 32u ? 0xFFFFFFFFu : ((1u << bw) - 1u);
    d[0] = fls_min((fls_shf_r(a0.x, b0.x, s) & mask) + base, lim);
    d[1] = fls_min((fls_shf_r(a0.y, b0.y, s) & mask) + base, lim);
    d[2] = fls_min((fls_shf_r(a0.z, b0.z, s) & mask) + base, lim);
    d[3] = fls_min((fls_shf_r(a0.w, b0.w, s) & mask) + base, lim);
    d[4] = fls_min((fls_shf_r(a1.x, b1.x, s) & mask) + base, lim);
    d[5] = fls_min((fls_shf_r(a1.y, b1.y, s) & mask) + base, lim);
    d[6] = fls_min((fls_shf_r(a1.z, b1.z, s) & mask) + base, lim);
    d[7] = fls_min((fls_shf_r(a1.w, b1.w, s) & mask) + base, lim);
  }
  // {run start, end 0 .. end 6} -> offsets[first string of the lane .. +7] (two aligned 128-bit stores)
  FLS_MEM static void store_offsets(uint32_t *o, uint32_t start, const uint32_t (&len)[8]) {
    fls_u4 a, b;
    a.x = start; a.y = a.x + len[0]; a.z = a.y + len[1]; a.w = a.z + len[2];
    b.x = a.w + len[3]; b.y = b.x + len[4]; b.z = b.y + len[5]; b.w = b.z + len[6];
    fls_st128(o, a);
    fls_st128(o + 4, b);
  }
  FLS_MEM static fls_vechdr header(const FlsChunkDev &k, const uint8_t *bsm, uint32_t v) {
    if (k.n_vec > 64u) return fls_load_hdr(k, 0, v);
    const uint64_t *p = (const uint64_t *)(bsm + HDRS) + 2u * v;
    fls_vechdr h;
    h.ep = (uint32_t)p[0]; h.meta = (uint32_t)(p[0] >> 32); h.base = p[1];
    return h;
  }
  // v_next: the vector this warp decodes next (FLS_NO_VECTOR if none): its index payload is bulk-copied into the
  // other half of the index buffer while v is decoded
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, uint32_t v_next, FlsWarp &w, const uint8_t *bsm, const uint8_t *) {
    const fls_vechdr h = header(k, bsm, v);
    uint32_t bw = FLS_META_BW(h.meta);
    if (bw > 32u) bw = 32u;
    const uint32_t *doff = (const uint32_t *)k.seg[2];
    const uint8_t *dbytes = k.seg[3];
    const uint32_t n_dict = k.aux32;
    const uint32_t vec_ep = k.n_vec <= 64u ? ((const uint32_t *)(bsm + VEP))[v] : fls_ldg((const uint32_t *)k.seg[4] + v);
    const uint32_t P = ((const uint32_t *)bsm)[0];
    const uint32_t maxlen = ((const uint32_t *)bsm)[1], minlen = ~((const uint32_t *)bsm)[2];
    // every entry has the same length 1, 2 or 4: offsets are arithmetic, strings merge into 8-byte pieces
    const uint32_t U = (P == 1u && maxlen == minlen && (maxlen == 1u || maxlen == 2u || maxlen == 4u)) ? maxlen : 0u;
    const uint32_t *lens = (const uint32_t *)(bsm + LENS);
    const uint8_t *tab = bsm + TAB;
    const uint32_t sh = ((const uint32_t *)bsm)[3];  // log2 of the table's entry stride: 4, or 7 when kept 8 times
    uint8_t *stg = w.sm;
    uint32_t ioff = IDX;  // where this vector's index payload is (half 0, half 1, or the whole buffer)
    const bool by_quarter = 128u * bw > 2u * HALF;  // too wide for the buffer: the packed rows of one quarter at a time
    if (w.pf_v == v) { ioff = w.pf_off; w.wait(); }
    else if (w.pf_v != FLS_NO_VECTOR) { w.wait(); w.pf_v = FLS_NO_VECTOR; }  // drain a copy nobody will use
    if (ioff != IDX || w.pf_v == v) { /* payload already staged */ }
    else if (bw && !by_quarter) { w.fence_async(); w.stage(IDX, k.seg[1] + 16ull * h.ep, 128u * bw); w.wait(); }
    w.pf_v = FLS_NO_VECTOR;
    if (v_next != FLS_NO_VECTOR && 128u * bw <= HALF) {
      const fls_vechdr hn = header(k, bsm, v_next);
      const uint32_t bwn = FLS_META_BW(hn.meta);
      if (bwn && 128u * bwn <= HALF) {
        w.pf_v = v_next;
        w.pf_off = ioff == IDX ? IDX + HALF : IDX;
        w.fence_async();
        w.stage(w.pf_off, k.seg[1] + 16ull * hn.ep, 128u * bwn);
      }
    }
    const uint8_t *idx = w.sm + ioff;
    uint8_t *pay = k.pay + k.pay_base;
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC;  // offs[i] = start of string i, offs[i + 1] = its end
    uint32_t qpos = vec_ep;  // payload offset (row-group relative) of the quarter's first byte
    const uint64_t rg_pay = k.aux64[0];  // payload bytes of this row group: no write may pass it (corrupt streams)
    if (U && !by_quarter && ((uintptr_t)(pay + vec_ep) & 15u) == 0u && (uint64_t)vec_ep + FLS_VEC * U <= rg_pay) {
      // Every entry is U = 1, 2 or 4 bytes long and the vector's payload starts on a 16-byte boundary: offsets are
      // arithmetic and a lane's 8 strings are 8U contiguous, aligned bytes — straight from registers to HBM, no
      // staging.  Dictionaries of up to 8 one-byte entries (flag columns) live in two registers and a lane's 8
      // bytes are two byte permutes of them.
      const bool tiny = U == 1u && n_dict <= 8u;
      uint32_t dlo = 0, dhi = 0;
      if (tiny) {
#pragma unroll
        for (uint32_t e = 0; e < 4u; ++e) {
          dlo |= (*(const uint32_t *)(tab + (e << sh)) & 0xFFu) << (8u * e);
          dhi |= (*(const uint32_t *)(tab + ((e + 4u) << sh)) & 0xFFu) << (8u * e);
        }
      }
      for (uint32_t qt = 0; qt < 4u; ++qt) {
        FLS_LANES(t) {
          uint32_t d[8];
          unpack8(idx, bw, 8u * qt + (t >> 2), t, (uint32_t)h.base, n_dict - 1u, d);
          const uint32_t s0 = 256u * qt + 8u * t;  // the lane's first string
          const uint32_t start = k.pay_base + vec_ep + s0 * U;
          fls_u4 a, b;
          a.x = start; a.y = start + U; a.z = start + 2u * U; a.w = start + 3u * U;
          b.x = start + 4u * U; b.y = start + 5u * U; b.z = start + 6u * U; b.w = start + 7u * U;
          fls_st128(offs + s0, a);
          fls_st128(offs + s0 + 4u, b);
          if (s0 == FLS_VEC - 8u) offs[FLS_VEC] = start + 8u * U;  // the vector's last end
          uint8_t *o = pay + vec_ep + s0 * U;
          if (tiny) {
            fls_st64(o, fls_prmt(dlo, dhi, d[0] | (d[1] << 4) | (d[2] << 8) | (d[3] << 12)),
                     fls_prmt(dlo, dhi, d[4] | (d[5] << 4) | (d[6] << 8) | (d[7] << 12)));
          } else {
            uint32_t x[8];
#pragma unroll
            for (uint32_t i = 0; i < 8u; ++i) x[i] = *(const uint32_t *)(tab + (d[i] << sh));  // the entry's bytes, zero above U
            if (U == 1u) {
              fls_st64(o, x[0] | (x[1] << 8) | (x[2] << 16) | (x[3] << 24), x[4] | (x[5] << 8) | (x[6] << 16) | (x[7] << 24));
            } else if (U == 2u) {
              fls_u4 y;
              y.x = x[0] | (x[1] << 16); y.y = x[2] | (x[3] << 16); y.z = x[4] | (x[5] << 16); y.w = x[6] | (x[7] << 16);
              fls_st128(o, y);
            } else {
              fls_u4 y0, y1;
              y0.x = x[0]; y0.y = x[1]; y0.z = x[2]; y0.w = x[3];
              y1.x = x[4]; y1.y = x[5]; y1.z = x[6]; y1.w = x[7];
              fls_st128(o, y0);
              fls_st128(o + 16, y1);
            }
          }
        }
      }
      return;
    }
    for (uint32_t qt = 0; qt < 4u; ++qt) {
      if (by_quarter) {
        // the quarter's indices are rows 8qt .. 8qt+7 = 32-bit words [k0, k1] of every packed lane (<= 9 word rows)
        const uint32_t k0 = (8u * qt * bw) >> 5, k1 = fls_min(((8u * qt + 8u) * bw - 1u) >> 5, bw - 1u);
        w.fence_async();
        w.stage(IDX, k.seg[1] + 16ull * h.ep + 128u * k0, 128u * (k1 - k0 + 1u));
        w.wait();
        idx = w.sm + IDX - 128u * k0;  // unpack8 addresses word rows absolutely; it reads rows k0 .. k1 only
      }
      // 1. the lane's 8 indices and lengths (kept in registers), exclusive scan -> the lane's first output byte
      FLS_VAR(uint32_t, run);
      FLS_VAR2(uint32_t, dd, 8);
      FLS_VAR2(uint32_t, ll, 8);
      FLS_LANES(t) {
        uint32_t d[8];
        unpack8(idx, bw, 8u * qt + (t >> 2), t, (uint32_t)h.base, n_dict - 1u, d);
        uint32_t tot = 0;
#pragma unroll
        for (uint32_t i = 0; i < 8u; ++i) {
          FLS_AT2(dd, i) = d[i];
          const uint32_t L = U ? U : (P ? lens[d[i]] : fls_ldg(doff + d[i] + 1) - fls_ldg(doff + d[i]));
          FLS_AT2(ll, i) = L;
          tot += L;
        }
        FLS_AT(run) = tot;
      }
      uint32_t qlen;
      FLS_EXCL_SCAN(run, qlen);
      if ((uint64_t)qpos + qlen > rg_pay) return;  // indices that do not add up to the declared payload: write nothing more
      uint8_t *g = pay + qpos;
      const uint32_t hh = (uint32_t)((uintptr_t)g & 15u);
      const bool staged = P && hh + qlen <= STG;
      // 2. end offsets, then the pieces of the lane's strings
      FLS_VAR(FlsPush, ps);
      w.store_wait();  // the previous quarter's bulk copy has read the staging buffer
      FLS_LANES(t) {
        uint32_t d[8], len[8];
#pragma unroll
        for (uint32_t i = 0; i < 8u; ++i) { d[i] = FLS_AT2(dd, i); len[i] = FLS_AT2(ll, i); }
        const uint32_t start = qpos + FLS_AT(run);
        store_offsets(offs + 256u * qt + 8u * t, k.pay_base + start, len);
        if (qt == 3u && t == 31u) offs[FLS_VEC] = k.pay_base + qpos + qlen;  // the vector's last end
        if (staged) {
          FlsPush p;
          p.init(stg, hh + FLS_AT(run));
          const uint8_t *tabr = tab + 8u * (t & 15u);  // this lane's copy of a replicated table
          if (U == 1u) push_uniform<1>(p, tab, sh, d);
          else if (U == 2u) push_uniform<2>(p, tab, sh, d);
          else if (U == 4u) push_uniform<4>(p, tab, sh, d);
          else if (sh == 7u) switch (P) {
            case 1: push_strings_rep<1>(p, tabr, d, len); break;
            case 2: push_strings_rep<2>(p, tabr, d, len); break;
            case 3: push_strings_rep<3>(p, tabr, d, len); break;
            case 4: push_strings_rep<4>(p, tabr, d, len); break;
            default:
              for (uint32_t i = 0; i < 8u; ++i)
                for (uint32_t pc = 0; pc < P; ++pc) p.push(piece_rep(tabr, d[i] * P + pc, pc, 8u * len[i]));
          }
          else switch (P) {
            case 1: push_strings<1>(p, tab, sh, d); break;
            case 2: push_strings<2>(p, tab, sh, d); break;
            case 3: push_strings<3>(p, tab, sh, d); break;
            case 4: push_strings<4>(p, tab, sh, d); break;
            default:
              for (uint32_t i = 0; i < 8u; ++i)
                for (uint32_t pc = 0; pc < P; ++pc) p.push(fls_lds128<fls_u4>(tab + ((d[i] * P + pc) << sh)));
          }
          FLS_AT(ps) = p;
        } else {
          // dictionary not cached, or a quarter larger than the staging buffer: byte-wise from global memory
          uint32_t pos = start;
          for (uint32_t i = 0; i < 8u; ++i) {
            const uint32_t a = fls_ldg(doff + d[i]), b = fls_ldg(doff + d[i] + 1);
            for (uint32_t c = a; c < b; ++c) pay[pos + (c - a)] = fls_ldg(dbytes + c);
            pos += b - a;
          }
        }
      }
      if (staged) {
        FLS_LANES(t) { FLS_AT(ps).finish(); }
        fls_stage_store(w, g, hh, qlen);
      }
      qpos += qlen;
    }
  }
};
typedef VecSTRDICTT<FLS_DICT_SHORT_WARPS, FLS_DICT_SHORT_STG> VecSTRDICT;  // entries of up to FLS_DICT_SHORT_MAXLEN bytes
typedef VecSTRDICTT<16u, 8448u> VecSTRDICTL;              // everything else

// ===================================================================== flag columns
// A STR_DICT chunk whose dictionary is up to 8 one-byte strings and whose vectors are all full (fls_dict_is_flags,
// fls_plan.h): aux64[1] = the entries' bytes.  One warp per vector, no table: string i of the vector is byte i of its
// 1024 payload bytes, offsets are arithmetic, a lane's 8 strings are two byte permutes of the dictionary.  The vector's
// payload position is 1024 v by construction (every string is one byte), whatever seg[4] says.
struct VecSTRFLAGS {
  static constexpr uint32_t SMEM = 4096u, BLOCK_SMEM = 0;  // the staged index payload, widths up to 32 bits
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr h = fls_load_hdr(k, 0, v);
    uint32_t bw = FLS_META_BW(h.meta);
    if (bw > 32u) bw = 32u;
    if (bw) { w.stage(0, k.seg[1] + 16ull * h.ep, 128u * bw); w.wait(); }
    const uint32_t dlo = (uint32_t)k.aux64[1], dhi = (uint32_t)(k.aux64[1] >> 32);
    const uint32_t lim = k.aux32 - 1u, first = k.pay_base + v * FLS_VEC;  // payload offset of the vector's first string
    uint8_t *pay = k.pay + first;
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC;
    const bool aligned = ((uintptr_t)pay & 7u) == 0u;
    for (uint32_t qt = 0; qt < 4u; ++qt) {
      FLS_LANES(t) {
        uint32_t d[8];
        VecSTRDICT::unpack8(w.sm, bw, 8u * qt + (t >> 2), t, (uint32_t)h.base, lim, d);
        const uint32_t s0 = 256u * qt + 8u * t, start = first + s0;
        fls_u4 a, b;
        a.x = start; a.y = start + 1u; a.z = start + 2u; a.w = start + 3u;
        b.x = start + 4u; b.y = start + 5u; b.z = start + 6u; b.w = start + 7u;
        fls_st128(offs + s0, a);
        fls_st128(offs + s0 + 4u, b);
        if (s0 == FLS_VEC - 8u) offs[FLS_VEC] = start + 8u;  // the vector's last end
        const uint32_t lo = fls_prmt(dlo, dhi, d[0] | (d[1] << 4) | (d[2] << 8) | (d[3] << 12));
        const uint32_t hi = fls_prmt(dlo, dhi, d[4] | (d[5] << 4) | (d[6] << 8) | (d[7] << 12));
        if (aligned) fls_st64(pay + s0, lo, hi);
        else {
#pragma unroll
          for (uint32_t i = 0; i < 4u; ++i) { pay[s0 + i] = (uint8_t)(lo >> (8u * i)); pay[s0 + 4u + i] = (uint8_t)(hi >> (8u * i)); }
        }
      }
    }
  }
};

// ===================================================================== FSST12 strings (12-bit codes, up to 4096 symbols)
// chunk: seg[0]/[1] = decoded end-offset stream, seg[2]/[3] = compressed end-offset stream (both DELTA u32),
//        seg[4] = symbol table (u64 symbol[aux32] | u8 len[aux32]; codes 0..255 = the bytes themselves, no escape),
//        seg[5] = compressed bytes.  A string's codes are packed LSB first, two per three bytes (an odd last code
//        takes two bytes), so string boundaries matter: a lane decodes string by string.
// One warp per vector.  Phase 1 walks both offset streams (decoded ends -> offsets[], compressed ends -> the warp's
// tile).  Phase 2, per quarter of 256 strings: the quarter's compressed bytes are bulk-copied into the slot, lane t
// owns the 8 consecutive strings 256q + 8t .. +7 = one contiguous run of output bytes, appended piece by piece
// (FlsPush) to a staging buffer that leaves as one bulk copy.  The 36 KB symbol table is read through L1 (a
// 4096-entry table does not fit the 8-bit kernel's replicated shared-memory layout).  A quarter whose compressed
// bytes or output do not fit the slot, or whose offsets are damaged, takes the byte-wise path.
// Slot: [0, 4224) offset-stream payload, later the compressed window | [4224, 8448) compressed ends (padded tile) |
//       [8448, 8448 + STG) staging.
struct VecSTRFSST12 {
  static constexpr uint32_t CW = DecDELTA<uint32_t>::STAGE;   // 4224: compressed window
  static constexpr uint32_t STG_OFF = DecDELTA<uint32_t>::SMEM, STG = 8448u;
  static constexpr uint32_t SMEM = STG_OFF + STG, BLOCK_SMEM = 0;
  static_assert(STG_OFF % 128u == 0 && SMEM % 128u == 0, "warp slots stay 128-byte aligned");
  FLS_MEM static uint32_t code_at(const uint8_t *in, uint32_t q) {  // code q of a string whose codes start at `in`
    const uint8_t *p = in + 3u * (q >> 1);
    return (q & 1u) ? ((uint32_t)p[1] >> 4) | ((uint32_t)p[2] << 4) : (uint32_t)p[0] | (((uint32_t)p[1] & 15u) << 8);
  }
  // byte-wise, straight from / to global memory: strings i0, i0 + step, .. < i1 of the vector
  FLS_MEM static void plain(const FlsChunkDev &k, const uint32_t *tile, const uint32_t *offs, uint32_t cv0, uint32_t dv0,
                            uint32_t i0, uint32_t i1, uint32_t step) {
    const uint64_t *sym = (const uint64_t *)k.seg[4];
    const uint32_t n_sym = k.aux32 < 4096u ? k.aux32 : 4096u;
    const uint8_t *len = k.seg[4] + 8ull * k.aux32, *comp = k.seg[5];
    uint8_t *pay = k.pay + k.pay_base;
    for (uint32_t i = i0; i < i1; i += step) {
      const uint32_t c0 = i ? tile[fls_tile_slot<uint32_t>(i - 1u)] : cv0, c1 = tile[fls_tile_slot<uint32_t>(i)];
      uint64_t d0 = (i ? fls_ldcg(offs + i) - k.pay_base : dv0), d1 = fls_ldcg(offs + i + 1u) - k.pay_base;
      if (c1 < c0 || c1 > k.aux64[1] || d1 < d0 || d1 > k.aux64[0]) continue;  // damaged offsets: nothing is written
      const uint32_t n_codes = 2u * (c1 - c0) / 3u;
      const uint8_t *in = comp + c0;
      for (uint32_t q = 0; q < n_codes && d0 < d1; ++q) {
        const uint8_t *p = in + 3u * (q >> 1);
        uint32_t c = (q & 1u) ? ((uint32_t)fls_ldg(p + 1) >> 4) | ((uint32_t)fls_ldg(p + 2) << 4)
                              : (uint32_t)fls_ldg(p) | (((uint32_t)fls_ldg(p + 1) & 15u) << 8);
        if (c >= n_sym) c = 0;
        uint64_t sv = fls_ldg(sym + c);
        for (uint32_t b = fls_min(8u, (uint32_t)fls_ldg(len + c)); b && d0 < d1; --b) { pay[d0++] = (uint8_t)sv; sv >>= 8; }
      }
    }
  }
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, FlsWarp &w, const uint8_t *) {
    const fls_vechdr hd = fls_load_hdr(k, 0, v), hc = fls_load_hdr(k, 2, v);
    uint32_t bwd = FLS_META_BW(hd.meta), bwc = FLS_META_BW(hc.meta);
    if (bwd > 32u) bwd = 32u;
    if (bwc > 32u) bwc = 32u;
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC;  // offs[i] = start of string i, offs[i + 1] its end
    const uint32_t *tile = (const uint32_t *)(w.sm + CW);
    DecDELTA<uint32_t>::to_tile(w, k.seg[1] + 16ull * hd.ep, bwd, hd.base);
    const uint32_t dv0 = ((const uint32_t *)w.sm)[0];  // lane 0's base = the vector's first decoded start
    FLS_LANES(t) { fls_tile_copy_out<uint32_t>(w.sm + CW, offs + 1, t, k.pay_base); }
    DecDELTA<uint32_t>::to_tile(w, k.seg[3] + 16ull * hc.ep, bwc, hc.base);
    const uint32_t cv0 = ((const uint32_t *)w.sm)[0];
    const uint64_t *sym = (const uint64_t *)k.seg[4];
    const uint32_t n_sym = k.aux32 < 4096u ? k.aux32 : 4096u;
    const uint8_t *len = k.seg[4] + 8ull * k.aux32, *comp = k.seg[5];
    uint8_t *pay = k.pay + k.pay_base;
    for (uint32_t qt = 0; qt < 4u; ++qt) {
      const uint32_t sq = 256u * qt;  // the quarter's first string
      const uint32_t cq0 = sq ? tile[fls_tile_slot<uint32_t>(sq - 1u)] : cv0, cq1 = tile[fls_tile_slot<uint32_t>(sq + 255u)];
      // decoded extent of the quarter: what phase 1 has just written (read back through L2)
      const uint32_t dq0 = sq ? fls_ldcg(offs + sq) - k.pay_base : dv0, dq1 = fls_ldcg(offs + sq + 256u) - k.pay_base;
      const uint32_t a0 = cq0 & ~15u, a1 = (cq1 + 15u) & ~15u;  // (the segment is padded to 16 bytes)
      uint8_t *g = pay + dq0;
      const uint32_t hh = (uint32_t)((uintptr_t)g & 15u), qlen = dq1 - dq0;
      const bool fits = cq0 <= cq1 && cq1 <= k.aux64[1] && a1 - a0 <= CW && dq0 <= dq1 && dq1 <= k.aux64[0] && hh + qlen <= STG;
      if (!fits) {
        FLS_LANES(t) { plain(k, tile, offs, cv0, dv0, sq + t, sq + 256u, 32u); }
        continue;
      }
      w.store_wait();  // the previous quarter's bulk copy has read the staging buffer
      if (a1 > a0) { w.fence_async(); w.stage(0, comp + a0, a1 - a0); w.wait(); }
      FLS_VAR(FlsPush, ps);
      FLS_LANES(t) {
        const uint32_t s0 = sq + 8u * t;
        uint32_t dd[9];  // the decoded boundaries of the lane's 8 strings
#pragma unroll
        for (uint32_t j = 0; j < 9u; ++j) dd[j] = s0 + j ? fls_ldcg(offs + s0 + j) - k.pay_base : dv0;
        bool ok = dd[0] >= dq0 && dd[8] <= dq1;
#pragma unroll
        for (uint32_t j = 0; j < 8u; ++j) ok = ok && dd[j] <= dd[j + 1u];
        FlsPush p;
        p.init(w.sm + STG_OFF, hh + (ok ? dd[0] - dq0 : 0u));
        for (uint32_t j = 0; j < 8u; ++j) {
          const uint32_t i = s0 + j;
          const uint32_t c0 = i ? tile[fls_tile_slot<uint32_t>(i - 1u)] : cv0, c1 = tile[fls_tile_slot<uint32_t>(i)];
          uint32_t left = ok ? dd[j + 1u] - dd[j] : 0u;  // damaged boundaries: the lane writes nothing
          if (c0 <= c1 && c0 >= cq0 && c1 <= cq1) {
            const uint8_t *in = w.sm + (c0 - a0);
            const uint32_t n_codes = 2u * (c1 - c0) / 3u;
            uint32_t q = 0;
            for (; q + 1u < n_codes && left; q += 2u) {  // two codes = three bytes; four table loads in flight
              const uint8_t *pp = in + 3u * (q >> 1);
              const uint32_t b1 = pp[1];
              uint32_t ca = (uint32_t)pp[0] | ((b1 & 15u) << 8), cb = (b1 >> 4) | ((uint32_t)pp[2] << 4);
              if (ca >= n_sym) ca = 0;
              if (cb >= n_sym) cb = 0;
              const uint64_t sa = fls_ldg(sym + ca), sb = fls_ldg(sym + cb);
              const uint32_t la = fls_min(fls_min(8u, (uint32_t)fls_ldg(len + ca)), left);
              const uint32_t lb = fls_min(fls_min(8u, (uint32_t)fls_ldg(len + cb)), left - la);
              p.push(fls_piece_entry(sa, la));
              p.push(fls_piece_entry(sb, lb));
              left -= la + lb;
            }
            if (q < n_codes && left) {  // an odd last code
              uint32_t c = code_at(in, q);
              if (c >= n_sym) c = 0;
              const uint32_t L = fls_min(fls_min(8u, (uint32_t)fls_ldg(len + c)), left);
              p.push(fls_piece_entry(fls_ldg(sym + c), L));
              left -= L;
            }
          }
          while (left) {  // codes that do not add up to the declared length: zeros, the run keeps its layout
            const uint32_t L = fls_min(8u, left);
            p.push(fls_piece_entry(0ull, L));
            left -= L;
          }
        }
        FLS_AT(ps) = p;
      }
      FLS_LANES(t) { FLS_AT(ps).finish(); }
      fls_stage_store_at(w, STG_OFF, g, hh, qlen);
    }
    w.store_wait();
  }
};

// ===================================================================== FSST strings
// chunk: seg[0]/[1] = decoded end-offset stream, seg[2]/[3] = compressed end-offset stream (both DELTA u32),
//        seg[4] = symbol table (u64 symbol[256] | u8 len[256]), seg[5] = compressed bytes (+16 slack).
// The strings of a vector are contiguous in both the compressed and the decoded space.  The vector's compressed
// bytes are bulk-copied into the warp's window; a vector is decoded in four quarters of 256 strings, lane t
// owning the 8 consecutive strings 256q + 8t .. +7 = one compressed run -> one output run.  A lane reads its run
// 16 bytes at a time; every code byte costs one conflict-free 128-bit table load (the block keeps 8 copies of the
// table, one per 16-byte bank group, lane t uses copy t & 7) and one FlsPush::push.  An escaped literal costs the
// same: its entry sits in a table of its own that the marker flag selects (no branch, no select).
#ifndef FLS_FSST_WARPS
#define FLS_FSST_WARPS 10
#define FLS_FSST_COPIES 8
#define FLS_FSST_STG 7424
#define FLS_FSST_CWIN 3456
#endif
struct VecSTRFSST {
  // The kernel is bound by instruction issue and latency, not by shared-memory bandwidth (profiles/r01_SUMMARY.md,
  // knock-out runs): shared memory is spent on WARPS, not on table copies or large passes.
  static constexpr uint32_t WARPS = FLS_FSST_WARPS;
  // A vector is decoded in passes of 256 strings — or 128 / 64 / 32 when a pass of the larger size would not fit:
  static constexpr uint32_t STG = FLS_FSST_STG;      // staging of one pass's output
  static constexpr uint32_t CWIN = FLS_FSST_CWIN;    // compressed bytes of one pass
  static constexpr uint32_t E16N = 1056u;    // 16-bit relative compressed end of every string, padded original order
  static constexpr uint32_t P1 = 2176u;      // phase 1: each staged offset stream (lane bases + deltas up to 16 bits wide)
  static constexpr uint32_t SMEM = STG + CWIN + 2u * E16N + 320u;  // + DS[40] + LS[40]
  // Three arrays of 256 entries x COPIES x 16 B (entry c of copy r at (c * COPIES + r) * 16; a lane reads copy
  // t % COPIES — 8 copies: the eight lanes of a 128-bit load phase never share a bank group):
  //   [symbols of table A | escaped literals (entry c = the byte c; the same for every chunk) | symbols of table B]
  // A code byte c that is (is not) an escaped literal reads entry (lit << 8 | c) of the 64 KB that start at table
  // A — or, for table B, that start at the literals with the `lit` bit flipped.  Two symbol tables: the next
  // chunk's is built while this one is in use, so nobody waits at a chunk boundary (with ONE 64 KB table holding
  // symbols and literals 13 % of the warps' time was the spin on ready[], profiles/r02_SUMMARY.md).
  static constexpr uint32_t COPIES = FLS_FSST_COPIES, SH = 7u;
  static constexpr uint32_t BLOCK_SMEM = 256u * 16u * COPIES;  // one array
#ifndef FLS_FSST_NBUF
#define FLS_FSST_NBUF 2  /* 1: one symbol table per block (64 KB with the literals), warps spin at chunk boundaries */
#endif
  static constexpr uint32_t NBUF = FLS_FSST_NBUF, CONST_SMEM = 0, TABLES_SMEM = (NBUF + 1u) * BLOCK_SMEM;
  FLS_MEM static constexpr uint32_t tab_off(uint32_t b) { return 2u * b * BLOCK_SMEM; }
  static constexpr bool PIPELINED = true;
  static_assert(COPIES == 8u, "the entry stride is 128 bytes");
  static_assert(2u * P1 <= STG + CWIN && 4224u <= STG + CWIN, "phase-1 buffers must fit the staging buffer + window");
  static_assert(SMEM % 128u == 0 && STG % 128u == 0, "warp slots stay 128-byte aligned");
  // the literals (called once per block, by all its threads; `tabs` = the start of the three arrays)
  FLS_MEM static void const_init(uint8_t *tabs, uint32_t tid, uint32_t nthreads) {
    for (uint32_t c = tid; c < 256u; c += nthreads) {
      const fls_u4 el = fls_piece_entry(c, 1u);
      for (uint32_t r = 0; r < COPIES; ++r) *(fls_u4 *)(tabs + BLOCK_SMEM + ((c * COPIES + r) << 4)) = el;
    }
  }
  // (all of a thread's symbol loads are issued before the first table store: one memory round trip per build)
  FLS_MEM static void block_init(const FlsChunkDev &k, uint8_t *bsm, uint32_t tid, uint32_t nthreads) {
    const uint64_t *sym = (const uint64_t *)k.seg[4];
    const uint8_t *len = k.seg[4] + 2048;
    constexpr uint32_t B = 8;
    for (uint32_t c0 = 0; c0 < 256u; c0 += B * nthreads) {
      uint64_t sv[B];
      uint32_t lv[B];
#pragma unroll
      for (uint32_t j = 0; j < B; ++j) {
        const uint32_t c = c0 + j * nthreads + tid;
        sv[j] = c < 255u ? fls_ldg(sym + c) : 0ull;
        lv[j] = c < 255u ? fls_min(8u, (uint32_t)fls_ldg(len + c)) : 0u;  // 255, the escape marker, produces nothing
      }
#pragma unroll
      for (uint32_t j = 0; j < B; ++j) {
        const uint32_t c = c0 + j * nthreads + tid;
        if (c < 256u) {
          const fls_u4 en = fls_piece_entry(sv[j], lv[j]);
          for (uint32_t r = 0; r < COPIES; ++r) *(fls_u4 *)(bsm + ((c * COPIES + r) << 4)) = en;
        }
      }
    }
  }
  // 0x80 in every byte of x that equals 0xFF (exact, byte by byte)
  FLS_MEM static uint32_t ff_bytes(uint32_t x) {
    const uint32_t y = ~x;  // zero byte <=> 0xFF
    return ~(((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y) & 0x80808080u;
  }
  // NW/2 x eight consecutive compressed bytes (32-bit words w[]) of a lane's run: all table loads in flight, then
  // all pushes, straight-line.  Byte j is a literal when byte j-1 is an escape marker (0xFF), `lit` standing for
  // byte -1; that is exact unless two 0xFF bytes are adjacent (a literal 0xFF), and such a window takes the
  // byte-by-byte path.  The marker flags shifted up by one bit are the literal flags, each in bit 0 of its own byte;
  // one byte permute puts {byte j, flags of byte j} side by side, which IS the 9-bit table index of byte j.
  // MASK: only bytes [s, e) of the window belong to the run (partial first / last window); the others are turned
  // into 0xFF bytes that are no markers — code 255 is the symbol tables' empty entry.
  // tabr = the lane's copy inside the 64 KB the index addresses, flip = 0 (table A) or 1 << 15 (table B).
  template <uint32_t NW, bool MASK>
  FLS_MEM static void window(FlsPush &p, const uint8_t *tabr, uint32_t flip, uint32_t (&w)[NW], uint32_t s, uint32_t e, bool &lit) {
    uint32_t ff[NW];
#pragma unroll
    for (uint32_t k = 0; k < NW; ++k) ff[k] = ff_bytes(w[k]);
    if (MASK) {
      const uint32_t s8 = 8u * s, e8 = 8u * e;
#pragma unroll
      for (uint32_t k = 0; k < NW; ++k) {
        // bytes [s, e) of the window that fall into word k: (all ones << bits below s) & ~(all ones << bits below e),
        // the shifts clamped at 32
        const uint32_t vm = fls_shl_c(~0u, fls_max(s8, 32u * k) - 32u * k) & ~fls_shl_c(~0u, fls_max(e8, 32u * k) - 32u * k);
        ff[k] &= vm;
        w[k] |= ~vm;
      }
    }
    const uint32_t c31 = lit ? 0x80000000u : 0u;  // "the byte before the window is a marker", where a word's last byte has it
    uint32_t lf[NW];  // literal flags: bit 8j = byte j of the word is a literal
#pragma unroll
    for (uint32_t k = 0; k < NW; ++k) lf[k] = fls_shf_l(k ? ff[k - 1u] : c31, ff[k], 1u);
    uint32_t adj = 0;
#pragma unroll
    for (uint32_t k = 0; k < NW; ++k) adj |= (lf[k] << 7) & ff[k];  // a marker that is itself a literal
    if (adj) {
      // adjacent 0xFF bytes: byte by byte
#pragma unroll
      for (uint32_t j = 0; j < 4u * NW; ++j) {  // (constant bounds: w[] stays in registers)
        if (MASK && (j < s || j >= e)) continue;
        const uint32_t c = (w[j >> 2] >> (8u * (j & 3u))) & 0xFFu;
        if (lit) { p.push_byte(c); lit = false; }
        else if (c == 255u) lit = true;
        else p.push(fls_lds128<fls_u4>(tabr + ((c << SH) ^ flip)));
      }
      return;
    }
    lit = (ff[NW - 1u] >> 31) != 0u;
    constexpr uint32_t IM = 0x1FFu << SH;  // the 9 index bits at the entry stride
    fls_u4 en[4u * NW];
#pragma unroll
    for (uint32_t k = 0; k < NW; ++k) {
      const uint32_t x01 = fls_prmt(w[k], lf[k], 0x5140u), x23 = fls_prmt(w[k], lf[k], 0x7362u);  // {b0, f0, b1, f1}, {b2, f2, b3, f3}
      en[4u * k] = fls_lds128<fls_u4>(tabr + (((x01 << SH) & IM) ^ flip));
      en[4u * k + 1u] = fls_lds128<fls_u4>(tabr + (((x01 >> (16u - SH)) & IM) ^ flip));
      en[4u * k + 2u] = fls_lds128<fls_u4>(tabr + (((x23 << SH) & IM) ^ flip));
      en[4u * k + 3u] = fls_lds128<fls_u4>(tabr + (((x23 >> (16u - SH)) & IM) ^ flip));
    }
    p.push_all(en);
  }
  // A lane's compressed run = bytes [a, b) of the staged window `cwin`, read 16 bytes at a time (8 at the end).
  FLS_MEM static void decode_run(FlsPush &p, const uint8_t *tabr, uint32_t flip, const uint8_t *cwin, uint32_t a, uint32_t b) {
    const uint32_t *cw = (const uint32_t *)cwin;
    bool lit = false;
    uint32_t i = a >> 3;  // 8-byte unit
    while (8u * i < b) {
      uint32_t w[4];
      w[0] = cw[2u * i]; w[1] = cw[2u * i + 1u];
      const uint32_t s = 8u * i < a ? a - 8u * i : 0u, left = b - 8u * i;
      // the variant is chosen by a vote of the lanes still in the loop, so they stay converged
      if (FLS_ANY(left > 8u)) {
        w[2] = cw[2u * i + 2u]; w[3] = cw[2u * i + 3u];  // the window has 16 bytes of slack
        if (FLS_ANY(s != 0u || left < 16u)) window<4, true>(p, tabr, flip, w, s, fls_min(16u, left), lit);
        else window<4, false>(p, tabr, flip, w, 0u, 16u, lit);
        i += 2u;
      } else {
        uint32_t w2[2] = {w[0], w[1]};
        window<2, true>(p, tabr, flip, w2, s, fls_min(8u, left), lit);
        i += 1u;
      }
    }
  }
  // Slow path (compressed bytes larger than the window, output larger than the staging buffer, offsets that do
  // not fit 16 bits): compressed bytes [c0, c1) -> o, byte-wise, straight from / to global memory.
  // Bounded on both sides: reads stay below the chunk's compressed size (aux64[1]), writes below `o_end`.
  FLS_MEM static void decode_bytes(const FlsChunkDev &k, uint32_t c0, uint32_t c1, uint8_t *o, const uint8_t *o_end) {
    const uint8_t *comp = k.seg[5];
    const uint64_t *gsym = (const uint64_t *)k.seg[4];
    const uint8_t *glen = k.seg[4] + 2048;
    c1 = (uint32_t)(c1 < k.aux64[1] ? c1 : k.aux64[1]);
    for (uint32_t i = c0; i < c1;) {
      const uint32_t c = fls_ldg(comp + i++);
      if (c == 255u) { if (i < c1 && o < o_end) *o++ = fls_ldg(comp + i++); else break; }
      else { uint64_t s = fls_ldg(gsym + c); for (uint32_t b = fls_min(8u, (uint32_t)fls_ldg(glen + c)); b && o < o_end; --b) { *o++ = (uint8_t)s; s >>= 8; } }
    }
  }
  // value (row, lane) of a packed u32 vector, 1 <= W <= 32: two loads and one funnel shift, no branch
  FLS_MEM static uint32_t extract32(const uint32_t *pk, uint32_t W, uint32_t row, uint32_t lane) {
    const uint32_t bit = row * W, k = bit >> 5;
    const uint32_t lo = pk[k * 32u + lane], hi = pk[fls_min(k + 1u, W - 1u) * 32u + lane];
    return fls_shf_r(lo, hi, bit) & (W >= 32u ? 0xFFFFFFFFu : ((1u << W) - 1u));
  }
  // DELTA u32 chain of lane t over a staged end-offset stream.  Writes {lane base, end 0 .. end 30} + add to the
  // 32 aligned slots line[0..31] (slot i = start of string i; end 31 is the next lane's base) and the 32 ends,
  // relative to `rel0`, as 16-bit values into e16 (padded original order).  Returns the lane's last end.
  template <bool LINE, bool E16>
  FLS_MEM static uint32_t walk_ends(const uint8_t *stage, uint32_t bw, uint32_t base, uint32_t t, uint32_t add, uint32_t rel0,
                                    uint32_t *line, uint16_t *e16) {
    const uint32_t *bases = (const uint32_t *)stage, *pk = (const uint32_t *)(stage + 128);
    uint32_t acc = bases[t];
    uint32_t *e16w = E16 ? (uint32_t *)(e16 + slot16(fls_lane_start(t))) : nullptr;
    // two groups of four steps per loop iteration: fully unrolled (8) the two walks are 12 KB of straight-line code
    // that the ten warps of a block, each in another phase, keep evicting from the instruction cache
    // (1 / 2 / 8 groups: 1.285 / 1.273 / 1.291 ms on the SF10 launch)
#ifndef FLS_WALK_UNROLL
#define FLS_WALK_UNROLL 2
#endif
    constexpr int WU = FLS_WALK_UNROLL;
#pragma unroll WU
    for (uint32_t r4 = 0; r4 < 32u; r4 += 4u) {
      uint32_t x[5];
      x[0] = acc;
#pragma unroll
      for (uint32_t j = 0; j < 4u; ++j)
        x[j + 1u] = x[j] + (bw ? extract32(pk, bw, fls_chain_row<32>(r4 + j), t) : 0u) + base;
      acc = x[4];
      if (LINE) {
        fls_u4 o;
        o.x = x[0] + add; o.y = x[1] + add; o.z = x[2] + add; o.w = x[3] + add;
        fls_st128(line + r4, o);
      }
      if (E16) {
        e16w[r4 / 2u] = ((x[1] - rel0) & 0xFFFFu) | ((x[2] - rel0) << 16);
        e16w[r4 / 2u + 1u] = ((x[3] - rel0) & 0xFFFFu) | ((x[4] - rel0) << 16);
      }
    }
    return acc;
  }
  // padded position of string e in the 16-bit end arrays: the 32 lanes of a walk write distinct banks
  FLS_MEM static uint32_t slot16(uint32_t e) { return e + ((e >> 6) << 1); }
  // start offset (relative) of string s: the end of string s - 1
  FLS_MEM static uint32_t start16(const uint16_t *e16, uint32_t s) { return s ? (uint32_t)e16[slot16(s - 1u)] : 0u; }
  FLS_MEM static void run(const FlsChunkDev &k, uint32_t v, uint32_t /*v_next*/, FlsWarp &w, const uint8_t *tab, const uint8_t *tabs) {
    // table A: the 64 KB [symbols A | literals]; table B: [literals | symbols B] with the literal bit flipped
    const uint32_t flip = tab == tabs ? 0u : 1u << 15;
    const uint8_t *tab64 = flip ? tabs + BLOCK_SMEM : tabs;
    const fls_vechdr hd = fls_load_hdr(k, 0, v), hc = fls_load_hdr(k, 2, v);
    uint32_t bwd = FLS_META_BW(hd.meta), bwc = FLS_META_BW(hc.meta);
    if (bwd > 32u) bwd = 32u;
    if (bwc > 32u) bwc = 32u;
    uint8_t *stg = w.sm;
    // Compressed ends: 16 bits each, all 1024 (the lanes' run boundaries are searched in them).  Decoded starts: only
    // of every 32nd string (DS[m] = start of string 32 m, relative to the vector's; every pass boundary is one); the
    // start of a lane's first string is read back from offsets[], which phase 1 has just written (an L2 hit).
    uint16_t *CE = (uint16_t *)(w.sm + STG + CWIN);
    uint32_t *DS = (uint32_t *)(CE + E16N);  // DS[0..32]
    uint32_t *LS = DS + 40;                  // LS[0..32] = first string of each lane, LS[34..35] = vector extents
    const uint8_t *comp = k.seg[5];
    uint8_t *pay = k.pay + k.pay_base;
    // ---- phase 1: both end-offset streams -> offsets[] (global) and 16-bit relative ends of every string (shared).
    //      Streams up to 16 bits wide are staged side by side; wider ones (strings of 64 KB and more, which end in
    //      the byte-wise path below anyway) one after the other.
    const bool both = bwd <= 16u && bwc <= 16u;
    w.store_wait();  // the previous vector's last bulk copy has read the staging buffer
    w.fence_async();
    w.stage(0, k.seg[1] + 16ull * hd.ep, 128u + 128u * bwd);
    if (both) w.stage(P1, k.seg[3] + 16ull * hc.ep, 128u + 128u * bwc);
    w.wait();
    uint32_t *offs = (uint32_t *)k.out + (size_t)v * FLS_VEC;  // offs[i] = start of string i
    const uint32_t dv0 = ((const uint32_t *)w.sm)[0];          // lane 0's base
    FLS_VAR(uint32_t, lc0); FLS_VAR(uint32_t, lc1); FLS_VAR(uint32_t, ld0);
    FLS_LANES(t) {
      FLS_AT(ld0) = ((const uint32_t *)w.sm)[t];
      const uint32_t dend = walk_ends<true, false>(w.sm, bwd, (uint32_t)hd.base, t, k.pay_base, dv0, offs + fls_lane_start(t), nullptr);
      DS[fls_lane_start(t) >> 5] = FLS_AT(ld0) - dv0;  // the lane's 32 strings start at a multiple of 32
      if (t == 31u) { offs[FLS_VEC] = dend + k.pay_base; LS[34] = dend; DS[32] = dend - dv0; }  // lane 31 owns the last strings
    }
    if (!both) { w.fence_async(); w.stage(0, k.seg[3] + 16ull * hc.ep, 128u + 128u * bwc); w.wait(); }
    const uint8_t *cst = w.sm + (both ? P1 : 0u);
    const uint32_t cv0 = ((const uint32_t *)cst)[0];
    FLS_LANES(t) {
      FLS_AT(lc0) = ((const uint32_t *)cst)[t];
      const uint32_t cend = walk_ends<false, true>(cst, bwc, (uint32_t)hc.base, t, 0u, cv0, nullptr, CE);
      FLS_AT(lc1) = cend;
      if (t == 31u) LS[35] = cend;
    }
    const uint32_t dv1 = LS[34], cv1 = LS[35];
    const uint8_t *pay_end = pay + k.aux64[0];
    // end offsets that run backwards or past the declared sizes (corrupt streams): nothing of this vector is written
    if (dv1 < dv0 || dv1 > k.aux64[0] || cv1 < cv0 || cv1 > k.aux64[1]) return;
    if (dv1 - dv0 >= 65536u || cv1 - cv0 >= 65536u) {
      // offsets do not fit 16 bits: every lane decodes the 32 strings of its own chain, byte-wise
      FLS_LANES(t) {
        if (FLS_AT(lc0) <= FLS_AT(lc1) && FLS_AT(ld0) <= k.aux64[0]) decode_bytes(k, FLS_AT(lc0), FLS_AT(lc1), pay + FLS_AT(ld0), pay_end);
      }
      return;
    }
    const uint32_t dspan = dv1 - dv0, cspan = cv1 - cv0;
    // ---- phase 2: quarter by quarter (256 strings); the quarter's compressed bytes are split into 32 runs of
    //      (nearly) equal length that start on string boundaries, one per lane; the compressed bytes of quarter
    //      q+1 are bulk-copied into the window as soon as quarter q has been read, while q is copied out
    // pass size: the largest of 256 / 128 / 64 / 32 strings for which every pass of this vector fits the staging
    // buffer and the window (lane t checks pass t of each size; the counts of misfits add up in one scan)
    FLS_VAR(uint32_t, bad);
    FLS_LANES(t) {
      uint32_t b = 0;
#pragma unroll
      for (uint32_t lv = 0; lv < 3u; ++lv) {
        const uint32_t ps = 256u >> lv;
        if (t < FLS_VEC / ps) {
          const uint32_t d = DS[((t + 1u) * ps) >> 5] - DS[(t * ps) >> 5], c = start16(CE, (t + 1u) * ps) - start16(CE, t * ps);
          if (d + 15u > STG || c + 31u > CWIN) b |= 1u << (8u * lv);
        }
      }
      FLS_AT(bad) = b;
    }
    uint32_t misfits;
    FLS_EXCL_SCAN(bad, misfits);
    const uint32_t lg = (misfits & 0xFFu) == 0u ? 0u : (misfits & 0xFF00u) == 0u ? 1u : (misfits & 0xFF0000u) == 0u ? 2u : 3u;
    const uint32_t PASS = 256u >> lg, NPASS = 4u << lg;
    uint32_t qc0 = 0, qc1 = start16(CE, PASS);  // compressed extent of the pass, relative to cv0
    uint32_t wbase = cv0 & ~15u;
    uint32_t span = cv0 + qc1 - wbase;
    // (the 16-bit ends of a well-formed vector ascend; `qc1 <= cspan` and the per-lane checks below keep a corrupt
    // one inside the window, the staging buffer and the row group's payload)
    bool windowed = span + 16u <= CWIN && qc1 <= cspan;
    const uint8_t *cwin = w.sm + STG;
    if (windowed && span) { w.fence_async(); w.stage(STG, comp + wbase, (span + 15u) & ~15u); }
    for (uint32_t qt = 0; qt < NPASS; ++qt) {
      if (windowed && span) w.wait();
      const uint32_t cur_wbase = wbase, cur_c0 = qc0, cur_c1 = qc1;
      const bool cur_windowed = windowed;
      const uint32_t S0 = PASS * qt, S1 = S0 + PASS;
      const uint32_t qd0 = DS[S0 >> 5], qlen = DS[S1 >> 5] - qd0;
      uint8_t *g = pay + dv0 + qd0;
      const uint32_t hh = (uint32_t)((uintptr_t)g & 15u);
      // first string of lane t: the first one that starts at or after t/32 of the quarter's compressed bytes
      FLS_LANES(t) {
        const uint32_t target = cur_c0 + (((cur_c1 - cur_c0) * t) >> 5);
        uint32_t lo = S0, hi = S1;
#ifndef FLS_BS_UNROLL
#define FLS_BS_UNROLL 9
#endif
        constexpr int BU = FLS_BS_UNROLL;
#pragma unroll BU
        for (uint32_t it = 0; it < 9u; ++it) {
          const uint32_t mid = (lo + hi) >> 1;
          const bool below = lo < hi && start16(CE, mid) < target;
          if (lo < hi) { if (below) lo = mid + 1u; else hi = mid; }
        }
        LS[t] = lo;
        if (t == 0) LS[32] = S1;
      }
      if (qd0 > dspan || qlen > dspan - qd0 || cur_c0 > cur_c1 || cur_c1 > cspan) return;  // corrupt end offsets
      const bool staged = cur_windowed && hh + qlen <= STG;
      FLS_VAR(FlsPush, ps);
      w.store_wait();  // the previous pass's bulk copy has read the staging buffer
      FLS_LANES(t) {
        const uint32_t sa = LS[t], sb = LS[t + 1u];
        uint32_t c0 = start16(CE, sa), c1 = start16(CE, sb);
        uint32_t d0 = fls_ldcg(offs + fls_min(sa, FLS_VEC)) - k.pay_base - dv0;
        // a lane's run lies inside the pass: always true for ascending ends, two compares otherwise
        c0 = fls_max(c0, cur_c0); c1 = fls_min(c1, cur_c1);
        d0 = fls_min(fls_max(d0, qd0), qd0 + qlen);
        if (staged) {
          FlsPush p;
          p.init(stg, hh + (d0 - qd0));
          if (c1 > c0) decode_run(p, tab64 + 16u * (t & (COPIES - 1u)), flip, cwin, cv0 + c0 - cur_wbase, cv0 + c1 - cur_wbase);
          FLS_AT(ps) = p;
        } else if (c1 > c0) {
          decode_bytes(k, cv0 + c0, cv0 + c1, pay + dv0 + d0, pay + dv0 + qd0 + qlen);
        }
      }
      if (qt + 1u < NPASS) {  // next pass's window
        qc0 = qc1;
        qc1 = start16(CE, PASS * qt + 2u * PASS);
        wbase = (cv0 + qc0) & ~15u;
        span = cv0 + qc1 - wbase;
        windowed = span + 16u <= CWIN && qc0 <= qc1 && qc1 <= cspan;
        if (windowed && span) { w.fence_async(); w.stage(STG, comp + wbase, (span + 15u) & ~15u); }
      }
      if (staged) {
        FLS_LANES(t) { FLS_AT(ps).finish(); }
        fls_stage_store(w, g, hh, qlen);
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
  const uint8_t *o_end = (uint8_t *)k.out + fls_ldg(doff + d + 1);  // an entry never writes past its own slot
  for (uint32_t i = fls_ldg(coff + d), e = fls_ldg(coff + d + 1); i < e && o < o_end;) {
    const uint32_t code = fls_ldg(in + i++);
    if (code == 255u) { *o++ = fls_ldg(in + i++); }
    else { uint64_t s = fls_ldg(sym + code); for (uint32_t b = fls_min(8u, (uint32_t)fls_ldg(len + code)); b && o < o_end; --b) { *o++ = (uint8_t)s; s >>= 8; } }
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
