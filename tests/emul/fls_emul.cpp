// fls_emul.cpp — HOST EMULATION of the CUDA per-vector decoders.  TEST INFRASTRUCTURE ONLY.
//
// Compiles duckdb-fastlane_b200/csrc/fls_decode.cuh / fls_kernels.cuh / fls_strings.cuh with FLS_HOST_EMUL so
// that the exact device source (same templates, same index arithmetic, same phases) runs as a 32-step loop per
// warp on the CPU.  Lets the GPU-less container check kernel logic against the oracle before GPU time is
// spent.  The product library never contains this path: libflsgpu.so is built without FLS_HOST_EMUL and has
// no CPU decode.
#define FLS_HOST_EMUL 1
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

#include "fls_plan.h"
#include "fls_strings.cuh"
#include "fls_validate.h"

static std::string g_err;
int fls_emul_force_any = 0;

template <class V, class = void> struct is_pipelined { static constexpr bool value = false; };
template <class V> struct is_pipelined<V, std::enable_if_t<V::PIPELINED>> { static constexpr bool value = true; };

template <class V> static void run_vectors(const FlsLaunch &l) {
  if constexpr (is_pipelined<V>::value) {
    // chunk kernels: [warp slot | constant area | tables | ...] as in fls_chunk_kernel; consecutive chunks alternate
    // between the table buffers, one warp decodes a chunk's vectors in order, each call prefetching the next vector
    std::vector<uint8_t> smem(V::SMEM + V::CONST_SMEM + V::TABLES_SMEM + 256);
    uint8_t *slot = (uint8_t *)(((uintptr_t)smem.data() + 127) & ~(uintptr_t)127);
    uint8_t *csm = slot + V::SMEM, *tabs = csm + V::CONST_SMEM;
    V::const_init(csm, 0, 1);
    uint32_t b = 0;
    for (const FlsChunkDev &k : l.chunks) {
      uint8_t *bsm = tabs + V::tab_off(b);
      b = (b + 1) % V::NBUF;
      V::block_init(k, bsm, 0, 1);
      FlsWarp w;
      w.init(slot, nullptr);
      for (uint32_t v = 0; v < k.n_vec; ++v) V::run(k, v, v + 1 < k.n_vec ? v + 1 : FLS_NO_VECTOR, w, bsm, csm);
    }
  } else {
    std::vector<uint8_t> smem(V::SMEM + V::BLOCK_SMEM + 256);
    uint8_t *slot = (uint8_t *)(((uintptr_t)smem.data() + 127) & ~(uintptr_t)127);
    uint8_t *bsm = slot + V::SMEM;
    for (const FlsChunkDev &k : l.chunks) {
      if constexpr (V::BLOCK_SMEM != 0) V::block_init(k, bsm, 0, 1);
      for (uint32_t v = 0; v < k.n_vec; ++v) {
        FlsWarp w;
        w.init(slot, nullptr);
        V::run(k, v, w, bsm);
      }
    }
  }
}
template <template <typename> class V> static bool run_by_width(const FlsLaunch &l) {
  switch (l.width) {
    case 1: run_vectors<V<uint8_t>>(l); return true;
    case 2: run_vectors<V<uint16_t>>(l); return true;
    case 4: run_vectors<V<uint32_t>>(l); return true;
    case 8: run_vectors<V<uint64_t>>(l); return true;
  }
  return false;
}

static bool run_launch(const FlsLaunch &l) {
  switch (l.kind) {
    case FLS_K_FFOR: return run_by_width<VecFFOR>(l);
    case FLS_K_DELTA: return run_by_width<VecDELTA>(l);
    case FLS_K_RLE: return run_by_width<VecRLE>(l);
    case FLS_K_DICT: return run_by_width<VecDICT>(l);
    case FLS_K_FFOR_PATCH: return run_by_width<VecFFORPATCH>(l);
    case FLS_K_CROSS_RLE: return run_by_width<VecCROSSRLE>(l);
    case FLS_K_ALP:
      if (l.width == 4) run_vectors<VecALP32>(l); else run_vectors<VecALP>(l);
      return true;
    case FLS_K_ALPRD: run_vectors<VecALPRD>(l); return true;
    case FLS_K_COPY:
      for (auto &k : l.chunks) memcpy(k.out, k.seg[0], k.aux64[1]);
      return true;
    case FLS_K_FILL:
      for (auto &k : l.chunks)
        for (uint64_t i = 0; i < (uint64_t)k.n_vec * FLS_VEC; ++i) memcpy((uint8_t *)k.out + i * k.aux32, &k.aux64[0], k.aux32);
      return true;
    case FLS_K_STR_OFFSETS: run_vectors<VecSTROFFS>(l); return true;
    case FLS_K_STR_COPY:
      for (auto &k : l.chunks) fls_str_copy_chunk_host(k);
      return true;
    case FLS_K_STR_CONST: run_vectors<VecSTRCONST>(l); return true;
    case FLS_K_STR_FSST12: run_vectors<VecSTRFSST12>(l); return true;
    case FLS_K_STR_FLAGS: run_vectors<VecSTRFLAGS>(l); return true;
    case FLS_K_STR_DICT: run_vectors<VecSTRDICT>(l); return true;
    case FLS_K_STR_DICT_LONG: run_vectors<VecSTRDICTL>(l); return true;
    case FLS_K_STR_FSST: run_vectors<VecSTRFSST>(l); return true;
    case FLS_K_STR_FSST_DICTPREP:
      for (auto &k : l.chunks) fls_fsst_dictprep_chunk_host(k);
      return true;
    case FLS_K_VALIDITY:
      for (auto &k : l.chunks) {
        if (k.seg[0]) memcpy(k.out, k.seg[0], 128ull * k.n_vec);
        else memset(k.out, 0xFF, 128ull * k.n_vec);
      }
      return true;
  }
  return false;
}

extern "C" {
const char *emul_error() { return g_err.c_str(); }
void emul_force_any(int on) { fls_emul_force_any = on; }

// Decode column `col` of row groups [g0,g1) of an in-memory image with the emulated kernels.
// Buffers sized by the caller exactly like libflsgpu's (see fls_plan.h: fls_column_need).
int emul_decode(const uint8_t *image, uint64_t size, uint32_t col, uint32_t g0, uint32_t g1, void *data,
                uint32_t *offsets, uint8_t *payload) {
  FlsTable t;
  g_err = fls_parse(image, size, t);
  if (!g_err.empty()) return 1;
  if (col >= t.n_cols || g0 > g1 || g1 > t.n_rg) { g_err = "bad range"; return 1; }
  for (uint32_t g = g0; g < g1; ++g) {  // the same gate libflsgpu applies before a descriptor reaches a kernel
    g_err = fls_validate_chunk(t, g, col, image, size);
    if (!g_err.empty()) return 2;
  }
  FlsColNeed need = fls_column_need(t, col, g0, g1, image);
  std::vector<uint8_t> scratch(need.scratch_bytes + 16);
  FlsColOut out;
  out.data = data; out.offsets = offsets; out.payload = payload; out.scratch = scratch.data();
  std::vector<FlsLaunch> ls;
  if (g1 > g0) g_err = fls_plan_column(t, col, g0, g1, image + t.rgs[g0].off, image, out, ls);
  if (!g_err.empty()) return 1;
  if (offsets) offsets[0] = 0;
  for (int pass = 0; pass < 2; ++pass)
    for (auto &l : ls) {
      if ((l.kind == FLS_K_STR_FSST_DICTPREP) != (pass == 0)) continue;
      if (!run_launch(l)) { g_err = "emul: unsupported launch kind"; return 1; }
    }
  return 0;
}
uint64_t emul_need(const uint8_t *image, uint64_t size, uint32_t col, uint32_t g0, uint32_t g1, uint64_t *out4) {
  FlsTable t;
  g_err = fls_parse(image, size, t);
  if (!g_err.empty()) return 1;
  if (col >= t.n_cols || g0 > g1 || g1 > t.n_rg) { g_err = "bad range"; return 1; }
  for (uint32_t g = g0; g < g1; ++g) {
    g_err = fls_validate_chunk(t, g, col, image, size);
    if (!g_err.empty()) return 2;
  }
  FlsColNeed n = fls_column_need(t, col, g0, g1, image);
  out4[0] = n.data_bytes; out4[1] = n.offsets_bytes; out4[2] = n.payload_bytes; out4[3] = n.scratch_bytes;
  return 0;
}
}
