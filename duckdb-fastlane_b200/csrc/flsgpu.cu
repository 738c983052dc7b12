// flsgpu.cu — libflsgpu.so: the C ABI of include/flsgpu.h over the sm_100a decode kernels.
//
// Replaces, for the reference's read path, FastLanesFacade::openFile + RowgroupReader::materialize()
// (/root/reference/src/fastlanes_facade.cpp:26-67) and the per-value copy-out of readNextChunk (:69-200):
// open parses the footer only; flsgpu_decode stages the compressed row groups of a range to HBM with pinned
// cudaMemcpyAsync (pipelined in slices: H2D | decode | D2H on three streams) and runs one kernel launch per
// (scheme, width) family per slice.  No CPU decode exists in this library.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/flsgpu.h"
#include "fls_plan.h"
#include "fls_strings.cuh"
#include "fls_validate.h"
#include "fls_filter.cuh"

// ---------------------------------------------------------------------------------------------------------
// kernel table
// ---------------------------------------------------------------------------------------------------------
namespace {

template <class V> cudaError_t vec_attr() {
  return cudaFuncSetAttribute(fls_vec_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fls_block_smem<V>());
}
template <class V> void vec_launch(const FlsChunkDev *d, uint32_t n, cudaStream_t s) {
  fls_vec_kernel<V><<<n * FLS_BLOCKS_PER_CHUNK, FLS_WARPS_PER_BLOCK * 32, fls_block_smem<V>(), s>>>(d);
}
template <class V> cudaError_t chunk_attr() {
  return cudaFuncSetAttribute(fls_chunk_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fls_chunk_smem<V>());
}
// persistent grid: one block per SM (fewer when the launch has fewer than two vectors per warp to hand out)
int g_num_sms = 148;
// SM partition (0 = off): the FSST grid takes g_split_fsst SMs (a block fills an SM's shared memory, so nothing else
// lands there) while the bandwidth-bound kernels and then the DICT grid run on the other g_split_dict SMs.
int g_split_fsst = 0, g_split_dict = 0;
template <class V> void chunk_launch(const FlsChunkDev *d, uint32_t n, cudaStream_t s, uint32_t cap = 0) {
  const uint32_t slots = n * FLS_RG_VECS;
  const uint32_t sms = cap ? std::min(cap, (uint32_t)g_num_sms) : (uint32_t)g_num_sms;
  const uint32_t grid = std::max(1u, std::min(sms, slots / (2u * V::WARPS)));
  fls_chunk_kernel<V><<<grid, V::WARPS * 32, fls_chunk_smem<V>(), s>>>(d, n);
}
template <template <typename> class V> cudaError_t vec_attr_all() {
  cudaError_t e;
  if ((e = vec_attr<V<uint8_t>>()) != cudaSuccess) return e;
  if ((e = vec_attr<V<uint16_t>>()) != cudaSuccess) return e;
  if ((e = vec_attr<V<uint32_t>>()) != cudaSuccess) return e;
  return vec_attr<V<uint64_t>>();
}
template <template <typename> class V> bool vec_launch_w(int width, const FlsChunkDev *d, uint32_t n, cudaStream_t s) {
  switch (width) {
    case 1: vec_launch<V<uint8_t>>(d, n, s); return true;
    case 2: vec_launch<V<uint16_t>>(d, n, s); return true;
    case 4: vec_launch<V<uint32_t>>(d, n, s); return true;
    case 8: vec_launch<V<uint64_t>>(d, n, s); return true;
  }
  return false;
}

cudaError_t set_kernel_attrs() {
  cudaError_t e;
  if ((e = vec_attr_all<VecFFOR>()) != cudaSuccess) return e;
  if ((e = vec_attr_all<VecDELTA>()) != cudaSuccess) return e;
  if ((e = vec_attr_all<VecRLE>()) != cudaSuccess) return e;
  if ((e = vec_attr_all<VecDICT>()) != cudaSuccess) return e;
  if ((e = vec_attr_all<VecFFORPATCH>()) != cudaSuccess) return e;
  if ((e = vec_attr_all<VecCROSSRLE>()) != cudaSuccess) return e;
  if ((e = vec_attr<VecALP>()) != cudaSuccess) return e;
  if ((e = vec_attr<VecALP32>()) != cudaSuccess) return e;
  if ((e = vec_attr<VecALPRD>()) != cudaSuccess) return e;
  if ((e = vec_attr<VecSTROFFS>()) != cudaSuccess) return e;
  if ((e = vec_attr<VecSTRCONST>()) != cudaSuccess) return e;
  if ((e = vec_attr<VecSTRFSST12>()) != cudaSuccess) return e;
  if ((e = vec_attr<VecSTRFLAGS>()) != cudaSuccess) return e;
  if ((e = chunk_attr<VecSTRDICT>()) != cudaSuccess) return e;
  if ((e = chunk_attr<VecSTRDICTL>()) != cudaSuccess) return e;
  return chunk_attr<VecSTRFSST>();
}

const char *const kKindName[FLS_K_N] = {"ffor",     "delta",     "rle",      "dict",     "alp",  "alp_rd", "copy",
                                        "fill",     "str_offsets", "str_copy", "str_const", "str_dict", "str_fsst",
                                        "fsst_dictprep", "validity", "ffor_patch", "cross_rle", "str_fsst12", "str_flags", "str_dict_long"};

struct LaunchRec {  // what is needed to (re)issue one launch once its descriptors are on the device
  int kind, width;
  uint32_t n;
  uint64_t rd, wr, vec;
  uint32_t cap = 0;  // persistent grids: at most this many blocks (0 = one per SM)
  size_t doff = 0;   // first descriptor of the launch in the batch's descriptor array
};

bool launch_kind(const LaunchRec &l, const FlsChunkDev *d, cudaStream_t s) {
  const uint32_t n = l.n;
  switch (l.kind) {
    case FLS_K_FFOR: return vec_launch_w<VecFFOR>(l.width, d, n, s);
    case FLS_K_DELTA: return vec_launch_w<VecDELTA>(l.width, d, n, s);
    case FLS_K_RLE: return vec_launch_w<VecRLE>(l.width, d, n, s);
    case FLS_K_DICT: return vec_launch_w<VecDICT>(l.width, d, n, s);
    case FLS_K_FFOR_PATCH: return vec_launch_w<VecFFORPATCH>(l.width, d, n, s);
    case FLS_K_CROSS_RLE: return vec_launch_w<VecCROSSRLE>(l.width, d, n, s);
    case FLS_K_ALP:
      if (l.width == 4) vec_launch<VecALP32>(d, n, s); else vec_launch<VecALP>(d, n, s);
      return true;
    case FLS_K_ALPRD: vec_launch<VecALPRD>(d, n, s); return true;
    case FLS_K_COPY: fls_copy_kernel<<<n * 32, 256, 0, s>>>(d, 32); return true;
    case FLS_K_FILL: fls_fill_kernel<<<n * 32, 256, 0, s>>>(d, 32); return true;
    case FLS_K_STR_OFFSETS: vec_launch<VecSTROFFS>(d, n, s); return true;
    case FLS_K_STR_COPY: fls_str_copy_kernel<<<n * 32, 256, 0, s>>>(d, 32); return true;
    case FLS_K_STR_CONST: vec_launch<VecSTRCONST>(d, n, s); return true;
    case FLS_K_STR_FSST12: vec_launch<VecSTRFSST12>(d, n, s); return true;
    case FLS_K_STR_FLAGS: vec_launch<VecSTRFLAGS>(d, n, s); return true;
    case FLS_K_STR_DICT: chunk_launch<VecSTRDICT>(d, n, s, l.cap); return true;
    case FLS_K_STR_DICT_LONG: chunk_launch<VecSTRDICTL>(d, n, s, l.cap); return true;
    case FLS_K_STR_FSST: chunk_launch<VecSTRFSST>(d, n, s, l.cap); return true;
    case FLS_K_STR_FSST_DICTPREP: fls_fsst_dictprep_kernel<<<n * 8, 128, 0, s>>>(d, 8); return true;
    case FLS_K_VALIDITY: fls_validity_kernel<<<n, 256, 0, s>>>(d); return true;
  }
  return false;
}

// out[j] = base + in[j]: a part's 32-bit string offsets, relative to the part's first payload byte, as 64-bit offsets
__global__ void fls_widen_offsets_kernel(const uint32_t *__restrict__ in, uint64_t *__restrict__ out, uint64_t n, uint64_t base) {
  for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n; j += (uint64_t)gridDim.x * blockDim.x)
    out[j] = base + in[j];
}
__global__ void fls_flush_kernel(uint4 *p, uint64_t n16, uint32_t v) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x)
    p[i] = make_uint4(v, v, v, v);
}

thread_local std::string g_open_err;

struct BufPool {  // size-keyed free lists; buffers live until the handle closes
  std::multimap<size_t, void *> free_dev, free_host;
  void *get(std::multimap<size_t, void *> &fl, size_t n) {
    auto it = fl.lower_bound(n);
    if (it != fl.end() && it->first <= n + n / 4 + 4096) {
      void *p = it->second;
      fl.erase(it);
      return p;
    }
    return nullptr;
  }
};

// Process-wide cache of the buffers closed handles leave behind: a query opens a handle, scans and closes it, and the
// next query's handle would otherwise pay cudaHostAlloc again (about 0.3 s per GB of page-locked memory — more than
// the scan itself).  Bounded (FLSGPU_CACHE_HOST_MB, default 24576; FLSGPU_CACHE_DEV_MB, default 32768);
// flsgpu_trim_cache() frees it.
struct GlobalCache {
  std::mutex mu;
  std::multimap<size_t, void *> host;
  std::map<int, std::multimap<size_t, void *>> dev;
  size_t host_bytes = 0, dev_bytes = 0;
  static size_t cap_mb(const char *name, size_t dflt) { const char *e = getenv(name); return (e ? strtoull(e, nullptr, 10) : dflt) << 20; }
  void *take_host(size_t n, size_t *got) {
    std::lock_guard<std::mutex> g(mu);
    auto it = host.lower_bound(n);
    if (it == host.end() || it->first > n + n / 4 + 4096) return nullptr;
    void *p = it->second;
    *got = it->first; host_bytes -= it->first;
    host.erase(it);
    return p;
  }
  void *take_dev(int device, size_t n, size_t *got) {
    std::lock_guard<std::mutex> g(mu);
    auto &fl = dev[device];
    auto it = fl.lower_bound(n);
    if (it == fl.end() || it->first > n + n / 4 + 4096) return nullptr;
    void *p = it->second;
    *got = it->first; dev_bytes -= it->first;
    fl.erase(it);
    return p;
  }
  bool give_host(void *p, size_t n) {
    std::lock_guard<std::mutex> g(mu);
    if (host_bytes + n > cap_mb("FLSGPU_CACHE_HOST_MB", 24576)) return false;
    host.emplace(n, p); host_bytes += n;
    return true;
  }
  bool give_dev(int device, void *p, size_t n) {
    std::lock_guard<std::mutex> g(mu);
    if (dev_bytes + n > cap_mb("FLSGPU_CACHE_DEV_MB", 32768)) return false;
    dev[device].emplace(n, p); dev_bytes += n;
    return true;
  }
  // a handle's six streams and its events outlive it too (a query = open + scan + close, on every scanning thread:
  // creating and destroying them was a fixed cost per query and thread); at most 64 sets / 4096 events are kept
  struct StreamSet { cudaStream_t s[6]; };
  std::map<int, std::vector<StreamSet>> streams;
  std::map<int, std::vector<cudaEvent_t>> events;
  std::map<int, bool> attrs_set;  // cudaFuncSetAttribute of every kernel: once per device
  bool take_streams(int device, StreamSet *out) {
    std::lock_guard<std::mutex> g(mu);
    auto &v = streams[device];
    if (v.empty()) return false;
    *out = v.back(); v.pop_back();
    return true;
  }
  bool give_streams(int device, const StreamSet &ss) {
    std::lock_guard<std::mutex> g(mu);
    auto &v = streams[device];
    if (v.size() >= 64) return false;
    v.push_back(ss);
    return true;
  }
  void take_events(int device, std::vector<cudaEvent_t> &into, size_t n) {
    std::lock_guard<std::mutex> g(mu);
    auto &v = events[device];
    while (n-- && !v.empty()) { into.push_back(v.back()); v.pop_back(); }
  }
  bool give_event(int device, cudaEvent_t e) {
    std::lock_guard<std::mutex> g(mu);
    auto &v = events[device];
    if (v.size() >= 4096) return false;
    v.push_back(e);
    return true;
  }
};
GlobalCache g_cache;

struct KStat {
  std::string name;
  uint64_t launches = 0, rd = 0, wr = 0, vectors = 0;
  double ms = 0;
};

}  // namespace

struct flsgpu {
  int device = 0;
  uint32_t flags = 0, slice_rgs = 0;
  FlsTable t;
  std::string path;
  const uint8_t *image = nullptr;  // host image (whole file)
  uint64_t size = 0;
  bool image_owned = false, image_registered = false, cuda_ready = false;
  bool image_mapped = false;  // opened by path: `image` is a read-only mapping of the file (pageable)
  int fd = -1;
  void *gds = nullptr;        // cuFile state (FLSGPU_OPEN_GDS)
  cudaStream_t s_h2d = nullptr, s_k = nullptr, s_d2h = nullptr;
  cudaStream_t s_kx[3] = {nullptr, nullptr, nullptr};  // side streams: independent kernels of a slice overlap
  uint8_t *d_res = nullptr;  // resident compressed row groups [res_g0, res_g1)
  uint64_t res_bytes = 0;
  uint32_t res_g0 = 0, res_g1 = 0;
  uint8_t *d_flush = nullptr;
  uint64_t flush_bytes = 0;
  std::string err;
  BufPool pool;
  std::vector<cudaEvent_t> ev_free;
  std::vector<KStat> kstats;
  std::vector<const char *> name_ptrs;
  std::vector<int32_t> types32;
  std::vector<uint8_t> validated;  // [rg * n_cols + col]: the chunk's footer entry and vector headers passed fls_validate_chunk
  // Plans of the last single-slice resident decodes: repeated scans of the same range into the same (pooled)
  // buffers reuse the descriptors already on the device instead of rebuilding and re-uploading them.  A scan that
  // walks a large table batch by batch (bench.py --scaling strong, the facade's cursor) keeps one plan per batch.
  struct PlanCache {
    bool valid = false;
    uint32_t g0 = 0, g1 = 0;
    uint64_t stamp = 0;
    std::vector<uint32_t> cols;
    std::vector<const void *> ptrs;
    std::vector<LaunchRec> recs;
    FlsChunkDev *d_desc = nullptr;
  };
  static constexpr size_t MAX_PLANS = 16;
  std::vector<PlanCache> plans;
  uint64_t plan_clock = 0;

  int fail(int code, const std::string &m) { err = m; return code; }
  int fail_cuda(const char *what, cudaError_t e) {
    err = std::string("CUDA: ") + what + ": " + cudaGetErrorString(e);
    return FLSGPU_E_CUDA;
  }
  cudaEvent_t ev_get() {  // nullptr when the runtime cannot create one (the caller bails)
    if (!ev_free.empty()) { cudaEvent_t e = ev_free.back(); ev_free.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return e;
  }
  // `prefer`: the buffer a cached plan of the same decode was built for — taken when it is free and fits, so that a
  // scan that walks a table batch by batch, again and again, finds each batch's descriptors still valid
  void *dev_alloc(size_t n, const void *prefer = nullptr) {
    n = (n + 255) & ~(size_t)255;
    if (!n) n = 256;
    if (prefer)
      for (auto it = pool.free_dev.lower_bound(n); it != pool.free_dev.end() && it->first <= n + n / 4 + 4096; ++it)
        if (it->second == prefer) { void *q = it->second; pool.free_dev.erase(it); return q; }
    void *p = pool.get(pool.free_dev, n);
    if (p) return p;
    size_t got = 0;
    if ((p = g_cache.take_dev(device, n, &got))) { dev_size[p] = got; return p; }
    if (cudaMalloc(&p, n) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    dev_size[p] = n;
    return p;
  }
  void *host_alloc(size_t n) {
    n = (n + 4095) & ~(size_t)4095;
    if (!n) n = 4096;
    void *p = pool.get(pool.free_host, n);
    if (p) return p;
    size_t got = 0;
    if ((p = g_cache.take_host(n, &got))) { host_size[p] = got; return p; }
    if (cudaHostAlloc(&p, n, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    host_size[p] = n;
    return p;
  }
  void dev_free(void *p) { if (p) pool.free_dev.emplace(dev_size[p], p); }
  void host_free(void *p) { if (p) pool.free_host.emplace(host_size[p], p); }
  std::map<void *, size_t> dev_size, host_size;
};

#define CK(call)                                          \
  do {                                                    \
    cudaError_t e_ = (call);                              \
    if (e_ != cudaSuccess) return h->fail_cuda(#call, e_); \
  } while (0)

struct flsgpu_batch {
  flsgpu *h = nullptr;
  uint32_t g0 = 0, g1 = 0, flags = 0;
  std::vector<uint32_t> cols;
  std::vector<flsgpu_column> out;      // what the caller sees
  std::vector<FlsColOut> dev;          // device buffers
  std::vector<FlsColNeed> need;
  std::vector<void *> dev_bufs, host_bufs;
  flsgpu_timing tm;
  // string columns whose payload reaches 4 GiB are decoded in PARTS of less than 4 GiB: inside a part the kernels'
  // 32-bit offsets are relative to the part's first payload byte (part k's offsets live 4 k entries further on in the
  // device array, so that a part's first entry can be 0 while the previous part's last end stays); a last kernel
  // widens them into the 64-bit offsets the caller sees
  struct StrPart { uint32_t g_first; uint64_t row_base, pay_base; };
  std::vector<std::vector<StrPart>> parts;   // per column; empty = not wide
  std::vector<uint64_t *> off64_dev;
  // filtered batches: row ids of the surviving rows (device, and host for FLSGPU_OUT_HOST)
  bool filtered = false;
  uint64_t *sel_dev = nullptr, *sel_host = nullptr;
  uint64_t n_sel = 0;
};

namespace {

// ---- GPUDirect Storage (cuFile), bound at run time: libflsgpu.so does not link libcufile ------------------------
// The handful of declarations used, as in cufile.h (CUDA 12.9): CUfileError_t {err, cu_err}; CU_FILE_SUCCESS = 0;
// descriptor type CU_FILE_HANDLE_TYPE_OPAQUE_FD = 1.
struct GdsApi {
  struct Err { int err; int cu_err; };
  struct Descr { int type; union { int fd; void *handle; } h; const void *fs_ops; };
  void *lib = nullptr, *fh = nullptr;
  Err (*DriverOpen)() = nullptr;
  Err (*HandleRegister)(void **, Descr *) = nullptr;
  void (*HandleDeregister)(void *) = nullptr;
  Err (*BufRegister)(const void *, size_t, int) = nullptr;
  Err (*BufDeregister)(const void *) = nullptr;
  ssize_t (*Read)(void *, void *, size_t, off_t, off_t) = nullptr;
};
int gds_open(flsgpu *h) {
  // Without the nvidia-fs kernel driver cuFile can only run its compatibility mode (POSIX reads into a bounce buffer:
  // what the mapped path already does, on more threads) — and on the B200 boxes of this pool, which have no
  // nvidia-fs, cuFileDriverOpen / cuFileRead did not return at all.  So: no driver, no GDS, said loudly.
  // FLSGPU_GDS_COMPAT=1 tries cuFile anyway.
  struct stat st;
  const char *force = getenv("FLSGPU_GDS_COMPAT");
  if (stat("/proc/driver/nvidia-fs", &st) != 0 && !(force && *force == '1'))
    return h->fail(FLSGPU_E_IO, "FLSGPU_OPEN_GDS: the nvidia-fs kernel driver is not loaded (no /proc/driver/nvidia-fs): "
                                "GPUDirect Storage is unavailable on this box; open without the flag");
  GdsApi *g = new GdsApi();
  g->lib = dlopen("libcufile.so.0", RTLD_NOW | RTLD_LOCAL);
  if (!g->lib) g->lib = dlopen("libcufile.so", RTLD_NOW | RTLD_LOCAL);
  if (!g->lib) { std::string m = std::string("FLSGPU_OPEN_GDS: cannot load libcufile: ") + dlerror(); delete g; return h->fail(FLSGPU_E_IO, m); }
  g->DriverOpen = (GdsApi::Err(*)())dlsym(g->lib, "cuFileDriverOpen");
  g->HandleRegister = (GdsApi::Err(*)(void **, GdsApi::Descr *))dlsym(g->lib, "cuFileHandleRegister");
  g->HandleDeregister = (void (*)(void *))dlsym(g->lib, "cuFileHandleDeregister");
  g->BufRegister = (GdsApi::Err(*)(const void *, size_t, int))dlsym(g->lib, "cuFileBufRegister");
  g->BufDeregister = (GdsApi::Err(*)(const void *))dlsym(g->lib, "cuFileBufDeregister");
  g->Read = (ssize_t(*)(void *, void *, size_t, off_t, off_t))dlsym(g->lib, "cuFileRead");
  if (!g->DriverOpen || !g->HandleRegister || !g->HandleDeregister || !g->Read) {
    dlclose(g->lib); delete g;
    return h->fail(FLSGPU_E_IO, "FLSGPU_OPEN_GDS: libcufile lacks the cuFile entry points");
  }
  GdsApi::Err e = g->DriverOpen();
  if (e.err != 0) {
    const int code = e.err;
    dlclose(g->lib); delete g;
    return h->fail(FLSGPU_E_IO, "FLSGPU_OPEN_GDS: cuFileDriverOpen failed (CUfileOpError " + std::to_string(code) + "): no GPUDirect Storage on this box");
  }
  GdsApi::Descr d;
  memset(&d, 0, sizeof(d));
  d.type = 1;
  d.h.fd = h->fd;
  e = g->HandleRegister(&g->fh, &d);
  if (e.err != 0) {
    const int code = e.err;
    dlclose(g->lib); delete g;
    return h->fail(FLSGPU_E_IO, "FLSGPU_OPEN_GDS: cuFileHandleRegister failed (CUfileOpError " + std::to_string(code) + ") for " + h->path);
  }
  h->gds = g;
  return 0;
}
void gds_close(flsgpu *h) {
  GdsApi *g = (GdsApi *)h->gds;
  if (!g) return;
  if (g->fh) g->HandleDeregister(g->fh);
  dlclose(g->lib);
  delete g;
  h->gds = nullptr;
}
// file bytes [off, off + n) -> device memory at dst (synchronous: cuFileRead returns when the bytes are there)
int gds_read(flsgpu *h, void *dst, uint64_t off, uint64_t n) {
  GdsApi *g = (GdsApi *)h->gds;
  uint64_t done = 0;
  while (done < n) {
    const ssize_t r = g->Read(g->fh, dst, (size_t)(n - done), (off_t)(off + done), (off_t)done);
    if (r <= 0) return h->fail(FLSGPU_E_IO, "cuFileRead failed (" + std::to_string((long long)r) + ") on " + h->path);
    done += (uint64_t)r;
  }
  return 0;
}
// pageable bytes -> page-locked bounce buffer, on a few threads (one thread copies out of the page cache at a fraction
// of what the PCIe link takes)
void parallel_copy(void *dst, const void *src, uint64_t n) {
  const uint64_t chunk = 8ull << 20;
  const unsigned nt = (unsigned)std::min<uint64_t>(4, (n + chunk - 1) / chunk);
  if (nt <= 1) { memcpy(dst, src, n); return; }
  std::vector<std::thread> th;
  for (unsigned i = 0; i < nt; ++i) {
    const uint64_t a = n * i / nt, b = n * (i + 1) / nt;
    th.emplace_back([=] { memcpy((uint8_t *)dst + a, (const uint8_t *)src + a, b - a); });
  }
  for (auto &t : th) t.join();
}

// Bring CUDA up for this handle and make the whole host image available (page-locked unless told otherwise).
int ensure_ready(flsgpu *h) {
  if (h->cuda_ready) { CK(cudaSetDevice(h->device)); return 0; }
  CK(cudaSetDevice(h->device));
  {
    GlobalCache::StreamSet ss;
    if (g_cache.take_streams(h->device, &ss)) {
      h->s_h2d = ss.s[0]; h->s_k = ss.s[1]; h->s_d2h = ss.s[2];
      for (int i = 0; i < 3; ++i) h->s_kx[i] = ss.s[3 + i];
    } else {
      CK(cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking));
      CK(cudaStreamCreateWithFlags(&h->s_k, cudaStreamNonBlocking));
      CK(cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking));
      for (auto &sx : h->s_kx) CK(cudaStreamCreateWithFlags(&sx, cudaStreamNonBlocking));
    }
    g_cache.take_events(h->device, h->ev_free, 64);
    bool need_attrs;
    { std::lock_guard<std::mutex> g(g_cache.mu); need_attrs = !g_cache.attrs_set[h->device]; }
    if (need_attrs) {
      CK(set_kernel_attrs());
      std::lock_guard<std::mutex> g(g_cache.mu);
      g_cache.attrs_set[h->device] = true;
    }
  }
  CK(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, h->device));
  if (const char *e = getenv("FLSGPU_SM_SPLIT")) {  // "fsst,dict" SM counts; "0" = no partition
    g_split_fsst = g_split_dict = 0;
    sscanf(e, "%d,%d", &g_split_fsst, &g_split_dict);
  }
  if (!h->image && !(h->flags & FLSGPU_OPEN_LOAD_FILE)) {
    // opened by path: map the file.  Footer, vector headers and dictionaries are read through the mapping (page
    // cache); the compressed bytes of each pipeline slice are copied into a page-locked bounce buffer right before
    // their H2D copy (stage_slice) — or read by cuFile straight into HBM — so a scan never waits for the whole file.
    h->fd = open(h->path.c_str(), O_RDONLY);
    void *p = h->fd >= 0 ? mmap(nullptr, h->size, PROT_READ, MAP_SHARED, h->fd, 0) : MAP_FAILED;
    if (p == MAP_FAILED) {
      if (h->fd >= 0) close(h->fd);
      h->fd = -1;
      return h->fail(FLSGPU_E_IO, "Failed to open FastLanes file: " + h->path);
    }
    h->image = (const uint8_t *)p;
    h->image_mapped = true;
    if (h->flags & FLSGPU_OPEN_GDS) { if (int rc = gds_open(h)) return rc; }
  } else if (!h->image) {  // FLSGPU_OPEN_LOAD_FILE: read the whole file into page-locked memory
    void *p = nullptr;
    const bool pin = !(h->flags & FLSGPU_OPEN_NO_PIN);
    if (pin) CK(cudaHostAlloc(&p, h->size, cudaHostAllocDefault));
    else if (!(p = malloc(h->size))) return h->fail(FLSGPU_E_NOMEM, "out of host memory");
    int fd = open(h->path.c_str(), O_RDONLY);
    uint64_t got = 0;
    while (fd >= 0 && got < h->size) {
      ssize_t r = pread(fd, (uint8_t *)p + got, std::min<uint64_t>(h->size - got, 1ull << 30), (off_t)got);
      if (r <= 0) break;
      got += (uint64_t)r;
    }
    if (fd >= 0) close(fd);
    if (got != h->size) {
      if (pin) cudaFreeHost(p); else free(p);
      return h->fail(FLSGPU_E_IO, "Failed to open FastLanes file: " + h->path);
    }
    h->image = (const uint8_t *)p;
    h->image_owned = true;
    h->image_registered = false;
    if (!pin) h->flags |= FLSGPU_OPEN_NO_PIN;
  } else if (!(h->flags & (FLSGPU_OPEN_PINNED_IMAGE | FLSGPU_OPEN_NO_PIN))) {
    // caller's pageable image: page-lock it in place (read-only registration is enough for H2D sources)
    cudaError_t e = cudaHostRegister((void *)h->image, h->size, cudaHostRegisterReadOnly);
    if (e != cudaSuccess) { cudaGetLastError(); e = cudaHostRegister((void *)h->image, h->size, cudaHostRegisterDefault); }
    if (e == cudaSuccess) h->image_registered = true;
    else cudaGetLastError();  // not fatal: copies fall back to the driver's staging path
  }
  h->cuda_ready = true;
  return 0;
}

// Byte ranges (relative to the file offset of row group s0) that a decode of columns `cols` over row groups [s0, s1)
// reads: the projected columns' chunks (and the chunk an EQUAL chunk names), each padded by 32 bytes for the kernels'
// 16-byte-granular reads, merged where the gap between two is small — a copy call costs about as much as 128 KB on
// the bus.  Every column projected = one range = the whole extent.
struct StageRange { uint64_t off, len; };
std::vector<StageRange> projected_ranges(const FlsTable &t, uint32_t s0, uint32_t s1, const std::vector<uint32_t> &cols, uint64_t ext) {
  std::vector<StageRange> r;
  std::vector<uint32_t> cs(cols);
  std::sort(cs.begin(), cs.end());
  cs.erase(std::unique(cs.begin(), cs.end()), cs.end());
  if (cs.size() >= t.n_cols) { r.push_back({0, ext}); return r; }
  auto add = [&](uint64_t a, uint64_t b) {
    b = std::min(ext, b + 32);
    a &= ~(uint64_t)15;
    if (!r.empty() && a <= r.back().off + r.back().len + (128u << 10)) {
      if (b > r.back().off + r.back().len) r.back().len = b - r.back().off;
    } else if (b > a) r.push_back({a, b - a});
  };
  for (uint32_t g = s0; g < s1; ++g) {
    const uint64_t base = t.rgs[g].off - t.rgs[s0].off;
    std::vector<std::pair<uint64_t, uint64_t>> ch;
    for (uint32_t c : cs) {
      const fls_colchunk *k = &t.chunk(g, c);
      if (k->scheme == FLS_S_EQUAL && k->aux32 < t.n_cols) k = &t.chunk(g, k->aux32);
      uint64_t a = ~0ull, b = 0;
      for (uint32_t sgi = 0; sgi < k->n_seg; ++sgi) {
        a = std::min<uint64_t>(a, k->seg[sgi].off);
        b = std::max<uint64_t>(b, k->seg[sgi].off + k->seg[sgi].size);
      }
      if (b > a) ch.emplace_back(base + a, base + b);
    }
    std::sort(ch.begin(), ch.end());
    for (auto &x : ch) add(x.first, x.second);
  }
  return r;
}

uint64_t range_extent(const FlsTable &t, uint32_t g0, uint32_t g1) {
  return g1 > g0 ? t.rgs[g1 - 1].off + t.rgs[g1 - 1].size - t.rgs[g0].off : 0;
}

flsgpu *new_handle(const flsgpu_opts *o) {
  flsgpu *h = new flsgpu();
  if (o) { h->device = o->device; h->flags = o->flags; h->slice_rgs = o->slice_rowgroups; }
  h->kstats.resize(FLS_K_N * 5);
  for (int k = 0; k < FLS_K_N; ++k)
    for (int w = 0; w < 5; ++w) {
      static const char *const wn[5] = {"", "_u8", "_u16", "_u32", "_u64"};
      h->kstats[k * 5 + w].name = std::string(kKindName[k]) + wn[w];
    }
  return h;
}
void finish_schema(flsgpu *h) {
  for (auto &n : h->t.names) h->name_ptrs.push_back(n.c_str());
  for (int ty : h->t.types) h->types32.push_back(ty);
}
int width_slot(int width) { return width == 1 ? 1 : width == 2 ? 2 : width == 4 ? 3 : width == 8 ? 4 : 0; }

}  // namespace

extern "C" {

int flsgpu_abi_version(void) { return FLSGPU_ABI_VERSION; }
int flsgpu_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int flsgpu_open(const char *path, const flsgpu_opts *opts, flsgpu_t **out) {
  if (!out || !path) { g_open_err = "null argument"; return FLSGPU_E_RANGE; }
  *out = nullptr;
  flsgpu *h = new_handle(opts);
  h->path = path;
  struct stat st;
  if (stat(path, &st) != 0 || !S_ISREG(st.st_mode)) {
    // same text the reference's bind raises (/root/reference/src/scan_fastlanes.cpp:42)
    g_open_err = std::string("Failed to open FastLanes file: ") + path;
    delete h;
    return FLSGPU_E_IO;
  }
  std::string e = fls_parse_file(path, h->t, &h->size);
  if (!e.empty()) {
    g_open_err = std::string("Failed to open FastLanes file: ") + path + " (" + e + ")";
    delete h;
    return FLSGPU_E_FORMAT;
  }
  finish_schema(h);
  *out = h;
  return FLSGPU_OK;
}

int flsgpu_open_image(const void *image, uint64_t size, const flsgpu_opts *opts, flsgpu_t **out) {
  if (!out || !image) { g_open_err = "null argument"; return FLSGPU_E_RANGE; }
  *out = nullptr;
  flsgpu *h = new_handle(opts);
  std::string e = fls_parse((const uint8_t *)image, size, h->t);
  if (!e.empty()) {
    g_open_err = "Failed to open FastLanes image (" + e + ")";
    delete h;
    return FLSGPU_E_FORMAT;
  }
  h->image = (const uint8_t *)image;
  h->size = size;
  finish_schema(h);
  *out = h;
  return FLSGPU_OK;
}

void flsgpu_close(flsgpu_t *h) {
  if (!h) return;
  if (h->cuda_ready) {
    cudaSetDevice(h->device);
    // wait for this handle's own work only: other handles (the scan's other threads) keep the device busy
    for (cudaStream_t st : {h->s_h2d, h->s_k, h->s_d2h, h->s_kx[0], h->s_kx[1], h->s_kx[2]}) if (st) cudaStreamSynchronize(st);
    // the pooled buffers go to the process-wide cache (up to its bounds) for the next handle
    for (auto &kv : h->pool.free_dev) { h->dev_size.erase(kv.second); if (!g_cache.give_dev(h->device, kv.second, kv.first)) cudaFree(kv.second); }
    for (auto &kv : h->pool.free_host) { h->host_size.erase(kv.second); if (!g_cache.give_host(kv.second, kv.first)) cudaFreeHost(kv.second); }
    for (auto &kv : h->dev_size) cudaFree(kv.first);       // (buffers of batches the caller never released)
    for (auto &kv : h->host_size) cudaFreeHost(kv.first);
    if (h->d_res) cudaFree(h->d_res);
    if (h->d_flush) cudaFree(h->d_flush);
    for (auto e : h->ev_free) if (!g_cache.give_event(h->device, e)) cudaEventDestroy(e);
    if (h->image_registered) cudaHostUnregister((void *)h->image);
    gds_close(h);
    if (h->image_mapped) munmap((void *)h->image, h->size);
    if (h->fd >= 0) close(h->fd);
    if (h->image_owned) { if (h->flags & FLSGPU_OPEN_NO_PIN) free((void *)h->image); else cudaFreeHost((void *)h->image); }
    GlobalCache::StreamSet ss = {{h->s_h2d, h->s_k, h->s_d2h, h->s_kx[0], h->s_kx[1], h->s_kx[2]}};
    if (!h->s_h2d || !h->s_k || !h->s_d2h || !h->s_kx[0] || !h->s_kx[1] || !h->s_kx[2] || !g_cache.give_streams(h->device, ss))
      for (cudaStream_t st : ss.s) if (st) cudaStreamDestroy(st);
  }
  delete h;
}

void flsgpu_trim_cache(void) {
  std::lock_guard<std::mutex> g(g_cache.mu);
  for (auto &kv : g_cache.host) cudaFreeHost(kv.second);
  g_cache.host.clear(); g_cache.host_bytes = 0;
  int cur = 0;
  cudaGetDevice(&cur);
  for (auto &d : g_cache.dev) {
    cudaSetDevice(d.first);
    for (auto &kv : d.second) cudaFree(kv.second);
  }
  g_cache.dev.clear(); g_cache.dev_bytes = 0;
  for (auto &d : g_cache.streams) {
    cudaSetDevice(d.first);
    for (auto &ss : d.second) for (cudaStream_t st : ss.s) cudaStreamDestroy(st);
  }
  g_cache.streams.clear();
  for (auto &d : g_cache.events) {
    cudaSetDevice(d.first);
    for (cudaEvent_t e : d.second) cudaEventDestroy(e);
  }
  g_cache.events.clear();
  cudaSetDevice(cur);
}

const char *flsgpu_last_error(const flsgpu_t *h) { return h ? h->err.c_str() : g_open_err.c_str(); }

uint32_t flsgpu_num_columns(const flsgpu_t *h) { return h->t.n_cols; }
uint32_t flsgpu_num_rowgroups(const flsgpu_t *h) { return h->t.n_rg; }
uint64_t flsgpu_num_rows(const flsgpu_t *h) { return h->t.n_rows; }
uint32_t flsgpu_rowgroup_rows(const flsgpu_t *h, uint32_t rg) { return rg < h->t.n_rg ? h->t.rgs[rg].n_tuples : 0; }

int flsgpu_schema(const flsgpu_t *h, uint32_t *n_cols, const char *const **names, const int32_t **types) {
  if (n_cols) *n_cols = h->t.n_cols;
  if (names) *names = h->name_ptrs.data();
  if (types) *types = h->types32.data();
  return FLSGPU_OK;
}

int flsgpu_chunk_info(const flsgpu_t *h, uint32_t rg, uint32_t col, int32_t *scheme, uint64_t *bytes) {
  if (rg >= h->t.n_rg || col >= h->t.n_cols) return FLSGPU_E_RANGE;
  const fls_colchunk &k = h->t.chunk(rg, col);
  if (scheme) *scheme = k.scheme;
  if (bytes) {
    *bytes = 0;
    for (uint32_t s = 0; s < k.n_seg; ++s) *bytes += k.seg[s].size;
  }
  return FLSGPU_OK;
}

int flsgpu_rowgroup_minmax(flsgpu_t *h, uint32_t rg, uint32_t col, int64_t *mn, int64_t *mx, int32_t *known) {
  if (!known || !mn || !mx) return h->fail(FLSGPU_E_RANGE, "null argument");
  *known = 0;
  if (rg >= h->t.n_rg || col >= h->t.n_cols) return h->fail(FLSGPU_E_RANGE, "bad row group / column id");
  const int type = h->t.types[col];
  const uint32_t tb = fls_type_bits(type);
  const bool is_signed = type == FLS_T_INT8 || type == FLS_T_INT16 || type == FLS_T_INT32 || type == FLS_T_INT64 || type == FLS_T_DATE;
  const bool is_unsigned = type == FLS_T_UINT8 || type == FLS_T_UINT16 || type == FLS_T_UINT32 || type == FLS_T_UINT64 || type == FLS_T_BOOLEAN;
  if (!tb || !(is_signed || is_unsigned)) return FLSGPU_OK;
  const fls_colchunk &k0 = h->t.chunk(rg, col);
  // EQUAL: the bounds of the chunk it repeats (same width; read in this column's own signedness)
  const bool alias = k0.scheme == FLS_S_EQUAL && k0.aux32 < h->t.n_cols && h->t.chunk(rg, k0.aux32).scheme != FLS_S_EQUAL &&
                     fls_type_bits(h->t.types[k0.aux32]) == tb;
  const fls_colchunk &k = alias ? h->t.chunk(rg, k0.aux32) : k0;
  const fls_rowgroup &r = h->t.rgs[rg];
  const uint64_t tmask = tb == 64 ? ~0ull : ((1ull << tb) - 1ull);
  auto as_i64 = [&](uint64_t v, int64_t *o) {  // value of the column's type -> int64; false if it does not fit
    if (is_signed) { *o = tb == 64 ? (int64_t)v : (int64_t)(v << (64 - tb)) >> (64 - tb); return true; }
    if (v > (uint64_t)INT64_MAX) return false;
    *o = (int64_t)v;
    return true;
  };
  if (k.scheme == FLS_S_CONSTANT) {
    int64_t v;
    if (!as_i64(k.aux64[0] & tmask, &v)) return FLSGPU_OK;
    *mn = *mx = v; *known = 1;
    return FLSGPU_OK;
  }
  const bool delta = k.scheme == FLS_S_DELTA, patch = k.scheme == FLS_S_FFOR_PATCH;
  int fd = -1;
  auto read_at = [&](uint64_t off, void *dst, size_t n) {  // bytes of the file: the host image, or the file itself
    if (h->image) { memcpy(dst, h->image + off, n); return true; }
    if (fd < 0) fd = open(h->path.c_str(), O_RDONLY);
    return fd >= 0 && pread(fd, dst, n, (off_t)off) == (ssize_t)n;
  };
  struct Closer { int &fd; ~Closer() { if (fd >= 0) close(fd); } } closer{fd};
  if (k.scheme == FLS_S_CROSS_RLE) {  // the run values are the column's distinct neighbours: their min / max are exact
    const uint64_t n = k.aux32;
    if (k.n_seg < 2 || !n || n > (1u << 22) || k.seg[0].size < n * (tb / 8)) return FLSGPU_OK;
    std::vector<uint8_t> vals(n * (tb / 8));
    if (!read_at(r.off + k.seg[0].off, vals.data(), vals.size())) return h->fail(FLSGPU_E_IO, "cannot read the run values");
    int64_t lo = INT64_MAX, hi = INT64_MIN;
    for (uint64_t i = 0; i < n; ++i) {
      uint64_t b = 0;
      memcpy(&b, vals.data() + i * (tb / 8), tb / 8);
      int64_t a;
      if (!as_i64(b, &a)) return FLSGPU_OK;
      lo = std::min(lo, a); hi = std::max(hi, a);
    }
    *mn = lo; *mx = hi; *known = 1;
    return FLSGPU_OK;
  }
  if ((k.scheme != FLS_S_FFOR && !delta && !patch) || k.n_seg < 2 || k.seg[0].size < 16ull * r.n_vec) return FLSGPU_OK;
  std::vector<uint64_t> hdr(2ull * r.n_vec);
  if (!read_at(r.off + k.seg[0].off, hdr.data(), hdr.size() * 8)) return h->fail(FLSGPU_E_IO, "cannot read vector headers");
  const int64_t tmax = is_signed ? (tb == 64 ? INT64_MAX : (int64_t)((1ull << (tb - 1)) - 1)) : (tb == 64 ? INT64_MAX : (int64_t)tmask);
  const int64_t tmin = is_signed ? -tmax - 1 : 0;
  __int128 lo = tmax, hi = tmin;
  for (uint32_t v = 0; v < r.n_vec; ++v) {
    const uint32_t bw = FLS_META_BW((uint32_t)(hdr[2 * v] >> 32));
    if (bw >= tb) return FLSGPU_OK;  // the vector spans the whole type
    const uint64_t base = hdr[2 * v + 1] & tmask, span = (1ull << bw) - 1ull;
    if (!delta) {
      // FFOR: values in [base, base + span]; the interval must not wrap in the column's own order
      int64_t a;
      if (!as_i64(base, &a)) return FLSGPU_OK;
      lo = std::min<__int128>(lo, a);
      hi = std::max<__int128>(hi, (__int128)a + span);
      if (patch) {  // ... and the exception values, wherever they lie
        const uint32_t ep = (uint32_t)hdr[2 * v], nx = FLS_META_N((uint32_t)(hdr[2 * v] >> 32));
        const uint64_t at = 16ull * ep + 128ull * bw, nb = (uint64_t)nx * (tb / 8);
        if (nx > FLS_VEC || k.seg[1].size < at + nb) return FLSGPU_OK;
        std::vector<uint8_t> xv(nb);
        if (nb && !read_at(r.off + k.seg[1].off + at, xv.data(), nb)) return FLSGPU_OK;
        for (uint32_t i = 0; i < nx; ++i) {
          uint64_t b = 0;
          memcpy(&b, xv.data() + (size_t)i * (tb / 8), tb / 8);
          int64_t x;
          if (!as_i64(b, &x)) return FLSGPU_OK;
          lo = std::min<__int128>(lo, x); hi = std::max<__int128>(hi, x);
        }
      }
    } else {
      // DELTA: element r (1..T) of a lane = lane base + r deltas, every delta in [dmin, dmin + span] (signed)
      const uint32_t ep = (uint32_t)hdr[2 * v];
      uint8_t lb[128];
      if (k.seg[1].size < 16ull * ep + 128u || !read_at(r.off + k.seg[1].off + 16ull * ep, lb, 128)) return FLSGPU_OK;
      const int64_t dmin = tb == 64 ? (int64_t)base : (int64_t)(base << (64 - tb)) >> (64 - tb);
      const __int128 dmax = (__int128)dmin + span;
      const __int128 down = std::min<__int128>(dmin, (__int128)tb * dmin), up = std::max<__int128>(dmax, (__int128)tb * dmax);
      for (uint32_t l = 0; l < 1024u / tb; ++l) {
        uint64_t b = 0;
        memcpy(&b, lb + l * (tb / 8), tb / 8);
        int64_t a;
        if (!as_i64(b, &a)) return FLSGPU_OK;
        lo = std::min<__int128>(lo, (__int128)a + down);
        hi = std::max<__int128>(hi, (__int128)a + up);
      }
    }
  }
  if (lo < tmin || hi > tmax) return FLSGPU_OK;  // the bounds leave the type: wrapping arithmetic, nothing to say
  *mn = (int64_t)lo; *mx = (int64_t)hi; *known = 1;
  return FLSGPU_OK;
}

int flsgpu_rowgroup_dict_contains(flsgpu_t *h, uint32_t rg, uint32_t col, const uint8_t *bytes, uint64_t len,
                                  int32_t *found, int32_t *known) {
  if (!known || !found || (!bytes && len)) return h->fail(FLSGPU_E_RANGE, "null argument");
  *known = 0; *found = 0;
  if (rg >= h->t.n_rg || col >= h->t.n_cols) return h->fail(FLSGPU_E_RANGE, "bad row group / column id");
  if (fls_type_bits(h->t.types[col])) return FLSGPU_OK;  // not a string column
  const fls_colchunk &k = h->t.chunk(rg, col);
  const fls_rowgroup &r = h->t.rgs[rg];
  int fd = -1;
  auto read_at = [&](uint64_t off, void *dst, size_t n) {
    if (!n) return true;
    if (h->image) { memcpy(dst, h->image + off, n); return true; }
    if (fd < 0) fd = open(h->path.c_str(), O_RDONLY);
    return fd >= 0 && pread(fd, dst, n, (off_t)off) == (ssize_t)n;
  };
  struct Closer { int &fd; ~Closer() { if (fd >= 0) close(fd); } } closer{fd};
  if (k.scheme == FLS_S_STR_CONSTANT) {
    if (k.n_seg < 1 || k.seg[0].size < k.aux32) return FLSGPU_OK;
    *known = 1;
    if (k.aux32 != len) return FLSGPU_OK;
    std::vector<uint8_t> v(len);
    if (!read_at(r.off + k.seg[0].off, v.data(), len)) { *known = 0; return h->fail(FLSGPU_E_IO, "cannot read the constant"); }
    *found = !len || memcmp(v.data(), bytes, len) == 0;
    return FLSGPU_OK;
  }
  // STR_DICT: seg[2] = entry offsets u32[n + 1], seg[3] = entry bytes; dictionaries beyond 64 MiB are left alone
  const uint64_t n = k.aux32;
  if (k.scheme != FLS_S_STR_DICT || k.n_seg < 4 || k.seg[2].size < 4ull * (n + 1) || k.seg[3].size > (64ull << 20)) return FLSGPU_OK;
  std::vector<uint32_t> off(n + 1);
  std::vector<uint8_t> dict(k.seg[3].size);
  if (!read_at(r.off + k.seg[2].off, off.data(), 4ull * (n + 1)) || !read_at(r.off + k.seg[3].off, dict.data(), dict.size()))
    return h->fail(FLSGPU_E_IO, "cannot read the dictionary");
  for (uint64_t i = 0; i < n; ++i)
    if (off[i] > off[i + 1] || off[i + 1] > dict.size()) return FLSGPU_OK;  // a dictionary that lies says nothing
  *known = 1;
  for (uint64_t i = 0; i < n && !*found; ++i)
    *found = off[i + 1] - off[i] == len && (!len || memcmp(dict.data() + off[i], bytes, len) == 0);
  return FLSGPU_OK;
}

int flsgpu_make_resident(flsgpu_t *h, uint32_t g0, uint32_t g1) {
  if (g0 > g1 || g1 > h->t.n_rg) return h->fail(FLSGPU_E_RANGE, "bad row-group range");
  if (int rc = ensure_ready(h)) return rc;
  if (h->d_res) { CK(cudaFree(h->d_res)); h->d_res = nullptr; h->res_g0 = h->res_g1 = 0; }
  const uint64_t bytes = range_extent(h->t, g0, g1);
  if (!bytes) return FLSGPU_OK;
  CK(cudaMalloc((void **)&h->d_res, bytes));
  if (h->gds) { if (int rc = gds_read(h, h->d_res, h->t.rgs[g0].off, bytes)) return rc; }
  else CK(cudaMemcpyAsync(h->d_res, h->image + h->t.rgs[g0].off, bytes, cudaMemcpyHostToDevice, h->s_h2d));
  CK(cudaStreamSynchronize(h->s_h2d));
  h->res_bytes = bytes;
  h->res_g0 = g0;
  h->res_g1 = g1;
  return FLSGPU_OK;
}

int flsgpu_decode(flsgpu_t *h, uint32_t g0, uint32_t g1, const uint32_t *cols, uint32_t n_cols, uint32_t flags,
                  flsgpu_batch_t **out) {
  if (!out) return h->fail(FLSGPU_E_RANGE, "null argument");
  *out = nullptr;
  const auto host_t0 = std::chrono::steady_clock::now();
  double plan_ms = 0;
  const FlsTable &t = h->t;
  if (g0 > g1 || g1 > t.n_rg) return h->fail(FLSGPU_E_RANGE, "bad row-group range");
  if (int rc = ensure_ready(h)) return rc;

  flsgpu_batch *b = new flsgpu_batch();
  b->h = h; b->g0 = g0; b->g1 = g1; b->flags = flags;
  if (cols) b->cols.assign(cols, cols + n_cols);
  else for (uint32_t c = 0; c < t.n_cols; ++c) b->cols.push_back(c);
  for (uint32_t c : b->cols)
    if (c >= t.n_cols) { delete b; return h->fail(FLSGPU_E_RANGE, "bad column id"); }
  const uint32_t nc = (uint32_t)b->cols.size();
  const bool to_host = (flags & FLSGPU_OUT_HOST) != 0;
  memset(&b->tm, 0, sizeof(b->tm));

  // Every exit after this point goes through bail(): work already queued on the handle's streams is drained before
  // the batch's buffers go back to the pool, and borrowed events go back to the free list.
  std::vector<cudaEvent_t> used_events;
  auto bail = [&](int code, const std::string &m) {
    cudaDeviceSynchronize();
    cudaGetLastError();
    for (auto e : used_events) if (e) h->ev_free.push_back(e);
    flsgpu_batch_release(b);
    return h->fail(code, m);
  };
#define CKB(call)                                                                                        \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) return bail(FLSGPU_E_CUDA, std::string("CUDA: " #call ": ") + cudaGetErrorString(e_)); \
  } while (0)

  // ---- the chunks of the range: footer entry + vector headers checked once per handle, before anything derived
  //      from them (sizes below, device addresses in the kernels) is used
  if (h->validated.empty()) h->validated.assign((size_t)t.n_rg * t.n_cols, 0);
  for (uint32_t g = g0; g < g1; ++g)
    for (uint32_t c : b->cols) {
      uint8_t &ok = h->validated[(size_t)g * t.n_cols + c];
      if (ok) continue;
      const std::string e = fls_validate_chunk(t, g, c, h->image, h->size);
      if (!e.empty()) return bail(FLSGPU_E_FORMAT, e);
      ok = 1;
    }

  // ---- output buffers
  b->out.resize(nc); b->dev.resize(nc); b->need.resize(nc); b->parts.resize(nc); b->off64_dev.assign(nc, nullptr);
  // a cached plan of this very decode (range, columns) names the buffers its descriptors point at: ask the pool for them
  const flsgpu::PlanCache *hint = nullptr;
  if (!to_host && !(h->flags & FLSGPU_OPEN_NO_PLAN_CACHE))
    for (auto &pc : h->plans)
      if (pc.valid && pc.g0 == g0 && pc.g1 == g1 && pc.cols == b->cols && pc.ptrs.size() == 2 + 4 * (size_t)nc &&
          (!hint || pc.stamp > hint->stamp))
        hint = &pc;  // (the most recently used one)
  auto hinted = [&](size_t k) -> const void * { return hint ? hint->ptrs[k] : nullptr; };
  const uint64_t rows = t.rows(g0, g1), prow = t.padded_rows(g0, g1);
  uint64_t scratch_total = 0;
  for (uint32_t i = 0; i < nc; ++i) {
    const uint32_t c = b->cols[i];
    FlsColNeed n = fls_column_need(t, c, g0, g1, h->image);
    b->need[i] = n;
    flsgpu_column &o = b->out[i];
    memset(&o, 0, sizeof(o));
    o.type = t.types[c];
    o.width = fls_type_bits(t.types[c]) / 8;
    o.rows = rows; o.padded_rows = prow;
    o.compressed_bytes = n.compressed_bytes; o.decoded_bytes = n.decoded_bytes;
    b->tm.compressed_bytes += n.compressed_bytes; b->tm.decoded_bytes += n.decoded_bytes;
    FlsColOut &d = b->dev[i];
    auto dalloc = [&](size_t sz, const void *prefer = nullptr) { void *p = h->dev_alloc(sz, prefer); b->dev_bufs.push_back(p); return p; };
    auto halloc = [&](size_t sz) { void *p = h->host_alloc(sz); b->host_bufs.push_back(p); return p; };
    if (o.width) {
      if (!(d.data = dalloc(n.data_bytes, hinted(2 + 4 * (size_t)i)))) return bail(FLSGPU_E_NOMEM, "out of device memory");
      o.data = d.data;
      if (to_host && !(o.data = halloc(n.data_bytes))) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
    } else {
      // 4 GiB of payload and more: parts of less than 4 GiB each, 64-bit offsets for the caller
      std::vector<flsgpu_batch::StrPart> &parts = b->parts[i];
      if (n.payload_bytes >= (1ull << 32)) {
        uint64_t in_part = 0, pay = 0, row = 0;
        for (uint32_t g = g0; g < g1; ++g) {
          const uint64_t rp = t.chunk(g, c).aux64[0];
          if (g == g0 || in_part + rp > 0xFFFF0000ull) { parts.push_back({g, row, pay}); in_part = 0; }
          in_part += rp; pay += rp;
          row += (uint64_t)t.rgs[g].n_vec * FLS_VEC;
        }
      }
      d.offsets = (uint32_t *)dalloc(n.offsets_bytes + 16 * parts.size(), hinted(3 + 4 * (size_t)i));
      d.payload = (uint8_t *)dalloc(n.payload_bytes, hinted(4 + 4 * (size_t)i));
      if (!d.offsets || !d.payload) return bail(FLSGPU_E_NOMEM, "out of device memory");
      o.offsets = d.offsets; o.payload = d.payload;
      o.payload_bytes = n.payload_bytes - 16;
      if (!parts.empty()) {
        if (!(b->off64_dev[i] = (uint64_t *)dalloc((prow + 1) * 8))) return bail(FLSGPU_E_NOMEM, "out of device memory");
        o.offsets = nullptr; o.offsets64 = b->off64_dev[i];
      }
      if (to_host) {
        if (parts.empty()) o.offsets = (uint32_t *)halloc(n.offsets_bytes);
        else o.offsets64 = (uint64_t *)halloc((prow + 1) * 8);
        o.payload = (uint8_t *)halloc(n.payload_bytes);
        if (!(o.offsets || o.offsets64) || !o.payload) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
      }
      scratch_total += n.scratch_bytes;
    }
    if (n.validity_bytes) {
      if (!(d.validity = (uint8_t *)dalloc(n.validity_bytes + 16, hinted(5 + 4 * (size_t)i)))) return bail(FLSGPU_E_NOMEM, "out of device memory");
      o.validity = (uint64_t *)d.validity;
      if (to_host && !(o.validity = (uint64_t *)halloc(n.validity_bytes + 16))) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
    }
  }
  if (g1 == g0) {  // an empty range: nothing to decode, but offsets[0] of a string column is still 0
    for (uint32_t i = 0; i < nc; ++i) {
      if (!b->dev[i].offsets) continue;
      CKB(cudaMemset(b->dev[i].offsets, 0, 4));
      if (to_host && b->out[i].offsets) b->out[i].offsets[0] = 0;
    }
    *out = b;
    return FLSGPU_OK;
  }

  // ---- compressed bytes: resident, or staged slice by slice
  const bool resident = h->d_res && g0 >= h->res_g0 && g1 <= h->res_g1;
  uint8_t *d_stage = nullptr;
  if (resident) d_stage = h->d_res + (t.rgs[g0].off - t.rgs[h->res_g0].off);
  else {
    d_stage = (uint8_t *)h->dev_alloc(range_extent(t, g0, g1));
    b->dev_bufs.push_back(d_stage);
    if (!d_stage) return bail(FLSGPU_E_NOMEM, "out of device memory");
  }
  uint8_t *d_scratch = nullptr;
  if (scratch_total) {
    d_scratch = (uint8_t *)h->dev_alloc(scratch_total + 16, hinted(1));
    b->dev_bufs.push_back(d_scratch);
    if (!d_scratch) return bail(FLSGPU_E_NOMEM, "out of device memory");
  }
  uint32_t slice = resident && !to_host ? (g1 - g0) : (h->slice_rgs ? h->slice_rgs : 64u);
  const uint32_t n_slices = (g1 - g0 + slice - 1) / slice;
  // plan cache (single-slice resident decodes only): key = range, columns and every buffer address
  const bool cacheable = n_slices == 1 && resident && !to_host && !(h->flags & FLSGPU_OPEN_NO_PLAN_CACHE);
  std::vector<const void *> key_ptrs;
  if (cacheable) {
    key_ptrs.push_back(d_stage); key_ptrs.push_back(d_scratch);
    for (uint32_t i = 0; i < nc; ++i) { key_ptrs.push_back(b->dev[i].data); key_ptrs.push_back(b->dev[i].offsets); key_ptrs.push_back(b->dev[i].payload); key_ptrs.push_back(b->dev[i].validity); }
  }
  flsgpu::PlanCache *plan = nullptr;
  if (cacheable)
    for (auto &pc : h->plans)
      if (pc.valid && pc.g0 == g0 && pc.g1 == g1 && pc.cols == b->cols && pc.ptrs == key_ptrs) { plan = &pc; break; }
  const bool plan_hit = plan != nullptr;
  // descriptors: at most 2 chunks per (row group, column)
  const size_t desc_cap = (size_t)(g1 - g0) * nc * 2;
  FlsChunkDev *h_desc = nullptr, *d_desc = nullptr;
  if (plan_hit) { d_desc = plan->d_desc; plan->stamp = ++h->plan_clock; }
  else {
    h_desc = (FlsChunkDev *)h->host_alloc(desc_cap * sizeof(FlsChunkDev));
    d_desc = (FlsChunkDev *)h->dev_alloc(desc_cap * sizeof(FlsChunkDev));
    b->host_bufs.push_back(h_desc);
    if (!h_desc || !d_desc) { h->dev_free(d_desc); return bail(FLSGPU_E_NOMEM, "out of memory for chunk descriptors"); }
    if (cacheable) {  // the cache takes the device descriptors over (least recently used entry goes)
      if (h->plans.size() < flsgpu::MAX_PLANS) { h->plans.emplace_back(); plan = &h->plans.back(); }
      else {
        plan = &h->plans[0];
        for (auto &pc : h->plans) if (pc.stamp < plan->stamp) plan = &pc;
        h->dev_free(plan->d_desc);
      }
      plan->valid = false;
      plan->d_desc = d_desc;
      plan->stamp = ++h->plan_clock;
    } else b->dev_bufs.push_back(d_desc);
  }

  struct EvPair { cudaEvent_t a, b; };
  std::vector<EvPair> ev_h2d, ev_k, ev_d2h;
  struct KEv { cudaEvent_t a, b; int slot; uint64_t rd, wr, vec; };
  std::vector<KEv> kev;
  const bool time_kernels = (h->flags & FLSGPU_OPEN_TIME_KERNELS) != 0;
  bool ev_fail = false;
  auto ev = [&]() { cudaEvent_t e = h->ev_get(); if (e) used_events.push_back(e); else ev_fail = true; return e; };
  cudaEvent_t ev_start = ev(), ev_end = ev();
  if (ev_fail) return bail(FLSGPU_E_CUDA, "CUDA: cannot create an event");

  CKB(cudaEventRecord(ev_start, h->s_h2d));
  CKB(cudaStreamWaitEvent(h->s_k, ev_start, 0));
  for (uint32_t i = 0; i < nc; ++i)
    if (b->dev[i].offsets) CKB(cudaMemsetAsync(b->dev[i].offsets, 0, 4, h->s_k));

  size_t desc_used = 0;
  std::vector<uint64_t> pay_done(nc, 0), scratch_done(nc, 0);
  uint64_t scratch_base = 0;
  std::vector<uint64_t> scratch_col_base(nc, 0);
  for (uint32_t i = 0; i < nc; ++i) { scratch_col_base[i] = scratch_base; scratch_base += b->need[i].scratch_bytes; }
  uint64_t rows_done = 0;
  std::string perr;
  for (uint32_t si = 0; si < n_slices; ++si) {
    const uint32_t s0 = g0 + si * slice, s1 = std::min(g1, s0 + slice);
    const uint64_t rel = t.rgs[s0].off - t.rgs[g0].off, ext = range_extent(t, s0, s1);
    const uint64_t srows = t.padded_rows(s0, s1);
    // plan the slice
    const auto plan_t0 = std::chrono::steady_clock::now();
    std::vector<FlsLaunch> ls;
    std::vector<LaunchRec> recs;
    std::vector<uint64_t> pay_slice(nc, 0);
    if (plan_hit) recs = plan->recs;
    for (uint32_t i = 0; i < nc && perr.empty() && !plan_hit; ++i) {
      const uint32_t c = b->cols[i];
      FlsColOut o = b->dev[i];
      const uint32_t wb = b->out[i].width;
      if (o.validity) o.validity += rows_done / 8;
      if (wb) o.data = (uint8_t *)o.data + rows_done * wb;
      else {
        FlsColNeed sn = fls_column_need(t, c, s0, s1, h->image);
        pay_slice[i] = sn.payload_bytes - 16;
        const auto &parts = b->parts[i];
        if (!parts.empty()) {  // a wide column: one planner call per (slice x part), each with its own 32-bit base
          for (size_t pk = 0; pk < parts.size() && perr.empty(); ++pk) {
            const uint32_t pa = std::max(s0, parts[pk].g_first), pb = std::min(s1, pk + 1 < parts.size() ? parts[pk + 1].g_first : g1);
            if (pa >= pb) continue;
            FlsColOut o2 = o;
            const uint64_t row_a = t.padded_rows(g0, pa);
            o2.offsets = b->dev[i].offsets + row_a + 4 * pk;  // (the kernels store offsets 16 bytes at a time: parts keep that alignment)
            o2.payload = b->dev[i].payload + parts[pk].pay_base;
            o2.pay_base0 = 0;
            for (uint32_t g = parts[pk].g_first; g < pa; ++g) o2.pay_base0 += t.chunk(g, c).aux64[0];
            if (o.validity) o2.validity = b->dev[i].validity + row_a / 8;
            FlsColNeed pn = fls_column_need(t, c, pa, pb, h->image);
            o2.scratch = d_scratch ? d_scratch + scratch_col_base[i] + scratch_done[i] : nullptr;
            scratch_done[i] += pn.scratch_bytes;
            if (pa == parts[pk].g_first) CKB(cudaMemsetAsync(o2.offsets, 0, 4, h->s_h2d));  // (ordered before the slice's kernels)
            perr = fls_plan_column(t, c, pa, pb, d_stage + (t.rgs[pa].off - t.rgs[g0].off), h->image, o2, ls);
          }
          continue;
        }
        o.offsets += rows_done;
        o.pay_base0 = pay_done[i];
        o.scratch = d_scratch ? d_scratch + scratch_col_base[i] + scratch_done[i] : nullptr;
        scratch_done[i] += sn.scratch_bytes;
      }
      perr = fls_plan_column(t, c, s0, s1, d_stage + rel, h->image, o, ls);
    }
    if (!perr.empty()) break;
    // dictionary preparation must run before the gathers that read it
    // ... and the longest kernels go first
    auto rank_of = [](int kind) { return kind == FLS_K_STR_FSST_DICTPREP ? 0 : kind == FLS_K_STR_FSST ? 1 : (kind == FLS_K_STR_DICT || kind == FLS_K_STR_DICT_LONG) ? 2 : 3; };
    std::stable_sort(ls.begin(), ls.end(), [&](const FlsLaunch &a, const FlsLaunch &b2) { return rank_of(a.kind) < rank_of(b2.kind); });
    const size_t desc0 = desc_used;
    for (auto &l : ls) {
      if (l.kind == FLS_K_STR_DICT || l.kind == FLS_K_STR_DICT_LONG || l.kind == FLS_K_STR_FSST) fls_sort_by_key(l);
      if (desc_used + l.chunks.size() > desc_cap) { perr = "internal: descriptor overflow"; break; }
      memcpy(h_desc + desc_used, l.chunks.data(), l.chunks.size() * sizeof(FlsChunkDev));
      recs.push_back(LaunchRec{l.kind, l.width, (uint32_t)l.chunks.size(), l.bytes_read, l.bytes_written, l.vectors, 0u, desc_used});
      desc_used += l.chunks.size();
    }
    if (!perr.empty()) break;
    if (cacheable && !plan_hit) {
      plan->g0 = g0; plan->g1 = g1; plan->cols = b->cols; plan->ptrs = key_ptrs; plan->recs = recs;
      plan->valid = true;
    }
    plan_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - plan_t0).count();
    // H2D: compressed bytes (unless resident) + descriptors
    EvPair eh = {ev(), ev()};
    CKB(cudaEventRecord(eh.a, h->s_h2d));
    if (!resident) {
      if (h->gds) {  // GPUDirect Storage: file -> HBM, no host bounce (cuFileRead is synchronous)
        if (int rc = gds_read(h, d_stage + rel, t.rgs[s0].off, ext)) { const std::string m = h->err; return bail(rc, m); }
        b->tm.h2d_bytes += ext;
      } else {
        // only what the projected columns' chunks occupy crosses the bus (the staging buffer keeps the file's layout,
        // so the planner's addresses do not change; the gaps are never read)
        const std::vector<StageRange> ranges = projected_ranges(t, s0, s1, b->cols, ext);
        uint8_t *bounce = nullptr;
        if (h->image_mapped) {  // mapped file: page cache -> page-locked bounce buffer -> HBM
          bounce = (uint8_t *)h->host_alloc(ext);
          if (!bounce) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
          b->host_bufs.push_back(bounce);
        }
        for (const StageRange &rg : ranges) {
          const uint8_t *src = h->image + t.rgs[s0].off + rg.off;
          if (bounce) { parallel_copy(bounce + rg.off, src, rg.len); src = bounce + rg.off; }
          CKB(cudaMemcpyAsync(d_stage + rel + rg.off, src, rg.len, cudaMemcpyHostToDevice, h->s_h2d));
          b->tm.h2d_bytes += rg.len;
        }
      }
    }
    if (!plan_hit)
      CKB(cudaMemcpyAsync(d_desc + desc0, h_desc + desc0, (desc_used - desc0) * sizeof(FlsChunkDev), cudaMemcpyHostToDevice,
                         h->s_h2d));
    CKB(cudaEventRecord(eh.b, h->s_h2d));
    ev_h2d.push_back(eh);
    // decode
    CKB(cudaStreamWaitEvent(h->s_k, eh.b, 0));
    EvPair ek = {ev(), ev()};
    CKB(cudaEventRecord(ek.a, h->s_k));
    // Without per-kernel timing the launches of a slice are independent (different columns) except that a
    // dictionary must be prepared before the gather that reads it: dictprep + str_dict stay on s_k, in order;
    // the long FSST kernel and the bandwidth-bound kernels go to side streams so that one kernel's tail overlaps
    // the next kernel's start.
    bool side_used[3] = {false, false, false};
    uint32_t rr = 1;
    // SM partition: FSST on its own SMs (side stream 0); everything else in order on s_k, the DICT grid last
    bool has_fsst = false, has_other = false;
    for (auto &l : recs) { if (l.kind == FLS_K_STR_FSST) has_fsst = true; else if (l.kind != FLS_K_STR_FSST_DICTPREP) has_other = true; }
    const bool split = !time_kernels && g_split_fsst > 0 && has_fsst && has_other;
    if (split) std::stable_sort(recs.begin(), recs.end(), [](const LaunchRec &a, const LaunchRec &b2) {
      auto rk = [](int kind) { return kind == FLS_K_STR_FSST ? 0 : kind == FLS_K_STR_FSST_DICTPREP ? 1 : (kind == FLS_K_STR_DICT || kind == FLS_K_STR_DICT_LONG) ? 3 : 2; };
      return rk(a.kind) < rk(b2.kind); });
    for (auto &l : recs) {
      KEv ke = {nullptr, nullptr, l.kind * 5 + width_slot(l.width), l.rd, l.wr, l.vec};
      cudaStream_t ks = h->s_k;
      l.cap = 0;
      if (split) {
        if (l.kind == FLS_K_STR_FSST) {
          l.cap = (uint32_t)g_split_fsst;
          ks = h->s_kx[0];
          if (!side_used[0]) { CKB(cudaStreamWaitEvent(ks, ek.a, 0)); side_used[0] = true; }
        } else if (l.kind == FLS_K_STR_DICT || l.kind == FLS_K_STR_DICT_LONG) l.cap = (uint32_t)g_split_dict;
      } else
      if (!time_kernels && recs.size() > 1 && l.kind != FLS_K_STR_FSST_DICTPREP && l.kind != FLS_K_STR_DICT && l.kind != FLS_K_STR_DICT_LONG) {
        const uint32_t si = l.kind == FLS_K_STR_FSST ? 0u : 1u + (rr++ & 1u);
        ks = h->s_kx[si];
        if (!side_used[si]) { CKB(cudaStreamWaitEvent(ks, ek.a, 0)); side_used[si] = true; }
      }
      if (time_kernels) { ke.a = ev(); ke.b = ev(); CKB(cudaEventRecord(ke.a, ks)); }
      if (!launch_kind(l, d_desc + l.doff, ks)) { perr = "internal: no kernel for launch kind"; break; }
      CKB(cudaGetLastError());
      if (time_kernels) CKB(cudaEventRecord(ke.b, ks));
      kev.push_back(ke);
      b->tm.kernel_launches++;
    }
    for (uint32_t si = 0; si < 3u; ++si)
      if (side_used[si]) {
        cudaEvent_t je = ev();
        CKB(cudaEventRecord(je, h->s_kx[si]));
        CKB(cudaStreamWaitEvent(h->s_k, je, 0));
      }
    if (!perr.empty()) break;
    // wide string columns: the parts' 32-bit offsets -> the caller's 64-bit offsets
    for (uint32_t i = 0; i < nc; ++i) {
      const auto &parts = b->parts[i];
      for (size_t pk = 0; pk < parts.size(); ++pk) {
        const uint32_t pa = std::max(s0, parts[pk].g_first), pb = std::min(s1, pk + 1 < parts.size() ? parts[pk + 1].g_first : g1);
        if (pa >= pb) continue;
        const uint64_t row_a = t.padded_rows(g0, pa), n = t.padded_rows(pa, pb) + 1;
        fls_widen_offsets_kernel<<<(unsigned)std::min<uint64_t>((n + 255) / 256, 148 * 16), 256, 0, h->s_k>>>(
            b->dev[i].offsets + row_a + 4 * pk, b->off64_dev[i] + row_a, n, parts[pk].pay_base);
        b->tm.kernel_launches++;
      }
    }
    CKB(cudaGetLastError());
    CKB(cudaEventRecord(ek.b, h->s_k));
    ev_k.push_back(ek);
    // D2H
    if (to_host) {
      CKB(cudaStreamWaitEvent(h->s_d2h, ek.b, 0));
      EvPair ed = {ev(), ev()};
      CKB(cudaEventRecord(ed.a, h->s_d2h));
      for (uint32_t i = 0; i < nc; ++i) {
        const uint32_t wb = b->out[i].width;
        if (b->dev[i].validity) {
          CKB(cudaMemcpyAsync((uint8_t *)b->out[i].validity + rows_done / 8, b->dev[i].validity + rows_done / 8, srows / 8,
                             cudaMemcpyDeviceToHost, h->s_d2h));
          b->tm.d2h_bytes += srows / 8;
        }
        if (wb) {
          CKB(cudaMemcpyAsync((uint8_t *)b->out[i].data + rows_done * wb, (uint8_t *)b->dev[i].data + rows_done * wb,
                             srows * wb, cudaMemcpyDeviceToHost, h->s_d2h));
          b->tm.d2h_bytes += srows * wb;
        } else {
          const uint64_t o0 = rows_done ? rows_done + 1 : 0, o1 = rows_done + srows + 1;
          if (b->off64_dev[i])
            CKB(cudaMemcpyAsync(b->out[i].offsets64 + o0, b->off64_dev[i] + o0, (o1 - o0) * 8, cudaMemcpyDeviceToHost, h->s_d2h));
          else
          CKB(cudaMemcpyAsync(b->out[i].offsets + o0, b->dev[i].offsets + o0, (o1 - o0) * 4, cudaMemcpyDeviceToHost, h->s_d2h));
          if (pay_slice[i])
            CKB(cudaMemcpyAsync(b->out[i].payload + pay_done[i], b->dev[i].payload + pay_done[i], pay_slice[i],
                               cudaMemcpyDeviceToHost, h->s_d2h));
          b->tm.d2h_bytes += (o1 - o0) * 4 + pay_slice[i];
        }
      }
      CKB(cudaEventRecord(ed.b, h->s_d2h));
      ev_d2h.push_back(ed);
    }
    for (uint32_t i = 0; i < nc; ++i) pay_done[i] += pay_slice[i];
    rows_done += srows;
    b->tm.slices++;
  }
  if (!perr.empty()) return bail(FLSGPU_E_FORMAT, perr);
  if (!ev_k.empty()) CKB(cudaStreamWaitEvent(h->s_d2h, ev_k.back().b, 0));
  CKB(cudaEventRecord(ev_end, h->s_d2h));
  cudaError_t se = cudaEventSynchronize(ev_end);
  if (se != cudaSuccess) return bail(FLSGPU_E_CUDA, std::string("CUDA: decode failed: ") + cudaGetErrorString(se));

  float ms = 0;
  cudaEventElapsedTime(&b->tm.ms_total, ev_start, ev_end);
  for (auto &p : ev_h2d) { cudaEventElapsedTime(&ms, p.a, p.b); b->tm.ms_h2d += ms; }
  for (auto &p : ev_k) { cudaEventElapsedTime(&ms, p.a, p.b); b->tm.ms_kernels += ms; }
  for (auto &p : ev_d2h) { cudaEventElapsedTime(&ms, p.a, p.b); b->tm.ms_d2h += ms; }
  if (time_kernels)
    for (auto &k : kev) {
      cudaEventElapsedTime(&ms, k.a, k.b);
      KStat &s = h->kstats[k.slot];
      s.launches++; s.ms += ms; s.rd += k.rd; s.wr += k.wr; s.vectors += k.vec;
    }
  for (auto e : used_events) h->ev_free.push_back(e);
  b->tm.ms_host_plan = (float)plan_ms;
  b->tm.ms_host_total = (float)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
  *out = b;
  return FLSGPU_OK;
#undef CKB
}

// ---------------------------------------------------------------------------------------------------------
// filtered decode: decode into HBM, select on the device, compact, and only then copy out
// ---------------------------------------------------------------------------------------------------------
int flsgpu_decode_filtered(flsgpu_t *h, uint32_t g0, uint32_t g1, const uint32_t *cols, uint32_t n_cols,
                           const flsgpu_predicate *preds, uint32_t n_preds, uint32_t flags, flsgpu_batch_t **out) {
  if (!out) return h->fail(FLSGPU_E_RANGE, "null argument");
  *out = nullptr;
  const FlsTable &t = h->t;
  if (n_preds > FLSGPU_MAX_PREDICATES || (n_preds && !preds)) return h->fail(FLSGPU_E_RANGE, "too many predicates");
  if (g0 > g1 || g1 > t.n_rg) return h->fail(FLSGPU_E_RANGE, "bad row-group range");
  const auto host_t0 = std::chrono::steady_clock::now();
  // projected columns first, then the predicate columns that are not projected
  std::vector<uint32_t> proj;
  if (cols) proj.assign(cols, cols + n_cols);
  else for (uint32_t c = 0; c < t.n_cols; ++c) proj.push_back(c);
  for (uint32_t c : proj) if (c >= t.n_cols) return h->fail(FLSGPU_E_RANGE, "bad column id");
  std::vector<uint32_t> dcols = proj;
  std::vector<uint32_t> pred_slot(n_preds);
  for (uint32_t k = 0; k < n_preds; ++k) {
    const flsgpu_predicate &p = preds[k];
    if (p.col >= t.n_cols || p.op > FLSGPU_CMP_GE) return h->fail(FLSGPU_E_RANGE, "bad predicate");
    const bool is_str = fls_type_bits(t.types[p.col]) == 0;
    if (is_str && (p.op > FLSGPU_CMP_NE || (p.text_len && !p.text))) return h->fail(FLSGPU_E_RANGE, "string predicates are = and <> only");
    auto it = std::find(dcols.begin(), dcols.end(), p.col);
    pred_slot[k] = (uint32_t)(it - dcols.begin());
    if (it == dcols.end()) dcols.push_back(p.col);
  }
  flsgpu_batch_t *b = nullptr;
  if (int rc = flsgpu_decode(h, g0, g1, dcols.data(), (uint32_t)dcols.size(), FLSGPU_OUT_DEVICE, &b)) return rc;
  const bool to_host = (flags & FLSGPU_OUT_HOST) != 0;
  const uint64_t prow = t.padded_rows(g0, g1);
  const uint32_t n_vec = (uint32_t)(prow / FLS_VEC);
  auto bail = [&](int code, const std::string &m) {
    cudaDeviceSynchronize();
    cudaGetLastError();
    flsgpu_batch_release(b);
    return h->fail(code, m);
  };
#define CKF(call)                                                                                        \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) return bail(FLSGPU_E_CUDA, std::string("CUDA: " #call ": ") + cudaGetErrorString(e_)); \
  } while (0)
  for (auto &pp : b->parts)
    if (!pp.empty()) return bail(FLSGPU_E_LIMIT, "a filtered decode cannot take a string column with 4 GiB of payload or more; split the row-group range");
  b->filtered = true;
  b->flags = flags;
  if (!n_vec) {
    for (auto &o : b->out) { o.rows = o.padded_rows = 0; }
    b->out.resize(proj.size()); b->dev.resize(proj.size()); b->cols = proj;
    *out = b;
    return FLSGPU_OK;
  }
  cudaStream_t s = h->s_k;
  auto dalloc = [&](size_t n) { void *p = h->dev_alloc(n); if (p) b->dev_bufs.push_back(p); return p; };
  auto halloc = [&](size_t n) { void *p = h->host_alloc(n); if (p) b->host_bufs.push_back(p); return p; };
  cudaEvent_t ev0 = h->ev_get(), ev1 = h->ev_get();
  if (!ev0 || !ev1) return bail(FLSGPU_E_CUDA, "CUDA: cannot create an event");
  CKF(cudaEventRecord(ev0, s));
  // ---- per-vector meta: valid rows and the logical row id of the vector's first row; predicate constants
  const uint32_t n_str = [&] { uint32_t n = 0; for (uint32_t c : proj) n += fls_type_bits(t.types[c]) == 0; return n; }();
  size_t text_bytes = 0;
  for (uint32_t k = 0; k < n_preds; ++k) text_bytes += (preds[k].text_len + 15) & ~(uint64_t)15;
  const size_t meta_bytes = (size_t)n_vec * 16 + text_bytes + 16;
  uint8_t *h_meta = (uint8_t *)halloc(meta_bytes), *d_meta = (uint8_t *)dalloc(meta_bytes);
  uint32_t *d_mask = (uint32_t *)dalloc((size_t)n_vec * 128);
  uint32_t *d_cnt = (uint32_t *)dalloc((size_t)n_vec * 4 * (1 + n_str));
  uint64_t *d_off = (uint64_t *)dalloc((size_t)n_vec * 8 * (1 + n_str));
  uint64_t *d_tot = (uint64_t *)dalloc(8 * (1 + n_str)), *h_tot = (uint64_t *)halloc(8 * (1 + n_str));
  if (!h_meta || !d_meta || !d_mask || !d_cnt || !d_off || !d_tot || !h_tot) return bail(FLSGPU_E_NOMEM, "out of memory for the selection");
  uint64_t *h_row0 = (uint64_t *)h_meta;
  uint32_t *h_valid = (uint32_t *)(h_meta + (size_t)n_vec * 8);
  uint8_t *h_text = h_meta + (size_t)n_vec * 16;
  {
    uint64_t v = 0, row = 0;
    for (uint32_t g = g0; g < g1; ++g) {
      uint64_t left = t.rgs[g].n_tuples;
      for (uint32_t i = 0; i < t.rgs[g].n_vec; ++i, ++v) {
        const uint32_t take = (uint32_t)std::min<uint64_t>(left, FLS_VEC);
        h_row0[v] = row; h_valid[v] = take;
        row += take; left -= take;
      }
    }
  }
  FlsPredSet ps;
  memset(&ps, 0, sizeof(ps));
  ps.n = n_preds;
  size_t tpos = 0;
  for (uint32_t k = 0; k < n_preds; ++k) {
    const flsgpu_predicate &p = preds[k];
    const FlsColOut &d = b->dev[pred_slot[k]];
    const int ty = t.types[p.col];
    FlsPredDev &q = ps.p[k];
    q.data = d.data; q.offsets = d.offsets; q.payload = d.payload;
    q.validity = (const uint32_t *)d.validity;  // a comparison with NULL is not true: the row goes
    q.width = fls_type_bits(ty) / 8; q.op = p.op; q.i64 = p.i64; q.f64 = p.f64;
    if (!q.width) {
      q.kind = 4;
      if (p.text_len) memcpy(h_text + tpos, p.text, p.text_len);
      q.text = d_meta + (size_t)n_vec * 16 + tpos; q.text_len = (uint32_t)p.text_len;
      if (p.text_len >> 32) q.text_len = 0xFFFFFFFFu;  // longer than any string of a batch: never equal
      tpos += (p.text_len + 15) & ~(uint64_t)15;
    } else if (ty == FLS_T_FLOAT) q.kind = 2;
    else if (ty == FLS_T_DOUBLE) q.kind = 3;
    else if (ty == FLS_T_BOOLEAN || ty == FLS_T_UINT8 || ty == FLS_T_UINT16 || ty == FLS_T_UINT32 || ty == FLS_T_UINT64) q.kind = 1;
    else q.kind = 0;
  }
  CKF(cudaMemcpyAsync(d_meta, h_meta, meta_bytes, cudaMemcpyHostToDevice, s));
  CKF(cudaMemsetAsync(d_tot, 0, 8 * (1 + n_str), s));
  const uint64_t *d_row0 = (const uint64_t *)d_meta;
  const uint32_t *d_valid = (const uint32_t *)(d_meta + (size_t)n_vec * 8);
  const uint32_t grid = (n_vec + 7) / 8;
  fls_filter::mask_kernel<<<grid, 256, 0, s>>>(ps, d_valid, n_vec, d_mask, d_cnt);
  fls_filter::scan_kernel<<<1, 1024, 0, s>>>(d_cnt, n_vec, d_off, d_tot);
  // string columns: bytes of the surviving strings per vector
  {
    uint32_t si = 0;
    for (size_t i = 0; i < proj.size(); ++i) {
      if (b->out[i].width) continue;
      ++si;
      fls_filter::str_bytes_kernel<<<grid, 256, 0, s>>>(b->dev[i].offsets, d_mask, n_vec, d_cnt + (size_t)si * n_vec);
      fls_filter::scan_kernel<<<1, 1024, 0, s>>>(d_cnt + (size_t)si * n_vec, n_vec, d_off + (size_t)si * n_vec, d_tot + si);
    }
  }
  CKF(cudaGetLastError());
  CKF(cudaMemcpyAsync(h_tot, d_tot, 8 * (1 + n_str), cudaMemcpyDeviceToHost, s));
  CKF(cudaStreamSynchronize(s));
  const uint64_t n_sel = h_tot[0];
  // ---- compact every projected column into exactly sized buffers
  std::vector<flsgpu_column> nout(proj.size());
  std::vector<FlsColOut> ndev(proj.size());
  uint64_t d2h = 0;
  b->n_sel = n_sel;
  b->sel_dev = (uint64_t *)dalloc(std::max<uint64_t>(n_sel, 1) * 8);
  if (!b->sel_dev) return bail(FLSGPU_E_NOMEM, "out of device memory");
  fls_filter::compact_rowid_kernel<<<grid, 256, 0, s>>>(d_row0, d_mask, d_off, n_vec, b->sel_dev);
  if (to_host) {
    b->sel_host = (uint64_t *)halloc(std::max<uint64_t>(n_sel, 1) * 8);
    if (!b->sel_host) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
  }
  uint32_t si = 0;
  for (size_t i = 0; i < proj.size(); ++i) {
    flsgpu_column o = b->out[i];
    FlsColOut nd;
    o.rows = o.padded_rows = n_sel;
    if (b->dev[i].validity) {
      const size_t vb = (n_sel + 63) / 64 * 8 + 16;
      if (!(nd.validity = (uint8_t *)dalloc(vb))) return bail(FLSGPU_E_NOMEM, "out of device memory");
      CKF(cudaMemsetAsync(nd.validity, 0, vb, s));
      fls_filter::compact_validity_kernel<<<grid, 256, 0, s>>>((const uint32_t *)b->dev[i].validity, d_mask, d_off, n_vec,
                                                              (unsigned long long *)nd.validity);
      o.validity = (uint64_t *)nd.validity;
      if (to_host) {
        if (!(o.validity = (uint64_t *)halloc(vb))) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
        CKF(cudaMemcpyAsync(o.validity, nd.validity, vb - 16, cudaMemcpyDeviceToHost, s));
        d2h += vb - 16;
      }
    }
    if (o.width) {
      nd.data = dalloc(std::max<uint64_t>(n_sel, 1) * o.width);
      if (!nd.data) return bail(FLSGPU_E_NOMEM, "out of device memory");
      switch (o.width) {
        case 1: fls_filter::compact_fixed_kernel<uint8_t><<<grid, 256, 0, s>>>((const uint8_t *)b->dev[i].data, d_mask, d_off, n_vec, (uint8_t *)nd.data); break;
        case 2: fls_filter::compact_fixed_kernel<uint16_t><<<grid, 256, 0, s>>>((const uint16_t *)b->dev[i].data, d_mask, d_off, n_vec, (uint16_t *)nd.data); break;
        case 4: fls_filter::compact_fixed_kernel<uint32_t><<<grid, 256, 0, s>>>((const uint32_t *)b->dev[i].data, d_mask, d_off, n_vec, (uint32_t *)nd.data); break;
        default: fls_filter::compact_fixed_kernel<uint64_t><<<grid, 256, 0, s>>>((const uint64_t *)b->dev[i].data, d_mask, d_off, n_vec, (uint64_t *)nd.data); break;
      }
      o.data = nd.data;
      o.decoded_bytes = n_sel * o.width;
      if (to_host) {
        if (!(o.data = halloc(std::max<uint64_t>(n_sel, 1) * o.width))) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
        if (n_sel) CKF(cudaMemcpyAsync(o.data, nd.data, n_sel * o.width, cudaMemcpyDeviceToHost, s));
        d2h += n_sel * o.width;
      }
    } else {
      ++si;
      const uint64_t pay = h_tot[si];
      if (pay >> 32) return bail(FLSGPU_E_LIMIT, "string payload of one batch must stay below 4 GiB; split the row-group range");
      nd.offsets = (uint32_t *)dalloc((n_sel + 1) * 4);
      nd.payload = (uint8_t *)dalloc(pay + 16);
      if (!nd.offsets || !nd.payload) return bail(FLSGPU_E_NOMEM, "out of device memory");
      CKF(cudaMemsetAsync(nd.offsets, 0, 4, s));
      fls_filter::compact_str_kernel<<<grid, 256, 0, s>>>(b->dev[i].offsets, b->dev[i].payload, d_mask, d_off, d_off + (size_t)si * n_vec,
                                                         n_vec, nd.offsets, nd.payload);
      o.offsets = nd.offsets; o.payload = nd.payload; o.payload_bytes = pay;
      o.decoded_bytes = (n_sel + 1) * 4 + pay;
      if (to_host) {
        o.offsets = (uint32_t *)halloc((n_sel + 1) * 4);
        o.payload = (uint8_t *)halloc(pay + 16);
        if (!o.offsets || !o.payload) return bail(FLSGPU_E_NOMEM, "out of pinned host memory");
        CKF(cudaMemcpyAsync(o.offsets, nd.offsets, (n_sel + 1) * 4, cudaMemcpyDeviceToHost, s));
        if (pay) CKF(cudaMemcpyAsync(o.payload, nd.payload, pay, cudaMemcpyDeviceToHost, s));
        d2h += (n_sel + 1) * 4 + pay;
      }
    }
    nout[i] = o; ndev[i] = nd;
  }
  if (to_host && n_sel) { CKF(cudaMemcpyAsync(b->sel_host, b->sel_dev, n_sel * 8, cudaMemcpyDeviceToHost, s)); d2h += n_sel * 8; }
  CKF(cudaGetLastError());
  CKF(cudaEventRecord(ev1, s));
  CKF(cudaEventSynchronize(ev1));
  float ms = 0;
  cudaEventElapsedTime(&ms, ev0, ev1);
  h->ev_free.push_back(ev0); h->ev_free.push_back(ev1);
  b->tm.ms_total += ms; b->tm.ms_kernels += ms; b->tm.d2h_bytes = d2h;
  b->tm.kernel_launches += 3 + (uint32_t)proj.size() + 2 * n_str;
  b->tm.ms_host_total = (float)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
  // the full-size decode buffers go back to the pool now; the batch keeps the compacted ones
  {
    std::vector<void *> keep;
    for (void *p : b->dev_bufs) {
      bool full = false;
      for (auto &d : b->dev) full = full || p == d.data || p == (void *)d.offsets || p == (void *)d.payload || p == (void *)d.validity;
      if (full) h->dev_free(p); else keep.push_back(p);
    }
    b->dev_bufs.swap(keep);
  }
  b->out.swap(nout); b->dev.swap(ndev); b->cols = proj;
  b->need.resize(proj.size());
  *out = b;
  return FLSGPU_OK;
#undef CKF
}
int flsgpu_batch_selection(const flsgpu_batch_t *b, const uint64_t **ids, uint64_t *n) {
  if (!b || !ids || !n) return FLSGPU_E_RANGE;
  if (!b->filtered) { *ids = nullptr; *n = b->out.empty() ? 0 : b->out[0].rows; return FLSGPU_OK; }
  *ids = (b->flags & FLSGPU_OUT_HOST) ? b->sel_host : b->sel_dev;
  *n = b->n_sel;
  return FLSGPU_OK;
}

uint32_t flsgpu_batch_num_columns(const flsgpu_batch_t *b) { return (uint32_t)b->out.size(); }
int flsgpu_batch_column(const flsgpu_batch_t *b, uint32_t i, flsgpu_column *out) {
  if (i >= b->out.size() || !out) return FLSGPU_E_RANGE;
  *out = b->out[i];
  return FLSGPU_OK;
}
int flsgpu_batch_timing(const flsgpu_batch_t *b, flsgpu_timing *out) {
  if (!out) return FLSGPU_E_RANGE;
  *out = b->tm;
  return FLSGPU_OK;
}
int flsgpu_batch_fetch(const flsgpu_batch_t *b, uint32_t i, void *values, uint32_t *offsets) {
  flsgpu *h = b->h;
  if (i >= b->out.size()) return h->fail(FLSGPU_E_RANGE, "bad column index");
  const flsgpu_column &o = b->out[i];
  const FlsColOut &d = b->dev[i];
  CK(cudaSetDevice(h->device));
  if (o.width) {
    if (values) CK(cudaMemcpy(values, d.data, o.padded_rows * o.width, cudaMemcpyDeviceToHost));
  } else {
    if (offsets && i < b->off64_dev.size() && b->off64_dev[i]) return h->fail(FLSGPU_E_RANGE, "the column's offsets are 64-bit: flsgpu_batch_fetch_offsets64");
    if (offsets) CK(cudaMemcpy(offsets, d.offsets, (o.padded_rows + 1) * 4, cudaMemcpyDeviceToHost));
    if (values && o.payload_bytes) CK(cudaMemcpy(values, d.payload, o.payload_bytes, cudaMemcpyDeviceToHost));
  }
  return FLSGPU_OK;
}
int flsgpu_batch_fetch_offsets64(const flsgpu_batch_t *b, uint32_t i, uint64_t *offsets64) {
  flsgpu *h = b->h;
  if (i >= b->out.size() || !offsets64 || i >= b->off64_dev.size() || !b->off64_dev[i]) return h->fail(FLSGPU_E_RANGE, "column has no 64-bit offsets");
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpy(offsets64, b->off64_dev[i], (b->out[i].padded_rows + 1) * 8, cudaMemcpyDeviceToHost));
  return FLSGPU_OK;
}
int flsgpu_batch_fetch_validity(const flsgpu_batch_t *b, uint32_t i, uint64_t *bits) {
  flsgpu *h = b->h;
  if (i >= b->out.size() || !bits || !b->dev[i].validity) return h->fail(FLSGPU_E_RANGE, "column has no validity bitmap");
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpy(bits, b->dev[i].validity, (b->out[i].padded_rows + 63) / 64 * 8, cudaMemcpyDeviceToHost));
  return FLSGPU_OK;
}
// Result gather over NVLink (optional; the reference merges per-thread DataChunks on the host,
// /root/reference/src/scanner/scan_fastlanes.cpp:89-120).  Peer access is switched on once per (source, destination)
// pair; without it cudaMemcpyPeerAsync stages through the host and still completes.
int flsgpu_batch_copy_to_device(const flsgpu_batch_t *b, uint32_t i, int32_t dst_device, void *dst_values, void *dst_offsets,
                                float *ms) {
  flsgpu *h = b->h;
  if (i >= b->out.size()) return h->fail(FLSGPU_E_RANGE, "bad column index");
  if (dst_device < 0 || dst_device >= flsgpu_device_count()) return h->fail(FLSGPU_E_RANGE, "no such destination device");
  const flsgpu_column &o = b->out[i];
  const FlsColOut &d = b->dev[i];
  if (!d.data && !d.offsets) return h->fail(FLSGPU_E_RANGE, "the batch holds no device copy of this column");
  CK(cudaSetDevice(h->device));
  if (dst_device != h->device) {
    int can = 0;
    CK(cudaDeviceCanAccessPeer(&can, h->device, dst_device));
    if (can) {
      const cudaError_t e = cudaDeviceEnablePeerAccess(dst_device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
      (void)cudaGetLastError();
    }
  }
  cudaEvent_t ea = nullptr, eb = nullptr;
  if (ms) {
    CK(cudaEventCreate(&ea));
    CK(cudaEventCreate(&eb));
    CK(cudaEventRecord(ea, h->s_d2h));
  }
  auto push = [&](void *dst, const void *src, size_t n) -> cudaError_t {
    if (!dst || !n) return cudaSuccess;
    return cudaMemcpyPeerAsync(dst, dst_device, src, h->device, n, h->s_d2h);
  };
  cudaError_t e = cudaSuccess;
  if (o.width) {
    e = push(dst_values, d.data, (size_t)o.padded_rows * o.width);
  } else {
    const bool wide = i < b->off64_dev.size() && b->off64_dev[i];
    e = wide ? push(dst_offsets, b->off64_dev[i], (size_t)(o.padded_rows + 1) * 8)
             : push(dst_offsets, d.offsets, (size_t)(o.padded_rows + 1) * 4);
    if (e == cudaSuccess) e = push(dst_values, d.payload, (size_t)o.payload_bytes);
  }
  if (e == cudaSuccess && ms) e = cudaEventRecord(eb, h->s_d2h);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->s_d2h);
  if (e == cudaSuccess && ms) e = cudaEventElapsedTime(ms, ea, eb);
  if (ea) cudaEventDestroy(ea);
  if (eb) cudaEventDestroy(eb);
  CK(e);
  return FLSGPU_OK;
}
void flsgpu_batch_release(flsgpu_batch_t *b) {
  if (!b) return;
  for (void *p : b->dev_bufs) b->h->dev_free(p);
  for (void *p : b->host_bufs) b->h->host_free(p);
  delete b;
}

uint32_t flsgpu_num_kernel_stats(const flsgpu_t *h) { return (uint32_t)h->kstats.size(); }
int flsgpu_kernel_stats(const flsgpu_t *h, uint32_t i, flsgpu_kernel_stat *out) {
  if (i >= h->kstats.size() || !out) return FLSGPU_E_RANGE;
  const KStat &s = h->kstats[i];
  out->name = s.name.c_str();
  out->launches = s.launches; out->ms = s.ms; out->bytes_read = s.rd; out->bytes_written = s.wr; out->vectors = s.vectors;
  return FLSGPU_OK;
}
void flsgpu_reset_kernel_stats(flsgpu_t *h) {
  for (auto &s : h->kstats) { s.launches = s.rd = s.wr = s.vectors = 0; s.ms = 0; }
}

int flsgpu_flush_l2(flsgpu_t *h, uint64_t bytes) {
  if (int rc = ensure_ready(h)) return rc;
  bytes = (bytes + 15) & ~15ull;
  if (h->flush_bytes < bytes) {
    if (h->d_flush) CK(cudaFree(h->d_flush));
    h->d_flush = nullptr;
    CK(cudaMalloc((void **)&h->d_flush, bytes));
    h->flush_bytes = bytes;
  }
  static uint32_t v = 0;
  fls_flush_kernel<<<148 * 8, 256, 0, h->s_k>>>((uint4 *)h->d_flush, bytes / 16, ++v);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->s_k));
  return FLSGPU_OK;
}

}  // extern "C"

// ---- selection without allocation (include/flsgpu.h: flsgpu_select_fixed_noalloc) ---------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc.  The selection kernels of flsgpu_decode_filtered run with their scratch in this file-scope array, which
// every GPU that loads the library holds until exit (FLSGPU_NOALLOC_SCRATCH_BYTES = 608.25 KiB).  In host code the
// array's NAME is a host-side stand-in, not a device address: the address each device gave it comes from
// cudaGetSymbolAddress, once per device with that device current, and that is what the kernels get.
static __device__ __align__(256) unsigned char g_noalloc_scratch[FLSGPU_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr uint32_t kNoallocMaxVec = FLSGPU_NOALLOC_MAX_ROWS / FLS_VEC;
// layout of g_noalloc_scratch (bytes); every part starts on a 256-byte boundary
constexpr size_t kNaTotal = 0;                                        // uint64: running total of the scan
constexpr size_t kNaValid = 256;                                      // uint32[n_vec]: real rows of each vector
constexpr size_t kNaCnt = kNaValid + (size_t)kNoallocMaxVec * 4;      // uint32[n_vec]: survivors of each vector
constexpr size_t kNaOff = kNaCnt + (size_t)kNoallocMaxVec * 4;        // uint64[n_vec]: exclusive scan of the counts
constexpr size_t kNaRow0 = kNaOff + (size_t)kNoallocMaxVec * 8;       // uint64[n_vec]: first row of each vector
constexpr size_t kNaMask = kNaRow0 + (size_t)kNoallocMaxVec * 8;      // uint32[n_vec * 32]: one bit per row
static_assert(kNaMask + (size_t)kNoallocMaxVec * 128 == FLSGPU_NOALLOC_SCRATCH_BYTES, "scratch layout and header disagree");
static_assert(kNaCnt % 256 == 0 && kNaOff % 256 == 0 && kNaRow0 % 256 == 0 && kNaMask % 256 == 0, "scratch parts are 256-byte aligned");

constexpr int kNaMaxDev = 64;
std::mutex g_na_mu;
unsigned char *g_na_ptr[kNaMaxDev];  // device address of g_noalloc_scratch on each device (null: not resolved yet)

// valid[v] = real rows of vector v, row0[v] = index of its first row
__global__ void __launch_bounds__(256) noalloc_meta_kernel(uint64_t n_rows, uint32_t n_vec, uint32_t *__restrict__ valid,
                                                           uint64_t *__restrict__ row0) {
  const uint32_t v = blockIdx.x * 256u + threadIdx.x;
  if (v >= n_vec) return;
  const uint64_t first = (uint64_t)v * FLS_VEC;
  valid[v] = (uint32_t)(n_rows - first < FLS_VEC ? n_rows - first : FLS_VEC);
  row0[v] = first;
}
__global__ void noalloc_count_kernel(const uint64_t *__restrict__ total, uint64_t *__restrict__ out) { *out = *total; }

int na_fail(int code, const std::string &m) { g_open_err = m; return code; }
}  // namespace

extern "C" int flsgpu_select_fixed_noalloc(int32_t device, const void *values, int32_t type, uint64_t n_rows, uint32_t op,
                                           int64_t i64, double f64, void *out_values, uint64_t *out_row_ids,
                                           uint64_t *out_count, void *stream) {
#define CKN(call)                                                                                   \
  do {                                                                                              \
    cudaError_t e_ = (call);                                                                        \
    if (e_ != cudaSuccess) return na_fail(FLSGPU_E_CUDA, std::string("CUDA: " #call ": ") + cudaGetErrorString(e_)); \
  } while (0)
  if (device < 0 || device >= kNaMaxDev) return na_fail(FLSGPU_E_RANGE, "flsgpu_select_fixed_noalloc: no such device");
  if (n_rows > FLSGPU_NOALLOC_MAX_ROWS)
    return na_fail(FLSGPU_E_LIMIT, "flsgpu_select_fixed_noalloc: " + std::to_string(n_rows) + " rows exceed FLSGPU_NOALLOC_MAX_ROWS (" +
                                       std::to_string(FLSGPU_NOALLOC_MAX_ROWS) + "): the scratch array is fixed; split the range");
  const uint32_t width = type > FLS_T_INVALID && type < FLS_T_N ? fls_type_bits(type) / 8 : 0;
  if (!width) return na_fail(FLSGPU_E_RANGE, "flsgpu_select_fixed_noalloc: not a fixed-width type");
  if (op > FLSGPU_CMP_GE) return na_fail(FLSGPU_E_RANGE, "flsgpu_select_fixed_noalloc: bad comparison");
  if (!out_count || (n_rows && !values)) return na_fail(FLSGPU_E_RANGE, "flsgpu_select_fixed_noalloc: null argument");
  CKN(cudaSetDevice(device));
  unsigned char *base;
  {
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_ptr[device]) {
      void *p = nullptr;
      CKN(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // this device's copy: the current device is `device`
      g_na_ptr[device] = (unsigned char *)p;
    }
    base = g_na_ptr[device];
  }
  cudaStream_t s = (cudaStream_t)stream;
  uint64_t *d_tot = (uint64_t *)(base + kNaTotal);
  uint32_t *d_valid = (uint32_t *)(base + kNaValid), *d_cnt = (uint32_t *)(base + kNaCnt), *d_mask = (uint32_t *)(base + kNaMask);
  uint64_t *d_off = (uint64_t *)(base + kNaOff), *d_row0 = (uint64_t *)(base + kNaRow0);
  // the scan adds to the total: it is the one word that must start at zero; everything else the kernels read they wrote
  // in this call (valid / row0 / cnt / off / mask entries [0, n_vec))
  CKN(cudaMemsetAsync(d_tot, 0, 8, s));
  const uint32_t n_vec = (uint32_t)((n_rows + FLS_VEC - 1) / FLS_VEC);
  if (n_vec) {
    FlsPredSet ps;
    memset(&ps, 0, sizeof(ps));
    ps.n = 1;
    FlsPredDev &q = ps.p[0];
    q.data = values; q.width = width; q.op = op; q.i64 = i64; q.f64 = f64;
    if (type == FLS_T_FLOAT) q.kind = 2;
    else if (type == FLS_T_DOUBLE) q.kind = 3;
    else if (type == FLS_T_BOOLEAN || type == FLS_T_UINT8 || type == FLS_T_UINT16 || type == FLS_T_UINT32 || type == FLS_T_UINT64) q.kind = 1;
    else q.kind = 0;
    const uint32_t grid = (n_vec + 7) / 8;
    noalloc_meta_kernel<<<(n_vec + 255) / 256, 256, 0, s>>>(n_rows, n_vec, d_valid, d_row0);
    fls_filter::mask_kernel<<<grid, 256, 0, s>>>(ps, d_valid, n_vec, d_mask, d_cnt);
    fls_filter::scan_kernel<<<1, 1024, 0, s>>>(d_cnt, n_vec, d_off, d_tot);
    if (out_row_ids) fls_filter::compact_rowid_kernel<<<grid, 256, 0, s>>>(d_row0, d_mask, d_off, n_vec, out_row_ids);
    if (out_values) switch (width) {
      case 1: fls_filter::compact_fixed_kernel<uint8_t><<<grid, 256, 0, s>>>((const uint8_t *)values, d_mask, d_off, n_vec, (uint8_t *)out_values); break;
      case 2: fls_filter::compact_fixed_kernel<uint16_t><<<grid, 256, 0, s>>>((const uint16_t *)values, d_mask, d_off, n_vec, (uint16_t *)out_values); break;
      case 4: fls_filter::compact_fixed_kernel<uint32_t><<<grid, 256, 0, s>>>((const uint32_t *)values, d_mask, d_off, n_vec, (uint32_t *)out_values); break;
      default: fls_filter::compact_fixed_kernel<uint64_t><<<grid, 256, 0, s>>>((const uint64_t *)values, d_mask, d_off, n_vec, (uint64_t *)out_values); break;
    }
  }
  noalloc_count_kernel<<<1, 1, 0, s>>>(d_tot, out_count);
  CKN(cudaGetLastError());
  return FLSGPU_OK;
#undef CKN
}
