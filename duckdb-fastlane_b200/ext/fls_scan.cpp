// fls_scan.cpp — see fls_scan.hpp.  Everything goes through the C ABI; there is no CPU decode here.
#include "fls_scan.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <future>

namespace flsgpu_host {

namespace {
struct Decoded {
  flsgpu_batch_t *batch = nullptr;
  uint32_t g0 = 0, g1 = 0, lane = 0;
  int rc = 0;
  std::string err;
};
}  // namespace

struct FlsScanner::Impl {
  flsgpu_t *h = nullptr;  // lane 0's handle: schema, row counts, pruning
  // One lane per device of the scan: its own handle (streams, buffer pools, pinned image) and at most ONE decode in
  // flight, so no handle ever has two callers.  Batches go to the lanes round-robin and come back in order: with N
  // devices, N batches are being decoded while the caller consumes one (one host thread +
  // context per device feeding one ordered queue).
  std::vector<flsgpu_t *> lanes;
  std::vector<std::vector<flsgpu_batch_t *>> to_release;  // per lane: served batches, released by the lane's next task
  uint32_t next_lane = 0;
  std::string err;
  std::vector<std::string> names;
  std::vector<int32_t> types;
  std::vector<uint32_t> proj;
  uint32_t n_rg = 0, g_begin = 0, g_end = 0, batch_rgs = 16;
  uint64_t n_rows = 0, rows_in_range = 0, served = 0;
  bool started = false;

  Decoded cur;                     // batch being served
  std::vector<flsgpu_column> cols; // its columns
  uint32_t rg = 0;                 // row group being served (absolute)
  uint64_t rg_row = 0;             // next row inside it
  uint64_t rg_base = 0;            // padded row index of the row group's first row inside the batch
  uint32_t next_g = 0;             // first row group not yet handed to the worker

  // Runs on the worker thread; it is the only caller of the handle while it runs (flsgpu.h: one caller at a time).
  std::vector<ZonePredicate> filt_src;   // keeps the predicates' text alive
  std::vector<flsgpu_predicate> filt;    // device filter (flsgpu_decode_filtered); empty = plain decode
  uint64_t f_row = 0;                    // filtered batches: next row of the batch to serve
  uint64_t scanned = 0, filtered_out = 0;
  Decoded decode(uint32_t lane, std::vector<flsgpu_batch_t *> release_first, uint32_t g0, uint32_t g1) {
    flsgpu_t *h = lanes[lane];  // (shadows the member: this task's handle)
    for (auto *b : release_first) flsgpu_batch_release(b);
    Decoded d;
    d.g0 = g0; d.g1 = g1; d.lane = lane;
    const uint32_t *pc = proj.empty() ? nullptr : proj.data();
    if (!filt.empty())
      d.rc = flsgpu_decode_filtered(h, g0, g1, pc, (uint32_t)proj.size(), filt.data(), (uint32_t)filt.size(), FLSGPU_OUT_HOST, &d.batch);
    else
      d.rc = flsgpu_decode(h, g0, g1, pc, (uint32_t)proj.size(), FLSGPU_OUT_HOST, &d.batch);
    if (d.rc) d.err = flsgpu_last_error(h);
    return d;
  }
  std::vector<uint8_t> skip;       // per row group: pruned by a zone predicate
  uint32_t n_skipped = 0;
  bool skipped(uint32_t g) const { return g < skip.size() && skip[g]; }
  std::deque<std::future<Decoded>> queue;  // decodes in flight, in scan order; at most one per lane
  // hand the next batches to the idle lanes (a served batch `old` of lane `old_lane` is released by that lane's next task)
  void launch_ahead(flsgpu_batch_t *old, uint32_t old_lane) {
    if (old) to_release[old_lane].push_back(old);
    while (queue.size() < lanes.size()) {
      while (next_g < g_end && skipped(next_g)) ++next_g;
      if (next_g >= g_end) break;
      // the next contiguous run of row groups that survive pruning, at most one batch long
      const uint32_t g0 = next_g;
      uint32_t g1 = g0 + 1;
      while (g1 < g_end && g1 - g0 < batch_rgs && !skipped(g1)) ++g1;
      next_g = g1;
      const uint32_t lane = next_lane;
      next_lane = (next_lane + 1) % (uint32_t)lanes.size();
      std::vector<flsgpu_batch_t *> rel;
      rel.swap(to_release[lane]);
      queue.push_back(std::async(std::launch::async, [this, lane, rel, g0, g1] { return decode(lane, rel, g0, g1); }));
    }
  }
  void drain() {
    while (!queue.empty()) {
      Decoded d = queue.front().get();
      queue.pop_front();
      if (d.batch) flsgpu_batch_release(d.batch);
    }
    if (cur.batch) { flsgpu_batch_release(cur.batch); cur.batch = nullptr; }
    for (auto &v : to_release) { for (auto *b : v) flsgpu_batch_release(b); v.clear(); }
  }
  // make `cur` the next batch; false at the end or on error
  bool advance_batch() {
    flsgpu_batch_t *old = cur.batch;
    const uint32_t old_lane = cur.lane;
    cur = Decoded();
    if (queue.empty()) {
      if (old) flsgpu_batch_release(old);
      if (!filt.empty() && getenv("FLS_SCAN_TRACE"))
        fprintf(stderr, "fls_filter: device filter kept %llu of %llu rows\n", (unsigned long long)(scanned - filtered_out),
                (unsigned long long)scanned);
      return false;
    }
    Decoded d = queue.front().get();
    queue.pop_front();
    if (d.rc) {
      err = d.err;
      if (old) to_release[old_lane].push_back(old);
      return false;
    }
    cur = d;
    const uint32_t nc = flsgpu_batch_num_columns(cur.batch);
    cols.resize(nc);
    for (uint32_t i = 0; i < nc; ++i) flsgpu_batch_column(cur.batch, i, &cols[i]);
    for (uint32_t i = 0; i < nc; ++i)
      if (!cols[i].width && !cols[i].offsets) {  // a batch with 4 GiB of strings in one column: the cursor serves 32-bit offsets
        err = "a batch of row groups holds 4 GiB or more of one string column: open the scanner with fewer row groups per batch";
        to_release[cur.lane].push_back(cur.batch);
        cur = Decoded();
        if (old) to_release[old_lane].push_back(old);
        return false;
      }
    rg = cur.g0; rg_row = 0; rg_base = 0; f_row = 0;
    if (!filt.empty()) {
      uint64_t in = 0;
      for (uint32_t g = cur.g0; g < cur.g1; ++g) in += flsgpu_rowgroup_rows(h, g);
      scanned += in;
      filtered_out += in - (nc ? cols[0].rows : 0);
    }
    launch_ahead(old, old_lane);
    return true;
  }
};

FlsScanner::FlsScanner() : p(new Impl()) {}
FlsScanner::~FlsScanner() { close(); }

bool FlsScanner::open(const std::string &path, int device, uint32_t batch_rowgroups) {
  return open(path, std::vector<int>(1, device), batch_rowgroups);
}

bool FlsScanner::open(const std::string &path, const std::vector<int> &devices, uint32_t batch_rowgroups) {
  close();
  if (devices.size() > 1 || (devices.size() == 1 && devices[0] != 0)) {  // named devices must exist (the default, device 0,
    const int n = flsgpu_device_count();                                  // is left to the first decode: bind needs no GPU)
    for (int d : devices)
      if (d < 0 || d >= n) {
        p->err = "CUDA device " + std::to_string(d) + " does not exist (" + std::to_string(n) + " visible)";
        return false;
      }
  }
  for (int device : devices.empty() ? std::vector<int>(1, 0) : devices) {
    flsgpu_opts o = {};
    o.device = device;
    // four pipeline slices per batch: the H2D copy, the kernels and the D2H copy of one batch overlap each other
    o.slice_rowgroups = std::max(1u, (batch_rowgroups ? batch_rowgroups : p->batch_rgs) / 4u);
    flsgpu_t *lane = nullptr;
    const int rc = flsgpu_open(path.c_str(), &o, &lane);
    if (rc != FLSGPU_OK) {
      p->err = flsgpu_last_error(nullptr);
      for (auto *l : p->lanes) flsgpu_close(l);
      p->lanes.clear();
      p->h = nullptr;
      return false;
    }
    p->lanes.push_back(lane);
  }
  p->h = p->lanes[0];
  p->to_release.assign(p->lanes.size(), {});
  p->next_lane = 0;
  p->err.clear();
  uint32_t nc = 0;
  const char *const *names = nullptr;
  const int32_t *types = nullptr;
  flsgpu_schema(p->h, &nc, &names, &types);
  p->names.assign(names, names + nc);
  p->types.assign(types, types + nc);
  p->n_rg = flsgpu_num_rowgroups(p->h);
  p->n_rows = flsgpu_num_rows(p->h);
  p->g_begin = 0; p->g_end = p->n_rg;
  p->rows_in_range = p->n_rows;
  p->proj.clear();
  if (batch_rowgroups) p->batch_rgs = batch_rowgroups;
  p->started = false;
  p->served = 0;
  p->skip.clear(); p->n_skipped = 0;
  p->filt.clear(); p->filt_src.clear(); p->scanned = p->filtered_out = 0;
  return true;
}

void FlsScanner::close() {
  if (!p->h) return;
  p->drain();
  for (auto *l : p->lanes) flsgpu_close(l);
  p->lanes.clear();
  p->h = nullptr;
  p->names.clear(); p->types.clear(); p->proj.clear();
  p->n_rg = 0; p->n_rows = 0; p->served = 0; p->started = false;
}

bool FlsScanner::isOpen() const { return p->h != nullptr; }
uint32_t FlsScanner::deviceCount() const { return (uint32_t)p->lanes.size(); }
const std::string &FlsScanner::error() const { return p->err; }
uint32_t FlsScanner::columnCount() const { return (uint32_t)p->names.size(); }
uint64_t FlsScanner::rowCount() const { return p->n_rows; }
uint32_t FlsScanner::rowGroupCount() const { return p->n_rg; }
const std::vector<std::string> &FlsScanner::columnNames() const { return p->names; }
const std::vector<int32_t> &FlsScanner::columnTypes() const { return p->types; }
uint64_t FlsScanner::rowsServed() const { return p->served; }
uint64_t FlsScanner::rowsInRange() const { return p->rows_in_range; }

bool FlsScanner::setProjection(const std::vector<uint32_t> &cols) {
  if (!p->h || p->started) { p->err = "projection must be set on an open scanner before the first chunk"; return false; }
  for (uint32_t c : cols)
    if (c >= p->names.size()) { p->err = "bad column id"; return false; }
  p->proj = cols;
  return true;
}

void FlsScanner::rewind() {
  if (!p->h) return;
  p->drain();
  p->cur = Decoded();
  p->started = false;
  p->err.clear();
  p->g_begin = 0; p->g_end = p->n_rg;
  p->rows_in_range = p->n_rows;
  p->skip.clear(); p->n_skipped = 0;
  p->next_lane = 0;
  p->scanned = p->filtered_out = 0; p->served = 0;
}

bool FlsScanner::setRowGroupRange(uint32_t g0, uint32_t g1) {
  if (!p->h || p->started || g0 > g1 || g1 > p->n_rg) { p->err = "bad row-group range"; return false; }
  p->g_begin = g0; p->g_end = g1;
  p->skip.clear(); p->n_skipped = 0;
  p->rows_in_range = 0;
  for (uint32_t g = g0; g < g1; ++g) p->rows_in_range += flsgpu_rowgroup_rows(p->h, g);
  return true;
}

uint32_t FlsScanner::pruneRowGroups(const std::vector<ZonePredicate> &preds) {
  if (!p->h || p->started) return 0;
  p->skip.assign(p->n_rg, 0);
  p->n_skipped = 0;
  for (uint32_t g = p->g_begin; g < p->g_end; ++g) {
    for (const ZonePredicate &z : preds) {
      int64_t mn = 0, mx = 0;
      int32_t known = 0;
      if (z.is_double) continue;  // floating-point chunks carry no bounds
      if (z.is_text) {
        int32_t found = 0;
        if (z.op != ZONE_EQ || z.col >= p->names.size() ||
            flsgpu_rowgroup_dict_contains(p->h, g, z.col, (const uint8_t *)z.text.data(), z.text.size(), &found, &known) != FLSGPU_OK ||
            !known || found)
          continue;
        p->skip[g] = 1;
        p->n_skipped++;
        p->rows_in_range -= flsgpu_rowgroup_rows(p->h, g);
        break;
      }
      if (z.col >= p->names.size() || flsgpu_rowgroup_minmax(p->h, g, z.col, &mn, &mx, &known) != FLSGPU_OK || !known) continue;
      bool none = false;  // no value in [mn, mx] satisfies the predicate
      switch (z.op) {
        case ZONE_EQ: none = z.value < mn || z.value > mx; break;
        case ZONE_NE: none = mn == mx && mn == z.value; break;
        case ZONE_LT: none = mn >= z.value; break;
        case ZONE_LE: none = mn > z.value; break;
        case ZONE_GT: none = mx <= z.value; break;
        case ZONE_GE: none = mx < z.value; break;
      }
      if (none) {
        p->skip[g] = 1;
        p->n_skipped++;
        p->rows_in_range -= flsgpu_rowgroup_rows(p->h, g);
        break;
      }
    }
  }
  if (getenv("FLS_SCAN_TRACE"))
    fprintf(stderr, "fls_scan: skipped %u of %u row groups\n", p->n_skipped, p->g_end - p->g_begin);
  return p->n_skipped;
}
uint32_t FlsScanner::rowGroupsSkipped() const { return p->n_skipped; }

bool FlsScanner::setDeviceFilter(const std::vector<ZonePredicate> &preds) {
  if (!p->h || p->started) { p->err = "the device filter must be set on an open scanner before the first chunk"; return false; }
  p->filt.clear();
  p->filt_src = preds;
  for (const ZonePredicate &z : p->filt_src) {
    if (p->filt.size() == FLSGPU_MAX_PREDICATES) break;
    if (z.col >= p->names.size()) continue;
    const int32_t ty = p->types[z.col];
    const bool str_col = ty == FLSGPU_T_STR || ty == FLSGPU_T_BYTE_ARRAY, fp_col = ty == FLSGPU_T_FLOAT || ty == FLSGPU_T_DOUBLE;
    if (z.is_text != str_col || z.is_double != fp_col) continue;           // constant of another kind: leave it to the caller
    if (z.is_text && z.op != ZONE_EQ && z.op != ZONE_NE) continue;
    if (ty == FLSGPU_T_UINT64 && z.value < 0) continue;                    // (an i64 constant cannot say what it means here)
    flsgpu_predicate q = {};
    q.col = z.col; q.op = (uint32_t)z.op;  // ZoneOp and flsgpu_cmp share their order
    q.i64 = z.value; q.f64 = z.dvalue;
    q.text = (const uint8_t *)z.text.data(); q.text_len = z.text.size();
    p->filt.push_back(q);
  }
  return true;
}
uint64_t FlsScanner::rowsFilteredOut() const { return p->filtered_out; }

bool FlsScanner::next(uint64_t max_rows, uint64_t &rows, std::vector<ColumnSlice> &out) {
  rows = 0;
  if (!p->h || !max_rows) return false;
  if (!p->started) {
    p->started = true;
    p->next_g = p->g_begin;
    p->launch_ahead(nullptr, 0);
    if (!p->advance_batch()) return false;
  }
  for (;;) {
    if (!p->cur.batch) return false;
    if (!p->filt.empty()) {  // a filtered batch is one run of surviving rows, no padding, no row-group structure
      const uint64_t have = p->cols.empty() ? 0 : p->cols[0].rows;
      if (p->f_row >= have) {
        if (!p->advance_batch()) return false;
        continue;
      }
      rows = std::min<uint64_t>(max_rows, have - p->f_row);
      out.resize(p->cols.size());
      for (size_t i = 0; i < p->cols.size(); ++i) {
        const flsgpu_column &c = p->cols[i];
        ColumnSlice &s = out[i];
        s.type = c.type; s.width = c.width;
        s.data = c.width ? (const uint8_t *)c.data + p->f_row * c.width : nullptr;
        s.offsets = c.width ? nullptr : c.offsets + p->f_row;
        s.payload = c.width ? nullptr : c.payload;
        s.validity = c.validity; s.validity_bit0 = p->f_row;
      }
      p->f_row += rows;
      p->served += rows;
      return true;
    }
    if (p->rg >= p->cur.g1) {
      if (!p->advance_batch()) return false;
      continue;
    }
    const uint64_t rg_rows = flsgpu_rowgroup_rows(p->h, p->rg);
    if (p->rg_row >= rg_rows) {  // next row group: skip the padding of the last vector
      p->rg_base += (rg_rows + 1023u) / 1024u * 1024u;
      p->rg++;
      p->rg_row = 0;
      continue;
    }
    rows = std::min<uint64_t>(max_rows, rg_rows - p->rg_row);
    const uint64_t at = p->rg_base + p->rg_row;
    out.resize(p->cols.size());
    for (size_t i = 0; i < p->cols.size(); ++i) {
      const flsgpu_column &c = p->cols[i];
      ColumnSlice &s = out[i];
      s.type = c.type; s.width = c.width;
      if (c.width) {
        s.data = (const uint8_t *)c.data + at * c.width;
        s.offsets = nullptr; s.payload = nullptr;
      } else {
        s.data = nullptr;
        s.offsets = c.offsets + at;
        s.payload = c.payload;
      }
      s.validity = c.validity; s.validity_bit0 = at;
    }
    p->rg_row += rows;
    p->served += rows;
    return true;
  }
}

}  // namespace flsgpu_host
