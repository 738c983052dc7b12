#!/usr/bin/env python
"""bench.py — read_fls decode throughput on B200 (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload): lineitem-shaped synthetic table, `--rowgroups` row groups
of 65,536 rows PER GPU (default 916 = SF10; N GPUs hold N different shards -> weak scaling, SF10*N in total).
A step = one read_fls pass over the rank's whole shard: every column of every row group decoded.

  value   decoded GB/s with the compressed row groups already resident in HBM (kernels + host planning, timed
          with CUDA events inside libflsgpu on its own streams), whole job = sum over ranks / max-over-ranks time
  e2e     the same pass through the C ABI from pinned HOST memory to pinned HOST memory: H2D of the compressed
          bytes, decode, D2H of the decoded columns, all inside the timed region
  roofline       the dominant kernel's algorithmic bytes / its CUDA-event time vs the measured HBM copy peak
  cpu_baseline   the CPU oracle (oracle/fls_oracle.c, kind "port": the reference's own FastLanes reader is not in
                 /root/reference) on a bounded sample of the same workload, on this box's host cores
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlane_b200"))

import numpy as np  # noqa: E402

GB = 1e9
RG_ROWS = 65536


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default="auto", choices=["auto", "weak", "strong"],
                    help="weak: --rowgroups row groups PER GPU (BASELINE config 4 at N=1: SF10); strong: ONE table of "
                         "--rowgroups row groups sharded over the GPUs (BASELINE config 5: SF100 = 9156); "
                         "auto = weak at N=1, strong at N>1")
    ap.add_argument("--rowgroups", type=int, default=0, help="0 = 916 (SF10) per GPU when weak, 9156 (SF100) in total when strong")
    ap.add_argument("--batch-rowgroups", type=int, default=916,
                    help="row groups per flsgpu_decode call (a rank's shard is scanned in equal batches of at most this)")
    ap.add_argument("--distinct", type=int, default=64, help="distinct row groups generated+encoded, then repeated")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed end-to-end steps (0 = min(steps, 5))")
    ap.add_argument("--cpu-sample-rowgroups", type=int, default=0, help="0 = auto (about 10-30 s of CPU work)")
    ap.add_argument("--inproc", action="store_true",
                    help="ONE process, --gpus devices behind one scan cursor (ext/fls_scan.cpp, the table function's `gpus := [..]`): "
                         "end-to-end line only, pinned host -> host through FlsScanner")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the bit-exactness checks around the timed loops")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", 1))
    if a.scaling == "auto":
        a.scaling = "weak" if world == 1 else "strong"
    if not a.rowgroups:
        a.rowgroups = 916 if a.scaling == "weak" else 9156
    return a


def measured_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full`
    capture of this same command (profiles/traffic.json, written from the .ncu-rep by profiles/ncu_summary.py);
    None if no capture of that kernel at this workload size is on record."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return t.get(kernel)
    except Exception:
        return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons of one GPU, sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.3 and len(r) >= 8] or [r for _, r in self.rows if len(r) >= 8]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[4 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons, "samples": len(rows),
                "power_w_max": max(float(r[2]) for r in rows)}


def dist_setup(args):
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_
        torch.cuda.set_device(local)
        dist_.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_
    return rank, world, local, dist


def bind_near_gpu(local):
    """Run this rank on the CPUs next to its GPU (NVML's affinity mask for the device) BEFORE any pinned buffer is
    allocated, so that the page-locked image and result buffers are first-touched on the GPU's own NUMA node
    (SURVEY.md section 8e).  Best effort: returns the CPU list used, or None when NVML / the cgroup says otherwise."""
    try:
        import pynvml
        import torch
        pr = torch.cuda.get_device_properties(local)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByPciBusId(("%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)).encode())
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (max(ncpu, 1024) + 63) // 64)
        near = {i for i in range(64 * len(words)) if (words[i // 64] >> (i % 64)) & 1}
        cpus = sorted(near & os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return cpus
    except Exception:
        pass
    return None


def shard_of(args, rank, world):
    """(first row group, row groups) of this rank's shard."""
    if args.scaling == "weak":
        return rank * args.rowgroups, args.rowgroups
    import shard
    g0, g1 = shard.rowgroup_range(rank, world, args.rowgroups)
    return g0, g1 - g0


def make_image(args, rank, world, pinned):
    """The rank's shard of the lineitem-shaped table, in (optionally pinned) host memory: `distinct` row groups are
    generated and encoded, then repeated round-robin (row group i of the shard = distinct row group i % distinct).
    Weak scaling: every rank its own data (seeded by rank).  Strong scaling: one table, every rank the same distinct
    set and its own contiguous range of it."""
    import synth
    import fls_writer as W
    _, n = shard_of(args, rank, world)
    d = min(args.distinct, n)
    seed_rank = rank if args.scaling == "weak" else 0
    cols = synth.lineitem(d * RG_ROWS, seed=7 + seed_rank, first_order=seed_rank * (1 << 28))
    small = synth.encode(cols, synth.LINEITEM_NAMES, synth.LINEITEM_TYPES)
    del cols
    keep = []

    def alloc(nb):
        if pinned:
            import torch
            t = torch.empty(nb, dtype=torch.uint8, pin_memory=True)
            keep.append(t)
            return t.numpy()
        return np.empty(nb, dtype=np.uint8)

    if d < n:
        img = W.tile_rowgroups(small, n, pinned_alloc=alloc)
    else:
        img = alloc(small.size)
        img[:] = small
    return img, keep, small


def synth_col(name):
    import synth
    return synth.LINEITEM_NAMES.index(name)


def batches_of(n_rg, cap):
    k = max(1, -(-n_rg // max(cap, 1)))
    return [(i * n_rg // k, (i + 1) * n_rg // k) for i in range(k)]


def oracle_pass(ora, g0, g1, threads):
    """One CPU read_fls pass (every column) over row groups [g0,g1); returns decoded bytes."""
    tot = 0
    for c in range(ora.n_cols):
        d = ora.decode(c, g0, g1, threads=threads, padded=True)
        tot += d.nbytes if not isinstance(d, tuple) else d[0].nbytes + d[1].nbytes
    return tot


class Expect:
    """Oracle decode of the `d` distinct row groups, and the check of a decoded batch of the tiled shard against it
    (row group g of the shard must equal distinct row group g % d: values, string lengths and string bytes)."""

    def __init__(self, small_img, threads):
        from oracle import oracle
        self.ora = oracle.OracleTable(small_img)
        self.d = self.ora.n_rg
        self.cols = [self.ora.decode(c, 0, self.d, threads=threads, padded=True) for c in range(self.ora.n_cols)]
        self.lens, self.paybase = {}, {}
        for c, e in enumerate(self.cols):
            if isinstance(e, tuple):
                self.lens[c] = np.diff(e[0].astype(np.int64)).reshape(self.d, RG_ROWS)
                self.paybase[c] = e[0][::RG_ROWS].astype(np.int64)  # payload offset of each distinct row group (+ the end)

    def check_distinct(self, got):
        """got: host arrays of row groups [0, d) of the shard, padded."""
        for c, e in enumerate(self.cols):
            g = got[c]
            if isinstance(e, tuple):
                if not (np.array_equal(g[0], e[0]) and np.array_equal(g[1][:len(e[1])], e[1])):
                    return "column %d (strings) differs from the oracle" % c
            elif not np.array_equal(g.view(np.uint8), e.view(np.uint8)):
                return "column %d differs from the oracle" % c
        return None

    def check_batch(self, got, g0, g1):
        """got: host arrays (padded) of shard row groups [g0, g1)."""
        d = self.d
        for c, e in enumerate(self.cols):
            g = got[c]
            if isinstance(e, tuple):
                offs, pay = g
                ln = np.diff(offs.astype(np.int64)).reshape(g1 - g0, RG_ROWS)
                for i in range(g1 - g0):
                    k = (g0 + i) % d
                    if not np.array_equal(ln[i], self.lens[c][k]):
                        return "column %d: string lengths of row group %d differ" % (c, g0 + i)
                    a = int(offs[i * RG_ROWS])
                    n = int(self.paybase[c][k + 1] - self.paybase[c][k])
                    if not np.array_equal(pay[a:a + n], e[1][int(self.paybase[c][k]):int(self.paybase[c][k]) + n]):
                        return "column %d: string bytes of row group %d differ" % (c, g0 + i)
            else:
                gv = g.view(np.uint8).reshape(g1 - g0, -1)
                ev = e.view(np.uint8).reshape(d, -1)
                for i in range(g1 - g0):
                    if not np.array_equal(gv[i], ev[(g0 + i) % d]):
                        return "column %d: row group %d differs" % (c, g0 + i)
        return None


def run_reference(args, rank, world, dist):
    """--impl reference: the CPU implementation of the path, all host threads, on a bounded sample per step.

    The reference's own FastLanes reader is absent from /root/reference, so this is the
    oracle port (oracle/fls_oracle.c) — kind "port"."""
    if rank != 0:
        return
    from oracle import oracle
    cores = os.cpu_count() or 1
    sample = args.cpu_sample_rowgroups or (64 if cores <= 32 else 256)
    a = argparse.Namespace(**vars(args))
    a.scaling, a.rowgroups, a.distinct = "weak", sample, min(args.distinct, sample)
    img, _, _ = make_image(a, 0, 1, pinned=False)
    ora = oracle.OracleTable(img)
    for _ in range(args.warmup):
        oracle_pass(ora, 0, ora.n_rg, cores)
    t0 = time.perf_counter()
    nbytes = 0
    for _ in range(args.steps):
        nbytes += oracle_pass(ora, 0, ora.n_rg, cores)
    dt = time.perf_counter() - t0
    val = nbytes / dt / GB
    desc = "%d row groups (%d rows) of the lineitem-shaped shard per step" % (sample, sample * RG_ROWS)
    emit(({
        "impl": "reference", "metric": "read_fls decoded GB/s", "value": val, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "int64/f64/u8", "data": "synthetic",
        "rows_per_s": sample * RG_ROWS * args.steps / dt,
        "config": {"workload": "lineitem-shaped synthetic, 16 columns: a bounded CPU sample of the GPU arm's table "
                               "(a rate in decoded GB/s, so comparable whatever the sample size)",
                   "cpu_sample_rowgroups": sample, "rowgroups_per_step": sample, "scaling": args.scaling},
        "cpu_baseline": {"value": val, "unit": "GB/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def pcie_ceiling(torch, local, d2h_bytes, h2d_bytes, barrier, allmax, world, reps=4):
    """A bare D2H copy with an H2D copy a fraction of its size running the other way (the byte ratio of the scan),
    every rank at once between the same barriers as the e2e loop.  Returns (decoded-equivalent GB/s over all ranks,
    bytes of one D2H copy): what the PCIe / host side would give the scan if decode cost nothing."""
    n_d = int(min(d2h_bytes, 2 << 30))
    n_h = max(1 << 20, int(n_d * (h2d_bytes / max(d2h_bytes, 1))))
    dev = torch.device("cuda", local)
    d_src = torch.empty(n_d, dtype=torch.uint8, device=dev)
    d_dst = torch.empty(n_h, dtype=torch.uint8, device=dev)
    h_dst = torch.empty(n_d, dtype=torch.uint8, pin_memory=True)
    h_src = torch.empty(n_h, dtype=torch.uint8, pin_memory=True)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def once():
        with torch.cuda.stream(s1):
            h_dst.copy_(d_src, non_blocking=True)
        with torch.cuda.stream(s2):
            d_dst.copy_(h_src, non_blocking=True)

    once()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    barrier()
    dt = allmax(time.perf_counter() - t0)
    del d_src, d_dst, h_dst, h_src
    return reps * n_d * world / dt / GB, n_d


def run_ours(args, rank, world, local, dist):
    import torch
    import flsgpu
    assert torch.cuda.is_available(), "bench.py --impl ours needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    t_setup = time.time()
    near = bind_near_gpu(local) if world > 1 else None
    img, keep, small = make_image(args, rank, world, pinned=True)
    setup_s = time.time() - t_setup
    cores = os.cpu_count() or 1

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def allred(x, op):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    allmax = lambda x: allred(x, dist.ReduceOp.MAX) if dist else x   # noqa: E731
    allmin = lambda x: allred(x, dist.ReduceOp.MIN) if dist else x   # noqa: E731
    allsum = lambda x: allred(x, dist.ReduceOp.SUM) if dist else x   # noqa: E731

    # `value` is measured the way a user runs the library: no per-kernel events, so independent kernels of the pass
    # overlap on side streams.  The per-kernel roofline numbers come from a second, instrumented pass below.
    rd = flsgpu.Reader.from_image(img, device=local, flags=flsgpu.OPEN_PINNED_IMAGE)
    n_rg, n_rows = rd.n_rowgroups, rd.n_rows
    ranges = batches_of(n_rg, args.batch_rowgroups)

    # ---------------- parity gate 1: the distinct row groups, GPU (through the C ABI, host output) vs the oracle
    parity_err, expect = None, None
    if not args.no_verify:
        expect = Expect(small, cores)
        got = rd.decode(0, expect.d, padded=True)
        parity_err = expect.check_distinct(got)
        del got
        if parity_err:
            raise SystemExit("bench.py: PARITY FAILURE before timing: " + parity_err)

    # ---------------- HBM-resident: `value`
    t0 = time.perf_counter()
    rd.make_resident()
    for (a, b_) in ranges:
        rd.decode_batch(a, b_, to_host=False).release()   # the cold pass: pools are allocated, plans built and uploaded
    torch.cuda.synchronize()
    cold_first_ms = (time.perf_counter() - t0) * 1e3
    for _ in range(args.warmup):
        for (a, b_) in ranges:
            rd.decode_batch(a, b_, to_host=False).release()
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    w0, c0 = time.perf_counter(), time.time()
    ev_ms, launches, dec_bytes, cmp_bytes, plan_ms, kern_ms = 0.0, 0, 0, 0, 0.0, 0.0
    last = []
    for step in range(args.steps):
        dec_bytes = cmp_bytes = 0
        for (a, b_) in ranges:
            b = rd.decode_batch(a, b_, to_host=False)
            ev_ms += b.timing.ms_total
            plan_ms += b.timing.ms_host_plan
            kern_ms += b.timing.ms_kernels
            launches += b.timing.kernel_launches
            dec_bytes += b.timing.decoded_bytes
            cmp_bytes += b.timing.compressed_bytes
            b.release()
    barrier()
    wall_ms = (time.perf_counter() - w0) * 1e3
    c1 = time.time()
    # one more pass, untimed, whose batches are kept for the check below (keeping the last TIMED pass's batches would hold
    # their buffers while the pass's later calls run, and those calls would then miss their cached plans)
    if expect is not None:
        for (a, b_) in ranges:
            last.append((a, b_, rd.decode_batch(a, b_, to_host=False)))
    clocks = sampler.stop(c0, c1) if sampler else None
    ms_step = allmax(ev_ms / args.steps)          # device time (CUDA events), max over ranks
    wall_step = allmax(wall_ms / args.steps)
    total_dec = allsum(float(dec_bytes))
    total_rows = allsum(float(n_rows))
    value = total_dec / (ms_step * 1e-3) / GB
    # ---------------- parity gate 2: every row group of one more pass of the same calls (device output), vs the tiled expectation
    for (a, b_, b) in last:
        if not parity_err:
            parity_err = expect.check_batch([b.fetch(i, padded=True) for i in range(b.n)], a, b_)
        b.release()
    del last

    # the same pass without the plan cache: descriptors rebuilt and uploaded every time (what a first scan pays)
    rdc = flsgpu.Reader.from_image(img, device=local, flags=flsgpu.OPEN_PINNED_IMAGE | flsgpu.OPEN_NO_PLAN_CACHE)
    rdc.make_resident()
    for (a, b_) in ranges:
        rdc.decode_batch(a, b_, to_host=False).release()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    unc_ev, unc_plan, unc_steps = 0.0, 0.0, max(2, min(args.steps, 5))
    for _ in range(unc_steps):
        for (a, b_) in ranges:
            b = rdc.decode_batch(a, b_, to_host=False)
            unc_ev += b.timing.ms_total
            unc_plan += b.timing.ms_host_plan
            b.release()
    torch.cuda.synchronize()
    unc_wall = (time.perf_counter() - t0) * 1e3 / unc_steps
    rdc.close()

    # ---------------- end to end: pinned host -> H2D -> decode -> D2H -> pinned host
    e2e_steps = args.e2e_steps or min(args.steps, 5)
    rd2 = flsgpu.Reader.from_image(img, device=local, flags=flsgpu.OPEN_PINNED_IMAGE)
    for _ in range(max(1, min(args.warmup, 2))):
        for (a, b_) in ranges:
            rd2.decode_batch(a, b_, to_host=True).release()
    barrier()
    e0 = time.perf_counter()
    e_ev_ms = 0.0
    for step in range(e2e_steps):
        h2d = d2h = 0
        for (a, b_) in ranges:
            b = rd2.decode_batch(a, b_, to_host=True)
            h2d += b.timing.h2d_bytes
            d2h += b.timing.d2h_bytes
            e_ev_ms += b.timing.ms_total
            if step == e2e_steps - 1 and expect is not None and not parity_err and (a, b_) == ranges[-1]:
                e2e_last = (a, b_, b)
            else:
                b.release()
    barrier()
    e2e_ms = allmax((time.perf_counter() - e0) * 1e3 / e2e_steps)   # wall clock: includes every host-side cost
    e2e_val = total_dec / (e2e_ms * 1e-3) / GB
    e2e_dev_ms = allmax(e_ev_ms / e2e_steps)
    # parity gate 3: the host output of the last end-to-end batch
    if expect is not None and not parity_err:
        a, b_, b = e2e_last
        parity_err = expect.check_batch([b.numpy(i, padded=True, copy=False) for i in range(b.n)], a, b_)
        b.release()
    # ---------------- end to end with a selective filter (late materialisation): l_shipdate in a window that keeps about
    #                  1 % of the rows; every column is decoded in HBM, only the surviving rows cross PCIe
    sd = synth_col("l_shipdate")
    with rd2.decode_batch(0, min(n_rg, 8), cols=[sd], to_host=True) as b:
        dates = np.sort(b.numpy(0))
    lo, hi = int(dates[int(len(dates) * 0.495)]), int(dates[int(len(dates) * 0.505)])
    preds = [(sd, ">=", lo), (sd, "<", hi)]
    filt_err = None
    for (a, b_) in ranges[:1]:                      # warm-up + check of the first batch against an unfiltered decode
        with rd2.decode_batch(a, b_, cols=[sd, synth_col("l_orderkey")], to_host=True) as full, \
                rd2.decode_filtered_batch(preds, a, b_, to_host=True) as fb:
            d_all, k_all = full.numpy(0, copy=False), full.numpy(1, copy=False)
            m = (d_all >= lo) & (d_all < hi)
            if not (np.array_equal(fb.numpy(sd, copy=False), d_all[m]) and
                    np.array_equal(fb.numpy(synth_col("l_orderkey"), copy=False), k_all[m])):
                filt_err = "filtered decode differs from the unfiltered decode + mask"
    barrier()
    f0 = time.perf_counter()
    f_rows = f_d2h = 0
    for step in range(e2e_steps):
        f_rows = f_d2h = 0
        for (a, b_) in ranges:
            with rd2.decode_filtered_batch(preds, a, b_, to_host=True) as fb:
                f_rows += int(fb.columns[0].rows)
                f_d2h += int(fb.timing.d2h_bytes)
    barrier()
    filt_ms = allmax((time.perf_counter() - f0) * 1e3 / e2e_steps)
    filt = {"value": total_dec / (filt_ms * 1e-3) / GB, "unit": "GB/s (decoded bytes scanned per second)", "ms_per_step": filt_ms,
            "predicate": "l_shipdate >= %d AND l_shipdate < %d (days since 1970-01-01)" % (lo, hi),
            "selectivity": allsum(float(f_rows)) / max(total_rows, 1.0), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": f_d2h,
            "check": filt_err or "first batch equal to the unfiltered decode + mask (l_shipdate, l_orderkey)"}
    if filt_err and not parity_err:
        parity_err = filt_err
    rd2.close()
    ceiling, ceil_bytes = pcie_ceiling(torch, local, d2h, h2d, barrier, allmax, world)

    # ---------------- roofline of the dominant kernel: the same pass with a CUDA-event pair around every launch
    #                  (kernels serialised on one stream, so each is timed alone), same number of steps
    rd.close()
    rd = flsgpu.Reader.from_image(img, device=local, flags=flsgpu.OPEN_PINNED_IMAGE | flsgpu.OPEN_TIME_KERNELS)
    rd.make_resident()
    for _ in range(args.warmup):
        for (a, b_) in ranges:
            rd.decode_batch(a, b_, to_host=False).release()
    rd.reset_kernel_stats()
    ser_ms = 0.0
    for _ in range(args.steps):
        for (a, b_) in ranges:
            b = rd.decode_batch(a, b_, to_host=False)
            ser_ms += b.timing.ms_total
            b.release()
    kstats = rd.kernel_stats()
    peak, peak_src = peaks()
    roof = None
    if kstats:
        top = max(kstats, key=lambda s: s["ms"])
        per_launch = (top["bytes_read"] + top["bytes_written"]) / top["launches"]
        ach = per_launch / (top["ms"] / top["launches"] * 1e-3) / GB
        tot_ms = sum(s["ms"] for s in kstats)
        tr = measured_traffic(top["name"]) or {}
        roof = {"bound": "hbm", "kernel": top["name"], "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": tr.get("dram_bytes_per_launch") if ranges[0][1] - ranges[0][0] == tr.get("rowgroups") else None,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": per_launch,
                "share_of_step": top["ms"] / tot_ms, "serialised_ms_per_step": ser_ms / args.steps,
                "whole_step": {"achieved": (dec_bytes + cmp_bytes) / (ev_ms / args.steps * 1e-3) / GB,
                               "frac": (dec_bytes + cmp_bytes) / (ev_ms / args.steps * 1e-3) / GB / peak},
                "kernels": [{"name": s["name"], "ms_per_step": s["ms"] / args.steps,
                             "GBps": (s["bytes_read"] + s["bytes_written"]) / (s["ms"] * 1e-3) / GB,
                             "frac": (s["bytes_read"] + s["bytes_written"]) / (s["ms"] * 1e-3) / GB / peak}
                            for s in sorted(kstats, key=lambda s: -s["ms"])]}
    rd.close()

    # ---------------- CPU baseline (rank 0, N = 1 only): the oracle on a bounded sample, 1 thread and all cores
    cpu = None
    sample = 0
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import oracle
        sample = args.cpu_sample_rowgroups or min(n_rg, 64)
        ora = oracle.OracleTable(img)
        oracle_pass(ora, 0, min(4, n_rg), cores)
        t0 = time.perf_counter()
        nb1 = oracle_pass(ora, 0, min(sample, 16), 1)
        v1 = nb1 / (time.perf_counter() - t0) / GB
        reps, t0, nb = 0, time.perf_counter(), 0
        while reps < 3 or (time.perf_counter() - t0 < 5 and reps < 50):
            nb += oracle_pass(ora, 0, sample, cores)
            reps += 1
        vn = nb / (time.perf_counter() - t0) / GB
        cpu = {"value": vn, "unit": "GB/s", "cores": cores, "kind": "port", "single_thread_value": v1,
               "sample": "%d row groups (%d rows) x %d passes, all %d host threads; 1-thread figure on %d row groups"
                         % (sample, sample * RG_ROWS, reps, cores, min(sample, 16))}

    parity_ok = allmin(0.0 if parity_err else 1.0) == 1.0
    if parity_err:
        sys.stderr.write("bench.py: rank %d PARITY FAILURE: %s\n" % (rank, parity_err))
    if rank == 0:
        total_rg = args.rowgroups * world if args.scaling == "weak" else args.rowgroups
        if args.no_verify:
            parity = "not checked (--no-verify)"
        elif parity_ok:
            parity = ("bit-exact: %d distinct row groups x 16 columns GPU vs oracle before timing; every row group of "
                      "one more pass of the same decode calls right after the timed loop (device output, fetched) and of the last end-to-end batch (host output) "
                      "equal to the oracle's decode of the distinct row group it repeats" % min(args.distinct, n_rg))
        else:
            parity = "FAILED"
        emit(({
            "metric": "read_fls decoded GB/s", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "int64/f64/u8", "data": "synthetic", "parity": parity,
            "rows_per_s": total_rows / (ms_step * 1e-3),
            "config": {"workload": ("lineitem-shaped synthetic SF%.0f per GPU (%d row groups x 65536 rows x 16 columns), "
                                    "SF%.0f in total over %d GPU(s)" % (n_rows / 6e6, n_rg, n_rows * world / 6e6, world))
                       if args.scaling == "weak" else
                       ("lineitem-shaped synthetic SF%.0f, ONE table of %d row groups x 65536 rows x 16 columns sharded "
                        "contiguously over %d GPU(s) (%d row groups on rank 0)" % (total_rg * RG_ROWS / 6e6, total_rg, world, n_rg)),
                       "scaling": args.scaling, "rowgroups_total": total_rg, "rowgroups_per_gpu": n_rg, "rows_per_gpu": n_rows,
                       "batches_per_step": len(ranges), "rowgroups_per_decode_call": ranges[0][1] - ranges[0][0],
                       "distinct_rowgroups": min(args.distinct, n_rg), "cpu_sample_rowgroups": sample,
                       "l2": "inputs larger than L2: %.2f GB compressed + %.2f GB decoded per step per GPU"
                             % (cmp_bytes / GB, dec_bytes / GB),
                       "parallelism": "row-group sharding, no collective on the data path"},
            "compressed_bytes_per_step": cmp_bytes, "decoded_bytes_per_step": dec_bytes,
            "wall_ms_per_step": wall_step, "setup_s": setup_s,
            "host_plan_ms_per_step": plan_ms / args.steps, "kernel_ms_per_step": kern_ms / args.steps,
            "cold": {"first_pass_wall_ms": cold_first_ms,
                     "uncached_plan_wall_ms_per_step": unc_wall, "uncached_plan_device_ms_per_step": unc_ev / unc_steps,
                     "uncached_host_plan_ms_per_step": unc_plan / unc_steps,
                     "note": "first_pass = make_resident + the first pass (buffer pools allocated, chunks validated, "
                             "plans built); uncached = warm pools, descriptors rebuilt and uploaded on every pass "
                             "(FLSGPU_OPEN_NO_PLAN_CACHE); `value` reuses cached plans"},
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms, "steps": e2e_steps, "device_ms_per_step": e2e_dev_ms,
                    "pcie_ceiling_GBps": ceiling, "frac": e2e_val / ceiling if ceiling else None,
                    "pcie_ceiling_how": "bare cudaMemcpyAsync D2H of %d bytes with an H2D copy %.3f its size running "
                                        "the other way, all ranks at once, x4, wall clock between barriers"
                                        % (ceil_bytes, h2d / max(d2h, 1)),
                    "cpus_near_gpu": len(near) if near else None,
                    "rows_per_s": total_rows / (e2e_ms * 1e-3), "filtered": filt},
            "roofline": roof, "cpu_baseline": cpu,
        }))
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    if not parity_ok and not args.no_verify:
        raise SystemExit(3)


def run_inproc(args):
    """In-process multi-GPU: the scan cursor the DuckDB table function sits on (FlsScanner, `gpus := [0, .., N-1]`),
    driven by ext/fls_scan_check --bench.  Strong scaling of one SF10 table (916 row groups) unless --rowgroups says
    otherwise: the file is written to /dev/shm, opened once per device (each handle pins its own mapping), batches of
    --batch-rowgroups (capped at 64) row groups go to the devices round-robin and come back in file order."""
    import synth
    import fls_writer as W
    n_rg = 916 if args.rowgroups in (0, 9156) else args.rowgroups   # one SF10 table unless --rowgroups names a size
    small = synth.lineitem_image(min(args.distinct, n_rg), seed=7)
    img = W.tile_rowgroups(small, n_rg) if min(args.distinct, n_rg) < n_rg else small
    path = "/dev/shm/flsgpu_bench_%d.fls" % os.getpid()
    img.tofile(path)
    try:
        check = os.path.join(ROOT, "duckdb-fastlane_b200", "ext", "fls_scan_check")
        batch = min(args.batch_rowgroups, 32)
        r = subprocess.run([check, path, "--gpus", ",".join(str(i) for i in range(args.gpus)), "--lanes", "2", "--batch", str(batch),
                            "--chunk", "65536", "--bench", str(max(args.steps, 1))], capture_output=True, text=True, timeout=3000)
        out = json.loads(r.stdout.strip().splitlines()[-1])
        if "error" in out:
            raise SystemExit("bench.py --inproc: " + out["error"])
    finally:
        os.unlink(path)
    v = out["decoded_bytes"] / out["sec_mean"] / GB
    emit({"metric": "read_fls decoded GB/s", "value": v, "unit": "GB/s", "n_gpus": args.gpus, "steps": out["passes"], "warmup": 1,
          "ms_per_step": out["sec_mean"] * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
          "dtype": "int64/f64/u8", "data": "synthetic", "mode": "inproc",
          "config": {"workload": "lineitem-shaped synthetic, ONE table of %d row groups x 65536 rows x 16 columns scanned through "
                                 "FlsScanner (the table function's cursor) with %d device(s) behind it; file in /dev/shm, "
                                 "host -> H2D -> decode -> D2H -> pinned host, every served run touched" % (n_rg, args.gpus),
                     "rowgroups_total": n_rg, "rowgroups_per_decode_call": batch, "parallelism": "batches round-robin over devices, two decodes in flight per device"},
          "e2e": {"value": v, "unit": "GB/s", "ms_per_step": out["sec_mean"] * 1e3, "best_ms": out["sec_best"] * 1e3,
                  "h2d_bytes_per_step": int(img.size), "d2h_bytes_per_step": int(out["decoded_bytes"])},
          "rows_per_s": out["rows"] / out["sec_mean"], "gpu_launches": None})


_RESULT = sys.stdout


def emit(obj):
    """The ONE JSON line of the contract, on the process's real stdout."""
    _RESULT.write(json.dumps(obj) + "\n")
    _RESULT.flush()


def main():
    global _RESULT
    args = parse_args()
    if os.environ.get("NCCL_DEBUG"):
        # NCCL prints its banner on stdout when NCCL_DEBUG is set: keep the real stdout for the result line and
        # point fd 1 at stderr for everything else
        sys.stdout.flush()
        _RESULT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
    import __graft_entry__ as g
    if args.impl == "reference":  # rank 0 alone runs the CPU arm; the other ranks exit without work
        g.build(load_product=False)   # this arm never loads libflsgpu.so
        run_reference(args, int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), None)
        return
    g.build()
    if args.inproc:
        run_inproc(args)
        return
    rank, world, local, dist = dist_setup(args)
    run_ours(args, rank, world, local, dist)


if __name__ == "__main__":
    main()
