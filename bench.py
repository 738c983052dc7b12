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
    ap.add_argument("--rowgroups", type=int, default=916, help="row groups per GPU (916 = lineitem SF10)")
    ap.add_argument("--distinct", type=int, default=64, help="distinct row groups generated+encoded, then repeated")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed end-to-end steps (0 = min(steps, 5))")
    ap.add_argument("--cpu-sample-rowgroups", type=int, default=0, help="0 = auto (about 10-30 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


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


def make_image(args, rank, pinned):
    """The rank's shard: `rowgroups` row groups of lineitem-shaped data, in (optionally pinned) host memory."""
    import synth
    import fls_writer as W
    d = min(args.distinct, args.rowgroups)
    cols = synth.lineitem(d * RG_ROWS, seed=7 + rank, first_order=rank * (1 << 28))
    small = synth.encode(cols, synth.LINEITEM_NAMES, synth.LINEITEM_TYPES)
    del cols
    keep = []

    def alloc(n):
        if pinned:
            import torch
            t = torch.empty(n, dtype=torch.uint8, pin_memory=True)
            keep.append(t)
            return t.numpy()
        return np.empty(n, dtype=np.uint8)

    if d < args.rowgroups:
        img = W.tile_rowgroups(small, args.rowgroups, pinned_alloc=alloc)
    else:
        img = alloc(small.size)
        img[:] = small
    return img, keep


def oracle_pass(ora, g0, g1, threads):
    """One CPU read_fls pass (every column) over row groups [g0,g1); returns decoded bytes."""
    tot = 0
    for c in range(ora.n_cols):
        d = ora.decode(c, g0, g1, threads=threads, padded=True)
        tot += d.nbytes if not isinstance(d, tuple) else d[0].nbytes + d[1].nbytes
    return tot


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
    a.rowgroups, a.distinct = sample, min(args.distinct, sample)
    img, _ = make_image(a, 0, pinned=False)
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
        "scaling": "weak", "vs_baseline": None, "dtype": "int64/f64/u8", "data": "synthetic",
        "rows_per_s": sample * RG_ROWS * args.steps / dt,
        "config": {"workload": "lineitem-shaped synthetic, 16 columns, bounded CPU sample", "rowgroups_per_step": sample},
        "cpu_baseline": {"value": val, "unit": "GB/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def run_ours(args, rank, world, local, dist):
    import torch
    import flsgpu
    assert torch.cuda.is_available(), "bench.py --impl ours needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    t_setup = time.time()
    near = bind_near_gpu(local) if world > 1 else None
    img, keep = make_image(args, rank, pinned=True)
    setup_s = time.time() - t_setup

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # `value` is measured the way a user runs the library: no per-kernel events, so independent kernels of the pass
    # overlap on side streams.  The per-kernel roofline numbers come from a second, instrumented pass below.
    rd = flsgpu.Reader.from_image(img, device=local, flags=flsgpu.OPEN_PINNED_IMAGE)
    n_rg, n_rows = rd.n_rowgroups, rd.n_rows

    # ---------------- HBM-resident: `value`
    rd.make_resident()
    for _ in range(args.warmup):
        rd.decode_batch(to_host=False).release()
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    w0, c0 = time.perf_counter(), time.time()
    ev_ms, launches, dec_bytes, cmp_bytes, plan_ms, kern_ms = 0.0, 0, 0, 0, 0.0, 0.0
    for _ in range(args.steps):
        b = rd.decode_batch(to_host=False)
        ev_ms += b.timing.ms_total
        plan_ms += b.timing.ms_host_plan
        kern_ms += b.timing.ms_kernels
        launches += b.timing.kernel_launches
        dec_bytes, cmp_bytes = b.timing.decoded_bytes, b.timing.compressed_bytes
        b.release()
    barrier()
    wall_ms = (time.perf_counter() - w0) * 1e3
    c1 = time.time()
    clocks = sampler.stop(c0, c1) if sampler else None
    ms_step = allmax(ev_ms / args.steps)          # device time (CUDA events), max over ranks
    wall_step = allmax(wall_ms / args.steps)
    total_dec = allsum(float(dec_bytes))
    total_rows = allsum(float(n_rows))
    value = total_dec / (ms_step * 1e-3) / GB

    # ---------------- end to end: pinned host -> H2D -> decode -> D2H -> pinned host
    e2e_steps = args.e2e_steps or min(args.steps, 5)
    rd2 = flsgpu.Reader.from_image(img, device=local, flags=flsgpu.OPEN_PINNED_IMAGE)
    for _ in range(max(1, min(args.warmup, 2))):
        rd2.decode_batch(to_host=True).release()
    barrier()
    e0 = time.perf_counter()
    h2d = d2h = 0
    e_ev_ms = 0.0
    for _ in range(e2e_steps):
        b = rd2.decode_batch(to_host=True)
        h2d, d2h = b.timing.h2d_bytes, b.timing.d2h_bytes
        e_ev_ms += b.timing.ms_total
        b.release()
    barrier()
    e2e_ms = allmax((time.perf_counter() - e0) * 1e3 / e2e_steps)   # wall clock: includes every host-side cost
    e2e_val = total_dec / (e2e_ms * 1e-3) / GB
    e2e_dev_ms = allmax(e_ev_ms / e2e_steps)
    rd2.close()

    # ---------------- roofline of the dominant kernel: the same pass with a CUDA-event pair around every launch
    #                  (kernels serialised on one stream, so each is timed alone), same number of steps
    rd.close()
    rd = flsgpu.Reader.from_image(img, device=local, flags=flsgpu.OPEN_PINNED_IMAGE | flsgpu.OPEN_TIME_KERNELS)
    rd.make_resident()
    for _ in range(args.warmup):
        rd.decode_batch(to_host=False).release()
    rd.reset_kernel_stats()
    ser_ms = 0.0
    for _ in range(args.steps):
        b = rd.decode_batch(to_host=False)
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
        roof = {"bound": "hbm", "kernel": top["name"], "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": (measured_traffic(top["name"]) or {}).get("dram_bytes_per_launch")
                if n_rg == (measured_traffic(top["name"]) or {}).get("rowgroups") else None,
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
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import oracle
        cores = os.cpu_count() or 1
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

    if rank == 0:
        emit(({
            "metric": "read_fls decoded GB/s", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64/f64/u8", "data": "synthetic",
            "rows_per_s": total_rows / (ms_step * 1e-3),
            "config": {"workload": "lineitem-shaped synthetic SF%.0f per GPU (%d row groups x 65536 rows x 16 columns), "
                                   "SF%.0f in total over %d GPU(s), row groups sharded contiguously per rank"
                                   % (n_rows / 6e6, n_rg, n_rows * world / 6e6, world),
                       "rowgroups_per_gpu": n_rg, "rows_per_gpu": n_rows, "distinct_rowgroups": min(args.distinct, n_rg),
                       "l2": "inputs larger than L2: %.2f GB compressed + %.2f GB decoded per step per GPU"
                             % (cmp_bytes / GB, dec_bytes / GB),
                       "parallelism": "row-group sharding, no collective on the data path"},
            "compressed_bytes_per_step": cmp_bytes, "decoded_bytes_per_step": dec_bytes,
            "wall_ms_per_step": wall_step, "setup_s": setup_s,
            "host_plan_ms_per_step": plan_ms / args.steps, "kernel_ms_per_step": kern_ms / args.steps,
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms, "steps": e2e_steps, "device_ms_per_step": e2e_dev_ms,
                    "cpus_near_gpu": len(near) if near else None,
                    "rows_per_s": total_rows / (e2e_ms * 1e-3)},
            "roofline": roof, "cpu_baseline": cpu,
        }))
    if dist:
        dist.barrier()
        dist.destroy_process_group()


_RESULT = sys.stdout


def emit(obj):
    """The ONE JSON line of the contract, on the process's real stdout."""
    _RESULT.write(json.dumps(obj) + "\n")
    _RESULT.flush()


def main():
    global _RESULT
    args = parse_args()
    # Libraries (NCCL prints its version banner when NCCL_DEBUG is set on the box) must not add lines to stdout:
    # keep the real stdout for the result line and point fd 1 at stderr for everything else.
    sys.stdout.flush()
    _RESULT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import __graft_entry__ as g
    g.build()
    if args.impl == "reference":  # rank 0 alone runs the CPU arm; the other ranks exit without work
        run_reference(args, int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), None)
        return
    rank, world, local, dist = dist_setup(args)
    run_ours(args, rank, world, local, dist)


if __name__ == "__main__":
    main()
