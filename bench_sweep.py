#!/usr/bin/env python
"""bench_sweep.py — BASELINE.json configs 2 and 3 on one B200 (not the driver's bench line; see bench.py for that).

  config 2  FFOR bit-unpack sweep: 2^28 int32 / int64 values at every bit width 0..32 / 0..64
  config 3  ALP (decimal-like prices, 0.1 % exceptions) and ALP-RD (high-entropy doubles), 2^28 values

One JSON line per case on stdout: decoded GB/s, Gvalues/s and the fraction of the HBM roofline
(compressed bytes read + decoded bytes written per launch / CUDA-event time of the kernel, against
MEASURED_PEAKS.json).  Inputs are HBM-resident and far larger than L2 (1-2 GB decoded per pass).  The distinct
data is 64 row groups (4 Mi values) per case, tiled to 4096 row groups; each case is first checked bit-for-bit
against the CPU oracle on the distinct part.

  python bench_sweep.py [--widths 0,1,7,32] [--values 268435456] [--steps 5] [--skip-alp] [--skip-ffor]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlane_b200"))

import numpy as np  # noqa: E402

RG_ROWS = 65536


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def run_case(name, col, ftype, force, n_values, steps, extra):
    import flsgpu
    import fls_writer as W
    from oracle import oracle
    small = W.write_table([col], names=["v"], types=[ftype], force=[force] if force is not None else None)
    ora = oracle.OracleTable(small)
    n_rg = max(ora.n_rg, n_values // RG_ROWS)
    img = W.tile_rowgroups(small, n_rg) if n_rg > ora.n_rg else small
    with flsgpu.Reader.from_image(img, device=0, flags=flsgpu.OPEN_TIME_KERNELS) as rd:
        # parity on the distinct row groups first
        with rd.decode_batch(0, ora.n_rg, to_host=True) as b:
            got = b.numpy(0)
        exp = ora.decode(0)
        assert np.array_equal(got.view(np.uint8), exp.view(np.uint8)), name + ": GPU != oracle"
        rd.make_resident()
        for _ in range(3):
            rd.decode_batch(to_host=False).release()
        rd.reset_kernel_stats()
        for _ in range(steps):
            b = rd.decode_batch(to_host=False)
            dec, cmp_ = b.timing.decoded_bytes, b.timing.compressed_bytes
            b.release()
        ks = rd.kernel_stats()
    top = max(ks, key=lambda s: s["ms"])
    ms = top["ms"] / top["launches"]
    gbs = (top["bytes_read"] + top["bytes_written"]) / top["launches"] / (ms * 1e-3) / 1e9
    out = {"case": name, "kernel": top["name"], "values": n_rg * RG_ROWS, "ms": ms,
           "decoded_GBps": dec / (ms * 1e-3) / 1e9, "Gvalues_per_s": n_rg * RG_ROWS / (ms * 1e-3) / 1e9,
           "compressed_bytes": cmp_, "decoded_bytes": dec, "hbm_GBps": gbs, "roofline_frac": gbs / peak(),
           "parity": "bit-exact vs oracle on %d distinct row groups" % ora.n_rg}
    out.update(extra)
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--values", type=int, default=1 << 28)
    ap.add_argument("--distinct-rowgroups", type=int, default=64)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--widths", default="")
    ap.add_argument("--skip-ffor", action="store_true")
    ap.add_argument("--skip-alp", action="store_true")
    a = ap.parse_args()
    import __graft_entry__ as g
    g.build()
    import synth
    import fls_writer as W
    n = a.distinct_rowgroups * RG_ROWS
    only = [int(x) for x in a.widths.split(",")] if a.widths else None
    if not a.skip_ffor:
        for bits, ftype in ((32, W.INT32), (64, W.INT64)):
            for w in range(bits + 1):
                if only is not None and w not in only:
                    continue
                run_case("ffor_i%d_w%d" % (bits, w), synth.ffor_values(bits, w, n), ftype, W.FFOR, a.values, a.steps,
                         {"bits": bits, "width": w})
    if not a.skip_alp:
        run_case("alp_prices", synth.alp_prices(n), W.DOUBLE, W.ALP, a.values, a.steps, {})
        run_case("alprd_entropy", synth.alp_entropy(n), W.DOUBLE, W.ALP_RD, a.values, a.steps, {})


if __name__ == "__main__":
    main()
