"""Per-column decode time of the lineitem-shaped shard, HBM-resident, one column per decode call, per-kernel CUDA
events (FLSGPU_OPEN_TIME_KERNELS).  Usage: python profiles/colbench.py [rowgroups] [passes] > out.jsonl"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlane_b200"))
sys.path.insert(0, ROOT)
import __graft_entry__  # noqa: E402

if not os.environ.get("FLSGPU_LIB"):
    __graft_entry__.build()
import flsgpu  # noqa: E402
import synth  # noqa: E402
import fls_writer as W  # noqa: E402

n_rg = int(sys.argv[1]) if len(sys.argv) > 1 else 916
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 5
only = sys.argv[3].split(",") if len(sys.argv) > 3 else None
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
small = synth.lineitem_image(64, seed=7)
img = W.tile_rowgroups(small, n_rg)
with flsgpu.Reader.from_image(img, device=0, flags=flsgpu.OPEN_TIME_KERNELS) as rd:
    rd.make_resident()
    for c in range(rd.n_cols):
        if only and rd.names[c] not in only:
            continue
        for _ in range(2):
            rd.decode_batch(0, rd.n_rowgroups, cols=[c]).release()
        rd.reset_kernel_stats()
        for _ in range(passes):
            rd.flush_l2()
            rd.decode_batch(0, rd.n_rowgroups, cols=[c]).release()
        for s in rd.kernel_stats():
            ms = s["ms"] / passes
            by = (s["bytes_read"] + s["bytes_written"]) / passes
            print(json.dumps({"col": rd.names[c], "kernel": s["name"], "ms": round(ms, 4), "read_MB": round(s["bytes_read"] / passes / 1e6, 1),
                              "written_MB": round(s["bytes_written"] / passes / 1e6, 1), "GBps": round(by / ms / 1e6, 1),
                              "frac": round(by / ms / 1e6 / peak, 3)}), flush=True)
