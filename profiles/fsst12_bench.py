"""STR_FSST12 (12-bit codes) against STR_FSST on the same strings: the lineitem-shaped l_comment column, 64 distinct
row groups tiled to `rowgroups`, HBM-resident, per-kernel CUDA events.  Usage: python profiles/fsst12_bench.py [rowgroups=916] [passes=5]"""
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
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
comment = synth.lineitem(64 * 65536, seed=7)[15]
for name, scheme in (("str_fsst", W.STR_FSST), ("str_fsst12", W.STR_FSST12)):
    small = W.write_table([comment], names=["l_comment"], force=[scheme])
    img = W.tile_rowgroups(small, n_rg)
    with flsgpu.Reader.from_image(img, device=0, flags=flsgpu.OPEN_TIME_KERNELS) as rd:
        rd.make_resident()
        for _ in range(2):
            rd.decode_batch(0, rd.n_rowgroups).release()
        rd.reset_kernel_stats()
        for _ in range(passes):
            rd.flush_l2()
            rd.decode_batch(0, rd.n_rowgroups).release()
        for s in rd.kernel_stats():
            ms = s["ms"] / passes
            by = (s["bytes_read"] + s["bytes_written"]) / passes
            print(json.dumps({"scheme": name, "kernel": s["name"], "ms": round(ms, 4), "compressed_MB": round(s["bytes_read"] / passes / 1e6, 1),
                              "written_MB": round(s["bytes_written"] / passes / 1e6, 1), "GBps": round(by / ms / 1e6, 1),
                              "frac": round(by / ms / 1e6 / peak, 3)}), flush=True)
