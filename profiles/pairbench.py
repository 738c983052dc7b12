"""Kernel times of ONE decode call over a set of columns of the lineitem-shaped SF10 shard (HBM-resident, per-kernel CUDA
events).  Usage: [FLSGPU_LIB=..] python profiles/pairbench.py l_shipinstruct,l_shipmode [rowgroups=916] [passes=10]"""
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

names = sys.argv[1].split(",")
n_rg = int(sys.argv[2]) if len(sys.argv) > 2 else 916
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 10
img = W.tile_rowgroups(synth.lineitem_image(64, seed=7), n_rg)
with flsgpu.Reader.from_image(img, device=0, flags=flsgpu.OPEN_TIME_KERNELS) as rd:
    cols = [rd.names.index(n) for n in names]
    rd.make_resident()
    for _ in range(2):
        rd.decode_batch(0, rd.n_rowgroups, cols=cols).release()
    rd.reset_kernel_stats()
    for _ in range(passes):
        rd.flush_l2()
        rd.decode_batch(0, rd.n_rowgroups, cols=cols).release()
    print(json.dumps({"cols": names, "kernels": {s["name"]: round(s["ms"] / passes, 4) for s in rd.kernel_stats()}}))
