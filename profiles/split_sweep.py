"""Whole-pass time of the lineitem-shaped SF10 shard (HBM-resident, all 16 columns in one decode call) as a function
of the SM partition FLSGPU_SM_SPLIT="fsst,dict" (csrc/flsgpu.cu: the FSST grid takes `fsst` SMs, the bandwidth-bound
kernels and then the DICT grid of `dict` blocks run on the others).  "0" = no partition.
Usage: python profiles/split_sweep.py [rowgroups] [passes] [split ...] > out.jsonl"""
import json
import os
import sys
import time

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
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 10
splits = sys.argv[3:] or ["0", "70,78", "76,72", "80,68", "84,64", "90,58", "100,48", "76,148", "84,148"]
small = synth.lineitem_image(64, seed=7)
img = W.tile_rowgroups(small, n_rg)
for sp in splits:
    os.environ["FLSGPU_SM_SPLIT"] = sp
    with flsgpu.Reader.from_image(img, device=0) as rd:
        rd.make_resident()
        for _ in range(3):
            rd.decode_batch(0, rd.n_rowgroups).release()
        ms = []
        for _ in range(passes):
            rd.flush_l2()
            b = rd.decode_batch(0, rd.n_rowgroups)
            ms.append(float(b.timing.ms_kernels))
            b.release()
        ms.sort()
        print(json.dumps({"split": sp, "ms_median": round(ms[len(ms) // 2], 4), "ms_min": round(ms[0], 4), "ms_max": round(ms[-1], 4)}),
              flush=True)
