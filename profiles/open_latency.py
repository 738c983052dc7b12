"""Fixed cost of one query on one scanning thread: flsgpu_open(path) -> first decode of a small range -> close, repeated
in one process (what the DuckDB table function pays per thread and query once CUDA is up).  File: lineitem-shaped SF1
(92 row groups) in /dev/shm.  Prints one JSON line: median milliseconds of open / first decode / close, and the same
for a handle that decodes 8 row groups of every column.
Usage: [FLSGPU_LIB=...] python profiles/open_latency.py [repeats=20]"""
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

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
img = W.tile_rowgroups(synth.lineitem_image(64, seed=7), 92)
path = "/dev/shm/flsgpu_open_latency_%d.fls" % os.getpid()
img.tofile(path)


def med(v):
    v = sorted(v)
    return round(v[len(v) // 2], 3)


def run(cols, g1):
    t_open, t_dec, t_close = [], [], []
    for i in range(reps + 2):
        t0 = time.perf_counter()
        rd = flsgpu.Reader.open(path)
        t1 = time.perf_counter()
        b = rd.decode_batch(0, g1, cols=cols, to_host=True)
        t2 = time.perf_counter()
        b.release()
        rd.close()
        t3 = time.perf_counter()
        if i >= 2:  # the first handles bring CUDA up and fill the buffer cache
            t_open.append((t1 - t0) * 1e3)
            t_dec.append((t2 - t1) * 1e3)
            t_close.append((t3 - t2) * 1e3)
    return {"open_ms": med(t_open), "first_decode_ms": med(t_dec), "close_ms": med(t_close)}


try:
    out = {"file": "lineitem-shaped SF1, 92 row groups, /dev/shm", "repeats": reps,
           "one_column_8_rowgroups": run([4], 8), "all_columns_8_rowgroups": run(None, 8),
           "one_column_all_rowgroups": run([4], 92)}
    print(json.dumps(out))
finally:
    os.unlink(path)
