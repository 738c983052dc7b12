"""In-process scan cursor (FlsScanner, what the DuckDB table function sits on) on ONE GPU: seconds per SF10 scan as a
function of decodes in flight per device (--lanes) and row groups per decode call (--batch).  File in /dev/shm.
Usage: python profiles/inproc_sweep.py [rowgroups=916] [passes=3] > out.jsonl"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlane_b200"))
sys.path.insert(0, ROOT)
import __graft_entry__  # noqa: E402

__graft_entry__.build()
import synth  # noqa: E402
import fls_writer as W  # noqa: E402

n_rg = int(sys.argv[1]) if len(sys.argv) > 1 else 916
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 3
img = W.tile_rowgroups(synth.lineitem_image(64, seed=7), n_rg)
path = "/dev/shm/flsgpu_inproc_sweep_%d.fls" % os.getpid()
img.tofile(path)
check = os.path.join(ROOT, "duckdb-fastlane_b200", "ext", "fls_scan_check")
try:
    for lanes, batch in ((1, 32), (2, 32), (3, 32), (4, 32), (2, 16), (3, 16), (4, 16), (2, 64), (4, 8), (6, 8)):
        for env_extra in ({}, ):
            env = dict(os.environ)
            env.update(env_extra)
            r = subprocess.run([check, path, "--gpus", "0", "--lanes", str(lanes), "--batch", str(batch), "--chunk", "65536",
                                "--bench", str(passes)], capture_output=True, text=True, timeout=600, env=env)
            try:
                out = json.loads(r.stdout.strip().splitlines()[-1])
            except Exception:
                out = {"error": (r.stdout + r.stderr)[-300:]}
            out.update({"lanes": lanes, "batch": batch})
            if "sec_mean" in out:
                out["GBps"] = round(out["decoded_bytes"] / out["sec_mean"] / 1e9, 1)
            print(json.dumps(out), flush=True)
finally:
    os.unlink(path)
