"""CPU: the footer parser and the header-only bound reader never crash on damaged files.  A table function's bind
runs on whatever bytes the user points it at (/root/reference/src/scan_fastlanes.cpp:36-45 opens the file before
anything else), so every truncation and every byte flip must end in a clean error (or in a file that still parses),
never in a fault.  The damaged images are opened in a child process: a crash there is a failed test here, not a dead
test run."""
import os
import subprocess
import sys

import numpy as np

import synth

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

CHILD = r"""
import sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import flsgpu
img = np.fromfile(sys.argv[3], dtype=np.uint8)
mode, seed, n = sys.argv[4], int(sys.argv[5]), int(sys.argv[6])
rng = np.random.default_rng(seed)
size = len(img)
foot_off = int(img[size - 24:size - 16].view(np.uint64)[0])
ok = bad = 0
for it in range(n):
    x = img.copy()
    if mode == "truncate":
        x = x[: int(rng.integers(0, size))]
    elif mode == "footer":
        for _ in range(int(rng.integers(1, 4))):
            x[int(rng.integers(foot_off, size))] = int(rng.integers(0, 256))
    elif mode == "footer_wide":  # 8 aligned bytes replaced by a large value: offsets and sizes that point far away
        at = foot_off + 8 * int(rng.integers(0, (size - 24 - foot_off) // 8))
        x[at:at + 8] = np.frombuffer(np.uint64(int(rng.integers(0, 2**63))).tobytes(), dtype=np.uint8)
    else:  # headers of the row groups: the bound reader parses them on the host
        for _ in range(8):
            x[int(rng.integers(16, foot_off))] = int(rng.integers(0, 256))
    try:
        with flsgpu.Reader.from_image(x) as rd:
            for g in range(min(rd.n_rowgroups, 4)):
                rd.rowgroup_rows(g)
                for c in range(rd.n_cols):
                    rd.chunk_info(g, c)
                    rd.rowgroup_minmax(g, c)
        ok += 1
    except flsgpu.FlsError:
        bad += 1
print("done", ok, bad)
"""


def _image(tmp_path):
    cols, names = synth.basic_usage(3 * 65536 + 1000)  # integers, doubles and strings; DELTA, FFOR, ALP, string schemes
    p = str(tmp_path / "ok.fls")
    synth.encode(cols, names, synth.basic_usage_types(), path=p)
    return p


def _run(path, mode, seed, n):
    r = subprocess.run([sys.executable, "-c", CHILD, os.path.join(ROOT, "duckdb-fastlane_b200"), ROOT, path, mode, str(seed), str(n)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, "%s: the reader crashed (rc %d)\n%s" % (mode, r.returncode, r.stderr[-2000:])
    tag, ok, bad = r.stdout.split()[-3:]
    assert tag == "done"
    return int(ok), int(bad)


def test_truncated_files_are_errors(tmp_path):
    ok, bad = _run(_image(tmp_path), "truncate", 1, 200)
    assert ok == 0 and bad == 200


def test_flipped_footer_bytes_never_crash(tmp_path):
    ok, bad = _run(_image(tmp_path), "footer", 2, 400)
    assert ok + bad == 400 and bad > 0


def test_wild_footer_offsets_never_crash(tmp_path):
    ok, bad = _run(_image(tmp_path), "footer_wide", 3, 400)
    assert ok + bad == 400 and bad > 0


def test_flipped_header_bytes_never_crash_the_bound_reader(tmp_path):
    ok, bad = _run(_image(tmp_path), "headers", 4, 200)
    assert ok + bad == 200
