"""Host side above the C ABI: the scan cursor that the DuckDB facade / table function mirror is built on
(duckdb-fastlane_b200/ext), driven the way the reference's FastLanesScan drives its facade
(/root/reference/src/scan_fastlanes.cpp:65-77: one <= 2048-row chunk per call until a call yields nothing).

ext/fls_scan_check prints position-weighted digests of every served value; the expected ones come from the CPU oracle.
The error path is the one the reference's sqllogictest pins (/root/reference/test/sql/fastlane.test:9-12)."""
import json
import os
import subprocess

import numpy as np
import pytest

import synth
from fls_writer import StrColumn  # noqa: F401
from oracle import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECK = os.path.join(ROOT, "duckdb-fastlane_b200", "ext", "fls_scan_check")


def wsum(x: np.ndarray):
    """sum of x_k * (k + 1) mod 2^64"""
    x = x.astype(np.uint64)
    with np.errstate(over="ignore"):
        return int((x * np.arange(1, x.size + 1, dtype=np.uint64)).sum(dtype=np.uint64))


def expected_digest(ora, c, g0=0, g1=None):
    g1 = ora.n_rg if g1 is None else g1
    vals, lens = [], []
    for g in range(g0, g1):  # row group by row group: padding rows of a short last vector are never served
        dec = ora.decode(c, g, g + 1)
        if isinstance(dec, tuple):
            offs, pay = dec
            vals.append(pay[int(offs[0]):int(offs[-1])])
            lens.append(np.diff(offs.astype(np.int64)))
        else:
            vals.append(dec.view(np.uint8))
    a = wsum(np.concatenate(vals)) if vals else 0
    b = wsum(np.concatenate(lens)) if lens else 0
    return "%016x:%016x" % (a, b)


def run_check(path, *args):
    r = subprocess.run([CHECK, path] + list(args), capture_output=True, text=True, timeout=300)
    return r.returncode, json.loads(r.stdout.strip().splitlines()[-1])


def test_missing_file_error_text():
    # bind-time failure text of the reference: "Failed to open FastLanes file: <path>"
    rc, out = run_check("nonexistent.fls")
    assert rc == 1 and out == {"error": "Failed to open FastLanes file: nonexistent.fls"}


def test_not_an_fls_file(tmp_path):
    p = tmp_path / "junk.fls"
    p.write_bytes(b"definitely not a FastLanes file" * 100)
    rc, out = run_check(str(p))
    assert rc == 1 and out["error"].startswith("Failed to open FastLanes file: " + str(p))


@pytest.fixture(scope="module")
def small_table(tmp_path_factory):
    # 150,000 rows: 2 full row groups + a short one whose last vector is padded
    cols, names = synth.basic_usage(150_000, seed=42)
    path = str(tmp_path_factory.mktemp("fls") / "basic.fls")
    img = synth.encode(cols, names, path=path)
    return path, oracle.OracleTable(img), names


@pytest.mark.gpu
def test_full_scan_matches_oracle(small_table):
    path, ora, names = small_table
    rc, out = run_check(path)
    assert rc == 0, out
    assert out["rows"] == ora.n_rows == 150_000
    assert out["chunks"] == sum(-(-ora.rows(g, g + 1) // 2048) for g in range(ora.n_rg))
    assert out["names"] == list(names)
    assert out["digests"] == [expected_digest(ora, c) for c in range(ora.n_cols)]


@pytest.mark.gpu
@pytest.mark.parametrize("args", [["--chunk", "1000"], ["--batch", "1"], ["--chunk", "65536", "--batch", "2"]])
def test_chunking_and_batching_do_not_change_values(small_table, args):
    path, ora, _ = small_table
    rc, out = run_check(path, *args)
    assert rc == 0 and out["rows"] == ora.n_rows
    assert out["digests"] == [expected_digest(ora, c) for c in range(ora.n_cols)]


@pytest.mark.gpu
def test_projection_and_rowgroup_range(small_table):
    path, ora, names = small_table
    rc, out = run_check(path, "--cols", "4,0,4", "--range", "1:3")
    assert rc == 0
    assert out["rows"] == ora.rows(1, 3)
    assert out["names"] == [names[4], names[0], names[4]]
    assert out["digests"] == [expected_digest(ora, c, 1, 3) for c in (4, 0, 4)]
    rc, out = run_check(path, "--cols", "99")
    assert rc == 1 and "bad column" in out["error"]


@pytest.mark.gpu
def test_several_devices_behind_one_cursor_serve_the_same_rows_in_order(small_table):
    """SURVEY.md section 8e / `gpus := [..]`: one handle and one decode in flight per listed device, batches dealt
    round-robin and served in file order.  On a one-GPU box the device is listed twice (two lanes, two handles)."""
    import flsgpu
    path, ora, _ = small_table
    want = [expected_digest(ora, c) for c in range(ora.n_cols)]
    n_dev = 1
    try:
        import torch
        n_dev = max(1, torch.cuda.device_count())
    except Exception:
        pass
    lists = ["0,0", "0,0,0"] + (["0,1", ",".join(str(i) for i in range(n_dev))] if n_dev > 1 else [])
    for gpus in lists:
        for batch in ("1", "2"):
            rc, out = run_check(path, "--gpus", gpus, "--batch", batch)
            assert rc == 0 and out["rows"] == ora.n_rows, (gpus, out)
            assert out["digests"] == want, gpus
    rc, out = run_check(path, "--gpus", "0,99", "--batch", "1")   # the second batch goes to the device that does not exist
    assert rc == 1 and "error" in out
