"""The drop-in boundary end to end: a real DuckDB v1.3.2 instance (libduckdb built from the reference tree's DuckDB
sources) with the GPU read path registered the way the extension registers it, driven with SQL through
ext/duckdb_sql_check.  Statements and expectations follow the reference's own sqllogictest
(/root/reference/test/sql/fastlane.test) — its data.fls fixture is absent, so a table of the same shape (1024 rows,
one VARCHAR column `data`) is minted here — plus typed multi-column scans checked against the source columns.

Skipped when the check binary has not been built (it needs the DuckDB sources: dev container only; once built it
travels to the GPU box with the snapshot)."""
import json
import os
import subprocess

import numpy as np
import pytest

import synth
from fls_writer import StrColumn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXT = os.path.join(ROOT, "duckdb-fastlane_b200", "ext")
CHECK = os.path.join(EXT, "duckdb_sql_check")

pytestmark = pytest.mark.skipif(not os.path.exists(CHECK), reason="ext/duckdb_sql_check not built (needs the DuckDB sources)")


def sql(*stmts):
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.path.join(EXT, "_duckdb") + ":" + os.path.join(ROOT, "duckdb-fastlane_b200") + ":" + \
        env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([CHECK] + list(stmts), capture_output=True, text=True, timeout=600, env=env)
    out = [json.loads(line) for line in r.stdout.strip().splitlines() if line.startswith("{")]
    assert len(out) == len(stmts), r.stdout + r.stderr
    return out


def test_missing_file_is_a_bind_error_with_the_reference_text():
    # fastlane.test:9-12 (no GPU needed: bind fails on the footer)
    (res,) = sql("SELECT * FROM scan_fastlanes('nonexistent.fls') LIMIT 0;")
    assert "Failed to open FastLanes file: nonexistent.fls" in res["error"]
    (res,) = sql("SELECT * FROM scan_fastlanes(42);")
    assert "error" in res


def test_bind_returns_the_typed_schema_without_decoding(tmp_path):
    cols, names = synth.basic_usage(3000, seed=1)
    path = str(tmp_path / "b.fls")
    synth.encode(cols, names, types=synth.basic_usage_types(), path=path)
    (res,) = sql("DESCRIBE SELECT * FROM read_fls('%s');" % path)   # bind only: works without a GPU
    assert [r[0] for r in res["rows"]] == list(names)
    assert [r[1] for r in res["rows"]] == ["INTEGER", "BIGINT", "DOUBLE", "DOUBLE", "VARCHAR", "BOOLEAN", "DATE"]


@pytest.fixture(scope="module")
def paper_like(tmp_path_factory):
    """1024 rows, one VARCHAR column `data` (shape of third_party/fastlanes/data/fls/data.fls)."""
    r = np.random.default_rng(5)
    words = [b"FastLanes", b"compression", b"layout", b"decoding", b"integers", b"per", b"second", b"scalar", b"code"]
    rows = []
    for i in range(1024):
        n = int(r.integers(3, 30))
        rows.append(b" ".join(words[int(j)] for j in r.integers(1 if i % 14 else 0, len(words), n)))
    path = str(tmp_path_factory.mktemp("fls") / "data.fls")
    synth.encode([StrColumn.from_list(rows)], ["data"], path=path)
    return path, rows


@pytest.mark.gpu
def test_reference_sqllogictest_statements(paper_like):
    path, rows = paper_like
    q = "scan_fastlanes('%s')" % path
    res = sql("SELECT COUNT(*) FROM %s;" % q,                                            # fastlane.test:15-18
              "SELECT COUNT(*) FROM %s WHERE LENGTH(data) > 0;" % q,                     # :21-25
              "SELECT COUNT(*) FROM (SELECT * FROM %s LIMIT 5);" % q,                    # :28-31
              "SELECT COUNT(*) FROM %s WHERE data LIKE '%%FastLanes%%';" % q,            # :34-38
              "SELECT COUNT(*), MIN(LENGTH(data)), MAX(LENGTH(data)) FROM %s;" % q,      # :41-45
              "SELECT COUNT(*) FROM (SELECT data as col1, data as col2, data as col3 FROM %s LIMIT 3);" % q,  # :49-56
              "SELECT col1, col2, col3 FROM (SELECT data as col1, data as col2, data as col3 FROM %s LIMIT 1);" % q)
    assert res[0]["rows"] == [["1024"]]
    assert res[1]["rows"] == [["1024"]]
    assert res[2]["rows"] == [["5"]]
    assert res[3]["rows"] == [[str(sum(b"FastLanes" in x for x in rows))]]
    assert res[4]["rows"] == [["1024", str(min(map(len, rows))), str(max(map(len, rows)))]]
    assert res[5]["rows"] == [["3"]]
    assert res[6]["rows"] == [[rows[0].decode()] * 3]
    assert res[0]["columns"] == ["count_star()"] and all("error" not in x for x in res)


@pytest.mark.gpu
def test_typed_scan_aggregates_projection_multifile_replacement(tmp_path):
    n = 150_000  # 2 full row groups + a short one
    cols, names = synth.basic_usage(n, seed=42)
    path = str(tmp_path / "basic.fls")
    synth.encode(cols, names, types=synth.basic_usage_types(), path=path)
    ids, big, score, rnd, user = cols[0], cols[1], cols[2], cols[3], cols[4].to_list()
    res = sql("SELECT count(*), sum(id), min(big), max(big) FROM read_fls('%s');" % path,
              "SELECT sum(score) FROM read_fls('%s');" % path,                      # projection: one column decoded
              "SELECT username FROM read_fls('%s') WHERE id = 77777;" % path,
              "SELECT count(*) FROM read_fls(['%s', '%s']);" % (path, path),        # LIST(VARCHAR) overload
              "SELECT count(*) FROM '%s';" % path,                                  # replacement scan on *.fls
              "SELECT id, username, rnd FROM scan_fastlanes('%s') ORDER BY id DESC LIMIT 2;" % path)
    assert all("error" not in x for x in res), res
    assert res[0]["rows"] == [[str(n), str(int(ids.astype(np.int64).sum())), str(int(big.min())), str(int(big.max()))]]
    assert abs(float(res[1]["rows"][0][0]) - float(score.sum())) <= 1e-6 * abs(float(score.sum()))
    k = int(np.nonzero(ids == 77777)[0][0])
    assert res[2]["rows"] == [[user[k].decode()]]
    assert res[3]["rows"] == [[str(2 * n)]]
    assert res[4]["rows"] == [[str(n)]]
    top = np.argsort(ids)[::-1][:2]
    assert [r[0] for r in res[5]["rows"]] == [str(int(ids[i])) for i in top]
    assert [r[1] for r in res[5]["rows"]] == [user[i].decode() for i in top]
    assert [float(r[2]) for r in res[5]["rows"]] == [float(rnd[i]) for i in top]
