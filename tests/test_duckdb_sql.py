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


def sql(*stmts, trace=None, env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    if trace is not None:
        env["FLS_SCAN_TRACE"] = "1"
    env["LD_LIBRARY_PATH"] = os.path.join(EXT, "_duckdb") + ":" + os.path.join(ROOT, "duckdb-fastlane_b200") + ":" + \
        env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([CHECK] + list(stmts), capture_output=True, text=True, timeout=600, env=env)
    out = [json.loads(line) for line in r.stdout.strip().splitlines() if line.startswith("{")]
    assert len(out) == len(stmts), r.stdout + r.stderr
    if trace is not None:
        # one entry per statement that scanned: {"skipped": row groups pruned, "of": row groups looked at,
        # "kept": rows the device filter handed back, "scanned": rows it looked at} summed over the scan's shards
        cur = None
        for line in r.stderr.splitlines():
            if line.startswith("sql_check: statement"):
                cur = None
            elif line.startswith(("fls_scan:", "fls_filter:")):
                if cur is None:
                    cur = {"skipped": 0, "of": 0, "kept": 0, "scanned": 0, "filtered": False}
                    trace.append(cur)
                w = line.split()
                if line.startswith("fls_scan:"):
                    cur["skipped"] += int(w[2]); cur["of"] += int(w[4])
                else:
                    cur["kept"] += int(w[4]); cur["scanned"] += int(w[6]); cur["filtered"] = True
    return out


def same_rows(a, b, rel=1e-9):
    """Result rows equal, floating-point cells up to the rounding of a differently ordered (parallel) summation."""
    if len(a) != len(b):
        return False
    for ra, rb in zip(a, b):
        if len(ra) != len(rb):
            return False
        for x, y in zip(ra, rb):
            if x == y:
                continue
            try:
                fx, fy = float(x), float(y)
            except ValueError:
                return False
            if abs(fx - fy) > rel * max(abs(fx), abs(fy), 1e-300):
                return False
    return True


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


# ---------------------------------------------------------------------------------------------- COPY ... TO (FORMAT FLS)
# the sink the reference declares in src/writer/write_fastlane_stream.cpp, serving the
# statement of examples/basic_usage.sql:27-36.  Writing is host code: these run without a GPU; the oracle reads back.
_COPY_SRC = ("SELECT i::INTEGER AS id, (i * 1000003 + 1234567890123)::BIGINT AS big, round(i * 0.37, 2)::DOUBLE AS score, "
             "'user_' || i AS username, (i % 2 = 0) AS is_active, DATE '2023-01-01' + (i % 365)::INTEGER AS join_date, "
             "(i % 7)::DECIMAL(9,2) AS amount FROM range(1, 150001) t(i)")


def test_copy_to_fls_writes_what_the_oracle_reads(tmp_path):
    from oracle import oracle
    path = str(tmp_path / "users.fls")
    res = sql("COPY (%s) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 50000)" % (_COPY_SRC, path),
              "DESCRIBE SELECT * FROM read_fls('%s')" % path)
    assert res[0]["rows"] == [["150000"]], res[0]
    assert [(r[0], r[1]) for r in res[1]["rows"]] == [
        ("id", "INTEGER"), ("big", "BIGINT"), ("score", "DOUBLE"), ("username", "VARCHAR"), ("is_active", "BOOLEAN"),
        ("join_date", "DATE"), ("amount", "VARCHAR")]            # DECIMAL is FALLBACK -> VARCHAR (type_mapping.cpp:58-60)
    ora = oracle.OracleTable(np.fromfile(path, dtype=np.uint8))
    assert ora.n_rows == 150000 and ora.n_rg == 3                # 50000 rounds up to 49 whole vectors per row group
    i = np.arange(1, 150001, dtype=np.int64)
    assert np.array_equal(ora.decode(0), i.astype(np.int32))
    assert np.array_equal(ora.decode(1), i * 1000003 + 1234567890123)
    assert np.array_equal(ora.decode(4).view(np.uint8), (i % 2 == 0).astype(np.uint8))
    assert np.array_equal(ora.decode(5), (19358 + i % 365).astype(np.int32))
    offs, pay = ora.decode(3)
    assert bytes(pay[offs[0]:offs[2]]) == b"user_1user_2" and bytes(pay[offs[149999]:offs[150000]]) == b"user_150000"
    offs, pay = ora.decode(6)
    assert bytes(pay[offs[0]:offs[2]]) == b"1.002.00"


def test_copy_to_fls_options_and_errors(tmp_path):
    p = str(tmp_path / "x.fls")
    res = sql("COPY (SELECT 1 AS a) TO '%s' (FORMAT FLS, FOO 3)" % p,
              "COPY (SELECT 1 AS a) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 2048, ROW_GROUP_SIZE_BYTES '1MB')" % p,
              "COPY (SELECT NULL::INTEGER AS a) TO '%s' (FORMAT FLS)" % p,                 # NULLs are written (validity bitmap)
              "COPY (SELECT 42 AS a) TO '%s'" % str(tmp_path / "y.fastlane"),      # format from the file extension
              "COPY (SELECT 42 AS a) TO '%s' (FORMAT FLS, CHUNK_SIZE 4096, ROW_GROUPS_PER_FILE 10)" % p)
    assert "Unknown option for FastLanes: FOO" in res[0]["error"]                  # write_fastlane_stream.cpp:104
    assert "mutually exclusive" in res[1]["error"]                                 # :85-87
    assert res[2]["rows"] == [["1"]]
    assert res[3]["rows"] == [["1"]] and res[4]["rows"] == [["1"]]


@pytest.mark.gpu
def test_copy_then_read_fls_round_trip(tmp_path):
    """examples/basic_usage.sql end to end: COPY a table out, scan it back on the GPU, compare with the source query."""
    path = str(tmp_path / "users.fls")
    agg = "count(*), sum(id), sum(big), sum(score), min(username), max(username), count(*) FILTER (is_active), max(join_date)"
    res = sql("COPY (%s) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 20480)" % (_COPY_SRC, path),
              "SELECT %s FROM (%s)" % (agg, _COPY_SRC),
              "SELECT %s FROM read_fls('%s')" % (agg, path),
              "SELECT count(*) FROM (SELECT * FROM (%s) EXCEPT SELECT * REPLACE (amount::DECIMAL(9,2) AS amount) FROM read_fls('%s'))"
              % (_COPY_SRC, path))
    assert same_rows(res[1]["rows"], res[2]["rows"]), (res[1], res[2])
    assert res[3]["rows"] == [["0"]], res[3]


@pytest.mark.gpu
def test_row_groups_are_skipped_by_their_compressed_bounds(tmp_path):
    """comparisons on integer / DATE columns skip the row groups whose FFOR / DELTA vector
    headers rule them out; the filter itself stays with DuckDB, so results equal the unfiltered source."""
    path = str(tmp_path / "sorted.fls")
    src = ("SELECT i::BIGINT AS k, (i % 1000)::INTEGER AS m, DATE '2020-01-01' + (i // 4096)::INTEGER AS d, "
           "'row_' || i AS s FROM range(0, 400000) t(i)")
    trace = []
    res = sql("COPY (%s) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 16384)" % (src, path),
              "SELECT count(*), sum(k), min(s), max(s) FROM read_fls('%s') WHERE k >= 100000 AND k < 120000" % path,
              "SELECT count(*), sum(k), min(s), max(s) FROM (%s) WHERE k >= 100000 AND k < 120000" % src,
              "SELECT count(*), sum(m) FROM read_fls('%s') WHERE d = DATE '2020-02-01'" % path,
              "SELECT count(*), sum(m) FROM (%s) WHERE d = DATE '2020-02-01'" % src,
              "SELECT count(*) FROM read_fls('%s') WHERE m = 7" % path,          # every row group holds m = 7: none skipped
              "SELECT count(*) FROM read_fls('%s') WHERE k > 1000000" % path,    # nothing qualifies: all skipped
              trace=trace)
    assert res[1]["rows"] == res[2]["rows"] and res[1]["rows"][0][0] == "20000"
    assert res[3]["rows"] == res[4]["rows"] and res[3]["rows"][0][0] == "4096"
    assert res[5]["rows"] == [["400"]] and res[6]["rows"] == [["0"]]
    skipped = [t["skipped"] for t in trace]
    total = [t["of"] for t in trace]
    assert total == [25, 25, 25, 25] and skipped == [23, 24, 0, 25], trace


@pytest.mark.gpu
def test_row_groups_are_skipped_by_their_dictionaries(tmp_path):
    """SURVEY.md section 8f rank 3, dictionary predicate: `column = 'text'` skips the row groups whose dictionary
    (or constant) does not hold the text; the filter stays with DuckDB, results equal the source's."""
    path = str(tmp_path / "modes.fls")
    modes = ["REG AIR", "AIR", "RAIL", "SHIP", "TRUCK", "MAIL", "FOB"]
    src = ("SELECT i::BIGINT AS k, (%s)[1 + ((i // 50000) + (i %% 2)) %% 7] AS mode FROM range(0, 400000) t(i)"
           % ("[" + ", ".join("'%s'" % m for m in modes) + "]"))
    trace = []
    res = sql("COPY (%s) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 16384)" % (src, path),
              "SELECT count(*), sum(k) FROM read_fls('%s') WHERE mode = 'SHIP'" % path,
              "SELECT count(*), sum(k) FROM (%s) WHERE mode = 'SHIP'" % src,
              "SELECT count(*) FROM read_fls('%s') WHERE mode = 'NOPE'" % path,
              "SELECT count(*) FROM read_fls('%s') WHERE 'AIR' = mode AND k < 60000" % path,
              "SELECT count(*) FROM (%s) WHERE 'AIR' = mode AND k < 60000" % src,
              "SELECT count(*) FROM read_fls('%s') WHERE mode <> 'SHIP'" % path,     # not an equality: nothing to prune by
              "SELECT count(*) FROM (%s) WHERE mode <> 'SHIP'" % src,
              trace=trace)
    assert res[1]["rows"] == res[2]["rows"] and int(res[1]["rows"][0][0]) > 0
    assert res[3]["rows"] == [["0"]]
    assert res[4]["rows"] == res[5]["rows"]
    assert res[6]["rows"] == res[7]["rows"]

    def holds(g, m):   # does row group g hold mode index m?
        lo, hi = 16384 * g, min(16384 * (g + 1), 400000)
        return any(((i // 50000) + (i % 2)) % 7 == m for i in (lo, lo + 1, hi - 2, hi - 1))
    n_rg = (400000 + 16383) // 16384
    want_ship = sum(not holds(g, 3) for g in range(n_rg))
    want_air = sum((not holds(g, 1)) or 16384 * g >= 60000 for g in range(n_rg))
    skipped = [t["skipped"] for t in trace]
    assert skipped == [want_ship, n_rg, want_air, 0], (trace, want_ship, want_air)
    assert 0 < want_ship < n_rg and 0 < want_air < n_rg and int(res[4]["rows"][0][0]) > 0


@pytest.mark.gpu
def test_device_filter_serves_only_surviving_rows(tmp_path):
    """SURVEY.md section 8f rank 3, second half (late materialisation): `column <op> constant` conjunctions run on the
    GPU after the decode (flsgpu_decode_filtered) and only the rows that pass are copied back; DuckDB re-applies the
    filter, so the answers equal the source query's and those of a scan with the device filter switched off."""
    path = str(tmp_path / "f.fls")
    src = ("SELECT i::BIGINT AS k, ((i * 7919) % 1000)::INTEGER AS m, DATE '2020-01-01' + ((i * 31) % 700)::INTEGER AS d, "
           "round(((i * 37) % 10007) * 0.01, 2)::DOUBLE AS price, (['AIR', 'RAIL', 'SHIP', 'TRUCK'])[1 + (i * 13) % 4] AS mode, "
           "'row_' || i AS s, (i % 3 = 0) AS flag FROM range(0, 300000) t(i)")
    qs = ["WHERE m < 10", "WHERE d BETWEEN DATE '2020-03-01' AND DATE '2020-03-07' AND mode = 'SHIP'",
          "WHERE price >= 99.5 AND mode <> 'AIR'", "WHERE flag = true AND k >= 299000", "WHERE m = 5000", "WHERE 17.25 = price"]
    agg = "count(*), sum(k), sum(m), min(s), max(s), sum(price), count(*) FILTER (flag)"
    stm = ["COPY (%s) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 32768)" % (src, path)]
    for q in qs:
        stm += ["SELECT %s FROM read_fls('%s') %s" % (agg, path, q), "SELECT %s FROM (%s) %s" % (agg, src, q)]
    trace = []
    res = sql(*stm, trace=trace)
    assert all("error" not in r for r in res), res
    for i in range(len(qs)):
        assert same_rows(res[1 + 2 * i]["rows"], res[2 + 2 * i]["rows"]), (qs[i], res[1 + 2 * i], res[2 + 2 * i])
    assert int(res[1]["rows"][0][0]) == 3000
    kept = [(t["kept"], t["scanned"]) for t in trace if t["filtered"]]
    assert len(kept) == len(qs) and kept[0] == (3000, 300000), trace       # the device handed back exactly the survivors
    counts = [int(res[1 + 2 * i]["rows"][0][0]) for i in range(len(qs))]
    assert all(k >= c for (k, _n), c in zip(kept, counts)), (kept, counts)   # a conjunct DuckDB keeps to itself may remain
    assert [k for k, _n in kept[:3]] == counts[:3] and min(counts[:4]) > 0, (kept, counts)
    off = sql(*[stm[0]] + stm[1::2], env_extra={"FLS_SCAN_NO_DEVICE_FILTER": "1"})
    assert all(same_rows(r["rows"], res[1 + 2 * i]["rows"]) for i, r in enumerate(off[1:]))


@pytest.mark.gpu
def test_nulls_round_trip_through_copy_and_read_fls(tmp_path):
    """NULLs written by COPY (validity bitmap, FORMAT.md) come back as NULLs: every aggregate that looks at them agrees
    with the source query, with and without the device filter."""
    path = str(tmp_path / "nulls.fls")
    src = ("SELECT i::BIGINT AS k, CASE WHEN i % 7 = 0 THEN NULL ELSE (i % 1000)::INTEGER END AS m, "
           "CASE WHEN i % 3 = 0 THEN NULL ELSE round(i * 0.25, 2)::DOUBLE END AS price, "
           "CASE WHEN i % 11 = 0 THEN NULL ELSE 'row_' || (i % 5000) END AS s, "
           "CASE WHEN i > 100000 AND i % 2 = 0 THEN NULL ELSE DATE '2021-01-01' + (i % 400)::INTEGER END AS d "
           "FROM range(0, 200000) t(i)")
    agg = ("count(*), count(m), count(price), count(s), count(d), sum(m), sum(price), min(s), max(s), max(d), "
           "count(*) FILTER (m IS NULL AND s IS NULL)")
    qs = ["", "WHERE m < 100", "WHERE s = 'row_77' OR m IS NULL", "WHERE price > 40000.0 AND d >= DATE '2021-06-01'", "WHERE m IS NULL"]
    stm = ["COPY (%s) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 40960)" % (src, path)]
    for q in qs:
        stm += ["SELECT %s FROM read_fls('%s') %s" % (agg, path, q), "SELECT %s FROM (%s) %s" % (agg, src, q)]
    stm += ["SELECT count(*) FROM (SELECT * FROM (%s) EXCEPT SELECT * FROM read_fls('%s'))" % (src, path)]
    res = sql(*stm)
    assert all("error" not in r for r in res), res
    for i in range(len(qs)):
        assert same_rows(res[1 + 2 * i]["rows"], res[2 + 2 * i]["rows"]), (qs[i], res[1 + 2 * i], res[2 + 2 * i])
    assert res[-1]["rows"] == [["0"]]
    assert res[1]["rows"][0][1] != res[1]["rows"][0][0]       # there are NULLs


def test_count_star_needs_no_decode_and_no_gpu(tmp_path):
    """count(*) asks for the row-id pseudo column alone: the table function serves the row counts from the footer (no
    CUDA device here) and never copies a file column of another type into the row-id vector — the first column of this
    file is a VARCHAR of long strings."""
    import fls_writer as W
    n = 65536 * 2 + 4321
    r = np.random.default_rng(3)
    words = [b"x" * 40, b"yy" * 30, b"a somewhat longer comment string for the heap"]
    col = W.StrColumn.from_list([words[i] for i in r.integers(0, 3, n)])
    a, b = str(tmp_path / "a.fls"), str(tmp_path / "b.fls")
    synth.encode([col, np.arange(n, dtype=np.int32)], ["s", "i"], path=a)
    synth.encode([W.StrColumn.from_list([b"q"] * 1000), np.arange(1000, dtype=np.int32)], ["s", "i"], path=b)
    res = sql("SELECT count(*) FROM read_fls('%s')" % a,
              "SET threads = 1",
              "SELECT count(*) FROM scan_fastlanes('%s')" % a,
              "SELECT count(*) FROM read_fls(['%s', '%s', '%s'])" % (a, b, a),
              "SELECT count(*) FROM (SELECT * FROM read_fls('%s') LIMIT 70000)" % a)
    assert res[0]["rows"] == [[str(n)]] and res[2]["rows"] == [[str(n)]], res
    assert res[3]["rows"] == [[str(2 * n + 1000)]], res[3]
    assert res[4].get("rows") == [["70000"]] or "error" in res[4], res[4]   # needs no column either; a GPU only if DuckDB asks for one


@pytest.mark.gpu
def test_gpus_and_device_named_parameters(tmp_path):
    """`gpus := [..]` / `device := n` on read_fls / scan_fastlanes (the reference registers only `file`,
    /root/reference/src/scan_fastlanes.cpp:80-81): several devices decode behind one ordered scan."""
    n = 200_000
    cols, names = synth.basic_usage(n, seed=9)
    path = str(tmp_path / "b.fls")
    synth.encode(cols, names, types=synth.basic_usage_types(), path=path)
    n_dev = 1
    try:
        import torch
        n_dev = max(1, torch.cuda.device_count())
    except Exception:
        pass
    gl = "[0, 1]" if n_dev > 1 else "[0, 0]"
    agg = "count(*), sum(id), sum(big), min(username), max(username), sum(score)"
    res = sql("SELECT %s FROM read_fls('%s')" % (agg, path),
              "SELECT %s FROM read_fls('%s', gpus := %s)" % (agg, path, gl),
              "SELECT %s FROM scan_fastlanes('%s', device := 0)" % (agg, path),
              "SELECT id FROM read_fls('%s', gpus := %s) WHERE id %% 50000 = 1 ORDER BY id" % (path, gl),
              "SELECT count(*) FROM read_fls('%s', gpus := [0, 77])" % path,
              "SELECT count(*) FROM read_fls('%s', device := -1)" % path,
              # the reference's stream scan registers `auto_detect` (src/scanner/scan_fastlanes.cpp:156) and never reads it
              "SELECT %s FROM scan_fastlanes('%s', auto_detect := true)" % (agg, path))
    assert "error" not in res[0] and same_rows(res[0]["rows"], res[1]["rows"]) and same_rows(res[0]["rows"], res[2]["rows"]), res[:3]
    assert [r[0] for r in res[3]["rows"]] == ["1", "50001", "100001", "150001"]
    assert "error" in res[4] and "error" in res[5]
    assert "error" not in res[6] and same_rows(res[0]["rows"], res[6]["rows"]), res[6]


# ------------------------------------------------------------------------------------------- the loadable extension
# fastlane.duckdb_extension (reference: CMakeLists.txt:32-33 build_loadable_extension): here the functions are NOT
# registered by the check driver — DuckDB's own loader opens the file, checks its metadata footer and calls fastlane_init.
EXTENSION = os.path.join(EXT, "fastlane.duckdb_extension")
needs_extension = pytest.mark.skipif(not os.path.exists(EXTENSION), reason="ext/fastlane.duckdb_extension not built")


@needs_extension
def test_load_extension_registers_the_functions(tmp_path):
    cols, names = synth.basic_usage(3000, seed=1)
    path = str(tmp_path / "b.fls")
    synth.encode(cols, names, types=synth.basic_usage_types(), path=path)
    res = sql("SELECT extension_name, loaded, extension_version FROM duckdb_extensions() WHERE extension_name = 'fastlane'",
              "SELECT * FROM scan_fastlanes('nonexistent.fls') LIMIT 0",                 # fastlane.test:9-12
              "DESCRIBE SELECT * FROM read_fls('%s')" % path,
              "COPY (SELECT 42 AS a) TO '%s' (FORMAT FLS)" % str(tmp_path / "x.fls"),
              env_extra={"FLS_SQL_CHECK_LOAD": EXTENSION})
    assert res[0]["rows"] == [["fastlane", "true", "b200-r02"]]
    assert "Failed to open FastLanes file: nonexistent.fls" in res[1]["error"]
    assert [r[1] for r in res[2]["rows"]] == ["INTEGER", "BIGINT", "DOUBLE", "DOUBLE", "VARCHAR", "BOOLEAN", "DATE"]
    assert res[3]["rows"] == [["1"]]


@needs_extension
@pytest.mark.gpu
def test_loaded_extension_scans_on_the_gpu(tmp_path):
    path = str(tmp_path / "users.fls")
    agg = "count(*), sum(id), sum(big), min(username), max(username), count(*) FILTER (is_active), max(join_date)"
    res = sql("COPY (%s) TO '%s' (FORMAT FLS, ROW_GROUP_SIZE 20480)" % (_COPY_SRC, path),
              "SELECT %s FROM (%s)" % (agg, _COPY_SRC),
              "SELECT %s FROM read_fls('%s')" % (agg, path),
              "SELECT %s FROM '%s' WHERE id < 1000" % (agg, path),
              "SELECT %s FROM (%s) WHERE id < 1000" % (agg, _COPY_SRC),
              env_extra={"FLS_SQL_CHECK_LOAD": EXTENSION})
    assert all("error" not in r for r in res), res
    assert res[1]["rows"] == res[2]["rows"] and res[3]["rows"] == res[4]["rows"]
