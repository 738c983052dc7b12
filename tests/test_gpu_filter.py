"""GPU: flsgpu_decode_filtered (late materialisation) against the oracle's decode of the
same file filtered with numpy.  The reference has filter pushdown switched off
(/root/reference/src/scanner/scan_fastlanes.cpp:154-155): the semantics checked here are DuckDB's comparison
operators (NaN = NaN, NaN greater than everything), which the engine re-applies on top of the scan."""
import numpy as np
import pytest

import flsgpu
import synth
from oracle import oracle

pytestmark = pytest.mark.gpu

N = synth.LINEITEM_NAMES.index


def _np_cmp(col, op, val):
    if isinstance(col, tuple):  # strings: = / <> only
        offs, pay = col
        lens = np.diff(offs.astype(np.int64))
        eq = lens == len(val)
        idx = np.nonzero(eq)[0]
        if len(val) and len(idx):
            pos = offs[idx].astype(np.int64)[:, None] + np.arange(len(val))[None, :]
            eq[idx] = (pay[pos] == np.frombuffer(val, np.uint8)[None, :]).all(axis=1)
        return eq if op == "=" else ~eq
    if col.dtype.kind == "f":
        a, b = col, col.dtype.type(val)
        an, bn = np.isnan(a), np.isnan(b)
        eq = (an & bn) | (a == b)
        lt = ~an & (bn | (a < b))
        return {"=": eq, "!=": ~eq, "<": lt, "<=": lt | eq, ">": ~(lt | eq), ">=": ~lt}[op]
    v = col.dtype.type(val)
    return {"=": col == v, "!=": col != v, "<": col < v, "<=": col <= v, ">": col > v, ">=": col >= v}[op]


def _take(col, mask):
    if isinstance(col, tuple):
        offs, pay = col
        idx = np.nonzero(mask)[0]
        lens = (offs[idx + 1] - offs[idx]).astype(np.int64)
        out_offs = np.zeros(len(idx) + 1, np.uint32)
        out_offs[1:] = np.cumsum(lens)
        parts = [pay[offs[i]:offs[i + 1]] for i in idx]
        return out_offs, (np.concatenate(parts) if parts else np.empty(0, np.uint8))
    return col[mask]


def _check(img, preds, cols=None, g0=0, g1=None):
    ora = oracle.OracleTable(img)
    g1 = ora.n_rg if g1 is None else g1
    want_cols = list(range(ora.n_cols)) if cols is None else cols
    mask = np.ones(ora.rows(g0, g1), bool)
    for (c, op, v) in preds:
        mask &= _np_cmp(ora.decode(c, g0, g1), op, v)
    with flsgpu.Reader.from_image(img) as rd:
        ids, got = rd.decode_filtered(preds, g0, g1, cols)
    assert np.array_equal(ids, np.nonzero(mask)[0].astype(np.uint64))
    for c, g in zip(want_cols, got):
        exp = _take(ora.decode(c, g0, g1), mask)
        if isinstance(exp, tuple):
            assert np.array_equal(g[0], exp[0]) and np.array_equal(g[1], exp[1]), "column %d" % c
        else:
            assert np.array_equal(g.view(np.uint8), exp.view(np.uint8)), "column %d" % c
    return int(mask.sum())


@pytest.fixture(scope="module")
def li():
    return synth.lineitem_image(3, tail_rows=777)


def test_date_range_all_columns(li):
    n = _check(li, [(N("l_shipdate"), ">=", 9000), (N("l_shipdate"), "<", 9030)])
    assert n > 0


def test_double_and_string_predicates_not_projected(li):
    n = _check(li, [(N("l_quantity"), "<", 10.0), (N("l_returnflag"), "=", b"R")], cols=[N("l_orderkey"), N("l_comment"), N("l_shipmode")])
    assert n > 0


def test_no_predicate_keeps_every_row(li):
    ora = oracle.OracleTable(li)
    assert _check(li, [], cols=[N("l_orderkey"), N("l_comment")]) == ora.rows(0, ora.n_rg)


def test_nothing_survives(li):
    assert _check(li, [(N("l_linenumber"), ">", 100)]) == 0
    assert _check(li, [(N("l_shipinstruct"), "=", b"no such instruction")], cols=[N("l_comment"), N("l_tax")]) == 0


def test_string_ne_int64_ge_and_subrange(li):
    ora = oracle.OracleTable(li)
    mid = int(np.median(ora.decode(N("l_orderkey"))))
    assert _check(li, [(N("l_shipmode"), "!=", b"AIR"), (N("l_orderkey"), ">=", mid)], g0=1, g1=3) > 0
    assert _check(li, [(N("l_comment"), "=", b"")], g0=2, g1=3, cols=[N("l_comment")]) >= 0


def test_every_operator_on_ints_and_doubles(li):
    for op in ("=", "!=", "<", "<=", ">", ">="):
        _check(li, [(N("l_linenumber"), op, 3)], cols=[N("l_linenumber"), N("l_extendedprice")], g0=2, g1=3)
        _check(li, [(N("l_discount"), op, 0.05)], cols=[N("l_discount")], g0=2, g1=3)


def test_nan_and_unsigned_semantics(tmp_path):
    import fls_writer as W
    r = np.random.default_rng(5)
    n = 70_000
    d = r.normal(size=n)
    d[r.integers(0, n, 500)] = np.nan
    d[r.integers(0, n, 50)] = np.inf
    u = r.integers(0, 1 << 63, n, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
    f = r.normal(size=n).astype(np.float32)
    f[::97] = np.nan
    s8 = r.integers(-128, 128, n).astype(np.int8)
    img = W.write_table([d, u, f, s8], names=["d", "u", "f", "s8"])
    for op in ("=", "!=", "<", "<=", ">", ">="):
        _check(img, [(0, op, float("nan"))])
        _check(img, [(0, op, 0.25)])
        _check(img, [(1, op, (1 << 63) + 12345)])
        _check(img, [(2, op, float("nan"))], cols=[2])
        _check(img, [(3, op, -5)], cols=[3, 0])
