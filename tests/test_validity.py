"""NULLs: the validity bitmap segment (FORMAT.md, chunk flag 1) written by flsw_append_rowgroup_v and read back by the
oracle (CPU) and by libflsgpu (GPU).  The reference ignores FastLanes' null maps
(/root/reference/src/fastlanes_facade.cpp:113-118,173-181 produce NULL only for unknown columns), so there is no
reference behaviour to match beyond "a NULL written is a NULL read"."""
import numpy as np
import pytest

import fls_writer as W
from fls_writer import StrColumn
from oracle import oracle


def _table(n=150_000, seed=3):
    r = np.random.default_rng(seed)
    i32 = r.integers(-1000, 1000, n).astype(np.int32)
    f64 = np.round(r.uniform(0, 1000, n), 2)
    strs = StrColumn.from_list([b"" if k % 5 == 0 else b"v%d" % (k % 977) for k in range(n)])
    i64 = np.arange(n, dtype=np.int64) * 3
    v_i32 = r.random(n) > 0.1
    v_f64 = r.random(n) > 0.5
    v_str = np.array([k % 5 != 0 for k in range(n)])
    v_f64[:70_000] = True            # the first row group of this column has no NULL: no bitmap is stored for it
    cols = [i32, f64, strs, i64]
    valid = [v_i32, v_f64, v_str, None]
    return cols, valid, W.write_table(cols, names=["a", "b", "s", "k"], validity=valid)


def test_oracle_reads_back_the_nulls_that_were_written():
    cols, valid, img = _table()
    ora = oracle.OracleTable(img)
    for c in range(3):
        assert np.array_equal(ora.validity(c), valid[c]), c
    assert ora.validity(3) is None
    assert ora.validity(1, 0, 1) is None and ora.validity(1, 1, 2) is not None   # per row group: stored only where needed
    a = ora.decode(0)
    assert np.array_equal(a[valid[0]], cols[0][valid[0]])
    b = ora.decode(1)
    assert np.array_equal(b[valid[1]].view(np.uint64), cols[1][valid[1]].view(np.uint64))
    offs, pay = ora.decode(2)
    want = cols[2].to_list()
    for k in (1, 2, 3, 149_999):
        assert bytes(pay[offs[k]:offs[k + 1]]) == want[k]


@pytest.mark.gpu
def test_gpu_validity_and_values_equal_the_oracle():
    import flsgpu
    cols, valid, img = _table()
    ora = oracle.OracleTable(img)
    with flsgpu.Reader.from_image(img) as rd:
        for to_host in (True, False):
            for (g0, g1) in ((0, ora.n_rg), (1, 3), (0, 1)):
                with rd.decode_batch(g0, g1, to_host=to_host) as b:
                    for c in range(4):
                        ev = ora.validity(c, g0, g1)
                        gv = b.validity(c)
                        assert (ev is None) == (gv is None), (c, g0, g1)
                        if ev is not None:
                            assert np.array_equal(gv, ev), (c, g0, g1)
                        exp = ora.decode(c, g0, g1)
                        got = b.numpy(c) if to_host else b.fetch(c)
                        if isinstance(exp, tuple):
                            assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1])
                        else:
                            assert np.array_equal(got.view(np.uint8), exp.view(np.uint8))


@pytest.mark.gpu
def test_gpu_filter_drops_nulls_and_compacts_validity():
    import flsgpu
    cols, valid, img = _table()
    ora = oracle.OracleTable(img)
    a, b_ = ora.decode(0), ora.decode(1)
    with flsgpu.Reader.from_image(img) as rd:
        # a comparison with NULL is not true: NULL rows of the predicate column go
        ids, got = rd.decode_filtered([(0, ">=", 0)], cols=[0, 1, 3])
        m = valid[0] & (a >= 0)
        assert np.array_equal(ids, np.nonzero(m)[0].astype(np.uint64))
        assert np.array_equal(got[0], a[m]) and np.array_equal(got[2], cols[3][m])
        with rd.decode_filtered_batch([(0, ">=", 0)], cols=[0, 1, 2]) as fb:
            assert np.array_equal(fb.validity(1), valid[1][m])      # the projected column keeps its own NULLs
            assert np.array_equal(fb.validity(2), valid[2][m])
            assert fb.validity(0).all()
        ids, got = rd.decode_filtered([(1, "<", 500.0), (2, "!=", b"v5")], cols=[3])
        offs, pay = ora.decode(2)
        is_v5 = np.array([bytes(pay[offs[k]:offs[k + 1]]) == b"v5" for k in range(len(a))])
        m2 = valid[1] & (b_ < 500.0) & valid[2] & ~is_v5
        assert np.array_equal(ids, np.nonzero(m2)[0].astype(np.uint64))
