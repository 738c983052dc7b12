"""GPU: one flsgpu_decode call over a string column with more than 4 GiB of payload.  The kernels' 32-bit offsets
cannot span it: the library decodes the column in parts of less than 4 GiB and hands back 64-bit offsets
(flsgpu_column.offsets64) — no caller-visible limit (round 1 returned FLSGPU_E_LIMIT).  The reference has no twin
(its facade copies one std::string per row, /root/reference/src/fastlanes_facade.cpp:157-172)."""
import numpy as np
import pytest

import flsgpu
import fls_writer as W
from fls_writer import StrColumn
from oracle import oracle

pytestmark = pytest.mark.gpu


def test_more_than_4_gib_of_strings_in_one_call():
    r = np.random.default_rng(4)
    n = 65536
    # 9 dictionary entries of 800..1200 bytes: a row group decodes to about 65 MB from a few KB
    pool = [bytes(r.integers(97, 123, int(L)).astype(np.uint8)) for L in r.integers(800, 1200, 9)]
    strs = StrColumn.from_list([pool[i] for i in r.integers(0, 9, n)])
    ids = np.arange(n, dtype=np.int64)
    small = W.write_table([strs, ids], names=["s", "k"])
    ora = oracle.OracleTable(small)
    eo, ep = ora.decode(0)
    per_rg = int(eo[-1])
    n_rg = (1 << 32) // per_rg + 3                       # a little over 4 GiB
    img = W.tile_rowgroups(small, n_rg)
    lens = np.diff(eo.astype(np.int64))
    for to_host in (False, True):
        with flsgpu.Reader.from_image(img) as rd:
            with rd.decode_batch(0, n_rg, to_host=to_host) as b:
                assert b.columns[0].offsets64 and not b.columns[0].offsets
                offs, pay = b.numpy(0, copy=False) if to_host else b.fetch(0)
                assert offs.dtype == np.uint64 and len(offs) == n_rg * n + 1 and int(offs[-1]) == n_rg * per_rg > (1 << 32)
                assert np.array_equal(np.diff(offs.astype(np.int64)), np.tile(lens, n_rg))
                for g in (0, 1, n_rg // 2, n_rg - 2, n_rg - 1):   # row groups on both sides of the 4 GiB boundary
                    assert np.array_equal(pay[g * per_rg:(g + 1) * per_rg], ep), g
                k = b.numpy(1) if to_host else b.fetch(1)
                assert np.array_equal(k, np.tile(ids, n_rg))
            # a range below 4 GiB keeps its 32-bit offsets
            with rd.decode_batch(3, 10, to_host=False) as b:
                assert b.columns[0].offsets and not b.columns[0].offsets64
                offs, pay = b.fetch(0)
                assert offs.dtype == np.uint32 and int(offs[-1]) == 7 * per_rg
            with pytest.raises(flsgpu.FlsError) as e:
                rd.decode_filtered([(1, "<", 5)])
            assert e.value.code == flsgpu.E_LIMIT
