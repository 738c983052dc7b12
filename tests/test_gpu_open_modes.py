"""GPU: the three ways a file opened by path reaches HBM — mapped + page-locked bounce buffer per pipeline slice (the
default), the whole file read into page-locked memory first (FLSGPU_OPEN_LOAD_FILE), and cuFile / GPUDirect Storage
(FLSGPU_OPEN_GDS).  All must decode the same bytes; GDS may instead refuse loudly when
the box has no cuFile driver — never fall back silently.  The reference has no twin (it opens the file through
FastLanes, /root/reference/src/fastlanes_facade.cpp:33-34)."""
import numpy as np
import pytest

import flsgpu
import synth
from oracle import oracle

pytestmark = pytest.mark.gpu


def _same(got, exp):
    if isinstance(exp, tuple):
        return np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1])
    return np.array_equal(got.view(np.uint8), exp.view(np.uint8))


def test_mapped_loaded_and_gds_opens_decode_the_same_bytes(tmp_path):
    path = str(tmp_path / "li.fls")
    img = synth.lineitem_image(5, tail_rows=333)
    img.tofile(path)
    ora = oracle.OracleTable(img)
    want = [ora.decode(c) for c in range(ora.n_cols)]
    for flags in (0, flsgpu.OPEN_LOAD_FILE):
        with flsgpu.Reader.open(path, flags=flags, slice_rowgroups=2) as rd:
            got = rd.decode()
            assert all(_same(g, e) for g, e in zip(got, want)), flags
            rd.make_resident()
            got = [rd.decode_batch().fetch(c) for c in range(3)]
            assert all(_same(g, e) for g, e in zip(got, want[:3]))


def test_gds_open_decodes_or_refuses_loudly(tmp_path):
    path = str(tmp_path / "li.fls")
    img = synth.lineitem_image(3, tail_rows=333)
    img.tofile(path)
    ora = oracle.OracleTable(img)
    want = [ora.decode(c) for c in range(ora.n_cols)]
    try:
        with flsgpu.Reader.open(path, flags=flsgpu.OPEN_GDS, slice_rowgroups=2) as rd:
            got = rd.decode()
            assert all(_same(g, e) for g, e in zip(got, want))
            print("GDS: cuFile read path decoded bit-exactly")
    except flsgpu.FlsError as e:
        assert e.code == flsgpu.E_IO and "FLSGPU_OPEN_GDS" in str(e), str(e)
        print("GDS refused:", e)
