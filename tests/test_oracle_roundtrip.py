"""CPU: decode(encode(x)) == x bit-for-bit through the oracle, and the host emulation of the device decoders
(tests/emul, the exact kernel source compiled with FLS_HOST_EMUL) == oracle.  No GPU needed."""
import numpy as np
import pytest

import cases
import emul
import fls_writer as W
from fls_writer import StrColumn
from oracle import oracle

ALL = cases.all_cases()


@pytest.mark.parametrize("case", ALL, ids=[c[0] for c in ALL])
def test_roundtrip_and_emulation(case):
    name, cols, types, force, min_bw = case
    img = cases.encode_case(case)
    ora = oracle.OracleTable(img)
    assert ora.n_rows == len(cols[0])
    for c, src in enumerate(cols):
        if force and force[c] is not None:
            for g in range(ora.n_rg):
                assert ora.scheme(g, c) == force[c], "encoder ignored the forced scheme"
        exp = cases.expected_bytes(src)
        got = ora.decode(c)
        em = emul.decode(img, c, ora.types[c], 0, ora.n_rg)
        if isinstance(src, StrColumn):
            assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1])
            n = ora.n_rows
            assert np.array_equal(em[0][:n + 1], got[0]) and np.array_equal(em[1][:len(got[1])], got[1])
            emul.force_any(True)  # every warp vote succeeds: the general (masked, escape-aware) variants run everywhere
            try:
                em = emul.decode(img, c, ora.types[c], 0, ora.n_rg)
            finally:
                emul.force_any(False)
            assert np.array_equal(em[0][:n + 1], got[0]) and np.array_equal(em[1][:len(got[1])], got[1])
        else:
            assert np.array_equal(got.view(np.uint8), exp)
            assert np.array_equal(em[:ora.n_rows].view(np.uint8), exp)


def test_ffor_width_is_exact():
    # the FFOR sweep generator makes every vector need exactly W bits: compressed bytes = 128*W + 16-byte header
    for bits, w in ((32, 0), (32, 13), (32, 32), (64, 1), (64, 47), (64, 64)):
        img = cases.encode_case(cases.ffor_case(bits, w, 64 * 1024))
        ora = oracle.OracleTable(img)
        assert ora.chunk_bytes(0, 0) == 64 * (128 * w + 16)


def test_rowgroup_range_decode():
    cols = [np.arange(200_000, dtype=np.int64) * 3, StrColumn.from_list([b"x" * (i % 7) for i in range(200_000)])]
    img = W.write_table(cols)
    ora = oracle.OracleTable(img)
    assert ora.n_rg == 4
    a = ora.decode(0, 1, 3)
    assert np.array_equal(a, cols[0][65536:3 * 65536])
    off, pay = ora.decode(1, 1, 3)
    exp = cases.expected_bytes(cols[1].slice(65536, 3 * 65536))
    assert np.array_equal(off, exp[0]) and np.array_equal(pay, exp[1])
    for threads in (1, 3):
        assert np.array_equal(ora.decode(0, threads=threads), cols[0])


def test_tile_rowgroups():
    cols = [np.arange(2 * 65536, dtype=np.int32)]
    img = W.write_table(cols)
    big = W.tile_rowgroups(img, 5)
    ora = oracle.OracleTable(big)
    assert ora.n_rg == 5 and ora.n_rows == 5 * 65536
    d = ora.decode(0)
    assert np.array_equal(d[4 * 65536:], cols[0][:65536])


def test_oracle_rejects_garbage():
    with pytest.raises(IOError):
        oracle.OracleTable(np.zeros(100, dtype=np.uint8))
    img = W.write_table([np.arange(10, dtype=np.int32)])
    bad = img.copy()
    bad[:8] = 0
    with pytest.raises(IOError):
        oracle.OracleTable(bad)


def test_small_row_groups_whole_vectors():
    """Row groups of fewer than 64 vectors (COPY ... ROW_GROUP_SIZE n): every one but the last holds whole vectors."""
    import synth
    import fls_writer as W
    cols, names = synth.basic_usage(30_000)
    for rg_rows in (1024, 3072, 16384):
        img = W.write_table(cols, names=names, types=synth.basic_usage_types(), rg_rows=rg_rows)
        ora = oracle.OracleTable(img)
        assert ora.n_rows == 30_000 and ora.n_rg == -(-30_000 // rg_rows)
        for c, src in enumerate(cols):
            exp = cases.expected_bytes(src)
            got = ora.decode(c)
            em = emul.decode(img, c, ora.types[c], 0, ora.n_rg)
            if isinstance(src, StrColumn):
                assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1])
                assert np.array_equal(em[0][:30_001], got[0]) and np.array_equal(em[1][:len(got[1])], got[1])
            else:
                assert np.array_equal(got.view(np.uint8), exp) and np.array_equal(em[:30_000].view(np.uint8), exp)
    # a partial vector can only end the file: appending after it is refused
    import ctypes as C
    lib = W.lib()
    w = lib.flsw_create(1, (C.c_char_p * 1)(b"a"), (C.c_int32 * 1)(W.INT32))
    a = np.arange(1500, dtype=np.int32)
    d, o = (C.c_void_p * 1)(a.ctypes.data), (C.c_void_p * 1)(None)
    assert lib.flsw_append_rowgroup(w, 1500, d, o) == 0
    assert lib.flsw_append_rowgroup(w, 1024, d, o) != 0 and b"partial vector" in lib.flsw_error(w)
    lib.flsw_destroy(w)
