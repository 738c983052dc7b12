"""CPU: libflsgpu.so loads, exports every symbol include/flsgpu.h declares, and the non-compute half of the ABI
(open / schema / chunk info / error text) behaves like the reference's bind.  No compute calls without a GPU."""
import ctypes as C
import os

import numpy as np
import pytest

import flsgpu
import fls_writer as W
import synth


def test_exports_every_declared_symbol():
    syms = flsgpu.header_symbols()
    assert len(syms) >= 20 and "flsgpu_decode" in syms and "flsgpu_open" in syms
    L = C.CDLL(flsgpu.lib_path())
    for s in syms:
        assert hasattr(L, s), s
    assert flsgpu.lib().flsgpu_abi_version() == 2


def test_missing_file_error_text_matches_reference():
    # /root/reference/test/sql/fastlane.test:9-12
    with pytest.raises(flsgpu.FlsError) as e:
        flsgpu.Reader.open("nonexistent.fls")
    assert "Failed to open FastLanes file: nonexistent.fls" in str(e.value)
    assert e.value.code == flsgpu.E_IO


def test_not_an_fls_file(tmp_path):
    p = tmp_path / "junk.fls"
    p.write_bytes(b"x" * 4096)
    with pytest.raises(flsgpu.FlsError) as e:
        flsgpu.Reader.open(str(p))
    assert e.value.code == flsgpu.E_FORMAT and "Failed to open FastLanes file" in str(e.value)


def test_schema_from_footer(tmp_path):
    cols, names = synth.basic_usage(70_000)
    p = str(tmp_path / "basic.fls")
    img = synth.encode(cols, names, synth.basic_usage_types(), path=p)
    for rd in (flsgpu.Reader.open(p), flsgpu.Reader.from_image(img)):
        with rd:
            assert rd.names == names
            assert rd.types == synth.basic_usage_types()
            assert rd.n_rows == 70_000 and rd.n_rowgroups == 2
            assert rd.rowgroup_rows(0) == 65536 and rd.rowgroup_rows(1) == 70_000 - 65536
            s, b = rd.chunk_info(0, 0)
            assert s == W.DELTA and b > 0
            with pytest.raises(flsgpu.FlsError):
                rd.chunk_info(9, 0)


def test_decode_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    img = synth.encode([np.arange(5000, dtype=np.int32)])
    with flsgpu.Reader.from_image(img) as rd:
        with pytest.raises(flsgpu.FlsError) as e:
            rd.decode()
        assert e.value.code == flsgpu.E_CUDA and "CUDA" in str(e.value)


def test_product_does_not_link_the_oracle():
    out = os.popen("ldd %s; nm -D --defined-only %s" % (flsgpu.lib_path(), flsgpu.lib_path())).read()
    assert "fls_oracle" not in out and "flso_" not in out and "emul_" not in out


def test_rowgroup_minmax_bounds_hold_without_a_gpu():
    """flsgpu_rowgroup_minmax reads FFOR / DELTA / CONSTANT bounds off the vector headers: no CUDA, and the bounds
    must contain every value of the row group (checked against the source columns)."""
    import numpy as np
    import synth
    import flsgpu
    import fls_writer as W
    n = 3 * 65536 + 1000
    cols = synth.lineitem(n, seed=5)
    img = synth.encode(cols, synth.LINEITEM_NAMES, synth.LINEITEM_TYPES)
    with flsgpu.Reader.from_image(img) as rd:
        known = 0
        for c, src in enumerate(cols):
            for g in range(rd.n_rowgroups):
                b = rd.rowgroup_minmax(g, c)
                if isinstance(src, np.ndarray) and src.dtype.kind in "iu":
                    part = src[g * 65536:(g + 1) * 65536].astype(np.int64)
                    scheme, _ = rd.chunk_info(g, c)
                    if scheme in (W.FFOR, W.DELTA, W.CONSTANT):
                        assert b is not None, (c, g, scheme)
                    if b is not None:
                        known += 1
                        assert b[0] <= part.min() and part.max() <= b[1], (c, g, b, part.min(), part.max())
                        if scheme == W.FFOR:   # one vector's bound is at most 2x its range: never wider than the type
                            assert b[1] - b[0] <= 2 * (int(part.max()) - int(part.min())) + 2
                else:
                    assert b is None            # doubles and strings have no integer bounds
        assert known >= 7 * rd.n_rowgroups      # 4 keys + 3 dates
    # negative values, unsigned columns, wide vectors
    r = np.random.default_rng(3)
    a = r.integers(-5000, 5000, 70000).astype(np.int32)
    u = r.integers(0, 2 ** 32, 70000, dtype=np.uint64).astype(np.uint32)
    s16 = np.sort(r.integers(-30000, 30000, 70000)).astype(np.int16)
    img = W.write_table([a, u, s16], force=[W.FFOR, W.FFOR, W.DELTA])
    with flsgpu.Reader.from_image(img) as rd:
        for g in range(2):
            sl = slice(g * 65536, (g + 1) * 65536)
            b = rd.rowgroup_minmax(g, 0)
            assert b is not None and b[0] <= a[sl].min() and a[sl].max() <= b[1] and b[0] >= -2 ** 31
            assert rd.rowgroup_minmax(g, 1) is None or rd.rowgroup_minmax(g, 1)[1] >= int(u[sl].max())
            b = rd.rowgroup_minmax(g, 2)
            assert b is None or (b[0] <= s16[sl].min() and s16[sl].max() <= b[1])


def test_rowgroup_dict_contains_without_a_gpu():
    """flsgpu_rowgroup_dict_contains answers `column = 'x'` from the row group's dictionary (STR_DICT) or its one
    value (STR_CONSTANT): the answer must agree with the source rows, row group by row group; other schemes say
    nothing."""
    modes = [b"REG AIR", b"AIR", b"RAIL", b"SHIP", b"TRUCK", b"MAIL", b"FOB", b""]
    r = np.random.default_rng(11)
    n = 3 * 65536 + 500
    idx = np.empty(n, dtype=np.int64)
    idx[:65536] = r.integers(0, 3, 65536)            # row group 0: REG AIR, AIR, RAIL
    idx[65536:2 * 65536] = r.integers(3, 8, 65536)   # row group 1: SHIP .. FOB and the empty string
    idx[2 * 65536:3 * 65536] = 5                     # row group 2: MAIL only -> a constant chunk
    idx[3 * 65536:] = r.integers(0, 8, 500)
    col = synth._pick(modes, idx)
    ids = np.arange(n, dtype=np.int64)
    img = W.write_table([ids, col], force=[None, W.STR_DICT])
    with flsgpu.Reader.from_image(img) as rd:
        for g in range(rd.n_rowgroups):
            assert rd.chunk_info(g, 1)[0] == W.STR_DICT
            here = set(idx[g * 65536:(g + 1) * 65536].tolist())
            for i, m in enumerate(modes):
                assert rd.rowgroup_dict_contains(g, 1, m) == (i in here), (g, m)
            assert rd.rowgroup_dict_contains(g, 1, "AI") is False          # a prefix of an entry is not an entry
            assert rd.rowgroup_dict_contains(g, 1, b"REG AIR\0") is False
            assert rd.rowgroup_dict_contains(g, 0, b"1") is None            # not a string column
        with pytest.raises(flsgpu.FlsError):
            rd.rowgroup_dict_contains(99, 1, b"AIR")
    # the encoder's own choice: a one-valued row group becomes STR_CONSTANT, long distinct text FSST (no answer)
    img = W.write_table([ids, col, synth.lineitem(n, seed=3)[15]])
    with flsgpu.Reader.from_image(img) as rd:
        schemes = [rd.chunk_info(g, 1)[0] for g in range(rd.n_rowgroups)]
        assert schemes[2] == W.STR_CONSTANT
        assert rd.rowgroup_dict_contains(2, 1, b"MAIL") is True and rd.rowgroup_dict_contains(2, 1, b"AIR") is False
        assert rd.rowgroup_dict_contains(2, 1, b"") is False
        assert rd.chunk_info(0, 2)[0] == W.STR_FSST and rd.rowgroup_dict_contains(0, 2, b"x") is None


def test_rowgroup_minmax_of_patched_run_length_and_aliased_chunks():
    """Zone maps of the operator variants: FFOR_PATCH = the frame plus the exception values, CROSS_RLE = min / max of the
    run values (exact), EQUAL = the bounds of the chunk it repeats."""
    import numpy as np
    import flsgpu
    import fls_writer as W
    r = np.random.default_rng(8)
    n = 70_000
    p = r.integers(100, 150, n).astype(np.int64)
    p[r.integers(0, n, 300)] = -(2 ** 40)
    p[r.integers(0, n, 300)] = 2 ** 41
    runs = np.repeat(r.integers(-500, 500, 200), 400)[:n].astype(np.int32)
    img = W.write_table([p, runs, p.copy()], force=[W.FFOR_PATCH, W.CROSS_RLE, None])
    with flsgpu.Reader.from_image(img) as rd:
        assert [rd.chunk_info(0, c)[0] for c in range(3)] == [W.FFOR_PATCH, W.CROSS_RLE, W.EQUAL]
        for g in range(rd.n_rowgroups):
            sl = slice(g * 65536, (g + 1) * 65536)
            for c, src in enumerate((p, runs, p)):
                b = rd.rowgroup_minmax(g, c)
                assert b is not None and b[0] <= src[sl].min() and src[sl].max() <= b[1], (g, c, b)
            assert rd.rowgroup_minmax(g, 1) == (int(runs[sl].min()), int(runs[sl].max()))
            assert rd.rowgroup_minmax(g, 2) == rd.rowgroup_minmax(g, 0)
