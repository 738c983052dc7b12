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


def test_projected_columns_alone_are_staged(tmp_path):
    """A decode of a few columns copies only their chunks' byte ranges to the device (projected_ranges, csrc/flsgpu.cu);
    the other columns' bytes in the staging buffer are never written and never read.  Every projection must still decode
    to the oracle's bytes — by path (mapped file, bounce buffer), from a pageable image and from a page-locked read of the
    file, in slices shorter than the range — including an EQUAL chunk whose twin column is not projected, and the H2D
    byte count must shrink with the projection."""
    import fls_writer as W
    r = np.random.default_rng(21)
    n = 65536 * 3 + 777
    a = r.integers(0, 1 << 20, n).astype(np.int64)
    d = np.round(r.uniform(0, 1000, n), 2)
    li = synth.lineitem(n, seed=4)
    cols = [a, li[15], d, a.copy(), li[14], d.copy(), li[8], r.integers(-9, 9, n).astype(np.int32)]
    img = W.write_table(cols)
    path = str(tmp_path / "p.fls")
    img.tofile(path)
    ora = oracle.OracleTable(img)
    projections = [[3], [5, 3], [1], [7, 0], [4, 6, 2], [6], list(range(8))]
    for mode in ("path", "image", "load"):
        if mode == "path":
            rd = flsgpu.Reader.open(path, slice_rowgroups=2)
        elif mode == "load":
            rd = flsgpu.Reader.open(path, flags=flsgpu.OPEN_LOAD_FILE, slice_rowgroups=2)
        else:
            rd = flsgpu.Reader.from_image(img, slice_rowgroups=2)
        with rd:
            h2d = []
            for p in projections:
                with rd.decode_batch(0, rd.n_rowgroups, cols=p, to_host=True) as b:
                    for i, c in enumerate(p):
                        assert _same(b.numpy(i), ora.decode(c)), (mode, p, c)
                    h2d.append(int(b.timing.h2d_bytes))
                # a sub-range, too
                with rd.decode_batch(1, 3, cols=p, to_host=True) as b:
                    for i, c in enumerate(p):
                        assert _same(b.numpy(i), ora.decode(c, 1, 3)), (mode, p, c, "sub-range")
            assert h2d[-1] >= img.size * 0.9 and h2d[0] < h2d[-1] / 3 and h2d[2] < h2d[-1], (mode, h2d)
