"""GPU: the CUDA decode path, called through the C ABI (libflsgpu.so), against the CPU oracle — bit-exact.

Every scheme x type x edge case of tests/cases.py, the three workload shapes, projection, row-group ranges,
slice pipelining, residency, and full-size properties."""
import numpy as np
import pytest

import cases
import flsgpu
import fls_writer as W
import synth
from fls_writer import StrColumn
from oracle import oracle

pytestmark = pytest.mark.gpu
ALL = cases.all_cases()


def check_table(img, rd=None, g0=0, g1=None, cols=None, **kw):
    ora = oracle.OracleTable(img)
    own = rd is None
    if own:
        rd = flsgpu.Reader.from_image(img, **kw)
    try:
        g1 = ora.n_rg if g1 is None else g1
        sel = list(range(ora.n_cols)) if cols is None else cols
        got = rd.decode(g0, g1, cols)
        assert len(got) == len(sel)
        for i, c in enumerate(sel):
            exp = ora.decode(c, g0, g1, threads=4)
            if isinstance(exp, tuple):
                assert np.array_equal(got[i][0], exp[0]), "offsets of column %d" % c
                assert np.array_equal(got[i][1], exp[1]), "payload of column %d" % c
            else:
                assert got[i].dtype == exp.dtype
                assert np.array_equal(got[i].view(np.uint8), exp.view(np.uint8)), "column %d" % c
    finally:
        if own:
            rd.close()
    return ora


@pytest.mark.parametrize("case", ALL, ids=[c[0] for c in ALL])
def test_case_bit_exact(case):
    name, cols, types, force, min_bw = case
    img = cases.encode_case(case)
    check_table(img)
    # and against the source values themselves
    with flsgpu.Reader.from_image(img) as rd:
        got = rd.decode()
    for c, src in enumerate(cols):
        exp = cases.expected_bytes(src)
        if isinstance(src, StrColumn):
            assert np.array_equal(got[c][0], exp[0]) and np.array_equal(got[c][1], exp[1])
        else:
            assert np.array_equal(got[c].view(np.uint8), exp)


def test_basic_usage_1m():
    # config 1: 1M-row int32/int64/double/varchar table (976 full vectors + 576-row tail, 16 row groups)
    cols, names = synth.basic_usage(1_000_000)
    img = synth.encode(cols, names, synth.basic_usage_types())
    ora = check_table(img)
    assert ora.n_rg == 16 and ora.n_rows == 1_000_000


def test_lineitem_shape_projection_ranges_slices():
    img = synth.lineitem_image(5, tail_rows=777)
    check_table(img)
    check_table(img, cols=[15, 0, 5])          # projection, any order
    check_table(img, g0=1, g1=4)               # interior range
    check_table(img, g0=5, g1=6)               # the partial row group alone
    check_table(img, g0=2, g1=2)               # empty range
    check_table(img, slice_rowgroups=1)        # one pipeline slice per row group
    check_table(img, slice_rowgroups=2, flags=flsgpu.OPEN_NO_PIN)
    with flsgpu.Reader.from_image(img, flags=flsgpu.OPEN_TIME_KERNELS) as rd:
        rd.make_resident(1, 5)
        check_table(img, rd, 1, 5)             # resident range
        check_table(img, rd, 0, 6)             # range not covered by residency -> staged
        check_table(img, rd, 2, 4, cols=[15])
        st = {s["name"]: s for s in rd.kernel_stats()}
        assert "str_fsst" in st and "alp_u64" in st and st["ffor_u64"]["launches"] >= 2


def test_small_row_groups():
    """Row groups of 1 / 3 / 16 vectors (COPY ... ROW_GROUP_SIZE n), with a partial last vector."""
    import fls_writer as W
    cols, names = synth.basic_usage(30_000)
    for rg_rows in (1024, 3072, 16384):
        img = W.write_table(cols, names=names, types=synth.basic_usage_types(), rg_rows=rg_rows)
        check_table(img)
        check_table(img, g0=1, g1=2)
        check_table(img, slice_rowgroups=2)


def test_open_by_path(tmp_path):
    p = str(tmp_path / "li.fls")
    img = synth.lineitem_image(2, tail_rows=5)
    img.tofile(p)
    with flsgpu.Reader.open(p) as rd:
        check_table(img, rd)


def test_device_output_and_timing():
    img = synth.lineitem_image(3)
    with flsgpu.Reader.from_image(img) as rd:
        rd.make_resident()
        with rd.decode_batch(to_host=False) as b:
            t = b.timing
            assert t.h2d_bytes == 0 and t.d2h_bytes == 0 and t.kernel_launches > 0 and t.ms_kernels > 0
            assert t.decoded_bytes > t.compressed_bytes > 0
            assert all(c.data or c.offsets for c in b.columns)
        with rd.decode_batch(to_host=True) as b:
            assert b.timing.d2h_bytes >= b.timing.decoded_bytes


def test_ffor_sweep_full_width_properties():
    # config 2 at reduced N per width but every width: sum / min / max of the decoded column match the source
    n = 16 * 65536
    for bits in (32, 64):
        for w in range(bits + 1):
            v = synth.ffor_values(bits, w, n)
            img = W.write_table([v], force=[W.FFOR], min_bw=[w])
            with flsgpu.Reader.from_image(img) as rd:
                got = rd.decode()[0]
            assert np.array_equal(got, v), (bits, w)


def test_large_single_column_properties():
    # 2^26 int64 values (1024 row groups): idempotence (two decodes identical) + exact equality with the source
    n = 1 << 26
    v = synth.ffor_values(64, 21, n)
    img = W.write_table([v], force=[W.FFOR], min_bw=[21])
    with flsgpu.Reader.from_image(img) as rd:
        a = rd.decode()[0]
        b = rd.decode()[0]
    assert np.array_equal(a, v) and np.array_equal(a, b)


def test_alp_large_bit_exact():
    n = 1 << 22
    for v in (synth.alp_prices(n), synth.alp_entropy(n)):
        img = W.write_table([v])
        with flsgpu.Reader.from_image(img) as rd:
            got = rd.decode()[0]
        assert np.array_equal(got.view(np.uint64), v.view(np.uint64))


def test_string_payload_limit_is_an_error_not_a_truncation():
    img = synth.encode([np.arange(10, dtype=np.int32)])
    with flsgpu.Reader.from_image(img) as rd:
        with pytest.raises(flsgpu.FlsError):
            rd.decode(0, 5)
        with pytest.raises(flsgpu.FlsError):
            rd.decode(cols=[3])


def test_alp_kernel_against_reference_golden_every_exponent_factor():
    # tests/golden/alp.npz: DuckDB's own DecodeValue(n, e, f) for all 190 (e, f) pairs.  One table with one column
    # per pair, the writer pinned to that pair: the CUDA kernel must reproduce the reference's bits.
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "alp.npz"))
    cols, efs, want = [], [], []
    for e in range(19):
        for f in range(e + 1):
            m = (g["dec_e"] == e) & (g["dec_f"] == f)
            bits = g["dec_bits"][m]
            cols.append(bits.view(np.float64)); efs.append((e, f)); want.append(bits)
    img = W.write_table(cols, force=[W.ALP] * len(cols), alp_ef=efs)
    with flsgpu.Reader.from_image(img) as rd:
        got = rd.decode()
    for c, bits in enumerate(want):
        assert np.array_equal(got[c].view(np.uint64), bits), efs[c]


def test_alpf_kernel_against_reference_golden_every_exponent_factor():
    # tests/golden/alpf.npz: DuckDB's own AlpCompression<float>::DecodeValue(n, e, f) for all 66 (e, f) pairs, f <= e <= 10
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "alpf.npz"))
    cols, efs, want = [], [], []
    for e in range(11):
        for f in range(e + 1):
            m = (g["dec_e"] == e) & (g["dec_f"] == f)
            order = np.argsort(g["dec_n"][m])
            ns, bits = g["dec_n"][m][order], g["dec_bits"][m][order]
            # a FLOAT vector's encoded integers must span less than 2^32: one column per run of close-by integers
            start = 0
            for i in range(1, len(ns) + 1):
                if i == len(ns) or int(ns[i]) - int(ns[start]) >= 2 ** 31:
                    grp = np.concatenate([bits[start:i], np.full(64 - (i - start), bits[i - 1], dtype=np.uint32)])  # 64 rows each
                    cols.append(grp.view(np.float32)); efs.append((e, f)); want.append(grp)
                    start = i
    img = W.write_table(cols, force=[W.ALP] * len(cols), alp_ef=efs)
    ora = oracle.OracleTable(img)
    assert all(ora.scheme(0, c) == W.ALP for c in range(len(cols)))
    with flsgpu.Reader.from_image(img) as rd:
        got = rd.decode()
    for c, bits in enumerate(want):
        assert np.array_equal(got[c].view(np.uint32), bits), efs[c]
    # whole vectors DuckDB's sampler + Compress + Decompress returned unchanged
    for bits in g["vec_in"]:
        img = W.write_table([bits.view(np.float32)], force=[W.ALP])
        with flsgpu.Reader.from_image(img) as rd:
            assert np.array_equal(rd.decode()[0].view(np.uint32), bits)


def test_fsst_kernel_on_reference_golden_strings():
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fsst.npz"))
    for name in ("comment", "shipinstruct", "user", "binary", "fixture_like"):
        col = StrColumn(g[name + "_raw"], g[name + "_off"])
        for f in (W.STR_FSST, W.STR_FSST_DICT):
            img = W.write_table([col], force=[f])
            with flsgpu.Reader.from_image(img) as rd:
                o, p = rd.decode()[0]
            assert np.array_equal(o.astype(np.uint64), g[name + "_off"]) and np.array_equal(p, g[name + "_raw"])


def test_resident_plan_cache_reuse_is_exact():
    # repeated device-output decodes of a resident range reuse the cached descriptors; results must not change,
    # also after a different range / projection was decoded in between
    img = synth.lineitem_image(4)
    ora = oracle.OracleTable(img)
    with flsgpu.Reader.from_image(img) as rd:
        rd.make_resident()
        for rnd in range(4):
            cols = None if rnd != 2 else [15, 3]
            g0, g1 = (0, 4) if rnd != 1 else (1, 3)
            with rd.decode_batch(g0, g1, cols, to_host=False) as b:
                assert rnd != 3 or b.timing.ms_host_plan < 0.05
                for i, c in enumerate(range(ora.n_cols) if cols is None else cols):
                    exp = ora.decode(c, g0, g1)
                    got = b.fetch(i)
                    if isinstance(exp, tuple):
                        assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1])
                    else:
                        assert np.array_equal(got.view(np.uint8), exp.view(np.uint8))
        # rounds 0 and 3 decode the same range into the same pooled buffers
        with rd.decode_batch(to_host=False) as b:
            assert b.timing.ms_host_plan < 0.05


def test_kernels_decode_streams_the_reference_encoders_wrote():
    # libfsst's own symbol tables + compressed bytes, DuckDB's own ALP-RD dictionaries / parts / exceptions, verbatim
    # in the container (writer inject knobs): the CUDA kernels consume bytes the reference tree produced
    for name, img, exp in cases.reference_minted_images():
        for kw in ({}, {"slice_rowgroups": 1}):
            with flsgpu.Reader.from_image(img, **kw) as rd:
                got = rd.decode()[0]
            if isinstance(exp, tuple):
                assert np.array_equal(got[0].astype(np.uint64), exp[0]) and np.array_equal(got[1], exp[1]), name
            else:
                assert np.array_equal(got.view(np.uint64), exp), name


def _concat(parts):
    """Concatenate per-shard decodes of one column: fixed width as is, strings with rebased offsets."""
    if isinstance(parts[0], tuple):
        offs, pays, base = [np.zeros(1, np.uint64)], [], 0
        for o, p in parts:
            offs.append(o[1:].astype(np.uint64) + base)
            pays.append(p)
            base += int(o[-1])
        return np.concatenate(offs), np.concatenate(pays)
    return np.concatenate(parts)


@pytest.mark.parametrize("n_shards", [2, 3, 4, 8])
def test_shards_reassemble_to_the_single_range_decode(n_shards):
    # SURVEY.md section 4 last bullet / section 8e: rank r decodes rowgroup_range(r, N, R) with its own handle (its own
    # device when the box has several); the concatenation must be the bytes of the one-range decode — string payload
    # bases across shard and pipeline-slice boundaries included.  slice_rowgroups is smaller than a shard.
    import shard
    import torch
    img = synth.lineitem_image(9, tail_rows=333)   # 10 row groups, the last one partial
    with flsgpu.Reader.from_image(img) as rd:
        whole = rd.decode()
        n_rg = rd.n_rowgroups
    ndev = torch.cuda.device_count()
    parts = []
    for r in range(n_shards):
        g0, g1 = shard.rowgroup_range(r, n_shards, n_rg)
        with flsgpu.Reader.from_image(img, device=r % ndev, slice_rowgroups=1 + r % 2) as rd:
            parts.append(rd.decode(g0, g1))
    for c in range(len(whole)):
        got = _concat([p[c] for p in parts])
        if isinstance(whole[c], tuple):
            assert np.array_equal(got[0], whole[c][0].astype(np.uint64)) and np.array_equal(got[1], whole[c][1]), c
        else:
            assert np.array_equal(got.view(np.uint8), whole[c].view(np.uint8)), c
