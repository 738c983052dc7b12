"""Seeded parity cases shared by the CPU (oracle / emulation) and GPU test files.

Each case is (name, columns, types, force, min_bw): a small table that pins one scheme x type x edge.
Edge cases follow SURVEY.md §4: tail vectors (n % 1024 != 0), W = 0 and W = T, ALP exceptions
(NaN / +-inf / -0.0 / huge), empty strings and strings longer than DuckDB's 12-byte inline threshold.
"""
import numpy as np

import fls_writer as W
import synth
from fls_writer import StrColumn

INT_DTYPES = {8: np.int8, 16: np.int16, 32: np.int32, 64: np.int64}
UINT_DTYPES = {8: np.uint8, 16: np.uint16, 32: np.uint32, 64: np.uint64}


def rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def ffor_case(bits, width, n, seed=None):
    v = synth.ffor_values(bits, width, n, seed)
    return ("ffor_i%d_w%d" % (bits, width), [v], None, [W.FFOR], [width])


def all_ffor_cases(n=3 * 1024 + 17):
    out = []
    for bits in (8, 16, 32, 64):
        for w in range(bits + 1):
            out.append(ffor_case(bits, w, n))
    return out


def delta_cases(n=70_000):
    r = rng(11)
    out = []
    for bits in (8, 16, 32, 64):
        dt = INT_DTYPES[bits]
        step = r.integers(0, 3 if bits == 8 else 50, n)
        v = np.cumsum(step).astype(np.int64)
        v = (v % (2 ** (bits - 1))).astype(dt) if bits < 64 else v * 1_000_003
        out.append(("delta_i%d" % bits, [v.astype(dt)], None, [W.DELTA], None))
    # descending + wrap-around deltas
    v = (np.arange(n, dtype=np.int64) * -7919 + 2 ** 40).astype(np.int64)
    out.append(("delta_i64_desc", [v], None, [W.DELTA], None))
    v = r.integers(-2 ** 31, 2 ** 31, n).astype(np.int32)
    out.append(("delta_i32_random_fullwidth", [v], None, [W.DELTA], None))
    v = np.full(n, 77, dtype=np.int32)
    out.append(("delta_i32_const_w0", [v], None, [W.DELTA], None))
    # 64-bit keys, one delta width per vector: 1 .. 16 bits take the register-resident path of DecDELTA::run64_narrow,
    # 17 .. 20 the generic one; the second column adds a negative frame of reference (descending runs)
    width = 1 + (np.arange(n) // 1024) % 20
    step = (r.integers(0, 2 ** 62, n) % (2 ** width.astype(np.int64))).astype(np.int64)
    step[np.arange(n) % 1024 == 5] = 0
    step[np.arange(n) % 1024 == 6] = 2 ** width[np.arange(n) % 1024 == 6].astype(np.int64) - 1
    up = np.cumsum(step).astype(np.int64) + 10 ** 12
    down = (2 ** 60 - np.cumsum(step + (np.arange(n) // 1024) * 3 - 1000)).astype(np.int64)
    out.append(("delta_i64_width_sweep", [up, down, up.astype(np.uint64)], None, [W.DELTA, W.DELTA, W.DELTA], None))
    return out


def rle_cases(n=70_000):
    r = rng(12)
    out = []
    for bits in (8, 16, 32, 64):
        dt = INT_DTYPES[bits]
        runs = r.integers(1, 400, n // 50)
        vals = r.integers(-100, 100, len(runs))
        v = np.repeat(vals, runs)[:n]
        if len(v) < n:
            v = np.concatenate([v, np.full(n - len(v), 5)])
        out.append(("rle_i%d" % bits, [v.astype(dt)], None, [W.RLE], None))
    d = np.repeat(r.uniform(0, 1, n // 100 + 1), 100)[:n]
    out.append(("rle_f64", [d], None, [W.RLE], None))
    return out


def dict_cases(n=70_000):
    r = rng(13)
    out = []
    for bits in (8, 16, 32, 64):
        dt = INT_DTYPES[bits]
        pool = r.integers(-2 ** (bits - 1), 2 ** (bits - 1) - 1, 37, dtype=np.int64).astype(dt)
        v = pool[r.integers(0, len(pool), n)]
        out.append(("dict_i%d" % bits, [v], None, [W.DICT], None))
    pool = r.uniform(-1e9, 1e9, 300)
    out.append(("dict_f64", [pool[r.integers(0, 300, n)]], None, [W.DICT], None))
    out.append(("dict_single", [np.full(n, -5, dtype=np.int32)], None, [W.DICT], None))
    return out


def misc_fixed_cases(n=70_000):
    r = rng(14)
    out = [
        ("uncompressed_i32", [r.integers(-2 ** 31, 2 ** 31, n).astype(np.int32)], None, [W.UNCOMPRESSED], None),
        ("uncompressed_f64", [r.normal(size=n)], None, [W.UNCOMPRESSED], None),
        ("uncompressed_u8", [r.integers(0, 256, n).astype(np.uint8)], None, [W.UNCOMPRESSED], None),
        ("constant_i64", [np.full(n, -123456789012, dtype=np.int64)], None, [W.CONSTANT], None),
        ("constant_i16", [np.full(n, -3, dtype=np.int16)], None, [W.CONSTANT], None),
        ("constant_f64", [np.full(n, 2.5)], None, [W.CONSTANT], None),
        ("bool_auto", [r.integers(0, 2, n) == 1], None, None, None),
        ("float32_auto", [r.normal(size=n).astype(np.float32)], None, None, None),
        ("u16_auto", [r.integers(0, 60000, n).astype(np.uint16)], None, None, None),
        ("u64_auto_high", [r.integers(2 ** 63, 2 ** 64 - 1, n, dtype=np.uint64)], None, None, None),
        ("date_auto", [(19000 + r.integers(0, 3000, n)).astype(np.int32)], [W.DATE], None, None),
        ("one_row", [np.array([42], dtype=np.int32)], None, None, None),
        ("exactly_one_vector", [np.arange(1024, dtype=np.int64)], None, None, None),
        ("exactly_one_rowgroup", [np.arange(65536, dtype=np.int32)], None, None, None),
        ("rowgroup_plus_one", [np.arange(65537, dtype=np.int32) * 3], None, None, None),
    ]
    return out


def variant_cases(n=70_000):
    """FFOR_PATCH (FastLanes SLPATCH; FREQUENCY when the frame is one value), CROSS_RLE and EQUAL."""
    r = rng(15)
    out = []
    for bits in (8, 16, 32, 64):
        dt = INT_DTYPES[bits]
        v = r.integers(0, 50, n).astype(np.int64)
        pos = r.integers(0, n, n // 100)
        v[pos] = r.integers(-2 ** (bits - 1), 2 ** (bits - 1) - 1, len(pos), dtype=np.int64)     # 1 % outliers, full width
        out.append(("slpatch_i%d" % bits, [v.astype(dt)], None, [W.FFOR_PATCH], None))
        f = np.full(n, -7, dtype=np.int64)
        pos = r.integers(0, n, n // 50)
        f[pos] = r.integers(0, 100, len(pos))                                                       # one frequent value
        out.append(("frequency_i%d" % bits, [f.astype(dt)], None, [W.FFOR_PATCH], None))
        runs = r.integers(1, 9000, 64)
        vals = r.integers(-100, 100, len(runs))
        c = np.repeat(vals, runs)[:n]
        c = np.concatenate([c, np.full(max(0, n - len(c)), 3)])
        out.append(("cross_rle_i%d" % bits, [c.astype(dt)], None, [W.CROSS_RLE], None))
    out.append(("slpatch_auto_pick", [np.where(r.random(n) < 0.002, 2 ** 40, r.integers(0, 9, n)).astype(np.int64)], None, None, None))
    out.append(("cross_rle_single_run", [np.full(n, 9, dtype=np.int32)], None, [W.CROSS_RLE], None))
    out.append(("cross_rle_f64", [np.repeat(r.normal(size=40), 2000)[:n]], None, [W.CROSS_RLE], None))
    out.append(("slpatch_all_exceptions_w0", [r.integers(-2 ** 62, 2 ** 62, 2048).astype(np.int64)], None, [W.FFOR_PATCH], None))
    a = r.integers(0, 1000, n).astype(np.int32)
    d = np.round(r.uniform(0, 100, n), 2)
    out.append(("equal_columns", [a, d, a.copy(), d.copy(), a.copy()], None, None, None))
    out.append(("equal_first_column_has_no_twin", [a, a.copy()], None, None, None))
    return out


SPECIALS = np.array([np.nan, np.inf, -np.inf, -0.0, 1e300, 0.1 + 0.2, 5e-324, -1e-310, 1.7976931348623157e308])


def alp_cases(n=70_000):
    r = rng(15)
    prices = synth.alp_prices(n, exc_rate=0.002)
    out = [("alp_prices", [prices], None, [W.ALP], None)]
    v = np.round(r.uniform(-500, 500, n), 1)
    out.append(("alp_1dec_signed", [v], None, [W.ALP], None))
    v = r.integers(-10 ** 6, 10 ** 6, n).astype(np.float64)
    out.append(("alp_integers", [v], None, [W.ALP], None))
    v = np.round(r.uniform(0, 1, n), 6)
    out.append(("alp_6dec", [v], None, [W.ALP], None))
    v = np.round(r.uniform(0, 1000, n), 2)
    v[r.choice(n, n // 3, replace=False)] = SPECIALS[r.integers(0, len(SPECIALS), n // 3)]
    out.append(("alp_many_exceptions", [v], None, [W.ALP], None))
    v = np.full(n, np.nan)
    out.append(("alp_all_nan", [v], None, [W.ALP], None))
    v = np.full(n, 12.5)
    out.append(("alp_constant_w0", [v], None, [W.ALP], None))
    v = np.round(r.uniform(0, 9e15, n))
    out.append(("alp_huge_ints_wide", [v], None, [W.ALP], None))
    # one integer width per vector, 30 .. 53 bits: 33 and up are decoded in two passes through the 4 KB slot
    bits = 30 + (np.arange(n) // 1024) % 24
    v = (r.integers(0, 2 ** 62, n) % (2 ** bits.astype(np.int64))).astype(np.float64)
    v[np.arange(n) % 1024 == 3] = 0.0
    v[np.arange(n) % 1024 == 4] = (2.0 ** bits[np.arange(n) % 1024 == 4]) - 1.0
    out.append(("alp_width_sweep_30_53", [v, -v], None, [W.ALP, W.ALP], None))
    out.append(("alprd_entropy", [synth.alp_entropy(n)], None, [W.ALP_RD], None))
    v = r.uniform(0, 1, n) * 1000.0
    out.append(("alprd_uniform", [v], None, [W.ALP_RD], None))
    v = r.integers(0, 2 ** 64 - 1, n, dtype=np.uint64).view(np.float64)
    out.append(("alprd_random_bits", [v], None, [W.ALP_RD], None))
    v = synth.alp_entropy(n)
    v[r.choice(n, 500, replace=False)] = SPECIALS[r.integers(0, len(SPECIALS), 500)]
    out.append(("alprd_with_specials", [v], None, [W.ALP_RD], None))
    # FLOAT columns: ALP in float arithmetic (AlpTypedConstants<float>), u32 stream
    f32 = np.float32
    out.append(("alpf_prices_2dec", [np.round(r.uniform(1, 1000, n), 2).astype(f32)], None, [W.ALP], None))
    out.append(("alpf_1dec_signed", [np.round(r.uniform(-500, 500, n), 1).astype(f32)], None, [W.ALP], None))
    out.append(("alpf_integers", [r.integers(-10 ** 5, 10 ** 5, n).astype(f32)], None, [W.ALP], None))
    v = np.round(r.uniform(0, 100, n), 2).astype(f32)
    sp = np.array([np.nan, np.inf, -np.inf, -0.0, 1e30, 0.1 + 0.2, 1e-45, 3.4e38], dtype=f32)
    v[r.choice(n, n // 4, replace=False)] = sp[r.integers(0, len(sp), n // 4)]
    out.append(("alpf_many_exceptions", [v], None, [W.ALP], None))
    out.append(("alpf_all_nan", [np.full(n, np.nan, dtype=f32)], None, [W.ALP], None))
    out.append(("alpf_constant_w0", [np.full(n, 2.5, dtype=f32)], None, [W.ALP], None))
    out.append(("float_auto_entropy", [r.normal(size=n).astype(f32)], None, None, None))   # (ALP does not pay: another scheme)
    out.append(("float_auto_prices", [np.round(r.uniform(1, 1000, n), 2).astype(f32)], None, None, None))
    out.append(("double_auto_prices", [prices], None, None, None))
    out.append(("double_auto_entropy", [synth.alp_entropy(n)], None, None, None))
    return out


def _rand_strings(r, n, lo, hi, alphabet=b"abcdefghijklmnopqrstuvwxyz ABCDE0123456789"):
    lens = r.integers(lo, hi + 1, n)
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    a = np.frombuffer(alphabet, dtype=np.uint8)
    return StrColumn(a[r.integers(0, len(a), int(off[-1]))], off)


def string_cases(n=70_000):
    r = rng(16)
    out = []
    li = synth.lineitem(n, seed=3)
    comment, shipmode, returnflag = li[15], li[14], li[8]
    for f, nm in ((W.STR_UNCOMPRESSED, "uncompressed"), (W.STR_FSST, "fsst")):
        out.append(("str_%s_comment" % nm, [comment], None, [f], None))
        out.append(("str_%s_random_0_40" % nm, [_rand_strings(r, n, 0, 40)], None, [f], None))
        out.append(("str_%s_bytes_all_values" % nm, [_rand_strings(r, 5000, 0, 20, bytes(range(256)))], None, [f], None))
        out.append(("str_%s_all_empty" % nm, [StrColumn(np.empty(0, np.uint8), np.zeros(3001, np.uint64))], None, [f], None))
    for f, nm in ((W.STR_DICT, "dict"), (W.STR_FSST_DICT, "fsst_dict")):
        out.append(("str_%s_shipmode" % nm, [shipmode], None, [f], None))
        out.append(("str_%s_returnflag" % nm, [returnflag], None, [f], None))
        pool = _rand_strings(r, 500, 0, 60).to_list()
        pick = r.integers(0, len(pool), n)
        out.append(("str_%s_500_entries" % nm, [StrColumn.from_list([pool[i] for i in pick])], None, [f], None))
    for L in (2, 4, 8):
        pool = [bytes(r.integers(65, 91, L).astype(np.uint8)) for _ in range(9)]
        out.append(("str_dict_uniform_len%d" % L, [StrColumn.from_list([pool[i] for i in r.integers(0, 9, n)])], None,
                    [W.STR_DICT], None))
    letters = [bytes([65 + i]) for i in range(20)]
    out.append(("str_dict_uniform_len1_20_entries", [StrColumn.from_list([letters[i] for i in r.integers(0, 20, n)])], None,
                [W.STR_DICT], None))
    # first row group has mixed lengths (odd payload size), second one is fixed-length: unaligned fast path
    mixed = [b"ab", b"c", b"defgh"]
    strs = [mixed[i] for i in r.integers(0, 3, 65536)] + [(b"XY", b"ZW")[i] for i in r.integers(0, 2, n - 65536 + 4000)]
    out.append(("str_dict_mixed_then_uniform", [StrColumn.from_list(strs)], None, [W.STR_DICT], None))
    out.append(("str_dict_entries_17_bytes", [li[13]], None, [W.STR_DICT], None))
    # quarters larger than the staging buffer: staged in 2 / 4 / 8 groups of lanes, byte-wise beyond
    for lo, hi in ((20, 40), (40, 90), (90, 190), (300, 700)):
        pool = _rand_strings(r, 40, lo, hi).to_list()
        out.append(("str_dict_40_entries_len_%d_%d" % (lo, hi), [StrColumn.from_list([pool[i] for i in r.integers(0, 40, 9000)])],
                    None, [W.STR_DICT], None))
    # flag columns (VecSTRFLAGS: up to 8 one-byte entries, full vectors): behind a row group with an odd payload size
    # (unaligned stores), with 8 entries including the bytes 0x00 and 0xFF, with a wide index stream, and a 9-entry
    # dictionary that must stay with the table kernel
    flags5 = [b"\x00", b"\xff", b"N", b"R", b"A"]
    strs = [mixed[i] for i in r.integers(0, 3, 65536)] + [flags5[i] for i in r.integers(0, 5, 65536 + 3000)]
    out.append(("str_flags_after_odd_payload", [StrColumn.from_list(strs)], None, [W.STR_DICT], None))
    flags8 = [bytes([c]) for c in b"ABCDEFG\x7f"]
    out.append(("str_flags_8_entries", [StrColumn.from_list([flags8[i] for i in r.integers(0, 8, 2 * 65536)])], None,
                [W.STR_DICT], None))
    out.append(("str_flags_8_entries_index_bw17", [StrColumn.from_list([flags8[i] for i in r.integers(0, 8, 65536 + 1024)])],
                None, [W.STR_DICT], [17]))
    flags9 = flags8 + [b"z"]
    out.append(("str_dict_9_one_byte_entries", [StrColumn.from_list([flags9[i] for i in r.integers(0, 9, 65536)])], None,
                [W.STR_DICT], None))
    out.append(("str_flags_one_entry", [StrColumn.from_list([b"F"] * 65536)], None, [W.STR_DICT], None))
    # index widths beyond the prefetch halves (<= 10 bits): whole-buffer staging (<= 20) and quarter-wise staging
    for bw in (11, 20, 21, 27, 32):
        out.append(("str_dict_shipmode_index_bw%d" % bw, [shipmode], None, [W.STR_DICT], [bw]))
    big = _rand_strings(r, 3000, 0, 12).to_list()
    out.append(("str_dict_3000_entries", [StrColumn.from_list([big[i] for i in r.integers(0, len(big), n)])], None,
                [W.STR_DICT], None))
    out.append(("str_dict_3000_entries_bw23", [StrColumn.from_list([big[i] for i in r.integers(0, len(big), n)])], None,
                [W.STR_DICT], [23]))
    # FSST12: 12-bit codes, up to 4096 symbols, no escape
    out.append(("str_fsst12_comment", [li[15]], None, [W.STR_FSST12], None))
    out.append(("str_fsst12_shipinstruct", [li[13]], None, [W.STR_FSST12], None))
    out.append(("str_fsst12_random_0_40", [_rand_strings(r, n, 0, 40)], None, [W.STR_FSST12], None))
    out.append(("str_fsst12_binary", [_rand_strings(r, 5000, 0, 90, alphabet=bytes(range(256)))], None, [W.STR_FSST12], None))
    out.append(("str_fsst12_all_empty", [StrColumn.from_list([b""] * 3000)], None, [W.STR_FSST12], None))
    # compressed bytes of one vector exceed the kernel's shared-memory window -> streaming fallback path
    out.append(("str_fsst_long_fallback", [_rand_strings(r, 9000, 60, 200)], None, [W.STR_FSST], None))
    # pass sizes 128 / 64 / 32 strings (a pass of 256 would not fit the staging buffer), and a mix within one column
    for lo, hi in ((35, 60), (70, 120), (150, 250)):
        out.append(("str_fsst_len_%d_%d" % (lo, hi), [_rand_strings(r, 20000, lo, hi)], None, [W.STR_FSST], None))
    mix = _rand_strings(r, 3000, 0, 12).to_list() + _rand_strings(r, 3000, 150, 260).to_list() + comment.to_list()[:6000]
    out.append(("str_fsst_mixed_lengths", [StrColumn.from_list([mix[i] for i in r.permutation(len(mix))])], None,
                [W.STR_FSST], None))
    out.append(("str_fsst_dict_long_entries", [StrColumn.from_list([pool2 for pool2 in
                [b"entry-%04d-" % (i % 300) + b"x" * (i % 300 % 90) for i in range(n)]])], None, [W.STR_FSST_DICT], None))
    out.append(("str_constant", [StrColumn.from_list([b"FastLanes"] * 5000)], None, [W.STR_CONSTANT], None))
    out.append(("str_constant_empty", [StrColumn.from_list([b""] * 2500)], None, [W.STR_CONSTANT], None))
    out.append(("str_auto_user", [synth.basic_usage(n)[0][4]], None, None, None))
    out.append(("str_auto_comment", [comment], None, None, None))
    out.append(("str_long_values", [_rand_strings(r, 3000, 100, 900)], None, None, None))
    return out


def all_cases():
    return (all_ffor_cases() + delta_cases() + rle_cases() + dict_cases() + misc_fixed_cases() + variant_cases() + alp_cases() +
            string_cases())


def encode_case(case):
    name, cols, types, force, min_bw = case
    return W.write_table(cols, types=types, force=force, min_bw=min_bw)


def expected_bytes(col):
    """What a decoder must produce for a source column: ndarray bytes, or (offsets u32, payload)."""
    if isinstance(col, StrColumn):
        off = (col.offsets - col.offsets[0]).astype(np.uint32)
        pay = col.data[int(col.offsets[0]):int(col.offsets[-1])]
        return off, pay
    return np.ascontiguousarray(col).view(np.uint8)


def reference_minted_images():
    """Tables whose FSST / ALP-RD chunks hold bytes the REFERENCE TREE's encoders produced (tests/golden/*.npz, minted by
    oracle/ref_twin.cpp from duckdb/third_party/fsst and duckdb's alprd.hpp), stored verbatim by the writer's inject
    knobs: [(name, image, expected)] with expected = (offsets u64, payload) or the doubles' bit patterns."""
    import os
    G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    out = []
    g = np.load(os.path.join(G, "fsst.npz"))
    for name in ("comment", "shipinstruct", "user", "binary", "fixture_like"):
        raw, off = g[name + "_raw"], g[name + "_off"]
        img = W.write_table([StrColumn(raw, off)],
                            inject={(0, 0): ("fsst", g[name + "_symtab"], g[name + "_comp"], g[name + "_coff"])})
        out.append(("libfsst_" + name, img, (off, raw)))
    g = np.load(os.path.join(G, "alprd.npz"))
    nv = len(g["in"])
    inj = {(0, v): ("alprd", g["rbw"][v], g["lbw"][v], g["ndict"][v], g["dict"][v], g["left"][v], g["right"][v],
                    g["nexc"][v], g["xpos"][v], g["xleft"][v]) for v in range(nv)}
    img = W.write_table([g["in"].reshape(-1).view(np.float64)], rg_rows=1024, inject=inj)
    out.append(("duckdb_alprd", img, g["in"].reshape(-1)))
    return out
