"""CPU: the oracle's ALP / ALP-RD / FSST arithmetic against golden vectors minted from the reference tree's own
code (DuckDB in-tree twins, tests/golden/make_golden.py).  These are the pins of the oracle: the FastLanes
library itself is absent from /root/reference, so FFOR/DELTA/RLE/DICT layouts stay "parity unpinned"."""
import ctypes as C
import os

import numpy as np
import pytest

import fls_writer as W
from fls_writer import StrColumn
from oracle import oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_alp_decode_value_every_exponent_factor():
    g = np.load(os.path.join(G, "alp.npz"))
    L = oracle.lib()
    assert len(g["dec_n"]) == 190 * 64
    for n, e, f, bits in zip(g["dec_n"], g["dec_e"], g["dec_f"], g["dec_bits"]):
        got = np.float64(L.flso_alp_decode_value(int(n), int(e), int(f))).view(np.uint64)
        assert got == bits, (n, e, f)


def test_alp_encode_value_matches_reference():
    g = np.load(os.path.join(G, "alp.npz"))
    L = W.lib()
    for v, e, f, n in zip(g["enc_v"].view(np.float64), g["enc_e"], g["enc_f"], g["enc_n"]):
        assert L.flsw_alp_encode_value(float(v), int(e), int(f)) == int(n), (v, e, f)


def test_alp_vectors_roundtrip_like_reference():
    # DuckDB's Compress+Decompress returns the input bits; so must encode -> oracle decode, for the same vectors
    g = np.load(os.path.join(G, "alp.npz"))
    assert np.array_equal(g["vec_in"], g["vec_out"])
    for i, bits in enumerate(g["vec_in"]):
        v = bits.view(np.float64)
        for ef in (None, (int(g["vec_e"][i]), int(g["vec_f"][i]))):  # own sampler, then DuckDB's chosen (e, f)
            img = W.write_table([v], force=[W.ALP], alp_ef=[ef])
            got = oracle.OracleTable(img).decode(0)
            assert np.array_equal(got.view(np.uint64), bits)
        # with DuckDB's (e,f) the encoded integers, base and width are DuckDB's too
        e, f = int(g["vec_e"][i]), int(g["vec_f"][i])
        enc = np.array([W.lib().flsw_alp_encode_value(float(x), e, f) for x in v], dtype=np.int64)
        dec = np.array([oracle.lib().flso_alp_decode_value(int(n), e, f) for n in enc])
        exc = dec.view(np.uint64) != bits
        assert exc.sum() == g["vec_nexc"][i]
        if (~exc).any():
            enc[exc] = enc[~exc][0]
            assert np.uint64(enc.min()) == g["vec_for"][i]
            assert np.array_equal((enc - enc.min()).astype(np.uint64), g["vec_enc"][i])


def test_alprd_formula_matches_reference():
    g = np.load(os.path.join(G, "alprd.npz"))
    L = oracle.lib()
    assert np.array_equal(g["in"], g["rt"])  # DuckDB's own Decompress returns the input
    for i in range(len(g["in"])):
        rbw, nx = int(g["rbw"][i]), int(g["nexc"][i])
        assert 48 <= rbw <= 63 and g["lbw"][i] <= 3 and g["ndict"][i] <= 8
        out = np.zeros(1024, np.uint64)
        d, l, r = (np.ascontiguousarray(g[k][i]) for k in ("dict", "left", "right"))
        xp, xl = np.ascontiguousarray(g["xpos"][i]), np.ascontiguousarray(g["xleft"][i])
        L.flso_alprd_decode(d.ctypes.data, l.ctypes.data, r.ctypes.data, rbw, 1024, xp.ctypes.data, xl.ctypes.data, nx,
                            out.ctypes.data)
        assert np.array_equal(out, g["in"][i])
        # and the writer/oracle pair on the same doubles
        img = W.write_table([g["in"][i].view(np.float64)], force=[W.ALP_RD])
        assert np.array_equal(oracle.OracleTable(img).decode(0).view(np.uint64), g["in"][i])


@pytest.mark.parametrize("name", ["comment", "shipinstruct", "user", "binary", "fixture_like"])
def test_fsst_decode_of_libfsst_output(name):
    g = np.load(os.path.join(G, "fsst.npz"))
    raw, off, sym, comp, coff = (np.ascontiguousarray(g[name + s]) for s in ("_raw", "_off", "_symtab", "_comp", "_coff"))
    L = oracle.lib()
    out = np.zeros(len(raw) + 64, np.uint8)
    pos = 0
    for i in range(len(off) - 1):
        n = L.flso_fsst_decode(sym.ctypes.data, comp.ctypes.data + int(coff[i]), int(coff[i + 1] - coff[i]),
                               out.ctypes.data + pos)
        assert n == off[i + 1] - off[i]
        pos += n
    assert np.array_equal(out[:pos], raw)
    # whole stream at once: string boundaries do not matter to the byte stream
    out2 = np.zeros(len(raw) + 64, np.uint8)
    n = L.flso_fsst_decode(sym.ctypes.data, comp.ctypes.data, len(comp), out2.ctypes.data)
    assert n == len(raw) and np.array_equal(out2[:n], raw)
    # the same strings through this repo's own FSST writer + oracle
    col = StrColumn(raw, off)
    img = W.write_table([col], force=[W.STR_FSST])
    o, p = oracle.OracleTable(img).decode(0)
    assert np.array_equal(o.astype(np.uint64), off) and np.array_equal(p, raw)


def test_reference_minted_streams_decode_through_the_container():
    """libfsst's symbol table + compressed bytes and DuckDB's ALP-RD streams, stored verbatim in FLS chunks: the oracle
    and the host emulation of the CUDA kernels (same device source) must both return the reference's input."""
    import cases
    import emul
    for name, img, exp in cases.reference_minted_images():
        ora = oracle.OracleTable(img)
        assert ora.scheme(0, 0) == (W.STR_FSST if isinstance(exp, tuple) else W.ALP_RD), name
        got = ora.decode(0)
        em = emul.decode(img, 0, ora.types[0], 0, ora.n_rg)
        if isinstance(exp, tuple):
            off, raw = exp
            n = len(off) - 1
            assert np.array_equal(got[0].astype(np.uint64), off) and np.array_equal(got[1], raw), name
            assert np.array_equal(em[0][:n + 1].astype(np.uint64), off) and np.array_equal(em[1][:len(raw)], raw), name
        else:
            assert np.array_equal(got.view(np.uint64), exp), name
            assert np.array_equal(em.view(np.uint64)[:len(exp)], exp), name


# ------------------------------------------------------------------------------------------------- ALP on FLOAT columns
def test_alpf_decode_value_every_exponent_factor():
    g = np.load(os.path.join(G, "alpf.npz"))
    L = oracle.lib()
    assert len(g["dec_n"]) == 66 * 64          # f <= e <= 10
    for n, e, f, bits in zip(g["dec_n"], g["dec_e"], g["dec_f"], g["dec_bits"]):
        got = np.float32(L.flso_alpf_decode_value(int(n), int(e), int(f))).view(np.uint32)
        assert got == bits, (n, e, f)


def test_alpf_encode_value_matches_reference():
    g = np.load(os.path.join(G, "alpf.npz"))
    L = W.lib()
    for v, e, f, n in zip(g["enc_v"].view(np.float32), g["enc_e"], g["enc_f"], g["enc_n"]):
        assert L.flsw_alpf_encode_value(float(v), int(e), int(f)) == int(n), (v, e, f)


def test_alpf_vectors_roundtrip_like_reference():
    import emul
    g = np.load(os.path.join(G, "alpf.npz"))
    assert np.array_equal(g["vec_in"], g["vec_out"])     # DuckDB's Compress + Decompress returns the input bits
    for i, bits in enumerate(g["vec_in"]):
        v = bits.view(np.float32)
        e, f = int(g["vec_e"][i]), int(g["vec_f"][i])
        for ef in (None, (e, f)):                         # own sampler, then DuckDB's chosen (e, f)
            img = W.write_table([v], force=[W.ALP], alp_ef=[ef])
            ora = oracle.OracleTable(img)
            assert ora.scheme(0, 0) == W.ALP
            assert np.array_equal(ora.decode(0).view(np.uint32), bits)
            assert np.array_equal(emul.decode(img, 0, ora.types[0], 0, 1)[:1024].view(np.uint32), bits)
        # with DuckDB's (e, f) the exception count is DuckDB's too
        enc = np.array([W.lib().flsw_alpf_encode_value(float(x), e, f) for x in v], dtype=np.int64)
        dec = np.array([oracle.lib().flso_alpf_decode_value(int(n), e, f) for n in enc], dtype=np.float32)
        assert int((dec.view(np.uint32) != bits).sum()) == int(g["vec_nexc"][i])
