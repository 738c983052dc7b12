"""Mint golden vectors from the REFERENCE TREE's own code (DuckDB's in-tree ALP / ALP-RD / FSST, same authors as
FastLanes) through oracle/_ref/libref_twin.so.  Run in the dev container only (needs /root/reference):

    python __graft_entry__.py && python tests/golden/make_golden.py

Outputs (committed, small): tests/golden/alp.npz, alprd.npz, fsst.npz.  tests/test_oracle_golden.py checks the
oracle (and the writer's encode arithmetic) against them on any box; tests/test_gpu_parity.py checks the CUDA
ALP kernel against alp.npz for every (exponent, factor) pair."""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlane_b200"))
import synth  # noqa: E402

L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_twin.so"))
L.twin_alp_encode.argtypes = [C.c_double, C.c_uint8, C.c_uint8]
L.twin_alp_encode.restype = C.c_int64
L.twin_alp_decode.argtypes = [C.c_int64, C.c_uint8, C.c_uint8]
L.twin_alp_decode.restype = C.c_double
vp = C.c_void_p
L.twin_alp_compress.argtypes = [vp, C.c_uint32] + [vp] * 7
L.twin_alp_roundtrip.argtypes = [vp, C.c_uint32, vp]
L.twin_alprd_compress.argtypes = [vp, C.c_uint32] + [vp] * 9
L.twin_fsst_compress.argtypes = [vp, vp, C.c_uint32, vp, vp, C.c_uint64, vp, vp]
L.twin_fsst_compress.restype = C.c_uint64


def p(a):
    return a.ctypes.data


def alp():
    r = np.random.Generator(np.random.PCG64(2024))
    # (1) DecodeValue for every (e, f), f <= e <= 18: 64 integers each, magnitudes spread over the int64 range
    es, fs, ns, ds = [], [], [], []
    for e in range(19):
        for f in range(e + 1):
            mags = r.integers(0, 62, 64)
            n = (r.integers(0, 2 ** 62, 64) >> (61 - mags)) * r.choice([-1, 1], 64)
            n[:4] = [0, 1, -1, 123456789]
            for x in n:
                es.append(e); fs.append(f); ns.append(int(x)); ds.append(L.twin_alp_decode(int(x), e, f))
    dec = dict(dec_e=np.array(es, np.uint8), dec_f=np.array(fs, np.uint8), dec_n=np.array(ns, np.int64),
               dec_bits=np.array(ds, np.float64).view(np.uint64))
    # (2) EncodeValue on decimal-like, integer, special and entropy doubles
    vals = np.concatenate([np.round(r.uniform(-1e5, 1e5, 400), 2), np.round(r.uniform(0, 1, 200), 6),
                           r.integers(-10 ** 9, 10 ** 9, 200).astype(np.float64), r.uniform(0, 1, 100),
                           np.array([np.nan, np.inf, -np.inf, -0.0, 0.0, 1e300, -1e300, 0.1 + 0.2, 5e-324, 9.2e18])])
    ee, ff, vv, nn = [], [], [], []
    for v in vals:
        for (e, f) in ((2, 0), (6, 0), (14, 12), (18, 18), (0, 0), (10, 5), (16, 3)):
            ee.append(e); ff.append(f); vv.append(v); nn.append(L.twin_alp_encode(float(v), e, f))
    enc = dict(enc_e=np.array(ee, np.uint8), enc_f=np.array(ff, np.uint8), enc_v=np.array(vv, np.float64).view(np.uint64),
               enc_n=np.array(nn, np.int64))
    # (3) whole vectors through DuckDB's sampler + Compress + Decompress
    vecs = [synth.alp_prices(1024, seed=5), np.round(r.uniform(0, 1000, 1024), 1), r.integers(0, 50, 1024).astype(np.float64),
            synth.alp_prices(1024, seed=6, exc_rate=0.05), np.round(r.uniform(-1, 1, 1024), 4)]
    vin, vout, ve, vf, vfor, vbw, venc, vnexc = [], [], [], [], [], [], [], []
    for v in vecs:
        v = np.ascontiguousarray(v)
        out = np.empty(1024)
        L.twin_alp_roundtrip(p(v), 1024, p(out))
        e, f, bw = np.zeros(1, np.uint8), np.zeros(1, np.uint8), np.zeros(1, np.uint8)
        fr = np.zeros(1, np.uint64)
        enc_i = np.zeros(1024, np.uint64)
        xp, xv = np.zeros(1024, np.uint16), np.zeros(1024)
        nx = L.twin_alp_compress(p(v), 1024, p(e), p(f), p(fr), p(bw), p(enc_i), p(xp), p(xv))
        vin.append(v.view(np.uint64)); vout.append(out.view(np.uint64)); ve.append(e[0]); vf.append(f[0])
        vfor.append(fr[0]); vbw.append(bw[0]); venc.append(enc_i); vnexc.append(nx)
    vec = dict(vec_in=np.array(vin), vec_out=np.array(vout), vec_e=np.array(ve), vec_f=np.array(vf),
               vec_for=np.array(vfor), vec_bw=np.array(vbw), vec_enc=np.array(venc), vec_nexc=np.array(vnexc))
    np.savez_compressed(os.path.join(HERE, "alp.npz"), **dec, **enc, **vec)


def alpf():
    """ALP on FLOAT columns: DuckDB's AlpCompression<float> (alp.hpp with AlpTypedConstants<float>)."""
    L.twin_alpf_encode.argtypes = [C.c_float, C.c_uint8, C.c_uint8]
    L.twin_alpf_encode.restype = C.c_int64
    L.twin_alpf_decode.argtypes = [C.c_int64, C.c_uint8, C.c_uint8]
    L.twin_alpf_decode.restype = C.c_float
    L.twin_alpf_roundtrip.argtypes = [vp, C.c_uint32, vp, vp, vp]
    r = np.random.Generator(np.random.PCG64(2025))
    es, fs, ns, ds = [], [], [], []
    for e in range(11):
        for f in range(e + 1):
            mags = r.integers(0, 40, 64)
            n = (r.integers(0, 2 ** 40, 64) >> (39 - mags)) * r.choice([-1, 1], 64)
            n[:4] = [0, 1, -1, 1234567]
            for x in n:
                es.append(e); fs.append(f); ns.append(int(x)); ds.append(L.twin_alpf_decode(int(x), e, f))
    dec = dict(dec_e=np.array(es, np.uint8), dec_f=np.array(fs, np.uint8), dec_n=np.array(ns, np.int64),
               dec_bits=np.array(ds, np.float32).view(np.uint32))
    vals = np.concatenate([np.round(r.uniform(-1e4, 1e4, 400), 2), np.round(r.uniform(0, 1, 200), 4),
                           r.integers(-10 ** 6, 10 ** 6, 200).astype(np.float64), r.uniform(0, 1, 100),
                           np.array([np.nan, np.inf, -np.inf, -0.0, 0.0, 1e30, -1e30, 0.1 + 0.2, 1e-45, 9.2e18])]).astype(np.float32)
    ee, ff, vv, nn = [], [], [], []
    for v in vals:
        for (e, f) in ((2, 0), (6, 0), (10, 8), (10, 10), (0, 0), (7, 5), (4, 1)):
            ee.append(e); ff.append(f); vv.append(v); nn.append(L.twin_alpf_encode(float(v), e, f))
    enc = dict(enc_e=np.array(ee, np.uint8), enc_f=np.array(ff, np.uint8), enc_v=np.array(vv, np.float32).view(np.uint32),
               enc_n=np.array(nn, np.int64))
    vecs = [np.round(r.uniform(1, 1000, 1024), 2), np.round(r.uniform(0, 100, 1024), 1), r.integers(0, 50, 1024).astype(np.float64),
            np.where(r.random(1024) < 0.05, r.normal(size=1024), np.round(r.uniform(0, 10, 1024), 3)), r.normal(size=1024)]
    vin, vout, ve, vf, vn = [], [], [], [], []
    for v in vecs:
        v = np.ascontiguousarray(v.astype(np.float32))
        out = np.empty(1024, np.float32)
        e, f = np.zeros(1, np.uint8), np.zeros(1, np.uint8)
        nx = L.twin_alpf_roundtrip(p(v), 1024, p(out), p(e), p(f))
        vin.append(v.view(np.uint32)); vout.append(out.view(np.uint32)); ve.append(e[0]); vf.append(f[0]); vn.append(nx)
    vec = dict(vec_in=np.array(vin), vec_out=np.array(vout), vec_e=np.array(ve), vec_f=np.array(vf), vec_nexc=np.array(vn))
    np.savez_compressed(os.path.join(HERE, "alpf.npz"), **dec, **enc, **vec)


def alprd():
    r = np.random.Generator(np.random.PCG64(2025))
    vecs = [synth.alp_entropy(1024, seed=9), r.uniform(0, 1000, 1024), r.normal(size=1024) * 1e-3,
            r.integers(0, 2 ** 64 - 1, 1024, dtype=np.uint64).view(np.float64)]
    out = {k: [] for k in ("in", "rbw", "lbw", "ndict", "dict", "left", "right", "nexc", "xpos", "xleft", "rt")}
    for v in vecs:
        b = np.ascontiguousarray(v).view(np.uint64)
        rbw, lbw, nd = np.zeros(1, np.uint8), np.zeros(1, np.uint8), np.zeros(1, np.uint8)
        d = np.zeros(8, np.uint16)
        left, right = np.zeros(1024, np.uint16), np.zeros(1024, np.uint64)
        xp, xl, rt = np.zeros(1024, np.uint16), np.zeros(1024, np.uint16), np.zeros(1024, np.uint64)
        nx = L.twin_alprd_compress(p(b), 1024, p(rbw), p(lbw), p(nd), p(d), p(left), p(right), p(xp), p(xl), p(rt))
        for k, x in zip(out, (b, rbw[0], lbw[0], nd[0], d, left, right, nx, xp, xl, rt)):
            out[k].append(x)
    np.savez_compressed(os.path.join(HERE, "alprd.npz"), **{k: np.array(v) for k, v in out.items()})


def fsst():
    li = synth.lineitem(3000, seed=21)
    r = np.random.Generator(np.random.PCG64(2026))
    binary = [bytes(r.integers(0, 256, int(r.integers(0, 24))).astype(np.uint8)) for _ in range(600)]
    sets = {"comment": li[15].to_list(), "shipinstruct": li[13].to_list()[:500],
            "user": synth.basic_usage(800)[0][4].to_list(), "binary": binary,
            "fixture_like": [b"FastLanes: Decoding >100 Billion Integers per Second with Scalar Code %d" % i for i in range(300)]}
    out = {}
    for name, strs in sets.items():
        n = len(strs)
        off = np.zeros(n + 1, np.uint64)
        off[1:] = np.cumsum([len(s) for s in strs])
        raw = np.frombuffer(b"".join(strs) + b"\0" * 16, np.uint8).copy()
        sym = np.zeros(2304, np.uint8)
        cap = 2 * int(off[-1]) + 16 * n + 1024
        comp = np.zeros(cap, np.uint8)
        coff = np.zeros(n + 1, np.uint64)
        rt = np.zeros(int(off[-1]) + 64, np.uint8)
        cb = L.twin_fsst_compress(p(raw), p(off), n, p(sym), p(comp), cap, p(coff), p(rt))
        assert cb > 0 and bytes(rt[:int(off[-1])]) == b"".join(strs)
        out[name + "_raw"] = raw[:int(off[-1])]
        out[name + "_off"] = off
        out[name + "_symtab"] = sym
        out[name + "_comp"] = comp[:cb]
        out[name + "_coff"] = coff
    np.savez_compressed(os.path.join(HERE, "fsst.npz"), **out)


if __name__ == "__main__":
    alp(); alpf(); alprd(); fsst()
    for f in ("alp.npz", "alpf.npz", "alprd.npz", "fsst.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
