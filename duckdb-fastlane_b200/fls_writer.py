"""ctypes binding of the FLS encoder/writer (csrc/fls_writer.cpp -> libflsw.so).

Mirrors the write half the reference declares but never implements
(/root/reference/src/include/fastlanes_facade.hpp:40-44: createFile / writeChunk / finalizeFile).
Used by tests and bench.py to mint .fls images; not on the timed decode path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

# enum fls_type (csrc/fls_format.h; order of /root/reference/src/type_mapping.cpp:11-62)
BOOLEAN, INT8, INT16, INT32, INT64, UINT8, UINT16, UINT32, UINT64, FLOAT, DOUBLE, STR, DATE, BYTE_ARRAY = range(1, 15)
# enum fls_scheme
(UNCOMPRESSED, CONSTANT, FFOR, DELTA, RLE, DICT, ALP, ALP_RD,
 STR_UNCOMPRESSED, STR_CONSTANT, STR_DICT, STR_FSST, STR_FSST_DICT, FFOR_PATCH, CROSS_RLE, EQUAL, STR_FSST12) = range(17)
SCHEME_NAMES = ["UNCOMPRESSED", "CONSTANT", "FFOR", "DELTA", "RLE", "DICT", "ALP", "ALP_RD",
                "STR_UNCOMPRESSED", "STR_CONSTANT", "STR_DICT", "STR_FSST", "STR_FSST_DICT", "FFOR_PATCH", "CROSS_RLE", "EQUAL", "STR_FSST12"]

TYPE_OF_DTYPE = {
    np.dtype(np.bool_): BOOLEAN, np.dtype(np.int8): INT8, np.dtype(np.int16): INT16, np.dtype(np.int32): INT32,
    np.dtype(np.int64): INT64, np.dtype(np.uint8): UINT8, np.dtype(np.uint16): UINT16,
    np.dtype(np.uint32): UINT32, np.dtype(np.uint64): UINT64, np.dtype(np.float32): FLOAT,
    np.dtype(np.float64): DOUBLE,
}
DTYPE_OF_TYPE = {v: k for k, v in TYPE_OF_DTYPE.items()}
DTYPE_OF_TYPE[DATE] = np.dtype(np.int32)

RG_ROWS = 65536

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "libflsw.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(path)
        L.flsw_create.restype = C.c_void_p
        L.flsw_create.argtypes = [C.c_uint32, C.POINTER(C.c_char_p), C.POINTER(C.c_int32)]
        L.flsw_force_scheme.argtypes = [C.c_void_p, C.c_uint32, C.c_int32, C.c_int32]
        L.flsw_append_rowgroup.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.flsw_append_rowgroup_v.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                             C.POINTER(C.c_void_p)]
        L.flsw_finish.argtypes = [C.c_void_p]
        L.flsw_data.restype = C.c_void_p
        L.flsw_data.argtypes = [C.c_void_p]
        L.flsw_size.restype = C.c_uint64
        L.flsw_size.argtypes = [C.c_void_p]
        L.flsw_error.restype = C.c_char_p
        L.flsw_error.argtypes = [C.c_void_p]
        L.flsw_write_file.argtypes = [C.c_void_p, C.c_char_p]
        L.flsw_destroy.argtypes = [C.c_void_p]
        L.flsw_force_alp_ef.argtypes = [C.c_void_p, C.c_uint32, C.c_int32, C.c_int32]
        L.flsw_alp_encode_value.argtypes = [C.c_double, C.c_int32, C.c_int32]
        L.flsw_alp_encode_value.restype = C.c_int64
        L.flsw_alpf_encode_value.argtypes = [C.c_float, C.c_int32, C.c_int32]
        L.flsw_alpf_encode_value.restype = C.c_int64
        L.flsw_inject_fsst.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        L.flsw_inject_alprd.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.flsw_tile.restype = C.c_void_p
        L.flsw_tile.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32]
        _lib = L
    return _lib


class StrColumn:
    """Arrow-style string column: payload bytes + u64 offsets[n+1]."""

    def __init__(self, data: np.ndarray, offsets: np.ndarray):
        self.data = np.ascontiguousarray(data, dtype=np.uint8)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.uint64)

    def __len__(self):
        return len(self.offsets) - 1

    @staticmethod
    def from_list(strs):
        bs = [s.encode() if isinstance(s, str) else bytes(s) for s in strs]
        off = np.zeros(len(bs) + 1, dtype=np.uint64)
        if bs:
            off[1:] = np.cumsum([len(b) for b in bs])
        return StrColumn(np.frombuffer(b"".join(bs), dtype=np.uint8), off)

    def to_list(self):
        d = self.data.tobytes()
        o = self.offsets
        return [d[int(o[i]):int(o[i + 1])] for i in range(len(self))]

    def slice(self, a, b):
        return StrColumn(self.data, self.offsets[a:b + 1])  # offsets stay absolute into data


def write_table(columns, names=None, types=None, force=None, min_bw=None, path=None, alp_ef=None, rg_rows=RG_ROWS,
                inject=None, validity=None) -> np.ndarray:
    """Encode a table into an FLS image (returned as a uint8 array; also written to `path` if given).

    columns: list of np.ndarray (fixed width) or StrColumn.   force/min_bw: optional per-column scheme / FFOR width.
    validity: optional per-column bool arrays (True = the row holds a value) or None: NULLs (flsw_append_rowgroup_v).
    rg_rows: rows per row group, a multiple of 1024 up to 65,536 (what COPY's ROW_GROUP_SIZE sets).
    inject: {(column, row group index): ("fsst", symtab, comp, comp_off) | ("alprd", rbw, lbw, ndict, dict8, left, right,
            nexc, xpos, xleft)} — streams another encoder produced, stored verbatim (flsw.h test knobs).
    """
    L = lib()
    n_cols = len(columns)
    n_rows = len(columns[0])
    names = names or [f"c{i}" for i in range(n_cols)]
    if types is None:
        types = [STR if isinstance(c, StrColumn) else TYPE_OF_DTYPE[c.dtype] for c in columns]
    cnames = (C.c_char_p * n_cols)(*[n.encode() for n in names])
    ctypes_ = (C.c_int32 * n_cols)(*types)
    w = L.flsw_create(n_cols, cnames, ctypes_)
    try:
        for i in range(n_cols):
            f = -1 if not force or force[i] is None else force[i]
            b = -1 if not min_bw or min_bw[i] is None else min_bw[i]
            if f >= 0 or b >= 0:
                L.flsw_force_scheme(w, i, f, b)
            if alp_ef and alp_ef[i] is not None:
                if L.flsw_force_alp_ef(w, i, alp_ef[i][0], alp_ef[i][1]):
                    raise ValueError("bad ALP exponent/factor")
        cols = [c if isinstance(c, StrColumn) else np.ascontiguousarray(c) for c in columns]
        if rg_rows % 1024 or not 0 < rg_rows <= RG_ROWS:
            raise ValueError("rg_rows must be a multiple of 1024 in [1024, 65536]")
        for r0 in range(0, max(n_rows, 1), rg_rows):
            r1 = min(n_rows, r0 + rg_rows)
            if r1 <= r0:
                break
            data = (C.c_void_p * n_cols)()
            offs = (C.c_void_p * n_cols)()
            keep = []
            for i, c in enumerate(cols):
                if isinstance(c, StrColumn):
                    o = np.ascontiguousarray(c.offsets[r0:r1 + 1])
                    keep.append(o)
                    data[i] = c.data.ctypes.data
                    offs[i] = o.ctypes.data
                else:
                    s = c[r0:r1]
                    keep.append(s)
                    data[i] = s.ctypes.data
                    offs[i] = None
            for i in range(n_cols):
                inj = (inject or {}).get((i, r0 // rg_rows))
                if inj is None:
                    continue
                arrs = [np.ascontiguousarray(a) for a in inj[1:] if isinstance(a, np.ndarray)]
                keep.append(arrs)
                if inj[0] == "fsst":
                    sym, comp, coff = (np.ascontiguousarray(a) for a in inj[1:4])
                    coff = coff.astype(np.uint64)
                    keep.append((sym, comp, coff))
                    rc = L.flsw_inject_fsst(w, i, sym.ctypes.data, comp.ctypes.data, coff.ctypes.data, len(coff) - 1)
                else:
                    rbw, lbw, ndict = (int(x) for x in inj[1:4])
                    d8 = np.ascontiguousarray(inj[4], dtype=np.uint16)
                    left = np.ascontiguousarray(inj[5], dtype=np.uint16).reshape(-1)
                    right = np.ascontiguousarray(inj[6], dtype=np.uint64).reshape(-1)
                    nexc = np.ascontiguousarray(np.atleast_1d(inj[7]), dtype=np.uint32)
                    xpos = np.ascontiguousarray(inj[8], dtype=np.uint16).reshape(-1)
                    xleft = np.ascontiguousarray(inj[9], dtype=np.uint16).reshape(-1)
                    keep.append((d8, left, right, nexc, xpos, xleft))
                    rc = L.flsw_inject_alprd(w, i, rbw, lbw, ndict, d8.ctypes.data, len(nexc), left.ctypes.data,
                                             right.ctypes.data, nexc.ctypes.data, xpos.ctypes.data, xleft.ctypes.data)
                if rc:
                    raise ValueError("bad injected streams for column %d" % i)
            vptr = None
            if validity and any(v is not None for v in validity):
                vptr = (C.c_void_p * n_cols)()
                for i, v in enumerate(validity):
                    if v is None:
                        continue
                    words = np.zeros((r1 - r0 + 63) // 64 * 8, dtype=np.uint8)
                    pk = np.packbits(np.asarray(v[r0:r1], dtype=bool), bitorder="little")
                    words[:len(pk)] = pk
                    keep.append(words)
                    vptr[i] = words.ctypes.data
            if vptr is not None:
                if L.flsw_append_rowgroup_v(w, r1 - r0, data, offs, vptr):
                    raise RuntimeError(L.flsw_error(w).decode())
            elif L.flsw_append_rowgroup(w, r1 - r0, data, offs):
                raise RuntimeError(L.flsw_error(w).decode())
        if L.flsw_finish(w):
            raise RuntimeError(L.flsw_error(w).decode())
        size = L.flsw_size(w)
        img = np.empty(size, dtype=np.uint8)
        C.memmove(img.ctypes.data, L.flsw_data(w), size)
        if path:
            img.tofile(path)
        return img
    finally:
        L.flsw_destroy(w)


def tile_rowgroups(image: np.ndarray, n_total: int, pinned_alloc=None) -> np.ndarray:
    """New image with `n_total` row groups: the (full) row groups of `image` repeated round-robin, each repeat a
    separate copy of the bytes."""
    L = lib()
    image = np.ascontiguousarray(image, dtype=np.uint8)
    w = L.flsw_tile(image.ctypes.data, image.size, n_total)
    if not w:
        raise RuntimeError("tile_rowgroups: image has no row groups or a partial one")
    try:
        size = L.flsw_size(w)
        out = pinned_alloc(size) if pinned_alloc else np.empty(size, dtype=np.uint8)
        C.memmove(out.ctypes.data, L.flsw_data(w), size)
        return out
    finally:
        L.flsw_destroy(w)
