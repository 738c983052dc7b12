"""ctypes binding of the CPU oracle (oracle/fls_oracle.c -> libfls_oracle.so).  TEST INFRASTRUCTURE ONLY.

May be imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None

_DT = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.uint8, 7: np.uint16, 8: np.uint32,
       9: np.uint64, 10: np.float32, 11: np.float64, 13: np.int32}


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "libfls_oracle.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `make -C oracle`")
        L = C.CDLL(path)
        L.flso_open.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p)]
        L.flso_close.argtypes = [C.c_void_p]
        for f in ("flso_num_columns", "flso_num_rowgroups"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = C.c_uint32
        L.flso_num_rows.argtypes = [C.c_void_p]
        L.flso_num_rows.restype = C.c_uint64
        L.flso_column_type.argtypes = [C.c_void_p, C.c_uint32]
        L.flso_column_name.argtypes = [C.c_void_p, C.c_uint32]
        L.flso_column_name.restype = C.c_char_p
        L.flso_rowgroup_vectors.argtypes = [C.c_void_p, C.c_uint32]
        L.flso_rowgroup_vectors.restype = C.c_uint32
        L.flso_rowgroup_rows.argtypes = [C.c_void_p, C.c_uint32]
        L.flso_rowgroup_rows.restype = C.c_uint32
        L.flso_chunk_scheme.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.flso_chunk_bytes.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.flso_chunk_bytes.restype = C.c_uint64
        L.flso_chunk_segment.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.flso_range_padded_rows.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.flso_range_padded_rows.restype = C.c_uint64
        L.flso_range_payload.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        L.flso_range_payload.restype = C.c_uint64
        L.flso_decode_validity.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.flso_decode_fixed.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]
        L.flso_decode_str.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
        L.flso_alp_decode_value.argtypes = [C.c_int64, C.c_uint, C.c_uint]
        L.flso_alp_decode_value.restype = C.c_double
        L.flso_alpf_decode_value.argtypes = [C.c_int64, C.c_uint, C.c_uint]
        L.flso_alpf_decode_value.restype = C.c_float
        L.flso_fsst_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.flso_fsst_decode.restype = C.c_uint64
        L.flso_alprd_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                        C.c_uint32, C.c_void_p]
        L.flso_alprd_decode.restype = None
        _lib = L
    return _lib


class OracleTable:
    """CPU restatement of read_fls + materialize (see fls_oracle.c header)."""

    def __init__(self, image: np.ndarray):
        self.image = np.ascontiguousarray(image, dtype=np.uint8)
        self.h = C.c_void_p()
        rc = lib().flso_open(self.image.ctypes.data, self.image.size, C.byref(self.h))
        if rc:
            raise IOError(f"oracle: not an FLS image (code {rc})")
        L = lib()
        self.n_cols = L.flso_num_columns(self.h)
        self.n_rg = L.flso_num_rowgroups(self.h)
        self.n_rows = L.flso_num_rows(self.h)
        self.types = [L.flso_column_type(self.h, c) for c in range(self.n_cols)]
        self.names = [L.flso_column_name(self.h, c).decode() for c in range(self.n_cols)]

    def close(self):
        if self.h:
            lib().flso_close(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def scheme(self, rg, col):
        return lib().flso_chunk_scheme(self.h, rg, col)

    def chunk_bytes(self, rg, col):
        return lib().flso_chunk_bytes(self.h, rg, col)

    def segments(self):
        """[(file offset, size, is_vector_header_table)] of every non-empty segment of every chunk."""
        out = []
        off, size = C.c_uint64(), C.c_uint64()
        no_stream = (0, 1, 9)  # UNCOMPRESSED, CONSTANT, STR_CONSTANT: blobs only
        for g in range(self.n_rg):
            for c in range(self.n_cols):
                sch = self.scheme(g, c)
                hdr_segs = () if sch in no_stream else ((0, 2) if sch == 11 else (0,))
                for s in range(8):
                    if not lib().flso_chunk_segment(self.h, g, c, s, C.byref(off), C.byref(size)):
                        break
                    if size.value:
                        out.append((off.value, size.value, s in hdr_segs))
        return out

    def compressed_bytes(self, col, g0=0, g1=None):
        g1 = self.n_rg if g1 is None else g1
        return sum(self.chunk_bytes(g, col) for g in range(g0, g1))

    def padded_rows(self, g0=0, g1=None):
        g1 = self.n_rg if g1 is None else g1
        return lib().flso_range_padded_rows(self.h, g0, g1)

    def rows(self, g0=0, g1=None):
        g1 = self.n_rg if g1 is None else g1
        return sum(lib().flso_rowgroup_rows(self.h, g) for g in range(g0, g1))

    def validity(self, col, g0=0, g1=None, padded=False):
        """bool per row (True = holds a value), or None when no chunk of the range carries a validity bitmap."""
        L = lib()
        g1 = self.n_rg if g1 is None else g1
        n_pad = L.flso_range_padded_rows(self.h, g0, g1)
        bits = np.empty(n_pad // 8, dtype=np.uint8)
        if not L.flso_decode_validity(self.h, col, g0, g1, bits.ctypes.data):
            return None
        v = np.unpackbits(bits, bitorder="little").astype(bool)
        return v if padded else v[:self.rows(g0, g1)]

    def decode(self, col, g0=0, g1=None, threads=1, padded=False):
        """Fixed width -> np.ndarray; strings -> (offsets u32[n+1], payload u8)."""
        L = lib()
        g1 = self.n_rg if g1 is None else g1
        n_pad = L.flso_range_padded_rows(self.h, g0, g1)
        n = n_pad if padded else self.rows(g0, g1)
        t = self.types[col]
        if t in _DT:
            out = np.empty(n_pad, dtype=_DT[t])
            rc = L.flso_decode_fixed(self.h, col, g0, g1, out.ctypes.data, threads)
            if rc:
                raise RuntimeError(f"oracle decode failed (code {rc})")
            return out[:n]
        pay = np.empty(L.flso_range_payload(self.h, col, g0, g1) + 16, dtype=np.uint8)
        offs = np.empty(n_pad + 1, dtype=np.uint32)
        rc = L.flso_decode_str(self.h, col, g0, g1, offs.ctypes.data, pay.ctypes.data, threads)
        if rc:
            raise RuntimeError(f"oracle string decode failed (code {rc})")
        return offs[:n + 1], pay[:int(offs[n])]
