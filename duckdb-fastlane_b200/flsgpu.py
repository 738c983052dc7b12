"""ctypes binding of libflsgpu.so (include/flsgpu.h) — the reference-facing C ABI of the B200 read_fls path.

Everything here goes through the C ABI; there is no Python or CPU decode.  If libflsgpu.so is missing, or a
decode is attempted without a CUDA device, the call raises.
"""
import ctypes as C
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_lib = None

OK, E_IO, E_FORMAT, E_RANGE, E_CUDA, E_LIMIT, E_NOMEM = range(7)
OPEN_PINNED_IMAGE, OPEN_NO_PIN, OPEN_TIME_KERNELS, OPEN_NO_PLAN_CACHE, OPEN_LOAD_FILE, OPEN_GDS = 1, 2, 4, 8, 16, 32
OUT_DEVICE, OUT_HOST = 0, 1
CMP = {"=": 0, "==": 0, "!=": 1, "<>": 1, "<": 2, "<=": 3, ">": 4, ">=": 5}

# enum fls_type -> numpy dtype of the decoded column
DTYPES = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.uint8, 7: np.uint16, 8: np.uint32,
          9: np.uint64, 10: np.float32, 11: np.float64, 13: np.int32}
TYPE_NAMES = {1: "BOOLEAN", 2: "TINYINT", 3: "SMALLINT", 4: "INTEGER", 5: "BIGINT", 6: "UTINYINT", 7: "USMALLINT",
              8: "UINTEGER", 9: "UBIGINT", 10: "FLOAT", 11: "DOUBLE", 12: "VARCHAR", 13: "DATE", 14: "BLOB"}


class Opts(C.Structure):
    _fields_ = [("device", C.c_int32), ("flags", C.c_uint32), ("slice_rowgroups", C.c_uint32), ("reserved", C.c_uint32)]


class Column(C.Structure):
    _fields_ = [("type", C.c_int32), ("width", C.c_uint32), ("rows", C.c_uint64), ("padded_rows", C.c_uint64),
                ("data", C.c_void_p), ("offsets", C.c_void_p), ("payload", C.c_void_p), ("payload_bytes", C.c_uint64),
                ("compressed_bytes", C.c_uint64), ("decoded_bytes", C.c_uint64), ("offsets64", C.c_void_p), ("validity", C.c_void_p)]


class Timing(C.Structure):
    _fields_ = [("ms_total", C.c_float), ("ms_h2d", C.c_float), ("ms_kernels", C.c_float), ("ms_d2h", C.c_float),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("compressed_bytes", C.c_uint64),
                ("decoded_bytes", C.c_uint64), ("kernel_launches", C.c_uint32), ("slices", C.c_uint32),
                ("ms_host_plan", C.c_float), ("ms_host_total", C.c_float)]


class Predicate(C.Structure):
    _fields_ = [("col", C.c_uint32), ("op", C.c_uint32), ("i64", C.c_int64), ("f64", C.c_double),
                ("text", C.c_char_p), ("text_len", C.c_uint64)]


class KernelStat(C.Structure):
    _fields_ = [("name", C.c_char_p), ("launches", C.c_uint64), ("ms", C.c_double), ("bytes_read", C.c_uint64),
                ("bytes_written", C.c_uint64), ("vectors", C.c_uint64)]


def header_symbols():
    """Every function name include/flsgpu.h declares."""
    txt = open(os.path.join(_ROOT, "include", "flsgpu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(flsgpu_[a-z0-9_]+)\s*\(", txt)))


def lib_path():
    # FLSGPU_LIB: another build of the same CUDA library (kernel experiments under profiles/); never a CPU path
    return os.environ.get("FLSGPU_LIB") or os.path.join(_HERE, "libflsgpu.so")


def lib():
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise RuntimeError(p + " is missing: build it with `python __graft_entry__.py` (nvcc, sm_100a). "
                           "There is no CPU fallback for read_fls.")
    L = C.CDLL(p)
    for s in header_symbols():
        getattr(L, s)  # AttributeError if the library does not export a declared symbol
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    L.flsgpu_open.argtypes = [C.c_char_p, C.POINTER(Opts), C.POINTER(vp)]
    L.flsgpu_open_image.argtypes = [vp, u64, C.POINTER(Opts), C.POINTER(vp)]
    L.flsgpu_close.argtypes = [vp]
    L.flsgpu_close.restype = None
    L.flsgpu_trim_cache.argtypes = []
    L.flsgpu_trim_cache.restype = None
    L.flsgpu_last_error.argtypes = [vp]
    L.flsgpu_last_error.restype = C.c_char_p
    for f in ("flsgpu_num_columns", "flsgpu_num_rowgroups"):
        getattr(L, f).argtypes = [vp]
        getattr(L, f).restype = u32
    L.flsgpu_num_rows.argtypes = [vp]
    L.flsgpu_num_rows.restype = u64
    L.flsgpu_rowgroup_rows.argtypes = [vp, u32]
    L.flsgpu_rowgroup_rows.restype = u32
    L.flsgpu_schema.argtypes = [vp, C.POINTER(u32), C.POINTER(C.POINTER(C.c_char_p)), C.POINTER(C.POINTER(C.c_int32))]
    L.flsgpu_chunk_info.argtypes = [vp, u32, u32, C.POINTER(C.c_int32), C.POINTER(u64)]
    L.flsgpu_rowgroup_minmax.argtypes = [vp, u32, u32, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
    L.flsgpu_rowgroup_dict_contains.argtypes = [vp, u32, u32, C.c_char_p, C.c_uint64, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.flsgpu_make_resident.argtypes = [vp, u32, u32]
    L.flsgpu_decode.argtypes = [vp, u32, u32, C.POINTER(u32), u32, u32, C.POINTER(vp)]
    L.flsgpu_decode_filtered.argtypes = [vp, u32, u32, C.POINTER(u32), u32, C.POINTER(Predicate), u32, u32, C.POINTER(vp)]
    L.flsgpu_batch_selection.argtypes = [vp, C.POINTER(C.POINTER(u64)), C.POINTER(u64)]
    L.flsgpu_batch_num_columns.argtypes = [vp]
    L.flsgpu_batch_num_columns.restype = u32
    L.flsgpu_batch_column.argtypes = [vp, u32, C.POINTER(Column)]
    L.flsgpu_batch_timing.argtypes = [vp, C.POINTER(Timing)]
    L.flsgpu_batch_fetch.argtypes = [vp, u32, vp, vp]
    L.flsgpu_batch_fetch_validity.argtypes = [vp, u32, vp]
    L.flsgpu_batch_fetch_offsets64.argtypes = [vp, u32, vp]
    L.flsgpu_batch_copy_to_device.argtypes = [vp, u32, C.c_int32, vp, vp, C.POINTER(C.c_float)]
    L.flsgpu_batch_release.argtypes = [vp]
    L.flsgpu_batch_release.restype = None
    L.flsgpu_num_kernel_stats.argtypes = [vp]
    L.flsgpu_num_kernel_stats.restype = u32
    L.flsgpu_kernel_stats.argtypes = [vp, u32, C.POINTER(KernelStat)]
    L.flsgpu_reset_kernel_stats.argtypes = [vp]
    L.flsgpu_reset_kernel_stats.restype = None
    L.flsgpu_flush_l2.argtypes = [vp, u64]
    L.flsgpu_select_fixed_noalloc.argtypes = [C.c_int32, vp, C.c_int32, u64, u32, C.c_int64, C.c_double, vp, vp, vp, vp]
    _lib = L
    return L


class FlsError(IOError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


NOALLOC_MAX_ROWS = 4096 * 1024


def select_fixed_noalloc(device, values, fls_type, n_rows, op, const, out_values, out_row_ids, out_count, stream=None):
    """flsgpu_select_fixed_noalloc (include/flsgpu.h): every pointer is a raw device address the caller owns; the
    kernels are launched on `stream` and not waited for.  Raises FlsError on a refused call."""
    i64 = int(const) if not isinstance(const, float) else 0
    f64 = float(const)
    if i64 >= 1 << 63:
        i64 -= 1 << 64
    rc = lib().flsgpu_select_fixed_noalloc(device, values, fls_type, n_rows, CMP[op] if isinstance(op, str) else op, i64, f64,
                                           out_values, out_row_ids, out_count, stream)
    if rc:
        raise FlsError(rc, (lib().flsgpu_last_error(None) or b"").decode("utf-8", "replace"))


class Batch:
    """A decoded row-group range.  Column buffers belong to the reader's pool until release()."""

    def __init__(self, reader, handle, to_host):
        self.reader, self.h, self.to_host = reader, handle, to_host
        L = lib()
        self.n = L.flsgpu_batch_num_columns(handle)
        self.columns = []
        for i in range(self.n):
            c = Column()
            L.flsgpu_batch_column(handle, i, C.byref(c))
            self.columns.append(c)
        t = Timing()
        L.flsgpu_batch_timing(handle, C.byref(t))
        self.timing = t

    def numpy(self, i, padded=False, copy=True):
        """Host view (copy by default) of column i: ndarray, or (offsets u32[n+1], payload u8) for strings."""
        assert self.to_host, "batch was decoded with OUT_DEVICE"
        c = self.columns[i]
        n = c.padded_rows if padded else c.rows
        if c.width:
            dt = np.dtype(DTYPES[c.type])
            a = np.ctypeslib.as_array((C.c_uint8 * (n * c.width)).from_address(c.data)).view(dt) if n else np.empty(0, dt)
            return a.copy() if copy else a
        if c.offsets64:   # 4 GiB of payload or more in this batch: 64-bit offsets
            offs = np.ctypeslib.as_array((C.c_uint64 * (n + 1)).from_address(c.offsets64))
        else:
            offs = np.ctypeslib.as_array((C.c_uint32 * (n + 1)).from_address(c.offsets))
        end = min(int(offs[n]), int(c.payload_bytes))  # a damaged file may decode to offsets past the payload buffer
        pay = np.ctypeslib.as_array((C.c_uint8 * end).from_address(c.payload)) if end else np.empty(0, np.uint8)
        return (offs.copy(), pay.copy()) if copy else (offs, pay)

    def fetch(self, i, padded=False):
        """Copy column i of a device-output batch to new host arrays (flsgpu_batch_fetch)."""
        c = self.columns[i]
        n = c.padded_rows if padded else c.rows
        if c.width:
            a = np.empty(c.padded_rows, dtype=DTYPES[c.type])
            rc = lib().flsgpu_batch_fetch(self.h, i, a.ctypes.data, None)
            self.reader._check(rc)
            return a[:n]
        pay = np.empty(max(int(c.payload_bytes), 1), dtype=np.uint8)
        if c.offsets64:
            offs = np.empty(c.padded_rows + 1, dtype=np.uint64)
            self.reader._check(lib().flsgpu_batch_fetch(self.h, i, pay.ctypes.data, None))
            self.reader._check(lib().flsgpu_batch_fetch_offsets64(self.h, i, offs.ctypes.data))
            return offs[:n + 1], pay[:min(int(offs[n]), len(pay))]
        offs = np.empty(c.padded_rows + 1, dtype=np.uint32)
        self.reader._check(lib().flsgpu_batch_fetch(self.h, i, pay.ctypes.data, offs.ctypes.data))
        return offs[:n + 1], pay[:min(int(offs[n]), len(pay))]

    def validity(self, i, padded=False):
        """bool per row of column i (True = holds a value), or None when the column has no validity bitmap."""
        c = self.columns[i]
        if not c.validity:
            return None
        nb = (c.padded_rows + 63) // 64 * 8
        if self.to_host:
            bits = np.ctypeslib.as_array((C.c_uint8 * nb).from_address(c.validity)).copy() if nb else np.empty(0, np.uint8)
        else:
            bits = np.empty(nb, dtype=np.uint8)
            self.reader._check(lib().flsgpu_batch_fetch_validity(self.h, i, bits.ctypes.data))
        v = np.unpackbits(bits, bitorder="little").astype(bool)
        return v[:c.padded_rows if padded else c.rows]

    def selection(self):
        """Row ids of a filtered host batch (None for an unfiltered one)."""
        ids, n = C.POINTER(C.c_uint64)(), C.c_uint64()
        lib().flsgpu_batch_selection(self.h, C.byref(ids), C.byref(n))
        if not ids:
            return None
        assert self.to_host
        return np.ctypeslib.as_array(ids, shape=(n.value,)).copy() if n.value else np.empty(0, np.uint64)

    def copy_to_device(self, i, dst_device, dst_values, dst_offsets=None):
        """flsgpu_batch_copy_to_device: column i of a device-output batch into caller buffers (raw device pointers) on
        `dst_device`.  Returns the copies' device time in ms."""
        ms = C.c_float()
        self.reader._check(lib().flsgpu_batch_copy_to_device(self.h, i, dst_device, dst_values, dst_offsets, C.byref(ms)))
        return float(ms.value)

    def release(self):
        if self.h:
            lib().flsgpu_batch_release(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.release()


class Reader:
    """read_fls over the C ABI: open -> schema -> decode(row-group range, columns)."""

    def __init__(self, handle, keep=None):
        self.h = handle
        self._keep = keep
        L = lib()
        self.n_cols = L.flsgpu_num_columns(handle)
        self.n_rowgroups = L.flsgpu_num_rowgroups(handle)
        self.n_rows = L.flsgpu_num_rows(handle)
        n = C.c_uint32()
        names = C.POINTER(C.c_char_p)()
        types = C.POINTER(C.c_int32)()
        L.flsgpu_schema(handle, C.byref(n), C.byref(names), C.byref(types))
        self.names = [names[i].decode("utf-8", "replace") for i in range(n.value)]
        self.types = [types[i] for i in range(n.value)]

    @staticmethod
    def _opts(device, flags, slice_rowgroups):
        return Opts(device, flags, slice_rowgroups, 0)

    @classmethod
    def open(cls, path, device=0, flags=0, slice_rowgroups=0):
        h = C.c_void_p()
        o = cls._opts(device, flags, slice_rowgroups)
        rc = lib().flsgpu_open(os.fsencode(path), C.byref(o), C.byref(h))
        if rc:
            raise FlsError(rc, lib().flsgpu_last_error(None).decode())
        return cls(h)

    @classmethod
    def from_image(cls, image, device=0, flags=0, slice_rowgroups=0):
        """image: C-contiguous uint8 ndarray (kept alive by the reader) or an (address, size) pair."""
        h = C.c_void_p()
        o = cls._opts(device, flags, slice_rowgroups)
        if isinstance(image, tuple):
            addr, size, keep = image[0], image[1], image
        else:
            image = np.ascontiguousarray(image, dtype=np.uint8)
            addr, size, keep = image.ctypes.data, image.size, image
        rc = lib().flsgpu_open_image(addr, size, C.byref(o), C.byref(h))
        if rc:
            raise FlsError(rc, lib().flsgpu_last_error(None).decode())
        return cls(h, keep)

    def _check(self, rc):
        if rc:
            raise FlsError(rc, lib().flsgpu_last_error(self.h).decode())

    def rowgroup_rows(self, g):
        return lib().flsgpu_rowgroup_rows(self.h, g)

    def chunk_info(self, g, c):
        s, b = C.c_int32(), C.c_uint64()
        self._check(lib().flsgpu_chunk_info(self.h, g, c, C.byref(s), C.byref(b)))
        return s.value, b.value

    def rowgroup_minmax(self, g, c):
        """(min, max) bounds of an integer / DATE column in row group g read off the compressed form, or None."""
        mn, mx, known = C.c_int64(), C.c_int64(), C.c_int32()
        self._check(lib().flsgpu_rowgroup_minmax(self.h, g, c, C.byref(mn), C.byref(mx), C.byref(known)))
        return (mn.value, mx.value) if known.value else None

    def rowgroup_dict_contains(self, g, c, text):
        """True / False: the dictionary of string column c in row group g holds `text` (bytes or str); None when the
        chunk's scheme cannot say."""
        raw = text.encode() if isinstance(text, str) else bytes(text)
        found, known = C.c_int32(), C.c_int32()
        self._check(lib().flsgpu_rowgroup_dict_contains(self.h, g, c, raw, len(raw), C.byref(found), C.byref(known)))
        return bool(found.value) if known.value else None

    def make_resident(self, g0=0, g1=None):
        self._check(lib().flsgpu_make_resident(self.h, g0, self.n_rowgroups if g1 is None else g1))

    def decode_batch(self, g0=0, g1=None, cols=None, to_host=False) -> Batch:
        g1 = self.n_rowgroups if g1 is None else g1
        b = C.c_void_p()
        if cols is None:
            arr, n = None, 0
        else:
            arr, n = (C.c_uint32 * len(cols))(*cols), len(cols)
        self._check(lib().flsgpu_decode(self.h, g0, g1, arr, n, OUT_HOST if to_host else OUT_DEVICE, C.byref(b)))
        return Batch(self, b, to_host)

    def decode_filtered_batch(self, preds, g0=0, g1=None, cols=None, to_host=True) -> Batch:
        """flsgpu_decode_filtered: preds = [(column id, "<" | "<=" | "=" | ..., constant)], a conjunction; the batch
        holds only the rows that pass.  The constant is an int, a float (FLOAT / DOUBLE columns) or bytes (strings)."""
        g1 = self.n_rowgroups if g1 is None else g1
        ps = (Predicate * max(len(preds), 1))()
        for i, (col, op, val) in enumerate(preds):
            ps[i].col, ps[i].op = col, CMP[op] if isinstance(op, str) else op
            if isinstance(val, (bytes, bytearray)):
                ps[i].text, ps[i].text_len = bytes(val), len(val)
            elif self.types[col] in (10, 11):
                ps[i].f64 = float(val)
            else:
                ps[i].i64 = int(val) if int(val) < (1 << 63) else int(val) - (1 << 64)
        b = C.c_void_p()
        if cols is None:
            arr, n = None, 0
        else:
            arr, n = (C.c_uint32 * len(cols))(*cols), len(cols)
        self._check(lib().flsgpu_decode_filtered(self.h, g0, g1, arr, n, ps, len(preds), OUT_HOST if to_host else OUT_DEVICE,
                                                 C.byref(b)))
        return Batch(self, b, to_host)

    def decode_filtered(self, preds, g0=0, g1=None, cols=None):
        """Host copies of the surviving rows: (row ids, [ndarray | (offsets, payload)])."""
        with self.decode_filtered_batch(preds, g0, g1, cols, to_host=True) as b:
            return b.selection(), [b.numpy(i) for i in range(b.n)]

    def decode(self, g0=0, g1=None, cols=None, to_host=True, padded=False):
        """Decode and return host copies: list of ndarray / (offsets, payload)."""
        with self.decode_batch(g0, g1, cols, to_host=True) as b:
            return [b.numpy(i, padded=padded) for i in range(b.n)]

    def kernel_stats(self):
        L = lib()
        out = []
        for i in range(L.flsgpu_num_kernel_stats(self.h)):
            s = KernelStat()
            L.flsgpu_kernel_stats(self.h, i, C.byref(s))
            if s.launches:
                out.append(dict(name=s.name.decode(), launches=s.launches, ms=s.ms, bytes_read=s.bytes_read,
                                bytes_written=s.bytes_written, vectors=s.vectors))
        return out

    def reset_kernel_stats(self):
        lib().flsgpu_reset_kernel_stats(self.h)

    def flush_l2(self, nbytes=256 << 20):
        self._check(lib().flsgpu_flush_l2(self.h, nbytes))

    def close(self):
        if self.h:
            lib().flsgpu_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
