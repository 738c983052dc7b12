"""ctypes binding of the host emulation of the CUDA decoders (tests/emul/fls_emul.cpp).  TEST ONLY."""
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
        L = C.CDLL(os.path.join(_HERE, "libfls_emul.so"))
        L.emul_error.restype = C.c_char_p
        L.emul_decode.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                  C.c_void_p]
        L.emul_need.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        _lib = L
    return _lib


def force_any(on):
    """Make every emulated warp vote succeed: all lanes take the most general code variant."""
    lib().emul_force_any(int(on))


def decode(image, col, ftype, g0, g1):
    """Returns padded output: fixed -> ndarray; strings -> (offsets[n_pad+1], payload)."""
    L = lib()
    need = np.zeros(4, dtype=np.uint64)
    if L.emul_need(image.ctypes.data, image.size, col, g0, g1, need.ctypes.data):
        raise RuntimeError(L.emul_error().decode())
    if ftype in _DT:
        out = np.zeros(int(need[0]) // np.dtype(_DT[ftype]).itemsize, dtype=_DT[ftype])
        rc = L.emul_decode(image.ctypes.data, image.size, col, g0, g1, out.ctypes.data, None, None)
        if rc:
            raise RuntimeError(L.emul_error().decode())
        return out
    offs = np.zeros(int(need[1]) // 4, dtype=np.uint32)
    pay = np.zeros(int(need[2]), dtype=np.uint8)
    rc = L.emul_decode(image.ctypes.data, image.size, col, g0, g1, None, offs.ctypes.data, pay.ctypes.data)
    if rc:
        raise RuntimeError(L.emul_error().decode())
    return offs, pay
