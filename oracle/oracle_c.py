"""TEST INFRASTRUCTURE: ctypes handle on oracle/_ref/liboracle_c.so (oracle/marlin_oracle.c, the
plain-C OpenMP restatement of the reference algorithm).  Used by tests/ and by bench.py's
cpu_baseline / --impl reference legs only."""
import ctypes
import os
from ctypes import c_int, c_int64, c_void_p

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "_ref", "liboracle_c.so")
_lib = None


def available() -> bool:
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        h = ctypes.CDLL(PATH)
        h.marlin_oracle_dequant.restype = c_int
        h.marlin_oracle_dequant.argtypes = [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int, c_int]
        h.marlin_oracle_gemm.restype = c_int
        h.marlin_oracle_gemm.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64,
                                         c_int64, c_int, c_int]
        h.marlin_oracle_threads.restype = c_int
        _lib = h
    return _lib


def threads() -> int:
    return lib().marlin_oracle_threads()


def _p(x):
    return None if x is None else c_void_p(x.ctypes.data)


def dequant(qweight, scales_marlin, k, n, group_size=128, dtype=0):
    """-> uint16 bit patterns [K][N] of W in T (fp16: view as float16; bf16: bf16 bits)."""
    q = np.ascontiguousarray(qweight, dtype=np.int32)
    s = np.ascontiguousarray(scales_marlin).view(np.uint16)
    out = np.empty((k, n), dtype=np.uint16)
    rc = lib().marlin_oracle_dequant(_p(q), _p(s), _p(out), n, k, group_size, dtype)
    if rc:
        raise ValueError(f"marlin_oracle_dequant: bad arguments (rc={rc})")
    return out


def gemm(a, qweight, scales_marlin, k, n, group_size=128, dtype=0, perm=None):
    """-> uint16 bit patterns [M][N] of C in T."""
    a16 = np.ascontiguousarray(a).view(np.uint16)
    q = np.ascontiguousarray(qweight, dtype=np.int32)
    s = np.ascontiguousarray(scales_marlin).view(np.uint16)
    p = None if perm is None else np.ascontiguousarray(perm, dtype=np.int32)
    m = a16.shape[0]
    out = np.empty((m, n), dtype=np.uint16)
    rc = lib().marlin_oracle_gemm(_p(a16), _p(q), _p(s), _p(p), _p(out), m, n, k, group_size, dtype)
    if rc:
        raise ValueError(f"marlin_oracle_gemm failed (rc={rc})")
    return out
