"""TEST INFRASTRUCTURE: ctypes handle on oracle/_ref/libmarlin_ref.so — the UNMODIFIED reference
Marlin (src/qgemm/gptq_marlin/*) compiled in place for sm_100 by oracle/Makefile, called through
oracle/ref_shim.cu.  Used by the GPU parity tests and bench.py --impl reference only."""
import ctypes
import os
from ctypes import c_char_p, c_int, c_int64, c_void_p

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "_ref", "libmarlin_ref.so")
_lib = None


def available() -> bool:
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        h = ctypes.CDLL(PATH)
        h.marlin_ref_gemm.restype = c_int
        h.marlin_ref_gemm.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int, c_int,
                                      c_void_p, c_void_p, c_void_p, c_char_p, c_int]
        h.marlin_ref_reduce_max_m.restype = c_int
        h.marlin_ref_reduce_max_m.argtypes = [c_int, c_int]
        _lib = h
    return _lib


def reduce_max_m(m: int) -> int:
    return lib().marlin_ref_reduce_max_m(m, 16)


def gemm(a, qweight, scales, workspace, c_tmp, out=None, group_size=128, perm=None, a_tmp=None,
         use_fp32_reduce=True, stream=None):
    """reference gptq_marlin_gemm<T> with the arguments W4A16GPTQMarlinLinear::prefill passes."""
    import torch
    m, k = a.shape
    n = qweight.shape[1] // 2
    if out is None:
        out = torch.empty((m, n), dtype=a.dtype, device=a.device)
    dt = 0 if a.dtype == torch.float16 else 1
    err = ctypes.create_string_buffer(1024)
    s = torch.cuda.current_stream() if stream is None else stream
    p = lambda t: None if t is None else c_void_p(t.data_ptr())
    rc = lib().marlin_ref_gemm(p(a), p(qweight), p(scales), None, p(perm), p(workspace), p(out), m, n, k,
                               group_size, dt, 1, 0 if perm is None else 1, 1 if use_fp32_reduce else 0,
                               p(a_tmp), p(c_tmp), c_void_p(s.cuda_stream), err, 1024)
    if rc == 1:
        raise ValueError(err.value.decode())
    if rc:
        raise RuntimeError(err.value.decode())
    return out
