"""ctypes binding of libb200_w4a16.so (C ABI: include/b200_w4a16.h).

PyTorch is used for device memory and streams only; every compute call goes through the C ABI
with raw device pointers, the way the reference's pybind module passes them
(reference: src/entry.cu:536-570).  There is NO fallback: if the library is missing or a call
fails, an exception is raised (ValueError / RuntimeError, matching the reference's
VALUE_CHECK / ERROR_CHECK -> pybind mapping, src/utils.cuh:54-66).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int32, c_int64, c_void_p, POINTER

FP16, BF16 = 0, 1  # dtype codes of the reference: cpmcu/llm_w4a16_gptq_marlin.py:12-15

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_NAME = "libb200_w4a16.so"


class B200W4A16Error(RuntimeError):
    pass


class FusedAR(ctypes.Structure):
    """b200_w4a16_ar: buffers of the all-reduce fused into a row-parallel GEMM launch"""
    _fields_ = [
        ("recv", c_void_p),
        ("flag", c_void_p),
        ("seen", c_void_p),
        ("slot_elems", c_int64),
        ("rank", c_int),
        ("world", c_int),
    ]


class Opts(ctypes.Structure):
    _fields_ = [
        ("bias", c_void_p),
        ("residual", c_void_p),
        ("c_tmp_bytes", c_int64),
        ("force_config", c_int),
        ("pdl", c_int),
        ("out_scale", c_float),
        ("gated_silu", c_int),
        ("qkv_q_dim", c_int),
        ("qkv_kv_dim", c_int),
        ("reserved", c_int * 1),
    ]


def library_path() -> str:
    return os.environ.get("B200_W4A16_LIB", os.path.join(_HERE, _LIB_NAME))


_lib = None

# name -> (restype, argtypes): every symbol include/b200_w4a16.h declares
SYMBOLS = {
    "b200_w4a16_gemm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                c_void_p]),
    "b200_w4a16_gemm_ex": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                   c_void_p, POINTER(Opts)]),
    "b200_w4a16_gemm_ar": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int,
                                   c_void_p, c_void_p, POINTER(Opts), POINTER(FusedAR)]),
    "b200_w4a16_linear_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_int, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p,
                                          c_void_p]),
    "b200_w4a16_linear_forward_ex": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                             c_int, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p,
                                             c_int64, c_void_p]),
    "b200_w4a16_dense_linear_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64,
                                                c_int64, c_float, c_int, c_int, c_void_p]),
    "b200_w4a16_dequant": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int, c_int, c_void_p]),
    "b200_w4a16_repack_gptq": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p]),
    "b200_w4a16_reduce_max_m": (c_int, [c_int64]),
    "b200_w4a16_reduce_bytes": (c_int64, [c_int64, c_int64, c_int64]),
    "b200_w4a16_reduce_bytes_fast": (c_int64, [c_int64, c_int64, c_int64]),
    "b200_w4a16_set_scratch_capacity": (c_int64, [c_int64]),
    "b200_w4a16_gemm_static_scratch": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int,
                                               c_int, c_void_p]),
    "b200_w4a16_workspace_ints": (c_int64, [c_int64]),
    "b200_w4a16_launch_count": (c_int64, []),
    "b200_w4a16_set_pdl": (c_int, [c_int]),
    "b200_w4a16_pick_config": (c_int, [c_int64, c_int64, c_int64, c_int, c_int]),
    "b200_w4a16_config_name": (c_char_p, [c_int]),
    "b200_w4a16_max_clusters": (c_int, [c_int, c_int64, c_int, c_int]),
    "b200_w4a16_last_error": (c_char_p, []),
    "b200_w4a16_version": (c_char_p, []),
    "b200_w4a16_debug_timeline": (None, [c_void_p, c_int64]),
    "b200_w4a16_allreduce_oneshot": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                                             c_void_p]),
    "b200_w4a16_allreduce": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                                     c_int, c_void_p]),
}


def lib() -> ctypes.CDLL:
    """Load the CUDA library (once).  Raises if it has not been built: there is no CPU path."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise B200W4A16Error(
                f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). cpmcu_b200 has no CPU or PyTorch fallback.")
        handle = ctypes.CDLL(path)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(handle, name)  # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def _raise(code: int):
    msg = lib().b200_w4a16_last_error().decode()
    if code == 1:
        raise ValueError(msg)
    raise RuntimeError(f"b200_w4a16 error {code}: {msg}")


def _ptr(t):
    return None if t is None else c_void_p(t.data_ptr())


def _stream(stream=None):
    import torch
    s = torch.cuda.current_stream() if stream is None else stream
    return c_void_p(s.cuda_stream)


def dtype_code(torch_dtype) -> int:
    import torch
    if torch_dtype == torch.float16:
        return FP16
    if torch_dtype == torch.bfloat16:
        return BF16
    raise ValueError(f"W4A16 supports float16 / bfloat16 activations, got {torch_dtype}")


def reduce_max_m(m: int) -> int:
    return lib().b200_w4a16_reduce_max_m(m)


def reduce_bytes(m: int, n: int, k: int) -> int:
    return lib().b200_w4a16_reduce_bytes(m, n, k)


def reduce_bytes_fast(m: int, n: int, k: int) -> int:
    """bytes of c_tmp full-width stream-K needs (reduce_bytes: what is touched inside the reference's contract)"""
    return lib().b200_w4a16_reduce_bytes_fast(m, n, k)


def set_scratch_capacity(nbytes: int) -> int:
    """process-wide promise about every c_tmp passed without a size; returns the previous value"""
    return lib().b200_w4a16_set_scratch_capacity(nbytes)


def workspace_ints(n: int) -> int:
    return lib().b200_w4a16_workspace_ints(n)


def launch_count() -> int:
    return lib().b200_w4a16_launch_count()


def set_pdl(enable: bool) -> bool:
    """process-wide default for programmatic dependent launch; returns the previous setting"""
    return bool(lib().b200_w4a16_set_pdl(1 if enable else 0))


def pick_config(m, n, k, group_size=128, dtype=FP16) -> int:
    return lib().b200_w4a16_pick_config(m, n, k, group_size, dtype)


def config_name(cfg: int) -> str:
    return lib().b200_w4a16_config_name(cfg).decode()


def gemm(a, qweight, scales, workspace, c_tmp, out=None, *, group_size=128, perm=None, a_tmp=None,
         bias=None, residual=None, out_scale=0.0, gated_silu=False, qkv_split=None, force_config=0,
         pdl=False, stream=None):
    """C[M,N] = A[M,K] . dequant(qweight, scales)  (+ residual, + bias, * T(out_scale)) on the current stream.

    a: [M,K] fp16/bf16 cuda, qweight: int32 [K/16, 2N] cuda (Marlin), scales: [K/g, N] (Marlin-permuted),
    workspace: int32 [(N/64)*16] zeroed, c_tmp: float32 scratch (its size is passed along: the library uses
    full-width stream-K when it holds reduce_bytes_fast(M,N,K) bytes and stays inside whatever it is given
    otherwise; may be None on the tcgen05 path, M > 32)."""
    import torch
    assert a.is_cuda and a.is_contiguous() and qweight.is_contiguous() and scales.is_contiguous()
    m, k = a.shape
    n = qweight.shape[1] // 2
    if out is None:
        out = torch.empty((m, n // 2 if gated_silu else n), dtype=a.dtype, device=a.device)
    o = Opts()
    o.bias = None if bias is None else bias.data_ptr()
    o.residual = None if residual is None else residual.data_ptr()
    o.c_tmp_bytes = 0 if c_tmp is None else c_tmp.numel() * c_tmp.element_size()
    o.out_scale = float(out_scale)
    o.gated_silu = 1 if gated_silu else 0
    if qkv_split is not None:   # (q_dim, kv_dim): c = [q: M x q_dim][k: M x kv_dim][v: M x kv_dim]
        o.qkv_q_dim, o.qkv_kv_dim = int(qkv_split[0]), int(qkv_split[1])
    o.force_config = force_config
    o.pdl = 1 if pdl else 0
    rc = lib().b200_w4a16_gemm_ex(_ptr(a), _ptr(qweight), _ptr(scales), None, _ptr(perm), _ptr(workspace),
                                  _ptr(out), m, n, k, group_size, dtype_code(a.dtype), 1,
                                  0 if perm is None else 1, _ptr(a_tmp), _ptr(c_tmp), _stream(stream),
                                  ctypes.byref(o))
    if rc:
        _raise(rc)
    return out


def gemm_static_scratch(a, qweight, scales, out, *, group_size=128, stream=None):
    """The M <= 32 GEMM through b200_w4a16_gemm_static_scratch: no workspace / c_tmp arguments, the library's file-scope
    device array serves as scratch and the call allocates nothing (so `out` is required).  No counterpart in the
    reference.  Calls on one device share the array: keep them on one stream."""
    m, k = a.shape
    n = qweight.shape[1] // 2
    rc = lib().b200_w4a16_gemm_static_scratch(_ptr(a), _ptr(qweight), _ptr(scales), _ptr(out), m, n, k, group_size,
                                              dtype_code(a.dtype), _stream(stream))
    if rc:
        _raise(rc)
    return out


def dense_linear(a, weight, out=None, *, bias=None, inplace=False, input_scale=1.0, pdl=False, stream=None):
    """out[M,N] = (a * T(input_scale)) @ weight.T (+ out when inplace) (+ bias): the un-quantised lm_head /
    draft-head linear at M <= 64, one HBM-streaming launch (reference: src/model/linear.cuh:8-37,76-105).

    a: [M,K] fp16/bf16 cuda, weight: [N,K] same dtype (nn.Linear layout)."""
    import torch
    assert a.is_cuda and a.is_contiguous() and weight.is_contiguous() and a.dtype == weight.dtype
    m, k = a.shape
    n = weight.shape[0]
    if weight.shape[1] != k:
        raise ValueError(f"dense_linear: weight is {tuple(weight.shape)}, input has K = {k}")
    if out is None:
        if inplace:
            raise ValueError("dense_linear: inplace needs `out`")
        out = torch.empty((m, n), dtype=a.dtype, device=a.device)
    rc = lib().b200_w4a16_dense_linear_forward(_ptr(a), _ptr(weight), _ptr(bias), _ptr(out), 1 if inplace else 0,
                                               m, n, k, float(input_scale), dtype_code(a.dtype), 1 if pdl else 0,
                                               _stream(stream))
    if rc:
        _raise(rc)
    return out


def dequant(qweight, scales, k, n, *, group_size=128, dtype=None, stream=None):
    """W[K,N] = RN(T(q-8) * T(scale)) — the dequantisation the GEMM kernels apply in registers."""
    import torch
    dt = scales.dtype if dtype is None else dtype
    out = torch.empty((k, n), dtype=dt, device=qweight.device)
    rc = lib().b200_w4a16_dequant(_ptr(qweight), _ptr(scales), _ptr(out), n, k, group_size, dtype_code(dt),
                                  _stream(stream))
    if rc:
        _raise(rc)
    return out


def repack_gptq(qweight_gptq, stream=None):
    """AutoGPTQ int32[K/8,N] -> Marlin int32[K/16,2N] on the device."""
    import torch
    k8, n = qweight_gptq.shape
    k = k8 * 8
    out = torch.empty((k // 16, 2 * n), dtype=torch.int32, device=qweight_gptq.device)
    rc = lib().b200_w4a16_repack_gptq(_ptr(qweight_gptq), _ptr(out), n, k, _stream(stream))
    if rc:
        _raise(rc)
    return out
