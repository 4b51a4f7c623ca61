"""GPU (-m gpu), collected LAST: b200_w4a16_gemm_static_scratch — the M <= 32 GEMM whose scratch is a file-scope
`static __device__` array inside the library, for callers that may not allocate (include/b200_w4a16.h).  The entry
point has no counterpart in the reference, so the check is against the
existing entry point b200_w4a16_gemm_ex on the same inputs, through the C ABI, on two calls in a row (the array keeps
the first call's contents), within the parity tolerance of tests/test_gpu_parity.py (max rel err <= 1e-2)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL_MAX_REL = 1e-2   # the bar of tests/test_gpu_parity.py (north_star)
DEV = "cuda"


def _b():
    import cpmcu_b200
    return cpmcu_b200


def problem(m, k, n, dtype, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
    sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).to(dtype)
    a = torch.randn((m, k), device=DEV, generator=g).to(dtype)
    return a, qw, sc


def existing_entry_point(a, qw, sc):
    b = _b()
    m, k = a.shape
    n = qw.shape[1] // 2
    ws = torch.zeros(b.workspace_ints(n), dtype=torch.int32, device=DEV)
    c_tmp = torch.empty(max(b.reduce_bytes_fast(m, n, k) // 4, b.reduce_max_m(m) * n, 1), dtype=torch.float32, device=DEV)
    out = b.gemm(a, qw, sc, ws, c_tmp)
    torch.cuda.synchronize()
    return out


def max_rel(x, ref):
    return float((x.float() - ref.float()).abs().max() / ref.float().abs().mean())


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m,k,n", [(1, 4096, 4608), (16, 4096, 32768), (32, 16384, 4096), (5, 1024, 512)])
def test_static_scratch_matches_existing_entry_point_twice(dtype, m, k, n):
    b = _b()
    a, qw, sc = problem(m, k, n, dtype, seed=m * 7 + n)
    ref = existing_entry_point(a, qw, sc)
    out1 = torch.full((m, n), float("nan"), dtype=dtype, device=DEV)
    out2 = torch.full((m, n), float("nan"), dtype=dtype, device=DEV)
    b.gemm_static_scratch(a, qw, sc, out1)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    b.gemm_static_scratch(a, qw, sc, out2)          # second call: the array holds the first call's partial tiles
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    for name, out in (("first", out1), ("second", out2)):
        assert torch.isfinite(out.float()).all(), f"{name} call left output unwritten"
        err = max_rel(out, ref)
        assert err <= TOL_MAX_REL, f"{name} call: max rel err {err:.3e} vs b200_w4a16_gemm_ex"
    assert torch.equal(out1, out2), "two calls in a row on the same inputs differ: stale scratch or counters"
    assert free_after == free_before, "the call changed the device's free memory: something allocated"


def test_static_scratch_different_shapes_back_to_back():
    """a wide problem right after a narrow one and back: nothing of the previous call's scratch leaks into the next"""
    b = _b()
    cases = [(4, 2048, 1024), (32, 4096, 32768), (4, 2048, 1024), (8, 4096, 4096)]
    for i, (m, k, n) in enumerate(cases):
        a, qw, sc = problem(m, k, n, torch.float16, seed=100 + i)
        out = torch.full((m, n), float("nan"), dtype=torch.float16, device=DEV)
        b.gemm_static_scratch(a, qw, sc, out)
        torch.cuda.synchronize()
        err = max_rel(out, existing_entry_point(a, qw, sc))
        assert err <= TOL_MAX_REL, (m, k, n, err)


def test_static_scratch_rejects_sizes_beyond_the_maximum():
    b = _b()
    a, qw, sc = problem(33, 1024, 512, torch.float16, seed=1)
    out = torch.empty((33, 512), dtype=torch.float16, device=DEV)
    with pytest.raises(ValueError, match="exceeds the maximum of 32"):
        b.gemm_static_scratch(a, qw, sc, out)
    n_big = 65536 + 64   # only the shape is inspected before the rejection: tiny dummy buffers suffice
    a1 = torch.zeros((1, 64), dtype=torch.float16, device=DEV)
    with pytest.raises(ValueError, match="exceeds the maximum of 65536"):
        b.capi.lib().b200_w4a16_gemm_static_scratch  # symbol exists
        _call_raw(b, a1, n_big, 64)


def _call_raw(b, a1, n, k):
    import ctypes
    rc = b.capi.lib().b200_w4a16_gemm_static_scratch(ctypes.c_void_p(a1.data_ptr()), ctypes.c_void_p(a1.data_ptr()),
                                                     ctypes.c_void_p(a1.data_ptr()), ctypes.c_void_p(a1.data_ptr()),
                                                     1, n, k, 128, 0, ctypes.c_void_p(0))
    if rc:
        b.capi._raise(rc)
