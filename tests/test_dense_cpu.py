"""CPU: the oracle's restatement of the un-quantised Linear / LMHead (src/model/linear.cuh:8-37,76-105) against
torch's CPU fp32 linear with the same rounding points.  The reference computes this op with cuBLAS (a
third-party library, CUDA 12.x, absent from /root/reference): parity is anchored on the call site's contract
(OP_T/OP_N, CUBLAS_COMPUTE_32F, beta = 1 when inplace, bias added afterwards in T)."""
import numpy as np
import pytest
import torch

from oracle import marlin_oracle as mo


@pytest.mark.parametrize("dtype,tdt", [(mo.FP16, torch.float16), (mo.BF16, torch.bfloat16)])
@pytest.mark.parametrize("inplace,bias,scale", [(False, False, 1.0), (True, True, 1.0), (False, True, 0.3), (True, False, 0.0625)])
def test_oracle_dense_vs_torch(dtype, tdt, inplace, bias, scale):
    g = torch.Generator().manual_seed(1)
    m, n, k = 5, 72, 256
    a = torch.randn((m, k), generator=g).to(tdt)
    w = (torch.randn((n, k), generator=g) * 0.1).to(tdt)
    b = torch.randn((n,), generator=g).to(tdt)
    old = torch.randn((m, n), generator=g).to(tdt)

    def st(t):
        return t.numpy() if dtype == mo.FP16 else t.view(torch.int16).numpy().view(np.uint16)

    got, exact = mo.dense_linear(st(a), st(w), dtype, bias=st(b) if bias else None, tgt=st(old), inplace=inplace,
                                 head_scale=scale)
    a_s = a if scale == 1.0 else (a * torch.tensor(scale).to(tdt))          # T * T -> T, one rounding
    acc = a_s.double() @ w.double().T
    if inplace:
        acc = acc + old.double()
    out = acc.to(tdt)
    if bias:
        out = (out.float() + b.float()).to(tdt)
    assert np.array_equal(st(out).view(np.uint16), np.asarray(got).view(np.uint16))
    assert np.allclose(exact, (acc + (b.double() if bias else 0)).numpy(), rtol=1e-12, atol=1e-12)
