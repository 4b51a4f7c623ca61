"""CPU: the plain-C oracle (oracle/marlin_oracle.c) against the numpy oracle and, through it, the
golden vectors of the reference's packer.  Two independent restatements must agree bit-for-bit on
the dequantisation and within fp32-accumulation noise on the GEMM."""
import subprocess
import os

import numpy as np
import pytest

from oracle import marlin_oracle as mo
from oracle import oracle_c

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def _built():
    if not oracle_c.available():
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle_c"])
    assert oracle_c.available()


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
@pytest.mark.parametrize("group", [128, -1])
def test_c_dequant_matches_numpy_bit_exact(dtype, group):
    k, n = 384, 192
    prob = mo.make_problem(1, k, n, group, dtype, seed=5)
    s = mo.to_f32(prob["scales"], dtype).copy()
    s[0, :8] = [6e-8, 1e-7, 3e-5, 6.1e-5, 1000.0, 8000.0, -0.01, 0.0]
    sm = mo.permute_scales(mo.to_storage(s, dtype), k, n, group)
    want = mo.dequant(prob["qweight"], sm, k, n, group, dtype)
    got = oracle_c.dequant(prob["qweight"], sm, k, n, group, dtype)
    assert np.array_equal(got, want.view(np.uint16))


def test_c_half_conversion_exhaustive():
    """all 63488 finite fp16 scales x 16 nibbles through the C conversion routines == numpy"""
    bits = np.arange(0x10000, dtype=np.uint32).astype(np.uint16)
    sbits = bits[np.isfinite(bits.view(np.float16))]
    n = (len(sbits) + 63) // 64 * 64
    s = np.zeros(n, dtype=np.uint16)
    s[:len(sbits)] = sbits
    k = 128
    wq = np.tile((np.arange(k, dtype=np.uint8) % 16)[:, None], (1, n))
    qw = mo.marlin_pack(wq)
    sm = mo.permute_scales(s.view(np.float16).reshape(1, n), k, n, 128)
    want = mo.dequant(qw, sm, k, n, 128, mo.FP16)
    got = oracle_c.dequant(qw, sm, k, n, 128, mo.FP16)
    assert np.array_equal(got, want.view(np.uint16))


@pytest.mark.parametrize("m,k,n", [(1, 128, 64), (3, 256, 192), (16, 1024, 320), (33, 512, 128)])
@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
@pytest.mark.parametrize("group", [128, -1])
def test_c_gemm_matches_numpy(m, k, n, dtype, group):
    prob = mo.make_problem(m, k, n, group, dtype, seed=m + n)
    _, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, group, dtype)
    got = oracle_c.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, group, dtype)
    got = got.view(np.float16) if dtype == mo.FP16 else got
    st = mo.error_stats(got, exact, dtype)
    assert st["max_rel"] <= 1e-2, st


def test_c_gemm_act_order():
    m, k, n = 4, 256, 128
    prob = mo.make_problem(m, k, n, 128, mo.FP16, seed=2)
    perm = np.random.default_rng(0).permutation(k).astype(np.int32)
    _, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16, perm=perm)
    got = oracle_c.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16, perm=perm)
    st = mo.error_stats(got.view(np.float16), exact, mo.FP16)
    assert st["max_rel"] <= 1e-2, st
