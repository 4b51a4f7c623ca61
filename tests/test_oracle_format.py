"""CPU: the oracle's layout functions against golden vectors produced by the reference's own
packer (tests/golden/make_format_golden.py imports scripts/model_convert/gptq2marlin.py)."""
import os

import numpy as np
import pytest

from oracle import marlin_oracle as mo

G = os.path.join(os.path.dirname(__file__), "golden")
PERMS = np.load(os.path.join(G, "format_perms.npz"))
CASES = np.load(os.path.join(G, "format_cases.npz"))
SHAPES = sorted({tuple(map(int, k.split("_")[-1].split("x"))) for k in CASES.files if k.startswith("gptq_")})


def test_perm_tables_match_reference_get_perms():
    assert np.array_equal(mo.PERM, PERMS["perm"])
    assert np.array_equal(mo.SCALE_PERM, PERMS["scale_perm"])
    assert np.array_equal(mo.SCALE_PERM_SINGLE, PERMS["scale_perm_single"])


@pytest.mark.parametrize("k,n", SHAPES)
def test_repack_matches_reference(k, n):
    gptq = CASES[f"gptq_{k}x{n}"]
    want = CASES[f"marlin_{k}x{n}"]
    got = mo.gptq_to_marlin_qweight(gptq)
    assert got.dtype == np.int32 and got.shape == (k // 16, 2 * n)
    assert np.array_equal(got, want)
    # and the inverse is a true inverse
    wq = mo.marlin_unpack(want)
    assert np.array_equal(wq, mo.gptq_unpack(gptq))
    assert np.array_equal(mo.gptq_pack(wq), gptq)


@pytest.mark.parametrize("k,n", SHAPES)
def test_scale_permutation_matches_reference(k, n):
    if f"scales_{k}x{n}" in CASES.files:
        s, sp = CASES[f"scales_{k}x{n}"], CASES[f"scales_marlin_{k}x{n}"]
        assert np.array_equal(mo.permute_scales(s, k, n, 128).view(np.uint16), sp.view(np.uint16))
        assert np.array_equal(mo.unpermute_scales(sp, k, n, 128).view(np.uint16), s.view(np.uint16))
    s, sp = CASES[f"chan_scales_{k}x{n}"], CASES[f"chan_scales_marlin_{k}x{n}"]
    assert np.array_equal(mo.permute_scales(s, k, n, -1).view(np.uint16), sp.view(np.uint16))
    assert np.array_equal(mo.unpermute_scales(sp, k, n, -1).view(np.uint16), s.view(np.uint16))


def test_fragment_order_is_the_mma_b_fragment():
    """lane (g,t), tile j, of a 16x64 block holds rows {2t,2t+1,2t+8,2t+9} x cols {16j+g, 16j+g+8}
    in nibble order [k0|n0, k8|n0, k0|n1, k8|n1, k1|n0, k9|n0, k1|n1, k9|n1]."""
    wq = (np.arange(16)[:, None] * 64 + np.arange(64)[None, :])          # value = 64*k + n (fits no nibble;
    idx = wq.reshape(1, 16, 4, 16).transpose(0, 2, 1, 3).reshape(-1)[mo.PERM]   # use the index map only)
    idx = idx.reshape(32, 4, 8)
    for lane in (0, 5, 17, 31):
        g, t = lane // 4, lane % 4
        for j in range(4):
            ks = [2 * t, 2 * t + 8, 2 * t, 2 * t + 8, 2 * t + 1, 2 * t + 9, 2 * t + 1, 2 * t + 9]
            ns = [g, g, g + 8, g + 8, g, g, g + 8, g + 8]
            want = [64 * k + 16 * j + n for k, n in zip(ks, ns)]
            assert idx[lane, j].tolist() == want


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
def test_dequant_word_is_q_minus_8_exact(dtype):
    """the lop3 / magic-number dequant (marlin_device_ops.cuh:91-137) yields q-8 exactly for all nibbles."""
    for nib in range(16):
        q = sum(nib << (4 * i) for i in range(8))
        assert np.array_equal(mo.dequant_word(q, dtype), np.full(8, nib - 8, dtype=np.float32))
    rng = np.random.default_rng(1)
    for q in rng.integers(0, 2**32, size=64, dtype=np.uint64):
        nibs = [(int(q) >> (4 * i)) & 0xF for i in range(8)]
        # register order: word -> {n0:(k0,k1)}, {n0:(k8,k9)} ; word>>8 -> {n1:(k0,k1)}, {n1:(k8,k9)}
        want = [nibs[0], nibs[4], nibs[1], nibs[5], nibs[2], nibs[6], nibs[3], nibs[7]]
        assert mo.dequant_word(int(q), dtype).tolist() == [float(v - 8) for v in want]


def test_dequant_single_rounding_all_scale_bit_patterns():
    """exhaustive KAT (SURVEY.md section 8(c) pin 1): for every finite fp16 scale and every nibble,
    W = RN16((q-8) * s) computed via exact fp32 product == via fp64 product."""
    bits = np.arange(0x10000, dtype=np.uint32).astype(np.uint16)
    s = bits.view(np.float16)
    finite = np.isfinite(s)
    s = s[finite]
    for nib in range(16):
        w32 = (np.float32(nib - 8) * s.astype(np.float32)).astype(np.float16)
        w64 = (np.float64(nib - 8) * s.astype(np.float64)).astype(np.float16)
        assert np.array_equal(w32.view(np.uint16), w64.view(np.uint16))


def test_reduce_max_m():
    # gptq_marlin_utils.cu:206-225
    assert [mo.determine_reduce_max_m(m) for m in (1, 16, 17, 32, 33, 48, 49, 64)] == [16, 16, 32, 32, 48, 48, 64, 64]
    assert mo.determine_reduce_max_m(65) == 128
    assert mo.determine_reduce_max_m(1024) == 1024
    assert mo.determine_reduce_max_m(4096) == 1024
