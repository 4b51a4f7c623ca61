"""CPU oracle for the W4A16 GPTQ-Marlin linear.   *** TEST INFRASTRUCTURE ONLY ***

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product (cpmcu_b200/) never does; it fails loudly
when its CUDA library is missing.

This is a numpy restatement of the reference algorithm (all file:line are relative to
the reference tree, jk3456a/CPM.cu @ 797eb93):

  * the Marlin weight / scale layout   scripts/model_convert/gptq2marlin.py:35-61, 99-134
  * int4 -> fp16/bf16 dequantisation    src/qgemm/gptq_marlin/marlin_device_ops.cuh:91-137
  * the group-scale multiply (1 RN)    src/qgemm/gptq_marlin/marlin_device_ops.cuh:294-303
  * fp32 accumulation, RN to T at end  src/qgemm/gptq_marlin/marlin_kernel_impl.cuh:944-956
  * channelwise scale after rounding   src/qgemm/gptq_marlin/marlin_kernel_impl.cuh:957-963
  * act-order column gather of A       src/qgemm/gptq_marlin/gptq_marlin_utils.cu:28-77
  * post-ops inplace-add / bias        src/model/w4a16_gptq_marlin/w4a16_gptq_marlin_linear.cuh:140-145

Parity pinning (see DESIGN.md "Oracle"):
  - layout functions: pinned against golden vectors produced by importing the reference's
    own packer (tests/golden/make_format_golden.py -> tests/golden/format_*.npz).
  - dequant + GEMM: the reference has no CPU implementation and no golden vectors for the
    GEMM; it is pinned against outputs of the reference Marlin kernel itself
    (oracle/_ref/libmarlin_ref.so run on a B200; tests/golden/make_gemm_golden.py ->
    tests/golden/gemm_ref_*.npz) and live on the GPU box in tests/test_gpu_parity.py.
"""
from __future__ import annotations

import numpy as np

FP16, BF16 = 0, 1          # dtype codes: cpmcu/llm_w4a16_gptq_marlin.py:12-15
TILE = 16                  # marlin.cuh:23
CHUNK_N = 64               # marlin.cuh:19  (min_thread_n; one 512-byte block = 16 k x 64 n)


# ----------------------------------------------------------------------------------------------
# layout
# ----------------------------------------------------------------------------------------------
def _fragment_gather_index() -> np.ndarray:
    """1024-entry gather index of one (16 k x 64 n) block.  gptq2marlin.py:35-53.

    The block is seen as 4 tiles of 16x16 (tile j = columns 16j..16j+15), each flattened
    row-major (16*row + col) at offset 256*j.  Output position = lane*32 + j*8 + slot:
    lane (g = lane//4, t = lane%4) of a warp gets, for every tile j, the eight nibbles of its
    mma.m16n8k16 B fragments: rows {2t, 2t+1, 2t+8, 2t+9} x cols {g, g+8}.  Slots are then
    interleaved [0,2,4,6,1,3,5,7] so that `q & 0x000f000f`, `q & 0x00f000f0` and the same on
    `q >> 8` peel (k, k+1) pairs as half2 (marlin_device_ops.cuh:91-112).
    """
    lane = np.arange(32)[:, None, None]
    tile = np.arange(4)[None, :, None]
    slot = np.arange(8)[None, None, :]
    e = np.array([0, 2, 4, 6, 1, 3, 5, 7])[slot]          # element held in nibble `slot`
    half, r = e // 4, e % 4                               # half: 0 -> col g, 1 -> col g+8
    t, g = lane % 4, lane // 4
    row = 2 * t + (r % 2) + 8 * (r // 2)
    col = g + 8 * half
    return (16 * row + col + 256 * tile).reshape(-1)


PERM = _fragment_gather_index()
# scale permutations, gptq2marlin.py:54-60
SCALE_PERM = np.array([i + 8 * j for i in range(8) for j in range(8)])            # grouped
SCALE_PERM_SINGLE = np.array(
    [2 * i + j for i in range(4) for j in (0, 1, 8, 9, 16, 17, 24, 25)])          # channelwise


def gptq_unpack(qweight: np.ndarray) -> np.ndarray:
    """AutoGPTQ int32[K/8][N] (nibble i of a word = row 8r+i) -> uint8[K][N].  gptq2marlin.py:111-118"""
    q = np.ascontiguousarray(qweight).view(np.uint32)
    k8, n = q.shape
    shifts = (4 * np.arange(8, dtype=np.uint32))[None, :, None]
    return ((q[:, None, :] >> shifts) & 0xF).astype(np.uint8).reshape(k8 * 8, n)


def gptq_pack(wq: np.ndarray) -> np.ndarray:
    """inverse of gptq_unpack (used to synthesise AutoGPTQ checkpoints for tests)."""
    k, n = wq.shape
    w = wq.astype(np.uint32).reshape(k // 8, 8, n)
    out = np.zeros((k // 8, n), dtype=np.uint32)
    for i in range(8):
        out |= w[:, i, :] << np.uint32(4 * i)
    return out.view(np.int32)


def marlin_pack(wq: np.ndarray) -> np.ndarray:
    """uint8[K][N] nibbles -> Marlin int32[K/16][2N].  gptq2marlin.py:120-132"""
    k, n = wq.shape
    assert k % TILE == 0 and n % CHUNK_N == 0, (k, n)
    tiles = wq.reshape(k // TILE, TILE, n // TILE, TILE).transpose(0, 2, 1, 3)
    flat = tiles.reshape(k // TILE, n * TILE)
    flat = flat.reshape(-1, PERM.size)[:, PERM].reshape(flat.shape).astype(np.uint32)
    words = np.zeros((k // TILE, n * TILE // 8), dtype=np.uint32)
    for i in range(8):
        words |= flat[:, i::8] << np.uint32(4 * i)
    return words.view(np.int32)


def marlin_unpack(qweight: np.ndarray) -> np.ndarray:
    """Marlin int32[K/16][2N] -> uint8[K][N] nibbles (inverse of marlin_pack)."""
    q = np.ascontiguousarray(qweight).view(np.uint32)
    k16, n2 = q.shape
    n = n2 // 2
    flat = np.empty((k16, n2 * 8), dtype=np.uint8)
    for i in range(8):
        flat[:, i::8] = (q >> np.uint32(4 * i)) & 0xF
    inv = np.empty_like(PERM)
    inv[PERM] = np.arange(PERM.size)
    flat = flat.reshape(-1, PERM.size)[:, inv].reshape(k16, n * TILE)
    return flat.reshape(k16, n // TILE, TILE, TILE).transpose(0, 2, 1, 3).reshape(k16 * TILE, n)


def gptq_to_marlin_qweight(qweight_gptq: np.ndarray) -> np.ndarray:
    """marlin_repack_qweight restated: gptq2marlin.py:109-134"""
    return marlin_pack(gptq_unpack(qweight_gptq))


def permute_scales(s: np.ndarray, size_k: int, size_n: int, group_size: int) -> np.ndarray:
    """gptq2marlin.py:99-107"""
    p = SCALE_PERM if (group_size < size_k and group_size != -1) else SCALE_PERM_SINGLE
    return np.ascontiguousarray(s.reshape(-1, p.size)[:, p].reshape(-1, size_n))


def unpermute_scales(s: np.ndarray, size_k: int, size_n: int, group_size: int,
                     kernel_view: bool = False) -> np.ndarray:
    """inverse of permute_scales.  kernel_view=True selects the layout the reference KERNEL reads
    (grouped whenever group_size != -1, marlin_kernel_impl.cuh:258-264), which differs from the
    converter's choice only in the degenerate case group_size == size_k."""
    grouped = (group_size != -1) if kernel_view else (group_size < size_k and group_size != -1)
    p = SCALE_PERM if grouped else SCALE_PERM_SINGLE
    inv = np.empty_like(p)
    inv[p] = np.arange(p.size)
    return np.ascontiguousarray(s.reshape(-1, p.size)[:, inv].reshape(-1, size_n))


# ----------------------------------------------------------------------------------------------
# numerics
# ----------------------------------------------------------------------------------------------
def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """round-to-nearest-even fp32 -> bf16 (returned as uint16 bit patterns)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    nan = (u & 0x7FFFFFFF) > 0x7F800000
    r = ((u + (0x7FFF + ((u >> 16) & 1))) >> 16).astype(np.uint16)
    r[nan] = 0x7FC0
    return r


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b).astype(np.uint32) << 16).view(np.float32)


def to_storage(x: np.ndarray, dtype: int) -> np.ndarray:
    """fp32/fp64 array -> storage array: float16 for FP16, uint16 bf16 bit patterns for BF16."""
    return x.astype(np.float16) if dtype == FP16 else f32_to_bf16_bits(x.astype(np.float32))


def to_f32(x: np.ndarray, dtype: int) -> np.ndarray:
    return x.astype(np.float32) if dtype == FP16 else bf16_bits_to_f32(x)


def dequant_word(q: int, dtype: int = FP16) -> np.ndarray:
    """Bit-level restatement of dequant<T, kU4B8>(q) ++ dequant(q >> 8) for one 32-bit word:
    returns the 8 values in register order [b0.lo.x, b0.lo.y, b0.hi.x, b0.hi.y, b1...] as fp32.
    marlin_device_ops.cuh:91-112 (fp16) / 114-137 (bf16), marlin_kernel_impl.cuh:703-719."""
    out = []
    for w in (q & 0xFFFFFFFF, (q & 0xFFFFFFFF) >> 8):
        if dtype == FP16:
            lo = (w & 0x000F000F) | 0x64006400
            hi = (w & 0x00F000F0) | 0x64006400
            lo_h = np.array([lo & 0xFFFF, lo >> 16], dtype=np.uint16).view(np.float16)
            hi_h = np.array([hi & 0xFFFF, hi >> 16], dtype=np.uint16).view(np.float16)
            sub = np.array([0x6408], dtype=np.uint16).view(np.float16)[0]
            mul = np.array([0x2C00], dtype=np.uint16).view(np.float16)[0]
            add = np.array([0xD480], dtype=np.uint16).view(np.float16)[0]
            a = (lo_h.astype(np.float32) - np.float32(sub)).astype(np.float16)
            b = (hi_h.astype(np.float64) * np.float64(mul) + np.float64(add)).astype(np.float16)
            out += [np.float32(a[0]), np.float32(a[1]), np.float32(b[0]), np.float32(b[1])]
        else:
            lo = (w & 0x000F000F) | 0x43004300
            hi = ((w >> 4) & 0x000F000F) | 0x43004300
            for v in (lo, hi):
                h = bf16_bits_to_f32(np.array([v & 0xFFFF, v >> 16], dtype=np.uint16))
                r = bf16_bits_to_f32(f32_to_bf16_bits(h * np.float32(1.0) + np.float32(-136.0)))
                out += [r[0], r[1]]
    return np.array(out, dtype=np.float32)


def dequant(qweight: np.ndarray, scales_marlin: np.ndarray, size_k: int, size_n: int,
            group_size: int, dtype: int = FP16) -> np.ndarray:
    """Marlin qweight + Marlin-permuted scales -> W[K][N] in storage dtype, rounded exactly
    where the reference rounds: W = RN_T( T(q-8) * T(s) ).  For channelwise (group_size == -1)
    the reference does NOT scale the weights (scale is applied to the rounded output), so this
    returns T(q-8) and the caller applies the scale (see gemm())."""
    wq = marlin_unpack(qweight).astype(np.float32) - 8.0           # exact small integers
    if group_size == -1:
        return to_storage(wq, dtype)
    s = unpermute_scales(np.asarray(scales_marlin), size_k, size_n, group_size, kernel_view=True)
    s32 = to_f32(s, dtype)                                          # [K/g][N]
    s_full = np.repeat(s32, group_size, axis=0)                     # [K][N]
    # product of two <=11-bit significands is exact in fp32 -> one RN to T, as HMUL2 does
    return to_storage(wq * s_full, dtype)


def gemm(a: np.ndarray, qweight: np.ndarray, scales_marlin: np.ndarray, size_k: int, size_n: int,
         group_size: int = 128, dtype: int = FP16, perm: np.ndarray | None = None,
         accumulate: str = "f64") -> tuple[np.ndarray, np.ndarray]:
    """C = RN_T( sum_k fp32(A[m,k]) * fp32(W[k,n]) ), fp32 accumulators in the reference
    (marlin_dtypes.cuh:24); we accumulate in fp64 ("f64", the order-free truth the tolerance is
    measured against) or fp32 ("f32").  Returns (C in storage dtype, C before the final rounding)."""
    a32 = to_f32(np.asarray(a), dtype)
    if perm is not None:                       # act-order: a_tmp[m,k] = A[m, perm[k]]
        a32 = a32[:, np.asarray(perm)]
    w = to_f32(dequant(qweight, scales_marlin, size_k, size_n, group_size, dtype), dtype)
    if accumulate == "f64":
        acc = a32.astype(np.float64) @ w.astype(np.float64)
    else:
        acc = a32 @ w
    if group_size == -1:
        # channelwise: the fp32 sum is rounded to T first, THEN multiplied by the scale in T
        # (marlin_kernel_impl.cuh:955-963) -> two roundings.
        s = to_f32(unpermute_scales(np.asarray(scales_marlin), size_k, size_n, -1), dtype)
        c1 = to_f32(to_storage(acc, dtype), dtype)
        exact = c1.astype(np.float64) * s.astype(np.float64)
        return to_storage(exact, dtype), exact
    return to_storage(acc, dtype), acc


def linear_forward(a, qweight, scales_marlin, size_k, size_n, group_size=128, dtype=FP16,
                   bias=None, tgt=None, inplace=False, out_scale=0.0):
    """W4A16GPTQMarlinLinear::prefill post-ops (w4a16_gptq_marlin_linear.cuh:104-146):
    inplace -> tgt = tgt + out (elementwise.cuh:17-23, computed in T), has_bias -> tgt += bias."""
    out, _ = gemm(a, qweight, scales_marlin, size_k, size_n, group_size, dtype)
    res = to_f32(out, dtype)
    if inplace:
        assert tgt is not None
        res = to_f32(to_storage(to_f32(np.asarray(tgt), dtype) + res, dtype), dtype)
    if bias is not None:
        res = to_f32(to_storage(res + to_f32(np.asarray(bias), dtype)[None, :], dtype), dtype)
    if out_scale:
        # the layer's residual scaling, applied to the stored output afterwards: out = out * T(v), both in T
        # (elementwise_scale_kernel, src/model/elementwise.cuh:35-41; call sites
        # w4a16_gptq_marlin_layer.cuh:57,71,79,93).  The fp32 product of two T values is exact -> one RN.
        sv = to_f32(to_storage(np.array([out_scale], dtype=np.float32), dtype), dtype)[0]
        res = to_f32(to_storage(res * sv, dtype), dtype)
    return to_storage(res, dtype)


def gated_silu_interleaved(gate_up, dtype=FP16):
    """gated_silu_interleaved_kernel (src/model/activation.cuh:7-18): src is gateup_proj's stored output
    T[M][2I] = [gate | up]; tgt[m][j] = T(g * (1 / (1 + expf(-g))) * u) with g, u = float(src), fp32 math."""
    x = to_f32(np.asarray(gate_up), dtype)
    half = x.shape[1] // 2
    g, u = x[:, :half].astype(np.float32), x[:, half:].astype(np.float32)
    sg = (np.float32(1.0) / (np.float32(1.0) + np.exp(-g, dtype=np.float32))).astype(np.float32)
    return to_storage((g * sg) * u, dtype)


def dense_linear(a, weight, dtype=FP16, bias=None, tgt=None, inplace=False, head_scale=1.0):
    """Un-quantised Linear<T>::prefill / LMHead<T>::prefill (src/model/linear.cuh:8-37,76-105) restated:
      LMHead: input * T(head_scale), rounded to T (elementwise_scale_kernel, elementwise.cuh:35-41);
      cublasGemmEx(OP_T, OP_N, CUBLAS_COMPUTE_32F): out = RN_T(sum_k a[m,k] w[n,k] + beta * out), beta = 1
      when inplace (one rounding); has_bias: out = RN_T(out + bias) (batched_add, elementwise.cuh:8-14).
    weight is [N][K] (`transposed = true`).  The sum is taken in fp64 (order-free truth; cuBLAS and the
    CUDA kernel accumulate in fp32 in their own orders).  Returns (out in storage dtype, out before the
    roundings of the GEMM result as real numbers)."""
    a32 = to_f32(np.asarray(a), dtype)
    if head_scale != 1.0:
        sv = to_f32(to_storage(np.array([head_scale], dtype=np.float32), dtype), dtype)[0]
        a32 = to_f32(to_storage(a32 * sv, dtype), dtype)   # fp32 product of two T values is exact -> one RN
    w32 = to_f32(np.asarray(weight), dtype)
    acc = a32.astype(np.float64) @ w32.astype(np.float64).T
    if inplace:
        assert tgt is not None
        acc = acc + to_f32(np.asarray(tgt), dtype).astype(np.float64)
    exact = acc
    res = to_f32(to_storage(acc, dtype), dtype)
    if bias is not None:
        b = to_f32(np.asarray(bias), dtype)[None, :]
        exact = exact + b
        res = to_f32(to_storage(res + b, dtype), dtype)
    return to_storage(res, dtype), exact


def determine_reduce_max_m(prob_m: int, max_par: int = 16) -> int:
    """gptq_marlin_utils.cu:206-225"""
    for i in (1, 2, 3, 4):
        if prob_m <= 16 * i:
            return 16 * i
    return 64 * min(-(-prob_m // 64), max_par)


# ----------------------------------------------------------------------------------------------
# synthetic problems (Wq ~ U{0..15}, scales ~ U(0.002, 0.02), A ~ N(0,1))
# ----------------------------------------------------------------------------------------------
def make_problem(m: int, k: int, n: int, group_size: int = 128, dtype: int = FP16, seed: int = 0):
    rng = np.random.default_rng(seed)
    wq = rng.integers(0, 16, size=(k, n), dtype=np.uint8)
    groups = 1 if group_size == -1 else k // group_size
    s = to_storage(rng.uniform(0.002, 0.02, size=(groups, n)).astype(np.float32), dtype)
    a = to_storage(rng.standard_normal((m, k)).astype(np.float32), dtype)
    return dict(a=a, wq=wq, qweight=marlin_pack(wq), scales=s,
                scales_marlin=permute_scales(s, k, n, group_size))


def error_stats(c: np.ndarray, ref: np.ndarray, dtype: int = FP16) -> dict:
    """`c` in storage dtype, `ref` as real numbers (fp32/fp64).  max relative error as SURVEY.md
    section 8(c) defines it: max |d| / max(|ref|, eps), with the floor eps = 1e-2 * mean|ref| so that
    outputs that cancel to ~0 are compared absolutely; plus mean / max absolute error."""
    c64 = to_f32(np.asarray(c), dtype).astype(np.float64)
    r = np.asarray(ref, dtype=np.float64)
    d = np.abs(c64 - r)
    floor = max(1e-2 * float(np.mean(np.abs(r))), 1e-6)
    return dict(max_rel=float(np.max(d / np.maximum(np.abs(r), floor))),
                mean_abs=float(np.mean(d)), max_abs=float(np.max(d)),
                ref_mean_abs=float(np.mean(np.abs(r))))
