"""GPU parity (-m gpu) of the un-quantised streaming linear (lm_head / draft heads at M <= 64,
SURVEY.md section 8(f)-2; reference: Linear<T>/LMHead<T>::prefill, src/model/linear.cuh:8-37,76-105),
called through the C ABI, against
  (1) the numpy oracle (oracle/marlin_oracle.py: dense_linear, fp64 sums + the reference's rounding points),
  (2) cuBLAS with fp32 accumulation via torch (what the reference's linear<T>() calls) at the full lm_head size.
Bar: max rel err <= 1e-2 (north_star's GEMM tolerance), mean abs err reported."""
import numpy as np
import pytest
import torch

from oracle import marlin_oracle as mo

pytestmark = pytest.mark.gpu

TOL_MAX_REL = 1e-2
DEV = "cuda"


def _b():
    import cpmcu_b200
    return cpmcu_b200


def t_store(x, dtype):
    if dtype == mo.FP16:
        return torch.from_numpy(np.ascontiguousarray(x)).to(DEV)
    return torch.from_numpy(np.ascontiguousarray(x).view(np.int16)).to(DEV).view(torch.bfloat16)


def t_load(t, dtype):
    if dtype == mo.FP16:
        return t.cpu().numpy()
    return t.view(torch.int16).cpu().numpy().view(np.uint16)


def assert_same_pipeline(got, want, stages, dtype):
    """`got` against the oracle's output computed with the SAME rounding points: the two may differ by a last-bit
    flip at each rounding stage (fp32 accumulation in kernel order vs the oracle's fp64 sums), i.e. by one ulp of
    the magnitudes in `stages` (real-valued intermediates), plus the fp32 summation noise itself."""
    g32, w32 = mo.to_f32(got, dtype), mo.to_f32(want, dtype)
    ulp = 2.0 ** -10 if dtype == mo.FP16 else 2.0 ** -7
    bound = ulp * (sum(np.abs(x) for x in stages) + np.abs(w32)) + 2e-5
    bad = np.abs(g32 - w32) > bound
    assert not bad.any(), (int(bad.sum()), float(np.abs(g32 - w32).max()))


def problem(m, n, k, dtype, seed):
    rng = np.random.default_rng(seed)
    a = mo.to_storage(rng.standard_normal((m, k)).astype(np.float32), dtype)
    w = mo.to_storage((rng.standard_normal((n, k)) * 0.05).astype(np.float32), dtype)
    bias = mo.to_storage(rng.standard_normal(n).astype(np.float32), dtype)
    old = mo.to_storage(rng.standard_normal((m, n)).astype(np.float32), dtype)
    return a, w, bias, old


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
@pytest.mark.parametrize("m", [1, 4, 8, 9, 16, 17, 32, 33, 60, 64])
@pytest.mark.parametrize("n,k", [(1000, 512), (264, 4096), (1002, 520), (64, 64), (8, 8)])
def test_dense_vs_oracle(dtype, m, n, k):
    """ragged N (not a multiple of the 64/128-row tile, nor of 4), K not a multiple of the 128-k stage."""
    b = _b()
    a, w, _, _ = problem(m, n, k, dtype, seed=m * 7 + n)
    out = b.dense_linear(t_store(a, dtype), t_store(w, dtype))
    torch.cuda.synchronize()
    _, exact = mo.dense_linear(a, w, dtype)
    st = mo.error_stats(t_load(out, dtype), exact, dtype)
    assert st["max_rel"] <= TOL_MAX_REL, st


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
@pytest.mark.parametrize("m", [1, 12, 24, 48])
@pytest.mark.parametrize("bias,inplace,scale", [(True, False, 1.0), (False, True, 1.0), (False, False, 0.0625),
                                                (True, True, 0.37)])
def test_dense_postops_vs_oracle(dtype, m, bias, inplace, scale):
    """bias (batched_add), inplace (cuBLAS beta = 1), LMHead head_scale: the reference's rounding points."""
    b = _b()
    n, k = 1160, 1024
    a, w, bs, old = problem(m, n, k, dtype, seed=11 + m)
    out = t_store(old, dtype).clone()
    got = b.dense_linear(t_store(a, dtype), t_store(w, dtype), out if inplace else None,
                         bias=t_store(bs, dtype) if bias else None, inplace=inplace, input_scale=scale)
    torch.cuda.synchronize()
    want, exact = mo.dense_linear(a, w, dtype, bias=bs if bias else None, tgt=old, inplace=inplace, head_scale=scale)
    # The reference rounds the GEMM (+ old output) to T and THEN adds the bias in T: where the terms cancel, the
    # error relative to the exact sum is unbounded by design.  So: agreement with the oracle's own rounded
    # pipeline, and north_star's tolerance against the exact value where no second rounding stage exists.
    pre = exact - (mo.to_f32(bs, dtype)[None, :].astype(np.float64) if bias else 0.0)
    gemm_mag = np.abs(pre) + (np.abs(mo.to_f32(old, dtype)) if inplace else 0.0)
    assert_same_pipeline(t_load(got, dtype), want, [pre, gemm_mag * 2.0 ** -6], dtype)
    if not bias and not inplace:
        st = mo.error_stats(t_load(got, dtype), exact, dtype)
        assert st["max_rel"] <= TOL_MAX_REL, st


@pytest.mark.parametrize("m", [1, 16, 64])
def test_lm_head_full_size_vs_cublas(m):
    """MiniCPM4-8B lm_head (73448 x 4096): against cuBLAS fp32-accumulate (the reference's
    linear<T>()), and the size-independent property that appending token rows does not change earlier rows."""
    b = _b()
    n, k = 73448, 4096
    g = torch.Generator(device=DEV).manual_seed(5)
    w = (torch.randn((n, k), generator=g, device=DEV, dtype=torch.float32) * 0.02).half()
    a = torch.randn((m, k), generator=g, device=DEV, dtype=torch.float32).half()
    out = b.dense_linear(a, w)
    ref = torch.nn.functional.linear(a.float(), w.float())      # fp32 cuBLAS, the order-free-enough truth here
    torch.cuda.synchronize()
    d = (out.float() - ref).abs()
    floor = max(1e-2 * float(ref.abs().mean()), 1e-6)
    max_rel = float((d / ref.abs().clamp_min(floor)).max())
    print(f"lm_head m={m}: max_rel={max_rel:.3e} mean_abs={float(d.mean()):.3e}")
    assert max_rel <= TOL_MAX_REL
    if m > 1:
        first = b.dense_linear(a[:1].contiguous(), w)
        torch.cuda.synchronize()
        # other token-block widths use other tile shapes (another fp32 summation order): last bit only
        df = (first[0].float() - out[0].float()).abs()
        assert bool((df <= 2.0 ** -10 * out[0].float().abs() + 2e-5).all()), "row 0 must not depend on the rows that follow"


def test_mirror_lmhead_and_errors():
    """host-side mirrors (same method names as src/model/linear.cuh) and the error convention."""
    b = _b()
    from cpmcu_b200.linear import LMHead, Linear
    k, n, m = 512, 1000, 5
    a, w, bs, _ = problem(m, n, k, mo.FP16, seed=2)
    head = LMHead(k, n, head_scale=0.25)
    head.load_to_storage("lm_head.weight", w)
    head.init_output_ptr(None, 64)
    out = head.prefill(None, m, t_store(a, mo.FP16))
    torch.cuda.synchronize()
    _, exact = mo.dense_linear(a, w, mo.FP16, head_scale=0.25)
    assert mo.error_stats(out.cpu().numpy(), exact, mo.FP16)["max_rel"] <= TOL_MAX_REL
    fc = Linear(k, n, has_bias=True)
    fc.load_to_storage("fc1.weight", w)
    fc.load_to_storage("fc1.bias", bs)
    with pytest.raises(ValueError):
        fc.load_to_storage("fc1.scales", bs)
    out = fc.prefill(None, m, t_store(a, mo.FP16))
    torch.cuda.synchronize()
    want, exact = mo.dense_linear(a, w, mo.FP16, bias=bs)
    assert_same_pipeline(out.cpu().numpy(), want, [exact - mo.to_f32(bs, mo.FP16)[None, :]], mo.FP16)
    big = torch.zeros((65, k), dtype=torch.float16, device=DEV)
    with pytest.raises(ValueError):
        b.dense_linear(big, t_store(w, mo.FP16))
    with pytest.raises(ValueError):
        b.dense_linear(torch.zeros((1, 12), dtype=torch.float16, device=DEV),
                       torch.zeros((8, 12), dtype=torch.float16, device=DEV))


def test_dense_graph_capture_and_pdl():
    """capturable into a CUDA graph (reference contract, src/entry.cu:540-562), with PDL behind another launch."""
    b = _b()
    k, n, m = 1024, 2048, 3
    a, w, _, _ = problem(m, n, k, mo.FP16, seed=9)
    ta, tw = t_store(a, mo.FP16), t_store(w, mo.FP16)
    out1 = torch.empty((m, n), dtype=torch.float16, device=DEV)
    out2 = torch.empty((m, n), dtype=torch.float16, device=DEV)
    b.dense_linear(ta, tw, out1)  # warm-up: descriptor cache, function attributes
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        with torch.cuda.graph(graph, stream=s):
            b.dense_linear(ta, tw, out1, pdl=True, stream=s)
            b.dense_linear(ta, tw, out2, pdl=True, stream=s)
    out1.zero_()
    out2.zero_()
    graph.replay()
    torch.cuda.synchronize()
    _, exact = mo.dense_linear(a, w, mo.FP16)
    assert mo.error_stats(out1.cpu().numpy(), exact, mo.FP16)["max_rel"] <= TOL_MAX_REL
    assert torch.equal(out1, out2)
