"""GPU parity (-m gpu): the CUDA path, called through the C ABI, against
  (1) the numpy oracle (oracle/marlin_oracle.py) on seeded inputs at sizes it finishes in seconds,
  (2) the UNMODIFIED reference Marlin kernel (oracle/_ref/libmarlin_ref.so) on the same inputs,
  (3) size-independent properties at BASELINE.json's full sizes.
Bar: dequant bit-exact; GEMM output max rel err <= 1e-2 (north_star), mean abs err reported."""
import numpy as np
import pytest
import torch

from oracle import marlin_oracle as mo
from oracle import ref_marlin

pytestmark = pytest.mark.gpu

TOL_MAX_REL = 1e-2  # north_star: "max rel err <= 1e-2, and report mean abs err"
DEV = "cuda"


def _b():
    import cpmcu_b200
    return cpmcu_b200


def t_store(x, dtype):
    """oracle storage array -> torch tensor on the GPU"""
    if dtype == mo.FP16:
        return torch.from_numpy(np.ascontiguousarray(x)).to(DEV)
    return torch.from_numpy(np.ascontiguousarray(x).view(np.int16)).to(DEV).view(torch.bfloat16)


def t_load(t, dtype):
    if dtype == mo.FP16:
        return t.cpu().numpy()
    return t.view(torch.int16).cpu().numpy().view(np.uint16)


def scratch(m, n, k, contract=False):
    """workspace + c_tmp.  Default: room for full-width stream-K (reduce_bytes_fast) and for the reference kernel;
    contract=True: EXACTLY the reference's rule, determine_reduce_max_m(M, 16) x N floats
    (gptq_marlin_utils.cu:206-225), followed by a guard region that must stay untouched."""
    b = _b()
    ws = torch.zeros(b.workspace_ints(n), dtype=torch.int32, device=DEV)
    ref_floats = b.reduce_max_m(m) * n
    if contract:
        buf = torch.full((ref_floats + GUARD,), GUARD_VAL, dtype=torch.float32, device=DEV)
        return ws, buf[:ref_floats]
    c_tmp = torch.empty(max(b.reduce_bytes_fast(m, n, k) // 4, ref_floats, 1), dtype=torch.float32, device=DEV)
    return ws, c_tmp


GUARD, GUARD_VAL = 1 << 16, -12345.0


def guard_intact(c_tmp):
    """the floats behind a contract-sized c_tmp (see scratch) were not written"""
    base = c_tmp._base if c_tmp._base is not None else c_tmp
    return bool((base[c_tmp.numel():] == GUARD_VAL).all())


def run_ours(prob, m, n, k, dtype, group=128, contract=False, **kw):
    b = _b()
    a = t_store(prob["a"], dtype)
    qw = torch.from_numpy(prob["qweight"]).to(DEV)
    sc = t_store(prob["scales_marlin"], dtype)
    ws, c_tmp = scratch(m, n, k, contract)
    out = b.gemm(a, qw, sc, ws, c_tmp, group_size=group, **kw)
    torch.cuda.synchronize()
    assert int(ws.abs().max()) == 0, "workspace must be returned zeroed (reference lock contract)"
    if contract:
        assert guard_intact(c_tmp), "wrote past determine_reduce_max_m(M) x N floats of c_tmp"
    return out, (a, qw, sc, ws, c_tmp)


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
@pytest.mark.parametrize("group", [128, -1])
def test_dequant_bit_exact(dtype, group):
    """int4 unpack + group dequant must be bit-exact vs the oracle (incl. subnormal / huge scales)."""
    b = _b()
    k, n = 512, 256
    prob = mo.make_problem(1, k, n, group, dtype, seed=3)
    s = mo.to_f32(prob["scales"], dtype).copy()
    s[0, :8] = [6e-8, 1e-7, 3e-5, 6.1e-5, 1000.0, 8000.0, -0.01, 0.0]      # subnormal .. overflow-ish .. negative
    s_st = mo.to_storage(s, dtype)
    sm = mo.permute_scales(s_st, k, n, group)
    want = mo.dequant(prob["qweight"], sm, k, n, group, dtype)
    got = b.dequant(torch.from_numpy(prob["qweight"]).to(DEV), t_store(sm, dtype), k, n, group_size=group)
    torch.cuda.synchronize()
    got = t_load(got, dtype)
    assert np.array_equal(got.view(np.uint16), want.view(np.uint16))


def test_dequant_exhaustive_scales_fp16():
    """16 nibbles x all 63488 finite fp16 scale bit patterns (SURVEY.md section 8(c) pin 1), on the device."""
    b = _b()
    bits = np.arange(0x10000, dtype=np.uint32).astype(np.uint16)
    finite = np.isfinite(bits.view(np.float16))
    sbits = bits[finite]
    n = (len(sbits) + 63) // 64 * 64
    s = np.zeros(n, dtype=np.uint16)
    s[:len(sbits)] = sbits
    k = 256                                                                        # two groups, same scales
    wq = np.tile(np.arange(16, dtype=np.uint8).repeat(16)[:, None], (1, n))        # every nibble in every column
    qw = mo.marlin_pack(wq)
    s16 = np.tile(s.view(np.float16).reshape(1, n), (2, 1))
    sm = mo.permute_scales(s16, k, n, 128)
    want = mo.dequant(qw, sm, k, n, 128, mo.FP16)
    got = b.dequant(torch.from_numpy(qw).to(DEV), torch.from_numpy(sm).to(DEV), k, n).cpu().numpy()
    assert np.array_equal(got.view(np.uint16), want.view(np.uint16))


SMALL = [  # (m, k, n): config 1 of BASELINE.json first, then ragged / edge shapes
    (1, 4096, 4096), (1, 128, 64), (3, 256, 192), (4, 1024, 320), (8, 512, 1088), (10, 2048, 576),
    (12, 4096, 4608), (16, 1024, 2048), (17, 640, 448), (32, 1152, 1024), (33, 512, 512), (60, 768, 640),
    (64, 1024, 1024), (100, 256, 256),
    # tcgen05 regime (M > 64): ragged token tiles, N not a multiple of the 128/256-column tile
    (65, 128, 64), (128, 512, 384), (257, 256, 448), (354, 1024, 640), (513, 512, 1152),
]


@pytest.mark.parametrize("m,k,n", SMALL)
@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
def test_gemm_vs_oracle_and_reference(m, k, n, dtype):
    prob = mo.make_problem(m, k, n, 128, dtype, seed=m * 7 + n)
    out, (a, qw, sc, ws, c_tmp) = run_ours(prob, m, n, k, dtype)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, dtype)
    st = mo.error_stats(t_load(out, dtype), exact, dtype)
    # half an ulp of T at the output is the floor of any implementation: 2^-11 (fp16), 2^-8 (bf16)
    assert st["max_rel"] <= TOL_MAX_REL, st
    if ref_marlin.available():
        ref = ref_marlin.gemm(a, qw, sc, ws, c_tmp)
        torch.cuda.synchronize()
        st_ref = mo.error_stats(t_load(ref, dtype), exact, dtype)
        st_x = mo.error_stats(t_load(out, dtype), mo.to_f32(t_load(ref, dtype), dtype), dtype)
        assert st_x["max_rel"] <= TOL_MAX_REL, ("ours vs reference Marlin", st_x)
        # we must be at least as close to the exact result as the reference kernel is (x2 slack)
        assert st["mean_abs"] <= 2 * st_ref["mean_abs"] + 1e-7, (st, st_ref)
    print(f"\n[m={m} k={k} n={n} dt={dtype}] vs exact: max_rel={st['max_rel']:.3e} mean_abs={st['mean_abs']:.3e}")


CONTRACT = [(1, 4096, 4096), (2, 4096, 4608), (4, 1024, 320), (5, 4096, 4608), (8, 512, 1088), (9, 4096, 4096),
            (12, 4096, 4608), (16, 4096, 4096), (16, 4096, 32768), (17, 640, 448), (24, 16384, 4096), (32, 1152, 1024)]


@pytest.mark.parametrize("m,k,n", CONTRACT)
def test_gemm_inside_reference_scratch_contract(m, k, n):
    """c_tmp sized EXACTLY by the reference's rule (determine_reduce_max_m(M, 16) x N floats: what
    w4a16_gptq_marlin_layer.cuh:34-37 allocates for chunk_length = M) followed by a guard region: same answer as
    with ample scratch (within tolerance of the exact result), nothing written behind the contract size, and the
    drop-in entry point without any size (b200_w4a16_gemm, what the gptq_marlin_gemm shim calls) behaves the same."""
    b = _b()
    from cpmcu_b200 import capi as L
    prob = mo.make_problem(m, k, n, 128, mo.FP16, seed=m + k + n)
    out, (a, qw, sc, ws, c_tmp) = run_ours(prob, m, n, k, mo.FP16, contract=True)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16)
    st = mo.error_stats(out.cpu().numpy(), exact, mo.FP16)
    assert st["max_rel"] <= TOL_MAX_REL, st
    assert L.reduce_bytes(m, n, k) <= L.reduce_max_m(m) * n * 4
    # raw entry point, no size information at all
    out2 = torch.empty_like(out)
    rc = L.lib().b200_w4a16_gemm(L._ptr(a), L._ptr(qw), L._ptr(sc), None, None, L._ptr(ws), L._ptr(out2), m, n, k, 128, 0, 1, 0,
                                 None, L._ptr(c_tmp), L._stream(None))
    assert rc == 0, L.lib().b200_w4a16_last_error()
    torch.cuda.synchronize()
    assert guard_intact(c_tmp) and int(ws.abs().max()) == 0
    assert torch.equal(out, out2)


@pytest.mark.parametrize("m,k,n", [(1, 1024, 512), (16, 2048, 1024), (40, 512, 256), (130, 512, 320), (600, 256, 256)])
@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
def test_gemm_channelwise(m, k, n, dtype):
    """group_size = -1: scale applied to the ROUNDED output (two roundings) — marlin_kernel_impl.cuh:957-963"""
    prob = mo.make_problem(m, k, n, -1, dtype, seed=11)
    out, (a, qw, sc, ws, c_tmp) = run_ours(prob, m, n, k, dtype, group=-1)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, -1, dtype)
    st = mo.error_stats(t_load(out, dtype), exact, dtype)
    assert st["max_rel"] <= TOL_MAX_REL, st
    if ref_marlin.available():
        ref = ref_marlin.gemm(a, qw, sc, ws, c_tmp, group_size=-1)
        torch.cuda.synchronize()
        st_x = mo.error_stats(t_load(out, dtype), mo.to_f32(t_load(ref, dtype), dtype), dtype)
        assert st_x["max_rel"] <= TOL_MAX_REL, st_x


def test_all_decode_configs_agree():
    """every kernel configuration (force_config) gives the same answer within tolerance, and the same
    configuration is bit-identical run to run (the split-K reduction order is fixed)."""
    b = _b()
    from cpmcu_b200 import capi as L
    m, k, n = 16, 2048, 1536
    prob = mo.make_problem(m, k, n, 128, mo.FP16, seed=5)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16)
    seen = 0
    for cfg in range(1, 200):
        name = L.config_name(cfg)
        if not name:
            break
        if not name.startswith("__half/nt2/") or not (name.endswith("true") or name.endswith("true/tmem")):
            continue
        out1, _ = run_ours(prob, m, n, k, mo.FP16, force_config=cfg)
        out2, _ = run_ours(prob, m, n, k, mo.FP16, force_config=cfg, pdl=True)
        assert torch.equal(out1, out2), f"{name}: not deterministic"
        # the same geometry squeezed into the reference's scratch contract (fewer CTAs or whole columns)
        out3, _ = run_ours(prob, m, n, k, mo.FP16, force_config=cfg, contract=True)
        st = mo.error_stats(out1.cpu().numpy(), exact, mo.FP16)
        st3 = mo.error_stats(out3.cpu().numpy(), exact, mo.FP16)
        assert st["max_rel"] <= TOL_MAX_REL and st3["max_rel"] <= TOL_MAX_REL, (name, st, st3)
        seen += 1
    assert seen >= 5


TMEM_TILES = ["tn4x64/kb8x16", "tn2x64/kb16x16"]   # geometries 7 and 8: 256 columns x 128 k, 128 columns x 256 k


def tmem_config(dtype, m, group, tile=TMEM_TILES[0]):
    """force_config id of the tensor-memory streaming kernel (names ending in '/tmem') for this dtype / width / scale kind"""
    from cpmcu_b200 import capi as L
    nt = 1 if m <= 8 else (2 if m <= 16 else 4)
    want = f"{'__half' if dtype == mo.FP16 else '__nv_bfloat16'}/nt{nt}/{tile}/w8/s5/{'true' if group != -1 else 'false'}/tmem"
    for cfg in range(1, 400):
        if L.config_name(cfg) == want:
            return cfg
    raise AssertionError(want)


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
@pytest.mark.parametrize("group", [128, -1])
@pytest.mark.parametrize("tile", TMEM_TILES)
@pytest.mark.parametrize("m,k,n", [(1, 1024, 512), (7, 2048, 448), (16, 1024, 1536), (20, 512, 256), (32, 4096, 832)])
def test_tensor_memory_streaming_kernel(dtype, group, m, k, n, tile):
    """w4a16_decode_tc_kernel (decode_tc_kernel.cuh): dequantised fragments written from registers into tensor memory
    as the A operand of tcgen05.mma, the token rows as the B operand, the accumulator in tensor memory.  Oracle-correct for
    grouped and channelwise scales, both tile shapes, ragged N (the last tile is partly out of range), split and whole columns
    (contract-sized scratch), bit-identical run to run and under PDL, workspace returned zeroed."""
    prob = mo.make_problem(m, k, n, group, dtype, seed=m * 3 + n)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, group, dtype)
    cfg = tmem_config(dtype, m, group, tile)
    out1, _ = run_ours(prob, m, n, k, dtype, group=group, force_config=cfg)
    out2, _ = run_ours(prob, m, n, k, dtype, group=group, force_config=cfg, pdl=True)
    assert torch.equal(out1, out2)
    st = mo.error_stats(t_load(out1, dtype), exact, dtype)
    assert st["max_rel"] <= TOL_MAX_REL, st
    out3, _ = run_ours(prob, m, n, k, dtype, group=group, force_config=cfg, contract=True)
    st3 = mo.error_stats(t_load(out3, dtype), exact, dtype)
    assert st3["max_rel"] <= TOL_MAX_REL, st3


def test_tensor_memory_streaming_kernel_full_size_and_postops():
    """the tensor-memory kernel at the full gate_up size against the mma.sync streaming kernel (same inputs: fp32
    accumulation in both, only the summation order differs), with the epilogue's post-ops and the q/k/v split write"""
    b = _b()
    from cpmcu_b200 import capi as L
    g = torch.Generator(device=DEV).manual_seed(11)
    for m, k, n in [(16, 4096, 32768), (32, 4096, 4608)]:
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
        sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).half()
        a = torch.randn((m, k), device=DEV, generator=g).half()
        res = torch.randn((m, n), device=DEV, generator=g).half()
        bias = torch.randn((n,), device=DEV, generator=g).half()
        ws, c_tmp = scratch(m, n, k)
        cfg = tmem_config(mo.FP16, m, 128, TMEM_TILES[1])
        assert "tmem" not in L.config_name(1)
        ref = b.gemm(a, qw, sc, ws, c_tmp, residual=res, bias=bias, out_scale=0.25,
                     force_config=next(c for c in range(1, 400) if L.config_name(c) == f"__half/nt{2 if m <= 16 else 4}/tn2x64/kb16x16/w8/s5/true"))
        out = b.gemm(a, qw, sc, ws, c_tmp, residual=res, bias=bias, out_scale=0.25, force_config=cfg)
        torch.cuda.synchronize()
        scale = float(ref.float().abs().mean())
        assert float((out.float() - ref.float()).abs().max()) <= 1e-2 * scale
        assert int(ws.abs().max()) == 0
    qa, qb = 4096, 256
    plain = b.gemm(a, qw, sc, ws, c_tmp, force_config=cfg)
    split = b.gemm(a, qw, sc, ws, c_tmp, force_config=cfg, qkv_split=(qa, qb))
    torch.cuda.synchronize()
    flat = split.reshape(-1)
    assert torch.equal(flat[: m * qa].view(m, qa), plain[:, :qa])
    assert torch.equal(flat[m * qa: m * (qa + qb)].view(m, qb), plain[:, qa:qa + qb])
    assert torch.equal(flat[m * (qa + qb):].view(m, qb), plain[:, qa + qb:])


def test_disabled_tiles_are_rejected():
    """the two single-CTA 256-column tiles failed intermittently in long PDL chains (profiles/r02/r_nsub2_tiles_stress.txt)
    and are refused with a value error instead of being launched"""
    b = _b()
    prob = mo.make_problem(40, 256, 256, 128, mo.FP16, seed=1)
    for cfg in (1001, 1003):
        with pytest.raises(ValueError, match="disabled"):
            run_ours(prob, 40, 256, 256, mo.FP16, force_config=cfg)


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
def test_all_tcgen05_tiles_agree(dtype):
    """every tile shape of the tcgen05 kernel (force_config 1001..) gives the oracle's answer, bit-identical
    run to run, also when forced onto a small M (ragged token tile, UMMA N rounded up to 16)."""
    from cpmcu_b200 import capi as L
    for (m, k, n) in [(40, 256, 192), (300, 512, 448)]:
        prob = mo.make_problem(m, k, n, 128, dtype, seed=m)
        c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, dtype)
        outs = []
        for cfg in (1002, 1004):   # 1001 / 1003 (256 columns per CTA) are disabled: DESIGN.md 6e
            assert L.config_name(cfg).startswith("tc/bn128/")
            out1, _ = run_ours(prob, m, n, k, dtype, force_config=cfg)
            out2, _ = run_ours(prob, m, n, k, dtype, force_config=cfg)
            assert torch.equal(out1, out2), f"{L.config_name(cfg)}: not deterministic"
            st = mo.error_stats(t_load(out1, dtype), exact, dtype)
            assert st["max_rel"] <= TOL_MAX_REL, (L.config_name(cfg), st)
            outs.append(out1)
        # same k order in every tile shape -> identical bits across shapes
        for o in outs[1:]:
            assert torch.equal(outs[0], o)


@pytest.mark.parametrize("m,k,n", [(33, 1024, 448), (64, 2048, 1152), (100, 512, 256), (128, 1024, 640)])
def test_tcgen05_splitk_clusters(m, k, n):
    """mid-M regime: tcgen05 tiles with the K range split over a cluster of 1/2/4/8 CTAs and reduced
    through distributed shared memory in rank order — oracle-correct, bit-identical run to run."""
    from cpmcu_b200 import capi as L
    prob = mo.make_problem(m, k, n, 128, mo.FP16, seed=m + k)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16)
    seen = 0
    for cfg in range(1005, 1013):
        name = L.config_name(cfg)
        assert "splitk" in name
        if "bm64" in name and m > 64:
            continue
        out1, _ = run_ours(prob, m, n, k, mo.FP16, force_config=cfg)
        out2, _ = run_ours(prob, m, n, k, mo.FP16, force_config=cfg, pdl=True)
        assert torch.equal(out1, out2), f"{name} cfg {cfg}: not deterministic"
        st = mo.error_stats(out1.cpu().numpy(), exact, mo.FP16)
        assert st["max_rel"] <= TOL_MAX_REL, (name, cfg, st)
        seen += 1
    assert seen >= 4
    # the heuristic's own choice for this shape is one of them
    assert "tc/" in L.config_name(L.pick_config(m, n, k))


def test_act_order_gather():
    """has_act_order + is_k_full: A's columns are gathered by perm, weights are already group-sorted."""
    b = _b()
    m, k, n = 5, 512, 256
    prob = mo.make_problem(m, k, n, 128, mo.FP16, seed=9)
    perm = np.random.default_rng(2).permutation(k).astype(np.int32)
    a = t_store(prob["a"], mo.FP16)
    qw = torch.from_numpy(prob["qweight"]).to(DEV)
    sc = t_store(prob["scales_marlin"], mo.FP16)
    ws, c_tmp = scratch(m, n, k)
    a_tmp = torch.empty_like(a)
    out = b.gemm(a, qw, sc, ws, c_tmp, perm=torch.from_numpy(perm).to(DEV), a_tmp=a_tmp)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16, perm=perm)
    st = mo.error_stats(out.cpu().numpy(), exact, mo.FP16)
    assert st["max_rel"] <= TOL_MAX_REL, st
    if ref_marlin.available():
        ref = ref_marlin.gemm(a, qw, sc, ws, c_tmp, perm=torch.from_numpy(perm).to(DEV), a_tmp=torch.empty_like(a))
        st_x = mo.error_stats(out.cpu().numpy(), ref.float().cpu().numpy(), mo.FP16)
        assert st_x["max_rel"] <= TOL_MAX_REL, st_x


def test_linear_operator_postops():
    """W4A16GPTQMarlinLinear.prefill: tgt / inplace / bias semantics of w4a16_gptq_marlin_linear.cuh:104-146"""
    b = _b()
    m, k, n = 6, 512, 384
    prob = mo.make_problem(m, k, n, 128, mo.FP16, seed=21)
    rng = np.random.default_rng(4)
    bias = rng.standard_normal(n).astype(np.float16)
    tgt0 = rng.standard_normal((m, n)).astype(np.float16)
    lin = b.W4A16GPTQMarlinLinear(k, n, 128, has_bias=True)
    lin.init_weight_ptr()
    lin.load_to_storage("x.qweight", torch.from_numpy(prob["qweight"]))
    lin.load_to_storage("x.scales", torch.from_numpy(prob["scales_marlin"]))
    lin.load_to_storage("x.bias", torch.from_numpy(bias))
    with pytest.raises(ValueError):
        lin.load_to_storage("x.qzeros", torch.zeros(1))
    with pytest.raises(ValueError):
        b.W4A16GPTQMarlinLinear(k, n, 64)
    a = torch.from_numpy(prob["a"]).to(DEV)
    _, c_tmp = scratch(m, n, k)
    tgt = torch.from_numpy(tgt0).to(DEV)
    lin.prefill(None, m, a, None, c_tmp, tgt=tgt, inplace=True)
    want = mo.linear_forward(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16,
                             bias=bias, tgt=tgt0, inplace=True)
    st = mo.error_stats(tgt.cpu().numpy(), want.astype(np.float64), mo.FP16)
    assert st["max_rel"] <= TOL_MAX_REL, st
    out = lin.prefill(None, m, a, None, c_tmp)            # tgt None -> self.output
    want2 = mo.linear_forward(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16, bias=bias)
    st2 = mo.error_stats(out.cpu().numpy(), want2.astype(np.float64), mo.FP16)
    assert st2["max_rel"] <= TOL_MAX_REL, st2
    assert int(lin.workspace.abs().max()) == 0


def test_linear_operator_postops_tcgen05():
    """the same post-op contract on the tcgen05 path (M > 64)"""
    b = _b()
    m, k, n = 200, 256, 320
    prob = mo.make_problem(m, k, n, 128, mo.FP16, seed=23)
    rng = np.random.default_rng(5)
    bias = rng.standard_normal(n).astype(np.float16)
    tgt0 = rng.standard_normal((m, n)).astype(np.float16)
    lin = b.W4A16GPTQMarlinLinear(k, n, 128, has_bias=True)
    lin.init_weight_ptr()
    lin.load_to_storage("x.qweight", torch.from_numpy(prob["qweight"]))
    lin.load_to_storage("x.scales", torch.from_numpy(prob["scales_marlin"]))
    lin.load_to_storage("x.bias", torch.from_numpy(bias))
    a = torch.from_numpy(prob["a"]).to(DEV)
    tgt = torch.from_numpy(tgt0).to(DEV)
    lin.prefill(None, m, a, None, None, tgt=tgt, inplace=True)      # no c_tmp needed above M = 64
    # the post-ops must be exactly the reference's rounding sequence applied to the kernel's own GEMM
    # output: RN(RN(tgt + gemm) + bias), bit for bit (w4a16_gptq_marlin_linear.cuh:140-145)
    ws, _ = scratch(m, n, k)
    g = b.gemm(a, lin.weight, lin.scales, ws, None)
    want = ((torch.from_numpy(tgt0).to(DEV) + g) + torch.from_numpy(bias).to(DEV))
    assert torch.equal(tgt, want)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, mo.FP16)
    st = mo.error_stats(g.cpu().numpy(), exact, mo.FP16)
    assert st["max_rel"] <= TOL_MAX_REL, st


@pytest.mark.parametrize("tdt", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m,force", [(1, 0), (5, 0), (12, 0), (24, 0), (32, 0), (64, 0), (200, 0),
                                     (700, 0)])
def test_out_scale_postop(tdt, m, force):
    """The layer's residual scaling fused as the last post-op (elementwise_scale on the o_proj / down_proj output,
    w4a16_gptq_marlin_layer.cuh:57,71,79,93; kernel elementwise.cuh:35-41): exactly the reference's rounding
    sequence applied to the kernel's own GEMM output, RN(RN(RN(res + gemm) + bias) * T(scale)), bit for bit,
    on every kernel family (persistent stream-K, tcgen05 split-K and whole-K)."""
    b = _b()
    k, n = 512, 640
    dtype = mo.FP16 if tdt == torch.float16 else mo.BF16
    prob = mo.make_problem(m, k, n, 128, dtype, seed=31 + m)
    a, qw, sc = t_store(prob["a"], dtype), torch.from_numpy(prob["qweight"]).to(DEV), t_store(prob["scales_marlin"], dtype)
    g = torch.Generator(device=DEV).manual_seed(m)
    res = torch.randn((m, n), generator=g, device=DEV, dtype=torch.float32).to(tdt)
    bias = torch.randn((n,), generator=g, device=DEV, dtype=torch.float32).to(tdt)
    scale = 1.4 / np.sqrt(32.0)       # MiniCPM4-8B: scale_depth / sqrt(num_hidden_layers)
    ws, c_tmp = scratch(m, n, k)
    plain = b.gemm(a, qw, sc, ws, c_tmp, force_config=force)
    got = b.gemm(a, qw, sc, ws, c_tmp, residual=res, bias=bias, out_scale=scale, force_config=force)
    only = b.gemm(a, qw, sc, ws, c_tmp, out_scale=scale, force_config=force)
    torch.cuda.synchronize()
    sv = torch.tensor(scale, dtype=torch.float32, device=DEV).to(tdt)
    assert torch.equal(got, ((res + plain) + bias) * sv)
    assert torch.equal(only, plain * sv)
    assert int(ws.abs().max()) == 0
    want = mo.linear_forward(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, dtype, out_scale=scale)
    st = mo.error_stats(t_load(only, dtype), mo.to_f32(want, dtype).astype(np.float64), dtype)
    assert st["max_rel"] <= TOL_MAX_REL, st


def gated_family_config(tdt, m, k):
    """the plain streaming configuration whose arithmetic the fused-activation kernel shares (geometry 0, or 1 when
    K is not a multiple of 128): the mma.sync kernel with the 256-column tile"""
    from cpmcu_b200 import capi as L
    nt = 1 if m <= 8 else (2 if m <= 16 else 4)
    tile = "tn4x64/kb8x16/w8/s5" if k % 128 == 0 else "tn4x64/kb4x16/w8/s8"
    want = f"{'__half' if tdt == torch.float16 else '__nv_bfloat16'}/nt{nt}/{tile}/true"
    return next(c for c in range(1, 400) if L.config_name(c) == want)


@pytest.mark.parametrize("tdt", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m", [1, 5, 8, 9, 16, 24, 32])
@pytest.mark.parametrize("k,inter", [(512, 384), (128, 256), (2048, 1024)])
def test_gated_silu_fused(tdt, m, k, inter):
    """gate_up GEMM + gated_silu_interleaved (activation.cuh:7-18; w4a16_gptq_marlin_ffn.cuh:70-72) in ONE launch:
    against the oracle's two-step pipeline (GEMM rounded to T, then the activation in fp32), and against the
    activation applied to the kernel's own plain gate_up output.  Small shapes: every column tile is split over
    several CTAs (k = 512, 2048) or whole (k = 128), so all reduction paths of the kernel emit gated pairs."""
    b = _b()
    n = 2 * inter
    dtype = mo.FP16 if tdt == torch.float16 else mo.BF16
    prob = mo.make_problem(m, k, n, 128, dtype, seed=41 + m)
    a, qw, sc = t_store(prob["a"], dtype), torch.from_numpy(prob["qweight"]).to(DEV), t_store(prob["scales_marlin"], dtype)
    ws, c_tmp = scratch(m, n, k)
    plain = b.gemm(a, qw, sc, ws, c_tmp, force_config=gated_family_config(tdt, m, k))
    got = b.gemm(a, qw, sc, ws, c_tmp, gated_silu=True)
    torch.cuda.synchronize()
    assert got.shape == (m, inter)
    assert int(ws.abs().max()) == 0
    g, u = plain[:, :inter].float(), plain[:, inter:].float()
    want_own = (g * torch.sigmoid(g) * u)
    d = (got.float() - want_own).abs()
    ulp = 2.0 ** -10 if dtype == mo.FP16 else 2.0 ** -7
    assert bool((d <= 1.5 * ulp * want_own.abs() + 1e-6).all()), float(d.max())   # fast-math exp / divide: last bit
    c, _ = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, dtype)
    want = mo.to_f32(mo.gated_silu_interleaved(c, dtype), dtype).astype(np.float64)
    st = mo.error_stats(t_load(got, dtype), want, dtype)
    assert st["max_rel"] <= 2 * TOL_MAX_REL, st     # two T roundings (GEMM, activation) against the oracle's pipeline


@pytest.mark.parametrize("m", [1, 16])
def test_gated_silu_full_size(m):
    """MiniCPM4-8B gate_up (4096 -> 2 x 16384) with the fused activation, against the activation of the plain output"""
    b = _b()
    k, inter = 4096, 16384
    n = 2 * inter
    g = torch.Generator(device=DEV).manual_seed(3)
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
    sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).half()
    a = (torch.randn((m, k), device=DEV, generator=g) * 0.1).half()
    ws, c_tmp = scratch(m, n, k)
    plain = b.gemm(a, qw, sc, ws, c_tmp, force_config=gated_family_config(torch.float16, m, k))
    got = b.gemm(a, qw, sc, ws, c_tmp, gated_silu=True)
    torch.cuda.synchronize()
    gg, uu = plain[:, :inter].float(), plain[:, inter:].float()
    want = gg * torch.sigmoid(gg) * uu
    d = (got.float() - want).abs()
    assert bool((d <= 1.5 * 2.0 ** -10 * want.abs() + 1e-6).all()), float(d.max())
    assert int(ws.abs().max()) == 0
    with pytest.raises(ValueError):
        b.gemm(torch.zeros((33, k), dtype=torch.float16, device=DEV), qw, sc, ws, c_tmp, gated_silu=True)


@pytest.mark.parametrize("m,k,n", [(64, 16384, 4096), (64, 4096, 32768), (128, 16384, 4096)])
def test_tcgen05_splitk_repeatable_under_load(m, k, n):
    """Regression for the co-resident split-K race: with two split-K CTAs of one launch on an SM,
    about 1.5 % of the launches at M = 64 returned a wrong 16..64-column block (a raw int4 slot was released
    before its shared-memory loads had completed).  Every one of 300 back-to-back launches (PDL on, a streaming
    launch in between) must be bit-identical and within tolerance of the exact result."""
    b = _b()
    g = torch.Generator(device=DEV).manual_seed(k + n)
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
    sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).half()
    a = torch.randn((m, k), device=DEV, generator=g).half()
    a2 = torch.randn((7, k), device=DEV, generator=g).half()
    ws, c_tmp = scratch(m, n, k)
    _, c_tmp7 = scratch(7, n, k)      # the streaming launch in between needs its own split-K scratch size
    if c_tmp7.numel() > c_tmp.numel():
        c_tmp = c_tmp7
    ref = b.gemm(a, qw, sc, ws, c_tmp).clone()
    exact = (a.double() @ b.dequant(qw, sc, k, n).double()).cpu().numpy()
    assert mo.error_stats(ref.cpu().numpy(), exact, mo.FP16)["max_rel"] <= TOL_MAX_REL
    out = torch.empty_like(ref)
    bad = 0
    for i in range(300):
        if i % 3 == 0:
            b.gemm(a2, qw, sc, ws, c_tmp)
        b.gemm(a, qw, sc, ws, c_tmp, out=out, pdl=True)
        bad += 0 if torch.equal(out, ref) else 1
    assert bad == 0, f"{bad}/300 launches differ"


def test_error_convention():
    b = _b()
    a = torch.zeros((1, 128), dtype=torch.float16, device=DEV)
    qw = torch.zeros((8, 2 * 96), dtype=torch.int32, device=DEV)       # N = 96: not a multiple of 64
    sc = torch.zeros((1, 96), dtype=torch.float16, device=DEV)
    ws = torch.zeros(64, dtype=torch.int32, device=DEV)
    ct = torch.zeros(1 << 16, dtype=torch.float32, device=DEV)
    with pytest.raises(ValueError, match="min_thread_n"):
        b.gemm(a, qw, sc, ws, ct)
    with pytest.raises(ValueError, match="group size"):
        b.gemm(a, torch.zeros((8, 128), dtype=torch.int32, device=DEV), sc, ws, ct, group_size=32)


FULL = [  # BASELINE.json config 2 shapes (MiniCPM4-8B) at tree-verify widths: ours vs reference Marlin
    (1, 4096, 4608), (4, 4096, 4096), (16, 4096, 32768), (64, 16384, 4096), (12, 4096, 32768), (60, 4096, 4608),
    # BASELINE.json config 3: chunked prefill (tcgen05 path), incl. the ragged last chunk 106850 % 2048 = 354
    (1024, 4096, 4608), (1024, 16384, 4096), (4096, 4096, 4096), (354, 4096, 32768), (2048, 4096, 32768),
]


@pytest.mark.parametrize("m,k,n", FULL)
def test_full_size_vs_reference_and_properties(m, k, n):
    b = _b()
    g = torch.Generator(device=DEV).manual_seed(m + n)
    qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
    sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).half()
    a = torch.randn((m, k), device=DEV, generator=g).half()
    ws, c_tmp = scratch(m, n, k)
    out = b.gemm(a, qw, sc, ws, c_tmp)
    # property 1: against fp64 matmul with the library's own bit-exact dequant (checked above)
    w = b.dequant(qw, sc, k, n)
    exact = (a.double() @ w.double()).cpu().numpy()
    st = mo.error_stats(out.cpu().numpy(), exact, mo.FP16)
    assert st["max_rel"] <= TOL_MAX_REL, st
    # property 2: linearity in A (exact for power-of-two scaling)
    # (outputs in fp16's subnormal range are excluded: there RN(2x) != 2 RN(x) legitimately)
    out2 = b.gemm((a * 2).half(), qw, sc, ws, c_tmp)
    normal = out.abs() > 2.0 ** -13
    bad = (out2 != (out * 2).half()) & normal
    assert int(bad.sum()) == 0, (int(bad.sum()), float((out2.float() - 2 * out.float()).abs().max()))
    # property 2b: bit-identical run to run (fixed split-K reduction order)
    assert torch.equal(out, b.gemm(a, qw, sc, ws, c_tmp))
    # property 3: zero activations -> zero output; workspace still clean
    outz = b.gemm(torch.zeros_like(a), qw, sc, ws, c_tmp)
    assert float(outz.abs().max()) == 0.0 and int(ws.abs().max()) == 0
    if ref_marlin.available():
        ref = ref_marlin.gemm(a, qw, sc, ws, c_tmp)
        st_x = mo.error_stats(out.cpu().numpy(), ref.float().cpu().numpy(), mo.FP16)
        st_r = mo.error_stats(ref.cpu().numpy(), exact, mo.FP16)
        assert st_x["max_rel"] <= TOL_MAX_REL, st_x
        print(f"\n[{m}x{k}x{n}] ours vs exact {st['max_rel']:.2e}/{st['mean_abs']:.2e}  "
              f"ref vs exact {st_r['max_rel']:.2e}/{st_r['mean_abs']:.2e}  ours vs ref {st_x['max_rel']:.2e}")


def test_split_column_reduction_is_repeatable_under_pdl_chains():
    """The ticket-based split-column reduction (every sharer publishes, the last ticket holder adds in slot order)
    must give the same bits whoever arrives last: 200 back-to-back PDL launches over shapes whose columns are
    split 2..9 ways, alternating widths so that successive launches overlap differently."""
    b = _b()
    g = torch.Generator(device=DEV).manual_seed(77)
    shapes = [(4096, 4608), (4096, 4096), (16384, 4096)]
    probs = []
    for k, n in shapes:
        qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
        sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).half()
        probs.append((k, n, qw, sc))
    for m in (1, 7, 16, 32):
        refs, acts = [], []
        ws, c_tmp = scratch(m, 4608, 16384)
        for k, n, qw, sc in probs:
            a = torch.randn((m, k), device=DEV, generator=g).half()
            acts.append(a)
            refs.append(b.gemm(a, qw, sc, ws, c_tmp).clone())
        outs = [torch.empty_like(r) for r in refs]
        bad = 0
        for it in range(200):
            i = it % 3
            k, n, qw, sc = probs[i]
            b.gemm(acts[i], qw, sc, ws, c_tmp, out=outs[i], pdl=True)
            if it % 7 == 6:
                torch.cuda.synchronize()
                bad += sum(0 if torch.equal(o, r) else 1 for o, r in zip(outs, refs))
        torch.cuda.synchronize()
        assert bad == 0 and int(ws.abs().max()) == 0, (m, bad)


@pytest.mark.parametrize("case", ["16x64", "32x128", "128x64", "256x192", "512x256", "1024x320", "k4096n4096", "k4096n32768"])
def test_repack_gptq_on_device(case):
    """b200_w4a16_repack_gptq (AutoGPTQ int32[K/8][N] -> Marlin int32[K/16][2N]) bit-exact against the REFERENCE
    packer's output (scripts/model_convert/gptq2marlin.py:109-134): the committed golden fixtures produced by
    importing the reference (tests/golden/make_format_golden.py), plus one BASELINE-size tensor against the oracle's
    restatement of the same packer (itself pinned to those fixtures by tests/test_oracle_format.py)."""
    import os
    b = _b()
    from cpmcu_b200 import capi as L
    if case.startswith("k"):
        k, n = (int(x) for x in case[1:].split("n"))
        wq = np.random.default_rng(5).integers(0, 16, size=(k, n), dtype=np.uint8)
        gptq = mo.gptq_pack(wq)
        want = mo.marlin_pack(wq)
    else:
        gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "format_cases.npz"))
        gptq, want = gold[f"gptq_{case}"], gold[f"marlin_{case}"]
    got = L.repack_gptq(torch.from_numpy(np.ascontiguousarray(gptq)).to(DEV))
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy().view(np.uint32), np.ascontiguousarray(want).view(np.uint32))


@pytest.mark.parametrize("tdt", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m", [1, 5, 16, 32, 40, 200, 600])
def test_qkv_split_write(tdt, m):
    """opts.qkv_q_dim / qkv_kv_dim: the fused q | k | v projection written as [q: M x a][k: M x b][v: M x b] — bit for
    bit what the reference's permute launch (src/model/attn.cuh:23-39, called at w4a16_gptq_marlin_attn.cuh:132 when
    num_tokens > 1) makes of the row-major output — on the streaming, tcgen05 split-K and tcgen05 whole-K kernels."""
    b = _b()
    k, qa, qb = 512, 384, 128           # 3 "heads" of 128 + 1 kv head of 128
    n = qa + 2 * qb
    dtype = mo.FP16 if tdt == torch.float16 else mo.BF16
    prob = mo.make_problem(m, k, n, 128, dtype, seed=77 + m)
    a, qw, sc = t_store(prob["a"], dtype), torch.from_numpy(prob["qweight"]).to(DEV), t_store(prob["scales_marlin"], dtype)
    ws, c_tmp = scratch(m, n, k)
    bias = torch.randn((n,), device=DEV).to(tdt)
    plain = b.gemm(a, qw, sc, ws, c_tmp, bias=bias)
    split = torch.empty(m * n, dtype=tdt, device=DEV)
    b.gemm(a, qw, sc, ws, c_tmp, out=split.view(m, n), bias=bias, qkv_split=(qa, qb))
    torch.cuda.synchronize()
    want = torch.cat([plain[:, :qa].reshape(-1), plain[:, qa:qa + qb].reshape(-1), plain[:, qa + qb:].reshape(-1)])
    assert torch.equal(split, want)
    assert int(ws.abs().max()) == 0
    with pytest.raises(ValueError):
        b.gemm(a, qw, sc, ws, c_tmp, qkv_split=(qa, qb + 8))


PAIR_CFG = 1013   # "tc/pair256/bm256": cta_group::2 CTA pairs


@pytest.mark.parametrize("dtype", [mo.FP16, mo.BF16])
@pytest.mark.parametrize("m,k,n", [(256, 256, 256), (40, 256, 192), (300, 512, 448), (354, 1024, 640), (513, 512, 1152),
                                   (1024, 1024, 2048)])
def test_tcgen05_cta_pair_tile(dtype, m, k, n):
    """cta_group::2: a CTA pair per 256 weight columns x 256 tokens (each CTA dequantises 128 columns and stages half of
    the token tile).  Oracle-correct on ragged token tiles (the pair splits the VALID tokens) and on N that is not a
    multiple of 256 (the second CTA's columns are zero-filled / masked), bit-identical run to run and bit-identical
    to the single-CTA tiles (same k order)."""
    from cpmcu_b200 import capi as L
    assert L.config_name(PAIR_CFG) == "tc/pair256/bm256"
    prob = mo.make_problem(m, k, n, 128, dtype, seed=m + n)
    c, exact = mo.gemm(prob["a"], prob["qweight"], prob["scales_marlin"], k, n, 128, dtype)
    out1, _ = run_ours(prob, m, n, k, dtype, force_config=PAIR_CFG)
    out2, _ = run_ours(prob, m, n, k, dtype, force_config=PAIR_CFG, pdl=True)
    assert torch.equal(out1, out2)
    st = mo.error_stats(t_load(out1, dtype), exact, dtype)
    assert st["max_rel"] <= TOL_MAX_REL, st
    single, _ = run_ours(prob, m, n, k, dtype, force_config=1002)
    assert torch.equal(out1, single)


@pytest.mark.parametrize("m", [1024, 354])
def test_tcgen05_pair_tile_repeatable_under_pdl_chains(m):
    """the chunked-prefill chain as the operator runs it (auto-picked tiles: CTA pairs where a wave of pair tiles exists,
    single-CTA tiles elsewhere; PDL on, so every kernel's weight stream and dequantisers start under the previous
    kernel's tail and cluster launches interleave with non-cluster ones): 250 passes over 2 layers x {qkv, o, gate_up,
    down} at the MiniCPM4-8B sizes, every output bit-identical to the single-CTA tile's on every pass.  (The pair tile
    had a race that corrupted a half-stage once in ~2700 launches: profiles/r02/o_pair_tile_race*.txt.)"""
    b = _b()
    from cpmcu_b200 import capi as L
    shapes = [(4096, 4608), (4096, 4096), (4096, 32768), (16384, 4096)]
    picked = [L.config_name(L.pick_config(m, n, k, 128, L.FP16)) for k, n in shapes]
    assert any("pair256" in p for p in picked), picked
    g = torch.Generator(device=DEV).manual_seed(77)
    layers = 2
    w = []
    for k, n in shapes:
        for _ in range(layers):
            qw = torch.randint(-2**31, 2**31 - 1, (k // 16, 2 * n), dtype=torch.int32, device=DEV, generator=g)
            sc = (torch.rand((k // 128, n), device=DEV, generator=g) * 0.018 + 0.002).half()
            a = torch.randn((m, k), device=DEV, generator=g).half()
            w.append((a, qw, sc, torch.zeros(b.workspace_ints(n), dtype=torch.int32, device=DEV)))
    c_tmp = torch.empty(b.reduce_max_m(m) * 32768, dtype=torch.float32, device=DEV)
    refs = [b.gemm(a, qw, sc, ws, c_tmp, force_config=1002) for a, qw, sc, ws in w]
    outs = [torch.empty_like(r) for r in refs]
    torch.cuda.synchronize()
    for rep in range(250):
        for (a, qw, sc, ws), o in zip(w, outs):
            b.gemm(a, qw, sc, ws, c_tmp, out=o, pdl=True)
        torch.cuda.synchronize()
        for i, (o, r) in enumerate(zip(outs, refs)):
            assert torch.equal(o, r), f"pass {rep}, GEMM {i} ({picked[i // layers]}): differs from the single-CTA tile"
            o.fill_(float("nan"))
