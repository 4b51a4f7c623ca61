"""CPU: the C-ABI library loads, exports every symbol include/b200_w4a16.h declares, and its host-only
helpers agree with the reference's sizing rules.  No compute calls (no GPU here)."""
import os
import re

import pytest

from oracle import marlin_oracle as mo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_w4a16.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_w4a16_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from cpmcu_b200 import capi
    if not os.path.exists(capi.library_path()):
        import __graft_entry__ as g
        g.build()
    return capi.lib()


def test_header_symbols_all_exported_and_bound(lib):
    from cpmcu_b200 import capi
    names = declared_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200_w4a16.h but not exported"
        assert n in capi.SYMBOLS, f"{n} has no ctypes signature in cpmcu_b200/capi.py"
    assert set(capi.SYMBOLS) <= set(names)


def test_host_helpers_match_reference_sizing(lib):
    # determine_reduce_max_m (gptq_marlin_utils.cu:206-225), workspace (w4a16_gptq_marlin_linear.cuh:76)
    for m in (1, 8, 16, 17, 32, 33, 48, 49, 64, 65, 354, 1024, 4096):
        assert lib.b200_w4a16_reduce_max_m(m) == mo.determine_reduce_max_m(m)
    assert lib.b200_w4a16_workspace_ints(4096) == 64 * 16
    assert lib.b200_w4a16_version().decode().startswith("b200-w4a16")
    assert lib.b200_w4a16_reduce_bytes(1, 100, 128) == -1


SHAPES_8B = ((4096, 4608), (4096, 4096), (4096, 32768), (16384, 4096))
SHAPES_70B = ((8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192))
# tensor-parallel shards of the 70B shapes (column-parallel N/tp, row-parallel K/tp) at tp = 2, 4, 8
SHARDS_70B = tuple((k, n // tp) for tp in (2, 4, 8) for k, n in (SHAPES_70B[0], SHAPES_70B[2])) + \
    tuple((k // tp, n) for tp in (2, 4, 8) for k, n in (SHAPES_70B[1], SHAPES_70B[3]))


def test_scratch_stays_inside_the_reference_contract(lib):
    """c_tmp: the reference's rule is determine_reduce_max_m(M, 16) x N floats (gptq_marlin_utils.cu:206-225;
    callers: w4a16_gptq_marlin_layer.cuh:34-37).  What the library touches without a declared capacity must fit in
    it for EVERY M; the fast path's need is reported separately and fits CPM.cu's real allocations
    (>= reduce_max_m(1024) * 2*hidden floats = 32 MiB, minicpm4_eagle.cuh:134-137)."""
    for k, n in SHAPES_8B + SHAPES_70B + SHARDS_70B + ((128, 64), (640, 448), (1152, 1024)):
        for m in list(range(1, 34)) + [48, 64, 65, 354, 1024, 4096]:
            need = lib.b200_w4a16_reduce_bytes(m, n, k)
            assert 0 <= need <= lib.b200_w4a16_reduce_max_m(m) * n * 4, (m, k, n, need)
            fast = lib.b200_w4a16_reduce_bytes_fast(m, n, k)
            assert 0 <= fast <= 32 << 20, (m, k, n, fast)
    prev = lib.b200_w4a16_set_scratch_capacity(1 << 27)
    assert prev == 0 and lib.b200_w4a16_set_scratch_capacity(0) == 1 << 27


def test_missing_library_fails_loudly(monkeypatch):
    from cpmcu_b200 import capi
    monkeypatch.setattr(capi, "_lib", None)
    monkeypatch.setenv("B200_W4A16_LIB", "/nonexistent/libb200_w4a16.so")
    with pytest.raises(capi.B200W4A16Error):
        capi.lib()


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum(lib):
    """b200_w4a16_gemm_static_scratch checks M and N against the static array's limits before it touches CUDA, so the
    rejection (value error + message) can be exercised without a GPU"""
    import ctypes
    null = ctypes.c_void_p(0)
    assert lib.b200_w4a16_gemm_static_scratch(null, null, null, null, 33, 512, 1024, 128, 0, null) == 1
    assert b"exceeds the maximum of 32" in lib.b200_w4a16_last_error()
    assert lib.b200_w4a16_gemm_static_scratch(null, null, null, null, 1, 65536 + 64, 1024, 128, 0, null) == 1
    assert b"exceeds the maximum of 65536" in lib.b200_w4a16_last_error()


def test_configuration_names_of_the_kernel_families(lib):
    """ids are stable within a build and every kernel family is reachable by name: the mma.sync streaming kernel, the
    tensor-memory one (names ending in /tmem), the tcgen05 tiles incl. the CTA pair; the withdrawn tiles keep their ids"""
    from cpmcu_b200 import capi
    names = [capi.config_name(c) or "" for c in list(range(1, 200)) + list(range(1001, 1014))]
    assert "__half/nt1/tn4x64/kb8x16/w8/s5/true" in names
    assert "__nv_bfloat16/nt4/tn2x64/kb16x16/w8/s5/true/tmem" in names
    assert "__half/nt2/tn4x64/kb8x16/w8/s5/false/tmem" in names
    assert capi.config_name(1001) == "tc/bn256/bm256" and capi.config_name(1003) == "tc/bn256/bm128"
    assert capi.config_name(1002) == "tc/bn128/bm256" and capi.config_name(1013) == "tc/pair256/bm256"
    # what the heuristic takes for the decode widths (host-only planning: no GPU needed)
    assert capi.config_name(capi.pick_config(1, 32768, 4096)) == "__half/nt1/tn4x64/kb8x16/w8/s5/true"
    assert capi.config_name(capi.pick_config(16, 32768, 4096)) == "__half/nt2/tn2x64/kb16x16/w8/s5/true"
    assert capi.config_name(capi.pick_config(24, 32768, 4096)) == "__half/nt4/tn2x64/kb16x16/w8/s5/true/tmem"
