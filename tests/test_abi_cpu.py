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
    # c_tmp: the reference's callers allocate it once for chunk_length tokens (>= 1024): at least
    # reduce_max_m(1024) * 2*hidden floats = 32 MiB (minicpm4_eagle.cuh:134-137); we must fit in that
    for m in (1, 4, 16, 64, 354, 1024):
        for k, n in ((4096, 4608), (4096, 4096), (4096, 32768), (16384, 4096)):
            need = lib.b200_w4a16_reduce_bytes(m, n, k)
            assert 0 <= need <= 16 << 20, (m, k, n, need)
    assert lib.b200_w4a16_reduce_bytes(1, 100, 128) == -1


def test_missing_library_fails_loudly(monkeypatch):
    from cpmcu_b200 import capi
    monkeypatch.setattr(capi, "_lib", None)
    monkeypatch.setenv("B200_W4A16_LIB", "/nonexistent/libb200_w4a16.so")
    with pytest.raises(capi.B200W4A16Error):
        capi.lib()
