"""cpmcu_b200 — B200-native W4A16 GPTQ-Marlin linear: the quantised-linear hot path of
jk3456a/CPM.cu (cpmcu/llm_w4a16_gptq_marlin.py -> src/entry.cu -> src/model/w4a16_gptq_marlin ->
src/qgemm/gptq_marlin) rebuilt as hand-written sm_100a CUDA behind a C ABI (include/b200_w4a16.h).

The package holds only what that path needs:
  csrc/            CUDA kernels + the C ABI            -> libb200_w4a16.so (built in-tree)
  capi.py          ctypes binding of the C ABI (fails loudly when the library is missing)
  linear.py        W4A16GPTQMarlinLinear (+ the un-quantised Linear / LMHead of the same model at decode
                   widths): host-side mirrors of the reference operators
"""
from .capi import (  # noqa: F401
    B200W4A16Error,
    FP16,
    BF16,
    gemm,
    dequant,
    dense_linear,
    lib,
    library_path,
    reduce_bytes,
    reduce_bytes_fast,
    set_scratch_capacity,
    gemm_static_scratch,
    reduce_max_m,
    workspace_ints,
    launch_count,
    set_pdl,
)
from .linear import W4A16GPTQMarlinLinear, Linear, LMHead  # noqa: F401
