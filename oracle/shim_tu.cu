// TEST INFRASTRUCTURE (oracle/): the ONE translation unit a CPM.cu maintainer compiles in place of
// src/qgemm/gptq_marlin/*.cu to route the quantised GEMM to libb200_w4a16.so (see
// include/gptq_marlin_gemm_shim.cuh and INTEGRATION.md).  Built by `make -C oracle cpmcu` against the reference
// tree where it lies; also compiled on its own by tests/test_shim_compile.py.
#include "gptq_marlin_gemm_shim.cuh"
