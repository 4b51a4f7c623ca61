/*
 * gptq_marlin_gemm_shim.cuh — drop-in replacement of CPM.cu's quantised GEMM entry point by libb200_w4a16.so.
 *
 * CPM.cu's operator (src/model/w4a16_gptq_marlin/w4a16_gptq_marlin_linear.cuh:104-146) calls the free function
 * template `gptq_marlin_gemm<T>` DECLARED in src/qgemm/gptq_marlin/gptq_marlin.cuh:11-27 and DEFINED (with explicit
 * instantiations for __half / __nv_bfloat16) in src/qgemm/gptq_marlin/gptq_marlin.cu:42-123; its callers size the
 * fp32 reduce scratch with `marlin::determine_reduce_max_m` (gptq_marlin.cuh:5-9, gptq_marlin_utils.cu:206-225;
 * callers: w4a16_gptq_marlin_layer.cuh:34-37, minicpm4_eagle.cuh:134-137).
 *
 * This header DEFINES both symbols on top of the C ABI of include/b200_w4a16.h.  Integration: compile ONE translation
 * unit that includes it (with -I<CPM.cu>/src -I<this repo>/include and the reference's -DENABLE_DTYPE_* flags) IN PLACE
 * OF src/qgemm/gptq_marlin/*.cu, and link -lb200_w4a16.  No other CPM.cu source changes: the operator, the layers,
 * entry.cu and the pybind surface compile against the reference's own gptq_marlin.cuh untouched.
 * oracle/Makefile (`make -C oracle cpmcu`) does exactly that with the reference tree where it lies and
 * tests/test_shim_compile.py keeps it compiling.
 *
 * Argument mapping (reference name -> C ABI):
 *   a, b_q_weight, b_scales, g_idx, perm, workspace, c, a_tmp, c_tmp, stream    passed through
 *   b_zeros                          must be unused: has_zp == false at the only call site (linear.cuh:128)
 *   b_q_type                         must be kU4B8 (ScalarType(0, 4, false, 8), linear.cuh:39-43); anything else throws
 *   size_m/n/k, group_size           passed through; num_groups, b_q_weight_size1 are redundant (derived) and ignored
 *   is_k_full, has_act_order         passed through (the library supports act-order only with is_k_full, like CPM.cu)
 *   use_fp32_reduce                  always true at the call site (linear.cuh:56); the library reduces in fp32 either way
 * Scratch contract: the library stays inside the reference's c_tmp rule (determine_reduce_max_m(M, 16) x N floats) unless
 * the integrator raises the bound once with b200_w4a16_set_scratch_capacity(bytes) — CPM.cu allocates
 * reduce_max_m(chunk_length) x 2 x intermediate floats (32-128 MiB), far more than the fast path needs.
 * Error convention: the reference's VALUE_CHECK / ERROR_CHECK (src/utils.cuh:54-66): std::invalid_argument /
 * std::runtime_error, which pybind turns into ValueError / RuntimeError.
 */
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <stdexcept>
#include <string>
#include <type_traits>

#include "b200_w4a16.h"
#include "qgemm/gptq_marlin/gptq_marlin.cuh"  // the reference's own declaration (and vllm::ScalarType)

namespace marlin {
// gptq_marlin_utils.cu:206-225 (max_par is 16 at every call site; other values follow the same rule)
int determine_reduce_max_m(int prob_m, int max_par) {
  if (max_par == 16) return b200_w4a16_reduce_max_m(prob_m);
  if (prob_m <= 16) return 16;
  if (prob_m <= 32) return 32;
  if (prob_m <= 48) return 48;
  if (prob_m <= 64) return 64;
  int par = (prob_m + 63) / 64;
  if (par > max_par) par = max_par;
  return 64 * par;
}
}  // namespace marlin

template <typename T>
void gptq_marlin_gemm(T* a, int32_t* b_q_weight, T* b_scales, int32_t* b_zeros, int32_t* g_idx, int32_t* perm,
                      int32_t* workspace, vllm::ScalarType const& b_q_type, int64_t size_m, int64_t size_n,
                      int64_t size_k, bool is_k_full, bool has_zp, bool use_fp32_reduce, T* c, int num_groups,
                      int group_size, int b_q_weight_size1, bool has_act_order, cudaStream_t stream, T* a_tmp,
                      float* c_tmp) {
  static_assert(std::is_same<T, __half>::value || std::is_same<T, __nv_bfloat16>::value,
                "gptq_marlin_gemm: T must be __half or __nv_bfloat16");
  (void)b_zeros;
  (void)num_groups;
  (void)b_q_weight_size1;
  (void)use_fp32_reduce;
  const vllm::ScalarType u4b8(static_cast<uint8_t>(0), static_cast<uint8_t>(4), false, static_cast<int32_t>(8), false);
  if (has_zp || !(b_q_type == u4b8))
    throw std::invalid_argument("b200_w4a16: only the uint4b8 weight type without zero points is supported (the type "
                                "W4A16GPTQMarlinLinear constructs), got " + b_q_type.str());
  const int dtype = std::is_same<T, __half>::value ? B200_W4A16_FP16 : B200_W4A16_BF16;
  const int rc = b200_w4a16_gemm(a, b_q_weight, b_scales, g_idx, perm, workspace, c, size_m, size_n, size_k, group_size,
                                 dtype, is_k_full ? 1 : 0, has_act_order ? 1 : 0, a_tmp, c_tmp, stream);
  if (rc == B200_W4A16_OK) return;
  const std::string msg = b200_w4a16_last_error();
  if (rc == B200_W4A16_EINVAL) throw std::invalid_argument(msg);
  throw std::runtime_error(msg);
}

// the explicit instantiations the reference emits in gptq_marlin.cu:89-123
#ifdef ENABLE_DTYPE_FP16
template void gptq_marlin_gemm<__half>(__half*, int32_t*, __half*, int32_t*, int32_t*, int32_t*, int32_t*,
                                       vllm::ScalarType const&, int64_t, int64_t, int64_t, bool, bool, bool, __half*, int,
                                       int, int, bool, cudaStream_t, __half*, float*);
#endif
#ifdef ENABLE_DTYPE_BF16
template void gptq_marlin_gemm<__nv_bfloat16>(__nv_bfloat16*, int32_t*, __nv_bfloat16*, int32_t*, int32_t*, int32_t*,
                                              int32_t*, vllm::ScalarType const&, int64_t, int64_t, int64_t, bool, bool,
                                              bool, __nv_bfloat16*, int, int, int, bool, cudaStream_t, __nv_bfloat16*,
                                              float*);
#endif
