// TEST INFRASTRUCTURE ONLY (oracle/): a C shim around the UNMODIFIED reference
// Marlin host entry point `gptq_marlin_gemm<T>` (reference:
// src/qgemm/gptq_marlin/gptq_marlin.cuh:11-27, gptq_marlin.cu:42-86).
//
// The reference's hot path is itself CUDA, so "the reference run here" means its
// own sources compiled (where they lie under /root/reference, never copied) for
// compute_100/sm_100 with the reference's flags (setup.py:317-334) and linked
// with this file into oracle/_ref/libmarlin_ref.so by oracle/Makefile. The .so
// travels to the GPU box; it is the parity oracle for the GEMM and the
// "kernel to beat" in bench.py --impl reference.  Nothing in the product
// (cpmcu_b200/) links or loads it.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>

#include "qgemm/gptq_marlin/core/scalar_type.hpp"
#include "qgemm/gptq_marlin/gptq_marlin.cuh"

namespace {
void copy_err(char* dst, int n, const char* msg) {
  if (dst && n > 0) {
    std::strncpy(dst, msg, n - 1);
    dst[n - 1] = 0;
  }
}
}  // namespace

extern "C" {

// Same argument meaning as W4A16GPTQMarlinLinear::prefill's call
// (reference: src/model/w4a16_gptq_marlin/w4a16_gptq_marlin_linear.cuh:115-138).
// dtype: 0 = fp16, 1 = bf16 (codes of cpmcu/llm_w4a16_gptq_marlin.py:12-15).
// returns 0 ok, 1 std::invalid_argument (VALUE_CHECK), 2 std::runtime_error.
int marlin_ref_gemm(const void* a, const int32_t* qweight, const void* scales,
                    const int32_t* g_idx, const int32_t* perm,
                    int32_t* workspace, void* c, int64_t m, int64_t n,
                    int64_t k, int group_size, int dtype, int is_k_full,
                    int has_act_order, int use_fp32_reduce, void* a_tmp,
                    float* c_tmp, cudaStream_t stream, char* err, int errlen) {
  // kU4B8: ScalarType(exponent 0, mantissa 4, signed false, bias 8), the only type
  // CPM.cu's operator constructs (w4a16_gptq_marlin_linear.cuh:39-43).
  const vllm::ScalarType wtype(static_cast<uint8_t>(0), static_cast<uint8_t>(4),
                               false, static_cast<int32_t>(8), false);
  int num_groups = group_size == -1 ? 1 : static_cast<int>(k / group_size);
  try {
    if (dtype == 0) {
      gptq_marlin_gemm<__half>(
          (__half*)a, (int32_t*)qweight, (__half*)scales, nullptr,
          (int32_t*)g_idx, (int32_t*)perm, workspace, wtype, m, n, k,
          is_k_full != 0, false, use_fp32_reduce != 0, (__half*)c, num_groups,
          group_size, static_cast<int>(2 * n), has_act_order != 0, stream,
          (__half*)a_tmp, c_tmp);
    } else if (dtype == 1) {
      gptq_marlin_gemm<__nv_bfloat16>(
          (__nv_bfloat16*)a, (int32_t*)qweight, (__nv_bfloat16*)scales, nullptr,
          (int32_t*)g_idx, (int32_t*)perm, workspace, wtype, m, n, k,
          is_k_full != 0, false, use_fp32_reduce != 0, (__nv_bfloat16*)c,
          num_groups, group_size, static_cast<int>(2 * n), has_act_order != 0,
          stream, (__nv_bfloat16*)a_tmp, c_tmp);
    } else {
      copy_err(err, errlen, "dtype must be 0 (fp16) or 1 (bf16)");
      return 1;
    }
  } catch (const std::invalid_argument& e) {
    copy_err(err, errlen, e.what());
    return 1;
  } catch (const std::exception& e) {
    copy_err(err, errlen, e.what());
    return 2;
  }
  return 0;
}

// reference: gptq_marlin_utils.cu:206-225
int marlin_ref_reduce_max_m(int prob_m, int max_par) {
  return marlin::determine_reduce_max_m(prob_m, max_par);
}

}  // extern "C"
