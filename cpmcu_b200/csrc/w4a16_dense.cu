// Host side of the un-quantised streaming linear (lm_head / FR-Spec head / EAGLE fc at M <= 64):
// argument checks with the reference's error convention, TMA descriptors, persistent PDL launch.
// (reference: linear<T>() and Linear<T>/LMHead<T>::prefill, src/model/linear.cuh:8-37,76-105)
#include <cstring>

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "../../include/b200_w4a16.h"
#include "dense_kernel.cuh"
#include "host_common.h"

namespace b200 {
namespace {

struct DenseCall {
  const void* a;
  const void* w;
  const void* bias;
  void* out;
  int m, n, k;
  int inplace;
  float scale;
  bool pdl;
  cudaStream_t stream;
};

template <typename T>
uint32_t scale_pair(float v);
template <>
uint32_t scale_pair<__half>(float v) {
  const __half h = __float2half_rn(v);  // T(v), as elementwise_scale_kernel rounds it (elementwise.cuh:39)
  uint16_t u;
  std::memcpy(&u, &h, 2);
  return static_cast<uint32_t>(u) | (static_cast<uint32_t>(u) << 16);
}
template <>
uint32_t scale_pair<__nv_bfloat16>(float v) {
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  uint16_t u;
  std::memcpy(&u, &h, 2);
  return static_cast<uint32_t>(u) | (static_cast<uint32_t>(u) << 16);
}

template <typename T, int NT, int TR, int KBLK, int STAGES, bool SCALED>
int launch_dense(const DenseCall& g) {
  using Cfg = DenseCfg<T, NT, TR, KBLK, STAGES>;
  auto kern = dense_stream_kernel<T, NT, TR, KBLK, STAGES, SCALED>;
  static bool attr_set = false;  // benign race: idempotent
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess)
      return fail(B200_W4A16_ECUDA, "cudaFuncSetAttribute(smem=%d): %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
    attr_set = true;
  }
  CUtensorMap tm_w, tm_a;
  {  // weight: T [N][K] (torch nn.Linear layout, `transposed = true` in the reference), [TR x 64] boxes
    const uint64_t dims[2] = {static_cast<uint64_t>(g.k), static_cast<uint64_t>(g.n)};
    const uint64_t strides[1] = {static_cast<uint64_t>(g.k) * 2};
    const uint32_t box[2] = {64, TR};
    if (int rc = get_tensor_map(&tm_w, g.w, Elem<T>::kTmaType, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))
      return rc;
  }
  {  // activations: T [M][K], rows past M zero-filled
    const uint64_t dims[2] = {static_cast<uint64_t>(g.k), static_cast<uint64_t>(g.m)};
    const uint64_t strides[1] = {static_cast<uint64_t>(g.k) * 2};
    const uint32_t box[2] = {64, NT * 8};
    if (int rc = get_tensor_map(&tm_a, g.a, Elem<T>::kTmaType, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))
      return rc;
  }
  DenseArgs args;
  args.out = g.out;
  args.bias = g.bias;
  args.m = g.m;
  args.n = g.n;
  args.k = g.k;
  args.tiles = (g.n + TR - 1) / TR;
  args.kblks = (g.k + Cfg::KB - 1) / Cfg::KB;
  args.inplace = g.inplace;
  args.scale2 = scale_pair<T>(g.scale);
  const int sms = sm_count();
  const int grid = args.tiles < sms ? args.tiles : sms;

  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = g.stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g.pdl ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tm_w, tm_a, args);
  if (e != cudaSuccess) return fail(B200_W4A16_ECUDA, "dense kernel launch: %s", cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

template <typename T, bool SCALED>
int dispatch_nt(const DenseCall& g) {
  // 64-row tiles while the token block is small; 128 / 256-row tiles beyond, so that the activation boxes
  // re-read from L2 every stage stay at most a quarter of the bytes of the weight boxes
  // few rows (EAGLE fc1 / fc2, 4096 outputs): 32-row tiles, so that more SMs than N / 64 take part
  const bool small_n = (g.n + 63) / 64 < 2 * sm_count();
  if (g.m <= 8) return small_n ? launch_dense<T, 1, 32, 128, 8, SCALED>(g) : launch_dense<T, 1, 64, 128, 8, SCALED>(g);
  if (g.m <= 16) return small_n ? launch_dense<T, 2, 32, 128, 8, SCALED>(g) : launch_dense<T, 2, 64, 128, 8, SCALED>(g);
  if (g.m <= 32) return launch_dense<T, 4, 128, 128, 4, SCALED>(g);
  return launch_dense<T, 8, 256, 64, 4, SCALED>(g);
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int b200_w4a16_dense_linear_forward(const void* input, const void* weight, const void* bias,
                                               void* output, int inplace, int64_t num_tokens, int64_t dim_out,
                                               int64_t dim_in, float input_scale, int dtype, int pdl,
                                               void* stream) {
  if (!input || !weight || !output) return fail(B200_W4A16_EINVAL, "dense_linear_forward: null pointer");
  if (!(num_tokens > 0 && dim_in > 0 && dim_out > 0))
    return fail(B200_W4A16_EINVAL, "dense_linear_forward: sizes must be positive (M=%lld, N=%lld, K=%lld)",
                (long long)num_tokens, (long long)dim_out, (long long)dim_in);
  if (num_tokens > 64)
    return fail(B200_W4A16_EINVAL,
                "dense_linear_forward streams the weight matrix for num_tokens <= 64 (decode / tree verify, the "
                "reference's lm_head buffer size, src/model/model.cuh:116); got %lld",
                (long long)num_tokens);
  if (dim_in % 8) return fail(B200_W4A16_EINVAL, "dense_linear_forward: dim_in %% 8 != 0 (16-byte rows), got %lld", (long long)dim_in);
  if (dim_in > (1ll << 30) || dim_out > (1ll << 30)) return fail(B200_W4A16_EINVAL, "dense_linear_forward: dimension too large");
  if (dtype != 0 && dtype != 1) return fail(B200_W4A16_EINVAL, "dtype must be 0 (fp16) or 1 (bf16)");
  if ((reinterpret_cast<uintptr_t>(input) | reinterpret_cast<uintptr_t>(weight)) & 15)
    return fail(B200_W4A16_EINVAL, "dense_linear_forward: input and weight must be 16-byte aligned");
  if (reinterpret_cast<uintptr_t>(output) & 7) return fail(B200_W4A16_EINVAL, "dense_linear_forward: output must be 8-byte aligned");
  DenseCall g;
  g.a = input;
  g.w = weight;
  g.bias = bias;
  g.out = output;
  g.m = static_cast<int>(num_tokens);
  g.n = static_cast<int>(dim_out);
  g.k = static_cast<int>(dim_in);
  g.inplace = inplace ? 1 : 0;
  g.scale = input_scale;
  g.pdl = pdl < 0 ? g_default_pdl.load(std::memory_order_relaxed) != 0 : pdl != 0;
  g.stream = reinterpret_cast<cudaStream_t>(stream);
  const bool scaled = input_scale != 1.0f;
  if (dtype == 0) return scaled ? dispatch_nt<__half, true>(g) : dispatch_nt<__half, false>(g);
  return scaled ? dispatch_nt<__nv_bfloat16, true>(g) : dispatch_nt<__nv_bfloat16, false>(g);
}
