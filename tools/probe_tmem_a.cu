// Probe: tcgen05.mma with the A operand in TENSOR MEMORY, written from registers in the mma fragment layout by
// tcgen05.st.16x128b.x2 (the layout the W4A16 dequantiser produces), B (tokens) in shared memory (K-major, SWIZZLE_128B).
// D[n][tok] = sum_k A[n][k] * B[tok][k], 128 x 16 x 16.  Prints the number of wrong elements.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/_probe_tmem_a tools/probe_tmem_a.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "../cpmcu_b200/csrc/common.cuh"
using namespace b200;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ void tmem_st_16x128b_x2(uint32_t taddr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
  asm volatile("tcgen05.st.sync.aligned.16x128b.x2.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ float a_val(int n, int k) { return float(((n * 7 + k * 3) % 13) - 6); }
__device__ __forceinline__ float b_val(int tok, int k) { return float(((tok * 5 + k) % 7) - 3) * 0.5f; }

__global__ void probe(float* d_out, int order) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar = sbase + 4096;
  volatile uint32_t* slot = reinterpret_cast<volatile uint32_t*>(smem + 4096 + 16);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // B: 16 token rows x 64 k (only k < 16 used), K-major, 128-byte rows, SWIZZLE_128B
  for (int i = threadIdx.x; i < 16 * 16; i += blockDim.x) {
    const int r = i / 16, k = i % 16;
    const int off = r * 128 + ((((k * 2) >> 4) ^ (r & 7)) << 4) + ((k * 2) & 15);
    *reinterpret_cast<__half*>(smem + off) = __float2half(b_val(r, k));
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(const_cast<uint32_t*>(slot)), 64);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *slot;
  const uint32_t tm_d = tm, tm_a = tm + 32;
  // A fragments: thread (g, t) of warp w, half h: rows 32w + 16h + {g, g + 8}, k pairs t and t + 4
  const int g = lane >> 2, t = lane & 3;
  for (int h = 0; h < 2; ++h) {
    const int n0 = 32 * warp + 16 * h + g;
    auto pk = [&](int n, int k) {
      __half2 v = __floats2half2_rn(a_val(n, k), a_val(n, k + 1));
      return *reinterpret_cast<uint32_t*>(&v);
    };
    const uint32_t w0 = pk(n0, 2 * t), w1 = pk(n0 + 8, 2 * t), w2 = pk(n0, 2 * t + 8), w3 = pk(n0 + 8, 2 * t + 8);
    const uint32_t taddr = tm_a + (static_cast<uint32_t>(32 * warp + 16 * h) << 16);
    if (order == 0) tmem_st_16x128b_x2(taddr, w0, w1, w2, w3);
    else tmem_st_16x128b_x2(taddr, w0, w2, w1, w3);
  }
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 0) {
    const uint32_t idesc = umma_idesc_f16<__half>(128, 16);
    tc_mma_f16_ts(tm_d, tm_a, umma_desc_k_sw128(sbase), idesc, 0u);
    tc_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  uint32_t v[32];
  tmem_ld_32x32(tm_d + (static_cast<uint32_t>(32 * warp) << 16), v);
  tmem_ld_wait();
  for (int i = 0; i < 16; ++i) d_out[(32 * warp + lane) * 16 + i] = __uint_as_float(v[i]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 64);
}

int main() {
  float* d;
  CK(cudaMalloc(&d, 128 * 16 * 4));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192));
  for (int order = 0; order < 2; ++order) {
    CK(cudaMemset(d, 0, 128 * 16 * 4));
    probe<<<1, 128, 8192>>>(d, order);
    CK(cudaDeviceSynchronize());
    float h[128 * 16];
    CK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
    int bad = 0, first = -1;
    for (int n = 0; n < 128; ++n)
      for (int tok = 0; tok < 16; ++tok) {
        float ref = 0.f;
        for (int k = 0; k < 16; ++k) ref += float(((n * 7 + k * 3) % 13) - 6) * (float(((tok * 5 + k) % 7) - 3) * 0.5f);
        if (h[n * 16 + tok] != ref) {
          if (first < 0) first = n * 16 + tok;
          ++bad;
        }
      }
    printf("register order %d: %d of 2048 elements wrong", order, bad);
    if (first >= 0) printf(" (first: n=%d tok=%d got %g)", first / 16, first % 16, h[first]);
    printf("\n");
  }
  return 0;
}
