// Consumer-loop microbenchmark: the streaming kernel's per-stage consumer body (the product's own Dequant<T>::word and
// mma_m16n8k16 from cpmcu_b200/csrc/common.cuh, the same shared-memory reads: 1 LDS.128 of scales + RPW LDS.128 of int4
// words + one ldmatrix per row) on data that already sits in shared memory — no TMA, no barriers, no HBM.
// It answers: how fast can 8 (or 16) consumer warps per SM chew through stages when data is always there?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/_ubench_consumer tools/ubench_consumer.cu
// Output: cycles per int4 word (32 lanes x 4 bytes = 128 B of weights) per SMSP, and the TB/s that rate would sustain
// on 148 SMs at the measured clock, for NT = 1, 2, 4 and for a register cap of 96 (the product's 2-CTA/SM bound) or none.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "../cpmcu_b200/csrc/common.cuh"

using namespace b200;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int STAGES = 5, RPW = 4, ITERS = 400;
__host__ __device__ constexpr int stage_bytes(int warps) { return 2048 + warps * RPW * 512 + 1024; }

// WARPS consumer warps; stage block = [A: 2 KB][W: WARPS * RPW * 512 B][S: 1 KB]
template <int NT, int WARPS, int MINB, int MODE = 0>   // MODE 1: no HMMA (xor fold), 2: B operand loop-invariant (no ldmatrix), 3: HMMA only (A = raw words)
__global__ void __launch_bounds__(WARPS * 32, MINB) consumer(float* out, long long* cycles, int iters) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int STAGE_BYTES = stage_bytes(WARPS);
  const uint32_t sbase = smem_u32(smem);
  for (int i = threadIdx.x; i < STAGES * STAGE_BYTES / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x9e3779b9u * (i + 1) ^ (i << 7);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  float acc[4][NT][4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int n = 0; n < NT; ++n)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[j][n][e] = 0.f;
  const uint32_t q_off = 2048 + threadIdx.x * 16u;
  const uint32_t sc_off = 2048 + WARPS * RPW * 512 + (lane >> 2) * 16;
  const uint32_t a_lane = (lane & 7) * 128 + (((lane >> 3) & 1) << 4);
  uint32_t stg = sbase;
  int s = 0;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const uint4 sv = lds128(stg + sc_off);
    const uint32_t sc[4] = {sv.x, sv.y, sv.z, sv.w};
    uint4 q[RPW];
#pragma unroll
    for (int i = 0; i < RPW; ++i) q[i] = lds128(stg + q_off + i * (WARPS * 512));
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      uint32_t bf[NT][2];
      if (MODE == 2) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) { bf[nt][0] = 0x3c003c00u + lane + nt; bf[nt][1] = 0x38003800u + lane; }
      } else if (NT == 1) {
        ldmatrix_x2(bf[0][0], bf[0][1], stg + a_lane + i * 32);
      } else {
#pragma unroll
        for (int np = 0; np < NT / 2; ++np)
          ldmatrix_x4(bf[2 * np][0], bf[2 * np][1], bf[2 * np + 1][0], bf[2 * np + 1][1], stg + a_lane + i * 32 + np * 512);
      }
      const uint32_t qq[4] = {q[i].x, q[i].y, q[i].z, q[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t w[4];
        if (MODE == 3) {
          w[0] = qq[j]; w[1] = qq[(j + 1) & 3]; w[2] = qq[(j + 2) & 3]; w[3] = qq[(j + 3) & 3];
        } else {
          Dequant<__half>::template word<true>(qq[j], sc[j], w);
        }
        if (MODE == 1) {
          acc[j][0][0] = __uint_as_float(__float_as_uint(acc[j][0][0]) ^ w[0] ^ w[1]);
          acc[j][0][1] = __uint_as_float(__float_as_uint(acc[j][0][1]) ^ w[2] ^ w[3] ^ bf[0][0]);
        } else {
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_m16n8k16<__half>(acc[j][nt], w, bf[nt][0], bf[nt][1]);
        }
      }
    }
    stg += STAGE_BYTES;
    if (++s == STAGES) {
      s = 0;
      stg = sbase;
    }
  }
  const long long t1 = clock64();
  float r = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int n = 0; n < NT; ++n)
#pragma unroll
      for (int e = 0; e < 4; ++e) r += acc[j][n][e];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int NT, int WARPS, int MINB, int MODE = 0>
void run(const char* name, double ghz) {
  float* out;
  long long* cyc;
  CK(cudaMalloc(&out, 148 * 1024 * 4));
  CK(cudaMalloc(&cyc, 148 * 8));
  auto k = consumer<NT, WARPS, MINB, MODE>;
  constexpr int STAGE_BYTES = stage_bytes(WARPS);
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, STAGES * STAGE_BYTES));
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, k));
  for (int rep = 0; rep < 2; ++rep) {
    k<<<148, WARPS * 32, STAGES * STAGE_BYTES>>>(out, cyc, ITERS);
    CK(cudaDeviceSynchronize());
  }
  long long h[148];
  CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += h[i];
  avg /= 148;
  const double words_per_smsp = double(ITERS) * RPW * 4 * (WARPS / 4);   // warp-words each SMSP's warps process
  const double clk = avg / words_per_smsp;
  printf("%-44s regs %3d  %6.2f clk/word/SMSP  -> %5.2f TB/s on 148 SMs at %.3f GHz\n", name, fa.numRegs, clk,
         128.0 * 592 * ghz / clk / 1e3, ghz);
  CK(cudaFree(out));
  CK(cudaFree(cyc));
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  const double ghz = p.clockRate / 1e6;
  printf("%s: consumer loop on shared-memory-resident stages (no TMA / barriers / HBM)\n", p.name);
  run<1, 8, 2>("NT=1 (M<=8),  8 warps, regs capped (2 CTA/SM)", ghz);
  run<1, 8, 1>("NT=1 (M<=8),  8 warps, regs free", ghz);
  run<1, 16, 1>("NT=1 (M<=8), 16 warps", ghz);
  run<2, 8, 2>("NT=2 (M<=16), 8 warps, regs capped", ghz);
  run<2, 8, 1>("NT=2 (M<=16), 8 warps, regs free", ghz);
  run<2, 16, 1>("NT=2 (M<=16), 16 warps", ghz);
  run<4, 8, 1>("NT=4 (M<=32), 8 warps, regs free", ghz);
  run<4, 16, 1>("NT=4 (M<=32), 16 warps", ghz);
  run<1, 8, 1, 1>("NT=1, 8 warps: dequant only (no HMMA)", ghz);
  run<1, 16, 1, 1>("NT=1, 16 warps: dequant only (no HMMA)", ghz);
  run<1, 8, 1, 2>("NT=1, 8 warps: B operand loop-invariant", ghz);
  run<2, 8, 1, 2>("NT=2, 8 warps: B operand loop-invariant", ghz);
  run<1, 8, 1, 3>("NT=1, 8 warps: HMMA only (A = raw words)", ghz);
  run<2, 8, 1, 3>("NT=2, 8 warps: HMMA only", ghz);
  run<4, 8, 1, 3>("NT=4, 8 warps: HMMA only", ghz);
  run<4, 16, 1, 3>("NT=4, 16 warps: HMMA only", ghz);
  return 0;
}
