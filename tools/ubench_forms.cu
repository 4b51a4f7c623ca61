// Per-instruction-form issue cost on one SMSP (B200): the exact SASS forms of the W4A16 dequant sequence, 8 independent
// chains per thread, 4 warps per SMSP, and each form interleaved 1:1 with LOP3 (does it share an issue port with the ALU pipe?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/_ubench_forms tools/ubench_forms.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)
constexpr int ITERS = 1024;

__device__ __forceinline__ __half2 h2(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
__device__ __forceinline__ uint32_t u32(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ uint32_t lop(uint32_t a, uint32_t m, uint32_t e) {
  uint32_t r;
  asm volatile("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(r) : "r"(a), "n"(0x000f000f), "r"(e));
  (void)m;
  return r;
}

// FORM 0: HADD2 R, R, imm   1: HFMA2 R, R, imm.H0_H0-style (mul 1/16, add -72)   2: HMUL2 R, R, S (full register)
//      3: HMUL2 R, R, S.H0_H0   4: HMUL2 R, R, S.H1_H1   5: LOP3 R, R, imm, R   6: SHF   7: HFMA2 R, R, R, R (3 registers)
template <int FORM, bool WITH_LOP>
__global__ void k(uint32_t* out, long long* cyc, uint32_t seed) {
  uint32_t x[8], z[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { x[i] = 0x3c003c00u + i + seed; z[i] = seed * (i + 3) + threadIdx.x; }
  const uint32_t sc = 0x3c003a00u + (seed & 1), ex = 0x64006400u + (seed & 2);
  const uint32_t sfull = 0x3c003c00u + (seed & 4);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (FORM == 0) x[i] = u32(__hsub2(h2(x[i]), h2(0x64086408u)));
        if (FORM == 1) x[i] = u32(__hfma2(h2(x[i]), h2(0x2c002c00u), h2(0xd480d480u)));
        if (FORM == 2) x[i] = u32(__hmul2(h2(x[i]), h2(sfull)));
        if (FORM == 3) x[i] = u32(__hmul2(h2(x[i]), __low2half2(h2(sc))));
        if (FORM == 4) x[i] = u32(__hmul2(h2(x[i]), __high2half2(h2(sc))));
        if (FORM == 5) x[i] = lop(x[i], 0, ex);
        if (FORM == 6) x[i] = (x[i] >> 8) | 1u;
        if (FORM == 7) x[i] = u32(__hfma2(h2(x[i]), h2(sfull), h2(z[(i + 1) & 7])));
        if (WITH_LOP) z[i] = lop(z[i], 0, ex);
      }
  }
  const long long t1 = clock64();
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) r ^= x[i] ^ z[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int FORM, bool WITH_LOP>
void run(const char* name) {
  uint32_t* out;
  long long* cyc;
  CK(cudaMalloc(&out, 148 * 512 * 4));
  CK(cudaMalloc(&cyc, 148 * 8));
  for (int rep = 0; rep < 2; ++rep) {
    k<FORM, WITH_LOP><<<148, 512>>>(out, cyc, 3u + rep);
    CK(cudaDeviceSynchronize());
  }
  long long h[148];
  CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += h[i];
  avg /= 148;
  // 4 warps per SMSP, each ITERS * 32 instructions of the form (plus 32 LOP3 when interleaved)
  printf("%-42s %5.2f clk per %s per SMSP\n", name, avg / (ITERS * 32.0 * 4), WITH_LOP ? "pair (form + LOP3)" : "instruction");
  CK(cudaFree(out));
  CK(cudaFree(cyc));
}

int main() {
  run<0, false>("HADD2 R, R, imm");
  run<1, false>("HFMA2 R, R, imm, imm");
  run<2, false>("HMUL2 R, R, S");
  run<3, false>("HMUL2 R, R, S.H0_H0");
  run<4, false>("HMUL2 R, R, S.H1_H1");
  run<7, false>("HFMA2 R, R, S, Z (3 registers)");
  run<5, false>("LOP3 R, R, imm, E");
  run<6, false>("SHF + LOP (2 ALU)");
  run<0, true>("HADD2 imm          + LOP3");
  run<1, true>("HFMA2 imm imm      + LOP3");
  run<2, true>("HMUL2 S            + LOP3");
  run<3, true>("HMUL2 S.H0_H0      + LOP3");
  run<7, true>("HFMA2 3-reg        + LOP3");
  return 0;
}
