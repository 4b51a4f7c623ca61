// Issue-rate microbenchmark behind DESIGN.md 4.1 ("what bounds the decode loop"): how many cycles per warp
// instruction one SM sub-partition (SMSP) spends on the instruction mix of the W4A16 streaming loop.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_ubench_issue tools/ubench_issue.cu && tools/_ubench_issue
// Kernels (all operands in registers, no memory traffic, 8 independent chains per thread so latency is hidden):
//   alu      LOP3 only                               -> cycles per ALU-pipe instruction
//   fma      HFMA2 only                              -> cycles per FMA-pipe instruction
//   mix      LOP3 and HFMA2 alternating, independent -> do the two pipes issue concurrently?
//   hmma     mma.sync.m16n8k16 f16 -> f32 only       -> cycles per HMMA
//   word     the loop body per int4 word: 4 LOP3 + SHF + 2 HADD2 + 2 HFMA2 + 4 HMUL2 (+ NT HMMA), exactly the
//            product's Dequant<__half>::word + mma sequence
// printed per warps-per-SMSP (1, 2, 4): cycles per instruction / per word, per SMSP.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int ITERS = 2048;

__device__ __forceinline__ uint32_t lop3_or_and(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm volatile("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t hfma2(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm volatile("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t hmul2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm volatile("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t hsub2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm volatile("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ void hmma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// mode 0: alu, 1: fma, 2: mix, 3: hmma, 4: word (NT = nt), 5: word without HMMA (outputs xor-folded, +1 LOP3 per word)
template <int MODE, int NT>
__global__ void bench(uint32_t* out, long long* cycles, uint32_t seed) {
  uint32_t x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = seed * (threadIdx.x + 1) + i * 0x9e3779b9u;
  uint32_t y[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) y[i] = 0x3c003c00u + i;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const uint32_t m0 = 0x000f000fu, m1 = 0x00f000f0u, magic = 0x64006400u;
  const uint32_t sixteenth = 0x2c002c00u, c72 = 0xd480d480u, c1032 = 0x64086408u;
  const uint32_t scale = 0x211f211fu + (seed & 3);
  uint32_t fold = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = lop3_or_and(x[i], m0 + r, magic);
    } else if (MODE == 1) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 8; ++i) y[i] = hfma2(y[i], sixteenth, c72);
    } else if (MODE == 2) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          x[i] = lop3_or_and(x[i], m0 + r, magic);
          y[i] = hfma2(y[i], sixteenth, c72);
        }
    } else if (MODE == 3) {
      const uint32_t a[4] = {y[0], y[1], y[2], y[3]};
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 4; ++i) hmma(acc[i], a, y[4], y[5]);
    } else {
      // 8 words per iteration
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint32_t q = x[i];
        const uint32_t hi = q >> 8;
        uint32_t w[4];
        w[0] = hsub2(lop3_or_and(q, m0, magic), c1032);
        w[1] = hfma2(lop3_or_and(q, m1, magic), sixteenth, c72);
        w[2] = hsub2(lop3_or_and(hi, m0, magic), c1032);
        w[3] = hfma2(lop3_or_and(hi, m1, magic), sixteenth, c72);
        w[0] = hmul2(w[0], scale);
        w[1] = hmul2(w[1], scale);
        w[2] = hmul2(w[2], scale);
        w[3] = hmul2(w[3], scale);
        if (MODE == 4) {
          const uint32_t a[4] = {w[0], w[2], w[1], w[3]};
#pragma unroll
          for (int t = 0; t < NT; ++t) hmma(acc[(i + t) & 3], a, y[t & 7], y[(t + 1) & 7]);
        } else {
          fold ^= w[0] ^ w[1] ^ w[2] ^ w[3];   // 1 LOP3 (3-input xor x2 -> ptxas fuses to 2 LOP3)
        }
        x[i] = q + 0x01010101u * (it & 1);     // keeps q loop-variant (1 IADD/IMAD per word, uniform part hoisted)
      }
    }
  }
  const long long t1 = clock64();
  uint32_t r = fold;
#pragma unroll
  for (int i = 0; i < 8; ++i) r ^= x[i] ^ y[i];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) r ^= __float_as_uint(acc[i][j]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE, int NT>
void run(const char* name, double units_per_iter, const char* unit) {
  uint32_t* out;
  long long* cyc;
  CK(cudaMalloc(&out, 148 * 1024 * 4));
  CK(cudaMalloc(&cyc, 148 * 8));
  printf("%-34s", name);
  for (int wps : {1, 2, 4}) {   // warps per SMSP: block = 4 * wps warps, one block per SM
    const int threads = 4 * wps * 32;
    bench<MODE, NT><<<148, threads>>>(out, cyc, 12345u);
    CK(cudaDeviceSynchronize());
    bench<MODE, NT><<<148, threads>>>(out, cyc, 777u);
    CK(cudaDeviceSynchronize());
    long long h[148];
    CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += h[i];
    avg /= 148;
    // per SMSP: wps warps each execute ITERS * units_per_iter units in `avg` cycles
    printf("  %d w/SMSP: %6.2f clk/%s", wps, avg / (ITERS * units_per_iter * wps), unit);
  }
  printf("\n");
  CK(cudaFree(out));
  CK(cudaFree(cyc));
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  printf("%s, %d SMs; cycles per unit PER SMSP (lower bound of the time one SM sub-partition needs)\n", p.name, p.multiProcessorCount);
  run<0, 0>("alu: LOP3 only", 32, "inst");
  run<1, 0>("fma: HFMA2 only", 32, "inst");
  run<2, 0>("mix: LOP3 + HFMA2 interleaved", 64, "inst");
  run<3, 0>("hmma: mma.sync m16n8k16 only", 16, "hmma");
  run<5, 0>("word: dequant only (13 + ~3)", 8, "word");
  run<4, 1>("word: dequant + 1 HMMA (M <= 8)", 8, "word");
  run<4, 2>("word: dequant + 2 HMMA (M <= 16)", 8, "word");
  run<4, 4>("word: dequant + 4 HMMA (M <= 32)", 8, "word");
  return 0;
}
