/* CPU oracle for the W4A16 GPTQ-Marlin linear, plain C.      *** TEST INFRASTRUCTURE ONLY ***
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library (oracle/_ref/liboracle_c.so).  The product (cpmcu_b200/) never does.
 *
 * A second, independent restatement of the reference algorithm (the first is marlin_oracle.py);
 * tests/test_oracle_c.py checks the two against each other and against the golden vectors made
 * by the reference's own packer.  file:line are relative to the reference tree (CPM.cu @ 797eb93):
 *   layout of one 512-byte block (16 k x 64 n)      scripts/model_convert/gptq2marlin.py:35-61, 109-134
 *   scale permutation                               scripts/model_convert/gptq2marlin.py:54-60, 99-107
 *   q - 8 exact, then ONE rounding: W = RN_T((q-8)*s) src/qgemm/gptq_marlin/marlin_device_ops.cuh:91-137, 294-303
 *   fp32 accumulation, RN to T at the end           src/qgemm/gptq_marlin/marlin_kernel_impl.cuh:944-956
 *   channelwise: scale the ROUNDED sum in T         src/qgemm/gptq_marlin/marlin_kernel_impl.cuh:957-963
 * Parallelism: OpenMP over 64-column chunks (independent outputs; each output is summed in
 * ascending k by one thread, so the result does not depend on the thread count).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- fp16 / bf16 <-> fp32, round-to-nearest-even, in portable integer arithmetic ---- */
static inline float bits_f32(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f32_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static inline float half_to_f32(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  const uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ffu;
  if (exp == 0) {                       /* zero / subnormal: man * 2^-24 (exact in fp32) */
    const float v = (float)man * 5.9604644775390625e-08f;
    return sign ? -v : v;
  }
  if (exp == 31) return bits_f32(sign | 0x7f800000u | (man << 13));
  return bits_f32(sign | ((exp + 112) << 23) | (man << 13));
}

static inline uint16_t f32_to_half(float f) {
  const uint32_t u = f32_bits(f);
  const uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
  const uint32_t a = u & 0x7fffffffu;
  if (a >= 0x7f800000u) return (uint16_t)(sign | 0x7c00u | ((a > 0x7f800000u) ? 0x200u : 0));
  if (a >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);          /* >= 65520 rounds to inf */
  if (a < 0x38800000u) {                                            /* below 2^-14: subnormal half */
    /* value * 2^24 rounded to nearest-even integer; rintf uses the current mode (RN) */
    const float scaled = bits_f32(a) * 16777216.0f;
    const uint32_t m = (uint32_t)rintf(scaled);
    return (uint16_t)(sign | m);                                    /* m == 0x400 is the smallest normal */
  }
  uint32_t r = a - 0x38000000u;                                     /* rebias exponent 127 -> 15 */
  const uint32_t lsb = (r >> 13) & 1u;
  r += 0x0fffu + lsb;
  return (uint16_t)(sign | (r >> 13));
}

static inline float bf16_to_f32(uint16_t b) { return bits_f32((uint32_t)b << 16); }
static inline uint16_t f32_to_bf16(float f) {
  const uint32_t u = f32_bits(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  return (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
}

static inline float to_f32(uint16_t v, int dtype) { return dtype == 0 ? half_to_f32(v) : bf16_to_f32(v); }
static inline uint16_t from_f32(float f, int dtype) { return dtype == 0 ? f32_to_half(f) : f32_to_bf16(f); }

/* position of column c (0..63) inside its 64-column chunk of a grouped scale row:
 * SCALE_PERM[8*i + j] = i + 8*j  (gptq2marlin.py:55-57) */
static inline int grouped_scale_pos(int c) { return 8 * (c & 7) + (c >> 3); }
/* position of column c (0..31) inside its 32-column block of a channelwise scale row:
 * SCALE_PERM_SINGLE[8*i + jj] = 2*i + {0,1,8,9,16,17,24,25}[jj]  (gptq2marlin.py:58-60) */
static inline int chan_scale_pos(int c) {
  const int i = (c & 7) >> 1, jj = ((c >> 3) << 1) | (c & 1);
  return 8 * i + jj;
}

/* Decode one 512-byte block (k16 row r, chunk ch) into q[16][64] nibbles (0..15).
 * Word (lane, j) nibble s holds row 2t + (s&1 ? 8 : 0) + (s&4 ? 1 : 0), col 16j + g + (s&2 ? 8 : 0),
 * g = lane/4, t = lane%4: the m16n8k16 fragment order with the [0,2,4,6,1,3,5,7] interleave. */
static void decode_block(const uint32_t* blk, uint8_t q[16][64]) {
  for (int lane = 0; lane < 32; ++lane) {
    const int g = lane >> 2, t = lane & 3;
    for (int j = 0; j < 4; ++j) {
      const uint32_t w = blk[lane * 4 + j];
      for (int s = 0; s < 8; ++s) {
        const int kk = 2 * t + ((s & 1) ? 8 : 0) + ((s & 4) ? 1 : 0);
        const int nn = 16 * j + g + ((s & 2) ? 8 : 0);
        q[kk][nn] = (uint8_t)((w >> (4 * s)) & 0xfu);
      }
    }
  }
}

/* W[K][N] in T: RN_T((q-8) * s) (grouped) or T(q-8) (channelwise: the scale goes on the output). */
int marlin_oracle_dequant(const int32_t* qweight, const uint16_t* scales, uint16_t* w_out, int64_t n,
                          int64_t k, int group_size, int dtype) {
  if (n <= 0 || k <= 0 || n % 64 || k % 16 || (group_size != -1 && (group_size <= 0 || k % group_size)))
    return 1;
  const int64_t chunks = n / 64;
#pragma omp parallel for schedule(static)
  for (int64_t ch = 0; ch < chunks; ++ch) {
    uint8_t q[16][64];
    for (int64_t r = 0; r < k / 16; ++r) {
      decode_block((const uint32_t*)qweight + (r * chunks + ch) * 128, q);
      for (int kk = 0; kk < 16; ++kk)
        for (int c = 0; c < 64; ++c) {
          float v = (float)((int)q[kk][c] - 8);
          if (group_size != -1) {
            const int64_t grp = (r * 16 + kk) / group_size;
            v *= to_f32(scales[grp * n + ch * 64 + grouped_scale_pos(c)], dtype); /* exact in fp32 */
          }
          w_out[(r * 16 + kk) * n + ch * 64 + c] = from_f32(v, dtype);
        }
    }
  }
  return 0;
}

/* C[M][N] = RN_T( sum_k fp32(A[m][k]) * fp32(W[k][n]) ), W as above; optional act-order gather of A's
 * columns (perm != NULL: a_tmp[m][k] = A[m][perm[k]], gptq_marlin_utils.cu:28-77). */
int marlin_oracle_gemm(const uint16_t* a, const int32_t* qweight, const uint16_t* scales,
                       const int32_t* perm, uint16_t* c, int64_t m, int64_t n, int64_t k,
                       int group_size, int dtype) {
  if (m <= 0 || n <= 0 || k <= 0 || n % 64 || k % 16 || (group_size != -1 && (group_size <= 0 || k % group_size)))
    return 1;
  float* af = (float*)malloc(sizeof(float) * (size_t)m * (size_t)k);
  if (!af) return 2;
  for (int64_t i = 0; i < m; ++i)
    for (int64_t j = 0; j < k; ++j) af[i * k + j] = to_f32(a[i * k + (perm ? perm[j] : j)], dtype);
  const int64_t chunks = n / 64;
  int rc = 0;
#pragma omp parallel
  {
    float* acc = (float*)malloc(sizeof(float) * (size_t)m * 64);
    float w[16][64];
    uint8_t q[16][64];
    if (!acc) {
#pragma omp atomic write
      rc = 2;
    }
#pragma omp for schedule(dynamic, 1)
    for (int64_t ch = 0; ch < chunks; ++ch) {
      if (!acc) continue;
      memset(acc, 0, sizeof(float) * (size_t)m * 64);
      for (int64_t r = 0; r < k / 16; ++r) {
        decode_block((const uint32_t*)qweight + (r * chunks + ch) * 128, q);
        for (int kk = 0; kk < 16; ++kk) {
          const int64_t grp = group_size == -1 ? 0 : (r * 16 + kk) / group_size;
          for (int cc = 0; cc < 64; ++cc) {
            float v = (float)((int)q[kk][cc] - 8);
            if (group_size != -1) {
              v *= to_f32(scales[grp * n + ch * 64 + grouped_scale_pos(cc)], dtype);
              v = to_f32(from_f32(v, dtype), dtype);   /* the single rounding of the weight path */
            }
            w[kk][cc] = v;
          }
        }
        for (int64_t i = 0; i < m; ++i) {
          float* ai = acc + i * 64;
          const float* arow = af + i * k + r * 16;
          for (int kk = 0; kk < 16; ++kk) {
            const float av = arow[kk];
            for (int cc = 0; cc < 64; ++cc) ai[cc] += av * w[kk][cc];
          }
        }
      }
      for (int64_t i = 0; i < m; ++i)
        for (int cc = 0; cc < 64; ++cc) {
          const int64_t col = ch * 64 + cc;
          uint16_t o = from_f32(acc[i * 64 + cc], dtype);
          if (group_size == -1) {
            const uint16_t s = scales[(col & ~31) | chan_scale_pos((int)(col & 31))];
            o = from_f32(to_f32(o, dtype) * to_f32(s, dtype), dtype);
          }
          c[i * n + col] = o;
        }
    }
    free(acc);
  }
  free(af);
  return rc;
}

int marlin_oracle_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
