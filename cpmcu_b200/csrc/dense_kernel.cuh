// Un-quantised fp16 / bf16 linear at decode / tree-verify widths (M <= 64): the lm_head and the
// FR-Spec / EAGLE heads of the W4A16 model (reference: Linear<T>::prefill and LMHead<T>::prefill,
// src/model/linear.cuh:8-37,76-105 — a cuBLAS GemmEx there).  At these widths the op is a pure HBM
// stream of the weight matrix (lm_head: 73448 x 4096 fp16 = 602 MB per verify step),
// so it reuses the structure of the W4A16 streaming kernel without the dequantiser:
//
//   * grid = one persistent CTA per SM; the row tiles (TR weight rows x whole K) are split evenly and
//     contiguously over the CTAs; no split-K, so no cross-CTA reduction.
//   * 1 TMA producer lane feeds an NS-deep ring: per stage KC boxes [TR x 64] of W (128B swizzle,
//     evict-first) and KC boxes [MT x 64] of the activations (rows past M zero-filled, evict-last; they
//     come out of L2).  W boxes of the first NS stages are requested BEFORE griddepcontrol.wait (PDL).
//   * 8 consumer warps: warp = (16-row tile of W, k-slice of the stage).  Per k16 step: ldmatrix.x4 of W as
//     the 16-row MMA operand, ldmatrix of the tokens as the 8-column operand (times the head scale when
//     SCALED, rounded to T like the reference's elementwise_scale, src/model/elementwise.cuh:35-41),
//     mma.sync m16n8k16 with fp32 accumulators (CUBLAS_COMPUTE_32F in the reference).
//   * end of a row tile: the k-slices add their partials in a fixed order in a shared fp32 tile, then
//     all consumer threads write T with coalesced 8-byte stores: out = RN_T(acc [+ old out when
//     inplace: cuBLAS beta = 1, one rounding]); bias added afterwards in T (batched_add).  The TMA ring
//     keeps streaming during the epilogue; the consumers are far from issue-bound and catch up.
#pragma once
#include "common.cuh"

namespace b200 {

struct DenseArgs {
  void* out;           // T[M][N]
  const void* bias;    // T[N] or null
  int m, n, k;
  int tiles;           // ceil(N / TR)
  int kblks;           // ceil(K / KB)
  int inplace;         // out += result
  uint32_t scale2;     // T(head_scale) twice (used when SCALED)
};

template <typename T, int NT, int TR, int KBLK, int STAGES>
struct DenseCfg {
  static constexpr int WARPS = 8;
  static constexpr int KB = KBLK;              // k per stage
  static constexpr int KC = KB / 64;           // 128-byte sub-tiles per stage
  static constexpr int MTILES = TR / 16;       // 16-row MMA tiles of W per stage
  static constexpr int MPW = MTILES >= WARPS ? MTILES / WARPS : 1;  // 16-row tiles per warp
  static constexpr int WK = MTILES >= WARPS ? 1 : WARPS / MTILES;   // warps sharing a row tile (split over k16 steps)
  static constexpr int KS = KB / 16 / WK;      // k16 steps per warp and stage
  static constexpr int MT = NT * 8;            // token rows held in smem / accumulators
  static constexpr int W_SUB = TR * 128;
  static constexpr int A_SUB = MT * 128;
  static constexpr int W_STAGE = KC * W_SUB;
  static constexpr int A_STAGE = KC * A_SUB;
  static constexpr int STAGE_TX = W_STAGE + A_STAGE;
  static constexpr int NS = STAGES;
  static constexpr int TRP = TR + 4;           // padded fp32 row of the reduction tile
  static constexpr int RED_BYTES = MT * TRP * 4;
  static constexpr int OFF_W = 0;
  static constexpr int OFF_A = OFF_W + NS * W_STAGE;
  static constexpr int OFF_RED = OFF_A + NS * A_STAGE;
  static constexpr int OFF_BAR = OFF_RED + (RED_BYTES + 15) / 16 * 16;
  static constexpr int SMEM_BYTES = OFF_BAR + 2 * NS * 8 + 16;
  static constexpr int THREADS = WARPS * 32 + 32;  // + producer warp
  static_assert((MTILES >= WARPS ? MTILES % WARPS == 0 : WARPS % MTILES == 0) && (KB / 16) % WK == 0 && KB % 64 == 0,
                "warp layout");
  static_assert(W_SUB % 1024 == 0 && A_SUB % 1024 == 0, "TMA swizzle alignment");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory");
};

template <typename T>
__device__ __forceinline__ uint32_t mul2(uint32_t a, uint32_t b);
template <>
__device__ __forceinline__ uint32_t mul2<__half>(uint32_t a, uint32_t b) {
  const __half2 r = __hmul2(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}
template <>
__device__ __forceinline__ uint32_t mul2<__nv_bfloat16>(uint32_t a, uint32_t b) {
  const __nv_bfloat162 r =
      __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}

template <typename T, int NT, int TR, int KBLK, int STAGES, bool SCALED>
__global__ void __launch_bounds__(DenseCfg<T, NT, TR, KBLK, STAGES>::THREADS, 1)
dense_stream_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_a,
                    const DenseArgs p) {
  using Cfg = DenseCfg<T, NT, TR, KBLK, STAGES>;
  constexpr int NS = Cfg::NS, KC = Cfg::KC, KB = Cfg::KB, WK = Cfg::WK, KS = Cfg::KS, TRP = Cfg::TRP;
  constexpr int MPW = Cfg::MPW;
  constexpr int NCONS = Cfg::WARPS * 32;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0u) __trap();  // 128B-swizzled boxes need the 1024-aligned window; checked, not assumed
  float* red = reinterpret_cast<float*>(smem + Cfg::OFF_RED);
  const uint32_t bar_full = sbase + Cfg::OFF_BAR;
  const uint32_t bar_empty = bar_full + NS * 8;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, Cfg::WARPS);
    }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();

  // even, contiguous split of the row tiles
  const int grid = gridDim.x;
  const int t0 = static_cast<int>((static_cast<long long>(blockIdx.x) * p.tiles) / grid);
  const int t1 = static_cast<int>((static_cast<long long>(blockIdx.x + 1) * p.tiles) / grid);
  const int my_units = (t1 - t0) * p.kblks;
  if (my_units == 0) return;

  if (warp == Cfg::WARPS) {
    // ===================================== producer =====================================
    if (lane == 0) {
      tma_prefetch_desc(&tm_w);
      tma_prefetch_desc(&tm_a);
      auto issue_w = [&](int s, int tile, int b) {
        const uint32_t full = bar_full + 8 * s;
        mbar_arrive_expect_tx(full, Cfg::STAGE_TX);
#pragma unroll
        for (int kc = 0; kc < KC; ++kc)
          tma_load_2d(sbase + Cfg::OFF_W + s * Cfg::W_STAGE + kc * Cfg::W_SUB, &tm_w, full, b * KB + kc * 64,
                      tile * TR, kEvictFirst);
      };
      auto issue_a = [&](int s, int b) {
        const uint32_t full = bar_full + 8 * s;
#pragma unroll
        for (int kc = 0; kc < KC; ++kc)
          tma_load_2d(sbase + Cfg::OFF_A + s * Cfg::A_STAGE + kc * Cfg::A_SUB, &tm_a, full, b * KB + kc * 64, 0,
                      kEvictLast);
      };
      const int npre = my_units < NS ? my_units : NS;
      int tile = t0, b = 0;
      for (int it = 0; it < npre; ++it) {  // the weights do not depend on the prior grid
        issue_w(it, tile, b);
        if (++b == p.kblks) { b = 0; ++tile; }
      }
      pdl_wait_prior_grids();              // the activations do
      {
        int bb = 0;
        for (int it = 0; it < npre; ++it) {
          issue_a(it, bb);
          if (++bb == p.kblks) bb = 0;
        }
      }
      int s = npre % NS;
      uint32_t ph = (npre == NS) ? 0u : 1u;
      for (int it = npre; it < my_units; ++it) {
        mbar_wait(bar_empty + 8 * s, ph);
        issue_w(s, tile, b);
        issue_a(s, b);
        if (++b == p.kblks) { b = 0; ++tile; }
        if (++s == NS) { s = 0; ph ^= 1u; }
      }
    }
    return;
  }

  // ======================================= consumers =======================================
  const int wm = (WK == 1) ? warp * MPW : warp % Cfg::MTILES;  // first 16-row tile of W of this warp
  const int wk = (WK == 1) ? 0 : warp / Cfg::MTILES;           // k-slice: k16 steps wk*KS .. wk*KS + KS - 1
  const int g = lane >> 2, t = lane & 3;

  float acc[MPW][NT][4];
#pragma unroll
  for (int mp = 0; mp < MPW; ++mp)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[mp][nt][e] = 0.f;

  // ldmatrix row addresses inside the 128B-swizzled sub-tiles (16-byte chunk c of row r sits at
  // r*128 + ((c ^ (r & 7)) << 4)); k16 step ks covers chunks 2*(ks & 3), +1 of sub-tile ks >> 2.
  //   W (.x4, 16 rows x 16 k): lanes 0-15 -> rows 0..15 at the low chunk, lanes 16-31 -> the high chunk
  //   tokens (.x2 / .x4): lanes 0-7 -> tokens 0..7 low chunk, 8-15 -> high chunk, 16-31 -> tokens 8..15
  const int w_row = wm * 16 + (lane & 15);
  const int w_hi = lane >> 4;
  const int a_row = (lane & 7) + ((NT >= 2 && (lane & 16)) ? 8 : 0);
  const int a_hi = (lane >> 3) & 1;

  int b = 0;
  int tile = t0;
  int s = 0;
  uint32_t ph = 0;
  bool waited = false;

  for (int it = 0; it < my_units; ++it) {
    mbar_wait(bar_full + 8 * s, ph);
    const uint32_t wst = sbase + Cfg::OFF_W + s * Cfg::W_STAGE;
    const uint32_t ast = sbase + Cfg::OFF_A + s * Cfg::A_STAGE;
#pragma unroll
    for (int i = 0; i < KS; ++i) {
      const int ks = wk * KS + i;       // k16 step inside the stage (wk is warp-uniform)
      const int sub = ks >> 2, c0 = (ks & 3) * 2;
      uint32_t wf[MPW][4];
#pragma unroll
      for (int mp = 0; mp < MPW; ++mp)  // +16 rows per tile: 2048 B, swizzle phase unchanged
        ldmatrix_x4(wf[mp][0], wf[mp][1], wf[mp][2], wf[mp][3],
                    wst + sub * Cfg::W_SUB + w_row * 128 + (((c0 + w_hi) ^ (w_row & 7)) << 4) + mp * 2048);
      uint32_t bf[NT][2];
      const uint32_t a_addr = ast + sub * Cfg::A_SUB + a_row * 128 + (((c0 + a_hi) ^ (a_row & 7)) << 4);
      if (NT == 1) {
        ldmatrix_x2(bf[0][0], bf[0][1], a_addr);
      } else {
#pragma unroll
        for (int np = 0; np < NT / 2; ++np)  // token rows +16 per pair: 2048 B, swizzle phase unchanged
          ldmatrix_x4(bf[2 * np][0], bf[2 * np][1], bf[2 * np + 1][0], bf[2 * np + 1][1], a_addr + np * 2048);
      }
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        if (SCALED) {
          bf[nt][0] = mul2<T>(bf[nt][0], p.scale2);
          bf[nt][1] = mul2<T>(bf[nt][1], p.scale2);
        }
#pragma unroll
        for (int mp = 0; mp < MPW; ++mp) mma_m16n8k16<T>(acc[mp][nt], wf[mp], bf[nt][0], bf[nt][1]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * s);
    if (++s == NS) { s = 0; ph ^= 1u; }

    if (++b == p.kblks) {
      // ---- row tile complete: k-slices add up in wk order (fixed -> deterministic), then write T ----
      named_bar_sync(1, NCONS);  // the previous tile's readers of `red` are done
#pragma unroll
      for (int r = 0; r < WK; ++r) {
        if (wk == r) {
#pragma unroll
          for (int mp = 0; mp < MPW; ++mp)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int n = (wm + mp) * 16 + g + ((e & 2) ? 8 : 0);
                const int tok = nt * 8 + 2 * t + (e & 1);
                float* dst = red + tok * TRP + n;
                *dst = (r == 0) ? acc[mp][nt][e] : (*dst + acc[mp][nt][e]);
                acc[mp][nt][e] = 0.f;
              }
        }
        named_bar_sync(1, NCONS);
      }
      if (!waited) {
        pdl_wait_prior_grids();  // `out` may be read or written by prior grids
        waited = true;
      }
      const int tid = static_cast<int>(threadIdx.x);
      const int row0 = tile * TR;
      const int items = p.m * (TR / 4);
      T* outp = reinterpret_cast<T*>(p.out);
      const T* biasp = reinterpret_cast<const T*>(p.bias);
      const bool vec = (p.n & 3) == 0;
      for (int e = tid; e < items; e += NCONS) {
        const int tok = e / (TR / 4), nn = (e - tok * (TR / 4)) * 4;
        const int n = row0 + nn;
        if (n >= p.n) continue;
        const float4 v4 = *reinterpret_cast<const float4*>(red + tok * TRP + nn);
        const float v[4] = {v4.x, v4.y, v4.z, v4.w};
        const size_t off = static_cast<size_t>(tok) * p.n + n;
        T o[4];
        if (vec) {
          if (p.inplace) {
            const uint2 old = *reinterpret_cast<const uint2*>(outp + off);
            const T* oh = reinterpret_cast<const T*>(&old);
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q] = Elem<T>::from_float(v[q] + Elem<T>::to_float(oh[q]));
          } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q] = Elem<T>::from_float(v[q]);
          }
          if (biasp) {
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q] = __hadd(o[q], biasp[n + q]);
          }
          uint2 pk;
          pk.x = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[0])) |
                 (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[1])) << 16);
          pk.y = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[2])) |
                 (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[3])) << 16);
          *reinterpret_cast<uint2*>(outp + off) = pk;
        } else {
          for (int q = 0; q < 4 && n + q < p.n; ++q) {
            T x = Elem<T>::from_float(p.inplace ? v[q] + Elem<T>::to_float(outp[off + q]) : v[q]);
            if (biasp) x = __hadd(x, biasp[n + q]);
            outp[off + q] = x;
          }
        }
      }
      b = 0;
      ++tile;
    }
  }
}

}  // namespace b200
