// HBM-bound regime (decode / speculative tree-verify, M <= 32): a persistent, warp-specialised
// weight-streaming kernel.
//
//   * one CTA per SM; ONE producer lane feeds a STAGES-deep ring of shared-memory stages with TMA
//     (cp.async.bulk.tensor): a [KB x TN] box of the Marlin int32 stream (evict-first: streamed
//     once), the matching [M x KB] activation box (128B-swizzled, zero-filled past M) and the
//     group-scale row.  full/empty mbarriers; the ring keeps >= 64 KB of HBM requests in flight per SM.
//   * WARPS consumer warps.  Each owns one 64-column chunk and 1/WK of the stage's k-rows; a lane's
//     16 bytes of a 512-byte Marlin block are the A-operand fragments of four m16n8k16 MMAs once
//     dequantised (weights are the 16-row M side, tokens the 8-column N side), so the int4 stream
//     goes LDS.128 -> registers -> lop3/HSUB2/HFMA2 (exact) -> HMUL2 by the group scale (the
//     reference's single rounding) -> mma.sync with fp32 accumulators.  Nothing is re-staged.
//   * work = (column tile c, k-block b) units in column-major order, split evenly over the CTAs
//     (|units_i - units_j| <= 1).  A CTA accumulates a run of k-blocks of one column in registers,
//     then flushes: warps sharing a column reduce through shared memory; a column whose K range is
//     split over several CTAs is reduced through fp32 slots in c_tmp in FIXED slot order by
//     whichever CTA arrives last (atomic ticket in `workspace`, reset to 0) — deterministic, no
//     spinning, no co-residency assumption (the reference serialises CTAs with a spin lock,
//     marlin_device_ops.cuh:347-377).
//   * PDL: the producer prefetches weights/scales for its first STAGES units BEFORE
//     griddepcontrol.wait, so back-to-back linears keep HBM busy across kernel boundaries.
#pragma once
#include "common.cuh"

namespace b200 {

struct DecodeArgs {
  void* c;               // T[M][N]
  float* c_tmp;          // split-K slots
  int32_t* workspace;    // tickets, one per column tile
  const void* scales;    // raw pointer (channelwise epilogue)
  const void* bias;      // T[N] or null
  const void* residual;  // T[M][N] or null
  int m, n, k;
  int ncols;      // column tiles = ceil(N / TN)
  int kblks;      // K / KB
  int slot_max;   // slots per column reserved in c_tmp
  unsigned long long* dbg_ts;  // tuning only: per-CTA phase timestamps or null
};

__device__ __forceinline__ void dbg_stamp1(const DecodeArgs& p, int slot) {
  if (p.dbg_ts != nullptr && blockIdx.x < 1024) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.dbg_ts[blockIdx.x * 8 + slot] = t;
  }
}

template <typename T, int NT, int CHUNKS, int KROWS, int WARPS, int STAGES, bool GROUPED>
struct DecodeCfg {
  static constexpr int TN = CHUNKS * 64;       // columns per tile
  static constexpr int KB = KROWS * 16;        // k per stage
  static constexpr int WK = WARPS / CHUNKS;    // warps sharing a chunk (split over k-rows)
  static constexpr int RPW = KROWS / WK;       // k-rows per warp per stage
  static constexpr int KC = KB / 64;           // 64-wide activation sub-tiles per stage
  static constexpr int MT = NT * 8;            // token rows held in smem / accumulators
  static constexpr int A_SUB = MT * 128;       // bytes of one activation sub-tile
  static constexpr int A_STAGE = KC * A_SUB;
  static constexpr int W_STAGE = KROWS * CHUNKS * 512;
  static constexpr int S_STAGE = GROUPED ? TN * 2 : 0;
  static constexpr int STAGE_TX = A_STAGE + W_STAGE + S_STAGE;
  static constexpr int TNP = TN + 4;           // padded fp32 row of the reduction tile
  static constexpr int RED_BYTES = MT * TNP * 4;
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W = OFF_A + STAGES * A_STAGE;
  static constexpr int OFF_S = OFF_W + STAGES * W_STAGE;
  static constexpr int OFF_RED = (OFF_S + STAGES * S_STAGE + 127) / 128 * 128;
  static constexpr int OFF_BAR = OFF_RED + RED_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 2 * STAGES * 8 + 16 + 1024;  // +1024: manual alignment
  static constexpr int THREADS = WARPS * 32 + 32;
  static_assert(WARPS % CHUNKS == 0 && KROWS % WK == 0, "warp layout");
  static_assert(KB % 64 == 0, "stage k must be a multiple of the 128-byte activation row");
  static_assert(A_SUB % 1024 == 0 && W_STAGE % 1024 == 0, "TMA swizzle alignment");
};

// unit partition helpers (64-bit: units * grid can exceed 2^31)
__device__ __forceinline__ int unit_begin(int cta, long long units, int grid) {
  return static_cast<int>((static_cast<long long>(cta) * units) / grid);
}
// the CTA whose range contains unit u: largest i with floor(i*U/G) <= u
__device__ __forceinline__ int cta_of_unit(long long u, long long units, int grid) {
  return static_cast<int>(((u + 1) * grid - 1) / units);
}

template <typename T, int NT, int CHUNKS, int KROWS, int WARPS, int STAGES, bool GROUPED>
__global__ void __launch_bounds__(WARPS * 32 + 32, (NT <= 2 && WARPS == 8) ? 2 : 1)
w4a16_decode_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_s,
                    const __grid_constant__ CUtensorMap tm_a, const DecodeArgs p) {
  using Cfg = DecodeCfg<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED>;
  using T2 = typename Elem<T>::T2;
  constexpr int TN = Cfg::TN, KB = Cfg::KB, WK = Cfg::WK, RPW = Cfg::RPW, KC = Cfg::KC;
  constexpr int MT = Cfg::MT, TNP = Cfg::TNP;
  constexpr int NCONS = WARPS * 32;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const uint32_t sbase = smem_u32(smem);
  float* red = reinterpret_cast<float*>(smem + Cfg::OFF_RED);
  const uint32_t bar_full = sbase + Cfg::OFF_BAR;
  const uint32_t bar_empty = bar_full + STAGES * 8;
  volatile int* s_ticket = reinterpret_cast<volatile int*>(smem + Cfg::OFF_BAR + 2 * STAGES * 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, WARPS);
    }
    fence_barrier_init();
  }
  __syncthreads();
  // let the next kernel in the stream start its own weight prefetch as early as possible
  pdl_launch_dependents();
  if (threadIdx.x == 0) dbg_stamp1(p, 0);

  const long long units = static_cast<long long>(p.ncols) * p.kblks;
  const int grid = gridDim.x;
  const int u0 = unit_begin(blockIdx.x, units, grid);
  const int u1 = unit_begin(blockIdx.x + 1, units, grid);
  const int my_units = u1 - u0;

  if (warp == WARPS) {
    // ===================================== producer =====================================
    if (lane == 0) {
      tma_prefetch_desc(&tm_w);
      tma_prefetch_desc(&tm_a);
      if (GROUPED) tma_prefetch_desc(&tm_s);
      auto issue_w = [&](int it) {
        const int u = u0 + it;
        const int c = u / p.kblks, b = u - c * p.kblks;
        const int s = it % STAGES;
        const uint32_t full = bar_full + 8 * s;
        mbar_arrive_expect_tx(full, Cfg::STAGE_TX);
        tma_load_3d(sbase + Cfg::OFF_W + s * Cfg::W_STAGE, &tm_w, full, 0, c * CHUNKS, b * KROWS,
                    kEvictFirst);
        if (GROUPED)
          tma_load_3d(sbase + Cfg::OFF_S + s * Cfg::S_STAGE, &tm_s, full, 0, c * CHUNKS,
                      (b * KB) / 128, kEvictLast);
      };
      auto issue_a = [&](int it) {
        const int u = u0 + it;
        const int b = u % p.kblks;
        const int s = it % STAGES;
        const uint32_t full = bar_full + 8 * s;
#pragma unroll
        for (int kc = 0; kc < KC; ++kc)
          tma_load_2d(sbase + Cfg::OFF_A + s * Cfg::A_STAGE + kc * Cfg::A_SUB, &tm_a, full,
                      b * KB + kc * 64, 0, kEvictLast);
      };
      const int npre = my_units < STAGES ? my_units : STAGES;
      for (int it = 0; it < npre; ++it) issue_w(it);   // weights do not depend on the prior grid
      pdl_wait_prior_grids();                          // activations do
      dbg_stamp1(p, 1);
      for (int it = 0; it < npre; ++it) issue_a(it);
      for (int it = npre; it < my_units; ++it) {
        mbar_wait(bar_empty + 8 * (it % STAGES), ((it / STAGES) & 1) ^ 1);
        issue_w(it);
        issue_a(it);
      }
    }
    return;
  }

  // ======================================= consumers =======================================
  pdl_wait_prior_grids();  // before any global write (C, c_tmp, workspace are shared with prior grids)
  if (my_units == 0) return;

  const int wn = warp % CHUNKS;  // chunk (64 columns) of the tile
  const int wk = warp / CHUNKS;  // k-row phase
  const int g = lane >> 2, t = lane & 3;

  float acc[4][NT][4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[j][nt][e] = 0.f;

  // ldmatrix source row/chunk of this lane inside a 128B-swizzled [MT][64] activation sub-tile
  //   NT==1 (.x2): lanes 0-7 -> k 0..7 of the k16 step, lanes 8-15 -> k 8..15 (16-31 ignored)
  //   NT>=2 (.x4): lanes 16-31 -> token rows 8..15 of the pair
  const int lm_row = (lane & 7) + ((NT >= 2 && (lane & 16)) ? 8 : 0);
  const int lm_half = (lane >> 3) & 1;

  int c = u0 / p.kblks;
  int b = u0 - c * p.kblks;
  int seg_first_b = b;

  for (int it = 0; it < my_units; ++it) {
    const int s = it % STAGES;
    mbar_wait(bar_full + 8 * s, (it / STAGES) & 1);
    if (it == 0 && threadIdx.x == 0) dbg_stamp1(p, 2);

    uint32_t s_lo[4], s_hi[4];
    if (GROUPED) {
      const uint4 sv = lds128(sbase + Cfg::OFF_S + s * Cfg::S_STAGE + wn * 128 + g * 16);
      broadcast_scale_pair(sv.x, s_lo[0], s_hi[0]);
      broadcast_scale_pair(sv.y, s_lo[1], s_hi[1]);
      broadcast_scale_pair(sv.z, s_lo[2], s_hi[2]);
      broadcast_scale_pair(sv.w, s_lo[3], s_hi[3]);
    }
    const uint32_t wst = sbase + Cfg::OFF_W + s * Cfg::W_STAGE;
    const uint32_t ast = sbase + Cfg::OFF_A + s * Cfg::A_STAGE;

    uint4 q[RPW];
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      const int r = wk + WK * i;
      q[i] = lds128(wst + (r * CHUNKS + wn) * 512 + lane * 16);
    }
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      const int r = wk + WK * i;  // k16 row inside the stage
      // activation fragments: tokens as the N side of the MMA
      uint32_t bf[NT][2];
      {
        const uint32_t sub = ast + (r >> 2) * Cfg::A_SUB;
        const int c16 = ((r & 3) << 1) + lm_half;
        if (NT == 1) {
          const uint32_t addr = sub + lm_row * 128 + ((c16 ^ (lm_row & 7)) << 4);
          ldmatrix_x2(bf[0][0], bf[0][1], addr);
        } else {
#pragma unroll
          for (int np = 0; np < NT / 2; ++np) {
            const int row = np * 16 + lm_row;
            const uint32_t addr = sub + row * 128 + ((c16 ^ (row & 7)) << 4);
            ldmatrix_x4(bf[2 * np][0], bf[2 * np][1], bf[2 * np + 1][0], bf[2 * np + 1][1], addr);
          }
        }
      }
      const uint32_t qq[4] = {q[i].x, q[i].y, q[i].z, q[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t w[4];
        Dequant<T>::template word<GROUPED>(qq[j], s_lo[j], s_hi[j], w);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) mma_m16n8k16<T>(acc[j][nt], w, bf[nt][0], bf[nt][1]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * s);

    // ---- end of this CTA's run inside column c?  flush ----
    const bool last_unit = (it == my_units - 1);
    if (b == p.kblks - 1 || last_unit) {
      const int seg_last_b = b;
      if (threadIdx.x == 0) dbg_stamp1(p, last_unit ? 5 : 3);
      // 1. reduce the WK warps sharing each chunk through shared memory
#pragma unroll
      for (int r = 0; r < WK; ++r) {
        if (wk == r) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int n = wn * 64 + j * 16 + g + ((e & 2) ? 8 : 0);
                const int tok = nt * 8 + 2 * t + (e & 1);
                float* dst = red + tok * TNP + n;
                *dst = (r == 0) ? acc[j][nt][e] : (*dst + acc[j][nt][e]);
                acc[j][nt][e] = 0.f;
              }
        }
        named_bar_sync(1, NCONS);
      }
      const int col0 = c * TN;
      const int tid = threadIdx.x;
      const bool whole = (seg_first_b == 0 && seg_last_b == p.kblks - 1);

      // converts 4 fp32 sums to T with the reference's rounding points and stores them
      auto emit = [&](int tok, int n, float4 v) {
        T o[4] = {Elem<T>::from_float(v.x), Elem<T>::from_float(v.y), Elem<T>::from_float(v.z),
                  Elem<T>::from_float(v.w)};
        if (!GROUPED) {  // channelwise: scale the ROUNDED sum in T (marlin_kernel_impl.cuh:957-963)
          const T* sc = reinterpret_cast<const T*>(p.scales);
#pragma unroll
          for (int e = 0; e < 4; ++e) o[e] = __hmul(o[e], sc[chan_scale_pos(n + e)]);
        }
        const size_t off = static_cast<size_t>(tok) * p.n + n;
        if (p.residual) {
          const T* rs = reinterpret_cast<const T*>(p.residual) + off;
#pragma unroll
          for (int e = 0; e < 4; ++e) o[e] = __hadd(rs[e], o[e]);
        }
        if (p.bias) {
          const T* bs = reinterpret_cast<const T*>(p.bias) + n;
#pragma unroll
          for (int e = 0; e < 4; ++e) o[e] = __hadd(o[e], bs[e]);
        }
        uint2 pk;
        pk.x = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[0])) |
               (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[1])) << 16);
        pk.y = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[2])) |
               (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[3])) << 16);
        *reinterpret_cast<uint2*>(reinterpret_cast<T*>(p.c) + off) = pk;
      };

      if (whole) {
        for (int e = tid * 4; e < MT * TN; e += NCONS * 4) {
          const int tok = e / TN, nn = e - tok * TN;
          if (tok < p.m && col0 + nn < p.n)
            emit(tok, col0 + nn, *reinterpret_cast<const float4*>(red + tok * TNP + nn));
        }
      } else {
        // 2. split column: publish this CTA's fp32 partial, take a ticket, last arrival reduces
        const long long ufirst = static_cast<long long>(c) * p.kblks;
        const int cta_first = cta_of_unit(ufirst, units, grid);
        const int cta_last = cta_of_unit(ufirst + p.kblks - 1, units, grid);
        const int nparts = cta_last - cta_first + 1;
        const int my_slot = static_cast<int>(blockIdx.x) - cta_first;
        float* slots = p.c_tmp + static_cast<size_t>(c) * p.slot_max * p.m * TN;
        float* mine = slots + static_cast<size_t>(my_slot) * p.m * TN;
        for (int e = tid * 4; e < p.m * TN; e += NCONS * 4) {
          const int tok = e / TN, nn = e - tok * TN;
          __stcg(reinterpret_cast<float4*>(mine + e),
                 *reinterpret_cast<const float4*>(red + tok * TNP + nn));
        }
        __threadfence();
        named_bar_sync(1, NCONS);
        if (tid == 0) *s_ticket = atomicAdd(p.workspace + c, 1);
        named_bar_sync(1, NCONS);
        if (*s_ticket == nparts - 1) {
          __threadfence();
          for (int e = tid * 4; e < p.m * TN; e += NCONS * 4) {
            const int tok = e / TN, nn = e - tok * TN;
            if (col0 + nn >= p.n) continue;
            float4 sum = __ldcg(reinterpret_cast<const float4*>(slots + e));
            for (int sl = 1; sl < nparts; ++sl) {
              const float4 v =
                  __ldcg(reinterpret_cast<const float4*>(slots + static_cast<size_t>(sl) * p.m * TN + e));
              sum.x += v.x;
              sum.y += v.y;
              sum.z += v.z;
              sum.w += v.w;
            }
            emit(tok, col0 + nn, sum);
          }
          if (tid == 0) p.workspace[c] = 0;  // leave the workspace zeroed (reference contract)
        }
      }
      named_bar_sync(1, NCONS);  // `red` and s_ticket are reusable
      if (threadIdx.x == 0) dbg_stamp1(p, last_unit ? 7 : 4);
      ++c;
      b = 0;
      seg_first_b = 0;
    } else {
      ++b;
    }
  }
}

}  // namespace b200
