// HBM-bound regime, second generation (decode / speculative tree-verify, M <= 32 per launch):
// small weight-streaming CTAs, several resident per SM, split-K across a THREAD-BLOCK CLUSTER and
// reduced through distributed shared memory.
//
//   grid    = column tiles x S, cluster = (S,1,1): the S CTAs of a cluster own the same TN output
//             columns and consecutive K ranges.  No global scratch, no atomics, no tickets: after
//             the K loop the partial tiles sit in each CTA's shared memory, one cluster barrier
//             publishes them, CTA r sums slice r of the tile over ranks 0..S-1 IN RANK ORDER
//             (bit-deterministic) with ld.shared::cluster and writes T outputs.  The reference
//             serialises the CTAs of a column through a global spin lock and fp32 round trips in L2
//             (marlin_device_ops.cuh:347-377, marlin_kernel_impl.cuh:886-929).
//   CTA     = 1 TMA producer lane + WARPS consumer warps, <= ~70 KB of shared memory so that 2-3 CTAs
//             (and, with PDL, CTAs of the NEXT linear) are resident per SM: the next kernel's weight
//             ring is already full when the previous kernel drains.
//   stream  = [KROWS x CHUNKS] boxes of the Marlin int32 stream by TMA (evict-first) + the group's
//             scale row + the [M x KB] activation box (128B swizzle, zero-filled past M); a lane's
//             16 bytes -> lop3/HSUB2/HFMA2 (exact) -> HMUL2 by the group scale (the reference's
//             single rounding) -> mma.sync m16n8k16 with the weights as the 16-row operand.
#pragma once
#include "common.cuh"

namespace b200 {

struct StreamArgs {
  void* c;               // T[M][N]
  const void* scales;    // raw pointer (channelwise epilogue)
  const void* bias;      // T[N] or null
  const void* residual;  // T[M][N] or null
  uint32_t out_scale;    // 0, or 0x10000 | bits of T(scale): last post-op
  int m, n, k;
  int kblks;             // K / KB
  int dbg_skip;          // tuning only: 1 = consumers release stages without computing (memory-pipeline ceiling)
  unsigned long long* dbg_ts;  // tuning only: per-CTA phase timestamps (8 x u64 per CTA, first 1024 CTAs) or null
};

__device__ __forceinline__ void dbg_stamp(const StreamArgs& p, int slot) {
  if (p.dbg_ts != nullptr && blockIdx.x < 1024) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.dbg_ts[blockIdx.x * 8 + slot] = t;
  }
}

template <typename T, int NT, int CHUNKS, int KROWS, int WARPS, int STAGES, bool GROUPED>
struct StreamCfg {
  static constexpr int TN = CHUNKS * 64;
  static constexpr int KB = KROWS * 16;
  static constexpr int WK = WARPS / CHUNKS;
  static constexpr int RPW = KROWS / WK;
  static constexpr int KC = KB / 64;
  static constexpr int MT = NT * 8;
  static constexpr int A_SUB = MT * 128;
  static constexpr int A_STAGE = KC * A_SUB;
  static constexpr int W_STAGE = KROWS * CHUNKS * 512;
  static constexpr int S_STAGE = GROUPED ? TN * 2 : 0;
  static constexpr int STAGE_TX = A_STAGE + W_STAGE + S_STAGE;
  static constexpr int TNP = TN + 4;
  static constexpr int RED_BYTES = MT * TNP * 4;
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W = OFF_A + STAGES * A_STAGE;
  static constexpr int OFF_S = OFF_W + STAGES * W_STAGE;
  // the fp32 partial tile ALIASES the rings: it is written only after every consumer warp has
  // drained its last stage (named barrier), when no TMA write is outstanding any more
  static constexpr int OFF_RED = 0;
  static constexpr int RING_BYTES = (OFF_S + STAGES * S_STAGE + 127) / 128 * 128;
  static constexpr int OFF_BAR = RING_BYTES > RED_BYTES ? RING_BYTES : (RED_BYTES + 127) / 128 * 128;
  static constexpr int SMEM_BYTES = OFF_BAR + 2 * STAGES * 8 + 16;
  static constexpr int THREADS = WARPS * 32 + 32;
  static_assert(WARPS % CHUNKS == 0 && KROWS % WK == 0, "warp layout");
  static_assert(KB % 64 == 0, "stage k must be a multiple of the 128-byte activation row");
  static_assert(A_SUB % 1024 == 0 && W_STAGE % 512 == 0, "TMA alignment");
};

template <typename T, int NT, int CHUNKS, int KROWS, int WARPS, int STAGES, bool GROUPED>
__global__ void __launch_bounds__(WARPS * 32 + 32, (WARPS >= 16 && NT <= 2) ? 2 : 0)
w4a16_stream_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_s,
                    const __grid_constant__ CUtensorMap tm_a, const StreamArgs p) {
  using Cfg = StreamCfg<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED>;
  constexpr int TN = Cfg::TN, KB = Cfg::KB, WK = Cfg::WK, RPW = Cfg::RPW, KC = Cfg::KC;
  constexpr int MT = Cfg::MT, TNP = Cfg::TNP;
  constexpr int NCONS = WARPS * 32;

  // identical carve-up in every CTA of the cluster (same dynamic smem base), so a local offset
  // maps to the same object in a peer CTA.  No static shared memory in this kernel: the dynamic
  // window starts 1024-aligned (needed by the 128B-swizzled activation boxes); checked, not assumed.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0u) __trap();
  float* red = reinterpret_cast<float*>(smem + Cfg::OFF_RED);
  const uint32_t bar_full = sbase + Cfg::OFF_BAR;
  const uint32_t bar_empty = bar_full + STAGES * 8;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int S = static_cast<int>(cluster_nctarank());
  const int rank = static_cast<int>(cluster_ctarank());
  const int c = blockIdx.x / S;  // column tile

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, WARPS);
    }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();
  if (threadIdx.x == 0) dbg_stamp(p, 0);

  // this CTA's K range, in k-blocks of KB
  const int kb0 = static_cast<int>((static_cast<long long>(rank) * p.kblks) / S);
  const int kb1 = static_cast<int>((static_cast<long long>(rank + 1) * p.kblks) / S);
  const int my_blks = kb1 - kb0;

  if (warp == WARPS) {
    // ===================================== producer =====================================
    if (lane == 0) {
      tma_prefetch_desc(&tm_w);
      tma_prefetch_desc(&tm_a);
      if (GROUPED) tma_prefetch_desc(&tm_s);
      auto issue_w = [&](int it) {
        const int b = kb0 + it;
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
        const int b = kb0 + it;
        const int s = it % STAGES;
        const uint32_t full = bar_full + 8 * s;
#pragma unroll
        for (int kc = 0; kc < KC; ++kc)
          tma_load_2d(sbase + Cfg::OFF_A + s * Cfg::A_STAGE + kc * Cfg::A_SUB, &tm_a, full,
                      b * KB + kc * 64, 0, kEvictLast);
      };
      const int npre = my_blks < STAGES ? my_blks : STAGES;
      for (int it = 0; it < npre; ++it) issue_w(it);  // weights do not depend on the prior grid
      pdl_wait_prior_grids();                         // activations do
      dbg_stamp(p, 1);
      for (int it = 0; it < npre; ++it) issue_a(it);
      for (int it = npre; it < my_blks; ++it) {
        mbar_wait(bar_empty + 8 * (it % STAGES), ((it / STAGES) & 1) ^ 1);
        issue_w(it);
        issue_a(it);
      }
    }
  } else {
    // ======================================= consumers =======================================
    const int wn = warp % CHUNKS;
    const int wk = warp / CHUNKS;
    const int g = lane >> 2, t = lane & 3;

    float acc[4][NT][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[j][nt][e] = 0.f;

    const int lm_row = (lane & 7) + ((NT >= 2 && (lane & 16)) ? 8 : 0);
    const int lm_half = (lane >> 3) & 1;

    for (int it = 0; it < my_blks; ++it) {
      const int s = it % STAGES;
      mbar_wait(bar_full + 8 * s, (it / STAGES) & 1);
      if (it == 0 && threadIdx.x == 0) dbg_stamp(p, 2);

      if (p.dbg_skip) {
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + 8 * s);
        continue;
      }
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
        const int r = wk + WK * i;
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
    }

    // ---- reduce the WK warps sharing each chunk through shared memory (fixed order) ----
    if (threadIdx.x == 0) dbg_stamp(p, 3);
    named_bar_sync(1, NCONS);  // the partial tile aliases the rings: every warp is past its last stage
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
            }
      }
      if (WK > 1) named_bar_sync(1, NCONS);
    }
  }

  // ---- cluster reduction: publish, then CTA `rank` finishes slice `rank` of the tile ----
  if (threadIdx.x == 0) dbg_stamp(p, 4);
  if (S > 1) cluster_sync_all(); else __syncthreads();
  pdl_wait_prior_grids();  // every thread: C / residual are shared with prior grids
  if (threadIdx.x == 0) dbg_stamp(p, 5);

  if (warp < WARPS) {
    const int col0 = c * TN;
    constexpr int V4_PER_ROW = TN / 4;
    const int total = p.m * V4_PER_ROW;                      // float4 items of the valid rows
    const int e0 = static_cast<int>((static_cast<long long>(rank) * total) / S);
    const int e1 = static_cast<int>((static_cast<long long>(rank + 1) * total) / S);
    const uint32_t red_addr = sbase + Cfg::OFF_RED;
    for (int e = e0 + threadIdx.x; e < e1; e += NCONS) {
      const int tok = e / V4_PER_ROW, nn = (e - tok * V4_PER_ROW) * 4;
      const uint32_t off = static_cast<uint32_t>(tok * TNP + nn) * 4u;
      float4 sum;
      if (S == 1) {
        sum = *reinterpret_cast<const float4*>(red + tok * TNP + nn);
      } else {
        sum = ld_dsmem_f4(red_addr + off, 0);
        for (int r = 1; r < S; ++r) {
          const float4 v = ld_dsmem_f4(red_addr + off, static_cast<uint32_t>(r));
          sum.x += v.x;
          sum.y += v.y;
          sum.z += v.z;
          sum.w += v.w;
        }
      }
      const int n = col0 + nn;
      if (n >= p.n) continue;
      T o[4] = {Elem<T>::from_float(sum.x), Elem<T>::from_float(sum.y), Elem<T>::from_float(sum.z),
                Elem<T>::from_float(sum.w)};
      if (!GROUPED) {  // channelwise: scale the ROUNDED sum in T (marlin_kernel_impl.cuh:957-963)
        const T* sc = reinterpret_cast<const T*>(p.scales);
#pragma unroll
        for (int i = 0; i < 4; ++i) o[i] = __hmul(o[i], sc[chan_scale_pos(n + i)]);
      }
      const size_t goff = static_cast<size_t>(tok) * p.n + n;
      if (p.residual) {
        const T* rs = reinterpret_cast<const T*>(p.residual) + goff;
#pragma unroll
        for (int i = 0; i < 4; ++i) o[i] = __hadd(rs[i], o[i]);
      }
      if (p.bias) {
        const T* bs = reinterpret_cast<const T*>(p.bias) + n;
#pragma unroll
        for (int i = 0; i < 4; ++i) o[i] = __hadd(o[i], bs[i]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) o[i] = apply_out_scale<T>(o[i], p.out_scale);
      uint2 pk;
      pk.x = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[0])) |
             (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[1])) << 16);
      pk.y = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[2])) |
             (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[3])) << 16);
      *reinterpret_cast<uint2*>(reinterpret_cast<T*>(p.c) + goff) = pk;
    }
  }
  if (threadIdx.x == 0) dbg_stamp(p, 6);
  // no CTA may exit while a peer can still read its shared memory
  if (S > 1) cluster_sync_all();
  if (threadIdx.x == 0) dbg_stamp(p, 7);
}

}  // namespace b200
