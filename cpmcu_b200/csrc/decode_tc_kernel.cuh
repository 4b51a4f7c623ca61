// w4a16_decode_tc_kernel: the streaming (decode / tree-verify) GEMM with the math on the 5th-generation tensor cores.
//
// Same outer structure as w4a16_decode_kernel (decode_kernel.cuh): persistent CTAs, stream-K units of
// (256-column tile, 128-k block), ONE TMA producer lane streaming int4 words + group scales + token rows through a
// stage-contiguous ring, PDL (weights are requested before griddepcontrol.wait), split columns reduced through
// fp32 slots and tickets (finish_segment), the flusher warp finishing all but the CTA's last segment.
//
// What changes is the consumer math.  mma.sync on sm_100a costs 16-17 issue cycles per m16n8k16 per SM sub-partition
// (profiles/r02/m_ubench_*): one per int4 word per 8 tokens, so the register-operand loop is tensor-bound from M = 16 on
// (36.6 clk per word at M <= 16, 68.5 at M <= 32 against 22 for the dequant alone).  Here the dequantised fragments go
// from registers straight into TENSOR MEMORY as the A operand of tcgen05.mma (tcgen05.st.16x128b.x2 takes
// Dequant<T>::word's registers as they are, common.cuh), the token rows are the B operand in shared memory (the same
// 128B-swizzled boxes the producer already loads), and the accumulator D[256 columns][tokens] lives in tensor memory:
// the cost per word no longer depends on M, no shared-memory round trip of the 4x larger dequantised weights, and the
// consumer warps hold no accumulators.
//
//   warp  0..7   dequantisers: warp w owns TMEM lane quarter q = w % 4 of the 128-column half s = w / 4 of the tile,
//                i.e. 32 weight columns = words j in {2(q&1), 2(q&1)+1} of Marlin chunk 2s + q/2, for ALL 8 k16-rows
//                of a stage: LDS.64 -> Dequant -> tcgen05.st.  Every 4 rows (a half-stage): wait::st, fence, arrive.
//   warp  8      TMA producer (one lane), as in decode_kernel.cuh
//   warp  9      flusher: finish_segment for all but the last segment
//   warp 10      UMMA issuer and owner of the TMEM allocation (256 columns: 2 CTAs fit an SM, PDL co-residency)
//
// TMEM columns: D at [0, 2 * NROWS) (half s at s * NROWS; NROWS = token rows, 16 or 32); A ring at 64 + 64 * b (two
// half-stage buffers b), half s at + 32 * s, k16-row r at + 8 * (r & 3).
#pragma once
#include "decode_kernel.cuh"

namespace b200 {

template <typename T, int NT, int CH, bool GROUPED>
struct DecodeTcCfg {
  static constexpr int KROWS = CH == 4 ? 8 : 16;    // 256 columns x 128 k, or 128 columns x 256 k per stage (16 KB of int4 both)
  using Base = DecodeCfg<T, NT, CH, KROWS, 8, 5, GROUPED>;
  static constexpr int SUBS = CH / 2;                // 128-column UMMA halves of the tile
  static constexpr int HALVES = KROWS / 4;           // hand-over units of 4 k16-rows per stage
  static constexpr int NROWS = NT * 8;               // token rows = UMMA N (16 or 32)
  static constexpr int WARPS = 8;
  static constexpr int THREADS = WARPS * 32 + 96;    // + producer, flusher, issuer
  static constexpr int OFF_XBAR = (Base::SMEM_BYTES + 7) / 8 * 8;   // a_full[2], a_empty[2], d_full, done
  static constexpr int OFF_SLOT = OFF_XBAR + 6 * 8;
  static constexpr int SMEM_BYTES = OFF_SLOT + 16;
  static constexpr int TMEM_COLS = 256;
  static constexpr int TM_A = SUBS * NROWS > 64 ? SUBS * NROWS : 64;   // first column of the A ring
  static constexpr int A_BUF = SUBS * 32;             // columns of one A buffer (4 k16-rows x 8 columns per half)
  static_assert(NT == 2 || NT == 4, "token rows: 16 or 32 (UMMA N is a multiple of 16)");
  static_assert(CH == 2 || CH == 4, "tile: 128 or 256 columns");
  static_assert(Base::KC == HALVES && TM_A + 2 * A_BUF <= TMEM_COLS, "one 64-k token box per hand-over unit; TMEM budget");
};

template <typename T, int NT, int CH, bool GROUPED>
__global__ void __launch_bounds__(DecodeTcCfg<T, NT, CH, GROUPED>::THREADS, 2)
w4a16_decode_tc_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_s,
                       const __grid_constant__ CUtensorMap tm_a, const DecodeArgs p) {
  using TC = DecodeTcCfg<T, NT, CH, GROUPED>;
  using Cfg = typename TC::Base;
  constexpr int TN = Cfg::TN, TNP = Cfg::TNP, NS = Cfg::NSTAGE, WARPS = TC::WARPS, NROWS = TC::NROWS;
  constexpr int NCONS = WARPS * 32;
  constexpr int CHUNKS = CH, KROWS = TC::KROWS, KB = Cfg::KB, SUBS = TC::SUBS, HALVES = TC::HALVES;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0u) __trap();
  float* red = reinterpret_cast<float*>(smem + Cfg::OFF_RED);
  const uint32_t bar_full = sbase + Cfg::OFF_BAR;
  const uint32_t bar_empty = bar_full + NS * 8;
  const uint32_t bar_red_full = bar_empty + NS * 8;
  const uint32_t bar_red_empty = bar_red_full + 8;
  const uint32_t bar_a_full = sbase + TC::OFF_XBAR;        // [2] dequantisers -> issuer: half-stage of A written
  const uint32_t bar_a_empty = bar_a_full + 16;            // [2] issuer (commit) -> dequantisers: A buffer drained
  const uint32_t bar_d_full = bar_a_empty + 16;            // issuer (commit) -> dequantisers: the segment's D is complete
  const uint32_t bar_done = bar_d_full + 8;                // dequantisers -> issuer: tensor memory no longer used
  volatile int* tk = reinterpret_cast<volatile int*>(smem + Cfg::OFF_TK);
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + TC::OFF_SLOT);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // provably warp-uniform
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, WARPS + 1);   // 8 dequant warps (their LDS are done) + the UMMAs' commit (token rows read)
    }
    mbar_init(bar_red_full, WARPS);
    mbar_init(bar_red_empty, 1);
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar_a_full + 8 * b, CH == 4 ? WARPS : WARPS / 2);   // the warps that write a hand-over unit
      mbar_init(bar_a_empty + 8 * b, 1);
    }
    mbar_init(bar_d_full, 1);
    mbar_init(bar_done, WARPS);
    fence_barrier_init();
  }
  if (warp == WARPS + 2) tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TC::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot;
  pdl_launch_dependents();
  if (threadIdx.x == 0) dbg_stamp1(p, 0);

  const long long units = static_cast<long long>(p.ncols) * p.kblks;
  const int grid = gridDim.x;
  const int u0 = p.colpar ? static_cast<int>((static_cast<long long>(blockIdx.x) * p.ncols) / grid) * p.kblks
                          : unit_begin(blockIdx.x, units, grid);
  const int u1 = p.colpar ? static_cast<int>((static_cast<long long>(blockIdx.x + 1) * p.ncols) / grid) * p.kblks
                          : unit_begin(blockIdx.x + 1, units, grid);
  const int my_units = u1 - u0;
  if (my_units == 0) {
    if (warp == WARPS + 2) tmem_dealloc(tm, TC::TMEM_COLS);
    return;
  }

  if (warp == WARPS) {
    // ===================================== producer =====================================
    if (elect_one()) {
      tma_prefetch_desc(&tm_w);
      tma_prefetch_desc(&tm_a);
      if (GROUPED) tma_prefetch_desc(&tm_s);
      auto issue_w = [&](int s, int c, int b) {
        const uint32_t full = bar_full + 8 * s;
        mbar_arrive_expect_tx(full, Cfg::STAGE_TX);
        tma_load_3d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_W, &tm_w, full, 0, c * CHUNKS, b * KROWS, kEvictFirst);
        if (GROUPED)
          tma_load_3d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_S, &tm_s, full, 0, c * CHUNKS, (b * KB) / 128, kEvictLast);
      };
      auto issue_a = [&](int s, int b) {
        const uint32_t full = bar_full + 8 * s;
#pragma unroll
        for (int kc = 0; kc < Cfg::KC; ++kc)
          tma_load_2d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_A + kc * Cfg::A_SUB, &tm_a, full, b * KB + kc * 64, 0,
                      kEvictLast);
      };
      const SegWalk walk(u0, u1, p.kblks);
      const int npre = my_units < NS ? my_units : NS;
      int si = 0, c, b, bend;
      walk.get(0, c, b, bend);
      auto advance = [&]() {
        if (++b == bend && si + 1 < walk.nseg) walk.get(++si, c, b, bend);
      };
      for (int it = 0; it < npre; ++it) {   // weights do not depend on the prior grid
        issue_w(it, c, b);
        advance();
      }
      pdl_wait_prior_grids();               // the token rows do
      dbg_stamp1(p, 1);
      {
        int sj = 0, cc, bb, be;
        walk.get(0, cc, bb, be);
        for (int it = 0; it < npre; ++it) {
          issue_a(it, bb);
          if (++bb == be && sj + 1 < walk.nseg) walk.get(++sj, cc, bb, be);
        }
      }
      int s = npre % NS;
      uint32_t ph = (npre == NS) ? 0u : 1u;
      for (int it = npre; it < my_units; ++it) {
        mbar_wait(bar_empty + 8 * s, ph);
        issue_w(s, c, b);
        issue_a(s, b);
        advance();
        if (++s == NS) { s = 0; ph ^= 1u; }
      }
    }
    return;
  }

  if (warp == WARPS + 1) {
    // ======================================= flusher =======================================
    pdl_wait_prior_grids();  // C, c_tmp and the counters are shared with prior grids
    const SegWalk walk(u0, u1, p.kblks);
    for (int si = 0; si + 1 < walk.nseg; ++si) {
      int c, b0, b1;
      walk.get(si, c, b0, b1);
      mbar_wait_backoff(bar_red_full, static_cast<uint32_t>(si) & 1u, 64);
      finish_segment<T, TN, TNP, Cfg::FLAGS_PER_TILE, GROUPED, false>(p, red, c, /*whole=*/b0 == 0 && b1 == p.kblks, units,
                                                                      grid, lane, 32, WarpSync(), tk, nullptr, 0);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_red_empty);
    }
    return;
  }

  if (warp == WARPS + 2) {
    // ================================== UMMA issuer ==================================
    // per half-stage (4 k16-rows): 8 UMMAs of 128 columns x NROWS tokens x 16 k, A from tensor memory, B = the token
    // rows of the stage; commits free the A buffer, the ring slot (second half) and publish D (segment end)
    constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO = 8 rows x 128 B | version | SWIZZLE_128B
    const uint32_t idesc = umma_idesc_f16<T>(128, NROWS);
    const uint32_t b_lo0 = (((sbase + Cfg::ST_A) & 0x3FFFFu) >> 4) | (1u << 16);
    const SegWalk walk(u0, u1, p.kblks);
    uint32_t s = 0, ph = 0, it = 0;
    for (int si = 0; si < walk.nseg; ++si) {
      int seg_c, seg_b0, seg_b1;
      walk.get(si, seg_c, seg_b0, seg_b1);
#pragma unroll 1
  #pragma unroll 1
    for (int ub = seg_b0; ub < seg_b1; ++ub, ++it) {
        mbar_wait(bar_full + 8 * s, ph);   // the token rows of the stage have landed
#pragma unroll
        for (int h = 0; h < HALVES; ++h) {
          const uint32_t b = h & 1;
          const uint32_t use = it * (HALVES / 2) + (h >> 1);   // how often buffer b has been filled before
          mbar_wait(bar_a_full + 8 * b, use & 1u);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t b_lo = b_lo0 + s * (Cfg::STAGE_BYTES >> 4) + h * (Cfg::A_SUB >> 4);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const uint64_t bdesc = (static_cast<uint64_t>(kDescHi) << 32) | (b_lo + r * 2);
              const uint32_t accum = (ub != seg_b0 || h != 0 || r != 0) ? 1u : 0u;
#pragma unroll
              for (int sub = 0; sub < SUBS; ++sub)
                tc_mma_f16_ts(tm + sub * NROWS, tm + TC::TM_A + b * TC::A_BUF + sub * 32 + r * 8, bdesc, idesc, accum);
            }
            tc_commit(bar_a_empty + 8 * b);
            if (h == HALVES - 1) {
              tc_commit(bar_empty + 8 * s);
              if (ub == seg_b1 - 1) tc_commit(bar_d_full);
            }
          }
          __syncwarp();
        }
        if (++s == NS) { s = 0; ph ^= 1u; }
      }
    }
    mbar_wait(bar_done, 0);   // every dequant warp has read its last accumulator
    tc_fence_after();
    tmem_dealloc(tm, TC::TMEM_COLS);
    return;
  }

  // ====================================== dequantisers ======================================
  // warp w: TMEM lane quarter q = w % 4; grp = w / 4 is the 128-column half of a 256-column tile (CH == 4: every
  // hand-over unit of the stage) or, for the 128-column tile, the parity of the hand-over units it converts (CH == 2)
  const int q = warp & 3;
  const int grp = warp >> 2;
  const int g = lane >> 2;
  const int sub = CH == 4 ? grp : 0;
  const int chunk = 2 * sub + (q >> 1);
  const int jb = 2 * (q & 1);   // first of the two words (16-column blocks) of the chunk this warp converts
  // per-lane offsets inside a stage block: 8 bytes (words jb, jb + 1) of k16-row r's 512-byte block of `chunk`; the
  // matching two scale words (word j = {s[16j + g], s[16j + g + 8]}, common.cuh)
  const uint32_t q_off = Cfg::ST_W + static_cast<uint32_t>(chunk * 512 + lane * 16 + jb * 4);
  const uint32_t sc_off = Cfg::ST_S + static_cast<uint32_t>(chunk * 128 + g * 16 + jb * 4);
  // tensor-memory address of this warp's A rows: lane 32q (+ 16 for the second word), column TM_A + 32 * sub
  const uint32_t ta0 = tm + (static_cast<uint32_t>(32 * q) << 16) + TC::TM_A + 32 * sub;
  constexpr int MYH = CH == 4 ? HALVES : HALVES / 2;   // hand-over units this warp converts per stage
  constexpr int DCOLS = CH == 4 ? NROWS : NROWS / 2;   // accumulator columns (tokens) this warp reads back
  const uint32_t td = tm + (static_cast<uint32_t>(32 * q) << 16) + (CH == 4 ? sub * NROWS : grp * DCOLS);

  const SegWalk walk(u0, u1, p.kblks);
  int s = 0;
  uint32_t ph = 0, it = 0;
  uint32_t stg = sbase, bfull = bar_full, bempty = bar_empty;

  mbar_wait_backoff(bar_full, 0, 64);  // under PDL this first wait spans the tail of the previous kernel
  if (threadIdx.x == 0) dbg_stamp1(p, 2);

  for (int si = 0; si < walk.nseg; ++si) {
    int seg_c, seg_b0, seg_b1;
    walk.get(si, seg_c, seg_b0, seg_b1);
#pragma unroll 1
    for (int ub = seg_b0; ub < seg_b1; ++ub, ++it) {
      mbar_wait(bfull, ph);
      uint2 qw[MYH][4];
#pragma unroll
      for (int i = 0; i < MYH; ++i) {
        const int h = CH == 4 ? i : 2 * i + grp;
#pragma unroll
        for (int r = 0; r < 4; ++r) qw[i][r] = lds64(stg + q_off + (4 * h + r) * (CHUNKS * 512));
      }
#pragma unroll
      for (int i = 0; i < MYH; ++i) {
        const int h = CH == 4 ? i : 2 * i + grp;     // hand-over unit: k16-rows 4h .. 4h + 3
        const uint32_t b = h & 1;
        const uint32_t use = it * (HALVES / 2) + (h >> 1);
        uint32_t sc[2] = {0u, 0u};
        if (GROUPED) {   // rows 8.. of a 256-k stage belong to the next scale group
          const uint2 sv = lds64(stg + sc_off + (Cfg::SG > 1 ? (h >> 1) * (TN * 2) : 0));
          sc[0] = sv.x;
          sc[1] = sv.y;
        }
        mbar_wait(bar_a_empty + 8 * b, (use & 1u) ^ 1u);   // the UMMAs that read this buffer last time are done
        tc_fence_after();
        const uint32_t ta = ta0 + b * TC::A_BUF;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const uint32_t two[2] = {qw[i][r].x, qw[i][r].y};
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {
            uint32_t w[4];
            Dequant<T>::template word<GROUPED>(two[jj], sc[jj], w);
            tmem_st_16x128b_x2(ta + (static_cast<uint32_t>(16 * jj) << 16) + r * 8, w[0], w[1], w[2], w[3]);
          }
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_a_full + 8 * b);
      }
      if (lane == 0) mbar_arrive(bempty);   // this warp's shared-memory reads of the stage are done (their values were used)
      stg += Cfg::STAGE_BYTES;
      bfull += 8;
      bempty += 8;
      if (++s == NS) {
        s = 0;
        ph ^= 1u;
        stg = sbase;
        bfull = bar_full;
        bempty = bar_empty;
      }
    }
    // ---- the segment is complete: accumulator -> fp32 tile [token][column] in shared memory ----
    mbar_wait(bar_red_empty, (static_cast<uint32_t>(si) & 1u) ^ 1u);  // previous tile consumed (passes the first time)
    mbar_wait(bar_d_full, static_cast<uint32_t>(si) & 1u);            // every UMMA of the segment has completed
    tc_fence_after();
    {
      uint32_t v[DCOLS];
      tmem_ld_cols<DCOLS>(td, v);
      tmem_ld_wait();
      const int n = 128 * sub + 32 * q + lane;
      const int tok0 = CH == 4 ? 0 : grp * DCOLS;
#pragma unroll
      for (int i = 0; i < DCOLS; ++i)
        if (tok0 + i < p.m) red[(tok0 + i) * TNP + n] = __uint_as_float(v[i]);
    }
    tc_fence_before();
    if (si == walk.nseg - 1) {
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_done);   // tensor memory may be returned
      if (threadIdx.x == 0) dbg_stamp1(p, 3);
      named_bar_sync(1, NCONS);
      pdl_wait_prior_grids();  // C, c_tmp and the counters are shared with prior grids
      finish_segment<T, TN, TNP, Cfg::FLAGS_PER_TILE, GROUPED, false>(
          p, red, seg_c, /*whole=*/seg_b0 == 0 && seg_b1 == p.kblks, units, grid, static_cast<int>(threadIdx.x), NCONS,
          ConsumerSync<NCONS>(), tk + 1, reinterpret_cast<float4*>(smem), (NS * Cfg::STAGE_BYTES) / 16);
      if (threadIdx.x == 0) dbg_stamp1(p, 7);
    } else {
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_red_full);
    }
  }
}

}  // namespace b200
