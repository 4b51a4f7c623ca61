// Tensor-bound regime (chunked prefill, and every M above the streaming kernel's range): a persistent,
// warp-specialised tcgen05 kernel.  Operand roles are SWAPPED relative to the reference's Marlin
// kernel: the dequantised WEIGHT tile is the 128-row UMMA "A" operand (M side = output columns n)
// and the TOKENS are the UMMA "B" operand (N side, up to 256 per instruction), so one dequantised
// tile feeds 128 x BM x 64 MACs and the accumulator D^T[n][token] lives in tensor memory.
//
//   warp 6      TMA producer: raw Marlin int4 boxes + group scales (ring Q, deep: the HBM stream)
//   warp 4      TMA producer: activation boxes [BM tokens x 64 k], 128B-swizzled, rows past M
//               zero-filled (ring A)
//   warps 7..   dequantisers: LDS.128 of a lane's 16 B -> lop3/HSUB2/HFMA2 (exact) -> HMUL2 by the
//               group scale (the reference's single rounding point, marlin_device_ops.cuh:294-303)
//               -> stmatrix.x4 straight into the canonical K-major SWIZZLE_128B UMMA layout: the
//               Marlin fragment order IS the stmatrix 8x8 fragment order, so a 32-bit word of the
//               stream becomes four 16-byte core-matrix rows with no shuffles (ring W)
//   warp 5      one elected lane issues tcgen05.mma (kind::f16, fp32 accumulate in TMEM);
//               tcgen05.commit releases the A/W slots and publishes the accumulator
//   warps 0..3  epilogue: tcgen05.ld (lane = output column n), RN to T, the operator's post-ops
//               with the reference's rounding points, stores to C[token][n]
// Tiles (nt, mt) are walked persistently, mt fastest, one CTA per SM; the accumulator is double
// buffered in TMEM when it fits (<= 256 columns) so the epilogue overlaps the next tile.
#pragma once
#ifndef B200_NSUB2_SLACK
#define B200_NSUB2_SLACK 0   // diagnosis of the withdrawn 256-column single-CTA tiles
#endif
#include "common.cuh"

namespace b200 {

struct PrefillArgs {
  void* c;               // T[M][N]
  const void* scales;    // raw pointer (channelwise epilogue)
  const void* bias;      // T[N] or null
  const void* residual;  // T[M][N] or null (may alias c)
  uint32_t out_scale;    // 0, or 0x10000 | bits of T(scale): last post-op
  int m, n, k;
  int n_tiles, m_tiles, kblks;
  int qkv_a, qkv_b;      // > 0: q | k | v written split (out_offset, common.cuh)
};

// PAIR: cta_group::2 — a cluster of two CTAs works on 256 weight columns x BM tokens: each CTA dequantises its own 128
// columns and stages HALF of the token tile, so the activation operand crosses the L2 -> SM fabric once per 256
// columns instead of once per 128 (what bounds the single-CTA tiles) with the accumulator still double
// buffered in each CTA's tensor memory.
template <typename T, int NSUB, int BM, int STAGES, int QSTAGES, int DW, bool GROUPED, bool SPLIT = false, bool PAIR = false>
struct PrefillCfg {
  static constexpr int BN = NSUB * 128;            // output columns per CTA tile
  static constexpr int CHUNKS = NSUB * 2;          // 64-column Marlin chunks per tile
  static constexpr int BK = 64;                    // k per stage = one 128-byte swizzled row
  static constexpr int A_ROWS = PAIR ? BM / 2 : BM;  // token rows this CTA stages (PAIR: its half; the box is always A_ROWS tall)
  static constexpr int A_STAGE = A_ROWS * 128;     // activations  [A_ROWS][64] T
  static constexpr int W_STAGE = BN * 128;         // dequantised  [BN][64] T
  static constexpr int Q_STAGE = 4 * CHUNKS * 512; // raw int4: 4 k16-rows x CHUNKS blocks of 512 B
  static constexpr int S_STAGE = GROUPED ? BN * 2 : 0;
  static constexpr int ACC_COLS = NSUB * BM;       // fp32 TMEM columns per accumulator
  static constexpr int ACC_STAGES = (!SPLIT && (512 / ACC_COLS) >= 2) ? 2 : 1;
  static constexpr int TMEM_COLS = ACC_COLS * ACC_STAGES;  // a power of two >= 32
  // split-K: fp32 partial tile [BM tokens][128 + 4] staged for the cluster reduction; it aliases the
  // A/W rings, which are dead once the last MMA of the CTA's only tile has been committed
  static constexpr int RED_PITCH = 128 + 4;
  static constexpr int RED_BYTES = SPLIT ? BM * RED_PITCH * 4 : 0;
  // dequant warps work in DGROUPS groups that take alternate k-blocks (group g: every k-block whose
  // running index has parity g), so a warp converts more k16-rows per barrier round trip
  static constexpr int DGROUPS =
      ((DW / 2) >= CHUNKS && (DW / 2) % CHUNKS == 0 && STAGES >= 2 && QSTAGES >= 2) ? 2 : 1;
  static constexpr int DWG = DW / DGROUPS;         // dequant warps per group = arrivals per stage
  static constexpr int WPC = DWG / CHUNKS;         // dequant warps of a group sharing a chunk
  static constexpr int RPW = 4 / WPC;              // k16-rows per dequant warp per stage
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W = OFF_A + STAGES * A_STAGE;
  static constexpr int OFF_Q = OFF_W + STAGES * W_STAGE;
  static constexpr int OFF_S = OFF_Q + QSTAGES * Q_STAGE;
  static constexpr int OFF_BAR = (OFF_S + QSTAGES * S_STAGE + 127) / 128 * 128;
  static constexpr int NBARS = 3 * STAGES + 2 * QSTAGES + 4 + (PAIR ? STAGES : 0);
  static constexpr int SMEM_BYTES = OFF_BAR + NBARS * 8 + 16;
  static constexpr int THREADS = (7 + DW) * 32;
  static_assert(DWG % CHUNKS == 0 && WPC >= 1 && 4 % WPC == 0, "dequant warp layout");
  static_assert(DGROUPS <= STAGES && DGROUPS <= QSTAGES, "the dequant groups walk the rings with stride DGROUPS (at most one wrap per step)");
  static_assert(TMEM_COLS >= 32 && TMEM_COLS <= 512 && (TMEM_COLS & (TMEM_COLS - 1)) == 0, "tensor memory allocation");
  static_assert(!SPLIT || (NSUB == 1 && RED_BYTES <= OFF_BAR), "split-K staging must fit in the dead rings (A, W and Q are all drained when the accumulator is complete)");
  static_assert(BM % 16 == 0 && BM <= 256, "UMMA N");
  static_assert(!PAIR || (NSUB == 1 && !SPLIT && BM % 32 == 0), "cta_group::2: one 128-row sub-tile per CTA, whole K");
  static_assert(SMEM_BYTES <= 232448, "shared memory budget (227 KB)");
};

template <typename T, int NSUB, int BM, int STAGES, int QSTAGES, int DW, bool GROUPED, bool SPLIT, bool PAIR = false>
__global__ void __launch_bounds__((7 + DW) * 32, SPLIT ? 2 : 1)
w4a16_prefill_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_s,
                     const __grid_constant__ CUtensorMap tm_a, const PrefillArgs p) {
  using Cfg = PrefillCfg<T, NSUB, BM, STAGES, QSTAGES, DW, GROUPED, SPLIT, PAIR>;
  constexpr int CHUNKS = Cfg::CHUNKS;
  const int prank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;   // 0: leader (issues the UMMAs)

  // no static shared memory in this kernel: the dynamic window starts 1024-aligned (SWIZZLE_128B operand
  // tiles need it); checked, not assumed
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0u) __trap();
  const uint32_t bar0 = sbase + Cfg::OFF_BAR;
  const uint32_t bar_a_full = bar0;
  const uint32_t bar_w_full = bar_a_full + STAGES * 8;
  const uint32_t bar_consumed = bar_w_full + STAGES * 8;
  const uint32_t bar_q_full = bar_consumed + STAGES * 8;
  const uint32_t bar_q_empty = bar_q_full + QSTAGES * 8;
  const uint32_t bar_acc_full = bar_q_empty + QSTAGES * 8;
  const uint32_t bar_acc_empty = bar_acc_full + 2 * 8;
  const uint32_t bar_w_peer = bar_acc_empty + 2 * 8;   // PAIR, leader's copy: the peer CTA's half of the W stage is written
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + Cfg::OFF_BAR + Cfg::NBARS * 8);

  // warp index through a shuffle: provably warp-uniform for ptxas, so the role branches are uniform branches and the
  // single-lane roles (UMMA issuer, TMA producers) keep their loop state and descriptors in uniform registers
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_a_full + 8 * s, 1);
      mbar_init(bar_w_full + 8 * s, Cfg::DWG);
      mbar_init(bar_consumed + 8 * s, 1);
      if (PAIR) mbar_init(bar_w_peer + 8 * s, 1);
    }
#pragma unroll
    for (int s = 0; s < QSTAGES; ++s) {
      mbar_init(bar_q_full + 8 * s, 1);
      mbar_init(bar_q_empty + 8 * s, Cfg::DWG);
    }
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar_acc_full + 8 * a, 1);
      mbar_init(bar_acc_empty + 8 * a, PAIR ? 8 : 4);
    }
    fence_barrier_init();
  }
  if (warp == 5) {
    if (PAIR) tmem_alloc_pair(smem_u32(const_cast<uint32_t*>(tmem_slot)), Cfg::TMEM_COLS);
    else tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), Cfg::TMEM_COLS);
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();   // both CTAs' barriers are initialised before either one arrives remotely
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // PDL: the next kernel may start its prologue now; of THIS kernel only the activation producer and
  // the epilogue touch memory the prior grid writes (A, C, residual), so only they wait for it — the
  // int4 stream, the dequantisers and the W ring run ahead during the previous kernel's tail
  pdl_launch_dependents();

  const int total_tiles = p.n_tiles * p.m_tiles;
  // work items: persistent walk over whole-K tiles, or (split-K) exactly one tile and one K slice
  // per CTA, the S CTAs of a cluster sharing the tile
  int tile_begin = blockIdx.x, tile_step = gridDim.x, kb_lo = 0, kb_hi = p.kblks;
  if (PAIR) {   // the two CTAs of a pair walk the same (256-column, token) tiles
    tile_begin = blockIdx.x >> 1;
    tile_step = gridDim.x >> 1;
  }
  int csize = 1, crank = 0;
  if (SPLIT) {
    csize = static_cast<int>(cluster_nctarank());
    crank = static_cast<int>(cluster_ctarank());
    tile_begin = blockIdx.x / csize;
    tile_step = total_tiles;
    kb_lo = static_cast<int>((static_cast<long long>(crank) * p.kblks) / csize);
    kb_hi = static_cast<int>((static_cast<long long>(crank + 1) * p.kblks) / csize);
  }

  if (warp == 4) {
    // ============================== TMA producer: activations (ring A) ==============================
    if (elect_one()) {
      tma_prefetch_desc(&tm_a);
      pdl_wait_prior_grids();
      int it = 0;
      for (int tile = tile_begin; tile < total_tiles; tile += tile_step) {
        const int mt = tile % p.m_tiles;
        for (int kb = kb_lo; kb < kb_hi; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait_backoff(bar_consumed + 8 * s, ((it / STAGES) & 1) ^ 1, 32);
          if (PAIR) {
            // one k-block of slack before the token slot is refilled (the dequantisers take two before they rewrite a
            // W slot, see there): the TMA data lands >= 0.6 us after the request, which already covered the window — the
            // refill never showed the race — and one k-block doubles that margin
#pragma unroll
            for (int d = 1; d <= 1; ++d) {
              const int it1 = it + d - STAGES;
              if (it1 >= 0) mbar_wait_backoff(bar_consumed + 8 * (it1 % STAGES), (it1 / STAGES) & 1, 32);
            }
            // tokens [m0 + rank * h, + h) with h = half of the tile's valid tokens rounded up to 16: the UMMA reads h rows
            // from each CTA, so accumulator column j is token m0 + j for ragged tiles too
            int vt = p.m - mt * BM;
            vt = vt > BM ? BM : vt;
            const int h = ((vt + 15) & ~15) >> 1;
            if (prank == 0) mbar_arrive_expect_tx(bar_a_full + 8 * s, 2 * Cfg::A_STAGE);
            tma_load_2d_pair(sbase + Cfg::OFF_A + s * Cfg::A_STAGE, &tm_a, bar_a_full + 8 * s, kb * 64, mt * BM + prank * h,
                             kEvictLast);
          } else {
            mbar_arrive_expect_tx(bar_a_full + 8 * s, Cfg::A_STAGE);
            tma_load_2d(sbase + Cfg::OFF_A + s * Cfg::A_STAGE, &tm_a, bar_a_full + 8 * s, kb * 64, mt * BM,
                        kEvictLast);
          }
        }
      }
    }
  } else if (warp == 6) {
    // ===================== TMA producer: raw int4 weights + group scales (ring Q) =====================
    // its own lane and its own, deeper ring: the HBM stream runs QSTAGES boxes ahead of the dequantisers
    // independently of the activation / tensor-core pipeline
    if (elect_one()) {
      tma_prefetch_desc(&tm_w);
      if (GROUPED) tma_prefetch_desc(&tm_s);
      int it = 0;
      for (int tile = tile_begin; tile < total_tiles; tile += tile_step) {
        const int nt = tile / p.m_tiles;
        for (int kb = kb_lo; kb < kb_hi; ++kb, ++it) {
          const int sq = it % QSTAGES;
          mbar_wait_backoff(bar_q_empty + 8 * sq, ((it / QSTAGES) & 1) ^ 1, 32);
          mbar_arrive_expect_tx(bar_q_full + 8 * sq, Cfg::Q_STAGE + Cfg::S_STAGE);
          const int ch0 = PAIR ? (2 * nt + prank) * CHUNKS : nt * CHUNKS;   // PAIR: nt counts 256-column tiles
          tma_load_3d(sbase + Cfg::OFF_Q + sq * Cfg::Q_STAGE, &tm_w, bar_q_full + 8 * sq, 0, ch0,
                      kb * 4, SPLIT ? kEvictFirst : kEvictNormal);
          if (GROUPED)
            tma_load_3d(sbase + Cfg::OFF_S + sq * Cfg::S_STAGE, &tm_s, bar_q_full + 8 * sq, 0,
                        ch0, kb >> 1, kEvictLast);
        }
      }
    }
  } else if (warp == 5 && prank == 0) {
    // ======================================= MMA issuer =======================================
    // (PAIR: the leader CTA only; one instruction drives both SMs' tensor cores.)  The warp stays convergent; one
    // elected lane issues.  Everything the UMMAs take is warp-uniform by construction (ring slot, descriptors = base +
    // slot * stride, TMEM address through a shuffle), so a k-block is 2 waits + 4 UTCHMMA + 1 commit and a few uniform
    // adds — the round-2 profile (profiles/r02/i_*) had this loop at ~140 instructions per k-block with a
    // register -> uniform-register waterfall before every UMMA, which capped the tensor pipe at 54 %.
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO = 8 rows x 128 B | version | SWIZZLE_128B
    const uint32_t a_lo0 = (((sbase + Cfg::OFF_A) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t w_lo0 = (((sbase + Cfg::OFF_W) & 0x3FFFFu) >> 4) | (1u << 16);
    uint32_t s = 0, ph = 0;
    int acc_it = 0;
    for (int tile = tile_begin; tile < total_tiles; tile += tile_step, ++acc_it) {
      const int mt = tile % p.m_tiles;
      const int acc = acc_it % Cfg::ACC_STAGES;
      mbar_wait(bar_acc_empty + 8 * acc, ((acc_it / Cfg::ACC_STAGES) & 1) ^ 1);
      tc_fence_after();
      int vt = p.m - mt * BM;
      vt = vt > BM ? BM : vt;
      const uint32_t idesc = umma_idesc_f16<T>(PAIR ? 256 : 128, static_cast<uint32_t>((vt + 15) & ~15));
      const uint32_t d0 = tm + acc * Cfg::ACC_COLS;
      for (int kb = kb_lo; kb < kb_hi; ++kb) {
        mbar_wait(bar_a_full + 8 * s, ph);
        mbar_wait(bar_w_full + 8 * s, ph);
        if (PAIR) mbar_wait(bar_w_peer + 8 * s, ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_lo = a_lo0 + s * (Cfg::A_STAGE >> 4);
          const uint32_t w_lo = w_lo0 + s * (Cfg::W_STAGE >> 4);
#pragma unroll
          for (int sub = 0; sub < NSUB; ++sub) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t adesc = (static_cast<uint64_t>(kDescHi) << 32) | (w_lo + sub * 1024 + kk * 2);
              const uint64_t bdesc = (static_cast<uint64_t>(kDescHi) << 32) | (a_lo + kk * 2);
              const uint32_t accum = (kk != 0 || kb != kb_lo) ? 1u : 0u;
              if (PAIR) tc_mma_f16_pair(d0 + sub * BM, adesc, bdesc, idesc, accum);
              else tc_mma_f16(d0 + sub * BM, adesc, bdesc, idesc, accum);
            }
          }
          if (PAIR) {   // frees the A / W slots and publishes the accumulator in BOTH CTAs
            tc_commit_pair(bar_consumed + 8 * s);
            if (kb == kb_hi - 1) tc_commit_pair(bar_acc_full + 8 * acc);
          } else {
            tc_commit(bar_consumed + 8 * s);
            if (kb == kb_hi - 1) tc_commit(bar_acc_full + 8 * acc);
          }
        }
        __syncwarp();
        if (++s == STAGES) {
          s = 0;
          ph ^= 1u;
        }
      }
    }
  } else if (PAIR && warp == 5) {
    // ============ PAIR, peer CTA: forwards "my half of W stage s is written" to the leader ============
    // The dequantisers arrive on their OWN CTA's w_full — a local arrive travels the same shared-memory pipeline as
    // their stmatrix stores, so the phase cannot complete before the stores have landed — and this otherwise idle warp
    // turns each completed phase into one plain remote arrive on the leader's w_peer.  Arriving remotely straight
    // from the dequant warps was a race: the remote arrive takes the SM-to-SM network and, about once in 2700
    // launches, reached the leader before the peer's last stores had reached its shared memory, so the pair's UMMA
    // read a stale half-stage (profiles/r02/o_pair_tile_race.txt); the `.release.cluster` form is safe but puts a
    // ~1 us cluster-scope fence in front of every arrive.
    int my_tiles = 0;
    for (int tile = tile_begin; tile < total_tiles; tile += tile_step) ++my_tiles;
    const int total_its = my_tiles * (kb_hi - kb_lo);
    uint32_t s = 0, ph = 0;
    for (int it = 0; it < total_its; ++it) {
      mbar_wait(bar_w_full + 8 * s, ph);
      if (elect_one()) mbar_arrive_leader(bar_w_peer + 8 * s);
      __syncwarp();
      if (++s == STAGES) {
        s = 0;
        ph ^= 1u;
      }
    }
  } else if (warp >= 7) {
    // ====================================== dequantisers ======================================
    const int d = warp - 7;
    const int grp = d / Cfg::DWG;   // dequant group: takes the k-blocks whose running index has this parity
    const int dl = d % Cfg::DWG;
    const int ch = dl % CHUNKS;     // 64-column chunk of the tile
    const int rp = dl / CHUNKS;     // k16-row phase
    const int g = lane >> 2;
    // stmatrix row address of this lane for word j of k16-row r = rp | WPC*i:
    //   matrix (lane>>3): rows +8 for odd matrices, k half (lane>>4); row low bits = lane & 7
    //   tile row nrow = ch*64 + j*16 + lrow -> (nrow>>7)*16384 + (nrow&127)*128, 16-byte k chunk
    //   c16 = 2r + lkh XOR-swizzled with nrow&7 = lrow&7: a lane constant, XOR (2*WPC*i)<<4, plus j*2048
    const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8;
    const int lkh = lane >> 4;
    const uint32_t w_lane = static_cast<uint32_t>((ch >> 1) * 16384 + ((ch & 1) * 64 + lrow) * 128 +
                                                  ((((2 * rp) | lkh) ^ (lrow & 7)) << 4));
    const uint32_t q_lane = static_cast<uint32_t>((rp * CHUNKS + ch) * 512 + lane * 16);
    const uint32_t s_lane = static_cast<uint32_t>(ch * 128 + g * 16);
    int my_tiles = 0;
    for (int tile = tile_begin; tile < total_tiles; tile += tile_step) ++my_tiles;
    const int total_its = my_tiles * (kb_hi - kb_lo);
    int sq = grp % QSTAGES, s = grp % STAGES;
    uint32_t qph = 0, wph = 1;  // parities: raw slot filled; W slot drained (passes the first lap)
    for (int it = grp; it < total_its; it += Cfg::DGROUPS) {
      mbar_wait(bar_q_full + 8 * sq, qph);
      uint32_t sc[4] = {0, 0, 0, 0};
      if (GROUPED) {
        const uint4 sv = lds128(sbase + Cfg::OFF_S + sq * Cfg::S_STAGE + s_lane);
        sc[0] = sv.x; sc[1] = sv.y; sc[2] = sv.z; sc[3] = sv.w;
      }
      uint4 q[Cfg::RPW];
#pragma unroll
      for (int i = 0; i < Cfg::RPW; ++i)
        q[i] = lds128(sbase + Cfg::OFF_Q + sq * Cfg::Q_STAGE + q_lane + i * (Cfg::WPC * CHUNKS * 512));
      mbar_wait(bar_consumed + 8 * s, wph);  // W slot drained by the tensor core
      if (PAIR || (NSUB == 2 && B200_NSUB2_SLACK)) {
        // CTA pairs: one more k-block of slack.  The multicast commit that frees slot s arrives in BOTH CTAs at one
        // point in time while the two tensor cores do not finish reading their shared memory at the same instant; a
        // dequantiser that rewrites its W slot right behind that arrive corrupted a half-stage about once in 2700
        // launches (profiles/r02/o_pair_tile_race.txt; the TMA refill of the token slot is a microsecond later and
        // never showed it).  The UMMAs of a tensor core complete in order, so once the NEXT k-block's commit has
        // arrived this slot's reads are over on both SMs (two k-blocks: one left 1 mismatch in 32000 launches).
        constexpr int kSlack = PAIR ? 2 : 1;
#pragma unroll
        for (int d = 1; d <= kSlack; ++d) {
          const int it1 = it + d - STAGES;   // the d-th k-block after the one that last used slot s
          if (it1 >= 0 && it1 < total_its) mbar_wait(bar_consumed + 8 * (it1 % STAGES), (it1 / STAGES) & 1);
        }
      }
      const uint32_t w_addr = sbase + Cfg::OFF_W + s * Cfg::W_STAGE;
#pragma unroll
        for (int i = 0; i < Cfg::RPW; ++i) {
          const uint32_t qq[4] = {q[i].x, q[i].y, q[i].z, q[i].w};
          const uint32_t row_addr = w_addr + (w_lane ^ static_cast<uint32_t>((2 * Cfg::WPC * i) << 4));
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t w[4];
            Dequant<T>::template word<GROUPED>(qq[j], sc[j], w);
            // w[0]: n=g, k 0..7 | w[1]: n=g+8, k 0..7 | w[2]: n=g, k 8..15 | w[3]: n=g+8, k 8..15
            stmatrix_x4(row_addr + j * 2048, w[0], w[1], w[2], w[3]);
          }
        }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async proxy
      __syncwarp();
      // the raw slot is released only after its words were USED (the dequantised tile is written): the
      // shared-memory loads are then complete by data dependence, not by an assumption about issue order
      // (releasing right after ISSUING the loads was the round-1 split-K race)
      if (lane == 0) mbar_arrive(bar_q_empty + 8 * sq);
      if (lane == 0) mbar_arrive(bar_w_full + 8 * s);   // PAIR, peer CTA: its warp 5 forwards the completed phase to the leader
      sq += Cfg::DGROUPS;
      if (sq >= QSTAGES) { sq -= QSTAGES; qph ^= 1u; }
      s += Cfg::DGROUPS;
      if (s >= STAGES) { s -= STAGES; wph ^= 1u; }
    }
  } else if (warp < 4) {
    // ======================================== epilogue ========================================
    const int qd = warp;  // TMEM lane quarter: this warp may touch lanes 32*qd .. 32*qd+31
    pdl_wait_prior_grids();
    int acc_it = 0;
    for (int tile = tile_begin; tile < total_tiles; tile += tile_step, ++acc_it) {
      const int mt = tile % p.m_tiles, nt = tile / p.m_tiles;
      const int acc = acc_it % Cfg::ACC_STAGES;
      mbar_wait_backoff(bar_acc_full + 8 * acc, (acc_it / Cfg::ACC_STAGES) & 1, 256);
      tc_fence_after();
      const int m0 = mt * BM;
      int vt = p.m - m0;
      vt = vt > BM ? BM : vt;
      T* cbase = reinterpret_cast<T*>(p.c);
      const T* rbase = reinterpret_cast<const T*>(p.residual);
#pragma unroll 1
      for (int sub = 0; sub < NSUB; ++sub) {
        const int n = (PAIR ? (2 * nt + prank) * 128 : nt * Cfg::BN + sub * 128) + qd * 32 + lane;
        const bool nok = n < p.n;
        T bias_v = Elem<T>::from_float(0.f), chan_s = Elem<T>::from_float(1.f);
        if (nok && p.bias) bias_v = reinterpret_cast<const T*>(p.bias)[n];
        if (!GROUPED && nok) chan_s = reinterpret_cast<const T*>(p.scales)[chan_scale_pos(n)];
        // this thread owns output column n for the whole tile: its base offset and token stride are tile constants
        // (q/k/v split included), and the post-op set is a kernel constant — the per-element work of the common case
        // is one conversion and one store (the round-2 profile had ~30 branchy instructions per element here, and the
        // issuer waiting for the accumulator 29 % of the time at K = 4096: profiles/r02/i_*)
        size_t col_off = 0;
        int tok_stride = 0;
        out_column(nok ? n : 0, p.n, p.m, p.qkv_a, p.qkv_b, &col_off, &tok_stride);
        T* cptr = cbase + col_off + static_cast<size_t>(m0) * tok_stride;
        const T* rptr = rbase ? rbase + col_off + static_cast<size_t>(m0) * tok_stride : nullptr;
        const bool plain = GROUPED && !rbase && !p.bias && p.out_scale == 0u;
#pragma unroll 1
        for (int c0 = 0; c0 < vt; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(qd * 32) << 16) + acc * Cfg::ACC_COLS +
                            sub * BM + c0,
                        v);
          tmem_ld_wait();
          if (SPLIT) {
            // fp32 partial -> staging tile [token][n]; lanes write consecutive n: conflict-free
            float* red = reinterpret_cast<float*>(smem + Cfg::OFF_A);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              red[(c0 + i) * Cfg::RED_PITCH + qd * 32 + lane] = __uint_as_float(v[i]);
          } else if (nok) {
            T* cp = cptr + static_cast<size_t>(c0) * tok_stride;
            if (plain && c0 + 32 <= vt) {
#pragma unroll
              for (int i = 0; i < 32; ++i) cp[static_cast<size_t>(i) * tok_stride] = Elem<T>::from_float(__uint_as_float(v[i]));
            } else {
              const T* rp = rptr ? rptr + static_cast<size_t>(c0) * tok_stride : nullptr;
              const int cnt = vt - c0;
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                if (i < cnt) {
                  T o = Elem<T>::from_float(__uint_as_float(v[i]));
                  if (!GROUPED) o = __hmul(o, chan_s);  // channelwise: scale the ROUNDED sum
                  if (rp) o = __hadd(rp[static_cast<size_t>(i) * tok_stride], o);
                  if (p.bias) o = __hadd(o, bias_v);
                  o = apply_out_scale<T>(o, p.out_scale);
                  cp[static_cast<size_t>(i) * tok_stride] = o;
                }
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR) mbar_arrive_leader(bar_acc_empty + 8 * acc);
        else mbar_arrive(bar_acc_empty + 8 * acc);
      }
    }
  }

  if (SPLIT) {
    // ---- cluster reduction over the K slices, in rank order (bit-deterministic) ----
    cluster_sync_all();
    if (tile_begin < total_tiles && warp < 4) {
      const int mt = tile_begin % p.m_tiles, nt = tile_begin / p.m_tiles;
      const int m0 = mt * BM;
      int vt = p.m - m0;
      vt = vt > BM ? BM : vt;
      const int total = vt * 32;  // float4 items of the valid rows
      const int e0 = static_cast<int>((static_cast<long long>(crank) * total) / csize);
      const int e1 = static_cast<int>((static_cast<long long>(crank + 1) * total) / csize);
      const uint32_t red_addr = sbase + Cfg::OFF_A;
      for (int e = e0 + static_cast<int>(threadIdx.x); e < e1; e += 128) {
        const int tok = e >> 5, nn = (e & 31) * 4;
        const uint32_t off = static_cast<uint32_t>(tok * Cfg::RED_PITCH + nn) * 4u;
        float4 sum = ld_dsmem_f4(red_addr + off, 0);
        for (int r = 1; r < csize; ++r) {
          const float4 v = ld_dsmem_f4(red_addr + off, static_cast<uint32_t>(r));
          sum.x += v.x;
          sum.y += v.y;
          sum.z += v.z;
          sum.w += v.w;
        }
        const int n = nt * Cfg::BN + nn;
        if (n >= p.n) continue;
        T o[4] = {Elem<T>::from_float(sum.x), Elem<T>::from_float(sum.y), Elem<T>::from_float(sum.z),
                  Elem<T>::from_float(sum.w)};
        if (!GROUPED) {
          const T* sc = reinterpret_cast<const T*>(p.scales);
#pragma unroll
          for (int i = 0; i < 4; ++i) o[i] = __hmul(o[i], sc[chan_scale_pos(n + i)]);
        }
        const size_t goff = out_offset(m0 + tok, n, p.n, p.m, p.qkv_a, p.qkv_b);
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
    cluster_sync_all();  // no CTA exits while a peer may still read its staging tile
  }

  // teardown: every role is done with tensor memory before it is returned
  tc_fence_before();
  if (PAIR) cluster_sync_all();   // neither CTA returns tensor memory or exits while the pair's UMMAs / arrives may be in flight
  else __syncthreads();
  if (warp == 5) {
    if (PAIR) tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace b200
