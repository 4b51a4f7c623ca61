// HBM-bound regime (decode / speculative tree-verify, M <= 32 per launch): a PERSISTENT, warp-specialised
// weight-streaming kernel with stream-K work distribution.
//
//   * grid = one CTA per SM.  Work = (column tile c, k-block b) units in column-major order, split
//     evenly over the CTAs (|units_i - units_j| <= 1): every SM streams the same number of bytes, for
//     any shape — no wave quantisation, no cluster placement lottery.
//   * roles: 1 TMA producer lane, WARPS consumer warps, 1 flusher warp.
//       producer  feeds a STAGES-deep ring with cp.async.bulk.tensor: a [KROWS x CHUNKS] box of the
//                 Marlin int32 stream (evict-first), the group's scale row, the [M x KB] activation box
//                 (128B-swizzled, zero-filled past M).  Weights/scales of the first STAGES units are
//                 requested BEFORE griddepcontrol.wait (PDL): they do not depend on the prior grid.
//       consumers LDS.128 -> lop3/HADD2/HFMA2 (exact q-8) -> HMUL2 by the group scale (the reference's
//                 single rounding) -> mma.sync m16n8k16, weights as the 16-row operand, fp32 accumulate.
//                 At the end of a column segment the WK warps sharing a chunk reduce into one fp32 tile
//                 in shared memory (fixed order), hand it to the flusher through an mbarrier and go
//                 straight on to the next unit.
//       flusher   turns finished tiles into output OFF the consumers' critical path; the CTA's LAST
//                 segment is finished by the consumer warps themselves (nothing left to overlap with).
//   * a column whose K range is split over several CTAs is reduced WITHOUT ANY WAITING: every sharer
//     writes its fp32 partial to its slot in c_tmp, fences, and takes a ticket from one counter per
//     column in `workspace` (atomicAdd).  The sharer that draws the LAST ticket knows every other partial
//     is visible: it adds them IN SLOT ORDER (bit-deterministic, whoever happens to be last), applies the
//     epilogue, writes T and puts the counter back to zero — the reference's workspace contract
//     (zero on entry, zero on return; the reference serialises the CTAs of a column through a spin lock
//     instead, marlin_device_ops.cuh:347-377).  No CTA ever spins on another CTA, so the kernel makes
//     progress under ANY schedule: partial residency, other streams, MPS — nothing has to be co-resident.
//   * small CTAs (<= ~110 KB of shared memory, <= 96 registers x 320 threads) so that with PDL the NEXT
//     linear's CTA is co-resident on every SM and has its ring full when this kernel drains.
#pragma once
#include "common.cuh"

namespace b200 {

struct DecodeArgs {
  void* c;               // T[M][N]
  float* c_tmp;          // split-K slots
  int32_t* workspace;    // one arrival counter per column tile (first int of the tile's FLAGS_PER_TILE)
  const void* scales;    // raw pointer (channelwise epilogue)
  const void* bias;      // T[N] or null
  const void* residual;  // T[M][N] or null
  uint32_t out_scale;    // 0, or 0x10000 | bits of T(scale): last post-op
  int m, n, k;
  int ncols;      // column tiles = ceil(N / TN)
  int kblks;      // K / KB
  int slot_max;   // slots per column reserved in c_tmp
  int colpar;     // 1: CTAs take whole columns (no split columns, no reduction through c_tmp)
  int gated;      // 1: tile = [TN/2 gate columns | the same TN/2 up columns]; c[M][N/2] = silu(gate) * up
  int qkv_a, qkv_b;            // > 0: write the q | k | v column ranges split (out_offset, common.cuh); 0: row-major
  // fused all-reduce of a row-parallel linear over NVLink peer memory (ar_world > 1; see the "fused all-reduce" block)
  void* ar_recv[8];            // rank w's receive buffer T[2 parities][world sources][ar_slot_elems], peer-mapped
  int* ar_flag[8];             // rank w's arrival counters int[world sources][kArTiles], monotonic, peer-mapped
  int* ar_seen;                // local int[kArTiles + 1]: epochs consumed per column tile; last int = sticky error word
  long long ar_slot_elems;
  int ar_world, ar_rank;
  unsigned long long* dbg_ts;  // tuning only: per-CTA phase timestamps or null
  int lab;                     // B200_DECODE_LAB builds only (experiments; wrong results): 1 = activations fetched for the
                               // first ring fill only, 2 = scales likewise, 4 = consumers skip the dequant + MMA stream
};
#ifndef B200_DECODE_LAB
#define B200_DECODE_LAB 0
#endif
constexpr int kArTiles = 1024;   // column tiles a fused all-reduce can track (N <= 64 * 1024)

// ---- fused all-reduce (tensor-parallel row-parallel linears: o, down) ----------------------------------------
// The CTA that completes a column tile (whole owner or last ticket holder) does not write C: it rounds the tile to T
// and PUSHES it into every rank's receive buffer (slot = its own rank, buffer parity = the tile's epoch), then raises
// that tile's counter on every rank with a system-scope release.  After its last unit every CTA turns reducer for the
// tiles c = blockIdx.x, blockIdx.x + grid, ...: it waits for the world counters of the tile (pushes from OTHER GPUs
// and from local CTAs that never wait on anything themselves, so this cannot deadlock), adds the world partials out of
// LOCAL memory in rank order in fp32 (bit-identical on every rank), applies the epilogue and writes C.  One launch: the
// transfer of early tiles overlaps the rest of the stream, there is no separate collective kernel and no grid barrier.
// Two buffer parities suffice: a rank can only be one call ahead of a peer (its call i+1 completes only after the
// peer's pushes of call i+1, which follow the peer's whole call i in stream order).
__device__ __forceinline__ int ld_acquire_sys_s32(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_sys_u4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
// push 4 fp32 sums of (tok, columns n..n+3), rounded to T (channelwise: scaled as the reference scales the rounded sum),
// into slot `rank` of every peer's receive buffer
template <typename T, bool GROUPED>
__device__ __forceinline__ void ar_push(const DecodeArgs& p, int parity, int tok, int n, float4 v) {
  T o[4] = {Elem<T>::from_float(v.x), Elem<T>::from_float(v.y), Elem<T>::from_float(v.z), Elem<T>::from_float(v.w)};
  if (!GROUPED) {
    const T* sc = reinterpret_cast<const T*>(p.scales);
#pragma unroll
    for (int e = 0; e < 4; ++e) o[e] = __hmul(o[e], sc[chan_scale_pos(n + e)]);
  }
  uint2 pk;
  pk.x = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[0])) | (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[1])) << 16);
  pk.y = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[2])) | (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[3])) << 16);
  const size_t off = (static_cast<size_t>(parity) * p.ar_world + p.ar_rank) * p.ar_slot_elems + static_cast<size_t>(tok) * p.n + n;
#pragma unroll
  for (int w = 0; w < 8; ++w)
    if (w < p.ar_world) *reinterpret_cast<uint2*>(reinterpret_cast<T*>(p.ar_recv[w]) + off) = pk;
}

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
  static constexpr int SG = KB > 128 ? KB / 128 : 1;  // scale groups (rows of the scale box) per stage
  static constexpr int S_STAGE = GROUPED ? SG * TN * 2 : 0;
  static constexpr int STAGE_TX = A_STAGE + W_STAGE + S_STAGE;
  static constexpr int TNP = TN + 4;           // padded fp32 row of the reduction tile
  static constexpr int RED_BYTES = MT * TNP * 4;
  // ring depth: STAGES for M <= 8; one / two fewer for the wider token tiles, whose activation boxes and
  // reduction tile are bigger, so that the CTA stays under half an SM's shared memory (PDL co-residency)
  static constexpr int NSTAGE = (NT == 1) ? STAGES : ((NT == 2 ? STAGES - 1 : STAGES - 2) < 2 ? 2 : (NT == 2 ? STAGES - 1 : STAGES - 2));
  // ring layout: one contiguous block per stage, [activations | int4 words | scales], so that a consumer needs ONE
  // running stage pointer (every other offset is an immediate of the shared-memory instruction)
  static constexpr int ST_A = 0;                     // 1024-aligned: 128B-swizzled boxes
  static constexpr int ST_W = ST_A + A_STAGE;
  static constexpr int ST_S = ST_W + W_STAGE;
  static constexpr int STAGE_BYTES = (ST_S + S_STAGE + 1023) / 1024 * 1024;
  static constexpr int OFF_RED = NSTAGE * STAGE_BYTES;
  static constexpr int OFF_BAR = OFF_RED + RED_BYTES;
  static constexpr int OFF_TK = OFF_BAR + (2 * NSTAGE + 2) * 8;  // two ticket broadcast words (flusher, consumers)
  static constexpr int SMEM_BYTES = OFF_TK + 16;
  static constexpr int THREADS = WARPS * 32 + 64;  // + producer warp + flusher warp
  // two CTAs per SM (PDL co-residency with the next linear) while the register file allows it
  static constexpr int MINB = (WARPS <= 8 && NT <= 2) ? 2 : 1;
  static constexpr int FLAGS_PER_TILE = CHUNKS * 16;  // ints of `workspace` owned by one column tile
  static_assert(WARPS % CHUNKS == 0 && KROWS % WK == 0, "warp layout");
  static_assert(KB % 64 == 0 && (KB <= 128 || (KB % 128 == 0 && 8 % WK == 0)), "stage k: a multiple of the 128-byte activation row; one group, or whole groups whose 8 k16-rows fall to one row index i of every warp");
  static_assert(A_SUB % 1024 == 0 && W_STAGE % 128 == 0, "TMA swizzle / box alignment");
};

// unit partition helpers (64-bit: units * grid can exceed 2^31)
__device__ __forceinline__ int unit_begin(int cta, long long units, int grid) {
  return static_cast<int>((static_cast<long long>(cta) * units) / grid);
}
// the CTA whose range contains unit u: largest i with floor(i*U/G) <= u
__device__ __forceinline__ int cta_of_unit(long long u, long long units, int grid) {
  return static_cast<int>(((u + 1) * grid - 1) / units);
}

// converts 4 fp32 sums to T with the reference's rounding points (RN to T; channelwise scale on the
// ROUNDED sum, marlin_kernel_impl.cuh:957-963; then residual, then bias) and stores them
template <typename T, bool GROUPED>
__device__ __forceinline__ void emit_out(const DecodeArgs& p, int tok, int n, float4 v) {
  T o[4] = {Elem<T>::from_float(v.x), Elem<T>::from_float(v.y), Elem<T>::from_float(v.z),
            Elem<T>::from_float(v.w)};
  if (!GROUPED) {
    const T* sc = reinterpret_cast<const T*>(p.scales);
#pragma unroll
    for (int e = 0; e < 4; ++e) o[e] = __hmul(o[e], sc[chan_scale_pos(n + e)]);
  }
  const size_t off = out_offset(tok, n, p.n, p.m, p.qkv_a, p.qkv_b);
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
#pragma unroll
  for (int e = 0; e < 4; ++e) o[e] = apply_out_scale<T>(o[e], p.out_scale);
  uint2 pk;
  pk.x = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[0])) |
         (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[1])) << 16);
  pk.y = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[2])) |
         (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[3])) << 16);
  *reinterpret_cast<uint2*>(reinterpret_cast<T*>(p.c) + off) = pk;
}

// fused FFN activation over a [gate | up] tile (reference: gated_silu_interleaved_kernel,
// src/model/activation.cuh:7-18, compiled with --use_fast_math there): gate and up sums are rounded to T
// first (they are gateup_proj's stored output in the reference), then T(g * sigmoid(g) * u)
template <typename T>
__device__ __forceinline__ void emit_gated(const DecodeArgs& p, int tok, int n_out, float4 gs, float4 us) {
  const float gv[4] = {gs.x, gs.y, gs.z, gs.w};
  const float uv[4] = {us.x, us.y, us.z, us.w};
  T o[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float g = Elem<T>::to_float(Elem<T>::from_float(gv[e]));
    const float u = Elem<T>::to_float(Elem<T>::from_float(uv[e]));
    const float sg = __fdividef(1.0f, 1.0f + __expf(-g));
    o[e] = Elem<T>::from_float(g * sg * u);
  }
  uint2 pk;
  pk.x = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[0])) |
         (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[1])) << 16);
  pk.y = static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[2])) |
         (static_cast<uint32_t>(*reinterpret_cast<uint16_t*>(&o[3])) << 16);
  *reinterpret_cast<uint2*>(reinterpret_cast<T*>(p.c) + static_cast<size_t>(tok) * (p.n >> 1) + n_out) = pk;
}

// the sharers of column tile c under the even unit split: CTAs cta_first..cta_last, slot = cta - cta_first
struct ColumnShare {
  int nparts, my_slot;
  float* slots;   // c_tmp slots of this column: [slot][m][TN] fp32
  int* counter;   // arrival counter of this column in `workspace`; zero between launches
  size_t slot_elems;
};
template <int TN, int FLAGS_PER_TILE>
__device__ __forceinline__ ColumnShare column_share(const DecodeArgs& p, int c, long long units, int grid) {
  const long long ufirst = static_cast<long long>(c) * p.kblks;
  const int cta_first = cta_of_unit(ufirst, units, grid);
  const int cta_last = cta_of_unit(ufirst + p.kblks - 1, units, grid);
  ColumnShare cs;
  cs.nparts = cta_last - cta_first + 1;
  cs.my_slot = static_cast<int>(blockIdx.x) - cta_first;
  cs.slot_elems = static_cast<size_t>(p.m) * TN;
  cs.slots = p.c_tmp + static_cast<size_t>(c) * p.slot_max * cs.slot_elems;
  cs.counter = p.workspace + c * FLAGS_PER_TILE;
  return cs;
}

struct WarpSync {
  __device__ __forceinline__ void operator()() const { __syncwarp(); }
};
template <int NTHR>
struct ConsumerSync {  // named barrier 1: the consumer warps only
  __device__ __forceinline__ void operator()() const { named_bar_sync(1, NTHR); }
};

// Turns the finished fp32 tile `red` ([token][TNP]) of column tile c into output.  Called by a whole thread
// group (the flusher warp, or all consumer warps for the CTA's last segment) with `tid` in [0, nthr).
//   whole column                  -> epilogue + store.
//   split column (nparts sharers) -> publish the partial, fence, take a ticket; the LAST ticket holder adds all
//                                    partials in slot order, stores, and zeroes the counter.  Nobody waits.
// `stage` (may be null): `cap` float4 of dead shared memory the last arriver may use to pull the other
// partials with every copy in flight at once; without it the partials are read straight from L2.
template <typename T, int TN, int TNP, int FLAGS_PER_TILE, bool GROUPED, bool GATED, typename Sync>
__device__ __forceinline__ void finish_segment(const DecodeArgs& p, const float* red, int c, bool whole,
                                               long long units, int grid, int tid, int nthr, Sync sync,
                                               volatile int* tk, float4* stage, int cap) {
  constexpr int RI = TN / 4;               // float4 items per token row
  const int items = p.m * RI;              // float4 items of the valid token rows
  const int col0 = c * TN;
  const int gcol0 = c * (TN / 2);          // gated: first output column of the tile
  auto own_item = [&](int item) {
    const int tok = item / RI, nn = (item - tok * RI) * 4;
    return *reinterpret_cast<const float4*>(red + tok * TNP + nn);
  };
  // fused all-reduce: the finished tile goes to every rank's receive buffer instead of C (see ar_push)
  const bool ar = !GATED && p.ar_world > 1;
  const int ar_parity = ar ? ((p.ar_seen[c] + 1) & 1) : 0;
  auto emit = [&](int tok, int n, float4 v) {
    if (ar) ar_push<T, GROUPED>(p, ar_parity, tok, n, v);
    else emit_out<T, GROUPED>(p, tok, n, v);
  };
  auto ar_signal = [&]() {   // called by the whole group after the tile has been pushed
    if (!ar) return;
    sync();   // the group's pushes are ordered before the releases below
    if (tid < p.ar_world)
      asm volatile("red.release.sys.global.add.s32 [%0], 1;" ::"l"(p.ar_flag[tid] + p.ar_rank * kArTiles + c) : "memory");
  };
  if (whole) {
    if (GATED) {
      for (int e = tid; e < p.m * (RI / 2); e += nthr) {
        const int tok = e / (RI / 2), jj = e - tok * (RI / 2);
        emit_gated<T>(p, tok, gcol0 + jj * 4, own_item(tok * RI + jj), own_item(tok * RI + RI / 2 + jj));
      }
    } else {
      for (int e = tid; e < items; e += nthr) {
        const int tok = e / RI, nn = (e - tok * RI) * 4;
        if (col0 + nn < p.n) emit(tok, col0 + nn, own_item(e));
      }
    }
    ar_signal();
    return;
  }
  const ColumnShare cs = column_share<TN, FLAGS_PER_TILE>(p, c, units, grid);
  // ---- publish: fp32 partial -> own slot; fence; ticket ----
  {
    float* mine = cs.slots + static_cast<size_t>(cs.my_slot) * cs.slot_elems;
    for (int e = tid; e < items; e += nthr) __stcg(reinterpret_cast<float4*>(mine) + e, own_item(e));
  }
  sync();   // the group's stores are ordered before thread 0's release below (cumulativity over the barrier)
  if (tid == 0) {
    int t;   // release: the partial is visible before the ticket; acquire: the earlier sharers' partials are visible after it
    asm volatile("atom.add.acq_rel.gpu.global.s32 %0, [%1], 1;" : "=r"(t) : "l"(cs.counter) : "memory");
    *tk = t;
    if (nthr > 32) dbg_stamp1(p, 4);
  }
  sync();
  if (*tk != cs.nparts - 1) return;   // somebody else arrives later and finishes the column
  // ---- last arriver: every other partial is visible (released before the other tickets, acquired with ours) ----
  if (tid == 0) *cs.counter = 0;      // all sharers have drawn: the workspace goes back to zero (reference contract)
  const int np1 = cs.nparts - 1;
  // few loads per thread (decode widths: M x TN / 4 items x the other sharers over the group's threads): straight from L2
  // into registers — every load is in flight at once anyway, and the shared-memory staging round (cp.async, wait, two
  // barriers) is pure latency on the kernel's critical tail
  const bool direct = items * np1 <= 8 * nthr;
  if (stage == nullptr || direct || cap < np1 * (GATED ? RI : 1)) {
    // (also when the flusher warp drew the last ticket: its shared memory is live)
    auto sum_item = [&](int item) {
      float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int sl = 0; sl < cs.nparts; ++sl) {   // slot order: bit-deterministic
        const float4 v = (sl == cs.my_slot) ? own_item(item)
                                            : __ldcg(reinterpret_cast<const float4*>(cs.slots + sl * cs.slot_elems) + item);
        if (sl == 0) {
          sum = v;
        } else {
          sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
        }
      }
      return sum;
    };
    if (GATED) {
      for (int e = tid; e < p.m * (RI / 2); e += nthr) {
        const int tok = e / (RI / 2), jj = e - tok * (RI / 2);
        emit_gated<T>(p, tok, gcol0 + jj * 4, sum_item(tok * RI + jj), sum_item(tok * RI + RI / 2 + jj));
      }
    } else {
      for (int e = tid; e < items; e += nthr) {
        const int tok = e / RI, nn = (e - tok * RI) * 4;
        if (col0 + nn < p.n) emit(tok, col0 + nn, sum_item(e));
      }
    }
    ar_signal();
    return;
  }
  // staged: rounds of `ib` items; all copies of a round in flight at once (16-byte cp.async, L2 -> shared).
  // Whole token rows per round when gated, so that a gate item and its up item are staged together.
  const int ib = GATED ? (cap / np1) / RI * RI : cap / np1;
  for (int i0 = 0; i0 < items; i0 += ib) {
    const int ni = (items - i0) < ib ? (items - i0) : ib;
    for (int sl = 0, row = 0; sl < cs.nparts; ++sl) {
      if (sl == cs.my_slot) continue;
      const float4* src = reinterpret_cast<const float4*>(cs.slots + sl * cs.slot_elems) + i0;
      const uint32_t dst = smem_u32(stage + row * ni);
      for (int e = tid; e < ni; e += nthr)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + e * 16), "l"(src + e) : "memory");
      ++row;
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (tid == 0 && nthr > 32) dbg_stamp1(p, 5);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    sync();
    if (tid == 0 && nthr > 32) dbg_stamp1(p, 6);
    auto sum_item = [&](int e) {   // e: item index inside the round
      float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int sl = 0, row = 0; sl < cs.nparts; ++sl) {   // slot order: bit-deterministic
        float4 v;
        if (sl == cs.my_slot) {
          v = own_item(i0 + e);
        } else {
          v = stage[row * ni + e];
          ++row;
        }
        if (sl == 0) {
          sum = v;
        } else {
          sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
        }
      }
      return sum;
    };
    if (GATED) {
      for (int e = tid; e < ni / 2; e += nthr) {
        const int r = e / (RI / 2), jj = e - r * (RI / 2);
        emit_gated<T>(p, i0 / RI + r, gcol0 + jj * 4, sum_item(r * RI + jj), sum_item(r * RI + RI / 2 + jj));
      }
    } else {
      for (int e = tid; e < ni; e += nthr) {
        const int item = i0 + e;
        const int tok = item / RI, nn = (item - tok * RI) * 4;
        if (col0 + nn < p.n) emit(tok, col0 + nn, sum_item(e));
      }
    }
    sync();
  }
  ar_signal();
}

// A CTA's contiguous unit range [u0, u1) cut into column segments, walked in the order
//     [partial END of its first column] [partial START of its last column] [its whole columns ...]
// The partial segments are the ones other CTAs share: doing them FIRST means their split-column reduction
// (publish, ticket, the last arriver's sum) happens in the flusher warp while every consumer streams on; what a
// CTA finishes with is a whole column whenever it owns one, which needs no other CTA — the kernel's tail is then a
// plain epilogue instead of a cross-CTA round trip.  The ticket scheme has no ordering requirement, so any walk
// order is valid.
struct SegWalk {
  int kblks, c0, c1;   // first / last column of the range
  int bf, bl;          // first k-block in c0, one past the last k-block in c1
  int head, tail;      // 1: the range begins / ends inside a column (and c0 != c1)
  int nseg;
  __device__ __forceinline__ SegWalk(int u0, int u1, int kb) : kblks(kb) {
    c0 = u0 / kb;
    bf = u0 - c0 * kb;
    c1 = (u1 - 1) / kb;
    bl = (u1 - 1) - c1 * kb + 1;
    if (c0 == c1) {
      head = tail = 0;
      nseg = 1;
    } else {
      head = bf != 0;
      tail = bl != kb;
      nseg = c1 - c0 + 1;
    }
  }
  // i-th segment of the walk: column c, k-blocks [b0, b1)
  __device__ __forceinline__ void get(int i, int& c, int& b0, int& b1) const {
    if (c0 == c1) { c = c0; b0 = bf; b1 = bl; return; }
    if (head && i == 0) { c = c0; b0 = bf; b1 = kblks; return; }
    if (tail && i == head) { c = c1; b0 = 0; b1 = bl; return; }
    c = c0 + head + (i - head - tail);   // whole columns, ascending
    b0 = 0;
    b1 = kblks;
  }
};

// GATED: the fused [gate | up] tile + SiLU(gate) * up epilogue; a template parameter (not a runtime flag) so
// that the plain kernels keep their register allocation and code exactly (a runtime flag cost 2 % at M <= 8)
template <typename T, int NT, int CHUNKS, int KROWS, int WARPS, int STAGES, bool GROUPED, bool GATED = false>
__global__ void __launch_bounds__(WARPS * 32 + 64, DecodeCfg<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED>::MINB)
w4a16_decode_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_s,
                    const __grid_constant__ CUtensorMap tm_a, const DecodeArgs p) {
  using Cfg = DecodeCfg<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED>;
  constexpr int TN = Cfg::TN, KB = Cfg::KB, WK = Cfg::WK, RPW = Cfg::RPW, KC = Cfg::KC;
  constexpr int TNP = Cfg::TNP;
  constexpr int NS = Cfg::NSTAGE;  // ring depth
  constexpr int NCONS = WARPS * 32;

  // no static shared memory in this kernel: the dynamic window starts 1024-aligned (needed by the
  // 128B-swizzled activation boxes); checked, not assumed
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0u) __trap();
  float* red = reinterpret_cast<float*>(smem + Cfg::OFF_RED);
  const uint32_t bar_full = sbase + Cfg::OFF_BAR;
  const uint32_t bar_empty = bar_full + NS * 8;
  const uint32_t bar_red_full = bar_empty + NS * 8;   // consumers -> flusher: tile complete
  const uint32_t bar_red_empty = bar_red_full + 8;        // flusher -> consumers: tile consumed
  volatile int* tk = reinterpret_cast<volatile int*>(smem + Cfg::OFF_TK);  // [0] flusher, [1] consumers

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, WARPS);
    }
    mbar_init(bar_red_full, WARPS);
    mbar_init(bar_red_empty, 1);
    fence_barrier_init();
  }
  __syncthreads();
  // let the next kernel in the stream start its own weight prefetch as early as possible
  pdl_launch_dependents();
  if (threadIdx.x == 0) dbg_stamp1(p, 0);

  const long long units = static_cast<long long>(p.ncols) * p.kblks;
  const int grid = gridDim.x;
  // stream-K: an even split of the (column, k-block) units; column-parallel: an even split of whole columns
  const int u0 = p.colpar ? static_cast<int>((static_cast<long long>(blockIdx.x) * p.ncols) / grid) * p.kblks
                          : unit_begin(blockIdx.x, units, grid);
  const int u1 = p.colpar ? static_cast<int>((static_cast<long long>(blockIdx.x + 1) * p.ncols) / grid) * p.kblks
                          : unit_begin(blockIdx.x + 1, units, grid);
  const int my_units = u1 - u0;
  if (my_units == 0) return;

  if (warp == WARPS) {
    // ===================================== producer =====================================
    if (lane == 0) {
      tma_prefetch_desc(&tm_w);
      tma_prefetch_desc(&tm_a);
      if (GROUPED) tma_prefetch_desc(&tm_s);
#if B200_DECODE_LAB
      bool lab_first = true;   // still inside the first ring fill
#endif
      auto issue_w = [&](int s, int c, int b) {
        const uint32_t full = bar_full + 8 * s;
#if B200_DECODE_LAB
        const bool skip_s = (p.lab & 2) && !lab_first, skip_a = (p.lab & 1) && !lab_first;
        mbar_arrive_expect_tx(full, Cfg::STAGE_TX - (skip_s ? Cfg::S_STAGE : 0) - (skip_a ? Cfg::A_STAGE : 0));
#else
        mbar_arrive_expect_tx(full, Cfg::STAGE_TX);
#endif
        if (GATED) {  // one 4-d box: CHUNKS/2 gate chunks + the same CHUNKS/2 up chunks (see launch_decode)
          tma_load_4d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_W, &tm_w, full, 0, c * (CHUNKS / 2), 0, b * KROWS,
                      kEvictFirst);
          if (GROUPED
#if B200_DECODE_LAB
              && !skip_s
#endif
          )
            tma_load_4d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_S, &tm_s, full, 0, c * (CHUNKS / 2), 0,
                        (b * KB) / 128, kEvictLast);
        } else {
          tma_load_3d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_W, &tm_w, full, 0, c * CHUNKS, b * KROWS,
                      kEvictFirst);
          if (GROUPED
#if B200_DECODE_LAB
              && !skip_s
#endif
          )
            tma_load_3d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_S, &tm_s, full, 0, c * CHUNKS,
                        (b * KB) / 128, kEvictLast);
        }
      };
      auto issue_a = [&](int s, int b) {
        const uint32_t full = bar_full + 8 * s;
#if B200_DECODE_LAB
        if ((p.lab & 1) && !lab_first) return;
#endif
#pragma unroll
        for (int kc = 0; kc < KC; ++kc)
          tma_load_2d(sbase + s * Cfg::STAGE_BYTES + Cfg::ST_A + kc * Cfg::A_SUB, &tm_a, full,
                      b * KB + kc * 64, 0, kEvictLast);
      };
      const SegWalk walk(u0, u1, p.kblks);
      const int npre = my_units < NS ? my_units : NS;
      // weights do not depend on the prior grid: the first ring fill is requested before griddepcontrol.wait
      int si = 0, c, b, bend;
      walk.get(0, c, b, bend);
      auto advance = [&]() {
        if (++b == bend && si + 1 < walk.nseg) walk.get(++si, c, b, bend);
      };
      for (int it = 0; it < npre; ++it) {
        issue_w(it, c, b);
        advance();
      }
      pdl_wait_prior_grids();               // activations do
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
      uint32_t ph = (npre == NS) ? 0u : 1u;  // parity of the empty-barrier phase to wait for
#if B200_DECODE_LAB
      lab_first = false;
#endif
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
    // walks the same column segments as the consumers and finishes all but the CTA's last one
    pdl_wait_prior_grids();  // C, c_tmp and the counters are shared with prior grids
    const SegWalk walk(u0, u1, p.kblks);
    for (int si = 0; si + 1 < walk.nseg; ++si) {   // the CTA's last segment is finished by the consumer warps themselves
      int c, b0, b1;
      walk.get(si, c, b0, b1);
      mbar_wait_backoff(bar_red_full, static_cast<uint32_t>(si) & 1u, 64);
      finish_segment<T, TN, TNP, Cfg::FLAGS_PER_TILE, GROUPED, GATED>(p, red, c, /*whole=*/b0 == 0 && b1 == p.kblks, units,
                                                                       grid, lane, 32, WarpSync(), tk, nullptr, 0);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_red_empty);
    }
    return;
  }

  // ======================================= consumers =======================================
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

  // per-lane byte offsets inside a stage, fixed for the whole kernel (three registers):
  //   weights: k16-row r = wk + WK*i, chunk wn -> block (r*CHUNKS + wn) = warp + i*WARPS, so this lane's
  //     16 B sit at threadIdx.x*16 + i*(WARPS*512)
  //   activations (ldmatrix row address inside the 128B-swizzled [MT][64] sub-tile of row r):
  //     NT==1 (.x2): lanes 0-7 -> k 0..7 of the k16 step, lanes 8-15 -> k 8..15 (16-31 ignored)
  //     NT>=2 (.x4): lanes 16-31 -> token rows 8..15 of the pair
  //     r = wk | (WK*i) (disjoint bits), so the address is a lane constant XOR / plus compile-time terms
  const int lm_row = (lane & 7) + ((NT >= 2 && (lane & 16)) ? 8 : 0);
  const int lm_half = (lane >> 3) & 1;
  const uint32_t q_lane = threadIdx.x * 16u;
  const uint32_t a_lane = static_cast<uint32_t>((wk >> 2) * Cfg::A_SUB + lm_row * 128 +
                                                ((((wk & 3) << 1) ^ lm_half ^ (lm_row & 7)) << 4));
  const uint32_t s_off = static_cast<uint32_t>(wn * 128 + g * 16);

  const SegWalk walk(u0, u1, p.kblks);
  int s = 0;
  uint32_t ph = 0;
  uint32_t stg = sbase;        // running pointers: current stage block, its full / empty barriers
  uint32_t bfull = bar_full, bempty = bar_empty;
  // per-lane offsets inside a stage block (lane constants): the int4 words, the scales, the activation rows of the
  // two swizzle phases this warp's rows alternate between
  const uint32_t q_off = Cfg::ST_W + q_lane;
  const uint32_t sc_off = Cfg::ST_S + s_off;
  // row RI = WK * i of the stage: swizzle phase (RI & 3) << 1 XOR-ed into the 16-byte chunk index, sub-tile RI >> 2;
  // loop-invariant, so computed once (equal values share a register)
  uint32_t a_off[RPW];
#pragma unroll
  for (int i = 0; i < RPW; ++i) {
    const int RI = WK * i;
    a_off[i] = Cfg::ST_A + (a_lane ^ static_cast<uint32_t>(((RI & 3) << 1) << 4)) + static_cast<uint32_t>((RI >> 2) * Cfg::A_SUB);
  }

  mbar_wait_backoff(bar_full, 0, 64);  // under PDL this first wait spans the tail of the previous kernel
  if (threadIdx.x == 0) dbg_stamp1(p, 2);

  for (int si = 0; si < walk.nseg; ++si) {
   int seg_c, seg_b0, seg_b1;
   walk.get(si, seg_c, seg_b0, seg_b1);
#pragma unroll 1
   for (int ub = seg_b0; ub < seg_b1; ++ub) {
    mbar_wait(bfull, ph);

    // scales of the stage's groups: held for the whole stage when there are at most two, else fetched as
    // each warp's row index i enters a new group (fewer live registers).  One register per {column g, g + 8} pair.
    constexpr bool LAZY_S = Cfg::SG > 2 && NT > 1;
    constexpr int SGL = LAZY_S ? 1 : Cfg::SG;
    uint32_t sc[SGL][4];
    auto load_scales = [&](int slot, int gi) {
      const uint4 sv = lds128(stg + sc_off + gi * (TN * 2));
      sc[slot][0] = sv.x; sc[slot][1] = sv.y; sc[slot][2] = sv.z; sc[slot][3] = sv.w;
    };
    if (GROUPED && !LAZY_S) {
#pragma unroll
      for (int gi = 0; gi < Cfg::SG; ++gi) load_scales(gi, gi);
    }

    uint4 q[RPW];
#pragma unroll
    for (int i = 0; i < RPW; ++i) q[i] = lds128(stg + q_off + i * (WARPS * 512));
#if B200_DECODE_LAB
    if (!(p.lab & 4))
#endif
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      // activation fragments: tokens as the N side of the MMA
      uint32_t bf[NT][2];
      const int RI = WK * i;  // compile-time after unrolling
      const uint32_t a_row = stg + a_off[i];
      if (NT == 1) {
        ldmatrix_x2(bf[0][0], bf[0][1], a_row);
      } else {
#pragma unroll
        for (int np = 0; np < NT / 2; ++np)  // token rows +16 per pair: 2048 B, swizzle phase unchanged
          ldmatrix_x4(bf[2 * np][0], bf[2 * np][1], bf[2 * np + 1][0], bf[2 * np + 1][1], a_row + np * 2048);
      }
      if (GROUPED && LAZY_S && (RI & 7) == 0) load_scales(0, RI >> 3);
      const uint32_t qq[4] = {q[i].x, q[i].y, q[i].z, q[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t w[4];
        const int gi = (Cfg::SG > 1 && !LAZY_S) ? (RI >> 3) : 0;  // rows 8.. of a multi-group stage use the next scale rows
        Dequant<T>::template word<GROUPED>(qq[j], GROUPED ? sc[gi][j] : 0u, w);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) mma_m16n8k16<T>(acc[j][nt], w, bf[nt][0], bf[nt][1]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bempty);
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
    // ---- the segment is complete: reduce the WK warps of every chunk into one tile, hand it over ----
    {
      mbar_wait(bar_red_empty, (static_cast<uint32_t>(si) & 1u) ^ 1u);  // previous tile consumed (passes the first time)
      // the WK warps sharing a chunk add their partials in wk order (fixed -> deterministic)
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
                if (tok < p.m) {
                  float* dst = red + tok * TNP + n;
                  *dst = (r == 0) ? acc[j][nt][e] : (*dst + acc[j][nt][e]);
                }
                acc[j][nt][e] = 0.f;
              }
        }
        if (r + 1 < WK) named_bar_sync(1, NCONS);
      }
      if (si == walk.nseg - 1) {
        // last segment of this CTA: nothing left to overlap with -> all consumer warps finish it; the ring is dead
        // by now (every stage consumed) and serves as the staging area of the reduction
        if (threadIdx.x == 0) dbg_stamp1(p, 3);
        named_bar_sync(1, NCONS);
        pdl_wait_prior_grids();  // C, c_tmp and the counters are shared with prior grids
        finish_segment<T, TN, TNP, Cfg::FLAGS_PER_TILE, GROUPED, GATED>(
            p, red, seg_c, /*whole=*/seg_b0 == 0 && seg_b1 == p.kblks, units, grid, static_cast<int>(threadIdx.x), NCONS,
            ConsumerSync<NCONS>(), tk + 1, reinterpret_cast<float4*>(smem), (NS * Cfg::STAGE_BYTES) / 16);
        if (threadIdx.x == 0) dbg_stamp1(p, 7);
      } else {
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_red_full);
      }
    }
  }

  // ---- fused all-reduce, second half: this CTA's share of the column tiles (see the block comment above) ----
  if (!GATED && p.ar_world > 1) {
    const int tid = static_cast<int>(threadIdx.x);
    const T* recv = reinterpret_cast<const T*>(p.ar_recv[p.ar_rank]);
    const int* flags = p.ar_flag[p.ar_rank];
    for (int c = blockIdx.x; c < p.ncols; c += gridDim.x) {
      const int target = p.ar_seen[c] + 1;
      named_bar_sync(1, NCONS);   // everybody has read the epoch before anybody advances it
      bool ok = true;
      if (tid < p.ar_world) {
        const int* f = flags + tid * kArTiles + c;
        unsigned long long t0 = 0;
        for (unsigned spin = 0; ld_acquire_sys_s32(f) - target < 0; ++spin) {
          if ((spin & 1023u) == 1023u) {   // a source that is 2 s late is treated as dead: error word, no sum
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            if (t - t0 > 2000000000ull) { ok = false; break; }
          }
        }
        if (!ok) atomicExch(p.ar_seen + kArTiles, 1);
      }
      named_bar_sync(1, NCONS);
      const size_t pbase = static_cast<size_t>(target & 1) * p.ar_world * p.ar_slot_elems;
      const int col0 = c * TN;
      for (int e = tid; e < p.m * (TN / 8); e += NCONS) {
        const int tok = e / (TN / 8), n = col0 + (e - tok * (TN / 8)) * 8;
        if (n >= p.n) continue;
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int w = 0; w < p.ar_world; ++w) {   // rank order: every rank computes the same bits
          const uint4 v = ld_relaxed_sys_u4(recv + pbase + static_cast<size_t>(w) * p.ar_slot_elems + static_cast<size_t>(tok) * p.n + n);
          const T* h = reinterpret_cast<const T*>(&v);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] += Elem<T>::to_float(h[i]);
        }
        // the partials already carry the channelwise scale (ar_push): the epilogue here is the grouped one
        emit_out<T, true>(p, tok, n, make_float4(acc[0], acc[1], acc[2], acc[3]));
        emit_out<T, true>(p, tok, n + 4, make_float4(acc[4], acc[5], acc[6], acc[7]));
      }
      named_bar_sync(1, NCONS);
      if (tid == 0) p.ar_seen[c] = target;
    }
  }
}

}  // namespace b200
