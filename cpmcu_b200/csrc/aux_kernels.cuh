// Small helper kernels around the GEMM: the dequantisation dump (bit-exact parity surface), the
// act-order column gather of A, and the on-device AutoGPTQ -> Marlin repack.
#pragma once
#include "common.cuh"

namespace b200 {

// W[K][N] (T) from the Marlin stream, with the SAME Dequant<T>::word the GEMM kernels use.
// One warp per 512-byte block (16 k x 64 n); grid-stride over blocks.
template <typename T, bool GROUPED>
__global__ void __launch_bounds__(256)
w4a16_dequant_kernel(const uint4* __restrict__ qweight, const T* __restrict__ scales,
                     T* __restrict__ w_out, int n, int k, int group_size) {
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int chunks = n / 64;
  const long long blocks = static_cast<long long>(k / 16) * chunks;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  for (long long blk = warp0; blk < blocks; blk += nwarps) {
    const int r = static_cast<int>(blk / chunks), ch = static_cast<int>(blk % chunks);
    const uint4 qv = qweight[blk * 32 + lane];
    uint32_t sc[4] = {0, 0, 0, 0};
    if (GROUPED) {
      const int grp = (r * 16) / group_size;
      const uint4 sv = *reinterpret_cast<const uint4*>(scales + static_cast<size_t>(grp) * n +
                                                       ch * 64 + g * 8);
      sc[0] = sv.x; sc[1] = sv.y; sc[2] = sv.z; sc[3] = sv.w;
    }
    const uint32_t qq[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t w[4];
      Dequant<T>::template word<GROUPED>(qq[j], sc[j], w);
      // w[0]: (k 2t,2t+1 | n g)   w[1]: (k 2t,2t+1 | n g+8)   w[2]: (k 2t+8,+9 | n g)   w[3]: (.. | n g+8)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int kk = r * 16 + 2 * t + ((e & 2) ? 8 : 0);
        const int nn = ch * 64 + j * 16 + g + ((e & 1) ? 8 : 0);
        const uint16_t lo = static_cast<uint16_t>(w[e] & 0xffffu);
        const uint16_t hi = static_cast<uint16_t>(w[e] >> 16);
        reinterpret_cast<uint16_t*>(w_out)[static_cast<size_t>(kk) * n + nn] = lo;
        reinterpret_cast<uint16_t*>(w_out)[static_cast<size_t>(kk + 1) * n + nn] = hi;
      }
    }
  }
}

// act-order: a_tmp[m][kk] = a[m][perm[kk]]   (reference semantics: gptq_marlin_utils.cu:28-77)
template <typename T>
__global__ void __launch_bounds__(256)
gather_columns_kernel(const T* __restrict__ a, const int32_t* __restrict__ perm, T* __restrict__ a_tmp,
                      int m, int k) {
  const long long total = static_cast<long long>(m) * k;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>(i / k), kk = static_cast<int>(i - static_cast<long long>(row) * k);
    a_tmp[i] = a[static_cast<size_t>(row) * k + perm[kk]];
  }
}

// AutoGPTQ int32[K/8][N] (nibble i of a word = row 8r+i) -> Marlin int32[K/16][2N]
// (format: scripts/model_convert/gptq2marlin.py:35-61,109-134).  One thread per output word.
__global__ void __launch_bounds__(256)
repack_gptq_kernel(const uint32_t* __restrict__ gptq, uint32_t* __restrict__ marlin, int n, int k) {
  const long long words = static_cast<long long>(k / 16) * 2 * n;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < words;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / (2 * n));
    const int w = static_cast<int>(i - static_cast<long long>(r) * 2 * n);
    const int ch = w >> 7, lane = (w >> 2) & 31, j = w & 3;
    const int g = lane >> 2, t = lane & 3;
    uint32_t out = 0;
#pragma unroll
    for (int slot = 0; slot < 8; ++slot) {
      // slot -> (k offset, column half): see common.cuh "dequantisation"
      const int kk = r * 16 + 2 * t + ((slot & 1) ? 8 : 0) + ((slot & 4) ? 1 : 0);
      const int nn = ch * 64 + j * 16 + g + ((slot & 2) ? 8 : 0);
      const uint32_t src = gptq[static_cast<size_t>(kk >> 3) * n + nn];
      out |= ((src >> (4 * (kk & 7))) & 0xfu) << (4 * slot);
    }
    marlin[i] = out;
  }
}

// One-shot all-reduce (sum) over NVLink peer memory for the row-parallel linears (o, down) of a
// tensor-parallel model: every rank's GEMM writes its [M][N] partial into its own SYMMETRIC buffer;
// this kernel (a) tells every peer "my partial is complete" with a fire-and-forget system-scope
// red.release on the peer's signal word, (b) waits until every peer has said so, (c) adds the W
// partials straight out of peer memory IN RANK ORDER in fp32 (every rank computes bit-identical
// sums) and writes T.  One NVLink round trip instead of NCCL's ring / tree protocol: [M][8192] fp16
// is 16 KB .. 1 MB, pure latency.  Signals are monotonic counters and the expected value lives in
// device memory (`epoch`, one int per block), so the launch is CUDA-graph capturable and replayable.
// Callers alternate between two symmetric buffers: entering call i+1 proves every rank has finished
// reading the buffer of call i, so no trailing barrier is needed.
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

constexpr int kArMaxWorld = 8;
constexpr int kArMaxBlocks = 128;
constexpr int kArErrIndex = 2 * kArMaxBlocks;        // epoch[kArErrIndex]: sticky error word of this rank
constexpr unsigned long long kArTimeoutNs = 2000000000ull;  // a peer that is 2 s late is treated as dead

struct AllReduceArgs {
  const void* in[kArMaxWorld];  // every rank's symmetric input buffer (peer-mapped addresses)
  void* res[kArMaxWorld];       // two-shot only: every rank's symmetric result buffer
  int* sig[kArMaxWorld];        // every rank's signal pad: int[2][kArMaxBlocks][kArMaxWorld]
  int* epoch;                   // local: int[2][kArMaxBlocks] barriers completed per phase bank and block, then
                                // int[1] error word (kArErrIndex): set when a peer never arrived
  void* out;                    // local T[elems]
  long long elems;
  int rank, world;
};

// all peers' block b tell each other "phase reached" (fire-and-forget red.release.sys on the peer's
// counter) and wait until every peer has said so; `phase` selects the counter bank.  Returns false when a
// peer did not arrive within kArTimeoutNs (a rank died or never launched): the caller then raises the error
// word, poisons its output and leaves the epoch alone — a late or dead peer is an ERROR, never a wrong sum.
__device__ __forceinline__ bool ar_block_barrier(const AllReduceArgs& a, int b, int phase, int target) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < a.world) {
    const int w = threadIdx.x;
    const int slot = (phase * kArMaxBlocks + b) * kArMaxWorld;
    asm volatile("red.release.sys.global.add.s32 [%0], 1;" ::"l"(a.sig[w] + slot + a.rank) : "memory");
    const int* mine = a.sig[a.rank] + slot + w;
    unsigned long long t0 = 0;
    for (unsigned spin = 0; ld_acquire_sys(mine) - target < 0; ++spin) {
      if ((spin & 1023u) == 1023u) {   // look at the clock once in a while
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t0 == 0) t0 = t;
        if (t - t0 > kArTimeoutNs) {
          s_ok = 0;
          break;
        }
      }
    }
  }
  __syncthreads();
  return s_ok != 0;
}

// fp32 sum, in rank order, of the 8 elements at offset i of every rank's input
template <typename T>
__device__ __forceinline__ uint4 ar_sum8(const AllReduceArgs& a, long long i) {
  using T2 = typename Elem<T>::T2;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  uint4 v[kArMaxWorld];
#pragma unroll
  for (int w = 0; w < kArMaxWorld; ++w)
    if (w < a.world) v[w] = ld_relaxed_sys_v4(reinterpret_cast<const T*>(a.in[w]) + i);
#pragma unroll
  for (int w = 0; w < kArMaxWorld; ++w)
    if (w < a.world) {
      const T2* h = reinterpret_cast<const T2*>(&v[w]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        acc[2 * e] += Elem<T>::to_float(h[e].x);
        acc[2 * e + 1] += Elem<T>::to_float(h[e].y);
      }
    }
  uint4 o;
  T* oh = reinterpret_cast<T*>(&o);
#pragma unroll
  for (int e = 0; e < 8; ++e) oh[e] = Elem<T>::from_float(acc[e]);
  return o;
}

// TWO_SHOT = false: every rank adds all W partials itself (one NVLink round trip; reads W x elems).
// TWO_SHOT = true : rank r adds only slice r (reduce-scatter out of peer memory) into its symmetric
//   result buffer, a second block-level barrier, then every rank copies the W result slices
//   (all-gather out of peer memory): reads 2 x elems instead of W x elems — for the large messages
//   (M = 64: 1 MB) at W >= 4, where the one-shot form is NVLink-bandwidth bound.
// Block b of every rank owns the same part of every slice, so the barriers pair block b with block b.
template <typename T, bool TWO_SHOT>
__global__ void __launch_bounds__(256)
allreduce_kernel(const AllReduceArgs a) {
  const int b = blockIdx.x;
  // PDL: the next kernel in the stream (the following linear) may start its prologue and weight
  // prefetch now; this kernel waits for the GEMM that produced its input before telling the peers
  // "my partial is complete" (both are no-ops unless the launches carry the attribute)
  pdl_launch_dependents();
  pdl_wait_prior_grids();
  const int target = a.epoch[b] + 1;
  const long long tid8 = (static_cast<long long>(b) * blockDim.x + threadIdx.x) * 8;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * 8;
  auto fail_out = [&]() {   // error word + NaN output for this block's share; the epoch is NOT advanced
    if (threadIdx.x == 0) atomicExch(a.epoch + kArErrIndex, 1);
    const uint4 nan4 = make_uint4(0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu);  // NaN in fp16 and bf16
    for (long long i = tid8; i < a.elems; i += stride) *reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.out) + i) = nan4;
  };
  if (!ar_block_barrier(a, b, 0, target)) {
    fail_out();
    return;
  }
  if (!TWO_SHOT) {
    for (long long i = tid8; i < a.elems; i += stride)
      *reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.out) + i) = ar_sum8<T>(a, i);
  } else {
    const long long slice = ((a.elems / 8 + a.world - 1) / a.world) * 8;  // elements per rank slice
    const long long lo = slice * a.rank;
    const long long hi = lo + slice < a.elems ? lo + slice : a.elems;
    for (long long i = lo + tid8; i < hi; i += stride)
      *reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.res[a.rank]) + i) = ar_sum8<T>(a, i);
    __syncthreads();  // the release in the barrier below is cumulative over the block's result stores
    const int target1 = a.epoch[kArMaxBlocks + b] + 1;  // the second bank only advances in two-shot calls
    if (!ar_block_barrier(a, b, 1, target1)) {
      fail_out();
      return;
    }
    if (threadIdx.x == 0) a.epoch[kArMaxBlocks + b] = target1;
    for (int w = 0; w < a.world; ++w) {
      const long long wlo = slice * w;
      const long long whi = wlo + slice < a.elems ? wlo + slice : a.elems;
      for (long long i = wlo + tid8; i < whi; i += stride)
        *reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.out) + i) =
            ld_relaxed_sys_v4(reinterpret_cast<const T*>(a.res[w]) + i);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) a.epoch[b] = target;
}

}  // namespace b200
