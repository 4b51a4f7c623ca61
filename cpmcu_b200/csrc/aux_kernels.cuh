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
    uint32_t s_lo[4] = {0, 0, 0, 0}, s_hi[4] = {0, 0, 0, 0};
    if (GROUPED) {
      const int grp = (r * 16) / group_size;
      const uint4 sv = *reinterpret_cast<const uint4*>(scales + static_cast<size_t>(grp) * n +
                                                       ch * 64 + g * 8);
      broadcast_scale_pair(sv.x, s_lo[0], s_hi[0]);
      broadcast_scale_pair(sv.y, s_lo[1], s_hi[1]);
      broadcast_scale_pair(sv.z, s_lo[2], s_hi[2]);
      broadcast_scale_pair(sv.w, s_lo[3], s_hi[3]);
    }
    const uint32_t qq[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t w[4];
      Dequant<T>::template word<GROUPED>(qq[j], s_lo[j], s_hi[j], w);
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

}  // namespace b200
