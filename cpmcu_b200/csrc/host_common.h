// Host-side arithmetic shared by the dispatcher and the kernel launchers.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace b200 {

// one GEMM launch request, shared by the dispatcher (w4a16_abi.cu) and the kernel launchers
struct GemmCall {
  const void* a;
  const int32_t* qweight;
  const void* scales;
  int32_t* workspace;
  void* c;
  float* c_tmp;
  const void* bias;
  const void* residual;
  float out_scale;  // 0 = none
  int qkv_a, qkv_b; // > 0: q | k | v written split (opts.qkv_q_dim / qkv_kv_dim)
  bool gated;       // fused SiLU(gate) * up epilogue over the [gate | up] columns
  int m, n, k;
  int group_size;
  int dtype;
  bool pdl;
  cudaStream_t stream;
  // fused all-reduce (b200_w4a16_gemm_ar): device pointers copied from the caller's host arrays; ar_world <= 1: off
  void* ar_recv[8];
  int* ar_flag[8];
  int* ar_seen;
  long long ar_slot_elems;
  int ar_world, ar_rank;
};

// defined in w4a16_abi.cu
uint32_t out_scale_bits(float s, int dtype);  // 0 = none, else 0x10000 | bits of T(s)
int fail(int code, const char* fmt, ...);
int sm_count();
extern std::atomic<long long> g_launches;
extern std::atomic<int> g_default_pdl;  // b200_w4a16_set_pdl
// rank 2..4 tiled TMA descriptor (cached); dims/strides innermost first (strides[] in bytes for dims 1..)
int get_tensor_map(CUtensorMap* out, const void* ptr, CUtensorMapDataType type, int rank,
                   const uint64_t* dims, const uint64_t* strides, const uint32_t* box,
                   CUtensorMapSwizzle swz);
// tuning aid (b200_w4a16_debug_timeline): the next per-launch timestamp block, or null
unsigned long long* debug_next_slot();

// ---- streaming (decode) regime: geometries of w4a16_decode_kernel and the launch plan ----
struct DecodeGeom { int chunks, krows, warps, stages; };
constexpr int kGeoms = 9;
// 0: the default (256 columns x 128 k per stage, 8 consumer warps)      1: the same tile with 64-k stages (K % 128 != 0)
// 2: 64 columns x 512 k, for whole-column launches                     3, 4: 16 consumer warps (more warps per scheduler)
// 5: 128 columns x 256 k                                              6: the default tile with a 10-deep ring (one CTA per SM)
// 7, 8: tiles 0 and 5 on tcgen05 / tensor memory (w4a16_decode_tc_kernel, decode_tc_kernel.cuh); token rows 16 or 32
constexpr DecodeGeom kGeom[kGeoms] = {{4, 8, 8, 5}, {4, 4, 8, 8}, {1, 32, 8, 4}, {4, 8, 16, 5}, {4, 16, 16, 3}, {2, 16, 8, 5}, {4, 8, 8, 10}, {4, 8, 8, 5}, {2, 16, 8, 5}};
constexpr int kTcGeom = 7, kTcGeom2 = 8;   // tensor-memory kernel: 256- and 128-column tiles
constexpr int kWholeColGeom = 2;

// Upper bound on the number of CTAs whose unit ranges intersect one column of `kblks` units when
// `units` units are split evenly over `grid` CTAs (every CTA gets floor or ceil of units/grid):
// one partial run at each end plus the whole runs between them.  This is the per-column slot
// stride reserved in c_tmp; the kernel computes the exact count per column itself.
inline int decode_slot_max(int kblks, long long units, int grid) {
  long long minrun = units / grid;
  if (minrun < 1) minrun = 1;
  return static_cast<int>((kblks - 1 + minrun - 1) / minrun) + 1;
}

struct DecodePlan {
  int geom;        // index into kGeom
  int grid;        // CTAs
  int colpar;      // 1: CTAs take whole columns (no split column: c_tmp and workspace untouched)
  int slot_max;    // fp32 slots per column tile reserved in c_tmp
  int ncols, kblks;
  long long need_bytes;  // bytes of c_tmp this plan touches
};

// defined in w4a16_decode_fp16.cu / w4a16_decode_bf16.cu: launch one slab (g.m <= 32) with the given plan
int launch_decode_fp16(const GemmCall& g, const DecodePlan& pl, int nt);
int launch_decode_bf16(const GemmCall& g, const DecodePlan& pl, int nt);

// defined in w4a16_prefill.cu: the tcgen05 regime
constexpr int kPrefillTiles = 7;  // (NSUB,BM) = (2,256) (1,256) (2,128) (1,128); split-K (1,64) (1,128); CTA pair (256 cols, 256)
constexpr int kPrefillPairTile = 6;
constexpr int kPrefillWholeK = 4; // tiles below this index are whole-K (persistent); the rest are split-K
void prefill_pick(const GemmCall& g, int* tile, int* S);
const char* prefill_tile_name(int tile);
int launch_prefill(const GemmCall& g, int tile, int S);
int prefill_max_clusters(const GemmCall& g, int tile, int S);

}  // namespace b200
