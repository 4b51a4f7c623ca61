// Host side of the tcgen05 regime: tile-shape choice, TMA descriptors, launch.
// (the reference's host logic for large M is the `par` M-loop of marlin_mm, gptq_marlin_mm.cu:161-206;
//  here one persistent launch covers any M.)
#include <cstdlib>
#include <cstring>

#include "../../include/b200_w4a16.h"
#include "host_common.h"
#include "prefill_kernel.cuh"

namespace b200 {

namespace {

struct TileShape {
  int nsub, bm;
  float weight;  // relative per-tile efficiency (bigger tiles amortise dequant + operand traffic)
  const char* name;
};
const TileShape kTiles[kPrefillTiles] = {
    // weights: measured relative throughput on B200 at M = 4096 (profiles/r01, 8 dequant warps per CTA)
    {2, 256, 0.83f, "tc/bn256/bm256"},
    {1, 256, 1.00f, "tc/bn128/bm256"},
    {2, 128, 0.73f, "tc/bn256/bm128"},
    {1, 128, 0.66f, "tc/bn128/bm128"},
    // split-K over a thread-block cluster (mid M: the N x M grid alone cannot fill the machine)
    {1, 64, 0.f, "tc/bn128/bm64/splitk"},
    {1, 128, 0.f, "tc/bn128/bm128/splitk"},
    // cta_group::2: a CTA pair per 256 columns x 256 tokens (whole K, persistent)
    {1, 256, 0.f, "tc/pair256/bm256"},
};

template <typename T, int NSUB, int BM, int STAGES, int QSTAGES, int DW, bool GROUPED, bool SPLIT, bool PAIR = false>
int launch_one(const GemmCall& g, int S, int* query) {
  using Cfg = PrefillCfg<T, NSUB, BM, STAGES, QSTAGES, DW, GROUPED, SPLIT, PAIR>;
  auto kern = w4a16_prefill_kernel<T, NSUB, BM, STAGES, QSTAGES, DW, GROUPED, SPLIT, PAIR>;
  static bool attr_set = false;  // benign race: idempotent
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess)
      return fail(B200_W4A16_ECUDA, "cudaFuncSetAttribute(smem=%d): %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
    attr_set = true;
  }
  if (query) {
    // clusters of S such CTAs the device can hold at once (cached per kernel and S)
    static int cache[4] = {0, 0, 0, 0};
    int li = 0;
    while ((1 << li) < S) ++li;
    if (cache[li] == 0) {
      cudaLaunchConfig_t qc;
      std::memset(&qc, 0, sizeof(qc));
      qc.gridDim = dim3(S * 64);
      qc.blockDim = dim3(Cfg::THREADS);
      qc.dynamicSmemBytes = Cfg::SMEM_BYTES;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = S;
      qa[0].val.clusterDim.y = 1;
      qa[0].val.clusterDim.z = 1;
      qc.attrs = qa;
      qc.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &qc) != cudaSuccess || n < 1) {
        (void)cudaGetLastError();
        n = -1;
      }
      cache[li] = n;
    }
    *query = cache[li];
    return 0;
  }
  CUtensorMap tm_w, tm_s, tm_a;
  {  // weights: int32 [K/16][N/64][128]; chunks past N/64 are zero-filled
    const uint64_t dims[3] = {128, static_cast<uint64_t>(g.n / 64), static_cast<uint64_t>(g.k / 16)};
    const uint64_t strides[2] = {512, static_cast<uint64_t>(g.n) * 8};
    const uint32_t box[3] = {128, Cfg::CHUNKS, 4};
    if (int rc = get_tensor_map(&tm_w, g.qweight, CU_TENSOR_MAP_DATA_TYPE_INT32, 3, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_NONE))
      return rc;
  }
  if (GROUPED) {  // scales: T [K/128][N/64][64]
    const uint64_t dims[3] = {64, static_cast<uint64_t>(g.n / 64), static_cast<uint64_t>(g.k / 128)};
    const uint64_t strides[2] = {128, static_cast<uint64_t>(g.n) * 2};
    const uint32_t box[3] = {64, Cfg::CHUNKS, 1};
    if (int rc = get_tensor_map(&tm_s, g.scales, Elem<T>::kTmaType, 3, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_NONE))
      return rc;
  } else {
    tm_s = tm_w;
  }
  {  // activations: T [M][K], [BM x 64] boxes, 128B swizzle, rows past M zero-filled
    const uint64_t dims[2] = {static_cast<uint64_t>(g.k), static_cast<uint64_t>(g.m)};
    const uint64_t strides[1] = {static_cast<uint64_t>(g.k) * 2};
    const uint32_t box[2] = {64, static_cast<uint32_t>(Cfg::A_ROWS)};
    if (int rc = get_tensor_map(&tm_a, g.a, Elem<T>::kTmaType, 2, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_128B))
      return rc;
  }
  PrefillArgs args;
  args.c = g.c;
  args.scales = g.scales;
  args.bias = g.bias;
  args.residual = g.residual;
  args.out_scale = out_scale_bits(g.out_scale, g.dtype);
  args.m = g.m;
  args.n = g.n;
  args.k = g.k;
  args.n_tiles = PAIR ? (g.n + 255) / 256 : (g.n + Cfg::BN - 1) / Cfg::BN;
  args.m_tiles = (g.m + BM - 1) / BM;
  args.kblks = g.k / 64;
  args.qkv_a = g.qkv_a;
  args.qkv_b = g.qkv_b;
  const long long tiles = static_cast<long long>(args.n_tiles) * args.m_tiles;
  int grid = static_cast<int>(tiles < sm_count() ? tiles : sm_count());
  if (SPLIT) {
    if (S > args.kblks) S = args.kblks;
    while (S & (S - 1)) --S;
    grid = static_cast<int>(tiles) * S;
  }
  if (PAIR) {
    const int pairs = sm_count() / 2;
    grid = 2 * static_cast<int>(tiles < pairs ? tiles : pairs);
  }

  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = g.stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if ((SPLIT && S > 1) || PAIR) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = PAIR ? 2 : S;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (g.pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tm_w, tm_s, tm_a, args);
  if (e != cudaSuccess) return fail(B200_W4A16_ECUDA, "tcgen05 kernel launch: %s", cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

template <typename T, bool GROUPED>
int launch_tile(const GemmCall& g, int tile, int S, int* query) {
  switch (tile) {
    // tiles 0 and 2 (256 columns per CTA as two 128-column UMMA halves) are DISABLED: in long PDL chains they returned
    // wrong tiles about once in 1500 launches (0) and hung once in ~10 000 (2) — profiles/r02/r_nsub2_tiles_stress.txt;
    // cause not found before the end of round 2.  The scorer never preferred them (weights 0.83 / 0.73);
    // their ids stay reserved so that the other configuration ids do not move.
    case 0:
    case 2: {
      static const bool lab = std::getenv("B200_W4A16_LAB_NSUB2") != nullptr;   // diagnosis only
      if (!lab)
        return fail(B200_W4A16_EINVAL, "tcgen05 tile %s is disabled (intermittent failures under PDL chains, DESIGN.md 6e)", kTiles[tile].name);
      return tile == 0 ? launch_one<T, 2, 256, 3, 3, 8, GROUPED, false>(g, 1, query)
                       : launch_one<T, 2, 128, 4, 3, 8, GROUPED, false>(g, 1, query);
    }
    case 1: return launch_one<T, 1, 256, 4, 4, 8, GROUPED, false>(g, 1, query);
    case 3: return launch_one<T, 1, 128, 6, 6, 8, GROUPED, false>(g, 1, query);
    case 4: return launch_one<T, 1, 64, 2, 10, 8, GROUPED, true>(g, S, query);
    case 5: return launch_one<T, 1, 128, 2, 8, 8, GROUPED, true>(g, S, query);
    case 6: return launch_one<T, 1, 256, 6, 4, 8, GROUPED, false, true>(g, 1, query);
  }
  return fail(B200_W4A16_EINVAL, "tcgen05 tile id %d out of range", tile);
}

}  // namespace

const char* prefill_tile_name(int tile) {
  return (tile >= 0 && tile < kPrefillTiles) ? kTiles[tile].name : "";
}

namespace {
int dispatch(const GemmCall& g, int tile, int S, int* query) {
  const bool grouped = g.group_size != -1;
  if (S < 1) S = 1;
  if (S > 8) S = 8;
  if (g.dtype == 0)
    return grouped ? launch_tile<__half, true>(g, tile, S, query) : launch_tile<__half, false>(g, tile, S, query);
  return grouped ? launch_tile<__nv_bfloat16, true>(g, tile, S, query)
                 : launch_tile<__nv_bfloat16, false>(g, tile, S, query);
}
}  // namespace

// shape -> (tile, cluster size).  Large M: the widest whole-K tile whose wave quantisation on this
// device does not cost more than its size gains.  Mid M (one or two token tiles): split-K clusters,
// as wide as still fits the device in ONE wave of clusters.
void prefill_pick(const GemmCall& g, int* tile_out, int* S_out) {
  const int m = g.m, n = g.n;
  const int sms = sm_count();
  if (m <= 256) {
    // split-K tiles: two CTAs are resident per SM (87-110 KB each); clusters pack GPCs imperfectly, so
    // budget ~90 % of the 2 x SMs slots (measured on B200: 32 clusters of 8 are co-resident, 36 are
    // not; cudaOccupancyMaxActiveClusters under-reports this kernel, profiles/r01/README.md)
    const int t = m <= 64 ? 4 : 5;
    const int bm = kTiles[t].bm;
    const int tiles = ((n + 127) / 128) * ((m + bm - 1) / bm);
    const int kblks = g.k / 64;
    const int cap = (2 * sms * 9) / 10;
    int s = 8;
    while (s > 1 && (kblks / s < 4 || tiles * s > cap)) s >>= 1;
    if (m <= 128 || tiles <= cap) {
      *tile_out = t;
      *S_out = s;
      return;
    }
  }
  // CTA pairs (256 columns x 256 tokens per pair, half the activation traffic per SM) once there is about a wave of
  // pair tiles for the 74 pairs: measured per shape and M in profiles/r02/j_pair_vs_single_sweep.txt
  {
    const long long pair_tiles = static_cast<long long>((n + 255) / 256) * ((m + 255) / 256);
    if (m > 256 && pair_tiles >= 64) {
      *tile_out = kPrefillPairTile;
      *S_out = 1;
      return;
    }
  }
  int best = 0;
  float best_score = -1.f;
  for (int t = 1; t < 4; t += 2) {   // the single-CTA whole-K tiles in service: (128, 256) and (128, 128)
    const int bn = kTiles[t].nsub * 128, bm = kTiles[t].bm;
    const long long nt = (n + bn - 1) / bn, mt = (m + bm - 1) / bm, tiles = nt * mt;
    const long long waves = (tiles + sms - 1) / sms;
    // useful fraction of the issued MMA work: wave quantisation x padding of the last token / column tile
    const float wave_eff = static_cast<float>(tiles) / static_cast<float>(waves * sms);
    const float pad_eff = (static_cast<float>(m) / static_cast<float>(mt * bm)) *
                          (static_cast<float>(n) / static_cast<float>(nt * bn));
    const float score = wave_eff * pad_eff * kTiles[t].weight;
    if (score > best_score) {
      best_score = score;
      best = t;
    }
  }
  *tile_out = best;
  *S_out = 1;
}

int launch_prefill(const GemmCall& g, int tile, int S) { return dispatch(g, tile, S, nullptr); }
int prefill_max_clusters(const GemmCall& g, int tile, int S) {
  int n = -1;
  if (dispatch(g, tile, S, &n) != 0) return -1;
  return n;
}

}  // namespace b200
