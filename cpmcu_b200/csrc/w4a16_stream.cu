// Host side of the streaming regime (M <= 32 per launch): geometry / cluster-size choice, TMA
// descriptors, cluster + PDL launch.  (reference host logic: gptq_marlin_mm.cu:34-207)
#include <cstdlib>
#include <cstring>

#include "../../include/b200_w4a16.h"
#include "host_common.h"
#include "stream_kernel.cuh"

namespace b200 {

unsigned long long* g_dbg_ts = nullptr;
long long g_dbg_ts_launches = 0;
std::atomic<long long> g_dbg_ts_next{0};

namespace {

struct Geom {
  int chunks, krows, warps, stages;
  const char* name;
};
const Geom kGeoms[kStreamGeoms] = {
    {2, 8, 8, 6, "st/tn128/kb128/w8"},   {1, 8, 8, 8, "st/tn64/kb128/w8"},   {4, 8, 8, 4, "st/tn256/kb128/w8"},
    {2, 8, 4, 6, "st/tn128/kb128/w4"},   {4, 4, 8, 6, "st/tn256/kb64/w8"},   {2, 4, 8, 8, "st/tn128/kb64/w8"},
    {4, 8, 16, 5, "st/tn256/kb128/w16"}, {2, 8, 16, 6, "st/tn128/kb128/w16"}, {4, 8, 8, 3, "st/tn256/kb128/w8/s3"},
    {1, 8, 4, 8, "st/tn64/kb128/w4"},
};

template <typename T, int NT, int CHUNKS, int KROWS, int WARPS, int STAGES, bool GROUPED>
int launch_one(const GemmCall& g, int S, int* query) {
  using Cfg = StreamCfg<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED>;
  auto kern = w4a16_stream_kernel<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED>;
  static bool attr_set = false;  // benign race: idempotent
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess)
      return fail(B200_W4A16_ECUDA, "cudaFuncSetAttribute(smem=%d): %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
    attr_set = true;
  }
  if (query) {
    // how many clusters of S such CTAs the device can hold at once (cached per kernel and S)
    static int cache[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const int li = S;
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
  if (g.k % Cfg::KB != 0)
    return fail(B200_W4A16_EINVAL, "streaming geometry with k-block %d does not divide K = %d", Cfg::KB, g.k);
  CUtensorMap tm_w, tm_s, tm_a;
  {  // weights: int32 [K/16][N/64][128]
    const uint64_t dims[3] = {128, static_cast<uint64_t>(g.n / 64), static_cast<uint64_t>(g.k / 16)};
    const uint64_t strides[2] = {512, static_cast<uint64_t>(g.n) * 8};
    const uint32_t box[3] = {128, CHUNKS, KROWS};
    if (int rc = get_tensor_map(&tm_w, g.qweight, CU_TENSOR_MAP_DATA_TYPE_INT32, 3, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_NONE))
      return rc;
  }
  if (GROUPED) {  // scales: T [K/128][N/64][64]
    const uint64_t dims[3] = {64, static_cast<uint64_t>(g.n / 64), static_cast<uint64_t>(g.k / 128)};
    const uint64_t strides[2] = {128, static_cast<uint64_t>(g.n) * 2};
    const uint32_t box[3] = {64, CHUNKS, 1};
    if (int rc = get_tensor_map(&tm_s, g.scales, Elem<T>::kTmaType, 3, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_NONE))
      return rc;
  } else {
    tm_s = tm_w;
  }
  {  // activations: T [M][K], 64-wide (128-byte) boxes, 128B swizzle, rows past M zero-filled
    const uint64_t dims[2] = {static_cast<uint64_t>(g.k), static_cast<uint64_t>(g.m)};
    const uint64_t strides[1] = {static_cast<uint64_t>(g.k) * 2};
    const uint32_t box[2] = {64, NT * 8};
    if (int rc = get_tensor_map(&tm_a, g.a, Elem<T>::kTmaType, 2, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_128B))
      return rc;
  }
  StreamArgs args;
  args.c = g.c;
  args.scales = g.scales;
  args.bias = g.bias;
  args.residual = g.residual;
  args.out_scale = out_scale_bits(g.out_scale, g.dtype);
  args.m = g.m;
  args.n = g.n;
  args.k = g.k;
  args.kblks = g.k / Cfg::KB;
  static const int dbg_skip = [] { const char* e = std::getenv("B200_W4A16_DBG_SKIP"); return (e && e[0] == '1') ? 1 : 0; }();
  args.dbg_skip = dbg_skip;
  args.dbg_ts = stream_debug_next_slot();
  if (S > args.kblks) S = args.kblks;
  const int ncols = (g.n + Cfg::TN - 1) / Cfg::TN;

  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(ncols * S);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = g.stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (S > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = S;
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
  if (e != cudaSuccess)
    return fail(B200_W4A16_ECUDA, "streaming kernel launch (grid %d, cluster %d, smem %d): %s", ncols * S, S,
                Cfg::SMEM_BYTES, cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

template <typename T, int NT, bool GROUPED>
int launch_geom(const GemmCall& g, int geom, int S, int* query) {
  switch (geom) {
    case 0: return launch_one<T, NT, 2, 8, 8, 6, GROUPED>(g, S, query);
    case 1: return launch_one<T, NT, 1, 8, 8, 8, GROUPED>(g, S, query);
    case 2: return launch_one<T, NT, 4, 8, 8, 4, GROUPED>(g, S, query);
    case 3: return launch_one<T, NT, 2, 8, 4, 6, GROUPED>(g, S, query);
    case 4: return launch_one<T, NT, 4, 4, 8, 6, GROUPED>(g, S, query);
    case 5: return launch_one<T, NT, 2, 4, 8, 8, GROUPED>(g, S, query);
    case 6: return launch_one<T, NT, 4, 8, 16, 5, GROUPED>(g, S, query);
    case 7: return launch_one<T, NT, 2, 8, 16, 6, GROUPED>(g, S, query);
    case 8: return launch_one<T, NT, 4, 8, 8, 3, GROUPED>(g, S, query);
    case 9: return launch_one<T, NT, 1, 8, 4, 8, GROUPED>(g, S, query);
  }
  return fail(B200_W4A16_EINVAL, "streaming geometry %d out of range", geom);
}

template <typename T, bool GROUPED>
int launch_nt(const GemmCall& g, int geom, int S, int* query) {
  if (g.m <= 8) return launch_geom<T, 1, GROUPED>(g, geom, S, query);
  if (g.m <= 16) return launch_geom<T, 2, GROUPED>(g, geom, S, query);
  return launch_geom<T, 4, GROUPED>(g, geom, S, query);
}

int dispatch(const GemmCall& g, int geom, int S, int* query) {
  const bool grouped = g.group_size != -1;
  if (S < 1) S = 1;
  if (S > 8) S = 8;
  if (g.dtype == 0)
    return grouped ? launch_nt<__half, true>(g, geom, S, query) : launch_nt<__half, false>(g, geom, S, query);
  return grouped ? launch_nt<__nv_bfloat16, true>(g, geom, S, query)
                 : launch_nt<__nv_bfloat16, false>(g, geom, S, query);
}

// clusters of S CTAs of this geometry the device holds at once (-1 = unknown)
int max_clusters(const GemmCall& g, int geom, int S) {
  int n = -1;
  if (dispatch(g, geom, S, &n) != 0) return -1;
  return n;
}

}  // namespace

const char* stream_geom_name(int geom) {
  return (geom >= 0 && geom < kStreamGeoms) ? kGeoms[geom].name : "";
}

// shape -> (geometry, cluster size).  Rules distilled from the B200 sweeps (profiles/r01): the whole
// grid must be resident at once (a second wave of clusters costs more than any tile-shape gain), and
// within that the split is as wide as possible: 4-warp CTAs of 128 columns for M <= 16, 8-warp CTAs
// for M <= 32, 256-column tiles when N alone already fills the machine.
void stream_pick(const GemmCall& g, int* geom, int* S) {
  const int nt = g.m <= 8 ? 1 : (g.m <= 16 ? 2 : 4);
  const bool k128 = (g.k % 128) == 0;
  const int sms = sm_count();
  const int ncols128 = (g.n + 127) / 128;
  int gm;
  int smax = 8;
  if (!k128) {
    gm = 5;  // tn128 / kb64
  } else if (ncols128 > sms + sms / 3) {
    // N alone gives >= ~200 column tiles (gate_up-like)
    if (nt == 1) { gm = 2; smax = 2; }        // tn256, pairs
    else if (nt == 2) { gm = 3; smax = 1; }   // tn128 / 4 warps
    else { gm = 2; smax = 1; }                // tn256
  } else {
    gm = nt == 4 ? 0 : 3;
    if (nt == 4) smax = 4;
  }
  const int tn = kGeoms[gm].chunks * 64, kb = kGeoms[gm].krows * 16;
  const int ncols = (g.n + tn - 1) / tn, kblks = g.k / kb;
  int s = smax;
  while (s > 1 && (kblks / s < 2)) s >>= 1;
  // the grid must fit in one wave of clusters
  while (s > 1) {
    const int fit = max_clusters(g, gm, s);
    if (fit < 0 || ncols <= fit) break;
    s >>= 1;
  }
  *geom = gm;
  *S = s;
}

// debug timeline: one 1024-CTA record block per launch, in launch order (null when switched off)
unsigned long long* stream_debug_next_slot() {
  if (!g_dbg_ts) return nullptr;
  const long long idx = g_dbg_ts_next.fetch_add(1, std::memory_order_relaxed);
  return idx < g_dbg_ts_launches ? g_dbg_ts + idx * 8192 : nullptr;
}

void stream_debug_timeline(void* buf, long long launches) {
  g_dbg_ts = static_cast<unsigned long long*>(buf);
  g_dbg_ts_launches = buf ? launches : 0;
  g_dbg_ts_next.store(0);
}

int launch_stream(const GemmCall& g, int geom, int S) { return dispatch(g, geom, S, nullptr); }
int stream_max_clusters(const GemmCall& g, int geom, int S) { return max_clusters(g, geom, S); }

}  // namespace b200
