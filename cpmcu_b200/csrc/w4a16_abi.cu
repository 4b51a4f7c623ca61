// C ABI + host-side dispatch of libb200_w4a16.so (see include/b200_w4a16.h).
//
// Host logic mirrors what the reference does on the host for this path
// (gptq_marlin.cu:42-86 validation, gptq_marlin_mm.cu:34-207 configuration + M loop) but for the
// sm_100a kernels: it validates, picks a kernel configuration from the shape, encodes (and caches)
// the TMA descriptors, and launches on the caller's stream.  No allocation, no host<->device sync,
// graph-capturable.
#include "../../include/b200_w4a16.h"

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>

#include "aux_kernels.cuh"
#include "decode_kernel.cuh"
#include "host_common.h"

namespace b200 {

thread_local std::string g_last_error;
std::atomic<long long> g_launches{0};
std::atomic<int> g_default_pdl{[] {
  const char* e = std::getenv("B200_W4A16_PDL");
  return (e && e[0] == '1') ? 1 : 0;
}()};

// which streaming kernel serves M <= 32 by default: 0 = persistent stream-K (decode_kernel.cuh),
// 1 = cluster split-K (stream_kernel.cuh); environment B200_W4A16_STREAM_V2 overrides
bool g_stream_v2 = [] {
  const char* e = std::getenv("B200_W4A16_STREAM_V2");
  return e ? (e[0] == '1') : false;
}();

// Per-shape mix of the two streaming kernels (B200_W4A16_STREAM_V2 unset): the cluster split-K kernel
// reduces through DSMEM (about 1 us) where the persistent stream-K kernel ends in a cross-CTA reduction
// through L2 (2.5-4 us), so it wins while a launch streams few bytes; the persistent kernel wins on the
// big shapes (no wave quantisation).  Thresholds in int4 weight bytes (K*N/2) for M <= 8 / 9..16 / 17..32,
// from the B200 measurements in profiles/r01 (b_bench_ours_stream_v2.json against g_bench_ours_final.json
// and the i_* mixed-chain runs); 0 = never.
long long g_mix_bytes[3] = {
    [] { const char* e = std::getenv("B200_W4A16_MIX_BYTES_M8"); return e ? std::atoll(e) : 0ll; }(),
    [] { const char* e = std::getenv("B200_W4A16_MIX_BYTES_M16"); return e ? std::atoll(e) : 0ll; }(),
    [] { const char* e = std::getenv("B200_W4A16_MIX_BYTES_M32"); return e ? std::atoll(e) : 0ll; }(),
};
bool g_stream_v2_forced = std::getenv("B200_W4A16_STREAM_V2") != nullptr;
bool use_cluster_streamer(int mm, long long n, long long k) {
  if (g_stream_v2_forced) return g_stream_v2;
  const long long bytes = n * k / 2;
  const long long thr = g_mix_bytes[mm <= 8 ? 0 : (mm <= 16 ? 1 : 2)];
  return bytes <= thr;
}

// cap on resident CTAs per SM of the first-generation persistent streaming kernel
int g_v1_occ = [] {
  const char* e = std::getenv("B200_W4A16_V1_OCC");
  const int v = e ? std::atoi(e) : 1;
  return v < 1 ? 1 : (v > 4 ? 4 : v);
}();

uint32_t out_scale_bits(float v, int dtype) {
  if (v == 0.0f) return 0u;
  uint16_t u;
  if (dtype == 0) {
    const __half h = __float2half_rn(v);  // T(v), as elementwise_scale_kernel rounds it (elementwise.cuh:39)
    std::memcpy(&u, &h, 2);
  } else {
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    std::memcpy(&u, &h, 2);
  }
  return 0x10000u | u;
}

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  // like the reference's VALUE_CHECK / ERROR_CHECK (src/utils.cuh:54-66) the message also goes to stderr
  fprintf(stderr, "b200_w4a16 %s: %s\n", code == B200_W4A16_EINVAL ? "value error" : "error", buf);
  return code;
}

// ------------------------------------------------------------------------------------------
// device properties (cached) and the driver entry point for descriptor encoding
// ------------------------------------------------------------------------------------------
int sm_count() {
  static int sms = [] {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) {
      (void)cudaGetLastError();
      return 148;  // B200; only reached by the sizing helpers on a machine without a device
    }
    return v;
  }();
  return sms;
}

using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                              const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                              CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                              CUtensorMapFloatOOBfill);
EncodeFn encode_fn() {
  static EncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeFn>(p);
  }();
  return fn;
}

struct MapKey {
  const void* ptr;
  uint64_t d0, d1, d2, d3, s1, s2, s3;
  uint32_t b0, b1, b2, b3, misc, pad;
  bool operator==(const MapKey& o) const { return std::memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
    return static_cast<size_t>(h);
  }
};
std::mutex g_map_mu;
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;

// rank 2, 3 or 4 tiled descriptor; dims/strides innermost first (strides[] in bytes for dims 1..)
int get_tensor_map(CUtensorMap* out, const void* ptr, CUtensorMapDataType type, int rank,
                   const uint64_t* dims, const uint64_t* strides, const uint32_t* box,
                   CUtensorMapSwizzle swz) {
  MapKey key;
  std::memset(&key, 0, sizeof(key));
  key.ptr = ptr;
  key.d0 = dims[0];
  key.d1 = dims[1];
  key.d2 = rank > 2 ? dims[2] : 0;
  key.d3 = rank > 3 ? dims[3] : 0;
  key.s3 = rank > 3 ? strides[2] : 0;
  key.b3 = rank > 3 ? box[3] : 0;
  key.s1 = strides[0];
  key.s2 = rank > 2 ? strides[1] : 0;
  key.b0 = box[0];
  key.b1 = box[1];
  key.b2 = rank > 2 ? box[2] : 0;
  key.misc = (static_cast<uint32_t>(type) << 8) | (static_cast<uint32_t>(swz) << 4) | rank;
  {
    std::lock_guard<std::mutex> lk(g_map_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) {
      *out = it->second;
      return 0;
    }
  }
  EncodeFn enc = encode_fn();
  if (!enc) return fail(B200_W4A16_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t gdim[4] = {dims[0], dims[1], rank > 2 ? dims[2] : 1, rank > 3 ? dims[3] : 1};
  cuuint64_t gstr[3] = {strides[0], rank > 2 ? strides[1] : 0, rank > 3 ? strides[2] : 0};
  cuuint32_t bx[4] = {box[0], box[1], rank > 2 ? box[2] : 1, rank > 3 ? box[3] : 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, type, rank, const_cast<void*>(ptr), gdim, gstr, bx, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(B200_W4A16_ECUDA,
                "cuTensorMapEncodeTiled failed (%d): ptr=%p rank=%d dims=[%llu,%llu,%llu] "
                "strides=[%llu,%llu] box=[%u,%u,%u]",
                static_cast<int>(r), ptr, rank, (unsigned long long)gdim[0],
                (unsigned long long)gdim[1], (unsigned long long)gdim[2], (unsigned long long)gstr[0],
                (unsigned long long)gstr[1], bx[0], bx[1], bx[2]);
  std::lock_guard<std::mutex> lk(g_map_mu);
  if (g_maps.size() > 65536) g_maps.clear();
  g_maps.emplace(key, *out);
  return 0;
}

// ------------------------------------------------------------------------------------------
// decode-regime configurations
// ------------------------------------------------------------------------------------------
struct DecodeConfig {
  const char* name;
  int nt, chunks, krows, warps, stages;
  bool grouped;
  int dtype;
  int (*launch)(const GemmCall&);
  int smem;
};

template <typename T, int NT, int CHUNKS, int KROWS, int WARPS, int STAGES, bool GROUPED, bool GATED = false>
int launch_decode(const GemmCall& g) {
  using Cfg = DecodeCfg<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED>;
  auto kern = w4a16_decode_kernel<T, NT, CHUNKS, KROWS, WARPS, STAGES, GROUPED, GATED>;
  if (g.gated != GATED) return fail(B200_W4A16_ERUNTIME, "internal: gated launch through a plain kernel");
  static bool attr_set = false;  // benign race: idempotent
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess)
      return fail(B200_W4A16_ECUDA, "cudaFuncSetAttribute(smem=%d): %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
    attr_set = true;
  }
  CUtensorMap tm_w, tm_s, tm_a;
  if (g.gated) {
    // fused gate_up: the tile's first CHUNKS/2 chunks come from the gate half [0, I), the others from the
    // same columns of the up half [I, 2I) — ONE 4-d box over the views int32 [K/16][2][I/64][128] and
    // T [K/128][2][I/64][64]; the bytes land exactly where the 3-d box of CHUNKS chunks would put them
    if ((CHUNKS & 1) || (g.n / 2) % (Cfg::TN / 2) != 0 || !GROUPED)
      return fail(B200_W4A16_EINVAL, "gated_silu needs group_size 128 and N/2 a multiple of %d, got N = %d", Cfg::TN / 2, g.n);
    const uint64_t ich = static_cast<uint64_t>(g.n / 2 / 64);
    {
      const uint64_t dims[4] = {128, ich, 2, static_cast<uint64_t>(g.k / 16)};
      const uint64_t strides[3] = {512, ich * 512, static_cast<uint64_t>(g.n) * 8};
      const uint32_t box[4] = {128, CHUNKS / 2, 2, KROWS};
      if (int rc = get_tensor_map(&tm_w, g.qweight, CU_TENSOR_MAP_DATA_TYPE_INT32, 4, dims, strides, box,
                                  CU_TENSOR_MAP_SWIZZLE_NONE))
        return rc;
    }
    {
      const uint64_t dims[4] = {64, ich, 2, static_cast<uint64_t>(g.k / 128)};
      const uint64_t strides[3] = {128, ich * 128, static_cast<uint64_t>(g.n) * 2};
      const uint32_t box[4] = {64, CHUNKS / 2, 2, static_cast<uint32_t>(Cfg::SG)};
      if (int rc = get_tensor_map(&tm_s, g.scales, Elem<T>::kTmaType, 4, dims, strides, box,
                                  CU_TENSOR_MAP_SWIZZLE_NONE))
        return rc;
    }
  } else {
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
    const uint32_t box[3] = {64, CHUNKS, static_cast<uint32_t>(Cfg::SG)};
    if (int rc = get_tensor_map(&tm_s, g.scales, Elem<T>::kTmaType, 3, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_NONE))
      return rc;
  } else {
    tm_s = tm_w;
  }
  }
  {  // activations: T [M][K], 64-wide (128-byte) boxes, 128B swizzle, rows past M zero-filled
    const uint64_t dims[2] = {static_cast<uint64_t>(g.k), static_cast<uint64_t>(g.m)};
    const uint64_t strides[1] = {static_cast<uint64_t>(g.k) * 2};
    const uint32_t box[2] = {64, NT * 8};
    if (int rc = get_tensor_map(&tm_a, g.a, Elem<T>::kTmaType, 2, dims, strides, box,
                                CU_TENSOR_MAP_SWIZZLE_128B))
      return rc;
  }
  DecodeArgs args;
  args.gated = g.gated ? 1 : 0;
  args.c = g.c;
  args.c_tmp = g.c_tmp;
  args.workspace = g.workspace;
  args.scales = g.scales;
  args.bias = g.bias;
  args.residual = g.residual;
  args.out_scale = out_scale_bits(g.out_scale, g.dtype);
  args.m = g.m;
  args.n = g.n;
  args.k = g.k;
  args.ncols = g.gated ? (g.n / 2) / (Cfg::TN / 2) : (g.n + Cfg::TN - 1) / Cfg::TN;
  args.kblks = g.k / Cfg::KB;
  const long long units = static_cast<long long>(args.ncols) * args.kblks;
  // persistent CTAs: one per SM (B200_W4A16_V1_OCC raises it for experiments, if resources allow)
  static int occ = 0;
  if (occ == 0) {
    int o = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kern, Cfg::THREADS, Cfg::SMEM_BYTES) != cudaSuccess || o < 1) o = 1;
    occ = o < g_v1_occ ? o : g_v1_occ;
  }
  const long long slots = static_cast<long long>(sm_count()) * occ;
  int grid = static_cast<int>(units < slots ? units : slots);
  // column-parallel geometry: CTAs take whole 64-column tiles (no split column, hence no cross-CTA
  // reduction tail); some SMs idle when there are fewer tiles than SMs
  args.colpar = (CHUNKS == 1 && KROWS >= 32) ? 1 : 0;
  if (args.colpar) {
    grid = static_cast<int>(args.ncols < slots ? args.ncols : slots);
    args.slot_max = 1;
  } else {
    // a column may be shared by at most 15 CTAs (16 ints of `workspace` per 64 columns: one flag per
    // sharer + the completion counter)
    while (grid > 1 && decode_slot_max(args.kblks, units, grid) > 15) --grid;
    args.slot_max = decode_slot_max(args.kblks, units, grid);
  }
  args.dbg_ts = stream_debug_next_slot();

  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = g.stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g.pdl ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tm_w, tm_s, tm_a, args);
  if (e != cudaSuccess) return fail(B200_W4A16_ECUDA, "decode kernel launch: %s", cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

#define DCFG(T, DT, NT, CH, KR, WP, ST, GR)                                              \
  {#T "/nt" #NT "/tn" #CH "x64/kb" #KR "x16/w" #WP "/s" #ST "/" #GR, NT, CH, KR, WP, ST, GR, DT, \
   &launch_decode<T, NT, CH, KR, WP, ST, GR>, DecodeCfg<T, NT, CH, KR, WP, ST, GR>::SMEM_BYTES}

// geometry families (id within family = index):  chunks, krows, warps, stages
#define DECODE_FAMILY(T, DT, NT, GR)                                                    \
  DCFG(T, DT, NT, 2, 8, 8, 8, GR), DCFG(T, DT, NT, 4, 8, 8, 5, GR), DCFG(T, DT, NT, 4, 4, 8, 8, GR), \
      DCFG(T, DT, NT, 8, 4, 8, 5, GR), DCFG(T, DT, NT, 4, 8, 16, 5, GR), DCFG(T, DT, NT, 2, 8, 16, 8, GR), \
      DCFG(T, DT, NT, 2, 16, 8, 5, GR), DCFG(T, DT, NT, 1, 16, 4, 8, GR), DCFG(T, DT, NT, 3, 16, 12, 3, GR), \
      DCFG(T, DT, NT, 1, 32, 8, 4, GR)
constexpr int kGeoms = 10;
constexpr int kColparGeom = 9;  // 64-column x 512-k stages, launched column-parallel (see launch_decode)

const DecodeConfig g_decode[] = {
    DECODE_FAMILY(__half, 0, 1, true),          DECODE_FAMILY(__half, 0, 2, true),
    DECODE_FAMILY(__half, 0, 4, true),          DECODE_FAMILY(__nv_bfloat16, 1, 1, true),
    DECODE_FAMILY(__nv_bfloat16, 1, 2, true),   DECODE_FAMILY(__nv_bfloat16, 1, 4, true),
    DECODE_FAMILY(__half, 0, 1, false),         DECODE_FAMILY(__half, 0, 2, false),
    DECODE_FAMILY(__half, 0, 4, false),         DECODE_FAMILY(__nv_bfloat16, 1, 1, false),
    DECODE_FAMILY(__nv_bfloat16, 1, 2, false),  DECODE_FAMILY(__nv_bfloat16, 1, 4, false),
};

// fused gate_up + SiLU(gate) * up: the two default geometries (pick_geometry) of every token-block width
template <typename T>
int launch_gated_t(const GemmCall& g, int nt, int geom) {
  if (geom == 1) {
    if (nt == 1) return launch_decode<T, 1, 4, 8, 8, 5, true, true>(g);
    if (nt == 2) return launch_decode<T, 2, 4, 8, 8, 5, true, true>(g);
    return launch_decode<T, 4, 4, 8, 8, 5, true, true>(g);
  }
  if (nt == 1) return launch_decode<T, 1, 4, 4, 8, 8, true, true>(g);
  if (nt == 2) return launch_decode<T, 2, 4, 4, 8, 8, true, true>(g);
  return launch_decode<T, 4, 4, 4, 8, 8, true, true>(g);
}
int launch_gated(const GemmCall& g, int nt, int geom) {
  if (geom != 1 && geom != 2) return fail(B200_W4A16_EINVAL, "gated_silu: unsupported geometry %d", geom);
  return g.dtype == 0 ? launch_gated_t<__half>(g, nt, geom) : launch_gated_t<__nv_bfloat16>(g, nt, geom);
}

// index into g_decode
int decode_index(int dtype, int nt, bool grouped, int geom) {
  const int nti = nt == 1 ? 0 : (nt == 2 ? 1 : 2);
  return ((grouped ? 0 : 6) + dtype * 3 + nti) * kGeoms + geom;
}

int nt_for_m(int m) { return m <= 8 ? 1 : (m <= 16 ? 2 : 4); }

// shape -> geometry of the persistent streaming kernel.  Measured on B200 (profiles/r01): the 256-column
// x 128-k stage with 8 consumer warps (4 k16-rows per warp and stage) has the fewest instructions per
// byte and leaves room for the next kernel's CTA on the same SM (PDL); stream-K balances any shape.
int pick_geometry(int m, int n, int k) {
  // column-parallel (no split column) against stream-K: an SM streams about 29 bytes/ns of int4 at
  // M <= 8 (24 at M <= 16, 17 beyond; in-loop rates of profiles/r01 g_timeline); stream-K keeps every SM
  // busy but ends in the cross-CTA reduction of the split columns (about 3.5 us at M <= 8, 5 us beyond)
  static const int colpar_mode = [] { const char* e = std::getenv("B200_W4A16_COLPAR"); return e ? std::atoi(e) : 0; }();
  if ((k % 512) == 0 && colpar_mode != 0) {
    const double rate = m <= 8 ? 29e3 : (m <= 16 ? 24e3 : 17e3);  // bytes per us and SM
    const double tail = m <= 8 ? 3.5 : 5.0;
    const double col_bytes = 32.0 * k, cols = n / 64.0;
    const int sms = sm_count();
    const double t_colpar = std::ceil(cols / sms) * col_bytes / rate;
    const double t_streamk = cols * col_bytes / (rate * sms) + tail;
    if (colpar_mode == 1 || t_colpar < t_streamk) return kColparGeom;
  }
  return (k % 128) == 0 ? 1 : 2;  // tn256/kb128/w8, or tn256/kb64/w8 when K is only a multiple of 64
}

// configuration ids: 1..count = streaming-kernel configurations, kPrefillBase.. = tcgen05 tile shapes
constexpr int kPrefillBase = 1001;
// cluster split-K streaming kernel: id = kStreamBase + geometry * 8 + (cluster size - 1)
constexpr int kStreamBase = 2001;
// largest M served by the streaming kernel; above it the tcgen05 kernel runs (split-K clusters up to
// M = 256, persistent whole-K tiles beyond).  Measured crossover on B200: profiles/r01 sweep_mid.
int kStreamMaxM = [] {
  const char* e = std::getenv("B200_W4A16_STREAM_MAX_M");
  const int v = e ? std::atoi(e) : 32;
  return v < 0 ? 0 : v;
}();
// tcgen05 ids: kPrefillBase + t (whole-K tiles, t < kPrefillWholeK), then
// kPrefillBase + kPrefillWholeK + (t - kPrefillWholeK) * 4 + log2(cluster size) for the split-K tiles
constexpr int kPrefillIds = kPrefillWholeK + (kPrefillTiles - kPrefillWholeK) * 4;
int prefill_id(int tile, int S) {
  if (tile < kPrefillWholeK) return kPrefillBase + tile;
  int l2 = 0;
  while ((1 << l2) < S) ++l2;
  return kPrefillBase + kPrefillWholeK + (tile - kPrefillWholeK) * 4 + l2;
}
void prefill_from_id(int id, int* tile, int* S) {
  const int r = id - kPrefillBase;
  if (r < kPrefillWholeK) {
    *tile = r;
    *S = 1;
  } else {
    *tile = kPrefillWholeK + (r - kPrefillWholeK) / 4;
    *S = 1 << ((r - kPrefillWholeK) & 3);
  }
}

struct Geom { int chunks, krows; };
const Geom kGeom[kGeoms] = {{2, 8}, {4, 8}, {4, 4}, {8, 4}, {4, 8}, {2, 8}, {2, 16}, {1, 16}, {3, 16}, {1, 32}};

}  // namespace b200

using namespace b200;

// ------------------------------------------------------------------------------------------
// exported functions
// ------------------------------------------------------------------------------------------
extern "C" {

const char* b200_w4a16_last_error(void) { return g_last_error.c_str(); }
const char* b200_w4a16_version(void) { return "b200-w4a16 0.1 (sm_100a)"; }
int64_t b200_w4a16_launch_count(void) { return g_launches.load(); }
void b200_w4a16_debug_timeline(void* device_buf, int64_t launches) { stream_debug_timeline(device_buf, launches); }
int b200_w4a16_set_pdl(int enable) { return g_default_pdl.exchange(enable ? 1 : 0); }

int b200_w4a16_reduce_max_m(int64_t m) {
  // mirror of marlin::determine_reduce_max_m(prob_m, max_par = 16), gptq_marlin_utils.cu:206-225
  if (m <= 16) return 16;
  if (m <= 32) return 32;
  if (m <= 48) return 48;
  if (m <= 64) return 64;
  int64_t par = (m + 63) / 64;
  if (par > 16) par = 16;
  return static_cast<int>(64 * par);
}

int64_t b200_w4a16_workspace_ints(int64_t n) { return (n / 64) * 16; }

int64_t b200_w4a16_reduce_bytes(int64_t m, int64_t n, int64_t k) {
  if (m <= 0 || n <= 0 || k <= 0 || n % 64 || k % 64) return -1;
  if (m > kStreamMaxM) return 0;  // the tcgen05 kernel accumulates in tensor memory: no scratch
  const int rows = static_cast<int>(m < 32 ? m : 32);
  // worst case over the geometries the heuristic (or force_config) may use
  int64_t worst = 0;
  for (int geo = 0; geo < kGeoms; ++geo) {
    const int tn = kGeom[geo].chunks * 64, kb = kGeom[geo].krows * 16;
    if (k % kb) continue;
    const long long ncols = (n + tn - 1) / tn, kblks = k / kb, units = ncols * kblks;
    // worst case over the occupancies launch_decode may pick (1 .. g_v1_occ CTAs per SM)
    for (int occ = 1; occ <= g_v1_occ; ++occ) {
      const long long slots = static_cast<long long>(sm_count()) * occ;
      const int grid = static_cast<int>(units < slots ? units : slots);
      const int64_t b2 = ncols * decode_slot_max(static_cast<int>(kblks), units, grid) * rows * tn * 4ll;
      if (b2 > worst) worst = b2;
    }
    const int grid = static_cast<int>(units < sm_count() ? units : sm_count());
    const int64_t bytes = ncols * decode_slot_max(static_cast<int>(kblks), units, grid) * rows * tn * 4ll;
    if (bytes > worst) worst = bytes;
  }
  return worst;
}

int b200_w4a16_pick_config(int64_t m, int64_t n, int64_t k, int group_size, int dtype) {
  if (m <= 0 || n <= 0 || k <= 0 || (dtype != 0 && dtype != 1)) return -1;
  if (m > kStreamMaxM) {
    GemmCall g;
    std::memset(&g, 0, sizeof(g));
    g.m = static_cast<int>(m);
    g.n = static_cast<int>(n);
    g.k = static_cast<int>(k);
    g.group_size = group_size;
    g.dtype = dtype;
    int tile = 0, S = 1;
    prefill_pick(g, &tile, &S);
    return prefill_id(tile, S);
  }
  const int mm = static_cast<int>(m < 32 ? m : 32);
  if (use_cluster_streamer(mm, n, k)) {
    GemmCall g;
    std::memset(&g, 0, sizeof(g));
    g.m = mm;
    g.n = static_cast<int>(n);
    g.k = static_cast<int>(k);
    g.group_size = group_size;
    g.dtype = dtype;
    int geom = 0, S = 1;
    stream_pick(g, &geom, &S);
    return kStreamBase + geom * 8 + (S - 1);
  }
  return 1 + decode_index(dtype, nt_for_m(mm), group_size != -1, pick_geometry(mm, (int)n, (int)k));
}

int b200_w4a16_max_clusters(int config, int64_t m, int group_size, int dtype) {
  GemmCall g;
  std::memset(&g, 0, sizeof(g));
  g.m = static_cast<int>(m);
  g.n = 4096;
  g.k = 4096;
  g.group_size = group_size;
  g.dtype = dtype;
  if (config >= kStreamBase && config < kStreamBase + kStreamGeoms * 8)
    return stream_max_clusters(g, (config - kStreamBase) >> 3, 1 + ((config - kStreamBase) & 7));
  if (config >= kPrefillBase && config < kPrefillBase + kPrefillIds) {
    int tile, S;
    prefill_from_id(config, &tile, &S);
    return prefill_max_clusters(g, tile, S);
  }
  return -1;
}

const char* b200_w4a16_config_name(int config) {
  const int count = static_cast<int>(sizeof(g_decode) / sizeof(g_decode[0]));
  if (config >= kPrefillBase && config < kPrefillBase + kPrefillIds) {
    int tile, S;
    prefill_from_id(config, &tile, &S);
    return prefill_tile_name(tile);
  }
  if (config >= kStreamBase && config < kStreamBase + kStreamGeoms * 8) return stream_geom_name((config - kStreamBase) >> 3);
  if (config < 1 || config > count) return "";
  return g_decode[config - 1].name;
}

int b200_w4a16_gemm_ex(const void* a, const int32_t* qweight, const void* scales, const int32_t* g_idx,
                       const int32_t* perm, int32_t* workspace, void* c, int64_t m, int64_t n, int64_t k,
                       int group_size, int dtype, int is_k_full, int has_act_order, void* a_tmp,
                       float* c_tmp, void* stream, const b200_w4a16_opts* opts) {
  (void)g_idx;
  // ---- validation: the reference's checks (gptq_marlin.cu:71-72, gptq_marlin_mm.cu:53-54,84-131) ----
  if (!(m > 0 && n > 0 && k > 0))
    return fail(B200_W4A16_EINVAL, "Invalid MNK = [%lld, %lld, %lld]", (long long)m, (long long)n, (long long)k);
  if (n % 64 != 0)
    return fail(B200_W4A16_EINVAL, "size_n = %lld, is not divisible by min_thread_n = 64", (long long)n);
  if (k % 64 != 0)
    return fail(B200_W4A16_EINVAL, "size_k = %lld, is not divisible by min_thread_k = 64", (long long)k);
  if (dtype != B200_W4A16_FP16 && dtype != B200_W4A16_BF16)
    return fail(B200_W4A16_EINVAL, "dtype must be 0 (fp16) or 1 (bf16), got %d", dtype);
  if (group_size != 128 && group_size != -1)
    return fail(B200_W4A16_EINVAL, "Unsupported group size %d (128 or -1)", group_size);
  if (group_size == 128 && k % 128 != 0)
    return fail(B200_W4A16_EINVAL, "prob_k = %lld is not divisible by group_size = 128", (long long)k);
  if (m > INT32_MAX || n > INT32_MAX || k > INT32_MAX || m * n > (1ll << 40))
    return fail(B200_W4A16_EINVAL, "problem too large: MNK = [%lld, %lld, %lld]", (long long)m, (long long)n, (long long)k);
  if (!a || !qweight || !c || !workspace || (group_size == 128 && !scales) || (group_size == -1 && !scales))
    return fail(B200_W4A16_EINVAL, "null pointer argument (a/qweight/scales/workspace/c)");
  if (has_act_order) {
    if (!is_k_full)
      return fail(B200_W4A16_EINVAL,
                  "has_act_order with is_k_full = false (g_idx-indexed scales) is not supported; "
                  "CPM.cu never takes that path (w4a16_gptq_marlin_linear.cuh:55,66) and its converter "
                  "rejects desc_act checkpoints (gptq2marlin.py:91)");
    if (group_size == -1)
      return fail(B200_W4A16_EINVAL, "group_size must not be -1 when has_act_order=true and is_k_full=true");
    if (!perm || !a_tmp) return fail(B200_W4A16_EINVAL, "has_act_order needs perm and a_tmp");
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t esz = 2;
  if (has_act_order) {
    const long long total = m * k;
    const int blocks = static_cast<int>(total / 256 + 1 < 4096 ? total / 256 + 1 : 4096);
    if (dtype == 0)
      gather_columns_kernel<__half><<<blocks, 256, 0, st>>>((const __half*)a, perm, (__half*)a_tmp, (int)m, (int)k);
    else
      gather_columns_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)a, perm, (__nv_bfloat16*)a_tmp, (int)m, (int)k);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    a = a_tmp;
  }

  const bool grouped = group_size != -1;
  const int count = static_cast<int>(sizeof(g_decode) / sizeof(g_decode[0]));
  const int forced = opts ? opts->force_config : 0;
  const bool gated = opts && opts->gated_silu != 0;
  if (gated) {
    // fused SiLU(gate) * up: served by the persistent streaming kernel (decode / tree-verify widths)
    if (m > 32 || forced != 0)
      return fail(B200_W4A16_EINVAL, "gated_silu is fused for M <= 32 with the default configuration (got M = %lld, force_config %d)", (long long)m, forced);
    if (group_size != 128 || (n % 256) != 0)
      return fail(B200_W4A16_EINVAL, "gated_silu needs group_size 128 and N (= 2 x intermediate) a multiple of 256, got N = %lld", (long long)n);
    if (opts->bias || opts->residual || opts->out_scale != 0.0f)
      return fail(B200_W4A16_EINVAL, "gated_silu does not combine with bias / residual / out_scale");
    if (has_act_order) return fail(B200_W4A16_EINVAL, "gated_silu does not combine with act-order");
  }
  // ---- tensor-bound regime: one persistent tcgen05 launch for any M ----
  if ((forced == 0 && m > kStreamMaxM) || (forced >= kPrefillBase && forced < kStreamBase)) {
    if (forced >= kPrefillBase + kPrefillIds)
      return fail(B200_W4A16_EINVAL, "force_config %d out of range", forced);
    GemmCall g;
    g.a = a;
    g.qweight = qweight;
    g.scales = scales;
    g.workspace = workspace;
    g.c = c;
    g.c_tmp = c_tmp;
    g.bias = opts ? opts->bias : nullptr;
    g.residual = opts ? opts->residual : nullptr;
    g.out_scale = opts ? opts->out_scale : 0.0f;
    g.gated = false;
    g.m = static_cast<int>(m);
    g.n = static_cast<int>(n);
    g.k = static_cast<int>(k);
    g.group_size = group_size;
    g.dtype = dtype;
    g.pdl = opts ? (opts->pdl != 0) : (g_default_pdl.load(std::memory_order_relaxed) != 0);
    g.stream = st;
    int tile, S;
    if (forced) prefill_from_id(forced, &tile, &S);
    else prefill_pick(g, &tile, &S);
    return launch_prefill(g, tile, S);
  }
  // ---- streaming regime: needs the fp32 split-K scratch ----
  {
    const int64_t need = b200_w4a16_reduce_bytes(m < kStreamMaxM ? m : kStreamMaxM, n, k);
    if (!c_tmp) return fail(B200_W4A16_EINVAL, "c_tmp (fp32 reduce scratch) is required");
    if (opts && opts->c_tmp_bytes > 0 && opts->c_tmp_bytes < need)
      return fail(B200_W4A16_EINVAL, "c_tmp too small: %lld bytes given, %lld needed for MNK = [%lld, %lld, %lld]",
                  (long long)opts->c_tmp_bytes, (long long)need, (long long)m, (long long)n, (long long)k);
  }
  // ---- M loop: slabs of <= 32 tokens through the streaming kernel ----
  for (int64_t m0 = 0; m0 < m; m0 += 32) {
    const int mm = static_cast<int>(m - m0 < 32 ? m - m0 : 32);
    int idx = -1;
    if (forced >= kStreamBase || (forced == 0 && !gated && use_cluster_streamer(mm, n, k))) {
      // cluster split-K streaming kernel
    } else if (opts && opts->force_config > 0) {
      idx = opts->force_config - 1;
      if (idx >= count) return fail(B200_W4A16_EINVAL, "force_config %d out of range", opts->force_config);
      const DecodeConfig& f = g_decode[idx];
      if (f.dtype != dtype || f.grouped != grouped || f.nt * 8 < mm || k % (f.krows * 16) != 0)
        return fail(B200_W4A16_EINVAL, "force_config %d (%s) does not fit m=%d k=%lld dtype=%d group=%d",
                    opts->force_config, f.name, mm, (long long)k, dtype, group_size);
    } else {
      idx = decode_index(dtype, nt_for_m(mm), grouped, pick_geometry(mm, (int)n, (int)k));
    }
    GemmCall g;
    g.a = static_cast<const char*>(a) + static_cast<size_t>(m0) * k * esz;
    g.qweight = qweight;
    g.scales = scales;
    g.workspace = workspace;
    g.c = static_cast<char*>(c) + static_cast<size_t>(m0) * n * esz;
    g.c_tmp = c_tmp;
    g.bias = opts ? opts->bias : nullptr;
    g.residual = (opts && opts->residual)
                     ? static_cast<const char*>(opts->residual) + static_cast<size_t>(m0) * n * esz
                     : nullptr;
    g.out_scale = opts ? opts->out_scale : 0.0f;
    g.gated = gated;
    g.m = mm;
    g.n = static_cast<int>(n);
    g.k = static_cast<int>(k);
    g.group_size = group_size;
    g.dtype = dtype;
    g.pdl = opts ? (opts->pdl != 0) : (g_default_pdl.load(std::memory_order_relaxed) != 0);
    g.stream = st;
    if (idx < 0) {
      int geom, S;
      if (forced >= kStreamBase) {
        geom = (forced - kStreamBase) >> 3;
        S = 1 + ((forced - kStreamBase) & 7);
        if (geom >= kStreamGeoms) return fail(B200_W4A16_EINVAL, "force_config %d out of range", forced);
      } else {
        stream_pick(g, &geom, &S);
      }
      if (int rc = launch_stream(g, geom, S)) return rc;
      continue;
    }
    if (gated) {
      if (int rc = launch_gated(g, nt_for_m(mm), pick_geometry(mm, (int)n, (int)k))) return rc;
      continue;
    }
    if (int rc = g_decode[idx].launch(g)) return rc;
  }
  return B200_W4A16_OK;
}

int b200_w4a16_gemm(const void* a, const int32_t* qweight, const void* scales, const int32_t* g_idx,
                    const int32_t* perm, int32_t* workspace, void* c, int64_t m, int64_t n, int64_t k,
                    int group_size, int dtype, int is_k_full, int has_act_order, void* a_tmp, float* c_tmp,
                    void* stream) {
  return b200_w4a16_gemm_ex(a, qweight, scales, g_idx, perm, workspace, c, m, n, k, group_size, dtype,
                            is_k_full, has_act_order, a_tmp, c_tmp, stream, nullptr);
}

int b200_w4a16_linear_forward(const void* input, const int32_t* qweight, const void* scales, const void* bias,
                              int32_t* workspace, void* output, void* tgt, int inplace, int64_t num_tokens,
                              int64_t dim_out, int64_t dim_in, int group_size, int dtype, void* a_tmp,
                              float* c_tmp, void* stream) {
  // W4A16GPTQMarlinLinear::prefill, w4a16_gptq_marlin_linear.cuh:104-146
  b200_w4a16_opts o;
  std::memset(&o, 0, sizeof(o));
  o.bias = bias;
  o.pdl = g_default_pdl.load(std::memory_order_relaxed);
  void* dst = output;
  if (tgt && inplace) {
    o.residual = tgt;
    dst = tgt;
  } else if (tgt) {
    dst = tgt;
  }
  if (!dst) return fail(B200_W4A16_EINVAL, "linear_forward: no output buffer");
  return b200_w4a16_gemm_ex(input, qweight, scales, nullptr, nullptr, workspace, dst, num_tokens, dim_out,
                            dim_in, group_size, dtype, /*is_k_full=*/1, /*has_act_order=*/0, a_tmp, c_tmp,
                            stream, &o);
}

int b200_w4a16_dequant(const int32_t* qweight, const void* scales, void* w_out, int64_t n, int64_t k,
                       int group_size, int dtype, void* stream) {
  if (!(n > 0 && k > 0) || n % 64 || k % 16)
    return fail(B200_W4A16_EINVAL, "dequant: need n %% 64 == 0 and k %% 16 == 0, got n=%lld k=%lld", (long long)n, (long long)k);
  if (group_size != 128 && group_size != -1) return fail(B200_W4A16_EINVAL, "Unsupported group size %d", group_size);
  if (group_size == 128 && k % 128) return fail(B200_W4A16_EINVAL, "dequant: k %% 128 != 0");
  if (dtype != 0 && dtype != 1) return fail(B200_W4A16_EINVAL, "dtype must be 0 or 1");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long blocks = (k / 16) * (n / 64);
  const int grid = static_cast<int>((blocks + 7) / 8 < 148 * 8 ? (blocks + 7) / 8 : 148 * 8);
  const uint4* q = reinterpret_cast<const uint4*>(qweight);
  if (dtype == 0) {
    if (group_size == -1)
      w4a16_dequant_kernel<__half, false><<<grid, 256, 0, st>>>(q, (const __half*)scales, (__half*)w_out, (int)n, (int)k, 1);
    else
      w4a16_dequant_kernel<__half, true><<<grid, 256, 0, st>>>(q, (const __half*)scales, (__half*)w_out, (int)n, (int)k, group_size);
  } else {
    if (group_size == -1)
      w4a16_dequant_kernel<__nv_bfloat16, false><<<grid, 256, 0, st>>>(q, (const __nv_bfloat16*)scales, (__nv_bfloat16*)w_out, (int)n, (int)k, 1);
    else
      w4a16_dequant_kernel<__nv_bfloat16, true><<<grid, 256, 0, st>>>(q, (const __nv_bfloat16*)scales, (__nv_bfloat16*)w_out, (int)n, (int)k, group_size);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B200_W4A16_ECUDA, "dequant launch: %s", cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

int b200_w4a16_repack_gptq(const int32_t* qweight_gptq, int32_t* qweight_marlin, int64_t n, int64_t k, void* stream) {
  if (!(n > 0 && k > 0) || n % 64 || k % 16)
    return fail(B200_W4A16_EINVAL, "repack: need n %% 64 == 0 and k %% 16 == 0, got n=%lld k=%lld", (long long)n, (long long)k);
  const long long words = (k / 16) * 2 * n;
  const int grid = static_cast<int>((words + 255) / 256 < 148 * 16 ? (words + 255) / 256 : 148 * 16);
  repack_gptq_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const uint32_t*>(qweight_gptq), reinterpret_cast<uint32_t*>(qweight_marlin), (int)n, (int)k);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B200_W4A16_ECUDA, "repack launch: %s", cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

int b200_w4a16_allreduce(const void* const* peer_in, void* const* peer_res, int32_t* const* peer_sig,
                         int32_t* epoch, void* out, int64_t elems, int dtype, int rank, int world, int two_shot,
                         void* stream) {
  if (world < 1 || world > kArMaxWorld || rank < 0 || rank >= world)
    return fail(B200_W4A16_EINVAL, "allreduce: world must be 1..%d and 0 <= rank < world (got %d, %d)", kArMaxWorld, world, rank);
  if (elems <= 0 || elems % 8) return fail(B200_W4A16_EINVAL, "allreduce: elems must be a positive multiple of 8, got %lld", (long long)elems);
  if (dtype != 0 && dtype != 1) return fail(B200_W4A16_EINVAL, "dtype must be 0 or 1");
  if (!peer_in || !peer_sig || !epoch || !out || (two_shot && !peer_res))
    return fail(B200_W4A16_EINVAL, "allreduce: null pointer argument");
  AllReduceArgs a;
  std::memset(&a, 0, sizeof(a));
  for (int w = 0; w < world; ++w) {
    a.in[w] = peer_in[w];
    a.sig[w] = peer_sig[w];
    a.res[w] = two_shot ? peer_res[w] : nullptr;
    if (!a.in[w] || !a.sig[w] || (two_shot && !a.res[w]))
      return fail(B200_W4A16_EINVAL, "allreduce: null peer pointer for rank %d", w);
  }
  a.epoch = epoch;
  a.out = out;
  a.elems = elems;
  a.rank = rank;
  a.world = world;
  // one-shot: blocks over the whole message; two-shot: blocks over one rank slice (block b of every
  // rank owns the same part of every slice)
  const long long span = two_shot ? ((elems / 8 + world - 1) / world) : elems / 8;
  long long blocks = (span + 255) / 256;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  if (blocks < 1) blocks = 1;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int nb = static_cast<int>(blocks);
  cudaLaunchConfig_t cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(nb);
  cfg.blockDim = dim3(256);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  // PDL on the collective is opt-in (B200_W4A16_AR_PDL=1): measured on 2 B200s it helps the tcgen05
  // regime (M = 64: 2.55 -> 2.40 ms) but lets three kernels crowd each SM in the streaming regime
  // (M = 1: 1.39 -> 1.89 ms), profiles/r01/README.md
  static const bool ar_pdl = [] { const char* e = std::getenv("B200_W4A16_AR_PDL"); return e && e[0] == '1'; }();
  cfg.numAttrs = (ar_pdl && g_default_pdl.load(std::memory_order_relaxed)) ? 1 : 0;
  cudaError_t le;
  if (dtype == 0) {
    le = two_shot ? cudaLaunchKernelEx(&cfg, allreduce_kernel<__half, true>, a)
                  : cudaLaunchKernelEx(&cfg, allreduce_kernel<__half, false>, a);
  } else {
    le = two_shot ? cudaLaunchKernelEx(&cfg, allreduce_kernel<__nv_bfloat16, true>, a)
                  : cudaLaunchKernelEx(&cfg, allreduce_kernel<__nv_bfloat16, false>, a);
  }
  if (le != cudaSuccess) return fail(B200_W4A16_ECUDA, "allreduce launch: %s", cudaGetErrorString(le));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B200_W4A16_ECUDA, "allreduce launch: %s", cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

int b200_w4a16_allreduce_oneshot(const void* const* peer_in, int32_t* const* peer_sig, int32_t* epoch, void* out,
                                 int64_t elems, int dtype, int rank, int world, void* stream) {
  return b200_w4a16_allreduce(peer_in, nullptr, peer_sig, epoch, out, elems, dtype, rank, world, 0, stream);
}

}  // extern "C"
