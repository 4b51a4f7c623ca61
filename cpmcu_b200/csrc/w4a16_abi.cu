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
#include "host_common.h"

namespace b200 {

thread_local std::string g_last_error;
std::atomic<long long> g_launches{0};
std::atomic<int> g_default_pdl{[] {
  const char* e = std::getenv("B200_W4A16_PDL");
  return (e && e[0] == '1') ? 1 : 0;
}()};

// Scratch capacity the caller guarantees for every c_tmp passed WITHOUT a size (b200_w4a16_gemm,
// b200_w4a16_linear_forward, the gptq_marlin_gemm shim): 0 = only the reference contract,
// determine_reduce_max_m(M, 16) x N floats (gptq_marlin_utils.cu:206-225).  b200_w4a16_set_scratch_capacity.
std::atomic<long long> g_scratch_capacity{0};

// tuning aid: per-launch timestamp blocks (b200_w4a16_debug_timeline)
unsigned long long* g_dbg_ts = nullptr;
long long g_dbg_ts_launches = 0;
std::atomic<long long> g_dbg_ts_next{0};
unsigned long long* debug_next_slot() {
  if (!g_dbg_ts) return nullptr;
  const long long idx = g_dbg_ts_next.fetch_add(1, std::memory_order_relaxed);
  return idx < g_dbg_ts_launches ? g_dbg_ts + idx * 8192 : nullptr;
}

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
// streaming (decode) regime: launch plan
// ------------------------------------------------------------------------------------------
int nt_for_m(int m) { return m <= 8 ? 1 : (m <= 16 ? 2 : 4); }

// largest M served by the streaming kernel; above it the tcgen05 kernel runs (split-K clusters up to
// M = 256, persistent whole-K tiles beyond).  Measured crossover on B200: profiles/r01 sweep_mid.
int kStreamMaxM = [] {
  const char* e = std::getenv("B200_W4A16_STREAM_MAX_M");
  const int v = e ? std::atoi(e) : 32;
  return v < 0 ? 0 : v;
}();

int reduce_max_m(long long m) {
  // mirror of marlin::determine_reduce_max_m(prob_m, max_par = 16), gptq_marlin_utils.cu:206-225
  if (m <= 16) return 16;
  if (m <= 32) return 32;
  if (m <= 48) return 48;
  if (m <= 64) return 64;
  long long par = (m + 63) / 64;
  if (par > 16) par = 16;
  return static_cast<int>(64 * par);
}

bool geom_fits(int geo, int n, int k, bool gated) {
  const int tn = kGeom[geo].chunks * 64, kb = kGeom[geo].krows * 16;
  if (k % kb) return false;
  if (gated && (geo > 1 || (n / 2) % (tn / 2) != 0)) return false;
  return true;
}

// One slab of mm <= 32 tokens.  `capacity` = bytes of c_tmp the caller guarantees.  Stream-K (every SM
// streams the same number of bytes; split columns are reduced through fp32 slots in c_tmp) whenever the
// slots fit; otherwise fewer CTAs (fewer sharers per column) or whole columns per CTA (no scratch at all),
// whichever leaves the shorter critical path.  geom < 0: the default geometry for the shape.
DecodePlan plan_decode(int mm, int n, int k, int geom, long long capacity, bool gated, bool tmem_ok = true) {
  const int sms = sm_count();
  auto dims = [&](int geo, DecodePlan* pl) {
    const int tn = kGeom[geo].chunks * 64, kb = kGeom[geo].krows * 16;
    pl->geom = geo;
    pl->ncols = gated ? (n / 2) / (tn / 2) : (n + tn - 1) / tn;
    pl->kblks = k / kb;
  };
  // critical path in bytes of int4 stream: the busiest CTA's units
  auto path_bytes = [&](const DecodePlan& pl) {
    const long long unit_bytes = static_cast<long long>(kGeom[pl.geom].chunks) * kGeom[pl.geom].krows * 512;
    const long long units = static_cast<long long>(pl.ncols) * pl.kblks;
    const long long per = pl.colpar ? static_cast<long long>((pl.ncols + pl.grid - 1) / pl.grid) * pl.kblks
                                    : (units + pl.grid - 1) / pl.grid;
    return per * unit_bytes;
  };
  auto full_grid = [&](const DecodePlan& pl) {
    const long long units = static_cast<long long>(pl.ncols) * pl.kblks;
    return static_cast<int>(units < sms ? units : sms);
  };
  auto stream_k = [&](int geo, DecodePlan* pl) {   // false: not even two slots per column fit
    dims(geo, pl);
    const int tn = kGeom[geo].chunks * 64;
    const long long units = static_cast<long long>(pl->ncols) * pl->kblks;
    const long long slot_bytes = static_cast<long long>(pl->ncols) * mm * tn * 4;
    const long long cap = capacity / slot_bytes;
    pl->colpar = 0;
    pl->grid = full_grid(*pl);
    if (pl->kblks == 1) {  // a column is one unit: never shared
      pl->slot_max = 1;
      pl->need_bytes = 0;
      return true;
    }
    if (cap < 2) return false;
    while (pl->grid > 1 && decode_slot_max(pl->kblks, units, pl->grid) > cap) --pl->grid;
    pl->slot_max = decode_slot_max(pl->kblks, units, pl->grid);
    if (pl->slot_max > cap) return false;
    pl->need_bytes = slot_bytes * pl->slot_max;
    return true;
  };
  auto whole_cols = [&](int geo, DecodePlan* pl) {
    dims(geo, pl);
    pl->colpar = 1;
    pl->grid = pl->ncols < sms ? pl->ncols : sms;
    pl->slot_max = 1;
    pl->need_bytes = 0;
  };
  // default geometry, measured on B200 (profiles/r02/c_decode_geometry_chains.txt, h_decode_m16_mixing.txt): the
  // 256-column x 128-k stage has the fewest instructions per byte (M <= 8).  For M > 8 a 256-column tile of a small-N
  // shape (qkv, o, down) is shared by 8+ CTAs and the last arriver of a column pulls 8 x M x 1 KB: the 128-column x 256-k
  // stage halves both.  ONE geometry per token-block width for every shape: alternating between two kernel functions
  // inside a layer costs more than the better per-shape choice gains (M = 16: 74 us per layer mixed, 67 us uniform).
  // For M > 16 the same 128-column x 256-k tile runs on tcgen05 with the A operand in tensor memory (geometry 8,
  // decode_tc_kernel.cuh) unless the launch needs the fused all-reduce: mma.sync costs 16-17 cycles per word per 8 tokens,
  // the tensor-memory route does not depend on M (MiniCPM4-8B layer at M = 32: 82 vs 95 us; at M = 16 it is 2 % ahead
  // on the 8B shapes and 6 % behind on the Llama-70B ones, so M <= 16 stays on mma.sync; profiles/r02/n_tmem_*).
  int g0 = (k % 128) == 0 ? 0 : 1;
  if (geom < 0 && !gated && mm > 8 && (k % 256) == 0) {
    static const int tmem_min_m = [] { const char* e = std::getenv("B200_W4A16_TMEM_MIN_M"); return e ? std::atoi(e) : 17; }();
    // ONE tile for every shape of a layer: alternating the 256- and the 128-column tensor-memory tiles per shape is 4 %
    // ahead at M = 16 in an 8-layer chain but not in the 32-layer bench graph, and far behind at M = 24 / 32
    // (100 / 113 us against 82 uniform; profiles/r02/n_tmem_*).  B200_W4A16_TMEM_MIN_M moves the threshold (tuning).
    g0 = (tmem_ok && mm >= tmem_min_m) ? kTcGeom2 : 5;
  }
  if (geom >= 0) g0 = geom;
  DecodePlan best;
  bool have = stream_k(g0, &best);
  if (have && best.grid == full_grid(best)) return best;
  // scratch too small for full-width stream-K: compare the shrunken grid with whole-column launches
  DecodePlan alt;
  whole_cols(g0, &alt);
  if (!have || path_bytes(alt) < path_bytes(best)) best = alt;
  if (geom < 0 && !gated && geom_fits(kWholeColGeom, n, k, gated)) {
    whole_cols(kWholeColGeom, &alt);
    if (path_bytes(alt) < path_bytes(best)) best = alt;
  }
  return best;
}

// configuration ids (force_config / pick_config): 1 + ((grouped ? 0 : 6) + dtype * 3 + nti) * kGeoms + geom for the
// streaming kernel, kPrefillBase.. for the tcgen05 tile shapes
constexpr int kDecodeIds = 12 * kGeoms;
int decode_id(int dtype, int nt, bool grouped, int geom) {
  const int nti = nt == 1 ? 0 : (nt == 2 ? 1 : 2);
  return 1 + ((grouped ? 0 : 6) + dtype * 3 + nti) * kGeoms + geom;
}
void decode_from_id(int id, int* dtype, int* nt, bool* grouped, int* geom) {
  const int r = id - 1;
  *geom = r % kGeoms;
  const int fam = r / kGeoms;
  *grouped = fam < 6;
  *dtype = (fam % 6) / 3;
  *nt = 1 << (fam % 3);
}
const char* decode_name(int id) {
  static thread_local char buf[128];
  int dtype, nt, geom;
  bool grouped;
  decode_from_id(id, &dtype, &nt, &grouped, &geom);
  snprintf(buf, sizeof(buf), "%s/nt%d/tn%dx64/kb%dx16/w%d/s%d/%s%s", dtype == 0 ? "__half" : "__nv_bfloat16", nt,
           kGeom[geom].chunks, kGeom[geom].krows, kGeom[geom].warps, kGeom[geom].stages, grouped ? "true" : "false",
           geom >= kTcGeom ? "/tmem" : "");
  return buf;
}

constexpr int kPrefillBase = 1001;
// ids: kPrefillBase + t for the whole-K tiles t < 4; then 4 ids (cluster size 1, 2, 4, 8) for each of the two split-K
// tiles; then the CTA-pair tile
constexpr int kPrefillSplitTiles = 2;
constexpr int kPrefillIds = kPrefillWholeK + kPrefillSplitTiles * 4 + 1;
int prefill_id(int tile, int S) {
  if (tile < kPrefillWholeK) return kPrefillBase + tile;
  if (tile == kPrefillPairTile) return kPrefillBase + kPrefillWholeK + kPrefillSplitTiles * 4;
  int l2 = 0;
  while ((1 << l2) < S) ++l2;
  return kPrefillBase + kPrefillWholeK + (tile - kPrefillWholeK) * 4 + l2;
}
void prefill_from_id(int id, int* tile, int* S) {
  const int r = id - kPrefillBase;
  if (r < kPrefillWholeK) {
    *tile = r;
    *S = 1;
  } else if (r == kPrefillWholeK + kPrefillSplitTiles * 4) {
    *tile = kPrefillPairTile;
    *S = 1;
  } else {
    *tile = kPrefillWholeK + (r - kPrefillWholeK) / 4;
    *S = 1 << ((r - kPrefillWholeK) & 3);
  }
}

}  // namespace b200

using namespace b200;

// ------------------------------------------------------------------------------------------
// exported functions
// ------------------------------------------------------------------------------------------
extern "C" {

const char* b200_w4a16_last_error(void) { return g_last_error.c_str(); }
const char* b200_w4a16_version(void) { return "b200-w4a16 0.2 (sm_100a)"; }
int64_t b200_w4a16_launch_count(void) { return g_launches.load(); }
void b200_w4a16_debug_timeline(void* device_buf, int64_t launches) {
  g_dbg_ts = static_cast<unsigned long long*>(device_buf);
  g_dbg_ts_launches = device_buf ? launches : 0;
  g_dbg_ts_next.store(0);
}
int b200_w4a16_set_pdl(int enable) { return g_default_pdl.exchange(enable ? 1 : 0); }
int64_t b200_w4a16_set_scratch_capacity(int64_t bytes) { return g_scratch_capacity.exchange(bytes < 0 ? 0 : bytes); }

int b200_w4a16_reduce_max_m(int64_t m) { return reduce_max_m(m); }

int64_t b200_w4a16_workspace_ints(int64_t n) { return (n / 64) * 16; }

// what the library touches when the caller guarantees only the reference contract
int64_t b200_w4a16_reduce_bytes(int64_t m, int64_t n, int64_t k) {
  if (m <= 0 || n <= 0 || k <= 0 || n % 64 || k % 64) return -1;
  if (m > kStreamMaxM) return 0;  // the tcgen05 kernel accumulates in tensor memory: no scratch
  const long long contract = static_cast<long long>(reduce_max_m(m)) * n * 4;
  return plan_decode(static_cast<int>(m), static_cast<int>(n), static_cast<int>(k), -1, contract, false).need_bytes;
}

// what full-width stream-K needs (worst case over the geometries force_config may select)
int64_t b200_w4a16_reduce_bytes_fast(int64_t m, int64_t n, int64_t k) {
  if (m <= 0 || n <= 0 || k <= 0 || n % 64 || k % 64) return -1;
  const int mm = static_cast<int>(m < 32 ? m : 32);
  int64_t worst = 0;
  for (int geo = 0; geo < kGeoms; ++geo) {
    if (!geom_fits(geo, static_cast<int>(n), static_cast<int>(k), false)) continue;
    const DecodePlan pl = plan_decode(mm, static_cast<int>(n), static_cast<int>(k), geo, INT64_MAX / 4, false);
    if (pl.need_bytes > worst) worst = pl.need_bytes;
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
  const int mm = static_cast<int>(m);
  const DecodePlan pl = plan_decode(mm, static_cast<int>(n), static_cast<int>(k), -1, INT64_MAX / 4, false);
  return decode_id(dtype, nt_for_m(mm), group_size != -1, pl.geom);
}

int b200_w4a16_max_clusters(int config, int64_t m, int group_size, int dtype) {
  GemmCall g;
  std::memset(&g, 0, sizeof(g));
  g.m = static_cast<int>(m);
  g.n = 4096;
  g.k = 4096;
  g.group_size = group_size;
  g.dtype = dtype;
  if (config >= kPrefillBase && config < kPrefillBase + kPrefillIds) {
    int tile, S;
    prefill_from_id(config, &tile, &S);
    return prefill_max_clusters(g, tile, S);
  }
  return -1;
}

const char* b200_w4a16_config_name(int config) {
  if (config >= kPrefillBase && config < kPrefillBase + kPrefillIds) {
    int tile, S;
    prefill_from_id(config, &tile, &S);
    return prefill_tile_name(tile);
  }
  if (config < 1 || config > kDecodeIds) return "";
  return decode_name(config);
}

static int gemm_impl(const void* a, const int32_t* qweight, const void* scales, const int32_t* g_idx,
                     const int32_t* perm, int32_t* workspace, void* c, int64_t m, int64_t n, int64_t k,
                     int group_size, int dtype, int is_k_full, int has_act_order, void* a_tmp,
                     float* c_tmp, void* stream, const b200_w4a16_opts* opts, const b200_w4a16_ar* ar) {
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
  if (!a || !qweight || !c || !workspace || !scales)
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
  const int forced = opts ? opts->force_config : 0;
  const bool gated = opts && opts->gated_silu != 0;
  if (gated) {
    // fused SiLU(gate) * up: served by the persistent streaming kernel (decode / tree-verify widths)
    if (m > kStreamMaxM || forced >= kPrefillBase)
      return fail(B200_W4A16_EINVAL, "gated_silu is fused for M <= %d (got M = %lld, force_config %d)", kStreamMaxM, (long long)m, forced);
    if (group_size != 128 || (n % 256) != 0)
      return fail(B200_W4A16_EINVAL, "gated_silu needs group_size 128 and N (= 2 x intermediate) a multiple of 256, got N = %lld", (long long)n);
    if (opts->bias || opts->residual || opts->out_scale != 0.0f)
      return fail(B200_W4A16_EINVAL, "gated_silu does not combine with bias / residual / out_scale");
    if (has_act_order) return fail(B200_W4A16_EINVAL, "gated_silu does not combine with act-order");
  }
  const int qa = opts ? opts->qkv_q_dim : 0, qb = opts ? opts->qkv_kv_dim : 0;
  if (qa || qb) {
    if (qa <= 0 || qb <= 0 || qa % 8 || qb % 8 || qa + 2ll * qb != n)
      return fail(B200_W4A16_EINVAL, "qkv split: need qkv_q_dim + 2 * qkv_kv_dim == N and both multiples of 8 (got %d, %d, N = %lld)", qa, qb, (long long)n);
    if (gated || (opts && opts->residual) || (forced > 0 && forced < kPrefillBase && m > 32))
      return fail(B200_W4A16_EINVAL, "qkv split does not combine with gated_silu / residual / a multi-slab forced launch");
  }
  GemmCall g;
  g.qkv_a = qa;
  g.qkv_b = qb;
  g.a = a;
  g.qweight = qweight;
  g.scales = scales;
  g.workspace = workspace;
  g.c = c;
  g.c_tmp = c_tmp;
  g.bias = opts ? opts->bias : nullptr;
  g.residual = opts ? opts->residual : nullptr;
  g.out_scale = opts ? opts->out_scale : 0.0f;
  g.gated = gated;
  g.m = static_cast<int>(m);
  g.n = static_cast<int>(n);
  g.k = static_cast<int>(k);
  g.group_size = group_size;
  g.dtype = dtype;
  g.pdl = opts ? (opts->pdl != 0) : (g_default_pdl.load(std::memory_order_relaxed) != 0);
  g.stream = st;
  std::memset(g.ar_recv, 0, sizeof(g.ar_recv));
  std::memset(g.ar_flag, 0, sizeof(g.ar_flag));
  g.ar_seen = nullptr;
  g.ar_slot_elems = 0;
  g.ar_world = 1;
  g.ar_rank = 0;
  if (ar && ar->world > 1) {
    if (ar->world > 8 || ar->rank < 0 || ar->rank >= ar->world)
      return fail(B200_W4A16_EINVAL, "gemm_ar: world must be 2..8 and 0 <= rank < world (got %d, %d)", ar->world, ar->rank);
    if (m > kStreamMaxM || forced >= kPrefillBase || gated || has_act_order)
      return fail(B200_W4A16_EINVAL, "gemm_ar: the fused all-reduce serves M <= %d without gated_silu / act-order (got M = %lld)", kStreamMaxM, (long long)m);
    if (!ar->recv || !ar->flag || !ar->seen || ar->slot_elems < m * n || ar->slot_elems % 8)
      return fail(B200_W4A16_EINVAL, "gemm_ar: null buffers or slot_elems (%lld) smaller than m * n", (long long)ar->slot_elems);
    for (int w = 0; w < ar->world; ++w) {
      if (!ar->recv[w] || !ar->flag[w]) return fail(B200_W4A16_EINVAL, "gemm_ar: null peer pointer for rank %d", w);
      g.ar_recv[w] = ar->recv[w];
      g.ar_flag[w] = ar->flag[w];
    }
    g.ar_seen = ar->seen;
    g.ar_slot_elems = ar->slot_elems;
    g.ar_world = ar->world;
    g.ar_rank = ar->rank;
  }
  // ---- tensor-bound regime: one persistent tcgen05 launch for any M ----
  if ((forced == 0 && m > kStreamMaxM) || forced >= kPrefillBase) {
    if (forced >= kPrefillBase + kPrefillIds)
      return fail(B200_W4A16_EINVAL, "force_config %d out of range", forced);
    int tile, S;
    if (forced) prefill_from_id(forced, &tile, &S);
    else prefill_pick(g, &tile, &S);
    return launch_prefill(g, tile, S);
  }
  // ---- streaming regime ----
  int geom = -1;
  if (forced > 0) {
    if (forced > kDecodeIds) return fail(B200_W4A16_EINVAL, "force_config %d out of range", forced);
    int fdt, fnt;
    bool fgr;
    decode_from_id(forced, &fdt, &fnt, &fgr, &geom);
    const int mm0 = static_cast<int>(m < 32 ? m : 32);
    if (fdt != dtype || fgr != grouped || fnt * 8 < mm0 || !geom_fits(geom, g.n, g.k, gated))
      return fail(B200_W4A16_EINVAL, "force_config %d (%s) does not fit m=%d k=%lld dtype=%d group=%d",
                  forced, decode_name(forced), mm0, (long long)k, dtype, group_size);
  }
  // scratch the caller guarantees: stated per call, or process-wide, or the reference's contract
  // (determine_reduce_max_m(M, 16) x N floats; callers: w4a16_gptq_marlin_layer.cuh:34-37)
  long long capacity = opts && opts->c_tmp_bytes > 0 ? opts->c_tmp_bytes : g_scratch_capacity.load(std::memory_order_relaxed);
  if (capacity <= 0) capacity = static_cast<long long>(reduce_max_m(m)) * n * 4;
  if (!c_tmp) capacity = 0;  // no scratch at all: whole-column launches only
  // ---- M loop: slabs of <= 32 tokens through the streaming kernel ----
  for (int64_t m0 = 0; m0 < m; m0 += 32) {
    const int mm = static_cast<int>(m - m0 < 32 ? m - m0 : 32);
    GemmCall s = g;
    s.a = static_cast<const char*>(a) + static_cast<size_t>(m0) * k * esz;
    s.c = static_cast<char*>(c) + static_cast<size_t>(m0) * (gated ? n / 2 : n) * esz;
    s.residual = g.residual ? static_cast<const char*>(g.residual) + static_cast<size_t>(m0) * n * esz : nullptr;
    s.m = mm;
    const DecodePlan pl = plan_decode(mm, s.n, s.k, geom, capacity, gated, /*tmem_ok=*/g.ar_world <= 1);
    const int nt = forced > 0 ? (1 << (((forced - 1) / kGeoms) % 3)) : nt_for_m(mm);
    if (int rc = (dtype == 0 ? launch_decode_fp16(s, pl, nt) : launch_decode_bf16(s, pl, nt))) return rc;
  }
  return B200_W4A16_OK;
}

int b200_w4a16_gemm_ex(const void* a, const int32_t* qweight, const void* scales, const int32_t* g_idx,
                       const int32_t* perm, int32_t* workspace, void* c, int64_t m, int64_t n, int64_t k,
                       int group_size, int dtype, int is_k_full, int has_act_order, void* a_tmp,
                       float* c_tmp, void* stream, const b200_w4a16_opts* opts) {
  return gemm_impl(a, qweight, scales, g_idx, perm, workspace, c, m, n, k, group_size, dtype, is_k_full, has_act_order,
                   a_tmp, c_tmp, stream, opts, nullptr);
}

int b200_w4a16_gemm_ar(const void* a, const int32_t* qweight, const void* scales, int32_t* workspace, void* c,
                       int64_t m, int64_t n, int64_t k, int group_size, int dtype, float* c_tmp, void* stream,
                       const b200_w4a16_opts* opts, const b200_w4a16_ar* ar) {
  if (!ar) return fail(B200_W4A16_EINVAL, "gemm_ar: ar is null");
  return gemm_impl(a, qweight, scales, nullptr, nullptr, workspace, c, m, n, k, group_size, dtype, 1, 0, nullptr, c_tmp,
                   stream, opts, ar);
}

int b200_w4a16_gemm(const void* a, const int32_t* qweight, const void* scales, const int32_t* g_idx,
                    const int32_t* perm, int32_t* workspace, void* c, int64_t m, int64_t n, int64_t k,
                    int group_size, int dtype, int is_k_full, int has_act_order, void* a_tmp, float* c_tmp,
                    void* stream) {
  return b200_w4a16_gemm_ex(a, qweight, scales, g_idx, perm, workspace, c, m, n, k, group_size, dtype,
                            is_k_full, has_act_order, a_tmp, c_tmp, stream, nullptr);
}

int b200_w4a16_linear_forward_ex(const void* input, const int32_t* qweight, const void* scales, const void* bias,
                                 int32_t* workspace, void* output, void* tgt, int inplace, int64_t num_tokens,
                                 int64_t dim_out, int64_t dim_in, int group_size, int dtype, void* a_tmp,
                                 float* c_tmp, int64_t c_tmp_bytes, void* stream) {
  // W4A16GPTQMarlinLinear::prefill, w4a16_gptq_marlin_linear.cuh:104-146
  b200_w4a16_opts o;
  std::memset(&o, 0, sizeof(o));
  o.bias = bias;
  o.pdl = g_default_pdl.load(std::memory_order_relaxed);
  o.c_tmp_bytes = c_tmp_bytes;
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

int b200_w4a16_linear_forward(const void* input, const int32_t* qweight, const void* scales, const void* bias,
                              int32_t* workspace, void* output, void* tgt, int inplace, int64_t num_tokens,
                              int64_t dim_out, int64_t dim_in, int group_size, int dtype, void* a_tmp,
                              float* c_tmp, void* stream) {
  return b200_w4a16_linear_forward_ex(input, qweight, scales, bias, workspace, output, tgt, inplace, num_tokens,
                                      dim_out, dim_in, group_size, dtype, a_tmp, c_tmp, 0, stream);
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

// ---------------------------------------------------------------------------------------------------------------
// b200_w4a16_gemm_static_scratch: the streaming GEMM with its scratch in a file-scope device array — for callers that
// may not allocate (include/b200_w4a16.h).  No counterpart in the reference.
// Layout of the array: [partial tiles: B200_W4A16_STATIC_SCRATCH_BYTES][ticket counters: ..._COUNTER_BYTES].
// In HOST code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address handed to the
// kernel comes from cudaGetSymbolAddress, resolved once per device with that device current.
static __device__ __align__(256) unsigned char
    g_static_scratch[B200_W4A16_STATIC_SCRATCH_BYTES + B200_W4A16_STATIC_COUNTER_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_ptr[kMaxDevices] = {};

// device address of the array on the CURRENT device (set-up: once per device); null + error string on failure
unsigned char* static_scratch_ptr() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    b200::fail(B200_W4A16_ECUDA, "gemm_static_scratch: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || !ptr) {
      b200::fail(B200_W4A16_ECUDA, "gemm_static_scratch: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    g_static_ptr[dev] = static_cast<unsigned char*>(ptr);
  }
  return g_static_ptr[dev];
}
}  // namespace

int b200_w4a16_gemm_static_scratch(const void* a, const int32_t* qweight, const void* scales, void* c, int64_t m,
                                   int64_t n, int64_t k, int group_size, int dtype, void* stream) {
  using namespace b200;
  if (m > B200_W4A16_STATIC_MAX_M)
    return fail(B200_W4A16_EINVAL, "gemm_static_scratch: M = %lld exceeds the maximum of %d (the static scratch serves the "
                "streaming kernel only; use b200_w4a16_gemm for wider inputs)", (long long)m, B200_W4A16_STATIC_MAX_M);
  if (n > B200_W4A16_STATIC_MAX_N)
    return fail(B200_W4A16_EINVAL, "gemm_static_scratch: N = %lld exceeds the maximum of %d (ticket counters of the static "
                "array)", (long long)n, B200_W4A16_STATIC_MAX_N);
  if (n > 0 && b200_w4a16_workspace_ints(n) * 4 > B200_W4A16_STATIC_COUNTER_BYTES)
    return fail(B200_W4A16_EINVAL, "gemm_static_scratch: N = %lld needs more ticket counters than the static array holds", (long long)n);
  unsigned char* base = static_scratch_ptr();
  if (!base) return B200_W4A16_ECUDA;
  float* c_tmp = reinterpret_cast<float*>(base);
  int32_t* counters = reinterpret_cast<int32_t*>(base + B200_W4A16_STATIC_SCRATCH_BYTES);
  // the array keeps the previous call's contents; the partial tiles are written before they are read, the counters must
  // start at zero (the kernel leaves them at zero, but a launch that faulted may not have)
  cudaError_t e = cudaMemsetAsync(counters, 0, B200_W4A16_STATIC_COUNTER_BYTES, reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail(B200_W4A16_ECUDA, "gemm_static_scratch: cudaMemsetAsync: %s", cudaGetErrorString(e));
  b200_w4a16_opts o;
  std::memset(&o, 0, sizeof(o));
  o.c_tmp_bytes = B200_W4A16_STATIC_SCRATCH_BYTES;   // plan_decode stays inside it
  o.pdl = 0;                                          // calls share the array: never overlap the previous launch
  return gemm_impl(a, qweight, scales, nullptr, nullptr, counters, c, m, n, k, group_size, dtype, 1, 0, nullptr, c_tmp,
                   stream, &o, nullptr);
}

}  // extern "C"
