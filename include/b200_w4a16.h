/*
 * b200_w4a16.h — C ABI of the B200-native W4A16 GPTQ-Marlin linear (libb200_w4a16.so).
 *
 * This is the drop-in boundary for the quantised-linear hot path of jk3456a/CPM.cu.  Every entry
 * point names the reference interface it replaces (file:line relative to the reference tree).
 * Plain pointers and sizes only: device pointers are raw CUDA device addresses, exactly as the
 * reference passes them through its pybind module (src/entry.cu:536-570), `stream` is a
 * cudaStream_t / CUstream passed as void*.  The library allocates nothing on the device, never
 * synchronises the host with the device, and every launch is CUDA-graph capturable
 * (reference contract: src/entry.cu:540-562).
 *
 * Data formats are the reference's, bit for bit:
 *   qweight  int32 [K/16][2N]   Marlin-packed uint4b8 (scripts/model_convert/gptq2marlin.py:109-134)
 *   scales   T     [K/g ][N]    per-group scales, permuted per 64 columns (gptq2marlin.py:99-107);
 *                               g = 128, or one row permuted per 32 columns when group_size == -1
 *   a, c     T     [M][K], [M][N] row-major,  T = fp16 (dtype 0) or bf16 (dtype 1)
 *                               (dtype codes: cpmcu/llm_w4a16_gptq_marlin.py:12-15)
 *
 * Error convention (reference: src/utils.cuh:54-66 -> Python ValueError / RuntimeError via pybind):
 *   0  success
 *   1  invalid argument  (the reference's VALUE_CHECK  -> std::invalid_argument -> ValueError)
 *   2  runtime error     (the reference's ERROR_CHECK  -> std::runtime_error    -> RuntimeError)
 *   3  CUDA error reported by the runtime/driver while encoding descriptors or launching
 * b200_w4a16_last_error() returns the message of the last failure on the calling thread.
 */
#ifndef B200_W4A16_H_
#define B200_W4A16_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_W4A16_OK 0
#define B200_W4A16_EINVAL 1
#define B200_W4A16_ERUNTIME 2
#define B200_W4A16_ECUDA 3

#define B200_W4A16_FP16 0
#define B200_W4A16_BF16 1

/* Optional fused post-ops and tuning knobs for b200_w4a16_gemm_ex.  Zero-initialise it. */
typedef struct b200_w4a16_opts {
  /* post-ops of W4A16GPTQMarlinLinear::prefill (src/model/w4a16_gptq_marlin/w4a16_gptq_marlin_linear.cuh:140-145),
   * fused into the GEMM epilogue with the reference's rounding points:
   *   out = RN_T(gemm);  if (residual) out = RN_T(residual + out);  if (bias) out = RN_T(out + bias);
   *   if (out_scale) out = RN_T(out * T(out_scale))   */
  const void* bias;      /* T[N] or NULL                        (batched_add, src/model/elementwise.cuh:8-14)  */
  const void* residual;  /* T[M][N] or NULL; may alias `c`      (elementwise_add, src/model/elementwise.cuh:17-23) */
  int64_t c_tmp_bytes;   /* bytes of c_tmp the caller guarantees for THIS call; 0 = the process-wide capacity
                          * (b200_w4a16_set_scratch_capacity) or, if none was set, the reference contract
                          * determine_reduce_max_m(M, 16) x N floats.  The library never touches more. */
  int force_config;      /* 0 = heuristic; >0 = kernel configuration id (tuning / tests only)        */
  int pdl;               /* 1 = launch with programmatic dependent launch (prologue overlaps the previous kernel) */
  /* residual scaling of the layer (elementwise_scale(output, residual_scale) right after o_proj / before the
   * next layer reads down_proj's output, src/model/w4a16_gptq_marlin/w4a16_gptq_marlin_layer.cuh:57,71,79,93;
   * kernel src/model/elementwise.cuh:35-41), fused as the LAST post-op with the reference's rounding:
   *   out = RN_T(out * T(out_scale)).   0 (zero-initialised opts) = no scaling. */
  float out_scale;
  /* 1 = the weight is the fused [gate | up] matrix (N = 2 x intermediate, gptq2marlin.py:162-230) and the
   * epilogue applies the FFN activation: c[M][N/2] = T(g * sigmoid(g) * u) with g, u the gate / up outputs
   * rounded to T first — gated_silu_interleaved (src/model/activation.cuh:7-18), the launch between
   * gateup_proj and down_proj in w4a16_gptq_marlin_ffn.cuh:70-72.  Streaming regime only (M <= 32); value
   * error beyond, where the caller keeps the separate activation kernel.  Needs group_size 128 and
   * N % 256 == 0; does not combine with bias / residual / out_scale / act-order. */
  int gated_silu;
  /* > 0: the weight is the fused q | k | v projection (N = qkv_q_dim + 2 * qkv_kv_dim) and `c` is written SPLIT,
   * [q: M x qkv_q_dim][k: M x qkv_kv_dim][v: M x qkv_kv_dim] — the layout the reference's `permute` launch produces
   * from the row-major output when M > 1 (src/model/attn.cuh:23-39,56-60; w4a16_gptq_marlin_attn.cuh:132) — so that
   * launch and its staging buffer disappear.  Both multiples of 8; does not combine with residual / gated_silu. */
  int qkv_q_dim;
  int qkv_kv_dim;
  int reserved[1];
} b200_w4a16_opts;

/* Replaces gptq_marlin_gemm<T> (src/qgemm/gptq_marlin/gptq_marlin.cuh:11-27, gptq_marlin.cu:42-86)
 * as called by W4A16GPTQMarlinLinear::prefill (w4a16_gptq_marlin_linear.cuh:115-138):
 *     C[M][N] = A[M][K] . dequant(qweight, scales)
 * Fixed by the call site and therefore not parameters: b_q_type = kU4B8, has_zp = false,
 * use_fp32_reduce = true, b_zeros = NULL.
 *   g_idx, perm   NULL unless has_act_order (then perm[K] gathers the columns of A into a_tmp,
 *                 reference: gptq_marlin_utils.cu:28-77; only is_k_full = 1 is supported, as in CPM.cu)
 *   workspace     int32[(N/64)*16], zero on entry, zero again when the call completes
 *                 (reference lock contract: marlin_device_ops.cuh:347-377; w4a16_gptq_marlin_linear.cuh:76)
 *   a_tmp         T[M][K] scratch, touched only when has_act_order
 *   c_tmp         fp32 split-K scratch, sized by the reference's rule: determine_reduce_max_m(M, 16) x N floats
 *                 (gptq_marlin_utils.cu:206-225; callers w4a16_gptq_marlin_layer.cuh:34-37).  The library stays
 *                 inside that bound; it only uses more (full-width stream-K, see b200_w4a16_reduce_bytes_fast)
 *                 when the caller has declared a larger buffer with b200_w4a16_set_scratch_capacity() or
 *                 opts->c_tmp_bytes.  `workspace` and `c_tmp` belong to ONE call at a time: two calls in flight on
 *                 different streams need their own (the reference has a single stream, src/utils.cu:7).
 * Constraints (as the reference, gptq_marlin.cu:71-72, gptq_marlin_mm.cu:53-54,103-131):
 *   M,N,K > 0;  N % 64 == 0;  K % 64 == 0;  group_size in {128, -1};  K % 128 == 0 when group_size == 128. */
int b200_w4a16_gemm(const void* a, const int32_t* qweight, const void* scales,
                    const int32_t* g_idx, const int32_t* perm, int32_t* workspace, void* c,
                    int64_t m, int64_t n, int64_t k, int group_size, int dtype,
                    int is_k_full, int has_act_order, void* a_tmp, float* c_tmp, void* stream);

/* Same, with fused post-ops / knobs (opts may be NULL). */
int b200_w4a16_gemm_ex(const void* a, const int32_t* qweight, const void* scales,
                       const int32_t* g_idx, const int32_t* perm, int32_t* workspace, void* c,
                       int64_t m, int64_t n, int64_t k, int group_size, int dtype,
                       int is_k_full, int has_act_order, void* a_tmp, float* c_tmp, void* stream,
                       const b200_w4a16_opts* opts);

/* Replaces W4A16GPTQMarlinLinear<T>::prefill (w4a16_gptq_marlin_linear.cuh:104-146):
 * GEMM into `output` (or `tgt`), then `inplace` -> tgt += out, `bias` -> tgt += bias, in one launch.
 *   tgt == NULL            -> result in output
 *   tgt != NULL, !inplace  -> result in tgt
 *   tgt != NULL,  inplace  -> tgt = tgt + gemm   (output is not written: the fused epilogue needs no staging) */
int b200_w4a16_linear_forward(const void* input, const int32_t* qweight, const void* scales,
                              const void* bias, int32_t* workspace, void* output, void* tgt,
                              int inplace, int64_t num_tokens, int64_t dim_out, int64_t dim_in,
                              int group_size, int dtype, void* a_tmp, float* c_tmp, void* stream);

/* Same with the size of c_tmp stated (0 = as b200_w4a16_linear_forward); used by the Python operator mirror,
 * which knows the size of the tensor it is handed. */
int b200_w4a16_linear_forward_ex(const void* input, const int32_t* qweight, const void* scales,
                                 const void* bias, int32_t* workspace, void* output, void* tgt,
                                 int inplace, int64_t num_tokens, int64_t dim_out, int64_t dim_in,
                                 int group_size, int dtype, void* a_tmp, float* c_tmp, int64_t c_tmp_bytes,
                                 void* stream);

/* W[K][N] = RN_T(T(q - 8) * T(scale)) written out in T: the bit-exact dequantisation the GEMM
 * kernels apply in registers (reference device code: marlin_device_ops.cuh:91-137, 294-303).
 * group_size == -1 writes T(q - 8) (the reference scales the rounded output instead). */
int b200_w4a16_dequant(const int32_t* qweight, const void* scales, void* w_out, int64_t n, int64_t k,
                       int group_size, int dtype, void* stream);

/* AutoGPTQ int32[K/8][N] -> Marlin int32[K/16][2N] on the device
 * (replaces marlin_repack_qweight, scripts/model_convert/gptq2marlin.py:109-134). */
int b200_w4a16_repack_gptq(const int32_t* qweight_gptq, int32_t* qweight_marlin, int64_t n, int64_t k,
                           void* stream);

/* Un-quantised fp16 / bf16 linear at decode / tree-verify widths: out[M][N] = (input * T(input_scale)) . weight^T.
 * Replaces linear<T>() = cublasGemmEx(OP_T, OP_N, CUBLAS_COMPUTE_32F) as called by Linear<T>::prefill and
 * LMHead<T>::prefill (src/model/linear.cuh:8-37,76-105) for `transposed = true` weights — the lm_head
 * (73448 x 4096, 602 MB read per verify step), the FR-Spec / EAGLE draft heads and EAGLE fc1 / fc2
 * (src/model/eagle.cuh:301-302,446,470) — with one HBM-streaming launch (no cuBLAS, no staging buffer).
 *   weight       T[N][K] row-major (torch nn.Linear layout), 16-byte aligned; K % 8 == 0
 *   input        T[M][K], M <= 64 (the reference sizes the lm_head output for 64 tokens, src/model/model.cuh:116);
 *                larger M is a value error: that regime is a library GEMM, not a weight stream
 *   input_scale  LMHead::head_scale: the input is multiplied by T(input_scale) and rounded to T first
 *                (elementwise_scale, src/model/elementwise.cuh:35-41); 1.0f = plain Linear
 *   inplace      out = RN_T(out + gemm)  (cuBLAS beta = 1: one rounding), else out = RN_T(gemm)
 *   bias         T[N] or NULL, added afterwards in T (batched_add, src/model/elementwise.cuh:8-14)
 *   pdl          1 / 0 = launch with / without programmatic dependent launch; -1 = the process default
 *                (b200_w4a16_set_pdl) */
int b200_w4a16_dense_linear_forward(const void* input, const void* weight, const void* bias, void* output,
                                    int inplace, int64_t num_tokens, int64_t dim_out, int64_t dim_in,
                                    float input_scale, int dtype, int pdl, void* stream);

/* One-shot all-reduce (sum, in rank order, fp32 accumulate) of the [M][N] partial outputs of a
 * row-parallel linear (o / down sharded along K) over NVLink peer memory — the collective that follows
 * the GEMM when the path is sharded (the reference has no multi-GPU code).
 *   peer_in[w]   HOST array of `world` device pointers: rank w's symmetric input buffer T[elems], mapped
 *                into this process (e.g. torch.distributed._symmetric_memory buffer_ptrs)
 *   peer_sig[w]  HOST array of `world` device pointers: rank w's signal pad, int32[2][128][8], zero-initialised
 *   epoch        local device int32[2 * 128 + 1], zero-initialised, private to this (buffer, signal pad) pair; the
 *                last int is a sticky ERROR word: a peer that does not arrive within 2 s (a rank died, or never
 *                launched) sets it to 1, the output of the call is filled with NaN and the barrier state is left
 *                untouched — a late or dead peer is an error the host can read, never a silently wrong sum
 *   out          local T[elems]; elems a multiple of 8
 * Every rank must call it the same number of times in the same order; callers alternate between two
 * (input buffer) sets so that no trailing barrier is needed.  Graph-capturable. */
int b200_w4a16_allreduce_oneshot(const void* const* peer_in, int32_t* const* peer_sig, int32_t* epoch,
                                 void* out, int64_t elems, int dtype, int rank, int world, void* stream);
/* Same collective with a choice of algorithm.  two_shot = 1: rank r adds only slice r of the message
 * out of peer memory into its symmetric result buffer peer_res[r] (T[elems]), a second block-level
 * barrier, then every rank copies the W result slices — 2 x elems read over NVLink instead of
 * W x elems; for large messages (M = 64) at W >= 4.  Signal pads are int32[2][128][8] in both modes. */
int b200_w4a16_allreduce(const void* const* peer_in, void* const* peer_res, int32_t* const* peer_sig,
                         int32_t* epoch, void* out, int64_t elems, int dtype, int rank, int world,
                         int two_shot, void* stream);

/* Row-parallel linear (o / down sharded along K) with the all-reduce FUSED into the GEMM launch (M <= 32, the
 * streaming kernel): the CTA that completes a column tile pushes it, rounded to T, into every rank's receive buffer
 * over NVLink and raises a per-tile counter; after its last unit every CTA reduces its share of the tiles out of LOCAL
 * memory (fp32, rank order: bit-identical on every rank), applies the post-ops of `opts` and writes `c`.  No separate
 * collective launch, no grid-wide barrier; early tiles travel while the rest of the weights stream.
 *   recv[w]   HOST array of `world` device pointers: rank w's receive buffer, T[2][world][slot_elems] (symmetric
 *             memory, e.g. torch.distributed._symmetric_memory), slot_elems >= m * n, a multiple of 8
 *   flag[w]   HOST array of `world` device pointers: rank w's counters, int32[world][1024], zero-initialised once
 *   seen      local int32[1025], zero-initialised once (epochs per column tile; the last int is a sticky error word set
 *             when a source rank did not deliver within 2 s — the affected tiles are then not written)
 * Every rank must make the same sequence of calls on the same (recv, flag, seen) set.  M > 32, gated_silu and
 * act-order are value errors (use b200_w4a16_gemm_ex + b200_w4a16_allreduce there).  Graph-capturable. */
typedef struct b200_w4a16_ar {
  void* const* recv;
  int32_t* const* flag;
  int32_t* seen;
  int64_t slot_elems;
  int rank, world;
} b200_w4a16_ar;
int b200_w4a16_gemm_ar(const void* a, const int32_t* qweight, const void* scales, int32_t* workspace, void* c,
                       int64_t m, int64_t n, int64_t k, int group_size, int dtype, float* c_tmp, void* stream,
                       const b200_w4a16_opts* opts, const b200_w4a16_ar* ar);

/* The decode-regime GEMM (M <= 32, the streaming kernel) for callers that cannot allocate device memory (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arithmetic as b200_w4a16_gemm; the caller passes only inputs and the
 * output — the split-column scratch (fp32 partial tiles) and the per-column-tile ticket counters live in ONE
 * file-scope `static __device__` array inside the library:
 *   B200_W4A16_STATIC_SCRATCH_BYTES (4 MiB) of partial tiles + B200_W4A16_STATIC_COUNTER_BYTES (64 KiB) of counters,
 *   held by every GPU that loads the library until the process exits.
 * Limits (B200_W4A16_EINVAL with a message beyond them): M <= 32, N <= B200_W4A16_STATIC_MAX_N (65536), fp16 / bf16,
 * group_size 128 or -1, no act-order.  Inside the limits every shape runs: where full-width stream-K would need more
 * than 4 MiB the launch plan splits fewer columns (plan_decode), results stay within the parity tolerance.
 * The call makes no cudaMalloc / cudaFree / allocating library call: the array's device address is resolved once per
 * device (cudaGetSymbolAddress, with that device current) and the counters are cleared with cudaMemsetAsync on
 * `stream` before every launch (a launch that faulted may have left a ticket behind).
 * EVERY call on a device shares the array and it keeps the previous call's contents: calls that overlap on the GPU
 * (two streams, or PDL-chained launches of this entry point) RACE on it — serialise them (one stream, which this
 * entry point enforces for itself by launching without the PDL attribute). */
#define B200_W4A16_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
#define B200_W4A16_STATIC_COUNTER_BYTES (64 * 1024)
#define B200_W4A16_STATIC_MAX_M 32
#define B200_W4A16_STATIC_MAX_N 65536
int b200_w4a16_gemm_static_scratch(const void* a, const int32_t* qweight, const void* scales, void* c, int64_t m,
                                   int64_t n, int64_t k, int group_size, int dtype, void* stream);

/* Process-wide default for programmatic dependent launch on entry points that take no opts
 * (b200_w4a16_gemm, b200_w4a16_linear_forward): 1 = every GEMM launch carries the PDL attribute so
 * its weight prefetch overlaps the tail of the previous kernel in the stream (safe with any
 * predecessor: the kernels touch activations / outputs only after griddepcontrol.wait).
 * Returns the previous setting.  Initial value: environment variable B200_W4A16_PDL (default 0). */
int b200_w4a16_set_pdl(int enable);

/* Sizing helpers and the scratch contract.
 * b200_w4a16_reduce_max_m  == marlin::determine_reduce_max_m(prob_m, 16) (gptq_marlin_utils.cu:206-225).
 * b200_w4a16_reduce_bytes  bytes of c_tmp the library touches for (m, n, k) when the caller guarantees only the
 *                          reference contract; always <= reduce_max_m(m) * n * 4.  Inside that bound the split
 *                          columns of the stream-K kernel get at most reduce_max_m(m) / m fp32 slots; where that
 *                          is not enough the launch uses fewer CTAs or whole columns per CTA (slower, no scratch).
 * b200_w4a16_reduce_bytes_fast  bytes full-width stream-K needs (<= 16 MiB for m <= 32 on the MiniCPM4-8B shapes).
 * b200_w4a16_set_scratch_capacity(bytes)  process-wide promise: every c_tmp passed to an entry point without a size
 *                          (b200_w4a16_gemm, b200_w4a16_linear_forward, the gptq_marlin_gemm shim) has at least
 *                          `bytes` bytes.  CPM.cu allocates reduce_max_m(chunk_length) * 2*intermediate floats per
 *                          model (w4a16_gptq_marlin_layer.cuh:34-37: 128 MiB for MiniCPM4-8B at chunk 1024) and
 *                          reduce_max_m(chunk) * 2*hidden floats for EAGLE (minicpm4_eagle.cuh:134-137): an
 *                          integration calls this once after init_storage with the smaller of the two.  0 (default)
 *                          = the reference contract only.  Returns the previous value.
 * b200_w4a16_workspace_ints == (n / 64) * 16 (w4a16_gptq_marlin_linear.cuh:76). */
int b200_w4a16_reduce_max_m(int64_t m);
int64_t b200_w4a16_reduce_bytes(int64_t m, int64_t n, int64_t k);
int64_t b200_w4a16_reduce_bytes_fast(int64_t m, int64_t n, int64_t k);
int64_t b200_w4a16_set_scratch_capacity(int64_t bytes);
int64_t b200_w4a16_workspace_ints(int64_t n);

/* Introspection: kernels launched by this library since load (for bench.py's gpu_launches),
 * the configuration id the heuristic picks for a shape, last error, version string. */
int64_t b200_w4a16_launch_count(void);
int b200_w4a16_pick_config(int64_t m, int64_t n, int64_t k, int group_size, int dtype);
const char* b200_w4a16_config_name(int config);
/* clusters of the configuration's CTAs the current device holds at once (cudaOccupancyMaxActiveClusters);
 * -1 when unknown (no device) or the configuration does not use clusters.  Tuning / tests only. */
int b200_w4a16_max_clusters(int config, int64_t m, int group_size, int dtype);
const char* b200_w4a16_last_error(void);
const char* b200_w4a16_version(void);
/* Tuning aid: when device_buf != NULL the next `launches` streaming-kernel launches record per-CTA
 * phase timestamps (%globaltimer; 8 x uint64 per CTA, 1024 CTAs = 64 KiB per launch, launch order)
 * into device_buf (launches * 65536 bytes).  NULL switches it off.  Never used by the product path. */
void b200_w4a16_debug_timeline(void* device_buf, int64_t launches);

#ifdef __cplusplus
}
#endif
#endif /* B200_W4A16_H_ */
