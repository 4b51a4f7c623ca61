// Shared device helpers for the sm_100a W4A16 kernels: mbarrier / TMA / PDL PTX wrappers, the
// legacy-path MMA used in the HBM-bound regime, and THE dequantisation routine (one definition,
// used by every kernel and by the dequant dump, so "bit-exact dequant" is one piece of code).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// ---------------------------------------------------------------------------------------------
// small PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}

// same, for waits that are expected to be long (a whole main loop): back off between polls so the
// spinning warp does not take issue slots from the warps doing the work
__device__ __forceinline__ void mbar_wait_backoff(uint32_t bar, uint32_t parity, uint32_t ns) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "nanosleep.u32 %2;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity), "r"(ns)
      : "memory");
}

// L2 eviction-priority policies for TMA loads (createpolicy encodings)
static constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
static constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
static constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2, uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2, int c3, uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// programmatic dependent launch: no-ops unless the launch carried the attribute
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait_prior_grids() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

template <int lut>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm volatile("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(lut));
  return r;
}

__device__ __forceinline__ void ldmatrix_x2(uint32_t& r0, uint32_t& r1, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];"
               : "=r"(r0), "=r"(r1)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3,
                                            uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

// thread-block clusters / distributed shared memory
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t local_addr, uint32_t rank) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(rank));
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(ra)
               : "memory");
  return v;
}

// 16-byte shared-memory load by 32-bit shared address (keeps the access an LDS: pointers derived
// from the manually aligned dynamic-smem base are generic to the compiler)
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}

// four 8x8 b16 matrices, register i of every lane -> matrix i, row addresses from lanes 8i..8i+7
__device__ __forceinline__ void stmatrix_x4(uint32_t addr, uint32_t r0, uint32_t r1, uint32_t r2,
                                            uint32_t r3) {
  asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(r0),
               "r"(r1), "r"(r2), "r"(r3)
               : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05 (5th-generation tensor cores): tensor memory, UMMA descriptors, issue / commit / load
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// arrive on an mbarrier when every tcgen05.mma issued so far by this thread has completed
// (implies tcgen05.fence::before_thread_sync)
// true in exactly one lane of a converged warp (the form ptxas keeps uniform-register operands for)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (fp16 / bf16 inputs, fp32 accumulate)
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// ---- cta_group::2: a CTA pair (cluster of 2) issues ONE UMMA over both SMs.  M = 256: each CTA supplies 128 rows of A
// from its own shared memory (same offsets in both) and holds the matching 128 accumulator lanes in its own tensor
// memory; the N columns of B come half from each CTA.  Only the leader (cluster rank 0) issues; barriers the peer
// feeds live in the leader's shared memory and are reached through shared::cluster addresses (own address with the
// rank bit cleared), tcgen05.commit multicasts its arrive to the same barrier offset in both CTAs.
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in cluster rank 0
__device__ __forceinline__ void tmem_alloc_pair(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {   // arrives on `bar`'s offset in BOTH CTAs
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"(static_cast<uint16_t>(3))
               : "memory");
}
__device__ __forceinline__ void tc_mma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at `bar`'s offset in cluster rank 0.  Plain remote arrive (the form CUTLASS's
// ClusterBarrier::arrive(cta_id) emits): what it publishes is the arriving warp's OWN-CTA shared memory, already made
// visible to that SM's async proxy by the warp's fence.proxy.async, and read only by that SM's tensor core after the
// leader's wait.  The `.release.cluster` form puts a cluster-scope fence (ERRBAR, ~1 us) in front of every arrive: with
// it the eight dequant warps spent half their time fencing (profiles/r02/i_*).
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerBitMask) : "memory");
}
// TMA load whose transaction bytes are credited to the LEADER's barrier (executed by both CTAs of the pair)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                                 uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}

// K-major operand tile in shared memory, 128-byte rows, SWIZZLE_128B, 8-row atoms 1024 B apart
// (bit layout: cute::UMMA::SmemDescriptor — start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46),
//  version=1 [46,48), layout_type [61,64) with SWIZZLE_128B = 2)
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(1) << 16;             // LBO: unused for swizzled K-major
  d |= static_cast<uint64_t>(1024 >> 4) << 32;     // SBO: 8 rows * 128 B
  d |= static_cast<uint64_t>(1) << 46;             // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(2) << 61;             // SWIZZLE_128B
  return d;
}
// instruction descriptor, kind::f16: D fp32, A/B both T, both K-major, shape M x N
// (cute::UMMA::InstrDescriptor: c_format [4,6), a_format [7,10), b_format [10,13), n>>3 [17,23), m>>4 [24,29))
template <typename T>
__device__ __forceinline__ uint32_t umma_idesc_f16(uint32_t m, uint32_t n);
template <>
__device__ __forceinline__ uint32_t umma_idesc_f16<__half>(uint32_t m, uint32_t n) {
  return (1u << 4) | (0u << 7) | (0u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}
template <>
__device__ __forceinline__ uint32_t umma_idesc_f16<__nv_bfloat16>(uint32_t m, uint32_t n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}
// 32 lanes x 32 consecutive fp32 columns: lane i of the warp reads TMEM lane (taddr.lane + i)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
        "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
        "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
        "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x N consecutive fp32 columns, N in {8, 16, 32}
template <int N>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&v)[N]);
template <>
__device__ __forceinline__ void tmem_ld_cols<8>(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
template <>
__device__ __forceinline__ void tmem_ld_cols<16>(uint32_t taddr, uint32_t (&v)[16]) { tmem_ld_32x16(taddr, v); }
template <>
__device__ __forceinline__ void tmem_ld_cols<32>(uint32_t taddr, uint32_t (&v)[32]) { tmem_ld_32x32(taddr, v); }
// ---- A operand in TENSOR MEMORY, written from registers (tools/probe_tmem_a.cu pins the layout on B200) ----
// One 16-row x 16-k block of a 16-bit A operand in the mma.sync fragment layout: thread (g = lane / 4, t = lane % 4)
// passes {(row g, k 2t..2t+1), (row g + 8, same k), (row g, k 2t+8..2t+9), (row g + 8, same k)} — exactly
// Dequant<T>::word's w[0..3] — and they land in TMEM lanes [lane base, + 16) x 8 columns in natural k order, which is
// what tcgen05.mma reads as rows of a K-major A operand.  The lane base must lie in the issuing warp's quarter
// (32 * (warp % 4) .. + 31); both halves (+ 0, + 16) are addressable.
__device__ __forceinline__ void tmem_st_16x128b_x2(uint32_t taddr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
  asm volatile("tcgen05.st.sync.aligned.16x128b.x2.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(r0), "r"(r1), "r"(r2),
               "r"(r3)
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// D[tmem] (+)= A[tmem: 128 lanes x 8 columns per k16] * B[smem desc], kind::f16, one CTA
__device__ __forceinline__ void tc_mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// element-type traits
// ---------------------------------------------------------------------------------------------
template <typename T>
struct Elem;
template <>
struct Elem<__half> {
  using T2 = __half2;
  static constexpr CUtensorMapDataType kTmaType = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  static __device__ __forceinline__ __half from_float(float f) { return __float2half_rn(f); }
  static __device__ __forceinline__ float to_float(__half h) { return __half2float(h); }
};
template <>
struct Elem<__nv_bfloat16> {
  using T2 = __nv_bfloat162;
  static constexpr CUtensorMapDataType kTmaType = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  static __device__ __forceinline__ __nv_bfloat16 from_float(float f) {
    return __float2bfloat16_rn(f);
  }
  static __device__ __forceinline__ float to_float(__nv_bfloat16 h) {
    return __bfloat162float(h);
  }
};

// last post-op of the epilogues: out = RN_T(out * T(scale)) (the layer's residual scaling, reference:
// elementwise_scale, src/model/elementwise.cuh:35-41); os = 0: none, else 0x10000 | bits of T(scale)
template <typename T>
__device__ __forceinline__ T apply_out_scale(T o, uint32_t os) {
  if (os) {
    const uint16_t b = static_cast<uint16_t>(os & 0xffffu);
    o = __hmul(o, *reinterpret_cast<const T*>(&b));
  }
  return o;
}

// Output element offset.  qa == 0: row-major C[tok][n].  qa > 0 (opts.qkv_q_dim / qkv_kv_dim): the fused q|k|v
// projection written SPLIT, [q: M x qa][k: M x qb][v: M x qb] — what the reference's permute kernel produces from the
// row-major qkv output when M > 1 (src/model/attn.cuh:23-39,56-60; w4a16_gptq_marlin_attn.cuh:132), without the launch.
__device__ __forceinline__ size_t out_offset(int tok, int n, int n_total, int m_total, int qa, int qb) {
  if (qa == 0) return static_cast<size_t>(tok) * n_total + n;
  if (n < qa) return static_cast<size_t>(tok) * qa + n;
  if (n < qa + qb) return static_cast<size_t>(m_total) * qa + static_cast<size_t>(tok) * qb + (n - qa);
  return static_cast<size_t>(m_total) * (qa + qb) + static_cast<size_t>(tok) * qb + (n - qa - qb);
}

// The same mapping seen from one output column: offset of (token 0, column n) and the token stride, so that
// out_offset(tok, n, ...) == *col_off + tok * *tok_stride.  An epilogue whose thread owns a column hoists this.
__device__ __forceinline__ void out_column(int n, int n_total, int m_total, int qa, int qb, size_t* col_off,
                                           int* tok_stride) {
  if (qa == 0) {
    *col_off = static_cast<size_t>(n);
    *tok_stride = n_total;
  } else if (n < qa) {
    *col_off = static_cast<size_t>(n);
    *tok_stride = qa;
  } else if (n < qa + qb) {
    *col_off = static_cast<size_t>(m_total) * qa + (n - qa);
    *tok_stride = qb;
  } else {
    *col_off = static_cast<size_t>(m_total) * (qa + qb) + (n - qa - qb);
    *tok_stride = qb;
  }
}

// D(16x8,f32) += A(16x16,T,row) * B(16x8,T,col): the register-operand tensor path.  In the
// HBM-bound regime the weights arrive as registers straight out of the dequantiser, which is the
// one case where mma.sync (not tcgen05, whose operands live in shared/tensor memory) is the right
// instruction on sm_100a: a round trip through shared memory would cost 4x the bytes of the int4
// stream (DESIGN.md "Why mma.sync below the ridge").
template <typename T>
__device__ __forceinline__ void mma_m16n8k16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                             uint32_t b1);
template <>
__device__ __forceinline__ void mma_m16n8k16<__half>(float (&d)[4], const uint32_t (&a)[4],
                                                     uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma_m16n8k16<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4],
                                                            uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// ---------------------------------------------------------------------------------------------
// dequantisation (the reference's device semantics: marlin_device_ops.cuh:91-137 for the exact
// integer -> float step, :294-303 for the single-rounding group-scale multiply)
//
// One 32-bit word of the Marlin stream holds, for lane (g = lane/4, t = lane%4) and 16-column
// tile j, the eight uint4b8 nibbles of rows k in {2t, 2t+1, 2t+8, 2t+9} x cols n in {g, g+8}:
//     nibble:   0        1        2        3        4        5        6        7
//     (k, n): (2t,g) (2t+8,g) (2t,g+8) (2t+8,g+8) (2t+1,g) (2t+9,g) (2t+1,g+8) (2t+9,g+8)
// dequant_word() returns them as the four 32-bit registers of a 16(n) x 16(k) row-major
// m16n8k16 A-operand fragment (weights as the M side, tokens as the N side):
//     w[0] = {W[2t  ][g  ], W[2t+1][g  ]}      w[1] = {W[2t  ][g+8], W[2t+1][g+8]}
//     w[2] = {W[2t+8][g  ], W[2t+9][g  ]}      w[3] = {W[2t+8][g+8], W[2t+9][g+8]}
// Integer step: the nibble is OR-ed into the mantissa of a magic constant (1024+q for fp16,
// 128+q for bf16) and the bias (+8 zero point) removed by one exact subtraction / FMA; every
// value q-8 in [-8, 7] is exact, so any correct method is bit-identical to the reference's.
// Scale step (SCALED): one multiply in T with round-to-nearest-even — the only rounding on the
// weight path and the reference's rounding point (HMUL2).
// ---------------------------------------------------------------------------------------------
template <typename T>
struct Dequant;

template <>
struct Dequant<__half> {
  // sc = {scale of column g (low half), scale of column g+8 (high half)}: the compiler folds the two broadcasts
  // into the .H0_H0 / .H1_H1 operand modifiers of HMUL2 (no PRMT, one register per scale pair)
  template <bool SCALED>
  static __device__ __forceinline__ void word(uint32_t q, uint32_t sc, uint32_t (&w)[4]) {
    constexpr uint32_t LO = 0x000f000f, HI = 0x00f000f0, EX = 0x64006400;
    constexpr uint32_t SUB = 0x64086408;  // 1032 = 1024 + 8
    constexpr uint32_t MUL = 0x2c002c00;  // 1/16
    constexpr uint32_t ADD = 0xd480d480;  // -72 = -(1024/16 + 8)
    const uint32_t q8 = q >> 8;
    uint32_t lo0 = lop3<(0xf0 & 0xcc) | 0xaa>(q, LO, EX);
    uint32_t hi0 = lop3<(0xf0 & 0xcc) | 0xaa>(q, HI, EX);
    uint32_t lo1 = lop3<(0xf0 & 0xcc) | 0xaa>(q8, LO, EX);
    uint32_t hi1 = lop3<(0xf0 & 0xcc) | 0xaa>(q8, HI, EX);
    __half2 v0 = __hsub2(as_h2(lo0), as_h2(SUB));
    __half2 v2 = __hfma2(as_h2(hi0), as_h2(MUL), as_h2(ADD));
    __half2 v1 = __hsub2(as_h2(lo1), as_h2(SUB));
    __half2 v3 = __hfma2(as_h2(hi1), as_h2(MUL), as_h2(ADD));
    if (SCALED) {
      const __half2 s_lo = __low2half2(as_h2(sc)), s_hi = __high2half2(as_h2(sc));
      v0 = __hmul2(v0, s_lo);
      v2 = __hmul2(v2, s_lo);
      v1 = __hmul2(v1, s_hi);
      v3 = __hmul2(v3, s_hi);
    }
    w[0] = as_u32(v0);
    w[1] = as_u32(v1);
    w[2] = as_u32(v2);
    w[3] = as_u32(v3);
  }
  static __device__ __forceinline__ __half2 as_h2(uint32_t u) {
    return *reinterpret_cast<__half2*>(&u);
  }
  static __device__ __forceinline__ uint32_t as_u32(__half2 h) {
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

template <>
struct Dequant<__nv_bfloat16> {
  template <bool SCALED>
  static __device__ __forceinline__ void word(uint32_t q, uint32_t sc, uint32_t (&w)[4]) {
    constexpr uint32_t MASK = 0x000f000f, EX = 0x43004300;  // 128 + q, ulp(128) = 1
    constexpr uint32_t ONE = 0x3f803f80;                    // 1.0
    constexpr uint32_t ADD = 0xc308c308;                    // -136 = -(128 + 8)
    uint32_t lo0 = lop3<(0xf0 & 0xcc) | 0xaa>(q, MASK, EX);
    uint32_t hi0 = lop3<(0xf0 & 0xcc) | 0xaa>(q >> 4, MASK, EX);
    uint32_t lo1 = lop3<(0xf0 & 0xcc) | 0xaa>(q >> 8, MASK, EX);
    uint32_t hi1 = lop3<(0xf0 & 0xcc) | 0xaa>(q >> 12, MASK, EX);
    __nv_bfloat162 v0 = __hfma2(as_b2(lo0), as_b2(ONE), as_b2(ADD));
    __nv_bfloat162 v2 = __hfma2(as_b2(hi0), as_b2(ONE), as_b2(ADD));
    __nv_bfloat162 v1 = __hfma2(as_b2(lo1), as_b2(ONE), as_b2(ADD));
    __nv_bfloat162 v3 = __hfma2(as_b2(hi1), as_b2(ONE), as_b2(ADD));
    if (SCALED) {
      const __nv_bfloat162 s_lo = __low2bfloat162(as_b2(sc)), s_hi = __high2bfloat162(as_b2(sc));
      v0 = __hmul2(v0, s_lo);
      v2 = __hmul2(v2, s_lo);
      v1 = __hmul2(v1, s_hi);
      v3 = __hmul2(v3, s_hi);
    }
    w[0] = as_u32(v0);
    w[1] = as_u32(v1);
    w[2] = as_u32(v2);
    w[3] = as_u32(v3);
  }
  static __device__ __forceinline__ __nv_bfloat162 as_b2(uint32_t u) {
    return *reinterpret_cast<__nv_bfloat162*>(&u);
  }
  static __device__ __forceinline__ uint32_t as_u32(__nv_bfloat162 h) {
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

// A lane's 16 bytes of a 64-column scale chunk are 8 scales for columns g + 8*jj, jj = 0..7
// (gptq2marlin.py:55-57), i.e. word j = {s[16j+g], s[16j+g+8]}: exactly the pair Dequant<T>::word takes.

// inverse of SCALE_PERM_SINGLE (gptq2marlin.py:58-60): position of column (n % 32) inside its
// 32-column block of a channelwise (group_size == -1) scale row
__device__ __forceinline__ int chan_scale_pos(int n) {
  const int c = n & 31;
  // SCALE_PERM_SINGLE[8*i + jj] = 2*i + {0,1,8,9,16,17,24,25}[jj]
  const int i = (c & 7) >> 1;
  const int jj = ((c >> 3) << 1) | (c & 1);
  return (n & ~31) | (8 * i + jj);
}

}  // namespace b200
