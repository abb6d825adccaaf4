// Dual-mode int4-weight GEMM for sm_100a: one persistent, TMA-fed, tcgen05 kernel template serves
//   MODE_I8  (draft,  W4A4):  acc_i32[m,n] = sum_k a_i4[m,k] * w_i4[n,k]   (exact), fp32 scale epilogue
//   MODE_F16 / MODE_BF16 (verify, W4A16): out[m,n] = sum_k x[m,k] * w_i4[n,k] * sW[n] (+bias), fp32 accumulate
// over the SAME packed [N, K/2] weight bytes (two's-complement nibbles, even k in the low nibble).
//
// Replaces torchao's rowwise_scaled_linear_cutlass_s4s4_unified (third-party/ao/torchao/csrc/cuda/
// rowwise_scaled_linear_cutlass/rowwise_scaled_linear_cutlass_unified.cuh:240-488, an SM80 mma.sync.s4
// kernel) and the BitBLAS Matmul + per-call `weight ^ 0x88` pass (quarot_nn/linear.py:102-124,156-211).
//
// Design (nothing here follows the reference's kernels):
//  * swap-AB: the 128 output channels of a tile are the UMMA M dimension, tokens are UMMA N (16..128).
//  * The weight operand never exists in shared memory in MMA form. A dedicated TMA warp streams packed
//    int4 tiles (128 channels x 128 bytes = 256 k, SWIZZLE_128B so that per-row reads are bank-conflict
//    free) into a deep mbarrier ring; 8 converter warps read one row each, widen in registers
//    (int4 -> int8 as nibble<<4, or int4 -> f16/bf16 by the magic-number trick) and write the result
//    straight into TENSOR MEMORY with tcgen05.st; tcgen05.mma then takes A from TMEM and B (tokens)
//    from shared memory. Shared-memory traffic per weight byte is therefore 1 write (TMA) + 1 read,
//    which keeps the kernel HBM-bound instead of smem-port bound.
//  * Activations: W4A4 packed int4 arrive by TMA and are widened into the canonical K-major
//    SWIZZLE_128B int8 layout by the same converter warps; W4A16 f16/bf16 activations are TMA'd
//    directly in that layout (no conversion).
//  * Persistent stream-K: grid = #SMs; the (tile, k-unit) space is cut into equal contiguous ranges,
//    so every SM streams the same number of weight bytes whatever N and K are. A tile whose k range
//    is shared by several CTAs is reduced through a workspace in ascending CTA order by the last
//    arriver (int32: exact; fp32: fixed order => deterministic). No element-wise atomics.
//  * Two accumulator buffers in TMEM let the 4 epilogue warps drain tile i while tile i+1 is computed.
//  * Programmatic dependent launch: the weight stream starts before griddepcontrol.wait; only the
//    activation loads wait for the producer kernel.
#include "common.cuh"

#include <cuda.h>
#include <stdlib.h>
#include <mutex>
#include <unordered_map>

namespace qspec {
namespace w4s {

constexpr int MODE_I8 = 0, MODE_F16 = 1, MODE_BF16 = 2, MODE_A8 = 3;   // A8: int8 activations x int4 weights (W4A8)
constexpr bool mode_is_int(int m) { return m == MODE_I8 || m == MODE_A8; }

constexpr int TM = 128;              // channels per tile (UMMA M)
constexpr int UK = 256;              // k per unit
constexpr int UB = 128;              // packed weight bytes per row per unit
constexpr int W_STAGE = TM * UB;     // 16 KB
constexpr int NCONV = 8;             // converter warps: (lane quarter q = w & 3, half h = w >> 2)
constexpr int NEPI = 4;              // epilogue warps
constexpr int WARP_MMA = NCONV + NEPI;
constexpr int WARP_TMA = WARP_MMA + 1;
constexpr int THREADS = (WARP_TMA + 1) * 32;   // 448
constexpr int MAX_STAGES = 12;
constexpr int MAX_TA = 14;           // TMEM operand slots of 32 columns (I8: 128 k, F16: 64 k)

// ---- PTX wrappers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
// Every wait of the pipeline goes through mbar_wait. -DQSPEC_GEMM_WATCHDOG (python -m qspec_b200.build --watchdog) makes a
// barrier that never completes trap after 2^26 failed polls (seconds) instead of hanging the GPU: used while a pipeline
// protocol is being changed, not in the product build.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
#ifdef QSPEC_GEMM_WATCHDOG
  uint32_t spins = 0;
#endif
  do {
#if defined(QSPEC_MBAR_HINT_NS) && QSPEC_MBAR_HINT_NS > 0
    // explicit suspend-time hint: a waiting warp sleeps in hardware until the phase completes (or the hint expires) instead
    // of re-issuing the poll every system-default time slice -- polling warps compete with the converters for issue slots
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity), "r"((uint32_t)QSPEC_MBAR_HINT_NS) : "memory");
#else
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
#endif
#ifdef QSPEC_GEMM_WATCHDOG
    if (!done && ++spins == (1u << 26)) asm volatile("trap;\n");
#endif
  } while (!done);
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster. Default semantics (release, cta
// scope) like CUTLASS's ClusterBarrier::arrive(cta_id): what the arrival orders -- tcgen05.st results and async-proxy
// visible shared memory -- is fenced explicitly (tcgen05.wait::st + fence::before_thread_sync, fence.proxy.async) before it.
// A cluster-scope release / acquire on these barriers costs a full fence (L1 invalidate) per slot hand-over: measured 2x
// slower kernels (profiles/r02b_pair_first.md).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t rank) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];\n" ::"r"(mapa_shared(bar, rank)) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;\n" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory"); }
// One lane of a CONVERGED warp. tcgen05.mma / commit / TMA take their operands from uniform registers: issued
// under elect.sync the compiler moves them there directly, whereas under `if (lane == 0)` it wraps every such
// instruction in an ELECT/R2UR.BROADCAST waterfall loop (~85 cycles per MMA, measured).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, P;\n"
      "}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}

// PAIR: the CTA pair (cta_group::2) allocates the same columns in both SMs' tensor memory; one warp of EACH CTA issues it.
template <bool PAIR>
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  if (PAIR) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(dst_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
  } else {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(dst_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  }
}
template <bool PAIR>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(ncols) : "memory");
  else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(ncols) : "memory");
}
// PAIR: the arrival is multicast to the barrier at the same offset in both CTAs of the pair.
template <bool PAIR>
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  if (PAIR)
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
  else
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]; A is 128 lanes x 8 columns (32 bytes of K per row).
// PAIR (cta_group::2): ONE instruction, issued by the leader CTA, drives both SMs' tensor cores: M = 256 = the 128 weight
// channels in EACH CTA's tensor memory (same columns), the token tile B split between the two CTAs' shared memory (each
// holds N/2 rows at the same offset), D = 128 lanes x N columns in each CTA. Measured on B200 (scripts/umma_probe.cu,
// profiles/r02_umma_probe.md): ~58 cycles per instruction for N <= 64 in BOTH forms, so the pair halves the tensor-pipe
// issue time per weight byte.
template <int MODE, bool PAIR>
__device__ __forceinline__ void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
#define QSPEC_UMMA(GRP, KND)                                                                                            \
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::" GRP ".kind::" KND " [%0], [%1], %2, %3, p;\n}\n" \
               ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory")
  if (MODE == MODE_I8 || MODE == MODE_A8) {
    if (PAIR) QSPEC_UMMA("2", "i8"); else QSPEC_UMMA("1", "i8");
  } else {
    if (PAIR) QSPEC_UMMA("2", "f16"); else QSPEC_UMMA("1", "f16");
  }
#undef QSPEC_UMMA
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};\n" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

// 2-D TMA tile load, completion on an mbarrier, with an L2 eviction-priority hint.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;\n" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];\n" ::"l"(map) : "memory");
}
constexpr uint64_t POLICY_EVICT_NORMAL = 0x1000000000000000ull;
constexpr uint64_t POLICY_EVICT_FIRST = 0x12F0000000000000ull;   // streamed-once weights
constexpr uint64_t POLICY_EVICT_LAST = 0x14F0000000000000ull;    // activations re-read by every CTA

// K-major SWIZZLE_128B shared-memory matrix descriptor (see cute::UMMA::SmemDescriptor):
//   [0,14) start>>4 | [32,46) SBO>>4 = 1024 B (8 rows x 128 B) | [46,48) version 1 | [61,64) layout 2 (SW128)
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format [4,6), a_format [7,10), b_format [10,13),
// K-major A and B, N>>3 at [17,23), M>>4 at [24,29).
template <int MODE, bool PAIR>
__device__ __forceinline__ uint32_t make_idesc(int np) {
  const uint32_t fmt = (MODE == MODE_I8 || MODE == MODE_A8) ? ((2u << 4) | (1u << 7) | (1u << 10))      // S32 += S8 * S8
                     : MODE == MODE_F16 ? ((1u << 4) | (0u << 7) | (0u << 10))     // F32 += F16 * F16
                                        : ((1u << 4) | (1u << 7) | (1u << 10));    // F32 += BF16 * BF16
  return fmt | ((uint32_t)(np >> 3) << 17) | ((uint32_t)((PAIR ? 2 * TM : TM) >> 4) << 24);
}

// ---- int4 widening -----------------------------------------------------------------------------
// 16 packed bytes (32 nibbles) -> 32 int8 (nibble << 4; value = 16*q) in 8 words. Inside each
// 8-nibble group the k order becomes (0,2,4,6,1,3,5,7); weights and activations both go through this
// routine, so the permutation cancels in the dot product. acc = 256 * sum(a*w) exactly.
__device__ __forceinline__ void widen_i8(const uint4& p, uint32_t (&r)[8]) {
  r[0] = (p.x << 4) & 0xf0f0f0f0u; r[1] = p.x & 0xf0f0f0f0u;
  r[2] = (p.y << 4) & 0xf0f0f0f0u; r[3] = p.y & 0xf0f0f0f0u;
  r[4] = (p.z << 4) & 0xf0f0f0f0u; r[5] = p.z & 0xf0f0f0f0u;
  r[6] = (p.w << 4) & 0xf0f0f0f0u; r[7] = p.w & 0xf0f0f0f0u;
}
// Same widening in NATURAL k order (W4A8: the int8 activations are used as they lie in memory): bytes k..k+3 of a word.
__device__ __forceinline__ void widen_i8_nat(const uint4& p, uint32_t (&r)[8]) {
  const uint32_t w[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t lo = (w[i] << 4) & 0xf0f0f0f0u, hi = w[i] & 0xf0f0f0f0u;   // nibbles 0,2,4,6 | 1,3,5,7
    r[2 * i] = __byte_perm(lo, hi, 0x5140);                                     // k 0,1,2,3
    r[2 * i + 1] = __byte_perm(lo, hi, 0x7362);                                 // k 4,5,6,7
  }
}
__device__ __forceinline__ uint32_t lop3_and_xor(uint32_t a, uint32_t mask, uint32_t x) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0x6a;\n" : "=r"(r) : "r"(a), "r"(mask), "r"(x));  // (a & b) ^ c
  return r;
}
// One packed word (8 two's-complement nibbles, k ascending from bit 0) -> 4 words of 2 x f16/bf16 in
// NATURAL k order (the activations are used as they lie in memory), exact.
template <bool BF16>
__device__ __forceinline__ void dequant8(uint32_t w, uint32_t* out) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t t = __byte_perm(w, 0u, 0x4040u + 0x0101u * i);     // [b_i, 0, b_i, 0]
    if (!BF16) {
      // f16: 0x6400 = 1024 (ulp 1). low half: 0x6408 ^ n = 1032 + s; high half: 0x6480 ^ (n << 4) = 1152 + 16 s
      const uint32_t u = lop3_and_xor(t, 0x00f0000fu, 0x64806408u);
      const uint32_t mul_bits = 0x2C003C00u, add_bits = 0xD480E408u;    // (1, 1/16), (-1032, -72)
      const __half2 v = __hfma2(*reinterpret_cast<const __half2*>(&u), *reinterpret_cast<const __half2*>(&mul_bits),
                                *reinterpret_cast<const __half2*>(&add_bits));
      out[i] = *reinterpret_cast<const uint32_t*>(&v);
    } else {
      // bf16: 0x4300 = 128 (ulp 1, 7 mantissa bits): 0x4308 ^ n = 136 + s in both halves
      const uint32_t lo = lop3_and_xor(t, 0x0000000fu, 0x43084308u);
      const uint32_t u = lop3_and_xor(t >> 4, 0x000f0000u, lo);
      const __nv_bfloat162 c136 = __floats2bfloat162_rn(136.f, 136.f);
      const __nv_bfloat162 v = __hsub2(*reinterpret_cast<const __nv_bfloat162*>(&u), c136);
      out[i] = *reinterpret_cast<const uint32_t*>(&v);
    }
  }
}
template <bool BF16>
__device__ __forceinline__ void dequant32(const uint4& p, uint32_t (&r)[16]) {
  dequant8<BF16>(p.x, r + 0);
  dequant8<BF16>(p.y, r + 4);
  dequant8<BF16>(p.z, r + 8);
  dequant8<BF16>(p.w, r + 12);
}

struct Params {
  const void* sA;        // [M]   (I8 mode)
  const void* sW;        // [N]
  const void* bias;      // [N] or null
  void* out;             // [M, N] f16/bf16 (null in acc mode)
  int32_t* acc_out;      // [M, N] raw int32 accumulators (I8 debug/parity entry point)
  uint32_t* ws_part;     // [2 * G][np * 128] 32-bit partial tiles
  unsigned* counters;    // [G] (indexed by the first CTA of a shared tile) zero-initialised, self-cleaning
  int cs;                // > 1: the cs CTAs of a tile form a thread-block cluster and reduce through distributed shared memory
  uint32_t* split_out;   // non-null: deferred reduction -- raw partial accumulators [split][M][N], summed by the consumer kernel
  int M, N, K;
  int np;                // tokens per tile: multiple of 16, <= 128
  int tiles_m, KB;       // token tiles; k units per tile
  int U, G;              // total units; grid size
  int stages;
  int dcols, ta, tmem_cols;   // accumulator buffer width, operand slots, TMEM allocation (columns)
  int dbufs;                  // accumulator buffers: 2 (the epilogue of tile i overlaps the main loop of tile i + 1), or 1 for 256-token tiles
  int out_bf16;
  int debug_exit;        // debug: 1 = exit after setup, 2 = no TMEM alloc + exit after setup, 3 = exit at entry
  int trace_cta;         // debug: second CTA whose marks are recorded
  int trigger;           // when to release the dependent grid: 0 after setup, 1 when the weight stream has been issued, 2 at the end
  unsigned long long* trace;   // debug timeline (null in production): [2 CTAs][4 roles][256] globaltimer ns
};

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;\n" : "=l"(t));
  return t;
}
__device__ __forceinline__ int cta_start(int c, int U, int G) { return (int)(((long long)c * U) / G); }
__device__ __forceinline__ int cta_of_unit(int u, int U, int G) {
  int c = (int)(((long long)u * G) / U);
  while (c + 1 < G && cta_start(c + 1, U, G) <= u) ++c;
  while (c > 0 && cta_start(c, U, G) > u) --c;
  return c;
}

template <int MODE>
__device__ __forceinline__ void finalize_store(const Params& p, uint32_t raw, int tok_g, int ch, float sa, float sw, float bs) {
  const size_t o = (size_t)tok_g * p.N + ch;
  if (MODE == MODE_I8 || MODE == MODE_A8) {
    const int32_t v = ((int32_t)raw) >> (MODE == MODE_A8 ? 4 : 8);   // weights (and W4A4 activations) carry a factor 16
    if (p.acc_out) { p.acc_out[o] = v; return; }
    float f = __fmul_rn(__fmul_rn(__int2float_rn(v), sa), sw);
    if (p.bias) f = __fadd_rn(f, bs);
    if (p.out_bf16) ((__nv_bfloat16*)p.out)[o] = __float2bfloat16_rn(f);
    else ((__half*)p.out)[o] = __float2half_rn(f);
  } else {
    float f = __fmul_rn(__uint_as_float(raw), sw);
    if (p.bias) f = __fadd_rn(f, bs);
    if (MODE == MODE_BF16) ((__nv_bfloat16*)p.out)[o] = __float2bfloat16_rn(f);
    else ((__half*)p.out)[o] = __float2half_rn(f);
  }
}

// finalize_store for the 16 tokens of one tcgen05.ld group (tokens tok0 + j, j < nt), with everything that does not depend
// on the token -- output kind, element type, bias -- decided once outside the loop. Same arithmetic, same rounding.
template <int MODE>
__device__ __forceinline__ void finalize16(const Params& p, const uint32_t (&r)[16], int nt, int tok0, int ch, const float* sa,
                                           float sw, float bs) {
  constexpr bool INT = MODE == MODE_I8 || MODE == MODE_A8;
  constexpr int SH = MODE == MODE_A8 ? 4 : 8;
  const size_t o0 = (size_t)tok0 * p.N + ch;
  const size_t ld = (size_t)p.N;
  if (INT && p.acc_out) {
    int32_t* dst = p.acc_out + o0;
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (j < nt) dst[j * ld] = ((int32_t)r[j]) >> SH;
    return;
  }
  const bool has_bias = p.bias != nullptr;
  auto value = [&](int j) -> float {
    float f = INT ? __fmul_rn(__fmul_rn(__int2float_rn(((int32_t)r[j]) >> SH), sa[j]), sw) : __fmul_rn(__uint_as_float(r[j]), sw);
    if (has_bias) f = __fadd_rn(f, bs);
    return f;
  };
  const bool bf = INT ? (p.out_bf16 != 0) : (MODE == MODE_BF16);
  if (bf) {
    __nv_bfloat16* dst = (__nv_bfloat16*)p.out + o0;
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (j < nt) dst[j * ld] = __float2bfloat16_rn(value(j));
  } else {
    __half* dst = (__half*)p.out + o0;
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (j < nt) dst[j * ld] = __float2half_rn(value(j));
  }
}

// ---- the kernel --------------------------------------------------------------------------------
// Shared-memory map (dynamic, 1024-aligned): [stages][16 KB packed W] | [stages][act stage] |
// (I8 only) [ta][np * 128 B] int8 token tiles. act stage = np*128 B (I8: packed int4) or np*512 B
// (F16: 4 K-major SW128 atoms of 64 k).
// Tensor-memory map: D0 at column 0, D1 at dcols, then `ta` operand slots of 32 columns; one slot is what
// 4 UMMAs consume (I8: 128 k, F16: 64 k). Decode tiles (np <= 64) take 256 columns and ~100 KB of shared
// memory so that two CTAs -- the tail of one GEMM and the head of the next -- share an SM.
struct Ring {                         // (index, phase) of an mbarrier ring, advanced without divisions
  int idx, n; uint32_t phase;
  __device__ __forceinline__ Ring(int n_, int start = 0) : idx(start), n(n_), phase(0) {}
  __device__ __forceinline__ void next() { if (++idx == n) { idx = 0; phase ^= 1u; } }
  __device__ __forceinline__ void advance(int k) { for (int i = 0; i < k; ++i) next(); }
};

#ifdef QSPEC_GEMM_PROF
__device__ __forceinline__ unsigned long long gtimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;\n" : "=l"(t));
  return t;
}
// absolute %globaltimer marks of CTA 0 (id < 16) and of the last CTA (16 + id)
#define MARK(id) do { if (p.trace && (blockIdx.x == 0 || blockIdx.x == p.trace_cta)) p.trace[2048 + 64 + (blockIdx.x == 0 ? 0 : 16) + (id)] = gtimer_ns(); } while (0)
#define MARK_CTA(base) do { if (p.trace) p.trace[(base) + blockIdx.x] = gtimer_ns(); } while (0)
#define PROF_DECL(N) long long pc_[N] = {}; long long tprev_ = clock64()
#define PROF(k) do { const long long tn_ = clock64(); pc_[k] += tn_ - tprev_; tprev_ = tn_; } while (0)
#define PROF_DUMP(cond, off, N) do { if (p.trace && (cond)) for (int k_ = 0; k_ < (N); ++k_) p.trace[(off) + k_] = (unsigned long long)pc_[k_]; } while (0)
#else
#define MARK(id)
#define MARK_CTA(base)
#define PROF_DECL(N)
#define PROF(k)
#define PROF_DUMP(cond, off, N)
#endif

// PAIR = true: the kernel is launched as clusters of two CTAs (one TPC) that run as ONE tensor-core unit (cta_group::2).
// Both CTAs walk the same (pair-tile, k-unit) range; CTA r of the pair owns weight tile 2 * tile_n + r (its own TMA stream,
// its own converters, its own operand slots and accumulators in ITS tensor memory) and HALF of the token rows of the
// activation tile in its shared memory. The leader (rank 0) issues every tcgen05.mma for both; its a_full / d_empty
// barriers collect the arrivals of both CTAs (the follower arrives remotely), its commits are multicast to both.
// p.U / p.G then count pair-units and pairs; blockIdx.x >> 1 is the pair.
template <int MODE, bool PAIR>
__global__ void __launch_bounds__(THREADS, 1)
w4_stream_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_act, const Params p) {
  constexpr bool I8 = MODE == MODE_I8 || MODE == MODE_A8; // integer accumulate, per-token activation scale in the epilogue
  constexpr bool A8 = MODE == MODE_A8;                    // int8 activations, used as TMA delivers them (like F16)
  constexpr bool ACONV = MODE == MODE_I8;                 // packed int4 activations, widened by the converters into sm_b
  constexpr int SPU = I8 ? 2 : 4;                         // TMEM operand slots per 256-k unit
  constexpr int ATOMS = A8 ? 2 : 4, ATOM_ELEMS = A8 ? 128 : 64;   // 128-byte K-major atoms of the TMA'd activation stage
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t w_full[MAX_STAGES], w_empty[MAX_STAGES];
  __shared__ __align__(8) uint64_t a_full[MAX_TA], a_empty[MAX_TA];
  __shared__ __align__(8) uint64_t d_full[2], d_empty[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ unsigned s_is_last[2];
  __shared__ __align__(8) uint64_t red_full;             // cluster reduction: every pushing thread of the other CTAs arrives here
  __shared__ float s_sa[128];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) { MARK(0); MARK_CTA(512); }
  if (p.debug_exit == 3) return;
  const int S = p.stages, np = p.np, KB = p.KB, TA = p.ta;
  const uint32_t prank = PAIR ? cluster_ctarank() : 0u;       // rank in the CTA pair; 0 = leader (issues the MMAs)
  const int cta = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;   // position in the (pair-)unit space
  const int npl = PAIR ? np >> 1 : np;                         // token rows of the tile that live in THIS CTA's shared memory
  auto ptile = [&](int tn) { return PAIR ? 2 * tn + (int)prank : tn; };   // weight tile of this CTA in (pair-)tile tn
  auto phys = [&](int c) { return PAIR ? 2 * c + (int)prank : c; };       // CTA of pair c with this CTA's rank
  const int u0 = cta_start(cta, p.U, p.G), u1 = cta_start(cta + 1, p.U, p.G);
  const int n_units = u1 - u0;
  const int act_stage = ACONV ? npl * 128 : A8 ? npl * 256 : npl * 512;
  const uint32_t tmem_a_col = (uint32_t)(p.dbufs * p.dcols);
  const bool dbl = p.dbufs == 2;                                   // double-buffered accumulator

  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* sm_w = base;
  unsigned char* sm_act = sm_w + (size_t)S * W_STAGE;
  unsigned char* sm_b = sm_act + (size_t)S * act_stage;   // I8 only
  unsigned char* sm_recv = sm_b + (ACONV ? (size_t)TA * npl * 128 : 0);   // cluster reduction: [cs-1][np][128] partial sums pushed by the peers

  const uint32_t stage_tx = (uint32_t)(W_STAGE + act_stage);
  int pt = u0 / KB, kb = u0 - pt * KB;                // (producer warp) tile / k-unit of the next unit to load
  int tile_n = pt / p.tiles_m, tile_m = pt - tile_n * p.tiles_m;
  auto step_unit = [&]() {
    if (++kb == KB) { kb = 0; if (++tile_m == p.tiles_m) { tile_m = 0; ++tile_n; } }
  };
  auto load_w = [&](int s) {
    if (elect_one()) {
      mbar_expect_tx(smem_u32(&w_full[s]), stage_tx);
      tma_load_2d(smem_u32(sm_w + (size_t)s * W_STAGE), &map_w, kb * UB, ptile(tile_n) * TM, smem_u32(&w_full[s]), (p.debug_exit & 16) ? POLICY_EVICT_NORMAL : POLICY_EVICT_FIRST);
    }
    __syncwarp();
  };
  auto load_act = [&](int s, int kb_, int tile_m_) {
    const uint32_t dst = smem_u32(sm_act + (size_t)s * act_stage);
    const int row0 = tile_m_ * np + (int)prank * npl;       // PAIR: this CTA's half of the token tile
    if (elect_one()) {
      if (ACONV) {
        tma_load_2d(dst, &map_act, kb_ * UB, row0, smem_u32(&w_full[s]), (p.debug_exit & 16) ? POLICY_EVICT_NORMAL : POLICY_EVICT_LAST);
      } else {
#pragma unroll
        for (int a = 0; a < ATOMS; ++a)
          tma_load_2d(dst + a * npl * 128, &map_act, kb_ * UK + a * ATOM_ELEMS, row0, smem_u32(&w_full[s]), (p.debug_exit & 16) ? POLICY_EVICT_NORMAL : POLICY_EVICT_LAST);
      }
    }
    __syncwarp();
  };
  // The producer warp initialises the barriers and puts the first ring of weight tiles in flight BEFORE the CTA-wide
  // setup barrier: tensor-memory allocation and the other warps' setup overlap the first HBM round trip.
  const int pre = n_units < S ? n_units : S;
  const int kb0 = kb, tile_m0 = tile_m;
  if (warp == WARP_TMA) {
    if (elect_one()) {
      tma_prefetch_desc(&map_w);
      tma_prefetch_desc(&map_act);
      for (int s = 0; s < S; ++s) {
        mbar_init(smem_u32(&w_full[s]), 1);
        mbar_init(smem_u32(&w_empty[s]), NCONV / 2 + (ACONV ? 0 : 1));
      }
      for (int s = 0; s < TA; ++s) {
        mbar_init(smem_u32(&a_full[s]), PAIR ? NCONV : NCONV / 2);     // PAIR (leader): the converters of both CTAs
        mbar_init(smem_u32(&a_empty[s]), 1);
      }
      for (int s = 0; s < 2; ++s) {
        mbar_init(smem_u32(&d_full[s]), 1);
        mbar_init(smem_u32(&d_empty[s]), PAIR ? 2 * NEPI : NEPI);      // PAIR (leader): the epilogue warps of both CTAs
      }
      mbar_init(smem_u32(&red_full), (uint32_t)(max(p.cs, 2) - 1) * NEPI * 32);
      fence_mbar_init();
    }
    __syncwarp();
    if (p.debug_exit == 0 || p.debug_exit >= 16)
      for (int i = 0; i < pre; ++i) { load_w(i); step_unit(); }   // weights depend on no earlier kernel
    if (lane == 0) MARK(2);
  }
  if (warp == WARP_MMA && p.debug_exit != 2) tmem_alloc<PAIR>(smem_u32(&tmem_base_smem), (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();      // both CTAs' barriers are initialised and both allocations are done before any remote arrive / MMA
  tc_fence_after();
  if (!PAIR && p.cs > 1) asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");   // red_full is initialised, this CTA runs
  const uint32_t tmem = tmem_base_smem;
  if (p.debug_exit == 2) return;
  if (p.debug_exit == 1) {
    if (PAIR) cluster_sync_all();
    if (warp == WARP_MMA) tmem_dealloc<PAIR>(tmem, (uint32_t)p.tmem_cols);
    return;
  }
  if (tid == 0) MARK(1);
  if (p.trigger == 0) pdl_launch_dependents();   // the next kernel may start its own prologue / weight stream now

  if (warp == WARP_TMA) {
    // =========================== TMA producer (whole warp converged, one elected lane issues) ==========
    // (the first `pre` weight tiles were put in flight during setup) activations need the preceding kernel's output
    pdl_wait();
    if (lane == 0) MARK(3);
    {
      int kb_ = kb0, tm_ = tile_m0;
      for (int i = 0; i < pre; ++i) {
        load_act(i, kb_, tm_);
        if (++kb_ == KB) { kb_ = 0; if (++tm_ == p.tiles_m) tm_ = 0; }
      }
    }
    Ring st(S);                                          // stage of unit `pre` is 0 again, one phase later
    for (int i = pre; i < n_units; ++i) {
      mbar_wait(smem_u32(&w_empty[st.idx]), st.phase);   // phase p of w_empty completes when use p of the stage is consumed
      load_w(st.idx);
      load_act(st.idx, kb, tile_m);
      step_unit();
      st.next();
    }
    if (p.trigger == 1) pdl_launch_dependents();         // every weight byte of this CTA is on its way
  } else if (warp == WARP_MMA) {
    // =========================== MMA issuer: one elected lane runs the whole loop ===========================
    // (inside an elect.sync region the compiler keeps descriptors / addresses in uniform registers)
    // PAIR: only the leader issues; the follower's MMA warp has nothing to do.
    if ((!PAIR || prank == 0) && elect_one()) {
      const uint32_t idesc = make_idesc<MODE, PAIR>(np);
      PROF_DECL(8);
      Ring st(S), sl(TA);
      int f = 0;
      for (int u = u0; u < u1;) {
        const int t = u / KB, ka = u - t * KB;
        const int kb_end = min(KB, ka + (u1 - u));
        const int db = dbl ? (f & 1) : 0;
        const uint32_t d_tmem = tmem + (db ? (uint32_t)p.dcols : 0u);
        PROF(3);
        mbar_wait(smem_u32(&d_empty[db]), (((uint32_t)(dbl ? f >> 1 : f)) & 1u) ^ 1u);
        PROF(0);
        for (int k = ka; k < kb_end; ++k) {
          if (!ACONV) mbar_wait(smem_u32(&w_full[st.idx]), st.phase);
#pragma unroll
          for (int sub = 0; sub < SPU; ++sub) {
            PROF(3);
            mbar_wait(smem_u32(&a_full[sl.idx]), sl.phase);
            PROF(1);
            tc_fence_after();
            const uint32_t a_tmem = tmem + tmem_a_col + sl.idx * 32;
            // I8: slot = half unit, tokens from the converted int8 ring; F16: slot = one 64-k atom of x
            // (PAIR: each CTA holds npl = np / 2 token rows at this offset; the instruction reads both halves)
            const uint32_t b0 = ACONV ? smem_u32(sm_b + (size_t)sl.idx * (npl * 128))
                                   : smem_u32(sm_act + (size_t)st.idx * act_stage + sub * (npl * 128));
            const uint64_t bdesc = make_desc_sw128(b0);
            const uint32_t first = (k > ka || sub > 0) ? 1u : 0u;
            umma_ts<MODE, PAIR>(d_tmem, a_tmem, bdesc, idesc, first);
            umma_ts<MODE, PAIR>(d_tmem, a_tmem + 8, bdesc + 2, idesc, 1u);     // +32 B along K = +2 in the (>>4) start field
            umma_ts<MODE, PAIR>(d_tmem, a_tmem + 16, bdesc + 4, idesc, 1u);
            umma_ts<MODE, PAIR>(d_tmem, a_tmem + 24, bdesc + 6, idesc, 1u);
            PROF(4);
            umma_commit<PAIR>(smem_u32(&a_empty[sl.idx]));
            if (!ACONV && sub == SPU - 1) umma_commit<PAIR>(smem_u32(&w_empty[st.idx]));
            if (sub == SPU - 1 && k == kb_end - 1) umma_commit<PAIR>(smem_u32(&d_full[db]));
            PROF(5);
            sl.next();
          }
          st.next();
        }
        ++f;
        u += kb_end - ka;
      }
      PROF_DUMP(blockIdx.x == 0, 2048 + 16, 8);
    }
    __syncwarp();
  } else if (warp < NCONV) {
    // =========================== converters ===========================
    // Two groups of four warps (one warp per TMEM lane quarter). Group g owns units i = g, g + 2, ...: a thread
    // converts its whole 128-byte row of the unit, so the per-unit synchronisation cost (two barrier waits, the
    // proxy fence, the tcgen05.st drain and two arrives) is paid once per 16 KB instead of once per 8 KB, and the
    // two groups work on consecutive units concurrently.
    const int q = warp & 3, g = warp >> 2;
    const int row = q * 32 + lane;
    // "operand slot filled": the barrier the MMA issuer waits on (PAIR: the leader's, for both CTAs)
    auto arrive_a_full = [&](int idx) {
      if (PAIR && prank != 0) mbar_arrive_cluster(smem_u32(&a_full[idx]), 0);
      else mbar_arrive(smem_u32(&a_full[idx]));
    };
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + tmem_a_col;
    const int gtid = q * 32 + lane;                     // 0..127 within this group
    PROF_DECL(8);
    Ring st(S), sl(TA);
    st.advance(g);
    sl.advance(SPU * g);                                 // first slot of unit g
    const int swz = row & 7;
    for (int i = g; i < n_units; i += 2) {
      PROF(7);
      mbar_wait(smem_u32(&w_full[st.idx]), st.phase);
      if (i == 0 && tid == 0) MARK(4);
      PROF(0);
      const unsigned char* wrow = sm_w + (size_t)st.idx * W_STAGE + row * 128;
      uint4 pk[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) pk[c] = *reinterpret_cast<const uint4*>(wrow + ((c ^ swz) << 4));
      if (A8) {
        // W4A8: weights widened in natural k order into the two half-unit slots; the int8 activations stay where TMA put them
        Ring s1 = sl;
        s1.next();
        mbar_wait(smem_u32(&a_empty[sl.idx]), sl.phase ^ 1u);
        mbar_wait(smem_u32(&a_empty[s1.idx]), s1.phase ^ 1u);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          uint32_t r[8];
          widen_i8_nat(pk[c], r);
          tmem_st8(lane_base + (c < 4 ? sl.idx : s1.idx) * 32 + (c & 3) * 8, r);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          arrive_a_full(sl.idx);
          arrive_a_full(s1.idx);
          mbar_arrive(smem_u32(&w_empty[st.idx]));     // the weight half of the stage; the MMA commit releases the activation half
        }
      } else if (ACONV) {
        // this thread's first packed-activation chunk is fetched together with the weights
        const unsigned char* act = sm_act + (size_t)st.idx * act_stage;
        // Decode tiles (<= 16 token rows per CTA): one chunk per thread. Larger tiles: up to AB chunks per thread (64 tokens per
        // group), loaded together so that their shared-memory latencies overlap -- one chunk per loop iteration cost ~200 cycles
        // per iteration, 830 cycles per unit at 64 tokens (profiles/r02c_gemm_trace.txt). Kept apart: the batched form costs
        // the decode-sized kernels ~10 % (16 more live registers across the weight widening; r02o bench, configs[1]).
        constexpr int AB = 4;
        const bool a_small = npl <= 16;
        const int ac = gtid & 7;                         // a thread's chunks: tokens (gtid >> 3) + 16 j, the same 16-byte column
        uint4 akb[AB];
        akb[0] = make_uint4(0, 0, 0, 0);
        if ((gtid >> 3) < npl) akb[0] = *reinterpret_cast<const uint4*>(act + (gtid >> 3) * 128 + ((ac ^ ((gtid >> 3) & 7)) << 4));
        if (!a_small) {
#pragma unroll
          for (int j = 1; j < AB; ++j) {
            const int tok = (gtid >> 3) + 16 * j;
            akb[j] = make_uint4(0, 0, 0, 0);
            if (tok < npl) akb[j] = *reinterpret_cast<const uint4*>(act + tok * 128 + ((ac ^ (tok & 7)) << 4));
          }
        }
        Ring s1 = sl;                                    // second half-unit slot
        s1.next();
        mbar_wait(smem_u32(&a_empty[sl.idx]), sl.phase ^ 1u);
        mbar_wait(smem_u32(&a_empty[s1.idx]), s1.phase ^ 1u);
        tc_fence_after();
        PROF(1);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          uint32_t r[8];
          widen_i8(pk[c], r);
          tmem_st8(lane_base + (c < 4 ? sl.idx : s1.idx) * 32 + (c & 3) * 8, r);
        }
        PROF(2);
        // packed activations of the unit (8 chunks per token) -> two int8 atoms [npl][128 B], SW128 (PAIR: this CTA's half of the tokens)
        unsigned char* bt0 = sm_b + (size_t)sl.idx * (npl * 128);
        unsigned char* bt1 = sm_b + (size_t)s1.idx * (npl * 128);
        {
          const int cc = ac & 3;
          unsigned char* bt = ac < 4 ? bt0 : bt1;
          if (a_small) {
            const int tok = gtid >> 3;
            if (tok < npl) {
              uint32_t r[8];
              widen_i8(akb[0], r);
              unsigned char* brow = bt + tok * 128;
              *reinterpret_cast<uint4*>(brow + (((2 * cc) ^ (tok & 7)) << 4)) = make_uint4(r[0], r[1], r[2], r[3]);
              *reinterpret_cast<uint4*>(brow + (((2 * cc + 1) ^ (tok & 7)) << 4)) = make_uint4(r[4], r[5], r[6], r[7]);
            }
          } else {
            for (int tb = 0; tb < npl; tb += 16 * AB) {
              if (tb) {
#pragma unroll
                for (int j = 0; j < AB; ++j) {
                  const int tok = tb + (gtid >> 3) + 16 * j;
                  if (tok < npl) akb[j] = *reinterpret_cast<const uint4*>(act + tok * 128 + ((ac ^ (tok & 7)) << 4));
                }
              }
#pragma unroll
              for (int j = 0; j < AB; ++j) {
                const int tok = tb + (gtid >> 3) + 16 * j;
                if (tok < npl) {
                  uint32_t r[8];
                  widen_i8(akb[j], r);
                  unsigned char* brow = bt + tok * 128;
                  *reinterpret_cast<uint4*>(brow + (((2 * cc) ^ (tok & 7)) << 4)) = make_uint4(r[0], r[1], r[2], r[3]);
                  *reinterpret_cast<uint4*>(brow + (((2 * cc + 1) ^ (tok & 7)) << 4)) = make_uint4(r[4], r[5], r[6], r[7]);
                }
              }
            }
          }
        }
        PROF(3);
        fence_proxy_async_smem();
        PROF(4);
        tmem_wait_st();
        PROF(5);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          arrive_a_full(sl.idx);
          arrive_a_full(s1.idx);
          mbar_arrive(smem_u32(&w_empty[st.idx]));
        }
        PROF(6);
      } else {
        Ring sq = sl;
#pragma unroll
        for (int sub = 0; sub < 4; ++sub) {
          mbar_wait(smem_u32(&a_empty[sq.idx]), sq.phase ^ 1u);
          tc_fence_after();
          PROF(1);
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            uint32_t r[16];
            dequant32<MODE == MODE_BF16>(pk[2 * sub + cc], r);
            tmem_st16(lane_base + sq.idx * 32 + cc * 16, r);
          }
          PROF(2);
          tmem_wait_st();
          PROF(5);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            arrive_a_full(sq.idx);
            if (sub == 3) mbar_arrive(smem_u32(&w_empty[st.idx]));
          }
          sq.next();
          PROF(6);
        }
      }
      sl.advance(2 * SPU);                               // this group's next unit is two units further
      st.advance(2);
    }
    if (tid == 0) MARK(5);
    PROF_DUMP(blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 4), 2048 + (warp == 0 ? 0 : 8), 8);
  } else {
    // =========================== epilogue (warps NCONV..NCONV+3) ===========================
    const int q = warp - NCONV;                          // == warp % 4: this warp's TMEM lane quarter
    const int etid = q * 32 + lane;
    const int chl = etid;
    pdl_wait();                                          // out / workspace may still be in use upstream
    int f = 0;
    auto load_scale = [&](int unit, float& sw_, float& bs_) {      // per-channel scale / bias of the tile holding `unit`
      sw_ = 0.f; bs_ = 0.f;
      const int ch_ = ptile((unit / KB) / p.tiles_m) * TM + chl;
      if (unit < u1 && ch_ < p.N && p.out) {
        sw_ = p.out_bf16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p.sW + ch_)) : __half2float(__ldg((const __half*)p.sW + ch_));
        if (p.bias) bs_ = p.out_bf16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p.bias + ch_)) : __half2float(__ldg((const __half*)p.bias + ch_));
      }
    };
    int npf = 0, pf_t[2] = {0, 0}, pf_cf[2] = {0, 0}, pf_cl[2] = {0, 0};   // this CTA's tiles that are shared with other CTAs
    // ---- stream-K tiles shared with other CTAs (at most two per CTA: its first and its last tile). The protocol is
    // runs right after the tile's accumulator has been drained (overlapping the main loop unless it is the CTA's last
    // tile): a counter atomic, then the last arriver of the tile reduces the partial tiles in ascending CTA order.
    // release: partial stores (all threads) -> CTA barrier -> acq_rel atomic at gpu scope (cumulative);
    // acquire: the same atomic -> CTA barrier -> ld.cg of the other CTAs' partial tiles.
    auto resolve_shared = [&](int i0, int i1) {
      named_bar_sync(1, NEPI * 32);
      if (etid >= i0 && etid < i1) {
        unsigned prev;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;\n" : "=r"(prev) : "l"(p.counters + phys(pf_cf[etid])) : "memory");
        s_is_last[etid] = (prev == (unsigned)(pf_cl[etid] - pf_cf[etid])) ? 1u : 0u;
      }
      named_bar_sync(1, NEPI * 32);
      if (etid == 0) MARK(8);
      for (int i = i0; i < i1; ++i) {
        if (!s_is_last[i]) continue;
        const int t = pf_t[i], c_first = pf_cf[i], c_last = pf_cl[i];
        const int tile_n = t / p.tiles_m, tile_m = t % p.tiles_m;
        const int n0 = ptile(tile_n) * TM, m0 = tile_m * np;
        const int m_here = min(p.M - m0, np);
        // (c_first .. c_last are positions in the (pair-)unit space; phys() maps them to the CTAs that hold THIS weight tile)
        // One work item = (token, 4 consecutive channels): 16-byte loads from every contributor's partial tile, all
        // independent (one L2 round trip for the whole reduction), summed in ascending CTA order. Only the first
        // contributor can have started before this tile (its slot 1); every later one starts inside it (slot 0).
        const size_t slot_words = (size_t)(np * TM);
        const uint32_t* first_slot = p.ws_part + (size_t)(2 * phys(c_first) + (cta_start(c_first, p.U, p.G) < t * KB ? 1 : 0)) * slot_words;
        // A thread's items are tokens tok_base + 4 j of ONE channel quad (128 threads = 4 tokens x 32 quads): the per-channel
        // scales are loaded once, and RB items are reduced together so that RB x 2 partial-tile loads are in flight per L2
        // round trip. (One item per iteration made the reduction a chain of m_here / 4 dependent round trips: ~8 us for a
        // 64-token tile, ~16 us for 128 tokens -- profiles/r02m_gsweep_tp.md -- against ~2.5 us at decode sizes.)
        constexpr int RB = 4;
        const int quad = etid & 31, tok_base = etid >> 5;
        const int chq = n0 + 4 * quad;
        float swe[4] = {0.f, 0.f, 0.f, 0.f}, bse[4] = {0.f, 0.f, 0.f, 0.f};
        if (p.out && chq < p.N) {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            swe[e] = p.out_bf16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p.sW + chq + e)) : __half2float(__ldg((const __half*)p.sW + chq + e));
            if (p.bias) bse[e] = p.out_bf16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p.bias + chq + e)) : __half2float(__ldg((const __half*)p.bias + chq + e));
          }
        }
        if (m_here <= 4) {
          // decode (one item per thread): four CONTRIBUTORS per round trip instead (a 4-token tile shared by 4 CTAs is one round)
          const int tok = tok_base;
          if (tok < m_here) {
            const size_t off = (size_t)tok * TM + 4 * quad;
            float sa = 0.f;
            if (I8 && p.out && chq < p.N)
              sa = p.out_bf16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p.sA + m0 + tok)) : __half2float(__ldg((const __half*)p.sA + m0 + tok));
            uint4 acc = __ldcg(reinterpret_cast<const uint4*>(first_slot + off));
            for (int c = c_first + 1; c <= c_last; c += 4) {
              uint4 v[4];
#pragma unroll
              for (int e = 0; e < 4; ++e)
                v[e] = (c + e <= c_last) ? __ldcg(reinterpret_cast<const uint4*>(p.ws_part + (size_t)(2 * phys(c + e)) * slot_words + off))
                                         : make_uint4(0, 0, 0, 0);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if (c + e <= c_last) {
                  if (I8) {
                    acc.x += v[e].x; acc.y += v[e].y; acc.z += v[e].z; acc.w += v[e].w;
                  } else {
                    acc.x = __float_as_uint(__fadd_rn(__uint_as_float(acc.x), __uint_as_float(v[e].x)));
                    acc.y = __float_as_uint(__fadd_rn(__uint_as_float(acc.y), __uint_as_float(v[e].y)));
                    acc.z = __float_as_uint(__fadd_rn(__uint_as_float(acc.z), __uint_as_float(v[e].z)));
                    acc.w = __float_as_uint(__fadd_rn(__uint_as_float(acc.w), __uint_as_float(v[e].w)));
                  }
                }
              }
            }
            if (chq < p.N) {
              const uint32_t a4[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) finalize_store<MODE>(p, a4[e], m0 + tok, chq + e, sa, swe[e], bse[e]);
            }
          }
        } else {
          for (int tok0 = tok_base; tok0 < m_here; tok0 += 4 * RB) {
            uint4 acc4[RB];
            float sa[RB];
  #pragma unroll
            for (int k = 0; k < RB; ++k) {
              const int tok = tok0 + 4 * k;
              const bool ok = tok < m_here;
              acc4[k] = ok ? __ldcg(reinterpret_cast<const uint4*>(first_slot + (size_t)tok * TM + 4 * quad)) : make_uint4(0, 0, 0, 0);
              sa[k] = 0.f;
              if (I8 && ok && p.out && chq < p.N)
                sa[k] = p.out_bf16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p.sA + m0 + tok)) : __half2float(__ldg((const __half*)p.sA + m0 + tok));
            }
            for (int c = c_first + 1; c <= c_last; c += 2) {
              uint4 v[2][RB];
  #pragma unroll
              for (int e = 0; e < 2; ++e) {
                const uint32_t* slot = p.ws_part + (size_t)(2 * phys(min(c + e, c_last))) * slot_words + 4 * quad;
  #pragma unroll
                for (int k = 0; k < RB; ++k) {
                  const int tok = tok0 + 4 * k;
                  v[e][k] = (c + e <= c_last && tok < m_here) ? __ldcg(reinterpret_cast<const uint4*>(slot + (size_t)tok * TM)) : make_uint4(0, 0, 0, 0);
                }
              }
  #pragma unroll
              for (int e = 0; e < 2; ++e) {
                if (c + e <= c_last) {                       // ascending CTA order, whatever the batching
  #pragma unroll
                  for (int k = 0; k < RB; ++k) {
                    if (I8) {
                      acc4[k].x += v[e][k].x; acc4[k].y += v[e][k].y; acc4[k].z += v[e][k].z; acc4[k].w += v[e][k].w;
                    } else {
                      acc4[k].x = __float_as_uint(__fadd_rn(__uint_as_float(acc4[k].x), __uint_as_float(v[e][k].x)));
                      acc4[k].y = __float_as_uint(__fadd_rn(__uint_as_float(acc4[k].y), __uint_as_float(v[e][k].y)));
                      acc4[k].z = __float_as_uint(__fadd_rn(__uint_as_float(acc4[k].z), __uint_as_float(v[e][k].z)));
                      acc4[k].w = __float_as_uint(__fadd_rn(__uint_as_float(acc4[k].w), __uint_as_float(v[e][k].w)));
                    }
                  }
                }
              }
            }
            if (chq < p.N) {
  #pragma unroll
              for (int k = 0; k < RB; ++k) {
                const int tok = tok0 + 4 * k;
                if (tok < m_here) {
                  const uint32_t a4[4] = {acc4[k].x, acc4[k].y, acc4[k].z, acc4[k].w};
  #pragma unroll
                  for (int e = 0; e < 4; ++e) finalize_store<MODE>(p, a4[e], m0 + tok, chq + e, sa[k], swe[e], bse[e]);
                }
              }
            }
          }
        }
        if (etid == 0) p.counters[phys(c_first)] = 0u;   // self-cleaning
      }
      if (etid == 0) MARK(11);
    };
    float sw_next, bs_next;
    load_scale(u0, sw_next, bs_next);
    int sa_tile = -1;                                    // token tile whose activation scales sit in s_sa
    for (int u = u0; u < u1;) {
      const int t = u / KB, ka = u - t * KB;
      const int kb_end = min(KB, ka + (u1 - u));
      const int tile_n = t / p.tiles_m, tile_m = t % p.tiles_m;
      const int n0 = ptile(tile_n) * TM, m0 = tile_m * np;
      const int m_here = min(p.M - m0, np);
      const int ch = n0 + chl;
      const bool ch_ok = ch < p.N;
      const bool full = (ka == 0 && kb_end == KB);
      const int db = dbl ? (f & 1) : 0;
      // everything a tile needs from global memory is requested one tile ahead of its use
      const float sw = sw_next, bs = bs_next;
      load_scale(u + (kb_end - ka), sw_next, bs_next);
      if (I8 && p.out && !p.split_out && sa_tile != tile_m) {   // decode (one token tile): loaded once per kernel
        named_bar_sync(1, NEPI * 32);                    // previous tile's readers are done with s_sa
        if (etid < m_here)
          s_sa[etid] = p.out_bf16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p.sA + m0 + etid))
                                  : __half2float(__ldg((const __half*)p.sA + m0 + etid));
        named_bar_sync(1, NEPI * 32);
        sa_tile = tile_m;
      }
      const int c_first = full ? 0 : cta_of_unit(t * KB, p.U, p.G);
      const int c_last = full ? 0 : cta_of_unit(t * KB + KB - 1, p.U, p.G);
      uint32_t* my_slot = p.ws_part + (size_t)(2 * blockIdx.x + (u0 < t * KB ? 1 : 0)) * (size_t)(np * TM);
      mbar_wait(smem_u32(&d_full[db]), ((uint32_t)(dbl ? f >> 1 : f)) & 1u);
      tc_fence_after();
      if (etid == 0) MARK(6 + 3 * min(f, 2));
      const uint32_t d_addr = tmem + ((uint32_t)(q * 32) << 16) + (db ? (uint32_t)p.dcols : 0u);
      const int split_rank = full ? 0 : cta - c_first;     // deferred reduction: this CTA's slice of the tile's k range
      const bool clus = !PAIR && p.cs > 1;                 // the tile is shared by exactly this cluster; rank 0 owns it
      const uint32_t crank = clus ? (uint32_t)(blockIdx.x % p.cs) : 0u;
      uint32_t push_base = 0;
      if (clus) {
        if (crank != 0) {
          if (f == 0) asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");   // the owner's barrier is initialised
          push_base = mapa_shared(smem_u32(sm_recv), 0) + (uint32_t)((crank - 1) * np * TM + chl) * 4u;
        } else {
          // acquire at cluster scope: the peers' pushes are visible once their arrivals are
          uint32_t done = 0;
          while (!done)
            asm volatile("{\n.reg .pred pp;\nmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 pp, [%1], 0;\nselp.u32 %0, 1, 0, pp;\n}\n"
                         : "=r"(done) : "r"(smem_u32(&red_full)) : "memory");
        }
      }
      // What happens to the accumulators is chosen per TILE, outside the token loop. With the five-way choice inside it every
      // token carried ~330 instructions and one 16-token group spanned ~80 KB of code: the epilogue ran at instruction-fetch
      // speed (profiling build, gate_up W4A4: 3.4 us per 16-token group on an otherwise idle SM, 12-17 us per tile at M = 64).
      const int variant = (p.debug_exit & 64) ? -1 : p.split_out ? 0 : full ? 1 : !clus ? 2 : crank != 0 ? 3 : 4;
      for (int c0 = 0; c0 < m_here; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(d_addr + c0, r);                       // warp-collective: never skipped per thread
        const int nt = m_here - c0;                      // this group's tokens are c0 + jj, jj < nt
        if (variant == 0) {                              // deferred reduction: raw k-slice partial sums
          if (ch_ok) {
            uint32_t* dst = p.split_out + ((size_t)split_rank * p.M + m0 + c0) * p.N + ch;
            const size_t ld = (size_t)p.N;
#pragma unroll
            for (int jj = 0; jj < 16; ++jj)
              if (jj < nt) dst[jj * ld] = r[jj];
          }
        } else if (variant == 1) {                       // whole tile in this CTA: epilogue and store
          if (ch_ok) finalize16<MODE>(p, r, nt, m0 + c0, ch, s_sa + c0, sw, bs);
        } else if (variant == 2) {                       // stream-K partial tile through the workspace
          uint32_t* dst = my_slot + (size_t)c0 * TM + chl;
#pragma unroll
          for (int jj = 0; jj < 16; ++jj)
            if (jj < nt) dst[jj * TM] = r[jj];
        } else if (variant == 3) {                       // cluster reduction, non-owner: push to the owner's shared memory
#pragma unroll
          for (int jj = 0; jj < 16; ++jj)
            if (jj < nt) st_dsmem_u32(push_base + (uint32_t)((c0 + jj) * TM) * 4u, r[jj]);
        } else if (variant == 4) {                       // cluster reduction, owner
#pragma unroll
          for (int jj = 0; jj < 16; ++jj) {
            const int tok = c0 + jj;
            if (jj < nt) {
              // ascending rank = ascending CTA order: the same sum as the workspace reduction
              uint32_t acc = r[jj];
              const uint32_t* rv = reinterpret_cast<const uint32_t*>(sm_recv) + tok * TM + chl;
              for (int pr = 0; pr < p.cs - 1; ++pr) {
                const uint32_t v = rv[(size_t)pr * np * TM];
                acc = I8 ? acc + v : __float_as_uint(__fadd_rn(__uint_as_float(acc), __uint_as_float(v)));
              }
              if (ch_ok) finalize_store<MODE>(p, acc, m0 + tok, ch, I8 ? s_sa[tok] : 0.f, sw, bs);
            }
          }
        }
      }
      if (clus && crank != 0)
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(mapa_shared(smem_u32(&red_full), 0)) : "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {                                     // accumulator drained (PAIR: the leader's MMA issuer waits for both CTAs)
        if (PAIR && prank != 0) mbar_arrive_cluster(smem_u32(&d_empty[db]), 0);
        else mbar_arrive(smem_u32(&d_empty[db]));
      }
      if (etid == 0) MARK(7 + 3 * min(f, 2));
      if (!full && !clus && !p.split_out && !(p.debug_exit & 32)) {   // remember the shared tile; the protocol runs after the loop
        pf_t[npf] = t; pf_cf[npf] = c_first; pf_cl[npf] = c_last;
        ++npf;
        resolve_shared(npf - 1, npf);
      }
      ++f;
      u += kb_end - ka;
    }
  }

  tc_fence_before();
  __syncthreads();
  // PAIR: neither CTA may leave (or free its tensor memory) while the leader's instructions still read the follower's
  // operands or multicast arrivals into its barriers
  if (PAIR) cluster_sync_all();
  if (p.trigger != 0) pdl_launch_dependents();
  if (tid == 0) { MARK(15); MARK_CTA(1024); }
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc<PAIR>(tmem, (uint32_t)p.tmem_cols);
  }
}

// ---- host side -----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    return (EncodeTiledFn)f;
  }();
  return fn;
}

// 2-D row-major tensor [rows][row_bytes], tiles of box_rows x 128 bytes, SWIZZLE_128B. Encoded maps are
// cached by (pointer, geometry): weights and arena activations keep their addresses across calls.
struct MapKey {
  const void* ptr; uint64_t rows, row_bytes; uint32_t box_rows, elem;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && rows == o.rows && row_bytes == o.row_bytes && box_rows == o.box_rows && elem == o.elem;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = (size_t)k.ptr;
    h = h * 1000003u ^ k.rows; h = h * 1000003u ^ k.row_bytes; h = h * 1000003u ^ k.box_rows; h = h * 1000003u ^ k.elem;
    return h;
  }
};

static int get_map(const void* ptr, uint64_t rows, uint64_t row_bytes, uint32_t box_rows, uint32_t elem_bytes,
                   CUtensorMap* out) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  const MapKey key{ptr, rows, row_bytes, box_rows, elem_bytes};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return QSPEC_OK; }
  }
  EncodeTiledFn fn = encode_fn();
  QSPEC_REQUIRE(fn, QSPEC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[2] = {row_bytes / elem_bytes, rows};
  const cuuint64_t strides[1] = {row_bytes};
  const cuuint32_t box[2] = {128u / elem_bytes, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(out, elem_bytes == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16, 2,
                        const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  QSPEC_REQUIRE(r == CUDA_SUCCESS, QSPEC_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) ptr=%p rows=%llu row_bytes=%llu box_rows=%u",
                (int)r, ptr, (unsigned long long)rows, (unsigned long long)row_bytes, box_rows);
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 8192) cache.clear();
  cache.emplace(key, *out);
  return QSPEC_OK;
}

// The counters live in a fixed region at the start of the workspace (one per CTA, whatever the shape), so that
// they stay zero across calls of different geometry; the partial tiles behind them are never assumed clean.
constexpr size_t COUNTER_BYTES = 4096;
static unsigned long long* g_trace = nullptr;
static unsigned g_trace_seq = 0;     // debug: launch i records into region i % 8 of the trace buffer (4096 words each)

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

// pair = 1: CTA pairs (cta_group::2). tiles_n, U and G then count PAIR-tiles (two adjacent 128-channel tiles), pair-units
// (2 x 16 KB of weights) and pairs; the launch has 2 G CTAs in clusters of two.
struct Plan { int np, tiles_m, tiles_n, KB, U, G, cs, split, stages, dcols, dbufs, ta, tmem_cols, pair; size_t smem, slot_bytes; };

// QSPEC_GEMM_PAIR: 1 (default) = CTA pairs whenever the shape allows it, 0 = never (the round-1 kernel; kept for A/B runs).
// QSPEC_GEMM_PAIR_MIN_M: smallest token count that uses pairs (default below). Read at every call: A/B runs in one process.
static bool pair_wanted(int M, int N) {
  // measured (profiles/r02d_pair_ab.md): pairs win from one full 128-token tile on (M = 256: -9 .. -20 %), and lose 10-35 %
  // at decode sizes, where the cluster launch and the remote hand-overs add to an already latency-bound kernel
  return env_int("QSPEC_GEMM_PAIR", 1) != 0 && N % (2 * TM) == 0 && M >= env_int("QSPEC_GEMM_PAIR_MIN_M", 128);
}

static Plan make_plan(int mode, int M, int N, int K, bool deferred = false) {
  Plan pl;
  const int mp = ((M + 15) / 16) * 16;
  pl.np = mp < 128 ? mp : 128;
  pl.pair = pair_wanted(M, N) ? 1 : 0;
  // 256-token tiles (CTA pairs, W4A16 only): ONE pass over the weights for M in (128, 256] -- each weight unit is converted
  // once and multiplied with tcgen05.mma N = 256 (128 cycles per instruction = the dense rate, scripts/umma_probe.cu) instead
  // of twice with N = 128. The accumulator then takes 256 of the 512 tensor-memory columns, single-buffered.
  // Measured (profiles/r02j_np256_ab.md): -7 % where every pair owns whole tiles (gate_up 75.7 -> 70.3 us, the 70B gate_up
  // 250.5 -> 244.6 us); shapes with fewer pair tiles than SM pairs, whose tiles are shared between pairs, got SLOWER in
  // isolation (qkv 23.8 -> 54.9 us, down 46.5 -> 70.4 us; cause not isolated: the 256 KB shared-tile reduction with a
  // single accumulator buffer is the suspect) => only for N / 256 >= 2 x SM pairs (see below). QSPEC_GEMM_NP256: 0 = never, 2 = always.
  {
    const int want = env_int("QSPEC_GEMM_NP256", 1);
    const int pairs = (sm_count() > 0 ? sm_count() : 148) / 2;
    // (r03c, after the shared-tile reduction was batched and the scheduler refitted: the two-pass form is ahead again on the 8B
    // gate_up, 67.7 vs 70.1 us -- 112 pair tiles leave 18 of 74 pairs idle in the one-pass form -- and behind on the 70B gate_up,
    // 246 vs 228 us => one-pass tiles from two full waves of pair tiles)
    if (pl.pair && !mode_is_int(mode) && M > 128 && (want == 2 || (want == 1 && N / (2 * TM) >= 2 * pairs)))
      pl.np = mp < 256 ? mp : 256;
  }
  pl.tiles_m = (M + pl.np - 1) / pl.np;
  pl.tiles_n = (N + TM - 1) / TM / (pl.pair ? 2 : 1);
  pl.KB = (K + UK - 1) / UK;
  pl.U = pl.tiles_m * pl.tiles_n * pl.KB;
  const int sms = (sm_count() > 0 ? sm_count() : 148) / (pl.pair ? 2 : 1);   // schedulable units: SMs, or SM pairs
  // measured on B200 (scripts/umma_probe.cu): a tcgen05.mma of this kernel's shape issues every ~58 cycles (N <= 64) in both
  // cta_group forms => per (pair-)unit 8 (int) / 16 (f16) instructions = 0.25 / 0.49 us at 1.9 GHz, plus the converters'
  // share; a pair-unit carries twice the bytes
  // ... and grows with the token tile (the converters widen np activation rows per unit, the MMAs get longer from N = 128):
  // measured slopes of the grid sweeps at 64 / 256 tokens (profiles/r02m_gsweep_tp.md): W4A4 0.46 us per unit at np = 64,
  // W4A16 0.93 (0.75 per pair-unit) at np = 128
  const double tu_np = pl.np > 16 ? (double)(pl.np - 16) : 0.0;
  // (256-token pair tiles: 1.9 us per pair-unit -- 55 % of the MMA rate: every pair pulls 128 KB of activations per unit from L2)
  const double tu_cta = mode_is_int(mode) ? (pl.pair ? 0.27 : 0.316) + 0.003 * tu_np
                        : pl.pair ? (pl.np > 128 ? 1.9 : 0.50 + 0.0022 * tu_np) : 0.606 + 0.0029 * tu_np;
  // every further contributor of a shared tile adds np x 128 words to the last arriver's reduction: us per token row
  // (decode tiles: fitted in round 1; larger tiles: r02m sweep, 3.7 us for 2 and 6.0 us for 4 contributors at np = 64)
  const double red_row = pl.np <= 32 ? 0.00512 : 0.02;
  const double tu_hbm = pl.pair ? 0.00512 : 0.00256;
  // Grid size. Default: every SM, units split evenly (stream-K). When a CTA's range is shorter than a tile it then
  // usually straddles two tiles and ends with TWO shared-tile reductions back to back (~2.5 us each, measured), which
  // for the small decode GEMMs costs more than it saves: prefer a slightly smaller grid whose ranges are aligned to the
  // tiles (q units per CTA with q | KB: one shared tile per CTA; q a multiple of KB: whole tiles, no reduction).
  // Cost model (us, measured on B200 with scripts/gpu_gemm_gsweep.py): a CTA needs tu per unit -- the tensor pipe's
  // 8 (W4A4) / 16 (W4A16) minimum-latency MMAs per unit, or its share of HBM when every SM streams --; a trailing
  // reduction of a shared tile through the workspace costs tp, through distributed shared memory (the CTAs of a
  // tile launched as a thread-block cluster, decode-sized M only) tc.
  pl.G = pl.U < sms ? pl.U : sms;
  pl.cs = 1;
  {
    const double tp = 2.5, tc = 1.0;
    auto tu = [&](int g) { const double cta = tu_cta, hbm = tu_hbm * g; return cta > hbm ? cta : hbm; };
    // Measured: launching the whole-SM CTAs as clusters costs ~3 us per GEMM in the decode loop (co-scheduling of 4 free
    // SMs of one GPC, no overlap with the predecessor's tail), more than the cheaper reduction saves: off by default.
    static const bool use_cluster = env_int("QSPEC_GEMM_CLUSTER", 0) != 0;
    const int q0 = (pl.U + pl.G - 1) / pl.G;
    const int tiles = pl.tiles_m * pl.tiles_n;
    const double share0 = (double)pl.G / tiles;            // CTAs per tile of the all-SM split
    const double red0 = tp + (share0 > 2.0 ? share0 - 1.0 : 1.0) * pl.np * red_row;   // a shared tile has at least one further contributor
    double best = q0 * tu(pl.G) + ((pl.U % pl.G == 0 && (pl.KB % q0 == 0 || q0 % pl.KB == 0)) ? (q0 % pl.KB == 0 ? 0.0 : red0)
                                                                                            : 2 * red0);
    for (int cs = 1; cs <= pl.KB; ++cs) {                  // cs CTAs per tile, each CTA inside one tile
      const int g = tiles * cs, q = (pl.KB + cs - 1) / cs;
      if (g > sms) break;
      // co-resident clusters of whole-SM CTAs: 148 CTAs in pairs, 132 in fours (8 GPCs of 16..20 SMs)
      const bool clus = use_cluster && !pl.pair && pl.KB % cs == 0 && pl.np <= 32 && pl.tiles_m == 1 &&
                        ((cs == 2 && g <= 148) || (cs == 4 && g <= 128));
      // the last arriver of a tile reloads cs - 1 partial tiles of np x 128 words at ~100 GB/s (one SM)
      const double c = q * tu(g) + (cs == 1 ? 0.0 : clus ? tc : tp + (cs - 1) * pl.np * red_row);
      if (c < best - 1e-9) { best = c; pl.G = g; pl.cs = clus ? cs : 1; }
    }
    for (int m = 2; m <= 8; ++m) {                         // q = m * KB (whole tiles)
      if (tiles % m) continue;
      const int g = tiles / m;
      if (g > sms) continue;
      const double c = (double)m * pl.KB * tu(g);
      if (c < best - 1e-9) { best = c; pl.G = g; pl.cs = 1; }
    }
  }
  if (deferred) {
    // Deferred reduction: the consumer kernel sums the k-slices, so a shared tile costs nothing here; every tile is
    // split the same way (split = CTAs per tile), or CTAs take whole tiles (split = 1).
    auto tu = [&](int g) { const double cta = tu_cta, hbm = tu_hbm * g; return cta > hbm ? cta : hbm; };
    const int tiles = pl.tiles_m * pl.tiles_n;
    double best = 1e30;
    for (int cs = 1; cs <= 8 && cs <= pl.KB; ++cs) {
      const int g = tiles * cs, q = (pl.KB + cs - 1) / cs;
      if (g > sms) break;
      const double c = q * tu(g) + 0.15 * (cs - 1);          // the consumer reads one more slice
      if (c < best - 1e-9) { best = c; pl.G = g; pl.cs = cs; }
    }
    for (int m = 2; m <= 16; ++m) {
      if (tiles % m) continue;
      const int g = tiles / m;
      if (g > sms) continue;
      const double c = (double)m * pl.KB * tu(g);
      if (c < best - 1e-9) { best = c; pl.G = g; pl.cs = 1; }
    }
    if (best > 1e29) { pl.G = 0; pl.cs = 0; }                // more tiles than SMs and no even split: not supported
  }
  const int split = deferred ? pl.cs : 1;
  if (deferred) pl.cs = 1;                                 // (cs doubles as the cluster size of the in-kernel reduction)
  pl.split = split;
  if (!deferred) {
    const char* e = getenv("QSPEC_GEMM_G");                // experiment: force the grid size (read at every call)
    const int g = e ? atoi(e) : 0;
    if (g >= 1 && g <= sms && g <= pl.U) { pl.G = g; pl.cs = 1; }
  }
  // decode tiles: 256 TMEM columns and <= ~110 KB of shared memory, so that two CTAs (consecutive GEMMs) fit an SM
  // QSPEC_GEMM_BIG=1 (experiment): give every CTA the whole SM (512 TMEM columns, deep ring) instead of sharing it
  const bool small = pl.np <= 64 && !env_int("QSPEC_GEMM_BIG", 1);
  pl.dcols = pl.np <= 32 ? 32 : pl.np <= 64 ? 64 : pl.np <= 128 ? 128 : 256;
  pl.dbufs = pl.dcols <= 128 ? 2 : 1;
  pl.tmem_cols = small ? 256 : 512;
  pl.ta = (pl.tmem_cols - pl.dbufs * pl.dcols) / 32;
  if (pl.ta > MAX_TA) pl.ta = MAX_TA;
  const int npl = pl.pair ? pl.np / 2 : pl.np;              // token rows per CTA
  const size_t act_stage = mode == MODE_I8 ? (size_t)npl * 128 : mode == MODE_A8 ? (size_t)npl * 256 : (size_t)npl * 512;
  const size_t fixed = 1024 + (mode == MODE_I8 ? (size_t)pl.ta * npl * 128 : 0);
  const size_t budget = small ? (size_t)110 * 1024 : (size_t)(227 * 1024 - 2048);
  int st = (int)((budget - fixed) / (W_STAGE + act_stage));
  if (st > MAX_STAGES) st = MAX_STAGES;
  if (!small && st > env_int("QSPEC_GEMM_STAGES", 6)) st = env_int("QSPEC_GEMM_STAGES", 6);
  if (st < 2) st = 2;
  pl.stages = st;
  pl.smem = fixed + (size_t)st * (W_STAGE + act_stage) + (pl.cs > 1 ? (size_t)(pl.cs - 1) * pl.np * TM * 4 : 0);
  pl.slot_bytes = (size_t)pl.np * TM * 4;
  return pl;
}

static size_t workspace_bytes(int mode, int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  const Plan pl = make_plan(mode, M, N, K);
  return COUNTER_BYTES + 2 * (size_t)pl.G * (pl.pair ? 2 : 1) * pl.slot_bytes;
}

template <int MODE, bool PAIR>
static int launch(const CUtensorMap& mw, const CUtensorMap& ma, const Params& p, const Plan& pl, cudaStream_t st) {
  auto kern = w4_stream_kernel<MODE, PAIR>;
  static bool configured = false;
  if (!configured) {
    QSPEC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(PAIR ? 2 * pl.G : pl.G);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = pl.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (env_int("QSPEC_GEMM_PDL", 1)) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (PAIR || pl.cs > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = PAIR ? 2u : (unsigned)pl.cs;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  QSPEC_CUDA(cudaLaunchKernelEx(&cfg, kern, mw, ma, p));
  count_launch(1);
  return QSPEC_OK;
}

static int run(int mode, const void* act, const void* sA, const int8_t* w, const void* sW, const void* bias, void* out,
               int32_t* acc_out, int M, int N, int K, int out_bf16, void* workspace, size_t ws_bytes, cudaStream_t st,
               const char* who, uint32_t* split_out = nullptr, size_t split_bytes = 0) {
  QSPEC_REQUIRE(act && w && (out || acc_out || split_out), QSPEC_ERR_INVALID_ARG, "%s: null pointer", who);
  QSPEC_REQUIRE(!out || sW, QSPEC_ERR_INVALID_ARG, "%s: null scale pointer", who);
  QSPEC_REQUIRE(!mode_is_int(mode) || !out || sA, QSPEC_ERR_INVALID_ARG, "%s: null activation-scale pointer", who);
  QSPEC_REQUIRE(M >= 0 && N > 0 && K > 0, QSPEC_ERR_INVALID_ARG, "%s: bad sizes M=%d N=%d K=%d", who, M, N, K);
  QSPEC_REQUIRE(K % 128 == 0 && K <= 65536 && N % 8 == 0, QSPEC_ERR_UNSUPPORTED,
                "%s: need K %% 128 == 0, K <= 65536, N %% 8 == 0 (got N=%d K=%d)", who, N, K);
  QSPEC_REQUIRE(((uintptr_t)act & 15) == 0 && ((uintptr_t)w & 15) == 0, QSPEC_ERR_UNSUPPORTED,
                "%s: operand pointers must be 16-byte aligned", who);
  if (M == 0) return QSPEC_OK;
  const size_t need = workspace_bytes(mode, M, N, K);
  QSPEC_REQUIRE(workspace && ws_bytes >= need, QSPEC_ERR_WORKSPACE, "%s: workspace too small (%zu < %zu)", who, ws_bytes, need);
  const Plan pl = make_plan(mode, M, N, K, split_out != nullptr);
  if (split_out) {
    QSPEC_REQUIRE(pl.split >= 1 && pl.tiles_m == 1, QSPEC_ERR_UNSUPPORTED, "%s: no even k-split for M=%d N=%d K=%d", who, M, N, K);
    QSPEC_REQUIRE(split_bytes >= (size_t)pl.split * M * N * 4, QSPEC_ERR_WORKSPACE, "%s: partial buffer too small (%zu < %zu)", who,
                  split_bytes, (size_t)pl.split * M * N * 4);
  }
  CUtensorMap mw, ma;
  int rc = get_map(w, (uint64_t)N, (uint64_t)K / 2, TM, 1, &mw);
  if (rc != QSPEC_OK) return rc;
  const uint32_t act_rows = (uint32_t)(pl.pair ? pl.np / 2 : pl.np);     // token rows per TMA box: each CTA of a pair loads its half
  rc = mode == MODE_I8 ? get_map(act, (uint64_t)M, (uint64_t)K / 2, act_rows, 1, &ma)
     : mode == MODE_A8 ? get_map(act, (uint64_t)M, (uint64_t)K, act_rows, 1, &ma)
                       : get_map(act, (uint64_t)M, (uint64_t)K * 2, act_rows, 2, &ma);
  if (rc != QSPEC_OK) return rc;
  Params p;
  p.sA = sA; p.sW = sW; p.bias = bias; p.out = out; p.acc_out = acc_out; p.split_out = split_out;
  p.counters = (unsigned*)workspace;
  p.ws_part = (uint32_t*)((unsigned char*)workspace + COUNTER_BYTES);
  p.trigger = env_int("QSPEC_GEMM_TRIGGER", 2);
  p.debug_exit = env_int("QSPEC_GEMM_DEBUG_EXIT", 0);
  p.trace_cta = env_int("QSPEC_GEMM_TRACE_CTA", pl.G - 1);
  p.trace = g_trace ? g_trace + (size_t)(g_trace_seq++ % 8u) * 4096 : nullptr;
  p.M = M; p.N = N; p.K = K; p.np = pl.np; p.tiles_m = pl.tiles_m; p.KB = pl.KB; p.U = pl.U; p.G = pl.G; p.cs = pl.cs;
  p.stages = pl.stages; p.dcols = pl.dcols; p.dbufs = pl.dbufs; p.ta = pl.ta; p.tmem_cols = pl.tmem_cols; p.out_bf16 = out_bf16;
  if (pl.pair) {
    switch (mode) {
      case MODE_I8: return launch<MODE_I8, true>(mw, ma, p, pl, st);
      case MODE_F16: return launch<MODE_F16, true>(mw, ma, p, pl, st);
      case MODE_A8: return launch<MODE_A8, true>(mw, ma, p, pl, st);
      default: return launch<MODE_BF16, true>(mw, ma, p, pl, st);
    }
  }
  switch (mode) {
    case MODE_I8: return launch<MODE_I8, false>(mw, ma, p, pl, st);
    case MODE_F16: return launch<MODE_F16, false>(mw, ma, p, pl, st);
    case MODE_A8: return launch<MODE_A8, false>(mw, ma, p, pl, st);
    default: return launch<MODE_BF16, false>(mw, ma, p, pl, st);
  }
}

}  // namespace w4s
}  // namespace qspec

using namespace qspec;

extern "C" void qspec_debug_set_gemm_trace(void* device_buffer) {
  w4s::g_trace = (unsigned long long*)device_buffer;
  w4s::g_trace_seq = 0;
}

extern "C" size_t qspec_gemm_w4a4_workspace_bytes(int M, int N, int K) { return w4s::workspace_bytes(w4s::MODE_I8, M, N, K); }

extern "C" int qspec_gemm_w4a4(const int8_t* a, const void* sA, const int8_t* w, const void* sW, const void* bias,
                               void* out, int M, int N, int K, int dtype, void* workspace, size_t workspace_bytes,
                               void* stream) {
  QSPEC_REQUIRE(out, QSPEC_ERR_INVALID_ARG, "gemm_w4a4: null output");
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "gemm_w4a4: dtype %d", dtype);
  return w4s::run(w4s::MODE_I8, a, sA, w, sW, bias, out, nullptr, M, N, K, dtype == QSPEC_DTYPE_BF16, workspace,
                  workspace_bytes, (cudaStream_t)stream, "gemm_w4a4");
}

extern "C" int qspec_gemm_w4a4_acc_i32(const int8_t* a, const int8_t* w, int32_t* acc, int M, int N, int K,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(acc, QSPEC_ERR_INVALID_ARG, "gemm_w4a4_acc_i32: null output");
  return w4s::run(w4s::MODE_I8, a, nullptr, w, nullptr, nullptr, nullptr, acc, M, N, K, 0, workspace, workspace_bytes,
                  (cudaStream_t)stream, "gemm_w4a4_acc_i32");
}

// W4A8: int8 activations [M, K] x the same packed int4 weights, exact int32 accumulation, the W4A4 epilogue.
// Replaces torchao's rowwise_scaled_linear_cutlass_s8s4 as called by Linear4bit.forward_w4a8 (quarot_nn/linear.py:86-99).
extern "C" size_t qspec_gemm_w4a8_workspace_bytes(int M, int N, int K) { return w4s::workspace_bytes(w4s::MODE_A8, M, N, K); }

extern "C" int qspec_gemm_w4a8(const int8_t* a, const void* sA, const int8_t* w, const void* sW, const void* bias,
                               void* out, int M, int N, int K, int dtype, void* workspace, size_t workspace_bytes,
                               void* stream) {
  QSPEC_REQUIRE(out, QSPEC_ERR_INVALID_ARG, "gemm_w4a8: null output");
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "gemm_w4a8: dtype %d", dtype);
  return w4s::run(w4s::MODE_A8, a, sA, w, sW, bias, out, nullptr, M, N, K, dtype == QSPEC_DTYPE_BF16, workspace,
                  workspace_bytes, (cudaStream_t)stream, "gemm_w4a8");
}

extern "C" int qspec_gemm_w4a8_acc_i32(const int8_t* a, const int8_t* w, int32_t* acc, int M, int N, int K,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(acc, QSPEC_ERR_INVALID_ARG, "gemm_w4a8_acc_i32: null output");
  return w4s::run(w4s::MODE_A8, a, nullptr, w, nullptr, nullptr, nullptr, acc, M, N, K, 0, workspace, workspace_bytes,
                  (cudaStream_t)stream, "gemm_w4a8_acc_i32");
}

// Introspection for tests and tuning scripts (no launch, works without a device: 148 SMs assumed then):
// out[0..7] = {token tile np, CTA pairs?, grid (CTAs or pairs), cluster size, deferred split, ring stages, token tiles, (pair-)tiles along N}
extern "C" int qspec_debug_gemm_plan(int mode, int M, int N, int K, int deferred, int* out) {
  if (!out || mode < 0 || mode > 3 || M <= 0 || N <= 0 || K <= 0 || K % 128 || N % 8) return QSPEC_ERR_INVALID_ARG;
  const w4s::Plan pl = w4s::make_plan(mode, M, N, K, deferred != 0);
  out[0] = pl.np; out[1] = pl.pair; out[2] = pl.G; out[3] = pl.cs; out[4] = pl.split; out[5] = pl.stages; out[6] = pl.tiles_m; out[7] = pl.tiles_n;
  return QSPEC_OK;
}

extern "C" int qspec_gemm_split_factor(int w4a16, int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0 || K % 128 || N % 8) return 0;
  const w4s::Plan pl = w4s::make_plan(w4a16 == 2 ? w4s::MODE_A8 : w4a16 ? w4s::MODE_F16 : w4s::MODE_I8, M, N, K, true);   // 2: int8-container W4A4
  return pl.tiles_m == 1 ? pl.split : 0;
}

extern "C" int qspec_gemm_w4a4_split(const int8_t* a, const int8_t* w, void* partials, size_t partial_bytes, int M, int N,
                                     int K, void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(partials, QSPEC_ERR_INVALID_ARG, "gemm_w4a4_split: null output");
  return w4s::run(w4s::MODE_I8, a, nullptr, w, nullptr, nullptr, nullptr, nullptr, M, N, K, 0, workspace, workspace_bytes,
                  (cudaStream_t)stream, "gemm_w4a4_split", (uint32_t*)partials, partial_bytes);
}

// W4A4 from int8-container activations (one int4 value per byte, [M, K] int8): the int8-activation instantiation (MODE_A8),
// partials = 16 * sum(a * w) (qspec_split_add_norm_* kind 2 shifts by 4 instead of 8).
extern "C" int qspec_gemm_w4a8_split(const int8_t* a, const int8_t* w, void* partials, size_t partial_bytes, int M, int N,
                                     int K, void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(partials, QSPEC_ERR_INVALID_ARG, "gemm_w4a8_split: null output");
  return w4s::run(w4s::MODE_A8, a, nullptr, w, nullptr, nullptr, nullptr, nullptr, M, N, K, 0, workspace, workspace_bytes,
                  (cudaStream_t)stream, "gemm_w4a8_split", (uint32_t*)partials, partial_bytes);
}

extern "C" int qspec_gemm_w4a16_split(const void* x, const int8_t* w, void* partials, size_t partial_bytes, int M, int N,
                                      int K, int dtype, void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(partials, QSPEC_ERR_INVALID_ARG, "gemm_w4a16_split: null output");
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "gemm_w4a16_split: dtype %d", dtype);
  return w4s::run(dtype == QSPEC_DTYPE_BF16 ? w4s::MODE_BF16 : w4s::MODE_F16, x, nullptr, w, nullptr, nullptr, nullptr, nullptr, M, N,
                  K, dtype == QSPEC_DTYPE_BF16, workspace, workspace_bytes, (cudaStream_t)stream, "gemm_w4a16_split",
                  (uint32_t*)partials, partial_bytes);
}

extern "C" size_t qspec_gemm_w4a16_workspace_bytes(int M, int N, int K) {
  return w4s::workspace_bytes(w4s::MODE_F16, M, N, K);
}

extern "C" int qspec_gemm_w4a16(const void* x, const int8_t* w, const void* sW, const void* bias, void* out, int M,
                                int N, int K, int dtype, void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(out && sW, QSPEC_ERR_INVALID_ARG, "gemm_w4a16: null pointer");
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "gemm_w4a16: dtype %d", dtype);
  return w4s::run(dtype == QSPEC_DTYPE_BF16 ? w4s::MODE_BF16 : w4s::MODE_F16, x, nullptr, w, sW, bias, out, nullptr, M, N,
                  K, dtype == QSPEC_DTYPE_BF16, workspace, workspace_bytes, (cudaStream_t)stream, "gemm_w4a16");
}
