// W4A4 (draft-mode) GEMM on 5th-gen tensor cores: exact int32 accumulation of int4 x int4.
// Replaces torchao's rowwise_scaled_linear_cutlass_s4s4_unified
// (third-party/ao/torchao/csrc/cuda/rowwise_scaled_linear_cutlass/rowwise_scaled_linear_cutlass_unified.cuh:240-488,
//  an SM80-only mma.sync.s4 kernel). sm_100a only.
//
// sm_100 has no int4 tensor-core type, so both operands are widened to int8 on the fly:
// byte = nibble << 4 (sign lands in bit 7, value = 16*q), which costs 3 ALU ops per 8 values and
// makes the int32 accumulator exactly 256 * sum(a*w) (|.| <= 64*K*256 < 2^31 for K <= 65536);
// the epilogue shifts right by 8.
//
// Layout / roles (one CTA = one 128-channel x NP-token tile, split-K over CTAs):
//   swap-AB: UMMA M = 128 output channels (weights are the A operand), UMMA N = NP tokens
//   (16..256), UMMA K = 32 int8; accumulator D[128 x NP] int32 lives in TMEM.
//   warps 0-3  producers: 128-bit global loads of packed int4 (weights and activations),
//              widen in registers, store the int8 tile in the canonical K-major SWIZZLE_128B
//              layout, fence.proxy.async, arrive on full[stage]; afterwards the same warps
//              run the epilogue (tcgen05.ld of their 32 TMEM lanes).
//   warp 4     TMEM allocator + single-thread tcgen05.mma.kind::i8 issuer; tcgen05.commit
//              releases shared-memory stages (empty[stage]) and signals the epilogue.
//   split-K    int32 atomics into a zeroed workspace (order-independent => deterministic);
//              the last-arriving CTA of a tile applies the epilogue and re-zeroes it.
#include "common.cuh"

namespace qspec {
namespace w4a4 {

constexpr int TM = 128;        // channels per CTA (UMMA M)
constexpr int BK = 128;        // k per stage (128 int8 = one 128-byte swizzle row)
constexpr int STAGES = 4;
constexpr int PF = 2;          // k-blocks of global loads kept in flight per producer thread
constexpr int NPROD = 128;
constexpr int THREADS = 160;

// ---- PTX wrappers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}

// K-major SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout):
//   [0,14) start >> 4 | [16,30) LBO >> 4 (unused for swizzled K-major) | [32,46) SBO >> 4 = 1024 B
//   (8 rows x 128 B) | [46,48) version = 1 | [61,64) layout = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::i8 instruction descriptor (cute::UMMA::InstrDescriptor): c=S32 (2) at [4,6), a=INT8 (1) at
// [7,10), b=INT8 (1) at [10,13), both K-major, N>>3 at [17,23), M>>4 at [24,29).
__host__ __device__ constexpr uint32_t make_idesc_i8(int np) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(np >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
}

// 16 packed bytes (32 nibbles) -> 32 int8 (nibble << 4) as two 16-byte chunks. The k order inside
// each 8-byte group is (0,2,4,6,1,3,5,7); both operands go through this routine, so it cancels.
__device__ __forceinline__ void widen32(const uint4& p, uint4& c0, uint4& c1) {
  c0.x = (p.x << 4) & 0xf0f0f0f0u; c0.y = p.x & 0xf0f0f0f0u;
  c0.z = (p.y << 4) & 0xf0f0f0f0u; c0.w = p.y & 0xf0f0f0f0u;
  c1.x = (p.z << 4) & 0xf0f0f0f0u; c1.y = p.z & 0xf0f0f0f0u;
  c1.z = (p.w << 4) & 0xf0f0f0f0u; c1.w = p.w & 0xf0f0f0f0u;
}
// store the two chunks of (row, packed-chunk c) into a [rows][128 B] SWIZZLE_128B tile
__device__ __forceinline__ void store_sw128(unsigned char* tile, int row, int c, const uint4& c0, const uint4& c1) {
  unsigned char* r = tile + row * 128;
  *reinterpret_cast<uint4*>(r + (((2 * c) ^ (row & 7)) << 4)) = c0;
  *reinterpret_cast<uint4*>(r + (((2 * c + 1) ^ (row & 7)) << 4)) = c1;
}

struct Params {
  const uint8_t* a;      // [M, K/2] packed activations
  const uint8_t* w;      // [N, K/2] packed weights
  const void* sA;        // [M]
  const void* sW;        // [N]
  const void* bias;      // [N] or null
  void* out;             // [M, N] T          (null in acc mode)
  int32_t* acc_out;      // [M, N] raw int32  (acc mode)
  int32_t* ws_acc;       // [M, N] zeroed split-K accumulators
  unsigned* counters;    // [tiles_m * tiles_n]
  int M, N, K;
  int np;                // UMMA N of full token tiles (<= 256, multiple of 16)
  int splitK, kb_per_split, kb_total;
  int bf16;
};

template <int BCH>   // activation 16-byte chunks per producer thread per k-block (NP*4/128 rounded up)
__global__ void __launch_bounds__(THREADS, 1)
gemm_w4a4_kernel(const Params p) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // carve: [STAGES][A 16 KB] [STAGES][B np*128] ; barriers in static smem
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ unsigned s_is_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile_n = blockIdx.x, split = blockIdx.y, tile_m = blockIdx.z;
  const int n0 = tile_n * TM;
  const int m0 = tile_m * p.np;
  const int m_here = min(p.M - m0, p.np);                 // valid tokens in this tile
  const int np = ((m_here + 15) >> 4) << 4;               // UMMA N for this tile
  const int kb_begin = split * p.kb_per_split;
  const int kb_end = min(kb_begin + p.kb_per_split, p.kb_total);
  const int nkb = kb_end - kb_begin;
  const size_t row_bytes = (size_t)p.K >> 1;

  // dynamic smem base rounded up to 1024 B (SWIZZLE_128B atoms are 1024-byte aligned)
  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* smA = base;
  unsigned char* smB = base + STAGES * (TM * 128);
  const int b_stage_bytes = p.np * 128;

  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < np) tmem_cols <<= 1;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&full_bar[s]), NPROD);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    mbar_init(smem_u32(&tmem_full_bar), 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_smem;

  if (warp < 4) {
    // =========================== producers ===========================
    uint4 wreg[PF][4];
    uint4 areg[PF][BCH];
    auto issue_loads = [&](int i, int slot) {     // i = local k-block index
      const int kb = kb_begin + i;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int idx = tid + NPROD * j;
        const int c = idx & 3, row = idx >> 2;
        if (n0 + row < p.N)
          wreg[slot][j] = __ldg(reinterpret_cast<const uint4*>(p.w + (size_t)(n0 + row) * row_bytes + (size_t)kb * 64 + 16 * c));
        else
          wreg[slot][j] = make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int j = 0; j < BCH; ++j) {
        const int idx = tid + NPROD * j;
        const int c = idx & 3, tok = idx >> 2;
        if (tok < m_here)
          areg[slot][j] = __ldg(reinterpret_cast<const uint4*>(p.a + (size_t)(m0 + tok) * row_bytes + (size_t)kb * 64 + 16 * c));
        else
          areg[slot][j] = make_uint4(0, 0, 0, 0);
      }
    };
#pragma unroll
    for (int s = 0; s < PF; ++s)
      if (s < nkb) issue_loads(s, s);

    for (int i0 = 0; i0 < nkb; i0 += PF) {
#pragma unroll
      for (int s = 0; s < PF; ++s) {
        const int i = i0 + s;
        if (i < nkb) {
          const int stage = i % STAGES;
          const uint32_t phase = (uint32_t)(i / STAGES) & 1u;
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          unsigned char* tA = smA + stage * (TM * 128);
          unsigned char* tB = smB + stage * b_stage_bytes;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int idx = tid + NPROD * j;
            uint4 c0, c1;
            widen32(wreg[s][j], c0, c1);
            store_sw128(tA, idx >> 2, idx & 3, c0, c1);
          }
#pragma unroll
          for (int j = 0; j < BCH; ++j) {
            const int idx = tid + NPROD * j;
            if ((idx >> 2) < np) {
              uint4 c0, c1;
              widen32(areg[s][j], c0, c1);
              store_sw128(tB, idx >> 2, idx & 3, c0, c1);
            }
          }
          fence_proxy_async_smem();
          mbar_arrive(smem_u32(&full_bar[stage]));
          if (i + PF < nkb) issue_loads(i + PF, s);
        }
      }
    }

    // =========================== epilogue ===========================
    mbar_wait(smem_u32(&tmem_full_bar), 0);
    tc_fence_after();
    const int ch = n0 + warp * 32 + lane;                 // this thread's output channel
    const bool ch_ok = ch < p.N;
    const uint32_t lane_addr = tmem_d + ((uint32_t)(warp * 32) << 16);
    const bool direct = (p.splitK == 1);
    float sw = 0.f, bs = 0.f;
    if (direct && ch_ok && p.out) {
      sw = p.bf16 ? __bfloat162float(((const __nv_bfloat16*)p.sW)[ch]) : __half2float(((const __half*)p.sW)[ch]);
      if (p.bias) bs = p.bf16 ? __bfloat162float(((const __nv_bfloat16*)p.bias)[ch]) : __half2float(((const __half*)p.bias)[ch]);
    }
    for (int c0 = 0; c0 < np; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(lane_addr + (uint32_t)c0, r);
      if (!ch_ok) continue;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int tok = c0 + j;
        if (tok < m_here) {
          const int32_t v = (int32_t)r[j];
          const size_t o = (size_t)(m0 + tok) * p.N + ch;
          if (!direct) {
            atomicAdd(p.ws_acc + o, v);
          } else if (p.acc_out) {
            p.acc_out[o] = v >> 8;
          } else {
            const float sa = p.bf16 ? __bfloat162float(((const __nv_bfloat16*)p.sA)[m0 + tok])
                                    : __half2float(((const __half*)p.sA)[m0 + tok]);
            float f = __fmul_rn(__fmul_rn(__int2float_rn(v >> 8), sa), sw);
            if (p.bias) f = __fadd_rn(f, bs);
            if (p.bf16) ((__nv_bfloat16*)p.out)[o] = __float2bfloat16_rn(f);
            else ((__half*)p.out)[o] = __float2half_rn(f);
          }
        }
      }
    }
    tc_fence_before();
  } else {
    // =========================== MMA issuer (warp 4) ===========================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_i8(np);
      for (int i = 0; i < nkb; ++i) {
        const int stage = i % STAGES;
        const uint32_t phase = (uint32_t)(i / STAGES) & 1u;
        mbar_wait(smem_u32(&full_bar[stage]), phase);
        tc_fence_after();
        const uint64_t adesc = make_desc_sw128(smem_u32(smA + stage * (TM * 128)));
        const uint64_t bdesc = make_desc_sw128(smem_u32(smB + stage * b_stage_bytes));
#pragma unroll
        for (int k = 0; k < BK / 32; ++k)      // +32 bytes along K = +2 in the (>>4) start field
          umma_i8(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (i > 0 || k > 0) ? 1u : 0u);
        umma_commit(smem_u32(&empty_bar[stage]));
      }
      umma_commit(smem_u32(&tmem_full_bar));
    }
    __syncwarp();
  }

  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem_d, tmem_cols);
  }
  if (p.splitK == 1) return;

  // ---- split-K: last-arriving CTA of the tile turns the summed accumulators into the output
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned id = tile_m * gridDim.x + tile_n;
    const unsigned prev = atomicAdd(p.counters + id, 1u);
    s_is_last = (prev == (unsigned)p.splitK - 1u);
    if (s_is_last) p.counters[id] = 0u;
  }
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();
  if (tid < TM) {
    const int ch = n0 + tid;
    if (ch < p.N) {
      float sw = 0.f, bs = 0.f;
      if (p.out) {
        sw = p.bf16 ? __bfloat162float(((const __nv_bfloat16*)p.sW)[ch]) : __half2float(((const __half*)p.sW)[ch]);
        if (p.bias) bs = p.bf16 ? __bfloat162float(((const __nv_bfloat16*)p.bias)[ch]) : __half2float(((const __half*)p.bias)[ch]);
      }
      for (int tok = 0; tok < m_here; ++tok) {
        const size_t o = (size_t)(m0 + tok) * p.N + ch;
        const int32_t v = __ldcg(p.ws_acc + o);
        p.ws_acc[o] = 0;
        if (p.acc_out) {
          p.acc_out[o] = v >> 8;
        } else {
          const float sa = p.bf16 ? __bfloat162float(((const __nv_bfloat16*)p.sA)[m0 + tok])
                                  : __half2float(((const __half*)p.sA)[m0 + tok]);
          float f = __fmul_rn(__fmul_rn(__int2float_rn(v >> 8), sa), sw);
          if (p.bias) f = __fadd_rn(f, bs);
          if (p.bf16) ((__nv_bfloat16*)p.out)[o] = __float2bfloat16_rn(f);
          else ((__half*)p.out)[o] = __float2half_rn(f);
        }
      }
    }
  }
}

struct Plan { int np, tiles_n, tiles_m, splitK, kb_per_split, kb_total, bch; size_t smem; };

static Plan make_plan(int M, int N, int K) {
  Plan pl;
  const int mp = ((M + 15) / 16) * 16;
  pl.np = mp < 256 ? mp : 256;
  pl.tiles_m = (M + pl.np - 1) / pl.np;
  pl.tiles_n = (N + TM - 1) / TM;
  pl.kb_total = K / BK;
  pl.bch = (pl.np * 4 + NPROD - 1) / NPROD;
  pl.bch = pl.bch <= 1 ? 1 : pl.bch <= 2 ? 2 : pl.bch <= 4 ? 4 : 8;
  pl.smem = (size_t)STAGES * (TM * 128 + pl.np * 128) + 1024;
  const int sms = sm_count() > 0 ? sm_count() : 148;
  int per_sm = (int)((227 * 1024) / (pl.smem + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 3) per_sm = 3;
  const int tiles = pl.tiles_n * pl.tiles_m;
  int want = (sms * per_sm + tiles - 1) / tiles;
  if (want < 1) want = 1;
  if (want > 32) want = 32;
  if (want > pl.kb_total) want = pl.kb_total;
  pl.kb_per_split = (pl.kb_total + want - 1) / want;
  pl.splitK = (pl.kb_total + pl.kb_per_split - 1) / pl.kb_per_split;
  return pl;
}

template <int BCH>
static int launch(const Params& p, const Plan& pl, cudaStream_t st) {
  auto kern = gemm_w4a4_kernel<BCH>;
  static size_t configured = 0;
  if (pl.smem > configured) {
    QSPEC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
    configured = pl.smem;
  }
  dim3 grid(pl.tiles_n, pl.splitK, pl.tiles_m);
  kern<<<grid, THREADS, pl.smem, st>>>(p);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

static int run(const int8_t* a, const void* sA, const int8_t* w, const void* sW, const void* bias,
               void* out, int32_t* acc_out, int M, int N, int K, int dtype, void* workspace,
               size_t workspace_bytes, cudaStream_t st, const char* who) {
  QSPEC_REQUIRE(a && w && (out || acc_out), QSPEC_ERR_INVALID_ARG, "%s: null pointer", who);
  QSPEC_REQUIRE(!out || (sA && sW), QSPEC_ERR_INVALID_ARG, "%s: null scale pointer", who);
  QSPEC_REQUIRE(M >= 0 && N > 0 && K > 0, QSPEC_ERR_INVALID_ARG, "%s: bad sizes M=%d N=%d K=%d", who, M, N, K);
  QSPEC_REQUIRE(K % 128 == 0 && K <= 65536 && N % 8 == 0, QSPEC_ERR_UNSUPPORTED,
                "%s: need K %% 128 == 0, K <= 65536, N %% 8 == 0 (got N=%d K=%d)", who, N, K);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "%s: dtype %d", who, dtype);
  if (M == 0) return QSPEC_OK;
  const size_t need = qspec_gemm_w4a4_workspace_bytes(M, N, K);
  QSPEC_REQUIRE(workspace && workspace_bytes >= need, QSPEC_ERR_WORKSPACE,
                "%s: workspace too small (%zu < %zu)", who, workspace_bytes, need);
  const Plan pl = make_plan(M, N, K);
  Params p;
  p.a = (const uint8_t*)a; p.w = (const uint8_t*)w; p.sA = sA; p.sW = sW; p.bias = bias;
  p.out = out; p.acc_out = acc_out;
  p.counters = (unsigned*)workspace;
  const size_t cbytes = (((size_t)pl.tiles_n * pl.tiles_m * sizeof(unsigned)) + 255) & ~(size_t)255;
  p.ws_acc = (int32_t*)((unsigned char*)workspace + cbytes);
  p.M = M; p.N = N; p.K = K; p.np = pl.np;
  p.splitK = pl.splitK; p.kb_per_split = pl.kb_per_split; p.kb_total = pl.kb_total;
  p.bf16 = dtype == QSPEC_DTYPE_BF16;
  switch (pl.bch) {
    case 1: return launch<1>(p, pl, st);
    case 2: return launch<2>(p, pl, st);
    case 4: return launch<4>(p, pl, st);
    default: return launch<8>(p, pl, st);
  }
}

}  // namespace w4a4
}  // namespace qspec

using namespace qspec;

extern "C" size_t qspec_gemm_w4a4_workspace_bytes(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  const w4a4::Plan pl = w4a4::make_plan(M, N, K);
  const size_t cbytes = (((size_t)pl.tiles_n * pl.tiles_m * sizeof(unsigned)) + 255) & ~(size_t)255;
  return cbytes + (size_t)M * N * sizeof(int32_t);
}

extern "C" int qspec_gemm_w4a4(const int8_t* a, const void* sA, const int8_t* w, const void* sW,
                               const void* bias, void* out, int M, int N, int K, int dtype,
                               void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(out, QSPEC_ERR_INVALID_ARG, "gemm_w4a4: null output");
  return w4a4::run(a, sA, w, sW, bias, out, nullptr, M, N, K, dtype, workspace, workspace_bytes,
                   (cudaStream_t)stream, "gemm_w4a4");
}

extern "C" int qspec_gemm_w4a4_acc_i32(const int8_t* a, const int8_t* w, int32_t* acc, int M, int N,
                                       int K, void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(acc, QSPEC_ERR_INVALID_ARG, "gemm_w4a4_acc_i32: null output");
  return w4a4::run(a, nullptr, w, nullptr, nullptr, nullptr, acc, M, N, K, QSPEC_DTYPE_F16, workspace,
                   workspace_bytes, (cudaStream_t)stream, "gemm_w4a4_acc_i32");
}
