// W4A16 (verify-mode) decode GEMM: out[M,N] = x[M,K] @ (int4(W)[N,K] * sW[N])^T (+bias).
// Replaces the reference's BitBLAS Matmul + per-call `weight ^ 0x88` pass
// (quarot_nn/linear.py:102-124,156-211). sm_100a.
//
// Regime: M <= 64 per pass, HBM-bound on the packed weights (0.5 B/weight). Design:
//  * swap-AB: the 16 rows of an m16n8k16 MMA are 16 output channels, the 8 columns are tokens,
//    so the packed weight word a thread loads IS its A fragment after an in-register
//    int4->f16/bf16 decode (two's-complement nibbles decoded directly; no XOR pass over W);
//  * weights and activations are staged with 16-byte cp.async into a 4-stage shared-memory
//    ring (weights [kc][128 rows][64 B] read conflict-free; activations XOR-swizzled);
//  * K is split across CTAs so every SM streams HBM even at N=4096; partial fp32 tiles go to a
//    workspace and the last-arriving CTA of a tile reduces them in split order (deterministic),
//    applies sW/bias and stores 16-byte vectors.
// Tensor-core choice: legacy mma.sync (HMMA) is kept on purpose here -- at M<=64 the kernel is
// bound by HBM and by the int4 decode ALU work, and a tcgen05 path would have to write the
// decoded bf16 weights to shared memory (2 B/weight, 4x the HBM stream), which exceeds the
// 128 B/clk/SM shared-memory port at the HBM roofline rate. See DESIGN.md.
#include "common.cuh"

namespace qspec {
namespace w4a16 {

constexpr int TN = 128;          // output channels per CTA
constexpr int KC = 128;          // k per chunk (64 packed bytes per row)
constexpr int KCS = 2;           // chunks per stage
constexpr int STAGES = 4;
constexpr int THREADS = 256;     // 8 warps: 4 row groups (32 rows) x 2 chunks of the stage

__host__ __device__ constexpr int w_stage_bytes() { return KCS * TN * 64; }
__host__ __device__ constexpr int x_stage_bytes(int MT) { return MT * KCS * 256; }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ uint32_t lop3_and_xor(uint32_t a, uint32_t mask, uint32_t x) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0x6a;\n" : "=r"(r) : "r"(a), "r"(mask), "r"(x));  // (a & b) ^ c
  return r;
}

// 8 two's-complement nibbles -> 4 packed pairs (n0,n4) (n1,n5) (n2,n6) (n3,n7), exact.
template <bool BF16>
__device__ __forceinline__ void decode8(uint32_t w, uint32_t (&h)[4]) {
  if (!BF16) {
    // f16: 0x6400 = 1024, ulp 1.  (n ^ 8) = s + 8  ->  1024 + s + 8
    const uint32_t w8 = w >> 8;
    const uint32_t a0 = lop3_and_xor(w, 0x000f000fu, 0x64086408u);   // 1032 + s
    const uint32_t a1 = lop3_and_xor(w, 0x00f000f0u, 0x64806480u);   // 1152 + 16 s
    const uint32_t a2 = lop3_and_xor(w8, 0x000f000fu, 0x64086408u);
    const uint32_t a3 = lop3_and_xor(w8, 0x00f000f0u, 0x64806480u);
    const __half2 c1032 = __floats2half2_rn(1032.f, 1032.f);
    const __half2 c16th = __floats2half2_rn(0.0625f, 0.0625f);
    const __half2 cm72 = __floats2half2_rn(-72.f, -72.f);
    __half2 r0 = __hsub2(*reinterpret_cast<const __half2*>(&a0), c1032);
    __half2 r1 = __hfma2(*reinterpret_cast<const __half2*>(&a1), c16th, cm72);
    __half2 r2 = __hsub2(*reinterpret_cast<const __half2*>(&a2), c1032);
    __half2 r3 = __hfma2(*reinterpret_cast<const __half2*>(&a3), c16th, cm72);
    h[0] = *reinterpret_cast<uint32_t*>(&r0);
    h[1] = *reinterpret_cast<uint32_t*>(&r1);
    h[2] = *reinterpret_cast<uint32_t*>(&r2);
    h[3] = *reinterpret_cast<uint32_t*>(&r3);
  } else {
    // bf16: 0x4300 = 128, ulp 1 (7 mantissa bits) -> 128 + s + 8
    const __nv_bfloat162 c136 = __floats2bfloat162_rn(136.f, 136.f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t a = lop3_and_xor(w >> (4 * i), 0x000f000fu, 0x43084308u);
      __nv_bfloat162 r = __hsub2(*reinterpret_cast<const __nv_bfloat162*>(&a), c136);
      h[i] = *reinterpret_cast<uint32_t*>(&r);
    }
  }
}

template <bool BF16>
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  if (!BF16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};\n"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};\n"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
}

struct Params {
  const uint8_t* w;      // [N, K/2]
  const void* x;         // [M, K]
  const void* sW;        // [N]
  const void* bias;      // [N] or null
  void* out;             // [M, N]
  float* partial;        // [splitK][tiles_n][MT][TN] fp32 (splitK > 1)
  unsigned* counters;    // [tiles_n]
  int M, N, K;
  int splitK, stages_per_split, stages_total;
};

// NT = number of 8-token n-tiles (MT = 8*NT tokens per pass).
template <int NT, bool BF16>
__global__ void __launch_bounds__(THREADS, (NT <= 2) ? 2 : 1)
gemm_w4a16_kernel(const Params p) {
  constexpr int MT = 8 * NT;
  constexpr int WSB = w_stage_bytes();
  constexpr int XSB = x_stage_bytes(MT);
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* sm_w = smem;                       // STAGES x WSB
  unsigned char* sm_x = smem + STAGES * WSB;        // STAGES x XSB
  __shared__ unsigned s_is_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int wr = warp & 3, wk = warp >> 2;
  const int tile_n = blockIdx.x, split = blockIdx.y;
  const int n0 = tile_n * TN;
  const int st_begin = split * p.stages_per_split;
  const int st_end = min(st_begin + p.stages_per_split, p.stages_total);
  const int nst = st_end - st_begin;
  const size_t row_bytes = (size_t)p.K >> 1;
  const uint32_t sm_w_u = (uint32_t)__cvta_generic_to_shared(sm_w);
  const uint32_t sm_x_u = (uint32_t)__cvta_generic_to_shared(sm_x);

  auto load_stage = [&](int st_local) {
    const int slot = st_local % STAGES;
    const int kbase = (st_begin + st_local) * (KC * KCS);
    // weights: KCS x TN rows x 4 chunks of 16 B
#pragma unroll
    for (int i = 0; i < (KCS * TN * 4) / THREADS; ++i) {
      const int idx = tid + i * THREADS;
      const int c = idx & 3, row = (idx >> 2) & (TN - 1), kc = idx >> 9;
      const int k = kbase + kc * KC;
      const bool ok = (n0 + row < p.N) && (k < p.K);
      const uint8_t* src = p.w + (ok ? ((size_t)(n0 + row) * row_bytes + (k >> 1) + 16 * c) : 0);
      cp_async16(sm_w_u + slot * WSB + ((kc * TN + row) << 6) + (c << 4), src, ok ? 16 : 0);
    }
    // activations: MT tokens x KCS x 16 chunks, chunk swizzle p = c ^ ((c>>3)<<1) ^ (token&1)
#pragma unroll
    for (int i = 0; i < (MT * KCS * 16 + THREADS - 1) / THREADS; ++i) {
      const int idx = tid + i * THREADS;
      if (idx < MT * KCS * 16) {
        const int c = idx & 15, kc = (idx >> 4) % KCS, tok = idx / (16 * KCS);
        const int k = kbase + kc * KC + c * 8;
        const bool ok = (tok < p.M) && (k < p.K);
        const unsigned char* src =
            (const unsigned char*)p.x + (ok ? ((size_t)tok * p.K + k) * 2 : 0);
        const int pc = c ^ (((c >> 3) & 1) << 1) ^ (tok & 1);
        cp_async16(sm_x_u + slot * XSB + tok * (KCS * 256) + kc * 256 + (pc << 4), src, ok ? 16 : 0);
      }
    }
  };

  float acc[2][NT][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[mt][nt][i] = 0.f;

  // prologue
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < nst) load_stage(s);
    cp_async_commit();
  }

  for (int st = 0; st < nst; ++st) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    // prefetch the stage that reuses the slot consumed in the previous iteration
    if (st + STAGES - 1 < nst) load_stage(st + STAGES - 1);
    cp_async_commit();

    const int slot = st % STAGES;
    const unsigned char* wst = sm_w + slot * WSB + ((wk * TN + wr * 32) << 6);
    const unsigned char* xst = sm_x + slot * XSB + wk * 256;
    // A words: rows (16mt + g) and (16mt + g + 8), 16 bytes at chunk t
    uint4 aw[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      aw[mt][0] = *reinterpret_cast<const uint4*>(wst + ((mt * 16 + g) << 6) + (t << 4));
      aw[mt][1] = *reinterpret_cast<const uint4*>(wst + ((mt * 16 + g + 8) << 6) + (t << 4));
    }
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      // B fragments for this 8-k group: chunk c = 4t + w of token 8nt + g
      uint32_t b[NT][4];
      const int c = 4 * t + w;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int tok = nt * 8 + g;
        const int pc = c ^ (((c >> 3) & 1) << 1) ^ (tok & 1);
        const uint4 r = *reinterpret_cast<const uint4*>(xst + tok * (KCS * 256) + (pc << 4));
        b[nt][0] = __byte_perm(r.x, r.z, 0x5410);   // (x0, x4)
        b[nt][1] = __byte_perm(r.x, r.z, 0x7632);   // (x1, x5)
        b[nt][2] = __byte_perm(r.y, r.w, 0x5410);   // (x2, x6)
        b[nt][3] = __byte_perm(r.y, r.w, 0x7632);   // (x3, x7)
      }
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        uint32_t lo[4], hi[4];
        decode8<BF16>(reinterpret_cast<const uint32_t*>(&aw[mt][0])[w], lo);   // row g
        decode8<BF16>(reinterpret_cast<const uint32_t*>(&aw[mt][1])[w], hi);   // row g+8
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          mma16816<BF16>(acc[mt][nt], lo[0], hi[0], lo[1], hi[1], b[nt][0], b[nt][1]);
          mma16816<BF16>(acc[mt][nt], lo[2], hi[2], lo[3], hi[3], b[nt][2], b[nt][3]);
        }
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();

  // ---- CTA tile reduce over the two chunk halves, through shared memory: tile[MT][TN] fp32
  float* tile = reinterpret_cast<float*>(smem);
  auto store_acc = [&](bool add) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int ch = wr * 32 + mt * 16 + g + ((i & 2) ? 8 : 0);
          const int tok = nt * 8 + 2 * t + (i & 1);
          float* q = tile + tok * TN + ch;
          *q = add ? __fadd_rn(*q, acc[mt][nt][i]) : acc[mt][nt][i];
        }
  };
  if (wk == 1) store_acc(false);
  __syncthreads();
  if (wk == 0) store_acc(true);
  __syncthreads();

  const int vec_per_tok = TN / 8;
  if (p.splitK > 1) {
    float* mine = p.partial + ((size_t)split * gridDim.x + tile_n) * (MT * TN);
    for (int i = tid; i < MT * TN / 4; i += THREADS)
      reinterpret_cast<float4*>(mine)[i] = reinterpret_cast<const float4*>(tile)[i];
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      const unsigned prev = atomicAdd(p.counters + tile_n, 1u);
      s_is_last = (prev == (unsigned)p.splitK - 1u);
    }
    __syncthreads();
    if (!s_is_last) return;
    __threadfence();
    for (int i = tid; i < MT * TN / 4; i += THREADS) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int s = 0; s < p.splitK; ++s) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(
                                    p.partial + ((size_t)s * gridDim.x + tile_n) * (MT * TN)) + i);
        if (s == 0) a = v;
        else { a.x = __fadd_rn(a.x, v.x); a.y = __fadd_rn(a.y, v.y); a.z = __fadd_rn(a.z, v.z); a.w = __fadd_rn(a.w, v.w); }
      }
      reinterpret_cast<float4*>(tile)[i] = a;
    }
    if (tid == 0) p.counters[tile_n] = 0u;          // self-cleaning
    __syncthreads();
  }

  // ---- epilogue: out[tok, ch] = T(acc * sW[ch] + bias[ch]), 8 channels (16 B) per thread
  for (int i = tid; i < MT * vec_per_tok; i += THREADS) {
    const int tok = i / vec_per_tok, cv = i % vec_per_tok;
    const int ch = n0 + cv * 8;
    if (tok >= p.M || ch >= p.N) continue;
    float v[8], s[8];
    const float4 v0 = reinterpret_cast<const float4*>(tile + tok * TN + cv * 8)[0];
    const float4 v1 = reinterpret_cast<const float4*>(tile + tok * TN + cv * 8)[1];
    v[0] = v0.x; v[1] = v0.y; v[2] = v0.z; v[3] = v0.w; v[4] = v1.x; v[5] = v1.y; v[6] = v1.z; v[7] = v1.w;
    unpack8<BF16>(__ldg(reinterpret_cast<const uint4*>((const unsigned char*)p.sW + (size_t)ch * 2)), s);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = __fmul_rn(v[j], s[j]);
    if (p.bias) {
      unpack8<BF16>(__ldg(reinterpret_cast<const uint4*>((const unsigned char*)p.bias + (size_t)ch * 2)), s);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = __fadd_rn(v[j], s[j]);
    }
    *reinterpret_cast<uint4*>((unsigned char*)p.out + ((size_t)tok * p.N + ch) * 2) = pack8<BF16>(v);
  }
}

struct Plan { int nt, tiles_n, splitK, stages_per_split, stages_total; size_t partial_bytes; };

static Plan make_plan(int M, int N, int K) {
  Plan pl;
  const int mt = M >= 64 ? 64 : M;
  pl.nt = mt <= 8 ? 1 : mt <= 16 ? 2 : mt <= 32 ? 4 : 8;
  pl.tiles_n = (N + TN - 1) / TN;
  pl.stages_total = (K + KC * KCS - 1) / (KC * KCS);
  const int sms = sm_count() > 0 ? sm_count() : 148;
  const int target = sms * (pl.nt <= 2 ? 2 : 1);
  int want = (target + pl.tiles_n - 1) / pl.tiles_n;      // splits that fill the machine once
  if (want < 1) want = 1;
  if (want > 16) want = 16;
  if (want > pl.stages_total) want = pl.stages_total;
  pl.stages_per_split = (pl.stages_total + want - 1) / want;
  pl.splitK = (pl.stages_total + pl.stages_per_split - 1) / pl.stages_per_split;
  pl.partial_bytes = pl.splitK > 1 ? (size_t)pl.splitK * pl.tiles_n * (8 * pl.nt) * TN * sizeof(float) : 0;
  return pl;
}

template <int NT, bool BF16>
static int launch(const Params& p, const Plan& pl, cudaStream_t st) {
  constexpr int MT = 8 * NT;
  const size_t smem = (size_t)STAGES * (w_stage_bytes() + x_stage_bytes(MT));
  auto kern = gemm_w4a16_kernel<NT, BF16>;
  static bool configured = false;
  if (!configured) {
    QSPEC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  dim3 grid(pl.tiles_n, pl.splitK);
  kern<<<grid, THREADS, smem, st>>>(p);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

}  // namespace w4a16
}  // namespace qspec

using namespace qspec;

extern "C" size_t qspec_gemm_w4a16_workspace_bytes(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  const w4a16::Plan pl = w4a16::make_plan(M, N, K);
  return pl.partial_bytes + (size_t)pl.tiles_n * sizeof(unsigned) + 256;
}

extern "C" int qspec_gemm_w4a16(const void* x, const int8_t* w, const void* sW, const void* bias,
                                void* out, int M, int N, int K, int dtype, void* workspace,
                                size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(x && w && sW && out, QSPEC_ERR_INVALID_ARG, "gemm_w4a16: null pointer");
  QSPEC_REQUIRE(M >= 0 && N > 0 && K > 0, QSPEC_ERR_INVALID_ARG, "gemm_w4a16: bad sizes M=%d N=%d K=%d", M, N, K);
  QSPEC_REQUIRE(K % 128 == 0 && N % 16 == 0, QSPEC_ERR_UNSUPPORTED,
                "gemm_w4a16: need K %% 128 == 0 and N %% 16 == 0 (got N=%d K=%d)", N, K);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED,
                "gemm_w4a16: dtype %d", dtype);
  if (M == 0) return QSPEC_OK;
  QSPEC_REQUIRE(workspace && workspace_bytes >= qspec_gemm_w4a16_workspace_bytes(M, N, K),
                QSPEC_ERR_WORKSPACE, "gemm_w4a16: workspace too small (%zu < %zu)", workspace_bytes,
                qspec_gemm_w4a16_workspace_bytes(M, N, K));
  cudaStream_t st = (cudaStream_t)stream;
  const w4a16::Plan pl = w4a16::make_plan(M, N, K);
  // counters first (256-byte aligned block), partials after
  unsigned* counters = (unsigned*)workspace;
  const size_t cbytes = (((size_t)pl.tiles_n * sizeof(unsigned)) + 255) & ~(size_t)255;
  float* partial = (float*)((unsigned char*)workspace + cbytes);
  const int chunk = 8 * pl.nt;
  for (int m0 = 0; m0 < M; m0 += chunk) {
    w4a16::Params p;
    p.w = (const uint8_t*)w;
    p.x = (const unsigned char*)x + (size_t)m0 * K * 2;
    p.sW = sW;
    p.bias = bias;
    p.out = (unsigned char*)out + (size_t)m0 * N * 2;
    p.partial = partial;
    p.counters = counters;
    p.M = (M - m0 < chunk) ? (M - m0) : chunk;
    p.N = N;
    p.K = K;
    p.splitK = pl.splitK;
    p.stages_per_split = pl.stages_per_split;
    p.stages_total = pl.stages_total;
    int rc;
    const bool bf = dtype == QSPEC_DTYPE_BF16;
    switch (pl.nt) {
      case 1: rc = bf ? w4a16::launch<1, true>(p, pl, st) : w4a16::launch<1, false>(p, pl, st); break;
      case 2: rc = bf ? w4a16::launch<2, true>(p, pl, st) : w4a16::launch<2, false>(p, pl, st); break;
      case 4: rc = bf ? w4a16::launch<4, true>(p, pl, st) : w4a16::launch<4, false>(p, pl, st); break;
      default: rc = bf ? w4a16::launch<8, true>(p, pl, st) : w4a16::launch<8, false>(p, pl, st); break;
    }
    if (rc != QSPEC_OK) return rc;
  }
  return QSPEC_OK;
}
