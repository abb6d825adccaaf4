// Activation-side kernels of the Qwen2.5-14B variant of the QSpec path (reference
// vllm/model_executor/models/qwen2_quarot.py): hidden 5120 = 40 heads x 128, intermediate 13824 = 108 x 128.
//   * down_proj rotation had108 (x) H_128 (qwen2_quarot.py:114-127 -> hadamard_utils.matmul_hadU_cuda, K = 108:
//     FHT_128 * 1/sqrt(13824) rounded to f16, then had108 @ . with fp32 accumulation, f16 out), optional SiLU*up prologue
//     in vLLM's SiluAndMul order [gate ; up] (:122) and the per-token int4 quantiser behind it (:125);
//   * o_proj rotation had40 across the heads (o_had_trans, :269-277, o_K = 40): f16(f16(had40 @ x) * 1/sqrt(40)) -- the
//     f16 matmul rounds once, the multiplication by the python float rounds again -- optional quantiser (:261);
//   * rotary embedding + the per-head FHT_128 of q and k (qk_had_trans, :279-288: fp32 FHT of the f16 rope output,
//     * 1/sqrt(128), f16) + KV-cache append.
// Numerics contract as everywhere in this library: explicit IEEE fp32 ops in a fixed order (butterflies in ascending stride,
// +-1 mixes accumulated in ascending k), bit-identical to oracle/qspec_oracle.py. sm_100a.
#include "common.cuh"

namespace qspec {

#include "hadk_masks.inc"

__device__ __forceinline__ float silu_mul_q(float g, float u) {
  const float s = __half2float(__float2half_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf(-g)))));
  return __half2float(__float2half_rn(__fmul_rn(s, u)));
}

// ------------------------------------------------------------------------------------------
// had108 (x) H_128: one 256-thread CTA per token row [108][128].
// Phase 1: 16 lanes per 128-block (2 blocks per warp pass): in-thread fht8, lane xor 1,2,4,8 -> strides 8..64; * scale,
// f16, into shared memory. Phase 2: thread (column pair jp = tid % 64, output group og = tid / 64) accumulates outputs
// o = 27 og .. 27 og + 26 over k = 0..107 in ascending order.
// ------------------------------------------------------------------------------------------
constexpr int QK = 108, QN = 128, QOPT = 27;

__global__ void __launch_bounds__(256)
hadamard_k108_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gate_up, int gate_first,
                     __half* __restrict__ out_f, uint8_t* __restrict__ out_q, __half* __restrict__ scale_out,
                     float scale, float clip) {
  pdl_enter();
  extern __shared__ __align__(16) unsigned char k108_dyn[];
  __half* sm = reinterpret_cast<__half*>(k108_dyn);                                       // [QK][QN] f16 rotated blocks (27 KB)
  float4* s_sign = reinterpret_cast<float4*>(k108_dyn + (size_t)QK * QN * sizeof(__half)); // [QK][4][7]: the +-1 multipliers (47 KB)
  __shared__ float red[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t row = blockIdx.x;
  constexpr int L = QK * QN, inter_v = L >> 3, nv = QN >> 3;      // 16 vectors per block
  // the sign table into shared memory (from constant memory it cost a scoreboard stall per k-step, profiles/r03i_qwen_notes.md);
  // it does not depend on the predecessor kernel, and is consumed after the __syncthreads() that closes phase 1
  for (int i = threadIdx.x; i < QK * 4 * 7; i += 256) s_sign[i] = __ldg(&d_had108_sign[0][0][0] + i);
  const uint4* xrow = x ? x + row * inter_v : nullptr;
  const uint4* gurow = gate_up ? gate_up + row * 2 * inter_v : nullptr;
  for (int u = warp; u < QK / 2; u += 8) {
    const int blk = u * 2 + (lane >> 4), v0 = lane & 15;
    const int vi = blk * nv + v0;
    float v[8];
    if (gurow) {
      float a[8], b[8];
      unpack_h8(__ldg(gurow + vi), a);
      unpack_h8(__ldg(gurow + inter_v + vi), b);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = gate_first ? silu_mul_q(a[i], b[i]) : silu_mul_q(b[i], a[i]);
    } else {
      unpack_h8(__ldg(xrow + vi), v);
    }
    fht8(v);
#pragma unroll
    for (int m = 1; m < 16; m <<= 1) fht_lane_stage(v, lane, m);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __fmul_rn(v[i], scale);
    reinterpret_cast<uint4*>(sm)[vi] = pack_h8(v);
  }
  __syncthreads();

  const int jp = threadIdx.x & 63, o0 = (threadIdx.x >> 6) * QOPT;
  // acc[o] += +-y in ascending k as one FFMA2 per (o, k) for the column pair (the sign is a +-1 multiplier from the table in shared
  // memory, seven 128-bit broadcast loads per k for the 27 outputs of this thread's group): bit-identical to the add / subtract chain, a fifth
  // of its instructions (r03g: the predicated-FADD form made this kernel ~26 us per row)
  float acc[QOPT][2];
  {
    const int og = threadIdx.x >> 6;
    unsigned long long a2[QOPT];
#pragma unroll
    for (int o = 0; o < QOPT; ++o) a2[o] = 0ull;
#pragma unroll 4
    for (int k = 0; k < QK; ++k) {
      const float2 y = __half22float2(reinterpret_cast<const __half2*>(sm)[(k * QN >> 1) + jp]);
      const unsigned long long y2 = pack_f2(y.x, y.y);
#pragma unroll
      for (int q = 0; q < 7; ++q) {
        const float4 sg = s_sign[(k * 4 + og) * 7 + q];
        a2[4 * q] = ffma2_rn(y2, pack_f2(sg.x, sg.x), a2[4 * q]);
        a2[4 * q + 1] = ffma2_rn(y2, pack_f2(sg.y, sg.y), a2[4 * q + 1]);
        a2[4 * q + 2] = ffma2_rn(y2, pack_f2(sg.z, sg.z), a2[4 * q + 2]);
        if (q < 6) a2[4 * q + 3] = ffma2_rn(y2, pack_f2(sg.w, sg.w), a2[4 * q + 3]);
      }
    }
#pragma unroll
    for (int o = 0; o < QOPT; ++o) {
      const float2 r = unpack_f2(a2[o]);
      acc[o][0] = r.x;
      acc[o][1] = r.y;
    }
  }
  float amax = 0.f;
#pragma unroll
  for (int o = 0; o < QOPT; ++o) {
    acc[o][0] = __half2float(__float2half_rn(acc[o][0]));
    acc[o][1] = __half2float(__float2half_rn(acc[o][1]));
    amax = fmaxf(amax, fmaxf(fabsf(acc[o][0]), fabsf(acc[o][1])));
    if (out_f) reinterpret_cast<__half2*>(out_f + row * L)[(((o0 + o) * QN) >> 1) + jp] = __floats2half2_rn(acc[o][0], acc[o][1]);
  }
  if (out_q == nullptr) return;
  amax = block_max(amax, red);
  const __half sh = sym_scale(amax, clip);
  const float sc = __half2float(sh);
#pragma unroll
  for (int o = 0; o < QOPT; ++o) {
    const int q0 = max(-8, min(7, __half2int_rn(__float2half_rn(__fdiv_rn(acc[o][0], sc)))));
    const int q1 = max(-8, min(7, __half2int_rn(__float2half_rn(__fdiv_rn(acc[o][1], sc)))));
    out_q[row * (L >> 1) + (((o0 + o) * QN) >> 1) + jp] = (uint8_t)((q0 & 0xF) | ((q1 & 0xF) << 4));
  }
  if (threadIdx.x == 0) scale_out[row] = sh;
}

// ------------------------------------------------------------------------------------------
// had40 across the heads: one 256-thread CTA per token row [40][128]. Thread (d = tid & 127, half = tid >> 7) computes the
// outputs h = 20 half .. 20 half + 19 of dim d, fp32 accumulate in ascending k; result f16(f16(acc) * scale).
// ------------------------------------------------------------------------------------------
constexpr int QH = 40, QD = 128;

__global__ void __launch_bounds__(256)
hadamard_heads40_kernel(const __half* __restrict__ x, uint4* __restrict__ out_f, uint32_t* __restrict__ out_q,
                        __half* __restrict__ scale_out, float scale, float clip) {
  pdl_enter();
  __shared__ __align__(16) __half sx[QH * QD];
  __shared__ float red[8];
  const size_t row = blockIdx.x;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * QH * QD);
  for (int i = threadIdx.x; i < QH * QD / 8; i += 256) reinterpret_cast<uint4*>(sx)[i] = __ldg(xr + i);
  __syncthreads();
  const int d = threadIdx.x & 127, h0 = (threadIdx.x >> 7) * 20;
  float acc[20];
#pragma unroll
  for (int o = 0; o < 20; ++o) acc[o] = 0.f;
  {
    // acc[o] += +-y in ascending k as one FFMA with a +-1 multiplier from constant memory (exactly the add / subtract)
    const int hg = threadIdx.x >> 7;
#pragma unroll 4
    for (int k = 0; k < QH; ++k) {
      const float y = __half2float(sx[k * QD + d]);
#pragma unroll
      for (int q = 0; q < 5; ++q) {
        const float4 sg = d_had40_sign[k][hg][q];
        acc[4 * q] = __fmaf_rn(y, sg.x, acc[4 * q]);
        acc[4 * q + 1] = __fmaf_rn(y, sg.y, acc[4 * q + 1]);
        acc[4 * q + 2] = __fmaf_rn(y, sg.z, acc[4 * q + 2]);
        acc[4 * q + 3] = __fmaf_rn(y, sg.w, acc[4 * q + 3]);
      }
    }
  }
  __syncthreads();                                       // every thread is done reading sx
#pragma unroll
  for (int o = 0; o < 20; ++o) {
    const float r1 = __half2float(__float2half_rn(acc[o]));            // the f16 matmul's output
    sx[(h0 + o) * QD + d] = __float2half_rn(__fmul_rn(r1, scale));     // * o_had_scale
  }
  __syncthreads();
  float amax = 0.f;
  for (int i = threadIdx.x; i < QH * QD / 8; i += 256) {
    const uint4 v = reinterpret_cast<const uint4*>(sx)[i];
    if (out_f) out_f[row * (QH * QD / 8) + i] = v;
    amax = habs_max8(v, amax);
  }
  if (out_q == nullptr) return;
  amax = block_max(amax, red);
  const __half sh = sym_scale(amax, clip);
  const float sc = __half2float(sh);
  for (int i = threadIdx.x; i < QH * QD / 8; i += 256)
    out_q[row * (QH * QD / 8) + i] = sym_quant8(reinterpret_cast<const uint4*>(sx)[i], sc);
  if (threadIdx.x == 0) scale_out[row] = sh;
}

// ------------------------------------------------------------------------------------------
// RoPE (neox) + FHT_128 of every q and k head + KV-cache append. One 256-thread CTA per token; 8 threads per head
// (thread j holds dims 8j..8j+7 of the lower half and 64+8j..64+8j+7 of the upper half). D == 128.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rope_qkhad_kvcache_kernel(__half* __restrict__ qkv, const __half* __restrict__ cos_sin, const int* __restrict__ positions,
                          const int* __restrict__ seq_ids, __half* __restrict__ cache_k, __half* __restrict__ cache_v,
                          int Hq, int Hkv, int S, float had_scale) {
  pdl_enter();
  constexpr int D = 128, HV = 8;
  const int tok = blockIdx.x;
  const int pos = positions[tok], seq = seq_ids[tok];
  if ((unsigned)pos >= (unsigned)S) return;
  __half* row = qkv + (size_t)tok * (Hq + 2 * Hkv) * D;
  const uint4* cs = reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * D);
  const int lane = threadIdx.x & 31;
  const int n_rot = (Hq + Hkv) * HV;
  for (int base = 0; base < n_rot; base += 256) {        // warp-uniform trip count: the shuffles below need every lane
    const int i = base + threadIdx.x;
    const bool active = i < n_rot;
    const int h = active ? i / HV : 0, j = i % HV;
    uint4* p1 = reinterpret_cast<uint4*>(row + (size_t)h * D) + j;
    uint4* p2 = p1 + HV;
    float x1[8], x2[8], c[8], s[8];
    if (active) {
      unpack_h8(*p1, x1); unpack_h8(*p2, x2);
      unpack_h8(__ldg(cs + j), c); unpack_h8(__ldg(cs + HV + j), s);
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) { x1[e] = x2[e] = 0.f; c[e] = 1.f; s[e] = 0.f; }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float a = __fsub_rn(__fmul_rn(x1[e], c[e]), __fmul_rn(x2[e], s[e]));
      const float b = __fadd_rn(__fmul_rn(x2[e], c[e]), __fmul_rn(x1[e], s[e]));
      x1[e] = __half2float(__float2half_rn(a));          // rotary_emb returns f16; the transform starts from that
      x2[e] = __half2float(__float2half_rn(b));
    }
    fht8(x1); fht8(x2);                                  // strides 1, 2, 4
#pragma unroll
    for (int m = 1; m < 8; m <<= 1) { fht_lane_stage(x1, lane, m); fht_lane_stage(x2, lane, m); }   // 8, 16, 32
#pragma unroll
    for (int e = 0; e < 8; ++e) {                        // stride 64
      const float a = x1[e], b = x2[e];
      x1[e] = __fmul_rn(__fadd_rn(a, b), had_scale);
      x2[e] = __fmul_rn(__fsub_rn(a, b), had_scale);
    }
    if (active) {
      const uint4 o1 = pack_h8(x1), o2 = pack_h8(x2);
      *p1 = o1; *p2 = o2;
      if (h >= Hq) {
        uint4* dst = reinterpret_cast<uint4*>(cache_k + (((size_t)seq * S + pos) * Hkv + (h - Hq)) * D) + j;
        dst[0] = o1; dst[HV] = o2;
      }
    }
  }
  const int n_v = Hkv * (D >> 3);
  for (int k = threadIdx.x; k < n_v; k += 256)
    reinterpret_cast<uint4*>(cache_v + ((size_t)seq * S + pos) * Hkv * D)[k] =
        reinterpret_cast<const uint4*>(row + (size_t)(Hq + Hkv) * D)[k];
}

}  // namespace qspec

using namespace qspec;

extern "C" int qspec_hadamard_k108(const void* x, const void* gate_up, int gate_first, void* out_f16, int8_t* out_q,
                                   void* scale_out, int rows, int n, float scale, float clip_ratio, void* stream) {
  QSPEC_REQUIRE((x || gate_up) && (out_f16 || out_q), QSPEC_ERR_INVALID_ARG, "hadamard_k108: null pointer");
  QSPEC_REQUIRE(!out_q || scale_out, QSPEC_ERR_INVALID_ARG, "hadamard_k108: out_q needs scale_out");
  QSPEC_REQUIRE(rows >= 0, QSPEC_ERR_INVALID_ARG, "hadamard_k108: rows=%d", rows);
  QSPEC_REQUIRE(n == QN, QSPEC_ERR_UNSUPPORTED, "hadamard_k108: n=%d, only 108 x 128 (intermediate size 13824) is built", n);
  if (rows == 0) return QSPEC_OK;
  constexpr size_t K108_SMEM = (size_t)QK * QN * sizeof(__half) + (size_t)QK * 4 * 7 * sizeof(float4);   // 27 KB + 47 KB
  static bool attr_set = false;
  if (!attr_set) {
    QSPEC_CUDA(cudaFuncSetAttribute(hadamard_k108_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K108_SMEM));
    attr_set = true;
  }
  launch_pdl(hadamard_k108_kernel, dim3(rows), dim3(256), K108_SMEM, (cudaStream_t)stream, (const uint4*)(gate_up ? nullptr : x),
             (const uint4*)gate_up, gate_first, (__half*)out_f16, (uint8_t*)out_q, (__half*)scale_out, scale, clip_ratio);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

extern "C" int qspec_hadamard_heads40(const void* x, void* out_f16, int8_t* out_q, void* scale_out, int rows, int H, int D,
                                      float scale, float clip_ratio, void* stream) {
  QSPEC_REQUIRE(x && (out_f16 || out_q), QSPEC_ERR_INVALID_ARG, "hadamard_heads40: null pointer");
  QSPEC_REQUIRE(!out_q || scale_out, QSPEC_ERR_INVALID_ARG, "hadamard_heads40: out_q needs scale_out");
  QSPEC_REQUIRE(rows >= 0, QSPEC_ERR_INVALID_ARG, "hadamard_heads40: rows=%d", rows);
  QSPEC_REQUIRE(H == QH && D == QD, QSPEC_ERR_UNSUPPORTED, "hadamard_heads40: only 40 heads x 128 is built (got %d x %d)", H, D);
  if (rows == 0) return QSPEC_OK;
  launch_pdl(hadamard_heads40_kernel, dim3(rows), dim3(256), 0, (cudaStream_t)stream, (const __half*)x, (uint4*)out_f16,
             (uint32_t*)out_q, (__half*)scale_out, scale, clip_ratio);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

extern "C" int qspec_rope_qkhad_kvcache(void* qkv, const void* cos_sin_cache, const int* positions, const int* seq_ids,
                                        void* cache_k, void* cache_v, int T, int Hq, int Hkv, int D, int S, float had_scale,
                                        void* stream) {
  QSPEC_REQUIRE(qkv && cos_sin_cache && positions && seq_ids && cache_k && cache_v, QSPEC_ERR_INVALID_ARG,
                "rope_qkhad_kvcache: null pointer");
  QSPEC_REQUIRE(T >= 0 && Hq > 0 && Hkv > 0 && S > 0, QSPEC_ERR_INVALID_ARG, "rope_qkhad_kvcache: bad sizes");
  QSPEC_REQUIRE(D == 128, QSPEC_ERR_UNSUPPORTED, "rope_qkhad_kvcache: head_dim %d, only 128 is built", D);
  if (T == 0) return QSPEC_OK;
  launch_pdl(rope_qkhad_kvcache_kernel, dim3(T), dim3(256), 0, (cudaStream_t)stream, (__half*)qkv, (const __half*)cos_sin_cache,
             positions, seq_ids, (__half*)cache_k, (__half*)cache_v, Hq, Hkv, S, had_scale);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}
