// Neighbours of the dual-mode linear path that the QSpec decode loop needs between qkv_proj and
// the o_proj rotation (vllm ops.rotary_embedding + reshape_and_cache +
// flash-attn in the reference, quarot_llama.py:207-221). Not part of the parity contract of the
// linear path: fp32 math, compared against a torch fp32 reference with a tolerance. sm_100a.
#include "common.cuh"

namespace qspec {

// RoPE (neox style, cos_sin_cache [max_pos, D] f16 = [cos(D/2) | sin(D/2)]) on q and k, and
// append of the rotated k and of v to the contiguous KV cache [B, S, Hkv, D].
// qkv [T, (Hq + 2 Hkv) * D]; q is rotated in place.
__global__ void __launch_bounds__(256)
rope_kvcache_kernel(__half* __restrict__ qkv, const __half* __restrict__ cos_sin,
                    const int* __restrict__ positions, const int* __restrict__ seq_ids,
                    __half* __restrict__ cache_k, __half* __restrict__ cache_v,
                    int Hq, int Hkv, int D, int S) {
  pdl_enter();
  const int tok = blockIdx.x;
  const int pos = positions[tok], seq = seq_ids[tok];
  const int half_v = D >> 4;                              // 8-element vectors per half head
  __half* row = qkv + (size_t)tok * (Hq + 2 * Hkv) * D;
  const uint4* cs = reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * D);
  const int n_rot = (Hq + Hkv) * half_v;
  const int n_v = Hkv * (D >> 3);
  for (int i = threadIdx.x; i < n_rot + n_v; i += blockDim.x) {
    if (i < n_rot) {
      const int h = i / half_v, j = i % half_v;
      uint4* p1 = reinterpret_cast<uint4*>(row + (size_t)h * D) + j;
      uint4* p2 = p1 + half_v;
      float x1[8], x2[8], c[8], s[8];
      unpack_h8(*p1, x1); unpack_h8(*p2, x2);
      unpack_h8(__ldg(cs + j), c); unpack_h8(__ldg(cs + half_v + j), s);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float a = __fsub_rn(__fmul_rn(x1[e], c[e]), __fmul_rn(x2[e], s[e]));
        const float b = __fadd_rn(__fmul_rn(x2[e], c[e]), __fmul_rn(x1[e], s[e]));
        x1[e] = a; x2[e] = b;
      }
      const uint4 o1 = pack_h8(x1), o2 = pack_h8(x2);
      if (h < Hq) { *p1 = o1; *p2 = o2; }
      else {
        uint4* dst = reinterpret_cast<uint4*>(cache_k + (((size_t)seq * S + pos) * Hkv + (h - Hq)) * D) + j;
        dst[0] = o1; dst[half_v] = o2;
        *p1 = o1; *p2 = o2;
      }
    } else {
      const int k = i - n_rot;
      const uint4 v = reinterpret_cast<const uint4*>(row + (size_t)(Hq + Hkv) * D)[k];
      reinterpret_cast<uint4*>(cache_v + ((size_t)seq * S + pos) * Hkv * D)[k] = v;
    }
  }
}

// Decode attention: one 128-thread CTA per (token, q head); keys 0..pos of the token's sequence.
// D == 128. scores live in dynamic shared memory (S floats).
__global__ void __launch_bounds__(128)
decode_attention_kernel(const __half* __restrict__ qkv, const __half* __restrict__ cache_k,
                        const __half* __restrict__ cache_v, const int* __restrict__ positions,
                        const int* __restrict__ seq_ids, __half* __restrict__ out,
                        int Hq, int Hkv, int S, float sm_scale) {
  pdl_enter();
  constexpr int D = 128;
  extern __shared__ float scores[];                       // [S]
  __shared__ float qs[D];
  __shared__ float red[4];
  __shared__ float accs[4][D];
  const int tok = blockIdx.x, h = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int pos = positions[tok], seq = seq_ids[tok];
  const int nkeys = pos + 1;
  const int hk = h / (Hq / Hkv);
  qs[tid] = __half2float(qkv[(size_t)tok * (Hq + 2 * Hkv) * D + (size_t)h * D + tid]) * sm_scale;
  __syncthreads();
  const __half* kbase = cache_k + ((size_t)seq * S * Hkv + hk) * D;
  const __half* vbase = cache_v + ((size_t)seq * S * Hkv + hk) * D;
  float lmax = -INFINITY;
  for (int j = tid; j < nkeys; j += 128) {
    const uint4* kr = reinterpret_cast<const uint4*>(kbase + (size_t)j * Hkv * D);
    float dot = 0.f;
#pragma unroll
    for (int v = 0; v < D / 8; ++v) {
      float f[8];
      unpack_h8(__ldg(kr + v), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) dot = fmaf(f[e], qs[v * 8 + e], dot);
    }
    scores[j] = dot;
    lmax = fmaxf(lmax, dot);
  }
  lmax = warp_xor_max(lmax);
  if (lane == 0) red[warp] = lmax;
  __syncthreads();
  const float gmax = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  float lsum = 0.f;
  for (int j = tid; j < nkeys; j += 128) {
    const float p = __expf(scores[j] - gmax);
    scores[j] = p;
    lsum += p;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) lsum += __shfl_xor_sync(0xffffffffu, lsum, off);
  __syncthreads();                                        // red reuse + scores visible
  if (lane == 0) red[warp] = lsum;
  __syncthreads();
  const float inv = 1.0f / (red[0] + red[1] + red[2] + red[3]);
  // P @ V: warp w takes keys w, w+4, ...; lane owns dims 4*lane .. 4*lane+3
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  for (int j = warp; j < nkeys; j += 4) {
    const uint2 raw = __ldg(reinterpret_cast<const uint2*>(vbase + (size_t)j * Hkv * D) + lane);
    const float2 v01 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
    const float2 v23 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
    const float p = scores[j];
    a0 = fmaf(p, v01.x, a0); a1 = fmaf(p, v01.y, a1); a2 = fmaf(p, v23.x, a2); a3 = fmaf(p, v23.y, a3);
  }
  accs[warp][4 * lane + 0] = a0; accs[warp][4 * lane + 1] = a1;
  accs[warp][4 * lane + 2] = a2; accs[warp][4 * lane + 3] = a3;
  __syncthreads();
  const float o = (accs[0][tid] + accs[1][tid] + accs[2][tid] + accs[3][tid]) * inv;
  out[((size_t)tok * Hq + h) * D + tid] = __float2half_rn(o);
}

}  // namespace qspec

using namespace qspec;

extern "C" int qspec_rope_kvcache(void* qkv, const void* cos_sin_cache, const int* positions,
                                  const int* seq_ids, void* cache_k, void* cache_v, int T, int Hq,
                                  int Hkv, int D, int S, void* stream) {
  QSPEC_REQUIRE(qkv && cos_sin_cache && positions && seq_ids && cache_k && cache_v, QSPEC_ERR_INVALID_ARG,
                "rope_kvcache: null pointer");
  QSPEC_REQUIRE(T >= 0 && Hq > 0 && Hkv > 0 && S > 0, QSPEC_ERR_INVALID_ARG, "rope_kvcache: bad sizes");
  QSPEC_REQUIRE(D % 16 == 0, QSPEC_ERR_UNSUPPORTED, "rope_kvcache: head_dim %d must be a multiple of 16", D);
  if (T == 0) return QSPEC_OK;
  launch_pdl(rope_kvcache_kernel, dim3(T), dim3(256), 0, (cudaStream_t)stream, (__half*)qkv, (const __half*)cos_sin_cache, positions,
                                                           seq_ids, (__half*)cache_k, (__half*)cache_v, Hq, Hkv, D, S);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

extern "C" int qspec_decode_attention(const void* qkv, const void* cache_k, const void* cache_v,
                                      const int* positions, const int* seq_ids, void* out, int T,
                                      int Hq, int Hkv, int D, int S, float sm_scale, void* stream) {
  QSPEC_REQUIRE(qkv && cache_k && cache_v && positions && seq_ids && out, QSPEC_ERR_INVALID_ARG,
                "decode_attention: null pointer");
  QSPEC_REQUIRE(D == 128 && Hq % Hkv == 0, QSPEC_ERR_UNSUPPORTED, "decode_attention: need head_dim 128 and Hq %% Hkv == 0");
  QSPEC_REQUIRE(S > 0 && S <= 8192, QSPEC_ERR_UNSUPPORTED, "decode_attention: cache length %d > 8192", S);
  if (T == 0) return QSPEC_OK;
  dim3 grid(T, Hq);
  launch_pdl(decode_attention_kernel, dim3(grid), dim3(128), (size_t)S * sizeof(float), (cudaStream_t)stream, 
      (const __half*)qkv, (const __half*)cache_k, (const __half*)cache_v, positions, seq_ids, (__half*)out, Hq, Hkv, S,
      sm_scale);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}
