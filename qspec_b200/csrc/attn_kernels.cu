// Neighbours of the dual-mode linear path that the QSpec decode loop needs between qkv_proj and
// the o_proj rotation (vllm ops.rotary_embedding + reshape_and_cache +
// flash-attn in the reference, quarot_llama.py:207-221). Not part of the parity contract of the
// linear path: fp32 math, compared against a torch fp32 reference with a tolerance. sm_100a.
#include "common.cuh"
#include <stdlib.h>
#include <mutex>

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
  if ((unsigned)pos >= (unsigned)S) return;               // last resort: the caller bounds positions by the cache length
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
  if ((unsigned)pos >= (unsigned)S) return;               // out of the cache / rotary table: nothing valid to attend to
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

// ------------------------------------------------------------------------------------------
// The three kernels above plus the o_proj rotation of hadamard_kernels.cu as ONE launch for decode-sized passes
// (quarot_llama.py:207-238: rotary embedding, cache append, attention, Hadamard over heads, quantiser).
// Grid = T x Hkv CTAs of 128 threads, two CTAs per SM; the Hkv CTAs of a token form a thread-block cluster.
// CTA (token, kv head):
//   0. one round trip for the positions, then everything that only depends on them is put in flight at once:
//      the key rows of the first 128 positions (one per thread, registers), bulk copies of the remaining key rows
//      and of the value rows into shared memory (cp.async.bulk, one 256-byte copy per row), the rotary inputs;
//      rotates the token's G = Hq/Hkv query heads (-> shared memory, qkv is not written) and the key of EVERY
//      token of this pass that belongs to the same sequence at a position <= its own (keys/values written to the
//      cache by other CTAs of this launch are not visible yet, so they are recomputed: same arithmetic, same
//      bits); its own key/value go to the cache;
//   1. scores for all G heads with one thread per key (the key row is read once for the whole GQA group), exact
//      softmax, P.V with one warp per key from the staged value rows (restaged in chunks if they do not fit);
//   2. pushes its G heads (f16) into the CTA of the cluster that owns the dim (distributed shared memory), one
//      cluster barrier, then each CTA runs the Hq-point FHT for its 128/Hkv dims in registers;
//   3. quantiser: CTA amax pushed to every CTA of the cluster, one more cluster barrier, pack.
// Every floating-point operation is the one the separate kernels do, in the same order (thread <-> key mapping,
// reduction trees, butterfly order), so the outputs are bit-identical to the four-kernel chain.
// ------------------------------------------------------------------------------------------
constexpr int ST_K_PITCH = 272;            // staged key rows are padded by 16 bytes: thread-per-key reads hit all banks
static size_t attn_block_smem_budget() {          // dynamic shared memory per CTA (QSPEC_ATTN_SMEM_KB overrides)
  static const size_t kb = [] {
    const char* e = getenv("QSPEC_ATTN_SMEM_KB");
    const long v = e ? atol(e) : 0;
    return (size_t)(v >= 32 && v <= 200 ? v : 96);  // 96 KB + static: two CTAs per SM, so 8-CTA clusters of a 16-token pass fit in one wave
  }();
  return kb * 1024;
}

// dynamic shared memory without the staging area
static inline size_t attn_block_fixed_smem(int G, int T, int S) {
  const size_t S8 = ((size_t)S + 7) & ~(size_t)7;
  return (size_t)G * S8 * sizeof(float) + (size_t)T * 128 * 2 * sizeof(__half) + S8 * sizeof(int16_t) +
         2 * (size_t)((T + 3) & ~3) * sizeof(int);
}
// bytes left for staging cache rows, or -1 when the fixed part does not leave at least 16 KB
static inline long attn_block_stage_bytes(int G, int T, int S) {
  const size_t fixed = attn_block_fixed_smem(G, T, S), budget = attn_block_smem_budget();
  if (fixed + 16 * 1024 > budget) return -1;
  return (long)((budget - fixed) & ~(size_t)15);
}

__device__ __forceinline__ void bulk_row_256(uint32_t dst, const void* src, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 256, [%2];\n"
               ::"r"(dst), "l"(src), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait_parity(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  }
}
// neox rotary embedding of 8 + 8 elements (chunk j of the lower and of the upper half of a head)
__device__ __forceinline__ void rope16(const uint4* __restrict__ p1, const uint4* __restrict__ cs, int j, int HV,
                                       float (&x1)[8], float (&x2)[8]) {
  float c[8], sn[8];
  unpack_h8(__ldg(p1), x1); unpack_h8(__ldg(p1 + HV), x2);
  unpack_h8(__ldg(cs + j), c); unpack_h8(__ldg(cs + HV + j), sn);
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float a = __fsub_rn(__fmul_rn(x1[e], c[e]), __fmul_rn(x2[e], sn[e]));
    const float b = __fadd_rn(__fmul_rn(x2[e], c[e]), __fmul_rn(x1[e], sn[e]));
    x1[e] = a; x2[e] = b;
  }
}

template <int G, int HKV>
__global__ void __launch_bounds__(128, 2)
attn_block_kernel(const __half* __restrict__ qkv, const __half* __restrict__ cos_sin,
                  const int* __restrict__ positions, const int* __restrict__ seq_ids,
                  __half* __restrict__ cache_k, __half* __restrict__ cache_v, __half* __restrict__ attn_out,
                  uint4* __restrict__ out_f, uint32_t* __restrict__ out_q, __half* __restrict__ scale_out,
                  int T, int S, int stage_bytes, float sm_scale, float had_scale, float clip) {
  pdl_enter();
  cluster_arrive_relaxed();                                                         // "this CTA has started" (waited for before the first push)
  constexpr int D = 128, H = G * HKV, DPC = D / HKV, HV = D / 16;
  extern __shared__ __align__(16) unsigned char dyn[];
  const int S8 = (S + 7) & ~7, T4 = (T + 3) & ~3;
  float* scores = reinterpret_cast<float*>(dyn);                                   // [G][S8]
  __half* inK = reinterpret_cast<__half*>(dyn + (size_t)G * S8 * sizeof(float));   // [T][D] rotated keys of this pass
  __half* inV = inK + (size_t)T * D;                                               // [T][D]
  int16_t* slot = reinterpret_cast<int16_t*>(inV + (size_t)T * D);                 // [S8]: token of this pass at a position, or -1
  int* inpos = reinterpret_cast<int*>(slot + S8);                                  // [T4]: position if the token is a key here, or -1
  int* inlist = inpos + T4;                                                        // [T4]: the tokens that are keys here
  unsigned char* stage = reinterpret_cast<unsigned char*>(inlist + T4);            // staged key rows, then staged value rows
  __shared__ __align__(16) float qs[G][D];
  __shared__ float accs[4][G][D];
  __shared__ float red[G][4];
  __shared__ float red1[8];
  __shared__ __align__(16) __half gath[H][DPC];                                    // all heads x this CTA's dims (pushed by the cluster)
  __shared__ float amax_all[HKV];
  __shared__ int n_in;
  __shared__ __align__(8) uint64_t stage_bar;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tok = blockIdx.x / HKV, hk = blockIdx.x % HKV;                          // hk == rank in the cluster
  // ---- round trip 1: this token's position and sequence, and those of the other tokens of the pass
  const int pos = __ldg(positions + tok), seq = __ldg(seq_ids + tok);
  // Last resort (QSpecEngine bounds the positions on the host): a position outside the cache / rotary table would index
  // shared and global memory out of range. The Hkv CTAs of the token's cluster all see it and leave together.
  if ((unsigned)pos >= (unsigned)S) return;
  int p_first = -1, s_first = -1;
  if (tid < T) { p_first = __ldg(positions + tid); s_first = __ldg(seq_ids + tid); }
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&stage_bar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    n_in = 0;
  }
  for (int i = tid; i < S8 / 8; i += 128) reinterpret_cast<uint4*>(slot)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);

  const int nkeys = pos + 1;
  const size_t row_elems = (size_t)(H + 2 * HKV) * D;
  const __half* kbase = cache_k + ((size_t)seq * S * HKV + hk) * D;
  const __half* vbase = cache_v + ((size_t)seq * S * HKV + hk) * D;
  // staging plan: key rows 128.. (the first 128 go to registers) get at most half of the area, value rows the rest
  const int kck = min(max(nkeys - 128, 0), (stage_bytes / 2) / ST_K_PITCH);
  const int kcv = min(nkeys, (stage_bytes - kck * ST_K_PITCH) / 256);
  unsigned char* stK = stage;
  unsigned char* stV = stage + (size_t)kck * ST_K_PITCH;
  if (tid == 0)
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"((uint32_t)(kck + kcv) * 256u) : "memory");
  __syncthreads();

  // ---- everything that depends on the position only, in flight together
  for (int j = tid; j < kcv; j += 128)
    bulk_row_256((uint32_t)__cvta_generic_to_shared(stV + (size_t)j * 256), vbase + (size_t)j * HKV * D, bar);
  for (int j = tid; j < kck; j += 128)
    bulk_row_256((uint32_t)__cvta_generic_to_shared(stK + (size_t)j * ST_K_PITCH), kbase + (size_t)(128 + j) * HKV * D, bar);
  uint4 kreg[D / 8];
  if (tid < nkeys) {
    const uint4* kr = reinterpret_cast<const uint4*>(kbase + (size_t)tid * HKV * D);
#pragma unroll
    for (int v = 0; v < D / 8; ++v) kreg[v] = __ldg(kr + v);
  }
  // which tokens of this pass are keys of this token
  for (int t2 = tid; t2 < T; t2 += 128) {
    const int p2 = t2 < 128 ? p_first : __ldg(positions + t2);
    const int s2 = t2 < 128 ? s_first : __ldg(seq_ids + t2);
    const bool key = s2 == seq && p2 >= 0 && p2 <= pos;
    inpos[t2] = key ? p2 : -1;
    if (key) {
      slot[p2] = (int16_t)t2;
      inlist[atomicAdd(&n_in, 1)] = t2;
    }
  }
  // rotary embedding of this token's query heads -> qs (f16-rounded like the in-place kernel, then * sm_scale)
  if (tid < G * HV) {
    const int g = tid / HV, j = tid % HV;
    float x1[8], x2[8];
    rope16(reinterpret_cast<const uint4*>(qkv + (size_t)tok * row_elems + (size_t)(hk * G + g) * D) + j,
           reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * D), j, HV, x1, x2);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      qs[g][j * 8 + e] = __fmul_rn(__half2float(__float2half_rn(x1[e])), sm_scale);
      qs[g][D / 2 + j * 8 + e] = __fmul_rn(__half2float(__float2half_rn(x2[e])), sm_scale);
    }
  }
  // this token's own key and value (always a key of itself) in the same round trip as the queries: warp 2 rotates
  // the key (lanes 0..7), warp 3 copies the value (lanes 0..15); both go to the cache and to the pass-local arrays
  if (warp == 2 && lane < HV) {
    const int j = lane;
    float x1[8], x2[8];
    rope16(reinterpret_cast<const uint4*>(qkv + (size_t)tok * row_elems + (size_t)(H + hk) * D) + j,
           reinterpret_cast<const uint4*>(cos_sin + (size_t)pos * D), j, HV, x1, x2);
    const uint4 o1 = pack_h8(x1), o2 = pack_h8(x2);
    uint4* dk = reinterpret_cast<uint4*>(inK + (size_t)tok * D) + j;
    dk[0] = o1; dk[HV] = o2;
    uint4* dst = reinterpret_cast<uint4*>(cache_k + (((size_t)seq * S + pos) * HKV + hk) * D) + j;
    dst[0] = o1; dst[HV] = o2;
  } else if (warp == 3 && lane < D / 8) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(qkv + (size_t)tok * row_elems + (size_t)(H + HKV + hk) * D) + lane);
    reinterpret_cast<uint4*>(inV + (size_t)tok * D)[lane] = v;
    reinterpret_cast<uint4*>(cache_v + (((size_t)seq * S + pos) * HKV + hk) * D)[lane] = v;
  }
  __syncthreads();
  // keys/values of the OTHER tokens of this pass that precede this one in its sequence (verify / prefill passes):
  // one warp per token; lanes 0..7 rotate the key, lanes 8..23 copy the value
  for (int i = warp; i < n_in; i += 4) {
    const int t2 = inlist[i];
    if (t2 == tok) continue;                                                        // warp-uniform; done above
    const int p2 = inpos[t2];
    const __half* row2 = qkv + (size_t)t2 * row_elems;
    if (lane < HV) {
      const int j = lane;
      float x1[8], x2[8];
      rope16(reinterpret_cast<const uint4*>(row2 + (size_t)(H + hk) * D) + j,
             reinterpret_cast<const uint4*>(cos_sin + (size_t)p2 * D), j, HV, x1, x2);
      uint4* dk = reinterpret_cast<uint4*>(inK + (size_t)t2 * D) + j;
      dk[0] = pack_h8(x1); dk[HV] = pack_h8(x2);
    } else if (lane < HV + D / 8) {
      const int k = lane - HV;
      reinterpret_cast<uint4*>(inV + (size_t)t2 * D)[k] = __ldg(reinterpret_cast<const uint4*>(row2 + (size_t)(H + HKV + hk) * D) + k);
    }
  }
  __syncthreads();
  mbar_wait_parity(bar, 0);

  // ---- 1a. scores: one thread per key, all G heads at once
  float lmax[G];
#pragma unroll
  for (int g = 0; g < G; ++g) lmax[g] = -INFINITY;
  for (int j = tid; j < nkeys; j += 128) {
    const int sl = slot[j];
    if (sl >= 0 || j >= 128) {
      const uint4* kr = sl >= 0 ? reinterpret_cast<const uint4*>(inK + (size_t)sl * D)
                      : j - 128 < kck ? reinterpret_cast<const uint4*>(stK + (size_t)(j - 128) * ST_K_PITCH)
                                      : reinterpret_cast<const uint4*>(kbase + (size_t)j * HKV * D);
#pragma unroll
      for (int v = 0; v < D / 8; ++v) kreg[v] = kr[v];
    }
    float dot[G];
#pragma unroll
    for (int g = 0; g < G; ++g) dot[g] = 0.f;
#pragma unroll
    for (int v = 0; v < D / 8; ++v) {
      float f[8];
      unpack_h8(kreg[v], f);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const float4 q0 = *reinterpret_cast<const float4*>(&qs[g][v * 8]);
        const float4 q1 = *reinterpret_cast<const float4*>(&qs[g][v * 8 + 4]);
        dot[g] = fmaf(f[0], q0.x, dot[g]); dot[g] = fmaf(f[1], q0.y, dot[g]);
        dot[g] = fmaf(f[2], q0.z, dot[g]); dot[g] = fmaf(f[3], q0.w, dot[g]);
        dot[g] = fmaf(f[4], q1.x, dot[g]); dot[g] = fmaf(f[5], q1.y, dot[g]);
        dot[g] = fmaf(f[6], q1.z, dot[g]); dot[g] = fmaf(f[7], q1.w, dot[g]);
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      scores[(size_t)g * S8 + j] = dot[g];
      lmax[g] = fmaxf(lmax[g], dot[g]);
    }
  }
#pragma unroll
  for (int g = 0; g < G; ++g) {
    lmax[g] = warp_xor_max(lmax[g]);
    if (lane == 0) red[g][warp] = lmax[g];
  }
  __syncthreads();
  // ---- 1b. softmax numerators and their sums
  float lsum[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const float gmax = fmaxf(fmaxf(red[g][0], red[g][1]), fmaxf(red[g][2], red[g][3]));
    lsum[g] = 0.f;
    for (int j = tid; j < nkeys; j += 128) {
      const float pexp = __expf(scores[(size_t)g * S8 + j] - gmax);
      scores[(size_t)g * S8 + j] = pexp;
      lsum[g] += pexp;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) lsum[g] += __shfl_xor_sync(0xffffffffu, lsum[g], off);
  }
  __syncthreads();
#pragma unroll
  for (int g = 0; g < G; ++g)
    if (lane == 0) red[g][warp] = lsum[g];
  __syncthreads();
  // ---- 1c. P.V: warp w takes keys w, w+4, ...; lane owns dims 4*lane .. 4*lane+3. Value rows come from the staged
  // chunk [c0, c1); the next chunk is staged when this one is done.
  float acc[G][4];
#pragma unroll
  for (int g = 0; g < G; ++g) { acc[g][0] = 0.f; acc[g][1] = 0.f; acc[g][2] = 0.f; acc[g][3] = 0.f; }
  for (int c0 = 0, chunk = 0; c0 < nkeys; c0 += kcv, ++chunk) {
    const int c1 = min(c0 + kcv, nkeys);
    if (chunk > 0) {
      __syncthreads();                                                              // the previous chunk has been consumed
      if (tid == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"((uint32_t)(c1 - c0) * 256u) : "memory");
      for (int j = c0 + tid; j < c1; j += 128)
        bulk_row_256((uint32_t)__cvta_generic_to_shared(stV + (size_t)(j - c0) * 256), vbase + (size_t)j * HKV * D, bar);
      mbar_wait_parity(bar, chunk & 1);
    }
#pragma unroll 4
    for (int j = c0 + ((warp - c0) & 3); j < c1; j += 4) {
      const int sl = slot[j];
      const uint2* vr = sl >= 0 ? reinterpret_cast<const uint2*>(inV + (size_t)sl * D)
                                : reinterpret_cast<const uint2*>(stV + (size_t)(j - c0) * 256);
      const uint2 raw = vr[lane];
      const float2 v01 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
      const float2 v23 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const float pj = scores[(size_t)g * S8 + j];
        acc[g][0] = fmaf(pj, v01.x, acc[g][0]); acc[g][1] = fmaf(pj, v01.y, acc[g][1]);
        acc[g][2] = fmaf(pj, v23.x, acc[g][2]); acc[g][3] = fmaf(pj, v23.y, acc[g][3]);
      }
    }
  }
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int i = 0; i < 4; ++i) accs[warp][g][4 * lane + i] = acc[g][i];
  }
  __syncthreads();

  // ---- 2. Hadamard over the heads: dim d of every head goes to the CTA that owns it (rank d / DPC)
  cluster_wait_acquire();                                                           // every CTA of the cluster is running
  {
    const uint32_t gath_addr = (uint32_t)__cvta_generic_to_shared(&gath[0][0]);
    const uint32_t owner = (uint32_t)(tid / DPC), col = (uint32_t)(tid % DPC);
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float inv = 1.0f / (red[g][0] + red[g][1] + red[g][2] + red[g][3]);
      const float o = (accs[0][g][tid] + accs[1][g][tid] + accs[2][g][tid] + accs[3][g][tid]) * inv;
      const __half oh = __float2half_rn(o);
      if (attn_out) attn_out[((size_t)tok * H + hk * G + g) * D + tid] = oh;
      st_dsmem_u16(mapa_shared(gath_addr + (uint32_t)(((hk * G + g) * DPC + col) * 2), owner), __half_as_ushort(oh));
    }
  }
  cluster_arrive_release();
  cluster_wait_acquire();
  float amax = 0.f;
  if (tid < DPC) {
    float v[H];
#pragma unroll
    for (int h = 0; h < H; ++h) v[h] = __half2float(gath[h][tid]);
#pragma unroll
    for (int bit = 1; bit < H; bit <<= 1) {
#pragma unroll
      for (int h = 0; h < H; ++h) {
        if ((h & bit) == 0) {
          const float a = v[h], b = v[h | bit];
          v[h] = __fadd_rn(a, b);
          v[h | bit] = __fsub_rn(a, b);
        }
      }
    }
#pragma unroll
    for (int h = 0; h < H; ++h) {
      const __half r = __float2half_rn(__fmul_rn(v[h], had_scale));
      gath[h][tid] = r;                                                             // in place: this thread owns the column
      amax = fmaxf(amax, fabsf(__half2float(r)));
    }
  }
  __syncthreads();
  const size_t row_v = (size_t)tok * H * (D / 8);
  constexpr int VPH = DPC / 8;                                                      // 16-byte vectors per head here
  if (out_f) {
    for (int i = tid; i < H * VPH; i += 128) {
      const int h = i / VPH, dc = i % VPH;
      out_f[row_v + (size_t)h * (D / 8) + hk * VPH + dc] = *reinterpret_cast<const uint4*>(&gath[h][dc * 8]);
    }
  }
  if (out_q) {
    // ---- 3. quantiser: row amax = max over the CTAs of the cluster
    amax = block_max(amax, red1);
    if (tid < HKV) st_dsmem_u32(mapa_shared((uint32_t)__cvta_generic_to_shared(&amax_all[hk]), (uint32_t)tid), __float_as_uint(amax));
    cluster_arrive_release();
    cluster_wait_acquire();
    float row_amax = 0.f;
#pragma unroll
    for (int r = 0; r < HKV; ++r) row_amax = fmaxf(row_amax, amax_all[r]);
    const __half sh = sym_scale(row_amax, clip);
    const float sc = __half2float(sh);
    for (int i = tid; i < H * VPH; i += 128) {
      const int h = i / VPH, dc = i % VPH;
      out_q[row_v + (size_t)h * (D / 8) + hk * VPH + dc] = sym_quant8(*reinterpret_cast<const uint4*>(&gath[h][dc * 8]), sc);
    }
    if (hk == 0 && tid == 0) scale_out[tok] = sh;
  }
}

template <int G, int HKV>
static int launch_attn_block(const void* qkv, const void* cos_sin, const int* positions, const int* seq_ids, void* ck,
                             void* cv, void* attn_out, void* out_f, int8_t* out_q, void* scale_out, int T, int S,
                             float sm_scale, float had_scale, float clip, cudaStream_t st) {
  auto kern = attn_block_kernel<G, HKV>;
  const long stage_bytes = attn_block_stage_bytes(G, T, S);
  QSPEC_REQUIRE(stage_bytes >= 0, QSPEC_ERR_UNSUPPORTED, "attn_block: T=%d tokens x cache length %d does not fit in shared memory", T, S);
  const size_t dyn = attn_block_fixed_smem(G, T, S) + (size_t)stage_bytes;
  const int kcap = (int)stage_bytes;
  static bool attr_set = false;                        // per instantiation
  if (!attr_set) {
    QSPEC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_block_smem_budget()));
    attr_set = true;
    if (getenv("QSPEC_ATTN_DEBUG")) {
      cudaLaunchConfig_t c2 = {};
      c2.gridDim = dim3((unsigned)T * HKV); c2.blockDim = dim3(128); c2.dynamicSmemBytes = dyn;
      cudaLaunchAttribute a2[1];
      a2[0].id = cudaLaunchAttributeClusterDimension;
      a2[0].val.clusterDim.x = HKV; a2[0].val.clusterDim.y = 1; a2[0].val.clusterDim.z = 1;
      c2.attrs = a2; c2.numAttrs = 1;
      int nc = -1;
      cudaOccupancyMaxActiveClusters(&nc, kern, &c2);
      fprintf(stderr, "attn_block<%d,%d>: dyn smem %zu, staging %d bytes, max active clusters %d\n", G, HKV, dyn, kcap, nc);
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)T * HKV);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = dyn;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = HKV;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  QSPEC_CUDA(cudaLaunchKernelEx(&cfg, kern, (const __half*)qkv, (const __half*)cos_sin, positions, seq_ids, (__half*)ck,
                                (__half*)cv, (__half*)attn_out, (uint4*)out_f, (uint32_t*)out_q, (__half*)scale_out, T, S,
                                kcap, sm_scale, had_scale, clip));
  count_launch(1);
  return QSPEC_OK;
}

// ------------------------------------------------------------------------------------------
// Attention for passes with MANY sequences (BASELINE configs[2]/[3]: 64 / 32 concurrent sequences; verify passes of
// (k + 1) tokens per sequence): one 128-thread CTA per (sequence, kv head) reads that head's key and value rows ONCE
// for all tokens of the sequence in the pass and all G query heads of the GQA group (decode_attention_kernel reads them
// once per (token, query head): 16x the cache traffic at G = 4, k + 1 = 4), and does the arithmetic on the tensor
// cores (mma.sync m16n8k16, f16 operands, fp32 accumulate) -- on CUDA cores the same loop is issue-bound at ~45-70 us
// for 64 sequences (profiles/r02g_notes.md).
//   * 16 query rows per chunk = R tokens x G heads (R = 16 / G). The 4 warps split the keys; every warp runs a
//     flash-decoding loop over 16-key blocks with an online softmax, then the warps' (max, sum, O) are merged in shared memory.
//   * No shared-memory staging of K / V: the fragments are loaded straight from the cache with 128-bit loads by permuting
//     the reduction index (QK^T: lane (gid, tig) takes dims [32 tig, 32 tig + 32) of key row gid -- both operands use the
//     same permutation, so the dot products are unchanged) and the output index (P.V: n-tile t of lane gid is dim
//     16 gid + t, so a lane needs 32 contiguous bytes of 4 value rows and ends up holding 32 contiguous output dims).
//   * Accuracy is fp32-like although the operands are f16: q and k ARE f16 (products exact, fp32 accumulate; the softmax
//     scale is applied to the fp32 scores), and the probabilities enter P.V as a two-term f16 split P = P_hi + P_lo.
// qkv holds the ROTATED queries (rope_kvcache ran before), the cache already contains this pass's keys / values.
// Outside the bit-exact contract: compared with a torch fp32 reference (tolerance in the test).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
  const __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}
// (hi, lo) f16 split of two fp32 values: hi = f16(x), lo = f16(x - hi)
__device__ __forceinline__ void split_h2(float x, float y, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(x, y);
  const float2 hf = __half22float2(h);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = pack_h2(x - hf.x, y - hf.y);
}

constexpr int SEQ_ATTN_WARPS = 4;                          // warps of a CTA that split the keys of one (sequence, kv head)

template <int G>
__global__ void __launch_bounds__(SEQ_ATTN_WARPS * 32, 4)
seq_attention_kernel(const __half* __restrict__ qkv, const __half* __restrict__ cache_k, const __half* __restrict__ cache_v,
                     const int* __restrict__ positions, const int* __restrict__ seq_ids, __half* __restrict__ out,
                     int T, int Hq, int Hkv, int S, float sm_scale) {
  pdl_enter();
  constexpr int D = 128, R = 16 / G, NW = SEQ_ATTN_WARPS;    // R tokens x G heads <= 16 query rows per chunk; NW warps split the keys
  __shared__ int s_tok[1024];                                // tokens of this sequence in the pass (T <= 1024)
  __shared__ int s_pos[1024];                                // ... and their positions (loaded in the same round trip as the scan)
  __shared__ int s_ntok;
  __shared__ float s_m[NW][16], s_l[NW][16];
  __shared__ __align__(16) uint4 s_qf[8][32];                // Q fragments of the chunk in mma order: [k-step][lane] = (a0, a1, a2, a3)
  extern __shared__ __align__(16) unsigned char seq_dyn[];
  float (*s_o)[16][D] = reinterpret_cast<float (*)[16][D]>(seq_dyn);   // [NW][16][D] fp32: per-warp partial outputs (64 KB)
  const int seq = blockIdx.x, hk = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, tig = lane & 3;
  if (tid == 0) s_ntok = 0;
  __syncthreads();
  for (int t = tid; t < T; t += NW * 32) {
    const int sq = __ldg(seq_ids + t), ps = __ldg(positions + t);
    if (sq == seq) {
      const int i = atomicAdd(&s_ntok, 1);
      s_tok[i] = t;
      s_pos[i] = ps;
    }
  }
  __syncthreads();
  const int ntok = s_ntok;
  if (ntok == 0) return;
  const size_t row_elems = (size_t)(Hq + 2 * Hkv) * D;
  const __half* kbase = cache_k + ((size_t)seq * S * Hkv + hk) * D;
  const __half* vbase = cache_v + ((size_t)seq * S * Hkv + hk) * D;
  const size_t kv_stride = (size_t)Hkv * D;                  // elements between consecutive positions
  for (int c0 = 0; c0 < ntok; c0 += R) {
    const int nr = min(R, ntok - c0);
    // rows gid and gid + 8 of this lane: (token slot, head) and the position limit of the row (-1 = padding row)
    int tokA = -1, tokB = -1, gA = 0, gB = 0;
    int posA = -1, posB = -1;
    if (gid < nr * G) { tokA = s_tok[c0 + gid / G]; posA = s_pos[c0 + gid / G]; gA = gid % G; }
    if (gid + 8 < nr * G) { tokB = s_tok[c0 + (gid + 8) / G]; posB = s_pos[c0 + (gid + 8) / G]; gB = (gid + 8) % G; }
    int pmax = max(posA, posB);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) pmax = max(pmax, __shfl_xor_sync(0xffffffffu, pmax, off));
    const int nkeys = min(pmax + 1, S);
    // ---- Q fragments: dims [32 tig, 32 tig + 32) of rows gid / gid + 8 (zero for padding rows). Identical for every warp:
    // warp 0 loads them and leaves them in shared memory in mma order (registers are needed for the accumulators).
    __syncthreads();                                         // the previous chunk is done with s_qf
    if (warp == 0) {
      const uint4 z = make_uint4(0, 0, 0, 0);
      const uint4* pa = tokA >= 0 ? reinterpret_cast<const uint4*>(qkv + (size_t)tokA * row_elems + (size_t)(hk * G + gA) * D + tig * 32) : nullptr;
      const uint4* pb = tokB >= 0 ? reinterpret_cast<const uint4*>(qkv + (size_t)tokB * row_elems + (size_t)(hk * G + gB) * D + tig * 32) : nullptr;
#pragma unroll
      for (int i = 0; i < 4; ++i) {                          // 16-byte piece i holds the words of k-steps 2 i and 2 i + 1
        const uint4 va = pa ? __ldg(pa + i) : z, vb = pb ? __ldg(pb + i) : z;
        s_qf[2 * i][lane] = make_uint4(va.x, vb.x, va.y, vb.y);
        s_qf[2 * i + 1][lane] = make_uint4(va.z, vb.z, va.w, vb.w);
      }
    }
    __syncthreads();
    float o[16][4];                                          // n-tile t: (row gid: dims 32 tig + t, 32 tig + 16 + t), (row gid + 8: same)
#pragma unroll
    for (int t = 0; t < 16; ++t) { o[t][0] = 0.f; o[t][1] = 0.f; o[t][2] = 0.f; o[t][3] = 0.f; }
    float mA = -INFINITY, mB = -INFINITY, lA = 0.f, lB = 0.f;
    // this warp's keys: blocks of 16, interleaved over the warps
    for (int kb = warp * 16; kb < nkeys; kb += NW * 16) {
      // ---- loads of the block: K rows kb + gid, kb + 8 + gid (64 B each), V rows kb + 2 tig (+1), kb + 8 + 2 tig (+1) (32 B each)
      // (requesting a warp's first block before the Q fragments are staged was tried: the 16 vectors do not fit the 128 registers
      // beside the 64 accumulators across the staging -- ~200 bytes of spills)
      uint4 k0[4], k1[4], v[4][2];
      {
        const int j0 = min(kb + gid, nkeys - 1), j1 = min(kb + 8 + gid, nkeys - 1);
        const uint4* p0 = reinterpret_cast<const uint4*>(kbase + (size_t)j0 * kv_stride + tig * 32);
        const uint4* p1 = reinterpret_cast<const uint4*>(kbase + (size_t)j1 * kv_stride + tig * 32);
#pragma unroll
        for (int i = 0; i < 4; ++i) { k0[i] = __ldg(p0 + i); k1[i] = __ldg(p1 + i); }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int jv = min(kb + (i >> 1) * 8 + 2 * tig + (i & 1), nkeys - 1);
          const uint4* pv = reinterpret_cast<const uint4*>(vbase + (size_t)jv * kv_stride + gid * 16);
          v[i][0] = __ldg(pv); v[i][1] = __ldg(pv + 1);
        }
      }
      // ---- S = Q K^T for the two 8-key tiles
      float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
      {
        const uint32_t* w0 = reinterpret_cast<const uint32_t*>(k0);
        const uint32_t* w1 = reinterpret_cast<const uint32_t*>(k1);
#pragma unroll
        for (int st = 0; st < 8; ++st) {
          const uint4 a = s_qf[st][lane];
          mma16816(s0, a.x, a.y, a.z, a.w, w0[2 * st], w0[2 * st + 1]);
          mma16816(s1, a.x, a.y, a.z, a.w, w1[2 * st], w1[2 * st + 1]);
        }
      }
      // ---- scale, mask (key j <= position of the row), online softmax
      const int j00 = kb + 2 * tig, j10 = kb + 8 + 2 * tig;
      float pA[4], pB[4];                                    // row gid: keys j00, j00 + 1, j10, j10 + 1; row gid + 8 likewise
      pA[0] = j00 <= posA ? s0[0] * sm_scale : -INFINITY; pA[1] = j00 + 1 <= posA ? s0[1] * sm_scale : -INFINITY;
      pA[2] = j10 <= posA ? s1[0] * sm_scale : -INFINITY; pA[3] = j10 + 1 <= posA ? s1[1] * sm_scale : -INFINITY;
      pB[0] = j00 <= posB ? s0[2] * sm_scale : -INFINITY; pB[1] = j00 + 1 <= posB ? s0[3] * sm_scale : -INFINITY;
      pB[2] = j10 <= posB ? s1[2] * sm_scale : -INFINITY; pB[3] = j10 + 1 <= posB ? s1[3] * sm_scale : -INFINITY;
      float bmA = fmaxf(fmaxf(pA[0], pA[1]), fmaxf(pA[2], pA[3])), bmB = fmaxf(fmaxf(pB[0], pB[1]), fmaxf(pB[2], pB[3]));
      bmA = fmaxf(bmA, __shfl_xor_sync(0xffffffffu, bmA, 1)); bmA = fmaxf(bmA, __shfl_xor_sync(0xffffffffu, bmA, 2));
      bmB = fmaxf(bmB, __shfl_xor_sync(0xffffffffu, bmB, 1)); bmB = fmaxf(bmB, __shfl_xor_sync(0xffffffffu, bmB, 2));
      const float nmA = fmaxf(mA, bmA), nmB = fmaxf(mB, bmB);
      const float fA = nmA > -INFINITY ? __expf(mA - nmA) : 1.f, fB = nmB > -INFINITY ? __expf(mB - nmB) : 1.f;   // exp(-inf) = 0 on the first block
      float rsA = 0.f, rsB = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        pA[i] = nmA > -INFINITY ? __expf(pA[i] - nmA) : 0.f;
        pB[i] = nmB > -INFINITY ? __expf(pB[i] - nmB) : 0.f;
        rsA += pA[i]; rsB += pB[i];
      }
      lA = lA * fA + rsA; lB = lB * fB + rsB;                // (per-lane partial sums; the quad is reduced at the end)
      mA = nmA; mB = nmB;
#pragma unroll
      for (int t = 0; t < 16; ++t) { o[t][0] *= fA; o[t][1] *= fA; o[t][2] *= fB; o[t][3] *= fB; }
      // ---- O += P V with P = P_hi + P_lo (two f16 terms)
      uint32_t ah[4], al[4];
      split_h2(pA[0], pA[1], ah[0], al[0]); split_h2(pB[0], pB[1], ah[1], al[1]);
      split_h2(pA[2], pA[3], ah[2], al[2]); split_h2(pB[2], pB[3], ah[3], al[3]);
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        const uint32_t sel = (t & 1) ? 0x7632u : 0x5410u;
        const uint32_t* x0 = reinterpret_cast<const uint32_t*>(&v[0][0]);   // 8 words: dims 16 gid + 2 w, + 1 of value row kb + 2 tig
        const uint32_t* x1 = reinterpret_cast<const uint32_t*>(&v[1][0]);
        const uint32_t* x2 = reinterpret_cast<const uint32_t*>(&v[2][0]);
        const uint32_t* x3 = reinterpret_cast<const uint32_t*>(&v[3][0]);
        const uint32_t b0 = __byte_perm(x0[t >> 1], x1[t >> 1], sel), b1 = __byte_perm(x2[t >> 1], x3[t >> 1], sel);
        mma16816(o[t], ah[0], ah[1], ah[2], ah[3], b0, b1);
        mma16816(o[t], al[0], al[1], al[2], al[3], b0, b1);
      }
    }
    // ---- quad-reduce the row sums, publish this warp's (m, l, O)
    lA += __shfl_xor_sync(0xffffffffu, lA, 1); lA += __shfl_xor_sync(0xffffffffu, lA, 2);
    lB += __shfl_xor_sync(0xffffffffu, lB, 1); lB += __shfl_xor_sync(0xffffffffu, lB, 2);
    __syncthreads();                                         // the previous chunk's merge is done with shared memory
    if (tig == 0) { s_m[warp][gid] = mA; s_l[warp][gid] = lA; s_m[warp][gid + 8] = mB; s_l[warp][gid + 8] = lB; }
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      s_o[warp][gid][32 * tig + t] = o[t][0];      s_o[warp][gid][32 * tig + 16 + t] = o[t][1];
      s_o[warp][gid + 8][32 * tig + t] = o[t][2];  s_o[warp][gid + 8][32 * tig + 16 + t] = o[t][3];
    }
    __syncthreads();
    // ---- merge the warps: thread = (row r, 16 dims); out = sum_w e^{m_w - m} O_w / sum_w e^{m_w - m} l_w
    for (int item = tid; item < 16 * 8; item += NW * 32) {
      const int r = item >> 3, dq = item & 7;
      if (r >= nr * G) continue;
      float m = -INFINITY;
#pragma unroll
      for (int w = 0; w < NW; ++w) m = fmaxf(m, s_m[w][r]);
      float sc[NW], l = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) { sc[w] = s_m[w][r] > -INFINITY ? __expf(s_m[w][r] - m) : 0.f; l += sc[w] * s_l[w][r]; }
      const float inv = l > 0.f ? 1.0f / l : 0.f;
      float acc[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) acc[e] = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w)
#pragma unroll
        for (int e = 0; e < 16; ++e) acc[e] = fmaf(sc[w], s_o[w][r][dq * 16 + e], acc[e]);
      float f[8];
      __half* dst = out + ((size_t)s_tok[c0 + r / G] * Hq + hk * G + r % G) * D + dq * 16;
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = acc[e] * inv;
      reinterpret_cast<uint4*>(dst)[0] = pack_h8(f);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = acc[8 + e] * inv;
      reinterpret_cast<uint4*>(dst)[1] = pack_h8(f);
    }
  }
}

template <int G>
static int launch_seq_attention(const void* qkv, const void* ck, const void* cv, const int* positions, const int* seq_ids,
                                void* out, int T, int B, int Hq, int Hkv, int S, float sm_scale, cudaStream_t st) {
  constexpr size_t dyn = (size_t)SEQ_ATTN_WARPS * 16 * 128 * sizeof(float);
  auto kern = seq_attention_kernel<G>;
  static bool attr_set = false;                              // per instantiation
  if (!attr_set) {
    QSPEC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    attr_set = true;
  }
  launch_pdl(kern, dim3(B, Hkv), dim3(SEQ_ATTN_WARPS * 32), dyn, st, (const __half*)qkv, (const __half*)ck, (const __half*)cv,
             positions, seq_ids, (__half*)out, T, Hq, Hkv, S, sm_scale);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

// ------------------------------------------------------------------------------------------
// Greedy sampling: arg max of every row of the lm_head logits [rows, V] (f16), lowest index among equal maxima.
// (The reference samples with vLLM's sampler; QSpec's draft/verify acceptance compares greedy tokens.)
// Grid (nchunk, rows): every CTA reduces one slice of a row, the last CTA of a row to finish (counter) reduces the
// slices. part_val/part_idx: [rows][nchunk]; counters: [rows], zero-initialised, self-cleaning.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void argmax_merge(float& bv, int& bi, float v, int i) {
  if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
}

__global__ void __launch_bounds__(256)
argmax_rows_kernel(const __half* __restrict__ logits, int V, long long ld, int nchunk, float* __restrict__ part_val,
                   int* __restrict__ part_idx, unsigned* __restrict__ counters, long long* __restrict__ out_idx,
                   float* __restrict__ out_val) {
  pdl_enter();
  __shared__ float sv[8];
  __shared__ int si[8];
  __shared__ unsigned s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row = blockIdx.y, chunk = blockIdx.x;
  const __half* x = logits + (size_t)row * ld;
  const int nvec = V >> 3;
  const int per = (nvec + nchunk - 1) / nchunk;
  const int v0 = chunk * per, v1 = min(v0 + per, nvec);
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int v = v0 + tid; v < v1; v += 256) {
    float f[8];
    unpack_h8(__ldg(reinterpret_cast<const uint4*>(x) + v), f);
#pragma unroll
    for (int e = 0; e < 8; ++e)
      if (f[e] > bv) { bv = f[e]; bi = v * 8 + e; }       // ascending index within a thread: '>' keeps the first
  }
  if (chunk == nchunk - 1)
    for (int i = (nvec << 3) + tid; i < V; i += 256) argmax_merge(bv, bi, __half2float(x[i]), i);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
    argmax_merge(bv, bi, __shfl_xor_sync(0xffffffffu, bv, off), __shfl_xor_sync(0xffffffffu, bi, off));
  if (lane == 0) { sv[warp] = bv; si[warp] = bi; }
  __syncthreads();
  if (warp == 0) {
    bv = lane < 8 ? sv[lane] : -INFINITY;
    bi = lane < 8 ? si[lane] : 0x7fffffff;
#pragma unroll
    for (int off = 4; off > 0; off >>= 1)
      argmax_merge(bv, bi, __shfl_xor_sync(0xffffffffu, bv, off), __shfl_xor_sync(0xffffffffu, bi, off));
    if (lane == 0) {
      part_val[(size_t)row * nchunk + chunk] = bv;
      part_idx[(size_t)row * nchunk + chunk] = bi;
      __threadfence();
      s_last = atomicAdd(counters + row, 1u) == (unsigned)(nchunk - 1) ? 1u : 0u;
    }
  }
  __syncthreads();
  if (!s_last || warp != 0) return;
  __threadfence();
  bv = -INFINITY; bi = 0x7fffffff;
  for (int c = lane; c < nchunk; c += 32)
    argmax_merge(bv, bi, __ldcg(part_val + (size_t)row * nchunk + c), __ldcg(part_idx + (size_t)row * nchunk + c));
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
    argmax_merge(bv, bi, __shfl_xor_sync(0xffffffffu, bv, off), __shfl_xor_sync(0xffffffffu, bi, off));
  if (lane == 0) {
    out_idx[row] = bi < V ? bi : 0;                     // a row of NaNs compares false everywhere: return a valid index
    if (out_val) out_val[row] = bv;
    counters[row] = 0u;
  }
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

extern "C" int qspec_seq_attention_supported(int T, int Hq, int Hkv, int D, int S) {
  if (D != 128 || Hkv <= 0 || Hq % Hkv || T <= 0 || T > 1024 || S <= 0) return 0;
  const int G = Hq / Hkv;
  return (G == 1 || G == 2 || G == 4 || G == 5 || G == 8) ? 1 : 0;
}

extern "C" int qspec_seq_attention(const void* qkv, const void* cache_k, const void* cache_v, const int* positions,
                                   const int* seq_ids, void* out, int T, int B, int Hq, int Hkv, int D, int S,
                                   float sm_scale, void* stream) {
  QSPEC_REQUIRE(qkv && cache_k && cache_v && positions && seq_ids && out, QSPEC_ERR_INVALID_ARG, "seq_attention: null pointer");
  QSPEC_REQUIRE(T >= 0 && B > 0, QSPEC_ERR_INVALID_ARG, "seq_attention: T=%d B=%d", T, B);
  if (T == 0) return QSPEC_OK;
  QSPEC_REQUIRE(qspec_seq_attention_supported(T, Hq, Hkv, D, S), QSPEC_ERR_UNSUPPORTED,
                "seq_attention: unsupported T=%d Hq=%d Hkv=%d D=%d S=%d (use decode_attention)", T, Hq, Hkv, D, S);
  cudaStream_t st = (cudaStream_t)stream;
  switch (Hq / Hkv) {
    case 1: return launch_seq_attention<1>(qkv, cache_k, cache_v, positions, seq_ids, out, T, B, Hq, Hkv, S, sm_scale, st);
    case 2: return launch_seq_attention<2>(qkv, cache_k, cache_v, positions, seq_ids, out, T, B, Hq, Hkv, S, sm_scale, st);
    case 4: return launch_seq_attention<4>(qkv, cache_k, cache_v, positions, seq_ids, out, T, B, Hq, Hkv, S, sm_scale, st);
    case 5: return launch_seq_attention<5>(qkv, cache_k, cache_v, positions, seq_ids, out, T, B, Hq, Hkv, S, sm_scale, st);
    default: return launch_seq_attention<8>(qkv, cache_k, cache_v, positions, seq_ids, out, T, B, Hq, Hkv, S, sm_scale, st);
  }
}

extern "C" int qspec_attn_block_supported(int T, int Hq, int Hkv, int D, int S) {
  if (D != 128 || Hkv <= 0 || Hq % Hkv || T <= 0 || T > 32767 || S <= 0) return 0;
  const int G = Hq / Hkv;
  if (!((G == 2 || G == 4 || G == 8) && (Hkv == 1 || Hkv == 2 || Hkv == 4 || Hkv == 8))) return 0;
  return attn_block_stage_bytes(G, T, S) >= 0 ? 1 : 0;
}

extern "C" int qspec_attn_block(const void* qkv, const void* cos_sin_cache, const int* positions, const int* seq_ids,
                                void* cache_k, void* cache_v, void* attn_out, void* out_f16, int8_t* out_q,
                                void* scale_out, int T, int Hq, int Hkv, int D, int S, float sm_scale,
                                float had_scale, float clip_ratio, void* stream) {
  QSPEC_REQUIRE(qkv && cos_sin_cache && positions && seq_ids && cache_k && cache_v && (out_f16 || out_q),
                QSPEC_ERR_INVALID_ARG, "attn_block: null pointer");
  QSPEC_REQUIRE(!out_q || scale_out, QSPEC_ERR_INVALID_ARG, "attn_block: out_q needs scale_out");
  QSPEC_REQUIRE(T >= 0, QSPEC_ERR_INVALID_ARG, "attn_block: T=%d", T);
  if (T == 0) return QSPEC_OK;
  QSPEC_REQUIRE(qspec_attn_block_supported(T, Hq, Hkv, D, S), QSPEC_ERR_UNSUPPORTED,
                "attn_block: unsupported T=%d Hq=%d Hkv=%d D=%d S=%d (use the separate kernels)", T, Hq, Hkv, D, S);
  const int G = Hq / Hkv;
  cudaStream_t st = (cudaStream_t)stream;
#define QSPEC_AB(GG, KK)                                                                                             \
  if (G == GG && Hkv == KK)                                                                                          \
    return launch_attn_block<GG, KK>(qkv, cos_sin_cache, positions, seq_ids, cache_k, cache_v, attn_out, out_f16,    \
                                     out_q, scale_out, T, S, sm_scale, had_scale, clip_ratio, st);
  QSPEC_AB(2, 1) QSPEC_AB(2, 2) QSPEC_AB(2, 4) QSPEC_AB(2, 8)
  QSPEC_AB(4, 1) QSPEC_AB(4, 2) QSPEC_AB(4, 4) QSPEC_AB(4, 8)
  QSPEC_AB(8, 1) QSPEC_AB(8, 2) QSPEC_AB(8, 4) QSPEC_AB(8, 8)
#undef QSPEC_AB
  QSPEC_REQUIRE(false, QSPEC_ERR_UNSUPPORTED, "attn_block: no kernel for G=%d Hkv=%d", G, Hkv);
}

extern "C" size_t qspec_argmax_rows_workspace_bytes(int rows) { return 4096 + (size_t)(rows > 0 ? rows : 0) * 64 * 8; }

extern "C" int qspec_argmax_rows(const void* logits, int rows, int V, long long row_stride, long long* out_idx,
                                 float* out_val, void* workspace, size_t workspace_bytes, void* stream) {
  QSPEC_REQUIRE(logits && out_idx && workspace, QSPEC_ERR_INVALID_ARG, "argmax_rows: null pointer");
  QSPEC_REQUIRE(rows >= 0 && V > 0 && row_stride >= V, QSPEC_ERR_INVALID_ARG, "argmax_rows: bad sizes %d x %d (stride %lld)", rows, V, row_stride);
  QSPEC_REQUIRE(((uintptr_t)logits & 15) == 0 && row_stride % 8 == 0, QSPEC_ERR_UNSUPPORTED,
                "argmax_rows: rows must be 16-byte aligned");
  QSPEC_REQUIRE(rows <= 1024, QSPEC_ERR_UNSUPPORTED, "argmax_rows: at most 1024 rows per call (got %d)", rows);
  QSPEC_REQUIRE(workspace_bytes >= qspec_argmax_rows_workspace_bytes(rows), QSPEC_ERR_WORKSPACE, "argmax_rows: workspace too small");
  if (rows == 0) return QSPEC_OK;
  // enough CTAs to fill the GPU about twice, at most 64 slices per row, at least ~2K elements per slice
  const int sms = sm_count() > 0 ? sm_count() : 148;
  int nchunk = (2 * sms + rows - 1) / rows;
  if (nchunk > 64) nchunk = 64;
  const int by_size = (V + 2047) / 2048;
  if (nchunk > by_size) nchunk = by_size;
  if (nchunk < 1) nchunk = 1;
  unsigned* counters = (unsigned*)workspace;                                  // first 4 KB: [1024], zero on first use (caller) and after every call
  float* pv = (float*)((unsigned char*)workspace + 4096);
  int* pi = (int*)(pv + (size_t)rows * 64);
  launch_pdl(argmax_rows_kernel, dim3(nchunk, rows), dim3(256), 0, (cudaStream_t)stream, (const __half*)logits, V, row_stride,
             nchunk, pv, pi, counters, out_idx, out_val);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

// ---- allocation-free variant of qspec_argmax_rows ---------------------------------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the scratch of the
// arg max -- 4 KB of per-row arrival counters + 64 (value, index) slice results per row -- lives in ONE file-scope device
// array of QSPEC_ARGMAX_STATIC_MAX_ROWS rows (520 KB, held by every GPU that loads the library until the process exits).
// Host code must not pass the symbol itself to a kernel (on the host the name is a stand-in, not a device address): the
// device address is looked up once per device with cudaGetSymbolAddress, with that device current. All calls share the
// array: calls that overlap on one GPU (two streams) race; calls in stream order do not. The counters are cleared at the
// start of every call (cudaMemsetAsync on the caller's stream: an aborted or faulted earlier call cannot poison this one);
// the slice results are written before they are read within a call, so they need no clearing.
constexpr int ARGMAX_STATIC_MAX_ROWS = 1024;
constexpr size_t ARGMAX_STATIC_BYTES = 4096 + (size_t)ARGMAX_STATIC_MAX_ROWS * 64 * 8;
static __device__ __align__(256) unsigned char g_argmax_static_scratch[ARGMAX_STATIC_BYTES];

static int argmax_static_scratch(void** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static void* ptrs[MAX_DEV] = {};
  int dev = -1;
  QSPEC_REQUIRE(cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < MAX_DEV, QSPEC_ERR_CUDA,
                "argmax_rows_static: no current CUDA device (or device index >= %d)", MAX_DEV);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_argmax_static_scratch);
    QSPEC_REQUIRE(e == cudaSuccess && p, QSPEC_ERR_CUDA, "argmax_rows_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                  cudaGetErrorString(e));
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return QSPEC_OK;
}

extern "C" int qspec_argmax_rows_static_max_rows(void) { return ARGMAX_STATIC_MAX_ROWS; }

extern "C" int qspec_argmax_rows_static(const void* logits, int rows, int V, long long row_stride, long long* out_idx,
                                        float* out_val, void* stream) {
  QSPEC_REQUIRE(logits && out_idx, QSPEC_ERR_INVALID_ARG, "argmax_rows_static: null pointer");
  QSPEC_REQUIRE(rows >= 0 && V > 0 && row_stride >= V, QSPEC_ERR_INVALID_ARG, "argmax_rows_static: bad sizes %d x %d (stride %lld)", rows, V,
                row_stride);
  QSPEC_REQUIRE(((uintptr_t)logits & 15) == 0 && row_stride % 8 == 0, QSPEC_ERR_UNSUPPORTED,
                "argmax_rows_static: rows must be 16-byte aligned");
  QSPEC_REQUIRE(rows <= ARGMAX_STATIC_MAX_ROWS, QSPEC_ERR_UNSUPPORTED,
                "argmax_rows_static: %d rows exceed the built-in scratch of %d rows; use qspec_argmax_rows with a caller workspace", rows,
                ARGMAX_STATIC_MAX_ROWS);
  if (rows == 0) return QSPEC_OK;
  void* scratch = nullptr;
  if (int rc = argmax_static_scratch(&scratch)) return rc;
  const int sms = sm_count() > 0 ? sm_count() : 148;     // same slicing as qspec_argmax_rows
  int nchunk = (2 * sms + rows - 1) / rows;
  if (nchunk > 64) nchunk = 64;
  const int by_size = (V + 2047) / 2048;
  if (nchunk > by_size) nchunk = by_size;
  if (nchunk < 1) nchunk = 1;
  unsigned* counters = (unsigned*)scratch;
  float* pv = (float*)((unsigned char*)scratch + 4096);
  int* pi = (int*)(pv + (size_t)ARGMAX_STATIC_MAX_ROWS * 64);
  const cudaError_t e = cudaMemsetAsync(counters, 0, (size_t)rows * sizeof(unsigned), (cudaStream_t)stream);
  QSPEC_REQUIRE(e == cudaSuccess, QSPEC_ERR_CUDA, "argmax_rows_static: clearing the counters failed: %s", cudaGetErrorString(e));
  launch_pdl(argmax_rows_kernel, dim3(nchunk, rows), dim3(256), 0, (cudaStream_t)stream, (const __half*)logits, V, row_stride,
             nchunk, pv, pi, counters, out_idx, out_val);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}
