// Online Hadamard kernels (B): plain FHT rows, the o_proj across-heads rotation with both
// transposes folded in, and the down_proj had28 (x) H_n rotation, each optionally fused with
// the per-token int4 quantiser that follows it (and, for down_proj, SiLU*up in front). sm_100a.
//
// Butterflies are fp32 (a+b, a-b) in ascending stride -- in-thread over one 128-bit vector
// (strides 1,2,4), then warp shuffles, then register slots -- the order of the reference
// kernel (third-party/fast-hadamard-transform/csrc/fast_hadamard_transform_common.h:86-124),
// so results are bit-identical to oracle/qspec_oracle.py:fht.
#include "common.cuh"
#include <stdlib.h>

namespace qspec {

// ------------------------------------------------------------------------------------------
// Warp-level FHT of SLOTS*32 (or fewer) vectors of 8 elements.
//   lanes_log2 = log2(#lanes that belong to one transform) (<= 5); slot stages follow.
// ------------------------------------------------------------------------------------------
template <int SLOTS>
__device__ __forceinline__ void warp_fht(float (&v)[SLOTS][8], int lane, int lanes_log2) {
#pragma unroll
  for (int s = 0; s < SLOTS; ++s) fht8(v[s]);
  for (int step = 0; step < lanes_log2; ++step) {
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) fht_lane_stage(v[s], lane, 1 << step);
  }
#pragma unroll
  for (int bit = 1; bit < SLOTS; bit <<= 1) {
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      if ((s & bit) == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float a = v[s][i], b = v[s | bit][i];
          v[s][i] = __fadd_rn(a, b);
          v[s | bit][i] = __fsub_rn(a, b);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// Plain FHT rows: faster_fast_hadamard_transform(x, scale, out)
// (fast_hadamard_transform.cpp:69-110). One warp per row (n >= 256) or per 256/n rows.
// ------------------------------------------------------------------------------------------
template <int SLOTS, bool BF16>
__global__ void __launch_bounds__(128)
fht_rows_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int rows, int n, float scale) {
  pdl_enter();
  const int lane = threadIdx.x & 31;
  const long long unit = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int nv = n >> 3;                        // vectors per row
  const int L = nv < 32 ? nv : 32;              // lanes per row
  int lanes_log2 = 0;
  while ((1 << lanes_log2) < L) ++lanes_log2;
  const int rows_per_unit = 32 / L;
  const long long row = unit * rows_per_unit + lane / L;
  const int v0 = lane % L;
  const bool active = row < rows;
  float v[SLOTS][8];
#pragma unroll
  for (int s = 0; s < SLOTS; ++s) {
    if (active) {
      unpack8<BF16>(__ldg(x + row * nv + v0 + s * 32), v[s]);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[s][i] = 0.f;
    }
  }
  warp_fht<SLOTS>(v, lane, lanes_log2);
  if (active) {
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[s][i] = __fmul_rn(v[s][i], scale);
      out[row * nv + v0 + s * 32] = pack8<BF16>(v[s]);
    }
  }
}

// ------------------------------------------------------------------------------------------
// o_proj rotation across heads with the transposes folded in (quarot_llama.py:230-238).
// One 256-thread CTA per token row [H, D]. lane = (h_low << DL) | dchunk_low; high head bits
// live in register slots; a warp "task" covers 2^DL adjacent 16-byte chunks of every head.
// ------------------------------------------------------------------------------------------
template <int SLOTS, int MAXT>
__global__ void __launch_bounds__(256)
hadamard_heads_kernel(const uint4* __restrict__ x, uint4* __restrict__ out_f,
                      uint32_t* __restrict__ out_q, __half* __restrict__ scale_out,
                      int H, int D, int LH, float scale, float clip, int c8) {
  pdl_enter();
  __shared__ float red[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int DL = 5 - LH;
  const int dchunks = D >> 3;
  const int tasks = dchunks >> DL;
  const int hl = lane >> DL, dl = lane & ((1 << DL) - 1);
  const size_t row = blockIdx.x;
  const size_t row_v = row * (size_t)H * dchunks;   // row offset in 8-element vectors

  uint4 res[MAXT][SLOTS];
  float amax = 0.f;
#pragma unroll
  for (int ti = 0; ti < MAXT; ++ti) {
    const int task = warp + ti * 8;
    if (task < tasks) {                              // warp-uniform
      const int dc = (task << DL) | dl;
      float v[SLOTS][8];
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
        const int h = (s << LH) | hl;
        unpack_h8(__ldg(x + row_v + (size_t)h * dchunks + dc), v[s]);
      }
      // head bits 0..LH-1 are lane bits DL..4
      for (int step = 0; step < LH; ++step) {
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) fht_lane_stage(v[s], lane, 1 << (DL + step));
      }
#pragma unroll
      for (int bit = 1; bit < SLOTS; bit <<= 1) {
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) {
          if ((s & bit) == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float a = v[s][i], b = v[s | bit][i];
              v[s][i] = __fadd_rn(a, b);
              v[s | bit][i] = __fsub_rn(a, b);
            }
          }
        }
      }
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[s][i] = __fmul_rn(v[s][i], scale);
        res[ti][s] = pack_h8(v[s]);
        amax = habs_max8(res[ti][s], amax);
        if (out_f) out_f[row_v + (size_t)((s << LH) | hl) * dchunks + dc] = res[ti][s];
      }
    }
  }
  if (out_q == nullptr) return;                      // uniform
  amax = block_max(amax, red);
  const __half sh = sym_scale(amax, clip);
  const float sc = __half2float(sh);
#pragma unroll
  for (int ti = 0; ti < MAXT; ++ti) {
    const int task = warp + ti * 8;
    if (task < tasks) {
      const int dc = (task << DL) | dl;
#pragma unroll
      for (int s = 0; s < SLOTS; ++s)
      {
        const size_t qi = row_v + (size_t)((s << LH) | hl) * dchunks + dc;
        if (c8) reinterpret_cast<uint2*>(out_q)[qi] = sym_quant8_c8(res[ti][s], sc);   // one int4 value per byte
        else out_q[qi] = sym_quant8(res[ti][s], sc);
      }
    }
  }
  if (threadIdx.x == 0) scale_out[row] = sh;
}

// ------------------------------------------------------------------------------------------
// down_proj rotation had28 (x) H_n (quarot/functional/hadamard.py:109-113), optional SiLU*up
// prologue (quarot_llama.py:279-284) and quantiser epilogue (:292-294).
// One 256-thread CTA per token row [28, n]. Phase 1: warp FHT of each n-block, scaled and
// rounded to f16 into shared memory. Phase 2: the 28x28 +-1 mix, fp32 accumulate in
// ascending k, two adjacent columns per thread.
// ------------------------------------------------------------------------------------------
// Row masks of the 28x28 Hadamard matrix used by the reference (get_had28 in
// third-party/QuaRot/quarot/functional/hadamard.py); bit k of kHad28Neg[o] set <=> H[o][k] = -1.
// Filled from tests/golden/had28.npy by scripts/gen_had28_masks.py; verified by tests.
#include "had28_masks.inc"
// The same signs as fp32 multipliers (+1 / -1) for the packed-FMA form of the 28x28 mix: [o][k], built from the masks at compile time.
struct Had28Signs { float v[28 * 28]; };
constexpr Had28Signs make_had28_signs() {
  Had28Signs t{};
  for (int o = 0; o < 28; ++o)
    for (int k = 0; k < 28; ++k) t.v[o * 28 + k] = ((had28_neg_mask(o) >> k) & 1u) ? -1.0f : 1.0f;
  return t;
}
__constant__ Had28Signs d_had28_sign = make_had28_signs();

__device__ __forceinline__ float silu_mul_h(float g, float u) {
  const float s = __half2float(__float2half_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf(-g)))));
  return __half2float(__float2half_rn(__fmul_rn(s, u)));
}

template <int SLOTS>
__device__ __forceinline__ void k28_phase1(const uint4* __restrict__ xrow,
                                           const uint4* __restrict__ gurow, int inter_v,
                                           __half* __restrict__ sm, int n, float scale) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nv = n >> 3;
  const int L = nv < 32 ? nv : 32;
  int lanes_log2 = 0;
  while ((1 << lanes_log2) < L) ++lanes_log2;
  const int rpu = 32 / L;                            // n-blocks per warp pass
  const int units = (28 + rpu - 1) / rpu;
  for (int u = warp; u < units; u += 8) {
    const int blk = u * rpu + lane / L;
    const int v0 = lane % L;
    const bool active = blk < 28;
    float v[SLOTS][8];
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int vi = blk * nv + v0 + s * 32;
      if (active && gurow) {
        float up[8], g[8];
        unpack_h8(__ldg(gurow + vi), up);
        unpack_h8(__ldg(gurow + inter_v + vi), g);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[s][i] = silu_mul_h(g[i], up[i]);
      } else if (active) {
        unpack_h8(__ldg(xrow + vi), v[s]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[s][i] = 0.f;
      }
    }
    warp_fht<SLOTS>(v, lane, lanes_log2);
    if (active) {
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[s][i] = __fmul_rn(v[s][i], scale);
        reinterpret_cast<uint4*>(sm)[blk * nv + v0 + s * 32] = pack_h8(v[s]);
      }
    }
  }
}

// OPT outputs per thread starting at o0 (runtime). STATIC: o0 == 0 and OPT == 28 with
// compile-time signs (the Llama-3-8B/70B full-width case).
template <int OPT, bool STATIC>
__device__ __forceinline__ void k28_mix(const __half* __restrict__ sm, int n, int jp, int o0,
                                        float (&acc)[OPT][2]) {
  // both columns of a pair in one FFMA2 per (o, k): acc = y * (+-1) + acc, i.e. the fp32 add / subtract in ascending k of the
  // definition, bit for bit (common.cuh: ffma2_rn)
  unsigned long long a2[OPT];
#pragma unroll
  for (int o = 0; o < OPT; ++o) a2[o] = 0ull;
#pragma unroll
  for (int k = 0; k < 28; ++k) {
    const float2 y = __half22float2(reinterpret_cast<const __half2*>(sm)[(k * n >> 1) + jp]);
    const unsigned long long y2 = pack_f2(y.x, y.y);
#pragma unroll
    for (int o = 0; o < OPT; ++o) {
      const float sg = STATIC ? (((had28_neg_mask(o) >> k) & 1u) ? -1.0f : 1.0f) : d_had28_sign.v[(o0 + o) * 28 + k];
      a2[o] = ffma2_rn(y2, pack_f2(sg, sg), a2[o]);
    }
  }
#pragma unroll
  for (int o = 0; o < OPT; ++o) {
    const float2 r = unpack_f2(a2[o]);
    acc[o][0] = r.x;
    acc[o][1] = r.y;
  }
}

template <int SLOTS, int OPT, bool STATIC>
__global__ void __launch_bounds__(256)
hadamard_k28_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gate_up,
                    __half* __restrict__ out_f, uint8_t* __restrict__ out_q,
                    __half* __restrict__ scale_out, int n, float scale, float clip, int c8) {
  pdl_enter();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* sm = reinterpret_cast<__half*>(smem_raw);
  __shared__ float red[8];
  const size_t row = blockIdx.x;
  const int L = 28 * n;
  const int inter_v = L >> 3;
  k28_phase1<SLOTS>(x ? x + row * inter_v : nullptr,
                    gate_up ? gate_up + row * 2 * inter_v : nullptr, inter_v, sm, n, scale);
  __syncthreads();

  const int P = n >> 1;                               // column pairs
  constexpr int G = 28 / OPT;                         // output groups
  // P >= 256: loop over pairs, all outputs per thread (G == 1). P < 256: G*P <= 256 threads busy.
  const int iters = (G == 1) ? (P + 255) / 256 : 1;
  // results are kept as f16-rounded floats until the row amax is known (iters <= 2)
  float keep[2][OPT][2];
  float amax = 0.f;
  int jp_i[2], o0 = 0;
  bool act_i[2];
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    act_i[it] = false;
    jp_i[it] = 0;
    if (it < iters) {
      int jp, og;
      if (G == 1) { jp = threadIdx.x + it * 256; og = 0; }
      else { jp = threadIdx.x % P; og = threadIdx.x / P; }
      const bool active = (jp < P) && (og < G);
      act_i[it] = active;
      jp_i[it] = jp;
      o0 = og * OPT;
      if (active) {
        k28_mix<OPT, STATIC>(sm, n, jp, o0, keep[it]);
#pragma unroll
        for (int o = 0; o < OPT; ++o) {
          keep[it][o][0] = __half2float(__float2half_rn(keep[it][o][0]));
          keep[it][o][1] = __half2float(__float2half_rn(keep[it][o][1]));
          amax = fmaxf(amax, fmaxf(fabsf(keep[it][o][0]), fabsf(keep[it][o][1])));
          if (out_f)
            reinterpret_cast<__half2*>(out_f + row * L)[(((o0 + o) * n) >> 1) + jp] =
                __floats2half2_rn(keep[it][o][0], keep[it][o][1]);
        }
      }
    }
  }
  if (out_q == nullptr) return;
  amax = block_max(amax, red);
  const __half sh = sym_scale(amax, clip);
  const float sc = __half2float(sh);
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    if (it < iters && act_i[it]) {
#pragma unroll
      for (int o = 0; o < OPT; ++o) {
        const int q0 = max(-8, min(7, __half2int_rn(__float2half_rn(__fdiv_rn(keep[it][o][0], sc)))));
        const int q1 = max(-8, min(7, __half2int_rn(__float2half_rn(__fdiv_rn(keep[it][o][1], sc)))));
        const size_t qi = row * (L >> 1) + (((o0 + o) * n) >> 1) + jp_i[it];
        if (c8) reinterpret_cast<uint16_t*>(out_q)[qi] = (uint16_t)((q0 & 0xFF) | ((q1 & 0xFF) << 8));   // one value per byte
        else out_q[qi] = (uint8_t)((q0 & 0xF) | ((q1 & 0xF) << 4));
      }
    }
  }
  if (threadIdx.x == 0) scale_out[row] = sh;
}

// ------------------------------------------------------------------------------------------
// Cluster version of the down_proj rotation for n = 256 / 512 / 1024: one thread-block CLUSTER of 7 CTAs per
// token row instead of one CTA, so that a 4-token decode step uses 28 SMs and every phase is 7x shorter.
// All exchange is PUSHED over distributed shared memory (remote stores are fire-and-forget; remote loads would
// put a DSMEM round trip on the critical path after every barrier). CTA c of the cluster
//   phase 1: SiLU*up + FHT_n of n-blocks 4c..4c+3 (one warp per block), f16, stored into EVERY CTA's copy of
//            the [28][n] rotated row; one cluster barrier;
//   phase 2: outputs o = 4c..4c+3 of the 28x28 +-1 mix for all n columns from local shared memory, fp32
//            accumulate in ascending k;
//   quant:   CTA amax stored into every CTA, second barrier, row amax locally, quantiser on its 4 x n outputs.
// The arithmetic (order included) is that of hadamard_k28_kernel, so the two are bit-identical.
// ------------------------------------------------------------------------------------------
constexpr int K28_CLUSTER = 7;
constexpr int K28_BPC = 4;                           // n-blocks (and mix outputs) per CTA

template <int SLOTS>
__global__ void __launch_bounds__(128)
hadamard_k28_cluster_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gate_up,
                            __half* __restrict__ out_f, uint8_t* __restrict__ out_q,
                            __half* __restrict__ scale_out, int n, float scale, float clip, int c8) {
  pdl_enter();
  cluster_arrive_relaxed();                           // "this CTA runs": waited for before the first remote store
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* sm = reinterpret_cast<__half*>(smem_raw);   // [28][n] f16: ALL rotated blocks of the row (pushed by the cluster)
  __shared__ float red[8];
  __shared__ float amax_all[K28_CLUSTER];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x % K28_CLUSTER;             // rank in cluster (cluster dims = (7,1,1))
  const size_t row = blockIdx.x / K28_CLUSTER;
  const int L = 28 * n, inter_v = L >> 3, nv = n >> 3;
  const uint32_t sm_base = (uint32_t)__cvta_generic_to_shared(sm);

  // ---- phase 1: warp w rotates n-block 4c + w and pushes it into every CTA of the cluster
  {
    const int blk = K28_BPC * c + warp;
    float v[SLOTS][8];
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const int vi = blk * nv + lane + s * 32;
      if (gate_up) {
        float up[8], g[8];
        unpack_h8(__ldg(gate_up + row * 2 * inter_v + vi), up);
        unpack_h8(__ldg(gate_up + row * 2 * inter_v + inter_v + vi), g);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[s][i] = silu_mul_h(g[i], up[i]);
      } else {
        unpack_h8(__ldg(x + row * inter_v + vi), v[s]);
      }
    }
    warp_fht<SLOTS>(v, lane, 5);
    uint4 pk[SLOTS];
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[s][i] = __fmul_rn(v[s][i], scale);
      pk[s] = pack_h8(v[s]);
    }
    cluster_wait_acquire();                           // every CTA of the cluster has started
#pragma unroll
    for (int r = 0; r < K28_CLUSTER; ++r) {
      const uint32_t dst = mapa_shared(sm_base + (uint32_t)(blk * n * 2), (uint32_t)r);
#pragma unroll
      for (int s = 0; s < SLOTS; ++s) st_dsmem_v4(dst + (uint32_t)(lane + s * 32) * 16u, pk[s]);
    }
  }
  cluster_arrive_release();
  cluster_wait_acquire();

  // ---- phase 2: outputs o0..o0+3, column pairs jp = tid, tid + 128, ...; all 28 blocks are local now
  const int o0 = K28_BPC * c;
  constexpr int PPT = SLOTS;                          // pairs per thread: n/2/128 = 1 (n=256), 2 (n=512), 4 (n=1024)
  float keep[PPT][K28_BPC][2];
  {
    // one FFMA2 per (pair, o, k) with the sign as a +-1 multiplier from constant memory: the fp32 add / subtract in ascending k
    // of the definition, bit for bit, in a quarter of the issue slots of the predicated FADD pairs it replaces (ncu r02w: the
    // kernel was issue-bound, 51 % of its instructions were these FADDs)
    unsigned long long k2[PPT][K28_BPC];
#pragma unroll
    for (int pp = 0; pp < PPT; ++pp)
#pragma unroll
      for (int o = 0; o < K28_BPC; ++o) k2[pp][o] = 0ull;
#pragma unroll
    for (int k = 0; k < 28; ++k) {
      float sg[K28_BPC];
#pragma unroll
      for (int o = 0; o < K28_BPC; ++o) sg[o] = d_had28_sign.v[(o0 + o) * 28 + k];
#pragma unroll
      for (int pp = 0; pp < PPT; ++pp) {
        const float2 y = __half22float2(reinterpret_cast<const __half2*>(sm)[((k * n) >> 1) + threadIdx.x + pp * 128]);
        const unsigned long long y2 = pack_f2(y.x, y.y);
#pragma unroll
        for (int o = 0; o < K28_BPC; ++o) k2[pp][o] = ffma2_rn(y2, pack_f2(sg[o], sg[o]), k2[pp][o]);
      }
    }
#pragma unroll
    for (int pp = 0; pp < PPT; ++pp)
#pragma unroll
      for (int o = 0; o < K28_BPC; ++o) {
        const float2 r = unpack_f2(k2[pp][o]);
        keep[pp][o][0] = r.x;
        keep[pp][o][1] = r.y;
      }
  }
  float amax = 0.f;
#pragma unroll
  for (int pp = 0; pp < PPT; ++pp) {
    const int jp = threadIdx.x + pp * 128;
#pragma unroll
    for (int o = 0; o < K28_BPC; ++o) {
      keep[pp][o][0] = __half2float(__float2half_rn(keep[pp][o][0]));
      keep[pp][o][1] = __half2float(__float2half_rn(keep[pp][o][1]));
      amax = fmaxf(amax, fmaxf(fabsf(keep[pp][o][0]), fabsf(keep[pp][o][1])));
      if (out_f)
        reinterpret_cast<__half2*>(out_f + row * L)[(((o0 + o) * n) >> 1) + jp] = __floats2half2_rn(keep[pp][o][0], keep[pp][o][1]);
    }
  }
  if (out_q == nullptr) return;                       // uniform over the cluster; nothing remote is outstanding
  // ---- quantiser: CTA amax pushed to every CTA, one more barrier, row amax from local shared memory
  amax = block_max(amax, red);
  if (threadIdx.x < K28_CLUSTER)
    st_dsmem_u32(mapa_shared((uint32_t)__cvta_generic_to_shared(&amax_all[c]), (uint32_t)threadIdx.x), __float_as_uint(amax));
  cluster_arrive_release();
  cluster_wait_acquire();
  float row_amax = 0.f;
#pragma unroll
  for (int r = 0; r < K28_CLUSTER; ++r) row_amax = fmaxf(row_amax, amax_all[r]);
  const __half sh = sym_scale(row_amax, clip);
  const float sc = __half2float(sh);
#pragma unroll
  for (int pp = 0; pp < PPT; ++pp) {
    const int jp = threadIdx.x + pp * 128;
#pragma unroll
    for (int o = 0; o < K28_BPC; ++o) {
      const int q0 = max(-8, min(7, __half2int_rn(__float2half_rn(__fdiv_rn(keep[pp][o][0], sc)))));
      const int q1 = max(-8, min(7, __half2int_rn(__float2half_rn(__fdiv_rn(keep[pp][o][1], sc)))));
      const size_t qi = row * (L >> 1) + (((o0 + o) * n) >> 1) + jp;
      if (c8) reinterpret_cast<uint16_t*>(out_q)[qi] = (uint16_t)((q0 & 0xFF) | ((q1 & 0xFF) << 8));     // one value per byte
      else out_q[qi] = (uint8_t)((q0 & 0xF) | ((q1 & 0xF) << 4));
    }
  }
  if (c == 0 && threadIdx.x == 0) scale_out[row] = sh;
}

}  // namespace qspec

using namespace qspec;

static inline bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

extern "C" int qspec_fht(const void* x, void* out, int rows, int n, float scale, int dtype,
                         void* stream) {
  QSPEC_REQUIRE(x && out, QSPEC_ERR_INVALID_ARG, "fht: null pointer");
  QSPEC_REQUIRE(rows >= 0, QSPEC_ERR_INVALID_ARG, "fht: rows=%d", rows);
  QSPEC_REQUIRE(is_pow2(n) && n >= 8 && n <= 2048, QSPEC_ERR_UNSUPPORTED,
                "fht: n=%d must be a power of two in [8, 2048]", n);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED,
                "fht: dtype %d", dtype);
  if (rows == 0) return QSPEC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int nv = n / 8;
  const int L = nv < 32 ? nv : 32;
  const int rpu = 32 / L;
  const long long units = ((long long)rows + rpu - 1) / rpu;
  const unsigned grid = (unsigned)((units + 3) / 4);
  const int slots = nv <= 32 ? 1 : nv / 32;
  const bool bf = dtype == QSPEC_DTYPE_BF16;
#define QSPEC_FHT_CASE(S)                                                                       \
  do {                                                                                          \
    if (bf) launch_pdl(fht_rows_kernel<S, true>, dim3(grid), dim3(128), 0, st, (const uint4*)x, (uint4*)out, rows, n, scale); \
    else launch_pdl(fht_rows_kernel<S, false>, dim3(grid), dim3(128), 0, st, (const uint4*)x, (uint4*)out, rows, n, scale);   \
  } while (0)
  switch (slots) {
    case 1: QSPEC_FHT_CASE(1); break;
    case 2: QSPEC_FHT_CASE(2); break;
    case 4: QSPEC_FHT_CASE(4); break;
    default: QSPEC_FHT_CASE(8); break;
  }
#undef QSPEC_FHT_CASE
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

static int hadamard_heads_impl(const void* x, void* out_f16, int8_t* out_q, void* scale_out, int rows, int H, int D,
                               float scale, float clip_ratio, void* stream, int c8) {
  QSPEC_REQUIRE(x && (out_f16 || out_q), QSPEC_ERR_INVALID_ARG, "hadamard_heads: null pointer");
  QSPEC_REQUIRE(!out_q || scale_out, QSPEC_ERR_INVALID_ARG, "hadamard_heads: out_q needs scale_out");
  QSPEC_REQUIRE(rows >= 0, QSPEC_ERR_INVALID_ARG, "hadamard_heads: rows=%d", rows);
  QSPEC_REQUIRE(is_pow2(H) && H >= 2 && H <= 64, QSPEC_ERR_UNSUPPORTED,
                "hadamard_heads: H=%d must be a power of two in [2, 64]", H);
  int a = 0;
  while ((1 << a) < H) ++a;
  const int LH = a < 4 ? a : 4;
  const int DL = 5 - LH;
  const int slots = 1 << (a - LH);
  QSPEC_REQUIRE(D % 16 == 0 && ((D / 8) % (1 << DL)) == 0 && (long long)H * D <= 16384,
                QSPEC_ERR_UNSUPPORTED, "hadamard_heads: unsupported H=%d D=%d", H, D);
  const int tasks = (D / 8) >> DL;
  QSPEC_REQUIRE(tasks <= 16, QSPEC_ERR_UNSUPPORTED, "hadamard_heads: D=%d too large for H=%d", D, H);
  if (rows == 0) return QSPEC_OK;
  cudaStream_t st = (cudaStream_t)stream;
#define QSPEC_HH_CASE(S, T)                                                                     \
  launch_pdl(hadamard_heads_kernel<S, T>, dim3(rows), dim3(256), 0, st, (const uint4*)x, (uint4*)out_f16,           \
      (uint32_t*)out_q, (__half*)scale_out, H, D, LH, scale, clip_ratio, c8)
  const int maxt = tasks <= 8 ? 1 : 2;
  if (slots == 1 && maxt == 1) QSPEC_HH_CASE(1, 1);
  else if (slots == 1) QSPEC_HH_CASE(1, 2);
  else if (slots == 2 && maxt == 1) QSPEC_HH_CASE(2, 1);
  else if (slots == 2) QSPEC_HH_CASE(2, 2);
  else if (maxt == 1) QSPEC_HH_CASE(4, 1);
  else QSPEC_HH_CASE(4, 2);
#undef QSPEC_HH_CASE
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

extern "C" int qspec_hadamard_heads(const void* x, void* out_f16, int8_t* out_q, void* scale_out,
                                    int rows, int H, int D, float scale, float clip_ratio,
                                    void* stream) {
  return hadamard_heads_impl(x, out_f16, out_q, scale_out, rows, H, D, scale, clip_ratio, stream, 0);
}
// Same, with out_q written one int4 value per byte ([rows, H*D] int8): the operand of qspec_gemm_w4a8*.
extern "C" int qspec_hadamard_heads_c8(const void* x, void* out_f16, int8_t* out_q, void* scale_out,
                                       int rows, int H, int D, float scale, float clip_ratio, void* stream) {
  return hadamard_heads_impl(x, out_f16, out_q, scale_out, rows, H, D, scale, clip_ratio, stream, 1);
}

template <int SLOTS, int OPT, bool STATIC>
static int launch_k28(const void* x, const void* gu, void* of, int8_t* oq, void* so, int rows,
                      int n, float scale, float clip, cudaStream_t st, int c8) {
  const size_t smem = (size_t)28 * n * sizeof(__half);
  auto kern = hadamard_k28_kernel<SLOTS, OPT, STATIC>;
  if (smem > 48 * 1024)
    QSPEC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  launch_pdl(kern, dim3(rows), dim3(256), smem, st, (const uint4*)x, (const uint4*)gu, (__half*)of, (uint8_t*)oq,
                                (__half*)so, n, scale, clip, c8);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

template <int SLOTS>
static int launch_k28_cluster(const void* x, const void* gu, void* of, int8_t* oq, void* so, int rows, int n,
                              float scale, float clip, cudaStream_t st, int c8) {
  auto kern = hadamard_k28_cluster_kernel<SLOTS>;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)rows * K28_CLUSTER);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = (size_t)28 * n * sizeof(__half);
  static bool attr_set = false;                        // per instantiation (n = 1024 needs 56 KB)
  if (!attr_set) {
    QSPEC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 28 * 1024 * (int)sizeof(__half)));
    attr_set = true;
  }
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = K28_CLUSTER;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  QSPEC_CUDA(cudaLaunchKernelEx(&cfg, kern, (const uint4*)x, (const uint4*)gu, (__half*)of, (uint8_t*)oq, (__half*)so, n,
                                scale, clip, c8));
  count_launch(1);
  return QSPEC_OK;
}

static int hadamard_k28_impl(const void* x, const void* gate_up, void* out_f16, int8_t* out_q, void* scale_out, int rows,
                             int n, float scale, float clip_ratio, void* stream, int c8) {
  QSPEC_REQUIRE((x || gate_up) && (out_f16 || out_q), QSPEC_ERR_INVALID_ARG,
                "hadamard_k28: null pointer");
  QSPEC_REQUIRE(!out_q || scale_out, QSPEC_ERR_INVALID_ARG, "hadamard_k28: out_q needs scale_out");
  QSPEC_REQUIRE(rows >= 0, QSPEC_ERR_INVALID_ARG, "hadamard_k28: rows=%d", rows);
  QSPEC_REQUIRE(is_pow2(n) && n >= 64 && n <= 1024, QSPEC_ERR_UNSUPPORTED,
                "hadamard_k28: n=%d must be a power of two in [64, 1024]", n);
  if (rows == 0) return QSPEC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const void* xx = gate_up ? nullptr : x;
  static const bool use_cluster = !(getenv("QSPEC_K28_CLUSTER") && getenv("QSPEC_K28_CLUSTER")[0] == '0');
  // The 7-CTA cluster form serves every row count: with many rows (64 / 256-token passes of configs[2]) it still wins in the
  // decode loop (19.9 vs 20.8 ms per 64-sequence step, profiles/r02f_notes.md) because its short CTAs overlap the
  // neighbouring GEMMs' heads and tails. QSPEC_K28_CLUSTER_ROWS caps the row count that takes it (experiment knob).
  static const int cluster_rows = getenv("QSPEC_K28_CLUSTER_ROWS") ? atoi(getenv("QSPEC_K28_CLUSTER_ROWS")) : 65536;
  const bool few_rows = rows <= cluster_rows;
  if (use_cluster && few_rows && n == 256) return launch_k28_cluster<1>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
  if (use_cluster && few_rows && n == 512) return launch_k28_cluster<2>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
  if (use_cluster && few_rows && n == 1024) return launch_k28_cluster<4>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
  switch (n) {
    case 1024: return launch_k28<4, 28, true>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
    case 512:  return launch_k28<2, 28, true>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
    case 256:  return launch_k28<1, 14, false>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
    case 128:  return launch_k28<1, 7, false>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
    default:   return launch_k28<1, 4, false>(xx, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, st, c8);
  }
}

extern "C" int qspec_hadamard_k28(const void* x, const void* gate_up, void* out_f16,
                                  int8_t* out_q, void* scale_out, int rows, int n, float scale,
                                  float clip_ratio, void* stream) {
  return hadamard_k28_impl(x, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, stream, 0);
}
// Same, with out_q written one int4 value per byte ([rows, 28 n] int8): the operand of qspec_gemm_w4a8*.
extern "C" int qspec_hadamard_k28_c8(const void* x, const void* gate_up, void* out_f16, int8_t* out_q, void* scale_out,
                                     int rows, int n, float scale, float clip_ratio, void* stream) {
  return hadamard_k28_impl(x, gate_up, out_f16, out_q, scale_out, rows, n, scale, clip_ratio, stream, 1);
}
