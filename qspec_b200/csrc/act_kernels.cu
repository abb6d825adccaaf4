// Activation-side kernels of the QSpec dual-mode path: (A) norm(+int4 quant), per-token
// symmetric int4 quant, and the small element-wise neighbours. sm_100a.
//
// Numerics contract: every fp32 operation below is an explicit IEEE round-to-nearest
// intrinsic (no FMA contraction) in a fixed order, so the results are bit-identical to
// oracle/qspec_oracle.py. The reference kernels they replace are cited per kernel.
#include "common.cuh"

namespace qspec {

// ------------------------------------------------------------------------------------------
// (A) mean-subtracting norm without gamma, optional fused per-token int4 quantisation.
// Reference: third-party/kernels/csrc/layernorm_kernels.cu:569-720 (generalLayerNorm_fuse_sum_i4).
// One 256-thread CTA per row; the row lives in registers as raw 128-bit vectors
// (thread t owns vectors t, t+256, ...), so HBM is read exactly once.
// ------------------------------------------------------------------------------------------
//
// Tensor-parallel form (TP=true), SURVEY 8(e): the row-parallel GEMM that precedes this norm leaves one partial
// sum per rank. Instead of an all-reduce launch between the two (RowParallelLinear.forward,
// layers/linear.py:1143-1144) the norm reads every rank's partial over NVLink peer memory, adds them in rank
// order in fp32 (one rounding to the storage type: the value the all-reduce would have left in `out`) and goes
// on as the single-GPU kernel does. Ranks meet through one arrival flag per (rank, peer): CTA 0 stores the
// epoch into each peer's flag array once its own partial is complete (after the grid-dependency wait), every
// CTA then waits for the peers' epochs. Two partial buffers are used in turn by the caller (attention block,
// MLP block); a rank can only overwrite one of them after it has passed the wait of the other, which its peers
// enter after they finished reading the first, so no second barrier is needed. The epoch lives in device
// memory (state[0]) so that a captured CUDA graph can be replayed.
// ------------------------------------------------------------------------------------------
struct TpArgs {
  const uint4* peer[QSPEC_TP_MAX];   // partial sums of every rank (own rank included)
  uint32_t* flags[QSPEC_TP_MAX];     // arrival flags of every rank, QSPEC_TP_MAX words each
  uint32_t* state;                   // local: {epoch, CTAs done, time-outs}
  int world, rank;
  // two-step form (qspec_tp_rs_add_norm_*): row r is summed by rank r % world only, which pushes the rounded sum into
  // every peer's `reduced` buffer and then publishes the epoch in the peer's row_flags[r]
  uint4* reduced[QSPEC_TP_MAX];      // every rank's [rows, K] buffer of reduced rows (null: one-step form)
  uint32_t* row_flags[QSPEC_TP_MAX]; // every rank's per-row "reduced row has landed" flags
  int phases;                        // 3 = product; 1 / 2 = owners' rows / the other rows as separate launches (one-GPU emulation)
};

// Deferred GEMM reduction (SPLIT=true): the o_proj / down_proj GEMM that precedes this norm left `split` raw partial
// accumulators per output element (qspec_gemm_w4a4_split / qspec_gemm_w4a16_split: k-slices of a tile computed by
// different CTAs) instead of reducing them itself through its workspace. delta is rebuilt here with the arithmetic of
// the GEMM epilogue -- W4A4: exact int32 sum, >> 8, * sA[token] * sW[channel]; W4A16: fp32 sum in ascending slice
// order, * sW[channel]; + bias; one rounding to the storage type -- so the result is bit-identical to the GEMM that
// reduces in place, without its three serialized L2 round trips per launch.
constexpr int SPLIT_MAX = 8;
struct SplitArgs {
  const uint32_t* part;      // [split][rows][K], split <= SPLIT_MAX
  const void* sA;            // [rows]  (W4A4 only)
  const void* sW;            // [K]
  const void* bias;          // [K] or null
  int split, w4a16;
  int shift;                 // integer partials: 8 (W4A4: weights and activations both carry a factor 16) or 4 (int8-container activations)
};

template <bool BF16>
__device__ __forceinline__ float ld_scale(const void* p, size_t i) {
  return BF16 ? __bfloat162float(__ldg((const __nv_bfloat16*)p + i)) : __half2float(__ldg((const __half*)p + i));
}

// 8 consecutive channels of one token: delta values as floats already rounded to the storage type
template <bool BF16>
__device__ __forceinline__ void split_delta8(const SplitArgs& sp, size_t row, size_t rows, int K, int v, float (&b)[8]) {
  const size_t slice = rows * (size_t)K;
  const uint4* src = reinterpret_cast<const uint4*>(sp.part + row * K + (size_t)v * 8);
  // every slice is requested before the first one is used: one L2 round trip for the whole sum
  uint4 lo[SPLIT_MAX], hi[SPLIT_MAX];
#pragma unroll
  for (int c = 0; c < SPLIT_MAX; ++c) {
    if (c < sp.split) {
      const uint4* s2 = reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(src) + (size_t)c * slice);
      lo[c] = __ldcg(s2); hi[c] = __ldcg(s2 + 1);
    }
  }
  uint32_t acc[8] = {lo[0].x, lo[0].y, lo[0].z, lo[0].w, hi[0].x, hi[0].y, hi[0].z, hi[0].w};
#pragma unroll
  for (int c = 1; c < SPLIT_MAX; ++c) {
    if (c < sp.split) {
      const uint32_t w[8] = {lo[c].x, lo[c].y, lo[c].z, lo[c].w, hi[c].x, hi[c].y, hi[c].z, hi[c].w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
        acc[i] = sp.w4a16 ? __float_as_uint(__fadd_rn(__uint_as_float(acc[i]), __uint_as_float(w[i]))) : acc[i] + w[i];
    }
  }
  const float sa = sp.w4a16 ? 0.f : ld_scale<BF16>(sp.sA, row);
  float sw[8], bs[8];
  unpack8<BF16>(__ldg(reinterpret_cast<const uint4*>(sp.sW) + v), sw);           // 8 channel scales in one load
  if (sp.bias) unpack8<BF16>(__ldg(reinterpret_cast<const uint4*>(sp.bias) + v), bs);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float f = sp.w4a16 ? __fmul_rn(__uint_as_float(acc[i]), sw[i])
                       : __fmul_rn(__fmul_rn(__int2float_rn(((int32_t)acc[i]) >> sp.shift), sa), sw[i]);
    if (sp.bias) f = __fadd_rn(f, bs[i]);
    b[i] = f;
  }
  unpack8<BF16>(pack8<BF16>(b), b);
}

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_v4(uint4* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t v;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(v));
  return v;
}

template <int MAXV, bool QUANT, bool BF16, bool TP, bool SPLIT, bool Q8 = false>
__global__ void __launch_bounds__(256)
norm_kernel(const uint4* x, const uint4* __restrict__ delta, uint4* x_out, uint4* __restrict__ out_f,
            uint32_t* __restrict__ out_q, __half* __restrict__ scale_out, __half* __restrict__ sum_out, int K, float eps,
            const TpArgs tp, const SplitArgs sp, const int c8) {   // c8: int4 values one per byte (int8 container) instead of nibbles
  pdl_enter();
  __shared__ float red[8];
  const int t = threadIdx.x;
  const int nv = K >> 3;
  const size_t row = blockIdx.x;
  const uint4* xr = x + row * nv;
  const float fK = (float)K;
  uint32_t epoch = 0;
  bool rs_owner = false, rs_consumer = false;           // two-step form: this CTA's row is summed here / on another rank
  if (TP) {
    epoch = *reinterpret_cast<volatile uint32_t*>(tp.state) + 1u;
    const bool rs = tp.reduced[0] != nullptr;
    if (blockIdx.x == 0 && t < tp.world && t != tp.rank && (!rs || (tp.phases & 1))) {
      __threadfence_system();                           // own partial (previous kernel) before the flag
      st_release_sys(tp.flags[t] + tp.rank, epoch);
    }
    if (rs) {
      const bool mine = (int)(row % (size_t)tp.world) == tp.rank;
      if (tp.phases != 3 && (tp.phases == 1) != mine) { // one-GPU emulation: not this launch's row
        if (tp.phases == 2 && t == 0) {
          const uint32_t done = atomicAdd(tp.state + 1, 1u);
          if (done == gridDim.x - 1) {
            tp.state[1] = 0;
            *reinterpret_cast<volatile uint32_t*>(tp.state) = epoch;
          }
        }
        return;
      }
      rs_owner = mine;
      rs_consumer = !mine;
    }
    if (!rs_consumer) {
      if (t < tp.world && t != tp.rank) {               // every rank's partial is complete
        const uint32_t* f = tp.flags[tp.rank] + t;
        const uint64_t t0 = global_timer_ns();
        while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
          if (global_timer_ns() - t0 > 30000000000ull) {  // 30 s: a peer is gone; report, do not hang
            atomicAdd(tp.state + 2, 1u);
            break;
          }
        }
      }
    } else if (t == 0) {                                // the owner's rounded sum of this row has landed in local memory
      const uint32_t* f = tp.row_flags[tp.rank] + row;
      const uint64_t t0 = global_timer_ns();
      while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
        if (global_timer_ns() - t0 > 30000000000ull) {
          atomicAdd(tp.state + 2, 1u);
          break;
        }
      }
    }
    __syncthreads();
  }

  uint4 raw[MAXV];
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    const int v = t + s * 256;
    if (v < nv) {
      raw[s] = xr[v];
      if (TP || SPLIT || delta) {
        // fused residual add (quarot_llama.py:380,389): x <- T(x + delta), written back, then normalised
        float a[8], b[8];
        unpack8<BF16>(raw[s], a);
        if (TP && rs_consumer) {
          unpack8<BF16>(ld_relaxed_sys_v4(tp.reduced[tp.rank] + row * nv + v), b);   // pushed by the owner; not through L1
        } else if (TP) {
          // every rank's vector is requested before the first one is used: one NVLink round trip, not world - 1
          uint4 pv[QSPEC_TP_MAX];
#pragma unroll
          for (int r = 0; r < QSPEC_TP_MAX; ++r) {
            if (r < tp.world) {
              const uint4* src = tp.peer[r] + row * nv + v;
              pv[r] = r == tp.rank ? *src : ld_relaxed_sys_v4(src);
            }
          }
#pragma unroll
          for (int r = 0; r < QSPEC_TP_MAX; ++r) {
            if (r < tp.world) {
              float c[8];
              unpack8<BF16>(pv[r], c);
#pragma unroll
              for (int i = 0; i < 8; ++i) b[i] = r == 0 ? c[i] : __fadd_rn(b[i], c[i]);
            }
          }
          const uint4 pb = pack8<BF16>(b);              // the all-reduced partial, in its storage type
          unpack8<BF16>(pb, b);
          if (rs_owner) {
#pragma unroll
            for (int r = 0; r < QSPEC_TP_MAX; ++r)
              if (r < tp.world && r != tp.rank) st_relaxed_sys_v4(tp.reduced[r] + row * nv + v, pb);
          }
        } else if (SPLIT) {
          split_delta8<BF16>(sp, row, gridDim.x, K, v, b);
        } else {
          unpack8<BF16>(__ldg(delta + row * nv + v), b);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = __fadd_rn(a[i], b[i]);
        raw[s] = pack8<BF16>(a);
        x_out[row * nv + v] = raw[s];
      }
    }
  }
  if (TP && rs_owner) {                                 // the pushed row before its flag (every thread's stores, then one flag per peer)
    __threadfence_system();
    __syncthreads();
    if (t < tp.world && t != tp.rank) st_release_sys(tp.row_flags[t] + row, epoch);
  }
  // pass 1: mean
  float part = 0.f;
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    if (t + s * 256 < nv) {
      float f[8];
      unpack8<BF16>(raw[s], f);
      float a = f[0];
#pragma unroll
      for (int i = 1; i < 8; ++i) a = __fadd_rn(a, f[i]);
      part = __fadd_rn(part, a);
    }
  }
  const float mean = __fdiv_rn(block_sum_256(part, red), fK);
  // pass 2: variance of (x - mean)
  part = 0.f;
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    if (t + s * 256 < nv) {
      float f[8];
      unpack8<BF16>(raw[s], f);
      float a = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = __fsub_rn(f[i], mean);
        const float sq = __fmul_rn(d, d);
        a = (i == 0) ? sq : __fadd_rn(a, sq);
      }
      part = __fadd_rn(part, a);
    }
  }
  const float var = block_sum_256(part, red);
  const float r = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(__fdiv_rn(var, fK), eps)));

  if (!QUANT) {
#pragma unroll
    for (int s = 0; s < MAXV; ++s) {
      const int v = t + s * 256;
      if (v < nv) {
        float f[8];
        unpack8<BF16>(raw[s], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = __fmul_rn(__fsub_rn(f[i], mean), r);
        out_f[row * nv + v] = pack8<BF16>(f);
      }
    }
  }
  if (TP && !(tp.reduced[0] != nullptr && tp.phases == 1)) {
    // the last CTA to get here publishes the epoch for the next launch (every CTA read it before this point)
    __syncthreads();
    if (t == 0) {
      const uint32_t done = atomicAdd(tp.state + 1, 1u);
      if (done == gridDim.x - 1) {
        tp.state[1] = 0;
        *reinterpret_cast<volatile uint32_t*>(tp.state) = epoch;
      }
    }
  }
  if (!QUANT) return;

  // pass 3: amax of the f16-rounded y and its sum
  float amax = 0.f;
  part = 0.f;
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    if (t + s * 256 < nv) {
      float f[8];
      unpack8<BF16>(raw[s], f);
      float a = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float y = __fmul_rn(__fsub_rn(f[i], mean), r);
        const float yh = __half2float(__float2half_rn(y));
        amax = fmaxf(amax, fabsf(yh));
        a = (i == 0) ? yh : __fadd_rn(a, yh);
      }
      part = __fadd_rn(part, a);
    }
  }
  const float ysum = block_sum_256(part, red);
  amax = fmaxf(block_max(amax, red), __half2float(__float2half_rn(1e-6f)));
  // int4: 7 / amax, clamp to [-8, 7], nibbles (layernorm_kernels.cu:685-716); Q8 (W4A8): 127 / amax, saturating
  // cvt.rni to int8, one byte per value (generalLayerNorm_fuse_sum_i8, layernorm_kernels.cu:447-468)
  const float qmax = Q8 ? 127.0f : 7.0f;
  const float qs = __fdiv_rn(qmax, amax);
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    const int v = t + s * 256;
    if (v < nv) {
      float f[8];
      unpack8<BF16>(raw[s], f);
      int q[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float y = __fmul_rn(__fsub_rn(f[i], mean), r);
        const float c = Q8 ? fminf(fmaxf(__fmul_rn(y, qs), -128.0f), 127.0f) : fminf(fmaxf(__fmul_rn(y, qs), -8.0f), 7.0f);
        q[i] = __float2int_rn(c);
      }
      if (Q8 || c8) {
        uint2 w;
        w.x = (uint32_t)(q[0] & 0xFF) | ((uint32_t)(q[1] & 0xFF) << 8) | ((uint32_t)(q[2] & 0xFF) << 16) | ((uint32_t)(q[3] & 0xFF) << 24);
        w.y = (uint32_t)(q[4] & 0xFF) | ((uint32_t)(q[5] & 0xFF) << 8) | ((uint32_t)(q[6] & 0xFF) << 16) | ((uint32_t)(q[7] & 0xFF) << 24);
        reinterpret_cast<uint2*>(out_q)[row * nv + v] = w;
      } else {
        out_q[row * nv + v] = pack_nibbles8(q);
      }
    }
  }
  if (t == 0) {
    scale_out[row] = __float2half_rn(__fdiv_rn(amax, qmax));
    if (sum_out) sum_out[row] = __float2half_rn(ysum);
  }
}

template <bool QUANT, bool BF16, bool TP = false, bool SPLIT = false, bool Q8 = false>
static int launch_norm(const void* x, const void* delta, void* x_out, void* out_f, int8_t* out_q, void* scale,
                       void* isum, int rows, int K, float eps, cudaStream_t st, const TpArgs& tp = TpArgs(),
                       const SplitArgs& sp = SplitArgs(), int c8 = 0) {
  const int slots = (K / 8 + 255) / 256;
#define QSPEC_NORM_CASE(MV)                                                                   \
  launch_pdl(norm_kernel<MV, QUANT, BF16, TP, SPLIT, Q8>, dim3(rows), dim3(256), 0, st, (const uint4*)x, (const uint4*)delta, \
      (uint4*)x_out, (uint4*)out_f, (uint32_t*)out_q, (__half*)scale, (__half*)isum, K, eps, tp, sp, c8)
  if (slots <= 1) QSPEC_NORM_CASE(1);
  else if (slots <= 2) QSPEC_NORM_CASE(2);
  else if (slots <= 4) QSPEC_NORM_CASE(4);
  else QSPEC_NORM_CASE(8);
#undef QSPEC_NORM_CASE
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

// ------------------------------------------------------------------------------------------
// Per-token symmetric int4 quantisation (fp16 arithmetic as in the reference).
// Reference: third-party/QuaRot/quarot/kernels/quant.cu:102-166 (rowAbsMaxQuantizeKernel).
// ------------------------------------------------------------------------------------------
template <int MAXV>
__global__ void __launch_bounds__(256)
sym_quant_kernel(const uint4* __restrict__ x, __half* __restrict__ scale_out,
                 uint32_t* __restrict__ out_q, int K, float clip) {
  pdl_enter();
  __shared__ float red[8];
  const int t = threadIdx.x;
  const int nv = K >> 3;
  const size_t row = blockIdx.x;
  const uint4* xr = x + row * nv;
  uint4 raw[MAXV];
  float amax = 0.f;
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    const int v = t + s * 256;
    if (v < nv) {
      raw[s] = __ldg(xr + v);
      amax = habs_max8(raw[s], amax);
    }
  }
  amax = block_max(amax, red);
  const __half sh = sym_scale(amax, clip);
  const float sc = __half2float(sh);
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    const int v = t + s * 256;
    if (v < nv) out_q[row * nv + v] = sym_quant8(raw[s], sc);
  }
  if (t == 0) scale_out[row] = sh;
}

// W4A8 variant: one int8 per value, scale = amax / 127 (quant.cu:188-223, rowAbsMaxQuantizeKernel_i8).
template <int MAXV>
__global__ void __launch_bounds__(256)
sym_quant_i8_kernel(const uint4* __restrict__ x, __half* __restrict__ scale_out, uint2* __restrict__ out_q, int K, float clip) {
  pdl_enter();
  __shared__ float red[8];
  const int t = threadIdx.x;
  const int nv = K >> 3;
  const size_t row = blockIdx.x;
  const uint4* xr = x + row * nv;
  uint4 raw[MAXV];
  float amax = 0.f;
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    const int v = t + s * 256;
    if (v < nv) {
      raw[s] = __ldg(xr + v);
      amax = habs_max8(raw[s], amax);
    }
  }
  amax = block_max(amax, red);
  const __half sh = sym_scale_i8(amax, clip);
  const float sc = __half2float(sh);
#pragma unroll
  for (int s = 0; s < MAXV; ++s) {
    const int v = t + s * 256;
    if (v < nv) out_q[row * nv + v] = sym_quant8_i8(raw[s], sc);
  }
  if (t == 0) scale_out[row] = sh;
}

// ------------------------------------------------------------------------------------------
// element-wise neighbours: residual add, SiLU(gate)*up
// ------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void add_kernel(const uint4* a, const uint4* b, uint4* out, size_t nvec) {
  pdl_enter();   // out may alias a or b
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x) {
    float fa[8], fb[8];
    unpack8<BF16>(a[i], fa);
    unpack8<BF16>(b[i], fb);
#pragma unroll
    for (int k = 0; k < 8; ++k) fa[k] = __fadd_rn(fa[k], fb[k]);
    out[i] = pack8<BF16>(fa);
  }
}

template <bool BF16>
__device__ __forceinline__ float round16(float v) {
  return BF16 ? __bfloat162float(__float2bfloat16_rn(v)) : __half2float(__float2half_rn(v));
}
// f16(f16(silu(g)) * u), quarot_llama.py:284 (torch: silu in fp32 rounded to T, then a T multiply)
template <bool BF16>
__device__ __forceinline__ float silu_mul1(float g, float u) {
  const float s = round16<BF16>(__fdiv_rn(g, __fadd_rn(1.0f, expf(-g))));
  return __fmul_rn(s, u);
}

template <bool BF16>
__global__ void silu_mul_kernel(const uint4* __restrict__ gate_up, uint4* __restrict__ out,
                                int rows, int inter) {
  pdl_enter();
  const int nv = inter >> 3;
  const size_t total = (size_t)rows * nv;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / nv, c = i % nv;
    float u[8], g[8];
    unpack8<BF16>(__ldg(gate_up + r * 2 * nv + c), u);
    unpack8<BF16>(__ldg(gate_up + r * 2 * nv + nv + c), g);
#pragma unroll
    for (int k = 0; k < 8; ++k) u[k] = silu_mul1<BF16>(g[k], u[k]);
    out[i] = pack8<BF16>(u);
  }
}

}  // namespace qspec

using namespace qspec;

extern "C" int qspec_norm_quant_i4(const void* x, int8_t* out_q, void* scale, void* input_sum,
                                   int rows, int K, float eps, void* stream) {
  QSPEC_REQUIRE(x && out_q && scale, QSPEC_ERR_INVALID_ARG, "norm_quant_i4: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "norm_quant_i4: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "norm_quant_i4: K=%d must be a multiple of 16 and <= 16384", K);
  if (rows == 0) return QSPEC_OK;
  return launch_norm<true, false>(x, nullptr, nullptr, nullptr, out_q, scale, input_sum, rows, K, eps,
                                  (cudaStream_t)stream);
}

extern "C" int qspec_add_norm_quant_i4(void* x, const void* delta, int8_t* out_q, void* scale, void* input_sum,
                                       int rows, int K, float eps, void* stream) {
  QSPEC_REQUIRE(x && delta && out_q && scale, QSPEC_ERR_INVALID_ARG, "add_norm_quant_i4: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "add_norm_quant_i4: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "add_norm_quant_i4: K=%d must be a multiple of 16 and <= 16384", K);
  if (rows == 0) return QSPEC_OK;
  return launch_norm<true, false>(x, delta, x, nullptr, out_q, scale, input_sum, rows, K, eps, (cudaStream_t)stream);
}

extern "C" int qspec_add_norm_f16(void* x, const void* delta, void* out, int rows, int K, float eps, int dtype,
                                  void* stream) {
  QSPEC_REQUIRE(x && delta && out, QSPEC_ERR_INVALID_ARG, "add_norm_f16: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "add_norm_f16: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 8 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "add_norm_f16: K=%d must be a multiple of 8 and <= 16384", K);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "add_norm_f16: dtype %d", dtype);
  if (rows == 0) return QSPEC_OK;
  if (dtype == QSPEC_DTYPE_BF16)
    return launch_norm<false, true>(x, delta, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream);
  return launch_norm<false, false>(x, delta, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream);
}

static int make_tp_args(TpArgs& tp, const void* const* peer_partials, void* const* peer_flags, void* state,
                        int world, int rank, const char* who) {
  QSPEC_REQUIRE(peer_partials && peer_flags && state, QSPEC_ERR_INVALID_ARG, "%s: null pointer", who);
  QSPEC_REQUIRE(world >= 1 && world <= QSPEC_TP_MAX && rank >= 0 && rank < world, QSPEC_ERR_INVALID_ARG,
                "%s: rank %d of %d (at most %d ranks)", who, rank, world, QSPEC_TP_MAX);
  tp = TpArgs();
  for (int r = 0; r < world; ++r) {
    QSPEC_REQUIRE(peer_partials[r] && peer_flags[r], QSPEC_ERR_INVALID_ARG, "%s: null peer pointer for rank %d", who, r);
    tp.peer[r] = (const uint4*)peer_partials[r];
    tp.flags[r] = (uint32_t*)peer_flags[r];
  }
  tp.state = (uint32_t*)state;
  tp.world = world;
  tp.rank = rank;
  return QSPEC_OK;
}

// ---- vocab-sharded greedy token over peer memory ---------------------------------------------
// Each rank holds (max logit, arg max) of its vocabulary shard per token. One launch per rank, no collective call: the rank
// PUSHES its pairs into slot [rank] of every peer's table (remote stores, one NVLink one-way trip), publishes an arrival
// flag per peer, waits for the peers' flags and takes the best pair of its own table in rank order (strictly greater wins:
// ties go to the lowest rank = the lowest global index, like an arg max over the concatenated logits). Two tables are used
// in turn (epoch parity): a rank can only be one epoch ahead of a peer that is still reading.
struct ArgmaxPeerArgs {
  uint2* slots[QSPEC_TP_MAX];        // every rank's table: [2][QSPEC_TP_MAX][cap] (f32 bits, global index)
  uint32_t* flags[QSPEC_TP_MAX];     // every rank's QSPEC_TP_MAX arrival flags
  uint32_t* state;                   // local {epoch, -, time-outs}
  int world, rank;
};

__global__ void __launch_bounds__(256)
tp_argmax_combine_kernel(const float* __restrict__ val, const long long* __restrict__ idx, long long offset, int rows, int cap,
                         const ArgmaxPeerArgs tp, int phases, long long* __restrict__ out) {
  pdl_enter();
  const int t = threadIdx.x;
  const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(tp.state) + 1u;
  const size_t table = (size_t)(epoch & 1u) * QSPEC_TP_MAX * cap;
  if (phases & 1) {
    for (int i = t; i < rows; i += 256) {
      const uint32_t vb = __float_as_uint(val[i]), ib = (uint32_t)(idx[i] + offset);
      for (int r = 0; r < tp.world; ++r) {
        uint2* dst = tp.slots[r] + table + (size_t)tp.rank * cap + i;
        asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};\n" ::"l"(dst), "r"(vb), "r"(ib) : "memory");
      }
    }
    __threadfence_system();                               // every thread's pushes before the flags
    __syncthreads();
    if (t < tp.world && t != tp.rank) st_release_sys(tp.flags[t] + tp.rank, epoch);
  }
  if (phases & 2) {
    if (t < tp.world && t != tp.rank) {
      const uint32_t* f = tp.flags[tp.rank] + t;
      const uint64_t t0 = global_timer_ns();
      while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
        if (global_timer_ns() - t0 > 30000000000ull) {    // 30 s: a peer is gone; report, do not hang
          atomicAdd(tp.state + 2, 1u);
          break;
        }
      }
    }
    __syncthreads();
    for (int i = t; i < rows; i += 256) {
      float bv = 0.f;
      uint32_t bi = 0;
      for (int r = 0; r < tp.world; ++r) {
        uint32_t vb, ib;
        asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];\n" : "=r"(vb), "=r"(ib) : "l"(tp.slots[tp.rank] + table + (size_t)r * cap + i) : "memory");
        const float v = __uint_as_float(vb);
        if (r == 0 || v > bv) { bv = v; bi = ib; }
      }
      out[i] = (long long)bi;
    }
    __syncthreads();
    if (t == 0) *reinterpret_cast<volatile uint32_t*>(tp.state) = epoch;
  }
}

extern "C" int qspec_tp_argmax_combine(const float* local_max, const long long* local_idx, long long index_offset, int rows,
                                       void* const* peer_slots, void* const* peer_flags, void* state, int capacity,
                                       int world, int rank, int phases, long long* out_tokens, void* stream) {
  QSPEC_REQUIRE(local_max && local_idx && peer_slots && peer_flags && state && out_tokens, QSPEC_ERR_INVALID_ARG,
                "tp_argmax_combine: null pointer");
  QSPEC_REQUIRE(world >= 1 && world <= QSPEC_TP_MAX && rank >= 0 && rank < world, QSPEC_ERR_INVALID_ARG,
                "tp_argmax_combine: rank %d of %d (at most %d ranks)", rank, world, QSPEC_TP_MAX);
  QSPEC_REQUIRE(rows > 0 && rows <= capacity, QSPEC_ERR_INVALID_ARG,
                "tp_argmax_combine: %d rows for a table of %d (every rank must arrive: rows = 0 is not allowed)", rows, capacity);
  QSPEC_REQUIRE(phases >= 1 && phases <= 3, QSPEC_ERR_INVALID_ARG, "tp_argmax_combine: phases %d", phases);
  ArgmaxPeerArgs tp = ArgmaxPeerArgs();
  for (int r = 0; r < world; ++r) {
    QSPEC_REQUIRE(peer_slots[r] && peer_flags[r], QSPEC_ERR_INVALID_ARG, "tp_argmax_combine: null peer pointer for rank %d", r);
    tp.slots[r] = (uint2*)peer_slots[r];
    tp.flags[r] = (uint32_t*)peer_flags[r];
  }
  tp.state = (uint32_t*)state;
  tp.world = world;
  tp.rank = rank;
  launch_pdl(tp_argmax_combine_kernel, dim3(1), dim3(256), 0, (cudaStream_t)stream, local_max, local_idx, index_offset, rows,
             capacity, tp, phases, out_tokens);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

static int tp_add_norm_quant_impl(void* x, const void* const* peer_partials, void* const* peer_flags, void* state, int world,
                                  int rank, int8_t* out_q, void* scale, void* input_sum, int rows, int K, float eps,
                                  void* stream, int c8) {
  QSPEC_REQUIRE(x && out_q && scale, QSPEC_ERR_INVALID_ARG, "tp_add_norm_quant_i4: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "tp_add_norm_quant_i4: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "tp_add_norm_quant_i4: K=%d must be a multiple of 16 and <= 16384", K);
  QSPEC_REQUIRE(rows > 0, QSPEC_ERR_INVALID_ARG, "tp_add_norm_quant_i4: every rank must arrive, rows=0 is not allowed");
  TpArgs tp;
  if (int rc = make_tp_args(tp, peer_partials, peer_flags, state, world, rank, "tp_add_norm_quant_i4")) return rc;
  return launch_norm<true, false, true>(x, nullptr, x, nullptr, out_q, scale, input_sum, rows, K, eps,
                                        (cudaStream_t)stream, tp, SplitArgs(), c8);
}

extern "C" int qspec_tp_add_norm_quant_i4(void* x, const void* const* peer_partials, void* const* peer_flags,
                                          void* state, int world, int rank, int8_t* out_q, void* scale,
                                          void* input_sum, int rows, int K, float eps, void* stream) {
  return tp_add_norm_quant_impl(x, peer_partials, peer_flags, state, world, rank, out_q, scale, input_sum, rows, K, eps, stream, 0);
}
extern "C" int qspec_tp_add_norm_quant_i4_c8(void* x, const void* const* peer_partials, void* const* peer_flags,
                                             void* state, int world, int rank, int8_t* out_q, void* scale,
                                             void* input_sum, int rows, int K, float eps, void* stream) {
  return tp_add_norm_quant_impl(x, peer_partials, peer_flags, state, world, rank, out_q, scale, input_sum, rows, K, eps, stream, 1);
}

extern "C" int qspec_tp_add_norm_f16(void* x, const void* const* peer_partials, void* const* peer_flags,
                                     void* state, int world, int rank, void* out, int rows, int K, float eps,
                                     int dtype, void* stream) {
  QSPEC_REQUIRE(x && out, QSPEC_ERR_INVALID_ARG, "tp_add_norm_f16: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "tp_add_norm_f16: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 8 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "tp_add_norm_f16: K=%d must be a multiple of 8 and <= 16384", K);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "tp_add_norm_f16: dtype %d", dtype);
  QSPEC_REQUIRE(rows > 0, QSPEC_ERR_INVALID_ARG, "tp_add_norm_f16: every rank must arrive, rows=0 is not allowed");
  TpArgs tp;
  if (int rc = make_tp_args(tp, peer_partials, peer_flags, state, world, rank, "tp_add_norm_f16")) return rc;
  if (dtype == QSPEC_DTYPE_BF16)
    return launch_norm<false, true, true>(x, nullptr, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream, tp);
  return launch_norm<false, false, true>(x, nullptr, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream, tp);
}

// ---- two-step form of the fused all-reduce: reduce-scatter by row + push (all-gather) inside the same launch -------------
// The one-step form makes every rank read every rank's partial: (world - 1) x rows x K x 2 bytes per rank and exchange, 14 MB at
// TP = 8 for a 256-token verify pass (~20 us of NVLink). Here row r is summed by rank r % world alone (rank order, fp32, one
// rounding: the same bits), which pushes the rounded row to every peer and publishes the epoch in the peer's flag of that row;
// the other ranks' CTAs of row r wait for that flag and read the row from LOCAL memory. Traffic per rank: 2 (world - 1) / world
// x rows x K x 2 bytes; one more one-way NVLink latency. Every CTA of a launch is resident (rows <= 1024 CTAs of 256 threads)
// and an owner's CTA waits for nothing a consumer CTA produces, so the launch cannot dead-lock on itself.
static int add_rs_args(TpArgs& tp, void* const* peer_reduced, void* const* peer_row_flags, int phases, const char* who) {
  QSPEC_REQUIRE(peer_reduced && peer_row_flags, QSPEC_ERR_INVALID_ARG, "%s: null pointer", who);
  QSPEC_REQUIRE(phases >= 1 && phases <= 3, QSPEC_ERR_INVALID_ARG, "%s: phases %d", who, phases);
  for (int r = 0; r < tp.world; ++r) {
    QSPEC_REQUIRE(peer_reduced[r] && peer_row_flags[r], QSPEC_ERR_INVALID_ARG, "%s: null peer pointer for rank %d", who, r);
    QSPEC_REQUIRE(((uintptr_t)peer_reduced[r] & 15) == 0, QSPEC_ERR_UNSUPPORTED, "%s: reduced buffers must be 16-byte aligned", who);
    tp.reduced[r] = (uint4*)peer_reduced[r];
    tp.row_flags[r] = (uint32_t*)peer_row_flags[r];
  }
  tp.phases = phases;
  return QSPEC_OK;
}

extern "C" int qspec_tp_rs_add_norm_quant_i4(void* x, const void* const* peer_partials, void* const* peer_flags,
                                             void* const* peer_reduced, void* const* peer_row_flags, void* state, int world,
                                             int rank, int phases, int8_t* out_q, void* scale, void* input_sum, int rows,
                                             int K, float eps, void* stream) {
  QSPEC_REQUIRE(x && out_q && scale, QSPEC_ERR_INVALID_ARG, "tp_rs_add_norm_quant_i4: null pointer");
  QSPEC_REQUIRE(rows > 0 && K > 0, QSPEC_ERR_INVALID_ARG, "tp_rs_add_norm_quant_i4: bad sizes %d x %d (every rank must arrive: rows = 0 is not allowed)", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "tp_rs_add_norm_quant_i4: K=%d must be a multiple of 16 and <= 16384", K);
  QSPEC_REQUIRE(rows <= 1024, QSPEC_ERR_UNSUPPORTED, "tp_rs_add_norm_quant_i4: at most 1024 rows per call (all CTAs must be resident), got %d", rows);
  TpArgs tp;
  if (int rc = make_tp_args(tp, peer_partials, peer_flags, state, world, rank, "tp_rs_add_norm_quant_i4")) return rc;
  if (int rc = add_rs_args(tp, peer_reduced, peer_row_flags, phases, "tp_rs_add_norm_quant_i4")) return rc;
  return launch_norm<true, false, true>(x, nullptr, x, nullptr, out_q, scale, input_sum, rows, K, eps, (cudaStream_t)stream, tp);
}

extern "C" int qspec_tp_rs_add_norm_f16(void* x, const void* const* peer_partials, void* const* peer_flags,
                                        void* const* peer_reduced, void* const* peer_row_flags, void* state, int world,
                                        int rank, int phases, void* out, int rows, int K, float eps, int dtype, void* stream) {
  QSPEC_REQUIRE(x && out, QSPEC_ERR_INVALID_ARG, "tp_rs_add_norm_f16: null pointer");
  QSPEC_REQUIRE(rows > 0 && K > 0, QSPEC_ERR_INVALID_ARG, "tp_rs_add_norm_f16: bad sizes %d x %d (every rank must arrive: rows = 0 is not allowed)", rows, K);
  QSPEC_REQUIRE(K % 8 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED, "tp_rs_add_norm_f16: K=%d must be a multiple of 8 and <= 16384", K);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "tp_rs_add_norm_f16: dtype %d", dtype);
  QSPEC_REQUIRE(rows <= 1024, QSPEC_ERR_UNSUPPORTED, "tp_rs_add_norm_f16: at most 1024 rows per call (all CTAs must be resident), got %d", rows);
  TpArgs tp;
  if (int rc = make_tp_args(tp, peer_partials, peer_flags, state, world, rank, "tp_rs_add_norm_f16")) return rc;
  if (int rc = add_rs_args(tp, peer_reduced, peer_row_flags, phases, "tp_rs_add_norm_f16")) return rc;
  if (dtype == QSPEC_DTYPE_BF16)
    return launch_norm<false, true, true>(x, nullptr, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream, tp);
  return launch_norm<false, false, true>(x, nullptr, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream, tp);
}

static int make_split_args(SplitArgs& sp, const void* partials, int split, int w4a16, const void* sA, const void* sW,
                           const void* bias, const char* who) {
  QSPEC_REQUIRE(partials && sW && (w4a16 == 1 || sA), QSPEC_ERR_INVALID_ARG, "%s: null pointer", who);
  QSPEC_REQUIRE(split >= 1 && split <= SPLIT_MAX, QSPEC_ERR_INVALID_ARG, "%s: split=%d (at most %d)", who, split, SPLIT_MAX);
  QSPEC_REQUIRE(((uintptr_t)partials & 15) == 0 && ((uintptr_t)sW & 15) == 0 && ((uintptr_t)bias & 15) == 0, QSPEC_ERR_UNSUPPORTED,
                "%s: partials, sW and bias must be 16-byte aligned", who);
  QSPEC_REQUIRE(w4a16 >= 0 && w4a16 <= 2, QSPEC_ERR_INVALID_ARG, "%s: partial kind %d (0 = W4A4, 1 = W4A16, 2 = W4A4 from int8-container activations)", who, w4a16);
  sp.part = (const uint32_t*)partials; sp.sA = sA; sp.sW = sW; sp.bias = bias; sp.split = split; sp.w4a16 = w4a16 == 1 ? 1 : 0;
  sp.shift = w4a16 == 2 ? 4 : 8;
  return QSPEC_OK;
}

static int split_add_norm_quant_impl(void* x, const void* partials, int split, int w4a16, const void* sA, const void* sW,
                                     const void* bias, int8_t* out_q, void* scale, void* input_sum, int rows, int K,
                                     float eps, void* stream, int c8) {
  QSPEC_REQUIRE(x && out_q && scale, QSPEC_ERR_INVALID_ARG, "split_add_norm_quant_i4: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "split_add_norm_quant_i4: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "split_add_norm_quant_i4: K=%d must be a multiple of 16 and <= 16384", K);
  if (rows == 0) return QSPEC_OK;
  SplitArgs sp;
  if (int rc = make_split_args(sp, partials, split, w4a16, sA, sW, bias, "split_add_norm_quant_i4")) return rc;
  return launch_norm<true, false, false, true>(x, nullptr, x, nullptr, out_q, scale, input_sum, rows, K, eps,
                                               (cudaStream_t)stream, TpArgs(), sp, c8);
}

extern "C" int qspec_split_add_norm_quant_i4(void* x, const void* partials, int split, int w4a16, const void* sA,
                                             const void* sW, const void* bias, int8_t* out_q, void* scale,
                                             void* input_sum, int rows, int K, float eps, void* stream) {
  return split_add_norm_quant_impl(x, partials, split, w4a16, sA, sW, bias, out_q, scale, input_sum, rows, K, eps, stream, 0);
}
// Same, with the quantised row written one int4 value per byte (out_q: [rows, K] int8): the operand of qspec_gemm_w4a8*.
extern "C" int qspec_split_add_norm_quant_i4_c8(void* x, const void* partials, int split, int w4a16, const void* sA,
                                                const void* sW, const void* bias, int8_t* out_q, void* scale,
                                                void* input_sum, int rows, int K, float eps, void* stream) {
  return split_add_norm_quant_impl(x, partials, split, w4a16, sA, sW, bias, out_q, scale, input_sum, rows, K, eps, stream, 1);
}

extern "C" int qspec_split_add_norm_f16(void* x, const void* partials, int split, int w4a16, const void* sA,
                                        const void* sW, const void* bias, void* out, int rows, int K, float eps,
                                        int dtype, void* stream) {
  QSPEC_REQUIRE(x && out, QSPEC_ERR_INVALID_ARG, "split_add_norm_f16: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "split_add_norm_f16: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 8 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "split_add_norm_f16: K=%d must be a multiple of 8 and <= 16384", K);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED, "split_add_norm_f16: dtype %d", dtype);
  if (rows == 0) return QSPEC_OK;
  SplitArgs sp;
  if (int rc = make_split_args(sp, partials, split, w4a16, sA, sW, bias, "split_add_norm_f16")) return rc;
  if (dtype == QSPEC_DTYPE_BF16)
    return launch_norm<false, true, false, true>(x, nullptr, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream, TpArgs(), sp);
  return launch_norm<false, false, false, true>(x, nullptr, x, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream, TpArgs(), sp);
}

extern "C" int qspec_norm_f16(const void* x, void* out, int rows, int K, float eps, int dtype,
                              void* stream) {
  QSPEC_REQUIRE(x && out, QSPEC_ERR_INVALID_ARG, "norm_f16: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "norm_f16: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 8 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "norm_f16: K=%d must be a multiple of 8 and <= 16384", K);
  QSPEC_REQUIRE(dtype == QSPEC_DTYPE_F16 || dtype == QSPEC_DTYPE_BF16, QSPEC_ERR_UNSUPPORTED,
                "norm_f16: dtype %d", dtype);
  if (rows == 0) return QSPEC_OK;
  if (dtype == QSPEC_DTYPE_BF16)
    return launch_norm<false, true>(x, nullptr, nullptr, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream);
  return launch_norm<false, false>(x, nullptr, nullptr, out, nullptr, nullptr, nullptr, rows, K, eps, (cudaStream_t)stream);
}

extern "C" int qspec_sym_quant_i4(const void* x, void* scale, int8_t* out_q, int rows, int K,
                                  float clip_ratio, void* stream) {
  QSPEC_REQUIRE(x && out_q && scale, QSPEC_ERR_INVALID_ARG, "sym_quant_i4: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "sym_quant_i4: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "sym_quant_i4: K=%d must be a multiple of 16 and <= 16384", K);
  if (rows == 0) return QSPEC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int slots = (K / 8 + 255) / 256;
  if (slots <= 2)
    launch_pdl(sym_quant_kernel<2>, dim3(rows), dim3(256), 0, st, (const uint4*)x, (__half*)scale, (uint32_t*)out_q, K, clip_ratio);
  else
    launch_pdl(sym_quant_kernel<8>, dim3(rows), dim3(256), 0, st, (const uint4*)x, (__half*)scale, (uint32_t*)out_q, K, clip_ratio);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

// ---- W4A8 activation side ----------------------------------------------
extern "C" int qspec_norm_quant_i8(const void* x, int8_t* out_q, void* scale, void* input_sum,
                                   int rows, int K, float eps, void* stream) {
  QSPEC_REQUIRE(x && out_q && scale, QSPEC_ERR_INVALID_ARG, "norm_quant_i8: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "norm_quant_i8: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "norm_quant_i8: K=%d must be a multiple of 16 and <= 16384", K);
  if (rows == 0) return QSPEC_OK;
  return launch_norm<true, false, false, false, true>(x, nullptr, nullptr, nullptr, out_q, scale, input_sum, rows, K, eps,
                                                      (cudaStream_t)stream);
}

extern "C" int qspec_sym_quant_i8(const void* x, void* scale, int8_t* out_q, int rows, int K,
                                  float clip_ratio, void* stream) {
  QSPEC_REQUIRE(x && out_q && scale, QSPEC_ERR_INVALID_ARG, "sym_quant_i8: null pointer");
  QSPEC_REQUIRE(rows >= 0 && K > 0, QSPEC_ERR_INVALID_ARG, "sym_quant_i8: bad sizes %d x %d", rows, K);
  QSPEC_REQUIRE(K % 16 == 0 && K <= 16384, QSPEC_ERR_UNSUPPORTED,
                "sym_quant_i8: K=%d must be a multiple of 16 and <= 16384", K);
  if (rows == 0) return QSPEC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int slots = (K / 8 + 255) / 256;
  if (slots <= 2)
    launch_pdl(sym_quant_i8_kernel<2>, dim3(rows), dim3(256), 0, st, (const uint4*)x, (__half*)scale, (uint2*)out_q, K, clip_ratio);
  else
    launch_pdl(sym_quant_i8_kernel<8>, dim3(rows), dim3(256), 0, st, (const uint4*)x, (__half*)scale, (uint2*)out_q, K, clip_ratio);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

extern "C" int qspec_add(const void* a, const void* b, void* out, size_t n, int dtype, void* stream) {
  QSPEC_REQUIRE(a && b && out, QSPEC_ERR_INVALID_ARG, "add: null pointer");
  QSPEC_REQUIRE(n % 8 == 0, QSPEC_ERR_UNSUPPORTED, "add: n=%zu must be a multiple of 8", n);
  if (n == 0) return QSPEC_OK;
  const size_t nvec = n / 8;
  const int grid = (int)((nvec + 255) / 256 < 148 * 8 ? (nvec + 255) / 256 : 148 * 8);
  if (dtype == QSPEC_DTYPE_BF16)
    launch_pdl(add_kernel<true>, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const uint4*)a, (const uint4*)b, (uint4*)out, nvec);
  else
    launch_pdl(add_kernel<false>, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const uint4*)a, (const uint4*)b, (uint4*)out, nvec);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}

extern "C" int qspec_silu_mul(const void* gate_up, void* out, int rows, int inter, int dtype,
                              void* stream) {
  QSPEC_REQUIRE(gate_up && out, QSPEC_ERR_INVALID_ARG, "silu_mul: null pointer");
  QSPEC_REQUIRE(rows >= 0 && inter > 0 && inter % 8 == 0, QSPEC_ERR_UNSUPPORTED,
                "silu_mul: inter=%d must be a positive multiple of 8", inter);
  if (rows == 0) return QSPEC_OK;
  const size_t total = (size_t)rows * (inter / 8);
  const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  if (dtype == QSPEC_DTYPE_BF16)
    launch_pdl(silu_mul_kernel<true>, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const uint4*)gate_up, (uint4*)out, rows, inter);
  else
    launch_pdl(silu_mul_kernel<false>, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const uint4*)gate_up, (uint4*)out, rows, inter);
  QSPEC_CHECK_LAUNCH(1);
  return QSPEC_OK;
}
