// Shared host/device helpers for the qspec_b200 C-ABI library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <atomic>

#include "../../include/qspec_b200.h"

namespace qspec {

// ---- host-side status plumbing --------------------------------------------------------------
void set_error(const char* fmt, ...);            // stores thread-local text for qspec_last_error()
void count_launch(int n = 1);                    // bumps qspec_launch_count()
int  sm_count();

#define QSPEC_REQUIRE(cond, code, ...)                                                       \
  do {                                                                                       \
    if (!(cond)) { ::qspec::set_error(__VA_ARGS__); return (code); }                         \
  } while (0)

#define QSPEC_CUDA(call)                                                                     \
  do {                                                                                       \
    cudaError_t _e = (call);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      ::qspec::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__,   \
                         __LINE__);                                                          \
      return QSPEC_ERR_CUDA;                                                                 \
    }                                                                                        \
  } while (0)

// Checks the launch that just happened (no sync).
#define QSPEC_CHECK_LAUNCH(nkernels)                                                         \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess) {                                                                 \
      ::qspec::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),         \
                         __FILE__, __LINE__);                                                \
      return QSPEC_ERR_CUDA;                                                                 \
    }                                                                                        \
    ::qspec::count_launch(nkernels);                                                         \
  } while (0)

// ---- programmatic dependent launch ----------------------------------------------------------
// Every kernel of this library is launched as a programmatic dependent of its predecessor in the stream and
// begins with pdl_enter(): it first lets ITS dependent start (so that the next GEMM can allocate tensor memory
// and stream its weights, which depend on nothing, while this kernel runs), then waits until the predecessor
// grid has completed and its writes are visible. A kernel that is launched this way and skipped the wait would
// race with its predecessor -- launch_pdl() is only to be used with kernels that call pdl_enter().
bool pdl_enabled();                               // QSPEC_PDL=0 turns the launch attribute off (debug)

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// ---- device helpers -------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ void pdl_enter() {
#ifdef QSPEC_PDL_WAIT_FIRST          // experiment: the dependent may only start once the predecessor has completed
  asm volatile("griddepcontrol.wait;\n" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
#else
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  asm volatile("griddepcontrol.wait;\n" ::: "memory");
#endif
}

// ---- thread-block clusters: distributed shared memory ----------------------------------------
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ uint32_t ld_dsmem_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared::cluster.u32 %0, [%1];\n" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint16_t ld_dsmem_u16(uint32_t addr) {
  uint16_t v;
  asm volatile("ld.shared::cluster.u16 %0, [%1];\n" : "=h"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void st_dsmem_u16(uint32_t addr, uint16_t v) {
  asm volatile("st.shared::cluster.u16 [%0], %1;\n" ::"r"(addr), "h"(v) : "memory");
}
__device__ __forceinline__ void st_dsmem_u32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared::cluster.u32 [%0], %1;\n" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void st_dsmem_v4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}

__device__ __forceinline__ float warp_xor_sum(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}
__device__ __forceinline__ float warp_xor_max(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}

// Fixed-order block sum for 256-thread blocks; mirrored by oracle/qspec_oracle.py:tree_sum_f32.
// `sm` points at >= 8 floats of shared memory. Every thread gets the total.
__device__ __forceinline__ float block_sum_256(float part, float* sm) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float w = warp_xor_sum(part);
  __syncthreads();
  if (lane == 0) sm[warp] = w;
  __syncthreads();
  float v = lane < 8 ? sm[lane] : 0.f;
  return warp_xor_sum(v);
}
// Block max (order independent). nwarps = blockDim.x / 32 <= 32.
__device__ __forceinline__ float block_max(float part, float* sm) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float w = warp_xor_max(part);
  __syncthreads();
  if (lane == 0) sm[warp] = w;
  __syncthreads();
  float v = lane < nw ? sm[lane] : 0.f;
  return warp_xor_max(v);
}

__device__ __forceinline__ void unpack_h8(const uint4& r, float (&f)[8]) {
  const __half2* h = reinterpret_cast<const __half2*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __half22float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ void unpack_b8(const uint4& r, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack_h8(const float (&f)[8]) {
  uint4 r;
  __half2* h = reinterpret_cast<__half2*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
  return r;
}
__device__ __forceinline__ uint4 pack_b8(const float (&f)[8]) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}
template <bool BF16>
__device__ __forceinline__ void unpack8(const uint4& r, float (&f)[8]) {
  if (BF16) unpack_b8(r, f); else unpack_h8(r, f);
}
template <bool BF16>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return BF16 ? pack_b8(f) : pack_h8(f);
}

// 8 ints in [-8,7] -> 4 bytes, even index in the low nibble.
// Two fp32 lanes in one 64-bit register pair for the packed FP32 pipe of sm_100 (fma.rn.f32x2 -> FFMA2: one issue slot for two
// IEEE round-to-nearest FMAs; each lane rounds exactly like fma.rn.f32). With a multiplier of +-1 it is an exact add / subtract:
// fma(y, +-1, acc) == __fadd_rn(acc, +-y) bit for bit, which is how the +-1 Hadamard mixes use it.
__device__ __forceinline__ unsigned long long pack_f2(float lo, float hi) {
  return ((unsigned long long)__float_as_uint(hi) << 32) | (unsigned long long)__float_as_uint(lo);
}
__device__ __forceinline__ float2 unpack_f2(unsigned long long v) {
  return make_float2(__uint_as_float((uint32_t)v), __uint_as_float((uint32_t)(v >> 32)));
}
__device__ __forceinline__ unsigned long long ffma2_rn(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;\n" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}

__device__ __forceinline__ uint32_t pack_nibbles8(const int (&q)[8]) {
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) r |= (uint32_t)(q[i] & 0xF) << (4 * i);
  return r;
}

// In-thread radix-2 butterflies over 8 values, strides 1,2,4 (ascending).
__device__ __forceinline__ void fht8(float (&v)[8]) {
#pragma unroll
  for (int s = 1; s < 8; s <<= 1) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int lo = j & (s - 1);
      const int idx = (j - lo) * 2 + lo;
      const float a = v[idx], b = v[idx + s];
      v[idx] = __fadd_rn(a, b);
      v[idx + s] = __fsub_rn(a, b);
    }
  }
}
// One cross-lane butterfly stage on 8 values: partner lane = lane ^ mask.
__device__ __forceinline__ void fht_lane_stage(float (&v)[8], int lane, int mask) {
  const bool upper = (lane & mask) != 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float o = __shfl_xor_sync(0xffffffffu, v[i], mask);
    v[i] = upper ? __fsub_rn(o, v[i]) : __fadd_rn(v[i], o);
  }
}

// ---- per-token symmetric int4 quantiser pieces (quant.cu:102-166), shared by the fused kernels
__device__ __forceinline__ float habs_max8(const uint4& r, float m) {
  const __half2* h = reinterpret_cast<const __half2*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(__habs2(h[i]));
    m = fmaxf(m, fmaxf(f.x, f.y));   // fp16 magnitudes are exact in fp32
  }
  return m;
}
// q = clamp(half2int_rn(f16(x / scale)), -8, 7) for 8 f16 values -> 4 packed bytes.
__device__ __forceinline__ uint32_t sym_quant8(const uint4& r, float sc) {
  float f[8];
  unpack_h8(r, f);
  int q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const __half qh = __float2half_rn(__fdiv_rn(f[i], sc));
    q[i] = max(-8, min(7, __half2int_rn(qh)));
  }
  return pack_nibbles8(q);
}
// The same 8 int4 values, one per BYTE (two's complement int8 in [-8, 7]): the "int8 container" the dual-mode GEMM takes through
// its int8-activation path (TMA delivers the operand as it lies: no widening per weight tile inside the GEMM).
__device__ __forceinline__ uint2 pack_bytes8(const int (&q)[8]) {
  uint2 w;
  w.x = (uint32_t)(q[0] & 0xFF) | ((uint32_t)(q[1] & 0xFF) << 8) | ((uint32_t)(q[2] & 0xFF) << 16) | ((uint32_t)(q[3] & 0xFF) << 24);
  w.y = (uint32_t)(q[4] & 0xFF) | ((uint32_t)(q[5] & 0xFF) << 8) | ((uint32_t)(q[6] & 0xFF) << 16) | ((uint32_t)(q[7] & 0xFF) << 24);
  return w;
}
__device__ __forceinline__ uint2 sym_quant8_c8(const uint4& r, float sc) {
  float f[8];
  unpack_h8(r, f);
  int q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const __half qh = __float2half_rn(__fdiv_rn(f[i], sc));
    q[i] = max(-8, min(7, __half2int_rn(qh)));
  }
  return pack_bytes8(q);
}
// scale = f16(f16(amax/7) * f16(clip))
__device__ __forceinline__ __half sym_scale(float amax, float clip) {
  const __half s0 = __float2half_rn(__fdiv_rn(amax, 7.0f));
  const __half hc = __float2half_rn(clip);
  return __float2half_rn(__fmul_rn(__half2float(s0), __half2float(hc)));
}

// W4A8 (quant.cu:188-245): q = clamp(half2int_rn(f16(x / scale)), -128, 127) for 8 f16 values -> 8 int8 bytes.
__device__ __forceinline__ uint2 sym_quant8_i8(const uint4& r, float sc) {
  float f[8];
  unpack_h8(r, f);
  uint32_t w[2] = {0u, 0u};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const __half qh = __float2half_rn(__fdiv_rn(f[i], sc));
    const int q = max(-128, min(127, __half2int_rn(qh)));
    w[i >> 2] |= (uint32_t)(q & 0xFF) << (8 * (i & 3));
  }
  return make_uint2(w[0], w[1]);
}
// scale = f16(f16(amax/127) * f16(clip))
__device__ __forceinline__ __half sym_scale_i8(float amax, float clip) {
  const __half s0 = __float2half_rn(__fdiv_rn(amax, 127.0f));
  const __half hc = __float2half_rn(clip);
  return __float2half_rn(__fmul_rn(__half2float(s0), __half2float(hc)));
}

#endif  // __CUDACC__
}  // namespace qspec
