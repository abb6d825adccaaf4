// tcgen05.mma throughput probe: cycles per UMMA (M=128, K=32 B) as a function of N, kind (i8 / f16) and the
// source of A (TMEM vs shared memory). Operands are whatever garbage is in TMEM/smem: only timing matters.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/umma_probe.cu -o build/umma_probe
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t a) {
  return (uint64_t)((a >> 4) & 0x3FFFu) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
template <int KIND, int TS>
__device__ __forceinline__ void umma(uint32_t d, uint32_t a_t, uint64_t a_s, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (TS) {
    if (KIND == 0) asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d), "r"(a_t), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d), "r"(a_t), "l"(b), "r"(idesc), "r"(acc) : "memory");
  } else {
    if (KIND == 0) asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a_s), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a_s), "l"(b), "r"(idesc), "r"(acc) : "memory");
  }
}
template <int KIND, int TS>
__global__ void __launch_bounds__(256, 1) probe(int n, int iters, int nacc, int st_traffic, long long* out) {
  __shared__ volatile int stop_flag;
  if (threadIdx.x == 0) stop_flag = 0;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 256) reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = tmem_base;
  const uint32_t fmt = KIND == 0 ? ((2u << 4) | (1u << 7) | (1u << 10)) : ((1u << 4));
  const uint32_t idesc = fmt | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
  if (threadIdx.x < 32) {
    if (elect_one()) {
      const uint64_t bdesc = make_desc_sw128(smem_u32(base));
      const uint64_t adesc = make_desc_sw128(smem_u32(base + 32 * 1024));
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const uint32_t d = tmem + (uint32_t)((i % nacc) * 256 / nacc);
        umma<KIND, TS>(d, tmem + 256 + (i & 3) * 8, adesc + 2 * (i & 3), bdesc + 2 * (i & 3), idesc, i >= nacc ? 1u : 0u);
      }
      const long long t1 = clock64();
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      uint32_t done;
      do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
      } while (!done);
      const long long t2 = clock64();
      out[0] = t1 - t0;
      out[1] = t2 - t0;
      stop_flag = 1;
    }
    __syncwarp();
  } else if (threadIdx.x >= 128 && st_traffic) {
    // 4 warps (lane quarters 0..3) hammer tcgen05.st into columns 384.. like the converter warps of the GEMM do
    const uint32_t q = (threadIdx.x >> 5) & 3;
    uint32_t r0 = threadIdx.x, cnt = 0;
    while (!stop_flag) {
      const uint32_t taddr = tmem + ((q * 32) << 16) + 384 + (cnt & 3) * 32;
#pragma unroll
      for (int c = 0; c < st_traffic; ++c)
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr + c * 8), "r"(r0) : "memory");
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      ++cnt; r0 += cnt;
    }
    if (threadIdx.x == 128) out[2] = cnt;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}
template <int KIND, int TS>
void run(const char* name, long long* d_out) {
  cudaFuncSetAttribute(probe<KIND, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int n : {16, 64}) for (int stt : {0, 1, 4}) {
    const int nacc = 1;
    long long h[3] = {0, 0, 0};
    const int iters = 2000;
    probe<KIND, TS><<<1, 256, 64 * 1024>>>(n, iters, nacc, stt, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s N=%d: %s\n", name, n, cudaGetErrorString(e)); return; }
    cudaMemcpy(h, d_out, 24, cudaMemcpyDeviceToHost);
    printf("%-12s N=%3d st_traffic=%d: complete %.1f cyc/UMMA; concurrent store loops %lld (x%d x 1 KB per warp)\n", name, n, stt, (double)h[1] / iters, h[2], stt);
  }
}
int main() {
  long long* d_out; cudaMalloc(&d_out, 24); cudaMemset(d_out, 0, 24);
  run<0, 1>("i8  A=TMEM", d_out);
  run<1, 1>("f16 A=TMEM", d_out);
  return 0;
}
