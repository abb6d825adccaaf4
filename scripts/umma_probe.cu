// tcgen05.mma issue-rate probe (round 2): cycles per UMMA (M=128 per CTA, K=32 B) as a function of
//   * N (tokens per instruction), kind (i8 / f16), source of A (TMEM vs shared memory),
//   * the number of INDEPENDENT accumulators the instructions rotate over (nacc = 1, 2, 4): is the ~75-cycle figure of
//     round 1 the latency of a dependent chain on one accumulator, or the issue interval of the pipe?
//   * cta_group::2: ONE instruction issued by the leader of a CTA pair drives both SMs' tensor cores (M = 256 = 128 weight
//     channels from each CTA's own TMEM, the token tile split between the two CTAs' shared memory).
// Operands are whatever garbage is in TMEM/smem: only timing matters. Every wait has a clock bail-out (no hangs).
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
__device__ __forceinline__ bool wait_bar(uint32_t bar, uint32_t parity) {   // false on time-out (~2 s)
  const long long t0 = clock64();
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (!done && clock64() - t0 > 4000000000ll) return false;
  }
  return true;
}
// KIND 0 = i8, 1 = f16. TS 1 = A from TMEM. G = cta_group (1 or 2)
template <int KIND, int TS, int G>
__device__ __forceinline__ void umma(uint32_t d, uint32_t a_t, uint64_t a_s, uint64_t b, uint32_t idesc, uint32_t acc) {
#define UMMA_ASM(GRP, KND, AOP, ACON)                                                                                   \
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::" GRP ".kind::" KND " [%0], " AOP ", %2, %3, p;\n}\n" \
               ::"r"(d), ACON, "l"(b), "r"(idesc), "r"(acc) : "memory")
  if (G == 1) {
    if (TS) { if (KIND == 0) UMMA_ASM("1", "i8", "[%1]", "r"(a_t)); else UMMA_ASM("1", "f16", "[%1]", "r"(a_t)); }
    else    { if (KIND == 0) UMMA_ASM("1", "i8", "%1", "l"(a_s));   else UMMA_ASM("1", "f16", "%1", "l"(a_s)); }
  } else {
    if (TS) { if (KIND == 0) UMMA_ASM("2", "i8", "[%1]", "r"(a_t)); else UMMA_ASM("2", "f16", "[%1]", "r"(a_t)); }
    else    { if (KIND == 0) UMMA_ASM("2", "i8", "%1", "l"(a_s));   else UMMA_ASM("2", "f16", "%1", "l"(a_s)); }
  }
#undef UMMA_ASM
}

template <int KIND, int TS, int G>
__global__ void __launch_bounds__(256, 1) probe(int n, int iters, int nacc, int st_traffic, long long* out) {
  __shared__ volatile int stop_flag;
  if (threadIdx.x == 0) stop_flag = 0;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  uint32_t crank = 0;
  if (G == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 256) reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x < 32) {
    if (G == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (G == 2) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = tmem_base;
  const uint32_t fmt = KIND == 0 ? ((2u << 4) | (1u << 7) | (1u << 10)) : ((1u << 4));
  const uint32_t idesc = fmt | ((uint32_t)(n >> 3) << 17) | (((G == 2 ? 256u : 128u) >> 4) << 24);
  // accumulator width in columns: n (cta_group::1) or n (cta_group::2: each CTA holds its 128 rows x n columns)
  const int acc_cols = n < 32 ? 32 : n;
  if (threadIdx.x < 32) {
    if (crank == 0 && elect_one()) {
      const uint64_t bdesc = make_desc_sw128(smem_u32(base));
      const uint64_t adesc = make_desc_sw128(smem_u32(base + 32 * 1024));
      const uint32_t a_col = (uint32_t)(nacc * acc_cols);             // operand slots behind the accumulators
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const uint32_t d = tmem + (uint32_t)((i % nacc) * acc_cols);
        umma<KIND, TS, G>(d, tmem + a_col + (i & 3) * 8, adesc + 2 * (i & 3), bdesc + 2 * (i & 3), idesc, i >= nacc ? 1u : 0u);
      }
      const long long t1 = clock64();
      if (G == 1)
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      else
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     ::"r"(smem_u32(&bar)), "h"((uint16_t)3) : "memory");
      const bool ok = wait_bar(smem_u32(&bar), 0);
      const long long t2 = clock64();
      out[0] = t1 - t0;
      out[1] = ok ? t2 - t0 : -1;
      stop_flag = 1;
    } else if (G == 2 && crank == 1 && elect_one()) {
      wait_bar(smem_u32(&bar), 0);                                    // the multicast commit arrives here too
      stop_flag = 1;
    }
    __syncwarp();
  } else if (threadIdx.x >= 128 && st_traffic) {
    // 4 warps (lane quarters 0..3) hammer tcgen05.st into columns 448.. like the converter warps of the GEMM do
    const uint32_t q = (threadIdx.x >> 5) & 3;
    uint32_t r0 = threadIdx.x, cnt = 0;
    const long long t0 = clock64();
    while (!stop_flag && clock64() - t0 < 6000000000ll) {
      const uint32_t taddr = tmem + ((q * 32) << 16) + 448 + (cnt & 1) * 32;
#pragma unroll
      for (int c = 0; c < st_traffic; ++c)
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr + c * 8), "r"(r0) : "memory");
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      ++cnt; r0 += cnt;
    }
    if (threadIdx.x == 128 && crank == 0) out[2] = cnt;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (G == 2) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
  }
  if (threadIdx.x < 32) {
    if (G == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
  }
}

template <int KIND, int TS, int G>
void run(const char* name, long long* d_out) {
  auto kern = probe<KIND, TS, G>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  for (int n : {16, 32, 64, 128, 256}) {
    if (G == 2 && n < 32) continue;                       // each CTA of the pair holds n/2 tokens; keep n/2 >= 16
    for (int nacc : {1, 2, 4}) {
      if (nacc * (n < 32 ? 32 : n) > 384) continue;       // accumulators + 32 operand columns + the store-traffic area must fit 512 columns
      for (int stt : {0, 4}) {
        long long h[3] = {0, 0, 0};
        const int iters = 2000;
        cudaMemset(d_out, 0, 24);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(G); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 96 * 1024;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = G; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, kern, n, iters, nacc, stt, d_out);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s N=%d: %s\n", name, n, cudaGetErrorString(e)); return; }
        cudaMemcpy(h, d_out, 24, cudaMemcpyDeviceToHost);
        printf("%-22s N=%3d nacc=%d st_traffic=%d: issue %.1f cyc/UMMA, complete %.1f cyc/UMMA%s; concurrent store loops %lld\n", name, n, nacc,
               stt, (double)h[0] / iters, (double)h[1] / iters, h[1] < 0 ? " (TIME-OUT)" : "", h[2]);
        fflush(stdout);
      }
    }
  }
}
int main() {
  long long* d_out; cudaMalloc(&d_out, 24);
  run<1, 1, 1>("f16 A=TMEM cta_group::1", d_out);
  run<0, 1, 1>("i8  A=TMEM cta_group::1", d_out);
  run<1, 0, 1>("f16 A=smem cta_group::1", d_out);
  run<1, 1, 2>("f16 A=TMEM cta_group::2", d_out);
  run<0, 1, 2>("i8  A=TMEM cta_group::2", d_out);
  run<1, 0, 2>("f16 A=smem cta_group::2", d_out);
  return 0;
}
