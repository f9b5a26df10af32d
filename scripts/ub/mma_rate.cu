// Micro-benchmark: cycles per tcgen05.mma (cta_group::1, M=128, SS operands from 128B-swizzled
// shared memory) as a function of N, kind, number of TMEM accumulators the issue stream
// alternates between, and CTAs per SM.  Answers whether the ~110-cycle cost per small-N MMA seen
// in conv_tc.cu is a hardware floor of back-to-back accumulation into ONE accumulator.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c)); }
__device__ __forceinline__ bool try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
template <int KIND>
__device__ __forceinline__ void umma(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout & 7) << 61;
  return d;
}

// nacc accumulators, `iters` rounds of `per` MMAs; MMA j of a round goes to accumulator
// (j % nacc); A tile advances over `adist` distinct 16 KB tiles (smem read pattern of taps)
template <int KIND>
__global__ void mma_rate(int n, int iters, int per, int nacc, int adist, long long *out) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (adist * 16384 + 32768) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t *>(raw + (base - smem_u32(raw)))[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const uint32_t cols = 512;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  long long t0 = 0, t1 = 0;
  if (warp == 0) {
    uint32_t pred = 0;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    const bool leader = pred != 0;
    const uint32_t idesc = (1u << 4) | (KIND == 0 ? 0u : (2u << 7) | (2u << 10)) |
                           ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t hi = make_desc(0, 1024, 2);
    const uint32_t b_addr = base + adist * 16384;
    const uint32_t astride = 512 / nacc;   // columns between accumulators
    uint32_t ph = 0;
    const uint64_t da0 = hi | (uint64_t)(base >> 4), db0 = hi | (uint64_t)(b_addr >> 4);
    const uint32_t d1 = tmem + (nacc > 1 ? astride : 0);
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int j0 = 0; j0 < per; j0 += 16) {
        if (leader) {
#pragma unroll
          for (int j = 0; j < 16; ++j)      // tap tile (j & 3), K step ((j >> 2) & 3): constants
            umma<KIND>((j & 1) ? d1 : tmem, da0 + (uint64_t)((j & 3) * 1024 + ((j >> 2) & 3) * 2),
                       db0 + (uint64_t)(((j >> 2) & 3) * 2), idesc, 1u);
        }
      }
      if (leader) umma_commit(smem_u32(&bar));
      while (!try_wait(smem_u32(&bar), ph)) {}
      ph ^= 1u;
    }
    t1 = clock64();
    if (leader) out[blockIdx.x] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(cols) : "memory");
}

int main() {
  long long *out;
  cudaMallocManaged(&out, 1024 * sizeof(long long));
  const int smem = 4 * 16384 + 32768 + 2048;
  cudaFuncSetAttribute(mma_rate<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(mma_rate<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int ns[] = {16, 32, 48, 64, 96, 128, 192, 256};
  printf("kind n nacc adist per  cycles/mma (1 CTA/SM, one issuing thread)\n");
  for (int kind = 0; kind < 2; ++kind)
    for (int n : ns)
      for (int nacc = 1; nacc <= 2; ++nacc)
        for (int per : {32, 256}) {
          if (nacc * n > 512) continue;
          const int iters = 2048 / per * 4, adist = 4;
          for (int rep = 0; rep < 2; ++rep) {
            if (kind == 0) mma_rate<0><<<148, 64, smem>>>(n, iters, per, nacc, adist, out);
            else mma_rate<1><<<148, 64, smem>>>(n, iters, per, nacc, adist, out);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          }
          printf("%s %3d %d %d %3d  %.1f\n", kind ? "tf32" : "f16 ", n, nacc, adist, per,
                 (double)out[0] / ((double)iters * per));
        }
  return 0;
}
