// Micro-benchmark: cost per iteration of a producer/consumer mbarrier ring between two warps
// (one elected lane each), the skeleton of the TMA->MMA pipeline in conv_tc.cu.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c)); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(b) : "memory"); }
__device__ __forceinline__ bool try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool test_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
template <int MODE>
__device__ __forceinline__ void wait(uint32_t bar, uint32_t parity) {
  if (MODE == 0) {  // as in conv_tc.cu: try_wait, then clock-bounded loop
    if (try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!try_wait(bar, parity)) if (clock64() - t0 > 4000000000ll) __trap();
  } else if (MODE == 1) {  // plain try_wait loop
    while (!try_wait(bar, parity)) {}
  } else {  // test_wait spin
    while (!test_wait(bar, parity)) {}
  }
}

__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int MODE, int TX>
__global__ void pingpong(int iters, int stages, long long *out) {
  __shared__ __align__(8) uint64_t bars[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t b0 = smem_u32(bars);
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2 * stages; ++s) mbar_init(b0 + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long t0 = clock64();
  if (warp == 0 && lane == 0) {          // producer: wait empty, arrive full
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      wait<MODE>(b0 + 8 * (stages + s), ph ^ 1u);
      if (TX & 1) mbar_expect_tx(b0 + 8 * s, 0); else mbar_arrive(b0 + 8 * s);
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {   // consumer: wait full, arrive empty
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      wait<MODE>(b0 + 8 * s, ph);
      if (TX & 4) tc_fence_after();
      if (TX & 2) umma_commit(b0 + 8 * (stages + s)); else mbar_arrive(b0 + 8 * (stages + s));
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = clock64() - t0;
}

template <int MODE, int TX>
void run(const char *name, int stages, int threads) {
  long long *d, h;
  cudaMalloc(&d, 8);
  const int iters = 20000;
  pingpong<MODE, TX><<<1, threads>>>(iters, stages, d);
  pingpong<MODE, TX><<<1, threads>>>(iters, stages, d);
  cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("%-28s stages %d threads %4d: %7.1f cycles/iter (%s)\n", name, stages, threads, (double)h / iters, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  for (int stages : {1, 2, 6}) {
    run<0, 0>("try_wait+clock, arrive", stages, 64);
    run<1, 0>("try_wait loop, arrive", stages, 64);
    run<2, 0>("test_wait spin, arrive", stages, 64);
    run<1, 1>("try_wait loop, expect_tx", stages, 64);
  }
  for (int stages : {1, 2, 8}) {
    run<1, 2>("try_wait, consumer commit", stages, 64);
    run<1, 6>("try_wait, fence+commit", stages, 64);
    run<1, 4>("try_wait, fence+arrive", stages, 64);
  }
  run<1, 0>("try_wait loop, arrive", 6, 320);
  run<2, 0>("test_wait spin, arrive", 6, 320);
  return 0;
}
