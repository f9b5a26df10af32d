// ffn_tc.cu -- fused ConvFFN of a DepthConvBlock (reference DCVC-DC/src/models/layers.py:166-179,
// called from DepthConvBlock.forward :199-209) as ONE two-GEMM tcgen05 kernel for sm_100a:
//
//     out = ( t + lrelu( W2 . lrelu( W1 . t + b1, s1 ) + b2, s2 ) ) * scale
//
// t is a binary16 NHWC activation with C in {32, 48, 64} channels, W1 is the 1x1 up-projection
// (C -> 4C), W2 the 1x1 down-projection (4C -> C).  Unfused this is two HBM-bound kernels that
// write and re-read the 4C-wide hidden tensor (10C of the 14C..16C bytes per pixel); here the
// hidden tile never leaves the SM:
//
//   TMA:   t tile (128 pixels x 64 K-padded binary16, 128B swizzle) -> smem ring of 2
//   MMA1:  acc1[128 x 4C] (TMEM, fp32) = t . W1^T        ceil(C/16) x UTCHMMA, N = 4C
//   EPI1:  16 warps: tcgen05.ld -> + b1 -> LeakyReLU -> binary16 -> smem H in the K-major
//          128B-swizzled layout an A operand needs (4C/64 chunks of 16 KB)
//   MMA2:  acc2[128 x C] (TMEM) = H . W2^T               4C/16 x UTCHMMA, N = C
//   EPI2:  4 warps: tcgen05.ld -> + b2 -> LeakyReLU -> + t (read back from the smem A tile)
//          -> * scale -> 128-bit global stores (fp32 or binary16 rows)
//
// Both weight matrices stay resident in shared memory (<= 64 KB).  Persistent CTAs, one per SM;
// the MMA warp issues MMA1(i+1) before MMA2(i) (H double-buffered; single H and the opposite
// order at C = 64, where two H buffers do not fit), so EPI1(i+1) overlaps MMA2(i) and EPI2(i).
// tcgen05.commit tracks every earlier MMA of the issuing thread, which orders the reuse of acc1
// and H without extra barriers (see the protocol notes at the roles).
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace {

constexpr int TILE_M = 128;
constexpr int EPI1_WARPS = 16, EPI2_WARPS = 4;   // EPI1: four warps per TMEM lane quadrant
constexpr int THREADS = 64 + 32 * (EPI1_WARPS + EPI2_WARPS);
constexpr uint32_t SPIN_LIMIT = 1u << 26;
constexpr uint32_t A_BYTES = TILE_M * 128;       // one t tile: 128 rows x 64 binary16
constexpr uint32_t ACC2_COL = 256;               // acc1 at column 0 (<= 256 columns), acc2 here

struct FfnArgs {
  dcvc_view out;
  const float *b1, *b2, *scale;
  float slope1, slope2;
  int c, n1, pixels, tiles, out_f16, kk1, nch, na;   // na: t-tile ring depth (2..4)
  int hbufs;                                         // 2: H double-buffered (c <= 48), 1: single
  uint32_t w1_off, w2_off, h_off, vec_off, bar_off;   // from the 1024-aligned base; A ring at 0
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
// bounded waits: a protocol bug must fault, not hang the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins)
    if (spins > SPIN_LIMIT) __trap();
}
__device__ __forceinline__ void mbar_wait_warp(uint32_t bar, uint32_t parity, int lane) {
  if (lane == 0) {
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins)   // try_wait suspends in hardware
      if (spins > SPIN_LIMIT) __trap();
  }
  __syncwarp();
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() {   // generic-proxy smem writes -> async proxy
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\tmov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t *r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait_ld() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
  const __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t *>(&h);
}
__device__ __forceinline__ float2 unpack_h2(uint32_t u) {
  return __half22float2(*reinterpret_cast<const __half2 *>(&u));
}
__device__ __forceinline__ float lrelu01(float v, float slope) { return fmaxf(v, v * slope); }
__device__ __forceinline__ uint32_t desc_lo32(uint32_t addr) { return (addr >> 4) | 0x10000u; }

__global__ void __launch_bounds__(THREADS, 1)
ffn_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w1,
              const __grid_constant__ CUtensorMap map_w2, const __grid_constant__ FfnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t *const base_g = smem_raw + (base - raw);
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bars = base + a.bar_off;
  // barriers: 0..3 a_full | 4 w_full | 5 acc1_full | 6 h_full | 7 acc2_full | 8 acc2_empty |
  // 9..12 a_empty
  auto bar = [&](int i) { return bars + 8u * i; };
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(bar(i), 1); mbar_init(bar(9 + i), EPI2_WARPS); }
    mbar_init(bar(4), 1); mbar_init(bar(5), 1);
    mbar_init(bar(6), EPI1_WARPS);
    mbar_init(bar(7), 1);
    mbar_init(bar(8), EPI2_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  {  // b1 | b2 | scale as smem vectors (broadcast reads in the epilogues)
    float *vec = reinterpret_cast<float *>(base_g + a.vec_off);
    for (int i = threadIdx.x; i < a.n1; i += blockDim.x) vec[i] = __ldg(a.b1 + i);
    for (int i = threadIdx.x; i < 64; i += blockDim.x) {
      vec[a.n1 + i] = i < a.c ? __ldg(a.b2 + i) : 0.f;
      vec[a.n1 + 64 + i] = (a.scale && i < a.c) ? __ldg(a.scale + i) : 1.f;
    }
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(&tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int total = a.tiles, step = gridDim.x;

  if (warp == 0) {
    // ===== TMA producer: weights once, then the t tiles through a ring of two
    const bool leader = elect_one();
    if (leader) {
      mbar_expect_tx(bar(4), (uint32_t)a.n1 * 128u + (uint32_t)a.nch * (uint32_t)a.c * 128u);
      tma_load_2d(base + a.w1_off, &map_w1, bar(4), 0, 0);
      for (int j = 0; j < a.nch; ++j)
        tma_load_2d(base + a.w2_off + (uint32_t)j * (uint32_t)a.c * 128u, &map_w2, bar(4), 0, j * a.c);
    }
    uint32_t s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < total; tile += step) {
      mbar_wait(bar(9 + s), ph ^ 1u);                     // EPI2 has read the residual of tile lt-na
      if (leader) {
        mbar_expect_tx(bar(s), A_BYTES);
        tma_load_2d(base + s * A_BYTES, &map_a, bar(s), 0, tile * TILE_M);
      }
      if (++s == (uint32_t)a.na) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1) {
    // ===== MMA issuer.  Order: MMA1(0) | MMA1(i+1), MMA2(i) ...
    //  * MMA1(i+1) overwrites acc1: EPI1(i) has read it (h_full(i) was waited on).
    //  * H is double-buffered: EPI1(i+1) fills H[(i+1)&1] while MMA2(i) reads H[i&1].  The MMA
    //    that last read H[(i+1)&1] is MMA2(i-1), issued before MMA1(i+1); EPI1(i+1) waits
    //    acc1_full(i+1), and a commit completes only when every earlier MMA of this thread
    //    has -- so MMA2(i-1) is done.
    //  * MMA2(i+1) overwrites acc2: waits acc2_empty (EPI2(i) has read it).
    const bool issue = elect_one();
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);      // SBO 1024 B, 128B swizzle
    const uint32_t idesc1 = (1u << 4) | ((uint32_t)(a.n1 >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
    const uint32_t idesc2 = (1u << 4) | ((uint32_t)(a.c >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
    const uint32_t a_lo0 = desc_lo32(base), w1_lo = desc_lo32(base + a.w1_off),
                   w2_lo = desc_lo32(base + a.w2_off), h_lo = desc_lo32(base + a.h_off);
    const uint32_t w2_step = ((uint32_t)a.c * 128u) >> 4, acc2 = tb + ACC2_COL;
    const int kk1 = a.kk1, nch = a.nch;
    mbar_wait(bar(4), 0u);
    uint32_t as = 0, aph = 0;
    const uint32_t na = (uint32_t)a.na;
    auto mma1 = [&]() {                                    // next tile of the ring
      mbar_wait(bar(as), aph);
      tc_fence_after();
      if (issue) {
        const uint32_t a_lo = a_lo0 + as * (A_BYTES >> 4);
        for (int kk = 0; kk < kk1; ++kk)
          umma_f16(tb, a_lo + 2 * kk, w1_lo + 2 * kk, hi, idesc1, kk ? 1u : 0u);
        umma_commit(bar(5));
      }
      if (++as == na) { as = 0; aph ^= 1u; }
    };
    uint32_t lt = 0;
    if ((int)blockIdx.x < total) mma1();
    const uint32_t h_stride = ((uint32_t)nch * 16384u) >> 4;
    const bool dbl = a.hbufs == 2;
    for (int tile = blockIdx.x; tile < total; tile += step, ++lt) {
      mbar_wait(bar(6), lt & 1u);                          // H[lt & 1] written, acc1 drained
      // two H buffers: MMA1(lt+1) first, so that EPI1(lt+1) runs under MMA2(lt); with a single
      // buffer MMA2(lt) must be issued (and, by commit order, finished) before acc1_full(lt+1)
      if (dbl && tile + step < total) mma1();
      mbar_wait(bar(8), (lt & 1u) ^ 1u);                   // acc2 drained by EPI2(lt-1)
      tc_fence_after();
      if (issue) {
        const uint32_t hb = h_lo + (dbl ? (lt & 1u) * h_stride : 0u);
        for (int j = 0; j < nch; ++j) {
          const uint32_t hj = hb + (uint32_t)j * (16384u >> 4), wj = w2_lo + (uint32_t)j * w2_step;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_f16(acc2, hj + 2 * kk, wj + 2 * kk, hi, idesc2, (j | kk) ? 1u : 0u);
        }
        umma_commit(bar(7));
      }
      if (!dbl && tile + step < total) mma1();
    }
  } else if (warp < 2 + EPI1_WARPS) {
    // ===== EPI1: acc1 -> + b1 -> LeakyReLU -> binary16 -> H (A operand of MMA2)
    const int e = warp - 2, q = warp & 3, part = e >> 2;
    const int ng = a.n1 >> 5;                              // 32-column groups
    const uint32_t r = (uint32_t)(q * 32 + lane), sw = r & 7u;
    uint8_t *const hrow0 = base_g + a.h_off + r * 128u;
    const uint32_t h_bytes = (uint32_t)a.nch * 16384u;
    const float4 *const b1v = reinterpret_cast<const float4 *>(base_g + a.vec_off);
    const float slope = a.slope1;
    uint32_t lt = 0;
    for (int tile = blockIdx.x; tile < total; tile += step, ++lt) {
      mbar_wait_warp(bar(5), lt & 1u, lane);
      tc_fence_after();
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
      uint8_t *const hrow = hrow0 + (a.hbufs == 2 ? (lt & 1u) * h_bytes : 0u);
      for (int g = part; g < ng; g += EPI1_WARPS / 4) {
        uint32_t u[32];
        tmem_ld16_nowait(trow + (uint32_t)g * 32u, u);
        tmem_ld16_nowait(trow + (uint32_t)g * 32u + 16u, u + 16);
        tmem_wait_ld();
        uint8_t *const chunk = hrow + (uint32_t)(g >> 1) * 16384u;
        const uint32_t i0 = (uint32_t)(g & 1) * 4u;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 ba = b1v[g * 8 + 2 * i], bb = b1v[g * 8 + 2 * i + 1];
          const uint32_t *w = u + 8 * i;
          uint4 pk;
          pk.x = pack_h2(lrelu01(__uint_as_float(w[0]) + ba.x, slope), lrelu01(__uint_as_float(w[1]) + ba.y, slope));
          pk.y = pack_h2(lrelu01(__uint_as_float(w[2]) + ba.z, slope), lrelu01(__uint_as_float(w[3]) + ba.w, slope));
          pk.z = pack_h2(lrelu01(__uint_as_float(w[4]) + bb.x, slope), lrelu01(__uint_as_float(w[5]) + bb.y, slope));
          pk.w = pack_h2(lrelu01(__uint_as_float(w[6]) + bb.z, slope), lrelu01(__uint_as_float(w[7]) + bb.w, slope));
          *reinterpret_cast<uint4 *>(chunk + (((i0 + (uint32_t)i) ^ sw) << 4)) = pk;
        }
      }
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(6));
    }
  } else {
    // ===== EPI2: acc2 -> + b2 -> LeakyReLU -> + t -> * scale -> global
    const int q = warp & 3;
    const uint32_t r = (uint32_t)(q * 32 + lane), sw = r & 7u;
    const float4 *const b2v = reinterpret_cast<const float4 *>(base_g + a.vec_off) + (a.n1 >> 2);
    const float4 *const scv = b2v + 16;
    const float slope = a.slope2;
    const int c = a.c;
    const bool has_scale = a.scale != nullptr;
    uint32_t lt = 0, s = 0, sph = 0;
    for (int tile = blockIdx.x; tile < total; tile += step, ++lt) {
      mbar_wait_warp(bar(7), lt & 1u, lane);
      tc_fence_after();
      const uint32_t trow = tmem_base + ACC2_COL + ((uint32_t)(q * 32) << 16);
      // the TMA write of the t tile is made visible to this warp by observing the stage's own
      // barrier; its phase cannot advance before this warp releases the stage below
      mbar_wait_warp(bar(s), sph, lane);
      const uint8_t *const arow = base_g + s * A_BYTES + r * 128u;
      const int64_t p = (int64_t)tile * TILE_M + r;
      const size_t o = (size_t)(p < a.pixels ? p : 0) * a.out.cstride + a.out.coff;
#pragma unroll
      for (int hp = 0; hp < 2; ++hp) {                     // columns 32 hp .. 32 hp + 31
        if (hp * 32 >= c) break;
        uint32_t u[32];
        tmem_ld16_nowait(trow + (uint32_t)hp * 32u, u);
        if (hp * 32 + 16 < c) tmem_ld16_nowait(trow + (uint32_t)hp * 32u + 16u, u + 16);
        tmem_wait_ld();
        if ((hp + 1) * 32 >= c) {                          // accumulator fully read
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(8));
        }
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
          const int i = hp * 4 + ii;                       // 16-byte slot: channels 8i..8i+7
          if (i * 8 >= c) continue;
          const uint4 rs = *reinterpret_cast<const uint4 *>(arow + (((uint32_t)i ^ sw) << 4));
          float v[8];
          const float4 ba = b2v[2 * i], bb = b2v[2 * i + 1];
          const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
          const uint32_t rw[4] = {rs.x, rs.y, rs.z, rs.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 t2 = unpack_h2(rw[j]);
            v[2 * j] = lrelu01(__uint_as_float(u[8 * ii + 2 * j]) + bias[2 * j], slope) + t2.x;
            v[2 * j + 1] = lrelu01(__uint_as_float(u[8 * ii + 2 * j + 1]) + bias[2 * j + 1], slope) + t2.y;
          }
          if (has_scale) {
            const float4 sa = scv[2 * i], sb = scv[2 * i + 1];
            v[0] *= sa.x; v[1] *= sa.y; v[2] *= sa.z; v[3] *= sa.w;
            v[4] *= sb.x; v[5] *= sb.y; v[6] *= sb.z; v[7] *= sb.w;
          }
          if (p < a.pixels) {
            if (a.out_f16) {
              uint4 pk;
              pk.x = pack_h2(v[0], v[1]); pk.y = pack_h2(v[2], v[3]);
              pk.z = pack_h2(v[4], v[5]); pk.w = pack_h2(v[6], v[7]);
              *reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(a.out.data) + o + 8 * i) = pk;
            } else {
              float4 *dst = reinterpret_cast<float4 *>(a.out.data + o + 8 * i);
              dst[0] = make_float4(v[0], v[1], v[2], v[3]);
              dst[1] = make_float4(v[4], v[5], v[6], v[7]);
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(9 + s));              // the A stage may be refilled
      if (++s == (uint32_t)a.na) { s = 0; sph ^= 1u; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
                 ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------------------------------------------ chunked variant
// ffn2_tc_kernel: the same function with the hidden dimension walked in chunks of HC = 64 or 128
// columns, so that
//   * C = 128 (hidden 512: W1 + W2 = 256 KB, acc1 alone would fill the TMEM) fits: per chunk a
//     128 x HC accumulator; the weights of a chunk (32 KB of W1, 32 KB of W2) stream through a
//     shared-memory ring from L2, loaded by a warp of their own;
//   * two chunks are in flight: acc1 and H are double-buffered per CHUNK and acc2 per TILE, so
//     MMA1(u+1) runs under EPI1(u), MMA2(u) under EPI1(u+1), EPI2(tile) under the next tile.
// Units u = (tile, chunk) in a flat order shared by all roles; buffer = u & 1.
//   MMA thread, iteration u:  MMA1(u) -> commit acc1_full[u&1];  then MMA2(u-1): wait h_full[(u-1)&1]
//   (EPI1(u-1) has drained acc1 and written H), wait acc2_empty on a tile's first chunk, issue,
//   commit acc2_full on its last.  Hazards, all closed by "a commit completes when every earlier
//   MMA of the issuing thread has":
//     acc1[b] rewritten by MMA1(u):  EPI1(u-2) done  <= h_full(u-2) was waited for MMA2(u-2);
//     H[b] rewritten by EPI1(u):     MMA2(u-2) done  <= issued before MMA1(u), EPI1(u) starts on
//                                                       acc1_full(u);
//     weight stage reuse:            commit -> wr_empty of the stage after the MMAs that read it.
constexpr int F2_EPI1_WARPS = 16, F2_EPI2_WARPS = 4;
constexpr int F2_THREADS = 96 + 32 * (F2_EPI1_WARPS + F2_EPI2_WARPS);
constexpr uint32_t ATOM = 16384;                 // 128 rows x 64 binary16, 128B swizzle

struct Ffn2Args {
  dcvc_view out;
  const float *b1, *b2, *scale;
  float slope1, slope2;
  int c, n1, pixels, tiles, out_f16;
  int hc, nhc, ka, kk1;        // hidden chunk (64 | 128), chunks per tile, K atoms of t, k16 steps of MMA1
  int na, nw, stream_w;        // t ring depth, weight ring depth (streamed), weights streamed?
  int acc2_col, acc2_stride;   // TMEM columns
  uint32_t w_off, h_off, vec_off, bar_off;   // from the 1024-aligned base; A ring at 0
};

__global__ void __launch_bounds__(F2_THREADS, 1)
ffn2_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w1,
               const __grid_constant__ CUtensorMap map_w2, const __grid_constant__ Ffn2Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t *const base_g = smem_raw + (base - raw);
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bars = base + a.bar_off;
  // barriers: 0..3 a_full | 4..7 a_empty | 8 w_full (resident) | 9..12 wr_full | 13..16 wr_empty |
  // 17,18 acc1_full | 19,20 h_full | 21,22 acc2_full | 23,24 acc2_empty
  auto bar = [&](int i) { return bars + 8u * i; };
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) {
      mbar_init(bar(i), 1); mbar_init(bar(4 + i), F2_EPI2_WARPS);
      mbar_init(bar(9 + i), 1); mbar_init(bar(13 + i), 1);
    }
    mbar_init(bar(8), 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(17 + i), 1); mbar_init(bar(19 + i), F2_EPI1_WARPS);
      mbar_init(bar(21 + i), 1); mbar_init(bar(23 + i), F2_EPI2_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  {  // b2 | scale as smem vectors (broadcast reads in EPI2); b1 is read through L1
    float *vec = reinterpret_cast<float *>(base_g + a.vec_off);
    for (int i = threadIdx.x; i < 128; i += blockDim.x) {
      vec[i] = i < a.c ? __ldg(a.b2 + i) : 0.f;
      vec[128 + i] = (a.scale && i < a.c) ? __ldg(a.scale + i) : 1.f;
    }
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(&tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int total = a.tiles, step = gridDim.x;
  const int nhc = a.nhc, ka = a.ka;
  const uint32_t a_bytes = (uint32_t)ka * ATOM;            // one t tile
  const uint32_t hatoms = (uint32_t)a.hc >> 6;             // atoms of one H chunk
  const uint32_t wblk = 2u * ATOM;                         // one streamed weight block (32 KB)

  if (warp == 0) {
    // ===== t producer (and the resident weights, once)
    const bool leader = elect_one();
    if (leader && !a.stream_w) {
      const uint32_t w1_bytes = (uint32_t)a.n1 * 128u, w2_bytes = (uint32_t)(a.n1 >> 6) * (uint32_t)a.c * 128u;
      mbar_expect_tx(bar(8), w1_bytes + w2_bytes);
      for (int j = 0; j < nhc; ++j)
        tma_load_2d(base + a.w_off + (uint32_t)j * (uint32_t)a.hc * 128u, &map_w1, bar(8), 0, j * a.hc);
      for (int q = 0; q < (a.n1 >> 6); ++q)
        tma_load_2d(base + a.w_off + w1_bytes + (uint32_t)q * (uint32_t)a.c * 128u, &map_w2, bar(8), 0, q * a.c);
    }
    uint32_t s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < total; tile += step) {
      mbar_wait(bar(4 + s), ph ^ 1u);                      // EPI2 has read the residual of tile lt-na
      if (leader) {
        mbar_expect_tx(bar(s), a_bytes);
        for (int k = 0; k < ka; ++k)
          tma_load_2d(base + s * a_bytes + (uint32_t)k * ATOM, &map_a, bar(s), 64 * k, tile * TILE_M);
      }
      if (++s == (uint32_t)a.na) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 2) {
    // ===== weight streamer (C = 128): blocks in the order the MMA thread consumes them:
    // W1(u), then W2(u-1).  A W1 block = chunk j's rows of both K atoms; a W2 block = the two
    // hidden atoms of chunk j.
    if (a.stream_w) {
      const bool leader = elect_one();
      uint32_t ws = 0, wph = 0;
      auto push = [&](bool first, int j) {
        mbar_wait(bar(13 + ws), wph ^ 1u);
        if (leader) {
          mbar_expect_tx(bar(9 + ws), wblk);
          const uint32_t dst = base + a.w_off + ws * wblk;
          if (first) {
            tma_load_2d(dst, &map_w1, bar(9 + ws), 0, j * a.hc);
            tma_load_2d(dst + ATOM, &map_w1, bar(9 + ws), 0, a.n1 + j * a.hc);
          } else {
            tma_load_2d(dst, &map_w2, bar(9 + ws), 0, (2 * j) * a.c);
            tma_load_2d(dst + ATOM, &map_w2, bar(9 + ws), 0, (2 * j + 1) * a.c);
          }
        }
        if (++ws == (uint32_t)a.nw) { ws = 0; wph ^= 1u; }
      };
      int prev = -1;
      for (int tile = blockIdx.x; tile < total; tile += step)
        for (int j = 0; j < nhc; ++j) {
          push(true, j);
          if (prev >= 0) push(false, prev);
          prev = j;
        }
      if (prev >= 0) push(false, prev);
    }
  } else if (warp == 1) {
    // ===== MMA issuer
    const bool issue = elect_one();
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);      // SBO 1024 B, 128B swizzle
    const uint32_t idesc1 = (1u << 4) | ((uint32_t)(a.hc >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
    const uint32_t idesc2 = (1u << 4) | ((uint32_t)(a.c >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
    const uint32_t a_lo0 = desc_lo32(base), w_lo = desc_lo32(base + a.w_off), h_lo = desc_lo32(base + a.h_off);
    const uint32_t w1_bytes = (uint32_t)a.n1 * 128u;
    const bool stream_w = a.stream_w != 0;
    const int kk1 = a.kk1;
    uint32_t ws = 0, wph = 0;
    if (!stream_w) mbar_wait(bar(8), 0u);
    // MMA2 of unit v = (tile index lt_v, chunk j_v)
    auto mma2 = [&](uint32_t v, uint32_t ltv, int jv) {
      mbar_wait(bar(19 + (v & 1u)), (v >> 1) & 1u);        // H[v&1] written, acc1[v&1] drained
      if (jv == 0) mbar_wait(bar(23 + (ltv & 1u)), ((ltv >> 1) & 1u) ^ 1u);   // acc2 drained
      uint32_t w2_lo;
      if (stream_w) {
        mbar_wait(bar(9 + ws), wph);
        w2_lo = w_lo + ws * (wblk >> 4);
      } else {
        w2_lo = w_lo + ((w1_bytes + (uint32_t)jv * hatoms * (uint32_t)a.c * 128u) >> 4);
      }
      tc_fence_after();
      if (issue) {
        const uint32_t acc2 = tb + (uint32_t)a.acc2_col + (ltv & 1u) * (uint32_t)a.acc2_stride;
        const uint32_t hb = h_lo + (v & 1u) * ((hatoms * ATOM) >> 4);
        const uint32_t w2_step = stream_w ? (ATOM >> 4) : (((uint32_t)a.c * 128u) >> 4);
        for (uint32_t q = 0; q < hatoms; ++q)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_f16(acc2, hb + q * (ATOM >> 4) + 2 * kk, w2_lo + q * w2_step + 2 * kk, hi, idesc2,
                     (jv | (int)q | kk) ? 1u : 0u);
        if (stream_w) umma_commit(bar(13 + ws));
        if (jv == nhc - 1) umma_commit(bar(21 + (ltv & 1u)));
      }
      if (stream_w && ++ws == (uint32_t)a.nw) { ws = 0; wph ^= 1u; }
    };
    uint32_t u = 0, lt = 0, as = 0, aph = 0;
    int prev_j = -1;
    uint32_t prev_lt = 0;
    for (int tile = blockIdx.x; tile < total; tile += step, ++lt) {
      mbar_wait(bar(as), aph);
      for (int j = 0; j < nhc; ++j, ++u) {
        uint32_t w1_lo;
        if (stream_w) {
          mbar_wait(bar(9 + ws), wph);
          w1_lo = w_lo + ws * (wblk >> 4);
        } else {
          w1_lo = w_lo + (((uint32_t)j * (uint32_t)a.hc * 128u) >> 4);
        }
        tc_fence_after();
        if (issue) {
          const uint32_t acc1 = tb + (u & 1u) * (uint32_t)a.hc;
          const uint32_t a_lo = a_lo0 + as * (a_bytes >> 4);
          for (int kk = 0; kk < kk1; ++kk) {
            const uint32_t at = (uint32_t)(kk >> 2) * (ATOM >> 4) + 2u * (uint32_t)(kk & 3);
            umma_f16(acc1, a_lo + at, w1_lo + at, hi, idesc1, kk ? 1u : 0u);
          }
          if (stream_w) umma_commit(bar(13 + ws));
          umma_commit(bar(17 + (u & 1u)));
        }
        if (stream_w && ++ws == (uint32_t)a.nw) { ws = 0; wph ^= 1u; }
        if (prev_j >= 0) mma2(u - 1, prev_lt, prev_j);
        prev_j = j; prev_lt = lt;
      }
      if (++as == (uint32_t)a.na) { as = 0; aph ^= 1u; }
    }
    if (prev_j >= 0) mma2(u - 1, prev_lt, prev_j);
  } else if (warp < 3 + F2_EPI1_WARPS) {
    // ===== EPI1: acc1[u&1] -> + b1 -> LeakyReLU -> binary16 -> H[u&1] (A operand of MMA2)
    const int e = warp - 3, q = warp & 3, part = e >> 2;
    const int ng = a.hc >> 5;                              // 32-column groups of a chunk
    const uint32_t r = (uint32_t)(q * 32 + lane), sw = r & 7u;
    uint8_t *const hrow0 = base_g + a.h_off + r * 128u;
    const float slope = a.slope1;
    uint32_t u = 0;
    for (int tile = blockIdx.x; tile < total; tile += step)
      for (int j = 0; j < nhc; ++j, ++u) {
        mbar_wait_warp(bar(17 + (u & 1u)), (u >> 1) & 1u, lane);
        tc_fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + (u & 1u) * (uint32_t)a.hc;
        uint8_t *const hrow = hrow0 + (u & 1u) * hatoms * ATOM;
        const float4 *const b1v = reinterpret_cast<const float4 *>(a.b1 + j * a.hc);
        for (int g = part; g < ng; g += F2_EPI1_WARPS / 4) {
          uint32_t v[32];
          tmem_ld16_nowait(trow + (uint32_t)g * 32u, v);
          tmem_ld16_nowait(trow + (uint32_t)g * 32u + 16u, v + 16);
          tmem_wait_ld();
          uint8_t *const chunk = hrow + (uint32_t)(g >> 1) * ATOM;
          const uint32_t i0 = (uint32_t)(g & 1) * 4u;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 ba = __ldg(b1v + g * 8 + 2 * i), bb = __ldg(b1v + g * 8 + 2 * i + 1);
            const uint32_t *w = v + 8 * i;
            uint4 pk;
            pk.x = pack_h2(lrelu01(__uint_as_float(w[0]) + ba.x, slope), lrelu01(__uint_as_float(w[1]) + ba.y, slope));
            pk.y = pack_h2(lrelu01(__uint_as_float(w[2]) + ba.z, slope), lrelu01(__uint_as_float(w[3]) + ba.w, slope));
            pk.z = pack_h2(lrelu01(__uint_as_float(w[4]) + bb.x, slope), lrelu01(__uint_as_float(w[5]) + bb.y, slope));
            pk.w = pack_h2(lrelu01(__uint_as_float(w[6]) + bb.z, slope), lrelu01(__uint_as_float(w[7]) + bb.w, slope));
            *reinterpret_cast<uint4 *>(chunk + (((i0 + (uint32_t)i) ^ sw) << 4)) = pk;
          }
        }
        fence_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(19 + (u & 1u)));
      }
  } else {
    // ===== EPI2: acc2[lt&1] -> + b2 -> LeakyReLU -> + t -> * scale -> global
    const int q = warp & 3;
    const uint32_t r = (uint32_t)(q * 32 + lane), sw = r & 7u;
    const float4 *const b2v = reinterpret_cast<const float4 *>(base_g + a.vec_off);
    const float4 *const scv = b2v + 32;
    const float slope = a.slope2;
    const int c = a.c;
    const bool has_scale = a.scale != nullptr;
    uint32_t lt = 0, s = 0, sph = 0;
    for (int tile = blockIdx.x; tile < total; tile += step, ++lt) {
      mbar_wait_warp(bar(21 + (lt & 1u)), (lt >> 1) & 1u, lane);
      tc_fence_after();
      const uint32_t trow = tmem_base + (uint32_t)a.acc2_col + (lt & 1u) * (uint32_t)a.acc2_stride +
                            ((uint32_t)(q * 32) << 16);
      mbar_wait_warp(bar(s), sph, lane);                   // the t tile's TMA write is visible
      const uint8_t *const arow = base_g + s * a_bytes + r * 128u;
      const int64_t p = (int64_t)tile * TILE_M + r;
      const size_t o = (size_t)(p < a.pixels ? p : 0) * a.out.cstride + a.out.coff;
      for (int hp = 0; hp * 32 < c; ++hp) {                // columns 32 hp .. 32 hp + 31
        uint32_t v32[32];
        tmem_ld16_nowait(trow + (uint32_t)hp * 32u, v32);
        if (hp * 32 + 16 < c) tmem_ld16_nowait(trow + (uint32_t)hp * 32u + 16u, v32 + 16);
        tmem_wait_ld();
        if ((hp + 1) * 32 >= c) {                          // accumulator fully read
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(23 + (lt & 1u)));
        }
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
          const int i = hp * 4 + ii;                       // 16-byte slot: channels 8i..8i+7
          if (i * 8 >= c) continue;
          const uint4 rs = *reinterpret_cast<const uint4 *>(arow + (uint32_t)(i >> 3) * ATOM +
                                                            ((((uint32_t)i & 7u) ^ sw) << 4));
          float v[8];
          const float4 ba = b2v[2 * i], bb = b2v[2 * i + 1];
          const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
          const uint32_t rw[4] = {rs.x, rs.y, rs.z, rs.w};
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const float2 t2 = unpack_h2(rw[jj]);
            v[2 * jj] = lrelu01(__uint_as_float(v32[8 * ii + 2 * jj]) + bias[2 * jj], slope) + t2.x;
            v[2 * jj + 1] = lrelu01(__uint_as_float(v32[8 * ii + 2 * jj + 1]) + bias[2 * jj + 1], slope) + t2.y;
          }
          if (has_scale) {
            const float4 sa = scv[2 * i], sb = scv[2 * i + 1];
            v[0] *= sa.x; v[1] *= sa.y; v[2] *= sa.z; v[3] *= sa.w;
            v[4] *= sb.x; v[5] *= sb.y; v[6] *= sb.z; v[7] *= sb.w;
          }
          if (p < a.pixels) {
            if (a.out_f16) {
              uint4 pk;
              pk.x = pack_h2(v[0], v[1]); pk.y = pack_h2(v[2], v[3]);
              pk.z = pack_h2(v[4], v[5]); pk.w = pack_h2(v[6], v[7]);
              *reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(a.out.data) + o + 8 * i) = pk;
            } else {
              float4 *dst = reinterpret_cast<float4 *>(a.out.data + o + 8 * i);
              dst[0] = make_float4(v[0], v[1], v[2], v[3]);
              dst[1] = make_float4(v[4], v[5], v[6], v[7]);
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(4 + s));              // the t stage may be refilled
      if (++s == (uint32_t)a.na) { s = 0; sph ^= 1u; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
                 ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

// binary16 matrix [rows][cols] with row pitch `pitch_bytes`, box (64 columns, box_rows), 128B swizzle
int map_2d(const void *ptr, uint64_t cols, uint64_t rows, uint64_t pitch_bytes, uint32_t box_rows,
           CUtensorMap *out) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(ptr), dims, strides,
                   box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled(ffn) failed: %d", (int)r);
  return DCVC_OK;
}

}  // namespace

// DCVC_FFN2: 0 = the chunked kernel only where the first cannot go (C = 128), 1 = everywhere
static int ffn2_policy() {
  static const int v = getenv("DCVC_FFN2") ? atoi(getenv("DCVC_FFN2")) : 0;
  return v;
}

extern "C" int dcvc_ffn_eligible(int channels, int64_t pixels) {
  static const int c128 = getenv("DCVC_FFN_C128") ? atoi(getenv("DCVC_FFN_C128")) : 1;
  return (channels == 32 || channels == 48 || channels == 64 || (channels == 128 && c128)) &&
         pixels >= 4 * TILE_M;
}

static int launch_ffn2(const dcvc_ffn_desc *d, int c, int64_t pixels, void *stream) {
  Ffn2Args a{};
  a.out = d->out; a.b1 = d->b1; a.b2 = d->b2; a.scale = d->scale;
  a.slope1 = d->slope1; a.slope2 = d->slope2;
  a.c = c; a.n1 = 4 * c; a.pixels = (int)pixels; a.tiles = cdiv(pixels, TILE_M);
  a.out_f16 = d->out_f16;
  a.hc = c >= 64 ? 128 : 64;
  a.nhc = a.n1 / a.hc;
  a.ka = c > 64 ? 2 : 1;
  a.kk1 = (c + 15) / 16;
  a.stream_w = c > 64;
  a.na = a.stream_w ? 2 : 4;
  a.nw = 3;
  a.acc2_col = 2 * a.hc;
  a.acc2_stride = (c + 31) / 32 * 32;
  const uint32_t a_bytes = (uint32_t)a.ka * ATOM;
  a.w_off = (uint32_t)a.na * a_bytes;
  const uint32_t w_bytes = a.stream_w ? (uint32_t)a.nw * 2u * ATOM
                                      : (uint32_t)a.n1 * 128u + (uint32_t)(a.n1 / 64) * (uint32_t)c * 128u;
  a.h_off = (a.w_off + w_bytes + 1023u) & ~1023u;
  a.vec_off = a.h_off + 2u * (uint32_t)(a.hc / 64) * ATOM;
  a.bar_off = a.vec_off + 1024u;
  const size_t smem = 1024 + (size_t)a.bar_off + 256;      // 25 barriers
  if (smem > 232448 - 64) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "ffn: smem %zu too large", smem);
  CUtensorMap ma, m1, m2;
  const __half *in = reinterpret_cast<const __half *>(d->in.data) + d->in.coff;
  int rc = map_2d(in, (uint64_t)c, (uint64_t)pixels, (uint64_t)d->in.cstride * 2, TILE_M, &ma);
  if (rc) return rc;
  if ((rc = map_2d(d->w1, 64, (uint64_t)a.ka * a.n1, 128, (uint32_t)a.hc, &m1))) return rc;
  if ((rc = map_2d(d->w2, 64, (uint64_t)(a.n1 / 64) * c, 128, (uint32_t)c, &m2))) return rc;
  int num_sms = 0;
  if ((rc = dcvc_device_info(nullptr, &num_sms))) return rc;
  rc = dcvc_once_per_device(DCVC_ONCE_FFN2_TC, [] {
    return cudaFuncSetAttribute(ffn2_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 64);
  });
  if (rc) return rc;
  const int grid = a.tiles < num_sms ? a.tiles : num_sms;
  ffn2_tc_kernel<<<grid, F2_THREADS, smem, (cudaStream_t)stream>>>(ma, m1, m2, a);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_ffn(const dcvc_ffn_desc *d, void *stream) {
  if (!d || !view_ok(&d->in) || !view_ok(&d->out) || !d->w1 || !d->w2 || !d->b1 || !d->b2)
    return dcvc_fail(DCVC_ERR_ARG, "ffn: bad arguments");
  const int c = d->in.c;
  const int64_t pixels = (int64_t)d->in.h * d->in.w;
  if (!dcvc_ffn_eligible(c, pixels) || d->out.c != c || d->out.h != d->in.h || d->out.w != d->in.w)
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "ffn: unsupported shape (c %d, %lld pixels)", c, (long long)pixels);
  auto al = [](const dcvc_view &v, int e) {      // whole 16-byte groups of e-byte elements
    const int m = 16 / e - 1;
    return ((v.cstride | v.coff) & m) == 0 && (reinterpret_cast<uintptr_t>(v.data) & 15u) == 0;
  };
  if (!al(d->in, 2) || !al(d->out, d->out_f16 ? 2 : 4) ||
      ((reinterpret_cast<uintptr_t>(d->w1) | reinterpret_cast<uintptr_t>(d->w2)) & 15u))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "ffn: views / weights must be 16-byte aligned");
  if (c > 64 || ffn2_policy()) return launch_ffn2(d, c, pixels, stream);
  FfnArgs a{};
  a.out = d->out; a.b1 = d->b1; a.b2 = d->b2; a.scale = d->scale;
  a.slope1 = d->slope1; a.slope2 = d->slope2;
  a.c = c; a.n1 = 4 * c; a.pixels = (int)pixels; a.tiles = cdiv(pixels, TILE_M);
  a.out_f16 = d->out_f16; a.kk1 = (c + 15) / 16; a.nch = a.n1 / 64;
  static const int na_env = getenv("DCVC_FFN_NA") ? atoi(getenv("DCVC_FFN_NA")) : 4;
  a.na = na_env < 2 ? 2 : na_env > 4 ? 4 : na_env;
  a.w1_off = (uint32_t)a.na * A_BYTES;
  a.w2_off = a.w1_off + (uint32_t)a.n1 * 128u;
  a.h_off = a.w2_off + (uint32_t)a.nch * (uint32_t)c * 128u;
  a.hbufs = c == 64 ? 1 : 2;                     // 64 (A) + 64 (weights) + 64 (H) KB at c = 64
  a.vec_off = a.h_off + (uint32_t)a.hbufs * (uint32_t)a.nch * 16384u;
  a.bar_off = a.vec_off + ((((uint32_t)(a.n1 + 128) * 4u) + 15u) & ~15u);
  const size_t smem = 1024 + (size_t)a.bar_off + 128;   // 13 barriers
  if (smem > 232448 - 64) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "ffn: smem %zu too large", smem);
  CUtensorMap ma, m1, m2;
  const __half *in = reinterpret_cast<const __half *>(d->in.data) + d->in.coff;
  int rc = map_2d(in, (uint64_t)c, (uint64_t)pixels, (uint64_t)d->in.cstride * 2, TILE_M, &ma);
  if (rc) return rc;
  if ((rc = map_2d(d->w1, 64, (uint64_t)a.n1, 128, (uint32_t)a.n1, &m1))) return rc;
  if ((rc = map_2d(d->w2, 64, (uint64_t)a.nch * c, 128, (uint32_t)c, &m2))) return rc;
  int num_sms = 0;
  if ((rc = dcvc_device_info(nullptr, &num_sms))) return rc;
  rc = dcvc_once_per_device(DCVC_ONCE_FFN_TC, [] {
    return cudaFuncSetAttribute(ffn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 64);
  });
  if (rc) return rc;
  const int grid = a.tiles < num_sms ? a.tiles : num_sms;
  ffn_tc_kernel<<<grid, THREADS, smem, (cudaStream_t)stream>>>(ma, m1, m2, a);
  DCVC_LAUNCHED();
  return DCVC_OK;
}
