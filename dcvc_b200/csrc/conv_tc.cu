// conv_tc.cu -- implicit-GEMM convolution on the 5th-generation tensor cores (sm_100a):
// tcgen05.mma kind::tf32 with fp32 accumulators in TMEM, operands staged in shared memory by
// TMA (tiled tensor maps: the im2col gather is a shifted 4-D box per filter tap, zero padding is
// TMA's out-of-bounds fill), mbarrier producer/consumer pipeline, warp-specialised roles.
//
// Replaces cuDNN convolutions of the reference (nn.Conv2d call sites, DCVC-DC/src/models/
// layers.py:18-39, video_net.py:58-251) for every shape with stride 1 or 2 whose NHWC view is
// 16-byte aligned; conv_direct.cu keeps the rest and the exact fp32 mode.
//
// GEMM view per CTA: M = 128 output pixels (a TH x TW patch), N = all output channels (padded
// to 16, <= 256), K = taps x Cin in chunks of KC in {32,16,8} floats = one 128/64/32-byte swizzle
// row per pixel.  A tile: TMA box [TH][TW][KC] -> 128 K-major rows; B tile: box [N][KC] of the
// packed weight matrix [N_pad][taps*Cin_pad].  One elected thread issues KC/8 MMAs per chunk.
//
// Roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> fused epilogue -> NHWC global stores).
#include <cuda.h>
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace {

constexpr int TILE_M = 128;
constexpr int MAX_THREADS = 64 + 32 * 16;   // warp 0 TMA, warp 1 MMA, then 4/8/16 epilogue warps
constexpr uint32_t SPIN_LIMIT = 1u << 26;   // failed try_wait polls before a wait gives up and traps

struct TcArgs {
  dcvc_view out, res1, res2, out2;
  const float *bias, *scale;
  int cout_w, n_tile, tiles_n, gate, shuffle, clamp01;
  float out_slope, gate_slope, out2_slope;
  int hout, wout, th, tw, tiles_x, total_tiles;
  int ksize, stride, pad, kc, chunks_per_tap, units, group, stages;
  uint32_t unit_a, unit_b, unit_tx, tmem_cols, acc_stride, sbo, layout_type, stage_off;
  uint64_t rcp_tiles_n, rcp_tiles_x;  // ceil(2^40 / d), see tile_coords()
  int fast;       // lean epilogue: 0 none, 1 plain, 2 gated, 3 planar depth-to-space (tc_fast_mode)
  // halo mode (stride-1 k>1 convs): A stage = whole input window, B ring per tap
  int halo, halo_sa, halo_sb, halo_pw, halo_bg;   // halo_bg: taps per B stage
  uint32_t halo_sbo;
  int bres;       // halo mode with the whole weight matrix resident in smem (loaded once per CTA)
  int tma_epi;    // 1: TMA-store epilogue (stage_off region = 8 KB per warp, vec_off = bias/scale)
  uint32_t vec_off;
  int epi_warps;  // 4, 8 or 16 epilogue warps (blockDim = 64 + 32 * epi_warps)
  int epi_sets;   // TMA-store epilogue: 2 = two sets of epi_warps / 2 warps, set s drains the tiles
                  // whose accumulator is s (every other tile of the CTA); 1 = all warps on every tile
  uint32_t epi_buf;  // bytes of one epilogue tile buffer (two per warp): 4096, or 2048 for binary16 rows
  int kcb;        // bytes of one K chunk row (128/64/32) = kc * element size
  int kk_last;    // MMAs of a tap's last chunk: its zero-padded tail is not multiplied at all
  int f16_in;     // A/B operands are binary16: kind::f16, K = 16 per instruction
  int f16_out;    // TMA-store epilogue writes binary16 (64-byte rows, 64B swizzle)
  int f16_res;    // res1 holds binary16 (fetched as 64-byte rows, 64B swizzle)
  int f16_res2;   // res2 holds binary16 (TMA-store epilogue: direct 16-byte row loads)
  int shuf_tma;   // TMA-store epilogue of a planar depth-to-space conv: one strided tensor map per
                  // sub-pixel (map_o, map_r, map_o2, map_o3), 32-column chunks never straddle one
  int out2_tma;   // TMA-store epilogue also stores out2 = lrelu(out) (element type of out) via map_o2
  int res_direct; // TMA-store epilogue: binary16 res1 by direct row loads issued ahead of the
                  // accumulator wait (registers), not by TMA into the outgoing tile
  int pdl;    // launched with programmatic stream serialisation (griddepcontrol.wait after the prologue)
  int debug;  // experiments: bit0 = no TMA, bit1 = no MMA, bit2 = no epilogue stores
};

constexpr int STG_PITCH = 36;                    // floats per staged row (32 + 4: conflict-free)
constexpr int STG_WARP_FLOATS = 32 * STG_PITCH;  // one epilogue warp's staging tile

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
// Loop invariants that come from the kernel parameter block: without this the compiler
// re-reads them from the constant bank inside the single-warp producer / MMA loops (it treats
// LDC as free to rematerialise), and every reload is ~50 cycles on a latency-bound chain.
__device__ __forceinline__ int keep(int v) { asm volatile("" : "+r"(v)); return v; }
__device__ __forceinline__ uint32_t keep(uint32_t v) { asm volatile("" : "+r"(v)); return v; }
__device__ __forceinline__ uint64_t keep(uint64_t v) { asm volatile("" : "+l"(v)); return v; }

// one lane of the (fully active) warp, the same one every time
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
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
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// bounded wait: a protocol bug must fault, not hang the GPU.  The retry loop is nothing but
// try_wait (which sleeps in hardware until the phase flips or a time slice ends) and a counter:
// reading the clock between polls costs ~140 cycles per hand-off (scripts/ub/mbar_pingpong.cu).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins)
    if (spins > SPIN_LIMIT) __trap();
}
// Whole-warp wait used by the epilogue warps: one lane polls (with a back-off) and the warp
// reconverges behind it.  32 lanes x 8-16 warps hammering try_wait on the tile barriers slows the
// shared-memory barrier unit for the two threads that drive the TMA/MMA pipeline.
__device__ __forceinline__ void mbar_wait_warp(uint32_t bar, uint32_t parity, int lane) {
  if (lane == 0) {
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins) {
      __nanosleep(40);
      if (spins > SPIN_LIMIT) __trap();
    }
  }
  __syncwarp();
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *map, uint32_t bar,
                                            int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
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
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(bar)
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float *v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t addr, uint32_t sbo_bytes,
                                                   uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;                               // LBO: unused for swizzled K-major
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;     // 8-row group stride
  d |= (uint64_t)1 << 46;                               // descriptor version (sm_100)
  d |= (uint64_t)(layout_type & 7) << 61;
  return d;
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}


// ---- TMA store side (epilogue): smem tile -> global through a tensor map, bulk async-groups
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar,
                                            int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, uint32_t src, int c0, int c1,
                                             int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_read() {   // all but the N newest groups have read smem
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {   // two floats -> half2 bits (RN)
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float2 unpack_h2(uint32_t h2) {
  float2 r;
  asm("{\n\t.reg .f16 lo, hi;\n\tmov.b32 {lo, hi}, %2;\n\tcvt.f32.f16 %0, lo;\n\tcvt.f32.f16 %1, hi;\n\t}"
      : "=f"(r.x), "=f"(r.y) : "r"(h2));
  return r;
}
__device__ __forceinline__ void fence_async_smem() {  // generic-proxy smem writes -> async proxy
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// ------------------------------------------------------------------ epilogue (coalesced phase)
// Code size matters here (the instruction cache is shared by three warp roles), so the general
// case is one out-of-line scalar routine and only the common fully-aligned case is vectorised.

// scalar fused epilogue for final channel c at final pixel (y, x)
__device__ __noinline__ void finish_store(const TcArgs &a, int y, int x, int c, float v) {
  v = lrelu(v, a.out_slope);
  if (a.res1.data) v += a.res1.data[vidx(a.res1, y, x) + c];
  if (a.res2.data) v += a.res2.data[vidx(a.res2, y, x) + c];
  if (a.scale) v *= a.scale[c];
  if (a.clamp01) v = fminf(fmaxf(v, 0.f), 1.f);
  a.out.data[vidx(a.out, y, x) + c] = v;
  if (a.out2.data) a.out2.data[vidx(a.out2, y, x) + c] = lrelu(v, a.out2_slope);
}

__device__ __forceinline__ bool view_vec4(const dcvc_view &v) {
  return ((v.cstride | v.coff) & 3) == 0 && (reinterpret_cast<uintptr_t>(v.data) & 15) == 0;
}

// phase 2 for one staged chunk (32 rows x ncols): lanes run along channels, so each group of
// 8 (or 4) lanes reads/writes one pixel's 128 (64) contiguous bytes.
// EPI: 0 plain, 1 gate (pairs), 2 shuffle (depth-to-space quads)
template <int EPI>
__device__ __noinline__ void epilogue_chunk(const TcArgs &a, const float *stg, int lane, int q,
                                            int oy0, int ox0, int n_first, int ncols) {
  const int vshift = ncols == 32 ? 3 : 2, vpr = 1 << vshift;
  const int c4 = lane & (vpr - 1), rstep = 32 >> vshift;
  const int n = n_first + 4 * c4;
  if (n >= a.cout_w) return;
  const int n_ok = min(4, a.cout_w - n);
  float4 b = make_float4(0.f, 0.f, 0.f, 0.f), sc = make_float4(1.f, 1.f, 1.f, 1.f);
  if (a.bias) {
    b.x = __ldg(a.bias + n);
    if (n_ok > 1) b.y = __ldg(a.bias + n + 1);
    if (n_ok > 2) b.z = __ldg(a.bias + n + 2);
    if (n_ok > 3) b.w = __ldg(a.bias + n + 3);
  }
  // fully vectorisable plain epilogue: every view 16-byte aligned and all four channels valid
  const bool fast = EPI == 0 && n_ok == 4 && view_vec4(a.out) &&
                    (!a.res1.data || view_vec4(a.res1)) && (!a.res2.data || view_vec4(a.res2)) &&
                    (!a.out2.data || view_vec4(a.out2)) &&
                    (!a.scale || (reinterpret_cast<uintptr_t>(a.scale) & 15) == 0);
  if (fast && a.scale) sc = __ldg(reinterpret_cast<const float4 *>(a.scale + n));
  const int tw_mask = a.tw - 1, tw_shift = 31 - __clz(a.tw);
#pragma unroll 2
  for (int r = lane >> vshift; r < 32; r += rstep) {
    const int row = q * 32 + r;
    const int oy = oy0 + (row >> tw_shift), ox = ox0 + (row & tw_mask);
    if (oy >= a.hout || ox >= a.wout) continue;
    float4 t = *reinterpret_cast<const float4 *>(stg + r * STG_PITCH + 4 * c4);
    t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
    if (fast) {
      t.x = lrelu(t.x, a.out_slope); t.y = lrelu(t.y, a.out_slope);
      t.z = lrelu(t.z, a.out_slope); t.w = lrelu(t.w, a.out_slope);
      if (a.res1.data) {
        const float4 u = *reinterpret_cast<const float4 *>(a.res1.data + vidx(a.res1, oy, ox) + n);
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      if (a.res2.data) {
        const float4 u = *reinterpret_cast<const float4 *>(a.res2.data + vidx(a.res2, oy, ox) + n);
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      t.x *= sc.x; t.y *= sc.y; t.z *= sc.z; t.w *= sc.w;
      if (a.clamp01) {
        t.x = fminf(fmaxf(t.x, 0.f), 1.f); t.y = fminf(fmaxf(t.y, 0.f), 1.f);
        t.z = fminf(fmaxf(t.z, 0.f), 1.f); t.w = fminf(fmaxf(t.w, 0.f), 1.f);
      }
      *reinterpret_cast<float4 *>(a.out.data + vidx(a.out, oy, ox) + n) = t;
      if (a.out2.data) {
        t.x = lrelu(t.x, a.out2_slope); t.y = lrelu(t.y, a.out2_slope);
        t.z = lrelu(t.z, a.out2_slope); t.w = lrelu(t.w, a.out2_slope);
        *reinterpret_cast<float4 *>(a.out2.data + vidx(a.out2, oy, ox) + n) = t;
      }
    } else if (EPI == 0) {
      const float w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll 1
      for (int j = 0; j < n_ok; ++j) finish_store(a, oy, ox, n + j, w[j]);
    } else if (EPI == 1) {
      if (n_ok >= 2) finish_store(a, oy, ox, n >> 1, t.x * lrelu(t.y, a.gate_slope));
      if (n_ok >= 4) finish_store(a, oy, ox, (n >> 1) + 1, t.z * lrelu(t.w, a.gate_slope));
    } else {
      // raw channel n+sp -> pixel (2oy + (sp>>1), 2ox + (sp&1)), channel n>>2
      const float w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll 1
      for (int sp = 0; sp < n_ok; ++sp)
        finish_store(a, 2 * oy + (sp >> 1), 2 * ox + (sp & 1), n >> 2, w[sp]);
    }
  }
}

// Lean phase 2 for the common case (host sets a.fast: no shuffle, every view and the bias/scale
// vectors 16-byte aligned, cout_w % 4 == 0, slopes in [0,1]).  `a` lives in the kernel parameter
// bank here (forceinline), so nothing is re-read through memory inside the row loop; ~35
// instructions per row of four channels.  GATE: raw column pairs (v, g) -> v * lrelu(g).
__device__ __forceinline__ float lrelu01(float v, float slope) { return fmaxf(v, v * slope); }

template <int MODE>
__device__ __forceinline__ void fast_chunk(const TcArgs &a, const float *stg, int lane, int q,
                                           int oy0, int ox0, int n_first, int ncols) {
  const int vshift = ncols == 32 ? 3 : 2;
  const int c4 = lane & ((1 << vshift) - 1), rstep = 32 >> vshift;
  const int n = n_first + 4 * c4;                      // raw GEMM column of this lane's float4
  if (n >= a.cout_w) return;
  constexpr bool GATE = MODE == 1, SHUF = MODE == 2;
  // SHUF: GEMM columns are ordered (subpixel sp, channel c); reference channel = 4c + sp
  const int c_out = a.cout_w >> 2;
  const int sp = SHUF ? n / c_out : 0;
  const int c = GATE ? n >> 1 : SHUF ? n - sp * c_out : n;   // first output channel of this lane
  float4 b = make_float4(0.f, 0.f, 0.f, 0.f), sc = make_float4(1.f, 1.f, 1.f, 1.f);
  if (a.bias) {
    if (SHUF) {
      b.x = __ldg(a.bias + 4 * c + sp);      b.y = __ldg(a.bias + 4 * c + 4 + sp);
      b.z = __ldg(a.bias + 4 * c + 8 + sp);  b.w = __ldg(a.bias + 4 * c + 12 + sp);
    } else {
      b = __ldg(reinterpret_cast<const float4 *>(a.bias + n));
    }
  }
  if (a.scale) {
    if (GATE) {
      const float2 s2 = __ldg(reinterpret_cast<const float2 *>(a.scale + c));
      sc.x = s2.x; sc.y = s2.y;
    } else {
      sc = __ldg(reinterpret_cast<const float4 *>(a.scale + c));
    }
  }
  float *const o = a.out.data + a.out.coff + c;
  const float *const r1 = a.res1.data ? a.res1.data + a.res1.coff + c : nullptr;
  const float *const r2 = a.res2.data ? a.res2.data + a.res2.coff + c : nullptr;
  float *const o2 = a.out2.data ? a.out2.data + a.out2.coff + c : nullptr;
  const int tw_mask = a.tw - 1, tw_shift = 31 - __clz(a.tw);
  const float *src = stg + (lane >> vshift) * STG_PITCH + 4 * c4;
  if (MODE == 0 && (r1 || r2)) {
    // residual epilogue: the loads of four rows are issued before any of them is consumed (the
    // stores may alias the residual views as far as the compiler knows, so it will not hoist
    // them itself, and one dependent global load per row leaves the warp latency-bound)
    for (int r = lane >> vshift; r < 32; r += 4 * rstep, src += 4 * rstep * STG_PITCH) {
      float4 u1[4], u2[4];
      int pix[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int row = q * 32 + r + j * rstep;
        const int oy = oy0 + (row >> tw_shift), ox = ox0 + (row & tw_mask);
        pix[j] = (oy < a.hout && ox < a.wout) ? oy * a.wout + ox : -1;
        u1[j] = u2[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (pix[j] >= 0 && r1)
          u1[j] = __ldg(reinterpret_cast<const float4 *>(r1 + (size_t)pix[j] * a.res1.cstride));
        if (pix[j] >= 0 && r2)
          u2[j] = __ldg(reinterpret_cast<const float4 *>(r2 + (size_t)pix[j] * a.res2.cstride));
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (pix[j] < 0) continue;
        float4 t = *reinterpret_cast<const float4 *>(src + j * rstep * STG_PITCH);
        t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
        if (a.out_slope != 1.f) {
          t.x = lrelu01(t.x, a.out_slope); t.y = lrelu01(t.y, a.out_slope);
          t.z = lrelu01(t.z, a.out_slope); t.w = lrelu01(t.w, a.out_slope);
        }
        // same order as the generic path: (v + res1) + res2
        t.x += u1[j].x; t.y += u1[j].y; t.z += u1[j].z; t.w += u1[j].w;
        if (r2) { t.x += u2[j].x; t.y += u2[j].y; t.z += u2[j].z; t.w += u2[j].w; }
        if (a.scale) { t.x *= sc.x; t.y *= sc.y; t.z *= sc.z; t.w *= sc.w; }
        if (a.clamp01) {
          t.x = fminf(fmaxf(t.x, 0.f), 1.f); t.y = fminf(fmaxf(t.y, 0.f), 1.f);
          t.z = fminf(fmaxf(t.z, 0.f), 1.f); t.w = fminf(fmaxf(t.w, 0.f), 1.f);
        }
        *reinterpret_cast<float4 *>(o + (size_t)pix[j] * a.out.cstride) = t;
        if (o2) {
          t.x = lrelu01(t.x, a.out2_slope); t.y = lrelu01(t.y, a.out2_slope);
          t.z = lrelu01(t.z, a.out2_slope); t.w = lrelu01(t.w, a.out2_slope);
          *reinterpret_cast<float4 *>(o2 + (size_t)pix[j] * a.out2.cstride) = t;
        }
      }
    }
    return;
  }
#pragma unroll 2
  for (int r = lane >> vshift; r < 32; r += rstep, src += rstep * STG_PITCH) {
    const int row = q * 32 + r;
    const int oy = oy0 + (row >> tw_shift), ox = ox0 + (row & tw_mask);
    if (oy >= a.hout || ox >= a.wout) continue;
    const int pix = SHUF ? (2 * oy + (sp >> 1)) * (2 * a.wout) + 2 * ox + (sp & 1) : oy * a.wout + ox;
    float4 t = *reinterpret_cast<const float4 *>(src);
    t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
    if (GATE) {
      float2 g = make_float2(t.x * lrelu01(t.y, a.gate_slope), t.z * lrelu01(t.w, a.gate_slope));
      if (a.out_slope != 1.f) { g.x = lrelu01(g.x, a.out_slope); g.y = lrelu01(g.y, a.out_slope); }
      if (r1) {
        const float2 u = __ldg(reinterpret_cast<const float2 *>(r1 + (size_t)pix * a.res1.cstride));
        g.x += u.x; g.y += u.y;
      }
      if (r2) {
        const float2 u = __ldg(reinterpret_cast<const float2 *>(r2 + (size_t)pix * a.res2.cstride));
        g.x += u.x; g.y += u.y;
      }
      g.x *= sc.x; g.y *= sc.y;
      if (a.clamp01) { g.x = fminf(fmaxf(g.x, 0.f), 1.f); g.y = fminf(fmaxf(g.y, 0.f), 1.f); }
      *reinterpret_cast<float2 *>(o + (size_t)pix * a.out.cstride) = g;
      if (o2) {
        g.x = lrelu01(g.x, a.out2_slope); g.y = lrelu01(g.y, a.out2_slope);
        *reinterpret_cast<float2 *>(o2 + (size_t)pix * a.out2.cstride) = g;
      }
    } else {
      if (a.out_slope != 1.f) {
        t.x = lrelu01(t.x, a.out_slope); t.y = lrelu01(t.y, a.out_slope);
        t.z = lrelu01(t.z, a.out_slope); t.w = lrelu01(t.w, a.out_slope);
      }
      if (r1) {
        const float4 u = __ldg(reinterpret_cast<const float4 *>(r1 + (size_t)pix * a.res1.cstride));
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      if (r2) {
        const float4 u = __ldg(reinterpret_cast<const float4 *>(r2 + (size_t)pix * a.res2.cstride));
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      if (a.scale) { t.x *= sc.x; t.y *= sc.y; t.z *= sc.z; t.w *= sc.w; }
      if (a.clamp01) {
        t.x = fminf(fmaxf(t.x, 0.f), 1.f); t.y = fminf(fmaxf(t.y, 0.f), 1.f);
        t.z = fminf(fmaxf(t.z, 0.f), 1.f); t.w = fminf(fmaxf(t.w, 0.f), 1.f);
      }
      if (SHUF && a.f16_out) {
        // binary16 destination (no residual / second output on this path: host-checked)
        *reinterpret_cast<uint2 *>(reinterpret_cast<uint16_t *>(a.out.data) + a.out.coff + c +
                                   (size_t)pix * a.out.cstride) =
            make_uint2(pack_h2(t.x, t.y), pack_h2(t.z, t.w));
        continue;
      }
      *reinterpret_cast<float4 *>(o + (size_t)pix * a.out.cstride) = t;
      if (o2) {
        t.x = lrelu01(t.x, a.out2_slope); t.y = lrelu01(t.y, a.out2_slope);
        t.z = lrelu01(t.z, a.out2_slope); t.w = lrelu01(t.w, a.out2_slope);
        *reinterpret_cast<float4 *>(o2 + (size_t)pix * a.out2.cstride) = t;
      }
    }
  }
}

// tile -> (m tile, n tile, tile row, tile column) with host-made reciprocals:
// q = (n * ceil(2^40 / d)) >> 40 is exact for n * d < 2^40
__device__ __forceinline__ void tile_coords(const TcArgs &a, int tile, int &mt, int &nt, int &ty,
                                            int &tx) {
  mt = (int)(((uint64_t)(uint32_t)tile * a.rcp_tiles_n) >> 40);
  nt = tile - mt * a.tiles_n;
  ty = (int)(((uint64_t)(uint32_t)mt * a.rcp_tiles_x) >> 40);
  tx = mt - ty * a.tiles_x;
}

// ------------------------------------------------------------------ MMA issue
// The issuing warp's loops are written so that every operand of tcgen05.mma is provably
// warp-uniform (derived from kernel parameters, blockIdx and loop counters only -- no value that
// went through an inline-asm "+r" or shared memory) and so that stepping a descriptor is ONE
// 32-bit add on its low word: the matrix descriptor is passed as (low word, constant high word).
// Measured (scripts/ub/mma_rate.cu): back-to-back UTCHMMA with M = 128 retire one per
// max(N/2, 32 + N/4) cycles (SS operands: 128 + N rows of 32 bytes through the 128 B/clk smem
// port), i.e. 45 cycles at N = 48 -- the previous loops spent 110-170 cycles per MMA on
// vector->uniform register moves (R2UR) and 64-bit descriptor arithmetic.
template <bool F16>
__device__ __forceinline__ void umma2(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo,
                                      uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
  if (F16)
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %6, 0;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
        ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %6, 0;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %5, p;\n\t}"
        ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
// high word of a shared-memory matrix descriptor (8-row group stride, version, swizzle mode);
// the low word is (address >> 4) | 1 << 16 (LBO = 1, unused for swizzled K-major operands)
__device__ __forceinline__ uint32_t desc_hi32(uint32_t sbo_bytes, uint32_t layout_type) {
  return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14) | ((layout_type & 7u) << 29);
}
__device__ __forceinline__ uint32_t desc_lo32(uint32_t addr) { return (addr >> 4) | 0x10000u; }

#ifdef DCVC_TC_TIMELINE
__device__ unsigned long long g_tl[2][4096][8];
#define TL(role, idx, k) do { if (blockIdx.x == 0 && (idx) < 4096) g_tl[role][idx][k] = clock64(); } while (0)
#else
#define TL(role, idx, k) do {} while (0)
#endif

struct MmaCtx {
  uint32_t base, bars, tmem, idesc, a_hi, b_hi;
  int stages;
  bool issue;
  __device__ __forceinline__ uint32_t full_bar(int s) const { return bars + 8u * s; }
  __device__ __forceinline__ uint32_t empty_bar(int s) const { return bars + 8u * (stages + s); }
  __device__ __forceinline__ uint32_t tfull_bar(int s) const { return bars + 8u * (2 * stages + s); }
  __device__ __forceinline__ uint32_t tempty_bar(int s) const { return bars + 8u * (2 * stages + 2 + s); }
};

// resident weights (halo window per tile, every tap's B tile in place): KS*KS*KK MMAs per tile
// with compile-time offsets from the window / weight bases
template <int KS, int KK, bool F16>
__device__ __forceinline__ void mma_bres(const TcArgs &a, const MmaCtx &c) {
  const int total = a.total_tiles, step = gridDim.x, sa_n = a.halo_sa;
  const uint32_t p4 = (uint32_t)a.kcb >> 4, pwp4 = (uint32_t)(a.halo_pw * a.kcb) >> 4,
                 ub4 = a.unit_b >> 4, ua4 = a.unit_a >> 4, acc_stride = a.acc_stride;
  const uint32_t a4_0 = desc_lo32(c.base), b4 = desc_lo32(c.base + sa_n * a.unit_a);
  uint32_t as = 0, aph = 0, pha = 0, a4 = a4_0;
  int sa = 0;
  mbar_wait(c.full_bar(sa_n), 0u);               // the weights, once
  int tli = 0;
  for (int tile = blockIdx.x; tile < total; tile += step, ++tli) {
    if (c.issue) TL(0, tli, 0);
    mbar_wait(c.tempty_bar(as), aph ^ 1u);
    if (c.issue) TL(0, tli, 1);
    mbar_wait(c.full_bar(sa), pha);
    if (c.issue) TL(0, tli, 2);
    tc_fence_after();
    const uint32_t tacc = c.tmem + as * acc_stride;
    if (c.issue) {
#pragma unroll(KS <= 3 ? KS : 1)
      for (int dy = 0; dy < KS; ++dy) {
        const uint32_t ra = a4 + dy * pwp4, rb = b4 + dy * KS * ub4;
#pragma unroll
        for (int dx = 0; dx < KS; ++dx)
#pragma unroll
          for (int kk = 0; kk < KK; ++kk)
            umma2<F16>(tacc, ra + dx * p4 + 2 * kk, c.a_hi, rb + dx * ub4 + 2 * kk, c.b_hi, c.idesc,
                       (dx | kk) ? 1u : (uint32_t)dy);
      }
      umma_commit(c.empty_bar(sa));
      umma_commit(c.tfull_bar(as));
      TL(0, tli, 3);
    }
    a4 += ua4;
    if (++sa == sa_n) { sa = 0; pha ^= 1u; a4 = a4_0; }
    if ((as ^= 1u) == 0) aph ^= 1u;
  }
}

// resident weights with several channel chunks per tap (unit = tap * chunks + chunk): one window
// per (tile, chunk); KK MMAs per full chunk, kk_last in the last one
template <int KS, int KK, bool F16>
__device__ __forceinline__ void mma_bres_mc(const TcArgs &a, const MmaCtx &c) {
  const int total = a.total_tiles, step = gridDim.x, sa_n = a.halo_sa, chunks = a.chunks_per_tap,
            kk_last = a.kk_last;
  const uint32_t p4 = (uint32_t)a.kcb >> 4, pwp4 = (uint32_t)(a.halo_pw * a.kcb) >> 4,
                 ub4 = a.unit_b >> 4, ua4 = a.unit_a >> 4, acc_stride = a.acc_stride;
  const uint32_t tap4 = ub4 * (uint32_t)chunks;        // B stride between taps
  const uint32_t a4_0 = desc_lo32(c.base), b4_0 = desc_lo32(c.base + sa_n * a.unit_a);
  uint32_t as = 0, aph = 0, pha = 0, a4 = a4_0;
  int sa = 0;
  mbar_wait(c.full_bar(sa_n), 0u);               // the weights, once
  for (int tile = blockIdx.x; tile < total; tile += step) {
    mbar_wait(c.tempty_bar(as), aph ^ 1u);
    const uint32_t tacc = c.tmem + as * acc_stride;
    uint32_t b4 = b4_0;
    for (int ch = 0; ch < chunks; ++ch, b4 += ub4) {
      mbar_wait(c.full_bar(sa), pha);
      tc_fence_after();
      const int kn = ch == chunks - 1 ? kk_last : KK;
      if (c.issue) {
#pragma unroll
        for (int dy = 0; dy < KS; ++dy) {
          const uint32_t ra = a4 + dy * pwp4, rb = b4 + dy * KS * tap4;
#pragma unroll
          for (int dx = 0; dx < KS; ++dx)
#pragma unroll
            for (int kk = 0; kk < KK; ++kk)
              if (kk < kn)
                umma2<F16>(tacc, ra + dx * p4 + 2 * kk, c.a_hi, rb + dx * tap4 + 2 * kk, c.b_hi,
                           c.idesc, (dy | dx | kk) ? 1u : (uint32_t)ch);
        }
        umma_commit(c.empty_bar(sa));
      }
      a4 += ua4;
      if (++sa == sa_n) { sa = 0; pha ^= 1u; a4 = a4_0; }
    }
    if (c.issue) umma_commit(c.tfull_bar(as));
    if ((as ^= 1u) == 0) aph ^= 1u;
  }
}

// halo windows with streamed weights: per (tile, channel chunk) one window, taps walk it;
// B stages carry `halo_bg` consecutive taps
template <int KK, bool F16>
__device__ __forceinline__ void mma_halo(const TcArgs &a, const MmaCtx &c) {
  const int total = a.total_tiles, step = gridDim.x, sa_n = a.halo_sa, sb_n = a.halo_sb;
  const int ks = a.ksize, taps = ks * ks, chunks = a.chunks_per_tap, bg = a.halo_bg;
  const int kk_last = a.kk_last;
  const uint32_t p4 = (uint32_t)a.kcb >> 4, row4 = ((uint32_t)(a.halo_pw * a.kcb) >> 4) - ks * p4,
                 ub4 = a.unit_b >> 4, ua4 = a.unit_a >> 4, sb4 = (uint32_t)(a.halo_bg * a.unit_b) >> 4,
                 acc_stride = a.acc_stride;
  const uint32_t a4_0 = desc_lo32(c.base), b4_0 = desc_lo32(c.base + sa_n * a.unit_a);
  uint32_t as = 0, aph = 0, pha = 0, phb = 0, a4 = a4_0, b4 = b4_0;
  int sa = 0, sb = 0;
  for (int tile = blockIdx.x; tile < total; tile += step) {
    mbar_wait(c.tempty_bar(as), aph ^ 1u);
    tc_fence_after();
    const uint32_t tacc = c.tmem + as * acc_stride;
    uint32_t accf = 0;
    for (int ch = 0; ch < chunks; ++ch) {
      mbar_wait(c.full_bar(sa), pha);
      tc_fence_after();
      const int kn = ch == chunks - 1 ? kk_last : KK;
      uint32_t a_t = a4;
      int dx = 0;
      for (int t0 = 0; t0 < taps; t0 += bg) {          // one B stage = `bg` consecutive taps
        const int sbar = sa_n + sb, cnt = min(bg, taps - t0);
        mbar_wait(c.full_bar(sbar), phb);
        tc_fence_after();
        uint32_t b_t = b4;
        for (int g = 0; g < cnt; ++g) {
          if (c.issue) {
#pragma unroll
            for (int kk = 0; kk < KK; ++kk)
              if (kk < kn)
                umma2<F16>(tacc, a_t + 2 * kk, c.a_hi, b_t + 2 * kk, c.b_hi, c.idesc,
                           kk ? 1u : accf);
          }
          accf = 1u;
          b_t += ub4;
          a_t += p4;
          if (++dx == ks) { dx = 0; a_t += row4; }
        }
        if (c.issue) umma_commit(c.empty_bar(sbar));
        b4 += sb4;
        if (++sb == sb_n) { sb = 0; phb ^= 1u; b4 = b4_0; }
      }
      if (c.issue) umma_commit(c.empty_bar(sa));
      a4 += ua4;
      if (++sa == sa_n) { sa = 0; pha ^= 1u; a4 = a4_0; }
    }
    if (c.issue) umma_commit(c.tfull_bar(as));
    if ((as ^= 1u) == 0) aph ^= 1u;
  }
}

// one stage = `group` (tap, channel-chunk) units of [A tile | B tile]
template <int KK, bool F16>
__device__ __forceinline__ void mma_plain(const TcArgs &a, const MmaCtx &c) {
  const int total = a.total_tiles, step = gridDim.x, stages = a.stages, group = a.group,
            units = a.units, last_ch = a.chunks_per_tap - 1, kk_last = a.kk_last;
  const int groups = (units + group - 1) / group;
  const uint32_t ua4 = a.unit_a >> 4, unit4 = (a.unit_a + a.unit_b) >> 4,
                 stage4 = (uint32_t)group * unit4, acc_stride = a.acc_stride;
  const uint32_t a4_0 = desc_lo32(c.base);
  uint32_t as = 0, aph = 0, ph = 0, s4 = a4_0;
  int s = 0;
  for (int tile = blockIdx.x; tile < total; tile += step) {
    mbar_wait(c.tempty_bar(as), aph ^ 1u);       // epilogue has drained this accumulator
    tc_fence_after();
    const uint32_t tacc = c.tmem + as * acc_stride;
    uint32_t accf = 0;
    int u = 0, ch = 0;
    for (int gi = 0; gi < groups; ++gi) {
      mbar_wait(c.full_bar(s), ph);
      tc_fence_after();
      const int cnt = min(group, units - u);
      uint32_t a_t = s4;
      for (int g = 0; g < cnt; ++g) {
        const int kn = ch == last_ch ? kk_last : KK;
        if (c.issue) {
#pragma unroll
          for (int kk = 0; kk < KK; ++kk)
            if (kk < kn)
              umma2<F16>(tacc, a_t + 2 * kk, c.a_hi, a_t + ua4 + 2 * kk, c.b_hi, c.idesc,
                         kk ? 1u : accf);
        }
        accf = 1u;
        a_t += unit4;
        ch = ch == last_ch ? 0 : ch + 1;
      }
      u += cnt;
      if (c.issue) umma_commit(c.empty_bar(s));   // frees the smem slot when these MMAs retire
      s4 += stage4;
      if (++s == stages) { s = 0; ph ^= 1u; s4 = a4_0; }
    }
    if (c.issue) umma_commit(c.tfull_bar(as));    // accumulator complete
    if ((as ^= 1u) == 0) aph ^= 1u;
  }
}

template <bool F16>
__device__ __forceinline__ void mma_role(const TcArgs &a, uint32_t base, uint32_t bars,
                                         uint32_t tmem, bool issue) {
  MmaCtx c;
  c.base = base; c.bars = bars; c.tmem = tmem; c.issue = issue; c.stages = a.stages;
  // instruction descriptor: D = F32 (bit 4), A/B format (bits 7, 10: 2 = TF32, 0 = F16),
  // K-major both, N >> 3 at bit 17, M >> 4 at bit 24
  c.idesc = (1u << 4) | (F16 ? 0u : (2u << 7) | (2u << 10)) | ((uint32_t)(a.n_tile >> 3) << 17) |
            ((uint32_t)(TILE_M >> 4) << 24);
  c.b_hi = desc_hi32(a.sbo, a.layout_type);
  c.a_hi = a.halo ? desc_hi32(a.halo_sbo, a.layout_type) : c.b_hi;
  const int kk_n = a.kcb >> 5;     // one MMA consumes 32 bytes of K per row: 8 TF32 / 16 binary16
  if (a.bres && a.chunks_per_tap > 1) {      // 3x3 only (host)
    if (kk_n == 4) mma_bres_mc<3, 4, F16>(a, c);
    else if (kk_n == 2) mma_bres_mc<3, 2, F16>(a, c);
    else mma_bres_mc<3, 1, F16>(a, c);
  } else if (a.bres) {
    const int key = a.ksize * 8 + a.kk_last;
    switch (key) {
      case 3 * 8 + 1: mma_bres<3, 1, F16>(a, c); break;
      case 3 * 8 + 2: mma_bres<3, 2, F16>(a, c); break;
      case 3 * 8 + 3: mma_bres<3, 3, F16>(a, c); break;
      case 3 * 8 + 4: mma_bres<3, 4, F16>(a, c); break;
      case 5 * 8 + 1: mma_bres<5, 1, F16>(a, c); break;
      case 7 * 8 + 1: mma_bres<7, 1, F16>(a, c); break;
      default: __trap();           // the host only selects resident weights for these shapes
    }
  } else if (a.halo) {
    if (kk_n == 4) mma_halo<4, F16>(a, c);
    else if (kk_n == 2) mma_halo<2, F16>(a, c);
    else mma_halo<1, F16>(a, c);
  } else {
    if (kk_n == 4) mma_plain<4, F16>(a, c);
    else if (kk_n == 2) mma_plain<2, F16>(a, c);
    else mma_plain<1, F16>(a, c);
  }
}

// ------------------------------------------------------------------ the kernel
// Persistent: each CTA walks tiles blockIdx.x, +gridDim.x, ...; the smem ring and the two TMEM
// accumulators keep the TMA producer, the MMA issuer and the epilogue warps of consecutive
// tiles overlapped.
__global__ void __launch_bounds__(MAX_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const __grid_constant__ CUtensorMap map_o, const __grid_constant__ CUtensorMap map_r,
               const __grid_constant__ CUtensorMap map_o2, const __grid_constant__ CUtensorMap map_o3,
               const __grid_constant__ TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the swizzled tiles
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t unit_bytes = keep(a.unit_a + a.unit_b);
  const uint32_t stage_bytes = keep(a.group * unit_bytes);
  const uint32_t bars = keep(base + a.stage_off - 1024u);   // full[s], empty[s], tfull[2], tempty[2], res
  float *staging = reinterpret_cast<float *>(smem_raw + (base - raw) + a.stage_off);
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int groups = keep((a.units + a.group - 1) / a.group);

  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (a.stages + s); };
  auto tfull_bar = [&](int s) { return bars + 8u * (2 * a.stages + s); };
  auto tempty_bar = [&](int s) { return bars + 8u * (2 * a.stages + 2 + s); };
  auto res_bar = [&](int s) { return bars + 8u * (2 * a.stages + 4 + s); };   // 2 per epilogue warp

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), a.epi_warps / a.epi_sets);
    }
    if (a.tma_epi)
      for (int s = 0; s < 2 * a.epi_warps; ++s) mbar_init(res_bar(s), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(&tmem_base_slot)), "r"(a.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // Programmatic dependent launch: the barrier initialisation and the TMEM allocation above may run
  // while the previous kernel of the stream is still draining; nothing below reads or writes
  // global memory before that kernel has completed and flushed (the bias / scale vectors
  // included: they may have been packed by a kernel launched just before this one).  The next
  // kernel's prologue may start as soon as it finds room.
  if (a.pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;");
  }
  if (a.tma_epi) {
    // bias / scale of every GEMM column, padded with 0 / 1, for broadcast reads in the epilogue
    float *vec = reinterpret_cast<float *>(smem_raw + (base - raw) + a.vec_off);
    const int n_all = a.tiles_n * ((a.n_tile + 31) & ~31);   // whole 32-column chunks
    const int c_sp = a.cout_w >> 2;     // depth-to-space: column = (subpixel, channel), bias[4c+sp]
    for (int i = threadIdx.x; i < n_all; i += blockDim.x) {
      const int sp = a.shuf_tma ? i / c_sp : 0, c = a.shuf_tma ? i - sp * c_sp : i;
      vec[i] = (a.bias && i < a.cout_w) ? __ldg(a.bias + (a.shuf_tma ? 4 * c + sp : i)) : 0.f;
      vec[n_all + i] = (a.scale && i < a.cout_w) ? __ldg(a.scale + c) : 1.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  // Warps 0 and 1 run their loops warp-wide (uniform control flow, operands in uniform
  // registers) and only the asynchronous instructions themselves are issued by one elected lane:
  // wrapping the whole loop in `if (lane == 0)` makes the compiler treat every operand as
  // thread-divergent and surround each UTCHMMA / UTMALDG with an elect-and-retry loop.
  if (warp == 0) {
    // ===== TMA producer =====
    const bool leader = elect_one();
    const int total = keep((int)(a.total_tiles)), step = keep((int)(gridDim.x)), kc = keep((int)(a.kc)), n_tile = keep((int)(a.n_tile));
    const int chunks = keep((int)(a.chunks_per_tap)), ksize = keep((int)(a.ksize)), pad = keep((int)(a.pad));
    const bool no_tma = (a.debug & 1) != 0;
    if (a.halo) {
      // ----- halo mode: one A load per (tile, channel chunk) = the (TH+k-1) x PW pixel window
      // all k*k taps read through shifted descriptors; B streams per tap through its own ring.
      // A of step i+1 is issued before the B taps of step i so it has a whole step to arrive.
      const int taps = keep((int)(ksize * ksize)), sa_n = keep((int)(a.halo_sa)), sb_n = keep((int)(a.halo_sb));
      const int bg = keep((int)(a.halo_bg));
      const uint32_t stage_b = keep((uint32_t)(a.halo_bg * a.unit_b));
      const uint32_t unit_a = a.unit_a, unit_b = a.unit_b, tx_a = a.unit_tx,
                     tx_b = (uint32_t)(n_tile * a.kcb), b_base = base + sa_n * unit_a;
      int sa = 0, sb = 0;
      uint32_t pha = 0, phb = 0;
      if (a.bres) {
        // ----- resident weights: every tap's B tile is loaded once and stays; the ring carries
        // input windows only, one TMA load per tile
        const int units = keep((int)(taps * chunks));
        if (leader) {
          mbar_expect_tx(full_bar(sa_n), no_tma ? 0u : (uint32_t)units * tx_b);
          if (!no_tma)
            for (int g = 0; g < units; ++g)               // unit g = tap * chunks + chunk
              tma_load_2d(b_base + g * unit_b, &map_b, full_bar(sa_n), g * kc, 0);
        }
        for (int tile = blockIdx.x; tile < total; tile += step) {
          int mt, nt, ty, tx;
          tile_coords(a, tile, mt, nt, ty, tx);
          for (int ch = 0; ch < chunks; ++ch) {
            mbar_wait(empty_bar(sa), pha ^ 1u);
            if (leader) {
              mbar_expect_tx(full_bar(sa), no_tma ? 0u : tx_a);
              if (!no_tma)
                tma_load_4d(base + sa * unit_a, &map_a, full_bar(sa), ch * kc, tx * a.tw - pad,
                            ty * a.th - pad, 0);
            }
            if (++sa == sa_n) { sa = 0; pha ^= 1u; }
          }
        }
      }
      int a_tile = a.bres ? total : blockIdx.x, a_ch = 0;   // iterator of the A loads
      auto issue_a = [&]() {
        if (a_tile >= total) return;
        int mt, nt, ty, tx;
        tile_coords(a, a_tile, mt, nt, ty, tx);
        mbar_wait(empty_bar(sa), pha ^ 1u);
        if (leader) {
          mbar_expect_tx(full_bar(sa), no_tma ? 0u : tx_a);
          if (!no_tma)
            tma_load_4d(base + sa * unit_a, &map_a, full_bar(sa), a_ch * kc, tx * a.tw - pad,
                        ty * a.th - pad, 0);
        }
        if (++sa == sa_n) { sa = 0; pha ^= 1u; }
        if (++a_ch == chunks) { a_ch = 0; a_tile += step; }
      };
      issue_a();
      for (int tile = a.bres ? total : (int)blockIdx.x; tile < total; tile += step) {
        int mt, nt, ty, tx;
        tile_coords(a, tile, mt, nt, ty, tx);
        const int n_base = nt * n_tile;
        for (int ch = 0; ch < chunks; ++ch) {
          issue_a();
          for (int t0 = 0; t0 < taps; t0 += bg) {          // one B stage = `bg` consecutive taps
            const int sbar = sa_n + sb, cnt = min(bg, taps - t0);
            mbar_wait(empty_bar(sbar), phb ^ 1u);
            if (leader) {
              mbar_expect_tx(full_bar(sbar), no_tma ? 0u : (uint32_t)cnt * tx_b);
              if (!no_tma)
                for (int g = 0; g < cnt; ++g)
                  tma_load_2d(b_base + sb * stage_b + g * unit_b, &map_b, full_bar(sbar),
                              ((t0 + g) * chunks + ch) * kc, n_base);
            }
            if (++sb == sb_n) { sb = 0; phb ^= 1u; }
          }
        }
      }
    } else {
      // ----- one stage = `group` (tap, channel-chunk) units, running counters instead of
      // divisions (a single instruction stream is latency-bound instruction by instruction)
      const int stages = keep((int)(a.stages)), group = keep((int)(a.group)), units = keep((int)(a.units)), stride = keep((int)(a.stride));
      const uint32_t unit_a = keep((uint32_t)(a.unit_a)), unit_tx = keep((uint32_t)(a.unit_tx));
      int s = 0;
      uint32_t ph = 0, stage_addr = base;
      for (int tile = blockIdx.x; tile < total; tile += step) {
        int mt, nt, ty, tx;
        tile_coords(a, tile, mt, nt, ty, tx);
        const int x0 = tx * a.tw * stride - pad, y0 = ty * a.th * stride - pad;
        const int n_base = nt * n_tile;
        int u = 0, dx = 0, dy = 0, ch = 0;
        for (int gi = 0; gi < groups; ++gi) {
          mbar_wait(empty_bar(s), ph ^ 1u);
          const int cnt = min(group, units - u);
          if (leader) mbar_expect_tx(full_bar(s), no_tma ? 0u : cnt * unit_tx);
          uint32_t sa = stage_addr;
          for (int g = 0; g < cnt; ++g, ++u, sa += unit_bytes) {
            if (leader && !no_tma) {
              tma_load_4d(sa, &map_a, full_bar(s), ch * kc, x0 + dx, y0 + dy, 0);
              tma_load_2d(sa + unit_a, &map_b, full_bar(s), u * kc, n_base);
            }
            if (++ch == chunks) {
              ch = 0;
              if (++dx == ksize) { dx = 0; ++dy; }
            }
          }
          stage_addr += stage_bytes;
          if (++s == stages) { s = 0; ph ^= 1u; stage_addr = base; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const bool issue = elect_one() && !(a.debug & 2);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);   // provably warp-uniform
    if (a.f16_in) mma_role<true>(a, base, bars, tb, issue);
    else mma_role<false>(a, base, bars, tb, issue);
  } else if (a.tma_epi) {
    // ===== epilogue, TMA-store flavour (plain convs): each warp owns TMEM lanes 32q..32q+31 =
    // 32 output pixels, and walks its 32-column chunks.  Thread = pixel row: tcgen05.ld gives it
    // 32 consecutive channels; bias / activation / residual / scale / clamp are applied in
    // registers, the row is parked in a 128B-swizzled 4 KB smem tile, and one elected lane sends
    // the tile out with a tensor-map store (clipping ragged edges and padded channels).  The
    // residual tile arrives the same way (TMA load into the very buffer the result will leave
    // from), so no per-element address arithmetic or global LD/ST is issued by the warp.
    const bool leader = elect_one();
    // two sets (epi_sets = 2): the epilogue of a narrow tile is a chain of fixed costs (barrier
    // waits, tile coordinates, fence, elected store: ~2000 cycles per tile against ~1450 of MMA
    // issue, scripts/tc_timeline.py), so set s takes the tiles of accumulator s only and each set
    // has two MMA periods per tile
    const int we = warp - 2, q = warp & 3, per_set = a.epi_warps / a.epi_sets, eset = we / per_set;
    const int part = (we - eset * per_set) >> 2, parts = per_set >> 2;
    const int estep = a.epi_sets * (int)gridDim.x;
    const uint32_t ebuf = a.epi_buf;
    const uint32_t buf0 = base + a.stage_off + (uint32_t)we * 2u * ebuf;
    uint8_t *const buf0_g = smem_raw + (base - raw) + a.stage_off + (size_t)we * 2u * ebuf;
    const float4 *const vec4 = reinterpret_cast<const float4 *>(smem_raw + (base - raw) + a.vec_off);
    const int n_all4 = (a.tiles_n * ((a.n_tile + 31) & ~31)) >> 2;
    const int nchunks = (a.n_tile + 31) >> 5;
    const int tw_mask = a.tw - 1, tw_shift = 31 - __clz(a.tw);
    const int bx = (q * 32) & tw_mask, by = (q * 32) >> tw_shift;
    const bool direct1 = a.res1.data != nullptr && a.res_direct;
    const bool has_res = a.res1.data != nullptr && !direct1, has_res2 = a.res2.data != nullptr;
    const bool pre2 = has_res2 && a.f16_res2;
    const int wb = a.tw < 32 ? a.tw : 32, wb_mask = wb - 1, wb_shift = 31 - __clz(wb);
    const uint32_t res_bytes = a.f16_res ? 2048u : 4096u;
    const bool mixed = has_res && (a.f16_res != a.f16_out);   // residual / result rows differ in pitch
    const uint32_t row_off = (uint32_t)lane * 128u, sw = (uint32_t)(lane & 7);
    const uint32_t row16 = (uint32_t)lane * 64u, sw4 = (uint32_t)(lane >> 1) & 3u;
    uint32_t rph0 = 0, rph1 = 0, b = 0, lt = (uint32_t)eset;
    int tile = blockIdx.x + eset * (int)gridDim.x, ci = part;
    if (has_res && leader && tile < a.total_tiles) {
      int mt, nt, ty, tx;
      tile_coords(a, tile, mt, nt, ty, tx);
      mbar_expect_tx(res_bar(2 * we), res_bytes);
      tma_load_3d(buf0, &map_r, res_bar(2 * we), nt * a.n_tile + ci * 32, tx * a.tw + bx,
                  ty * a.th + by);
    }
    while (tile < a.total_tiles) {
      int mt, nt, ty, tx;
      tile_coords(a, tile, mt, nt, ty, tx);
      const int n0 = ci * 32, c0 = nt * a.n_tile + n0, x0 = tx * a.tw + bx, y0 = ty * a.th + by;
      const uint32_t as = lt & 1u, aph = (lt >> 1) & 1u;
      // binary16 residual rows of this thread's pixel: requested before the accumulator wait, so
      // their latency hides behind the MMAs of the tile
      uint4 pr1[4], pr2[4];
      if (direct1 || pre2) {
        const int px = x0 + (lane & wb_mask), py = y0 + (lane >> wb_shift);
        const bool inb = px < a.wout && py < a.hout;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          pr1[j] = pr2[j] = make_uint4(0u, 0u, 0u, 0u);
          if (inb && c0 + 8 * j < a.cout_w) {
            if (direct1)
              pr1[j] = __ldg(reinterpret_cast<const uint4 *>(
                           reinterpret_cast<const uint16_t *>(a.res1.data) +
                           ((size_t)py * a.res1.w + px) * a.res1.cstride + a.res1.coff + c0) + j);
            if (pre2)
              pr2[j] = __ldg(reinterpret_cast<const uint4 *>(
                           reinterpret_cast<const uint16_t *>(a.res2.data) +
                           ((size_t)py * a.res2.w + px) * a.res2.cstride + a.res2.coff + c0) + j);
          }
        }
      }
      if (we == 0 && lane == 0) TL(1, (int)lt, 0);
      if (ci == part) {                                   // first chunk of this tile for the warp
        mbar_wait_warp(tfull_bar(as), aph, lane);
        tc_fence_after();
      }
      const int tl_i = (int)lt;
      if (we == 0 && lane == 0) TL(1, tl_i, 1);
      const uint32_t trow = tmem_base + as * a.acc_stride + ((uint32_t)(q * 32) << 16);
      float v[32];
      if (!(a.debug & 16)) {
        tmem_ld16(trow + (uint32_t)n0, v);
        if (a.n_tile - n0 >= 32) tmem_ld16(trow + (uint32_t)n0 + 16u, v + 16);
      }
      // next chunk of this warp (possibly in its next tile)
      int ntile = tile, nci = ci + parts;
      if (nci >= nchunks) { nci = part; ntile += estep; }
      if ((direct1 || pre2) && ntile < a.total_tiles) {
        // the NEXT chunk's residual rows towards L2 now: their direct loads are issued just
        // before the next accumulator wait, too late to hide a DRAM round trip behind it
        int mt2, nt2, ty2, tx2;
        tile_coords(a, ntile, mt2, nt2, ty2, tx2);
        const int px2 = tx2 * a.tw + bx + (lane & wb_mask), py2 = ty2 * a.th + by + (lane >> wb_shift);
        const int cn = nt2 * a.n_tile + nci * 32;
        if (px2 < a.wout && py2 < a.hout && cn < a.cout_w) {
          if (direct1) {
            const uint16_t *rp = reinterpret_cast<const uint16_t *>(a.res1.data) +
                                 ((size_t)py2 * a.res1.w + px2) * a.res1.cstride + a.res1.coff + cn;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + 16));
          }
          if (pre2) {
            const uint16_t *rp = reinterpret_cast<const uint16_t *>(a.res2.data) +
                                 ((size_t)py2 * a.res2.w + px2) * a.res2.cstride + a.res2.coff + cn;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + 16));
          }
        }
      }
      if (nci == part) {                                  // accumulator fully read: hand it back
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(as));
        lt += (uint32_t)a.epi_sets;
      }
      const uint32_t buf = buf0 + b * ebuf;
      uint8_t *const buf_g = buf0_g + b * ebuf;
      if (we == 0 && lane == 0) TL(1, tl_i, 2);
      if (!(a.debug & 4)) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bb = vec4[(c0 >> 2) + j];
          v[4 * j] += bb.x; v[4 * j + 1] += bb.y; v[4 * j + 2] += bb.z; v[4 * j + 3] += bb.w;
        }
        if (a.out_slope != 1.f) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = lrelu01(v[j], a.out_slope);
        }
        if (we == 0 && lane == 0) TL(1, tl_i, 4);
        if (direct1) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float2 f;
            f = unpack_h2(pr1[j].x); v[8 * j] += f.x;     v[8 * j + 1] += f.y;
            f = unpack_h2(pr1[j].y); v[8 * j + 2] += f.x; v[8 * j + 3] += f.y;
            f = unpack_h2(pr1[j].z); v[8 * j + 4] += f.x; v[8 * j + 5] += f.y;
            f = unpack_h2(pr1[j].w); v[8 * j + 6] += f.x; v[8 * j + 7] += f.y;
          }
        }
        if (has_res) {
          mbar_wait_warp(res_bar(2 * we + b), b ? rph1 : rph0, lane);
          if (b) rph1 ^= 1u; else rph0 ^= 1u;
          if (a.f16_res) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 u = *reinterpret_cast<const uint4 *>(buf_g + row16 + (((uint32_t)j ^ sw4) << 4));
              float2 f;
              f = unpack_h2(u.x); v[8 * j] += f.x;     v[8 * j + 1] += f.y;
              f = unpack_h2(u.y); v[8 * j + 2] += f.x; v[8 * j + 3] += f.y;
              f = unpack_h2(u.z); v[8 * j + 4] += f.x; v[8 * j + 5] += f.y;
              f = unpack_h2(u.w); v[8 * j + 6] += f.x; v[8 * j + 7] += f.y;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 u = *reinterpret_cast<const float4 *>(buf_g + row_off + (((uint32_t)j ^ sw) << 4));
              v[4 * j] += u.x; v[4 * j + 1] += u.y; v[4 * j + 2] += u.z; v[4 * j + 3] += u.w;
            }
          }
          // rows of the two element sizes overlap other lanes' rows: everyone reads first
          if (mixed) __syncwarp();
        }
        if (has_res2) {
          // second residual (fp32 view): the thread's own pixel row, eight 16-byte loads; rows
          // outside the image or the real channels are skipped (their results are clipped by the
          // tensor-map store anyway)
          const int px = x0 + (lane & wb_mask), py = y0 + (lane >> wb_shift);
          if (px < a.wout && py < a.hout) {
            const size_t off = ((size_t)py * a.res2.w + px) * a.res2.cstride + a.res2.coff + c0;
            if (a.f16_res2) {
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (c0 + 8 * j < a.cout_w) {
                  const uint4 u = pr2[j];
                  float2 f;
                  f = unpack_h2(u.x); v[8 * j] += f.x;     v[8 * j + 1] += f.y;
                  f = unpack_h2(u.y); v[8 * j + 2] += f.x; v[8 * j + 3] += f.y;
                  f = unpack_h2(u.z); v[8 * j + 4] += f.x; v[8 * j + 5] += f.y;
                  f = unpack_h2(u.w); v[8 * j + 6] += f.x; v[8 * j + 7] += f.y;
                }
            } else {
              const float4 *r2 = reinterpret_cast<const float4 *>(a.res2.data + off);
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (c0 + 4 * j < a.cout_w) {
                  const float4 u = __ldg(r2 + j);
                  v[4 * j] += u.x; v[4 * j + 1] += u.y; v[4 * j + 2] += u.z; v[4 * j + 3] += u.w;
                }
            }
          }
        }
        if (a.scale) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 s4 = vec4[n_all4 + (c0 >> 2) + j];
            v[4 * j] *= s4.x; v[4 * j + 1] *= s4.y; v[4 * j + 2] *= s4.z; v[4 * j + 3] *= s4.w;
          }
        }
        if (a.clamp01) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fminf(fmaxf(v[j], 0.f), 1.f);
        }
        auto park = [&](uint8_t *dst) {
          if (a.f16_out) {
            // binary16 rows of 64 bytes, 64B swizzle: 16-byte chunk j of row r sits at
            // j ^ ((r >> 1) & 3) (address bits [4,6) ^= bits [7,9))
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 pk;
              pk.x = pack_h2(v[8 * j], v[8 * j + 1]);     pk.y = pack_h2(v[8 * j + 2], v[8 * j + 3]);
              pk.z = pack_h2(v[8 * j + 4], v[8 * j + 5]); pk.w = pack_h2(v[8 * j + 6], v[8 * j + 7]);
              *reinterpret_cast<uint4 *>(dst + row16 + (((uint32_t)j ^ sw4) << 4)) = pk;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *reinterpret_cast<float4 *>(dst + row_off + (((uint32_t)j ^ sw) << 4)) =
                  make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
        };
        if (we == 0 && lane == 0) TL(1, tl_i, 5);
        park(buf_g);
        fence_async_smem();
        __syncwarp();
        if (we == 0 && lane == 0) TL(1, tl_i, 6);
        if (leader) {
          if (a.shuf_tma) {
            const int c_sp = a.cout_w >> 2, sp = c0 / c_sp;
            if (sp < 4)                                   // (padding columns have no destination)
              tma_store_3d(sp == 0 ? &map_o : sp == 1 ? &map_r : sp == 2 ? &map_o2 : &map_o3, buf,
                           c0 - sp * c_sp, x0, y0);
          } else {
            tma_store_3d(&map_o, buf, c0, x0, y0);
          }
          bulk_commit();
        }
        if (a.out2_tma) {
          // second output lrelu(out) (the next ResBlock's activated input) from the warp's OTHER
          // tile buffer; this mode never toggles `b`, so buffer b carries out, b^1 carries out2
          if (leader) bulk_wait_read<1>();                // out2 store of the previous chunk has left
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = lrelu01(v[j], a.out2_slope);
          park(buf0_g + (b ^ 1u) * ebuf);
          fence_async_smem();
          __syncwarp();
          if (leader) {
            tma_store_3d(&map_o2, buf0 + (b ^ 1u) * ebuf, c0, x0, y0);
            bulk_commit();
          }
        }
      } else if (has_res) {                               // experiment: no stores, still drain loads
        mbar_wait_warp(res_bar(2 * we + b), b ? rph1 : rph0, lane);
        if (b) rph1 ^= 1u; else rph0 ^= 1u;
      }
      if (we == 0 && lane == 0) TL(1, tl_i, 7);
      if (leader) {
        bulk_wait_read<1>();                              // the store issued a chunk ago has left smem
        if (has_res && ntile < a.total_tiles) {
          int mt2, nt2, ty2, tx2;
          tile_coords(a, ntile, mt2, nt2, ty2, tx2);
          mbar_expect_tx(res_bar(2 * we + (b ^ 1u)), res_bytes);
          tma_load_3d(buf0 + (b ^ 1u) * ebuf, &map_r, res_bar(2 * we + (b ^ 1u)),
                      nt2 * a.n_tile + nci * 32, tx2 * a.tw + bx, ty2 * a.th + by);
        }
      }
      __syncwarp();
      if (we == 0 && lane == 0) TL(1, tl_i, 3);
      if (!a.out2_tma) b ^= 1u;
      tile = ntile;
      ci = nci;
    }
    if (leader) bulk_wait_all();
  } else {
    // ===== epilogue: warps 2..2+epi_warps; warp w owns TMEM lanes 32*(w%4).., column chunks
    // ci with ci % (epi_warps/4) == part.  Phase 1: thread = pixel row, tcgen05.ld 32 columns, park them in a
    // padded smem tile; phase 2: lanes along channels -> coalesced fused epilogue.
    const int q = warp & 3, part = (warp - 2) >> 2, parts = a.epi_warps >> 2;
    float *stg = staging + (warp - 2) * STG_WARP_FLOATS;
    const int nchunks = (a.n_tile + 31) >> 5;
    uint32_t lt = 0;
    for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++lt) {
      int mt, nt, ty, tx;
      tile_coords(a, tile, mt, nt, ty, tx);
      const int ox0 = tx * a.tw, oy0 = ty * a.th, n_base = nt * a.n_tile;
      const uint32_t as = lt & 1u, aph = (lt >> 1) & 1u;
      mbar_wait_warp(tfull_bar(as), aph, lane);
      tc_fence_after();
      const uint32_t trow = tmem_base + as * a.acc_stride + ((uint32_t)(q * 32) << 16);
      int last = -1;
      for (int ci = part; ci < nchunks; ci += parts) last = ci;
      if (last < 0) {                                  // no columns for this warp: just release
        if (lane == 0) mbar_arrive(tempty_bar(as));
        continue;
      }
      for (int ci = part; ci < nchunks; ci += parts) {
        const int n0 = ci * 32;
        const int ncols = min(32, a.n_tile - n0);      // 16 or 32
        float v[32];
        if (!(a.debug & 16)) {
          tmem_ld16(trow + (uint32_t)n0, v);           // warp-collective
          if (ncols == 32) tmem_ld16(trow + (uint32_t)n0 + 16u, v + 16);
        }
        if (ci == last) {                              // accumulator fully read: hand it back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(as));
        }
        if (n_base + n0 >= a.cout_w || (a.debug & 4)) continue;   // warp-uniform
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (4 * j < ncols)
            *reinterpret_cast<float4 *>(stg + lane * STG_PITCH + 4 * j) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
        if (a.fast == 1) fast_chunk<0>(a, stg, lane, q, oy0, ox0, n_base + n0, ncols);
        else if (a.fast == 2) fast_chunk<1>(a, stg, lane, q, oy0, ox0, n_base + n0, ncols);
        else if (a.fast == 3) fast_chunk<2>(a, stg, lane, q, oy0, ox0, n_base + n0, ncols);
        else if (a.shuffle) epilogue_chunk<2>(a, stg, lane, q, oy0, ox0, n_base + n0, ncols);
        else if (a.gate) epilogue_chunk<1>(a, stg, lane, q, oy0, ox0, n_base + n0, ncols);
        else epilogue_chunk<0>(a, stg, lane, q, oy0, ox0, n_base + n0, ncols);
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
                 ::"r"(tmem_base), "r"(a.tmem_cols)
                 : "memory");
  }
}

// ------------------------------------------------------------------ host: tensor maps
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

struct MapKey {
  uint64_t ptr, d0, d1, d2, s1, s2;
  uint32_t b0, b1, b2, e1, e2, swz, rank;
  bool operator==(const MapKey &o) const {
    return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && d2 == o.d2 && s1 == o.s1 && s2 == o.s2 &&
           b0 == o.b0 && b1 == o.b1 && b2 == o.b2 && e1 == o.e1 && e2 == o.e2 && swz == o.swz &&
           rank == o.rank;
  }
};
struct MapHash {
  size_t operator()(const MapKey &k) const {
    uint64_t h = k.ptr * 0x9E3779B97F4A7C15ull;
    const uint64_t f[] = {k.d0, k.d1, k.d2, k.s1, k.s2, k.b0, k.b1, k.b2, k.e1, k.e2, k.swz, k.rank};
    for (uint64_t v : f) h = (h ^ v) * 0x100000001B3ull;
    return (size_t)h;
  }
};
std::unordered_map<MapKey, CUtensorMap, MapHash> g_maps;
std::mutex g_maps_mu;

CUtensorMapSwizzle swizzle_of(int row_bytes) {
  return row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                                        : CU_TENSOR_MAP_SWIZZLE_32B;
}
// element size / tensor-map type of a view; the MapKey carries the type in the high bits of rank
inline uint64_t esz_of(bool f16) { return f16 ? 2 : 4; }
inline CUtensorMapDataType dt_of(bool f16) {
  return f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
}

// activation map: dims (C, W, H, 1), box (KC, TW', TH', 1), traversal strides (1, s, s, 1)
int make_map_a(const dcvc_view &in, int kc, int th, int tw, int stride, bool f16, CUtensorMap *out) {
  MapKey k{};
  const uint64_t esz = esz_of(f16);
  k.ptr = (uint64_t)in.data + (uint64_t)in.coff * esz;
  k.d0 = in.c; k.d1 = in.w; k.d2 = in.h;
  k.s1 = (uint64_t)in.cstride * esz; k.s2 = (uint64_t)in.cstride * esz * in.w;
  k.b0 = kc; k.b1 = stride == 1 ? tw : tw * 2 - 1; k.b2 = stride == 1 ? th : th * 2 - 1;
  k.e1 = stride; k.e2 = stride; k.swz = swizzle_of(kc * (int)esz); k.rank = 4 | (f16 ? 0x100 : 0);
  std::lock_guard<std::mutex> g(g_maps_mu);
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return DCVC_OK; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[4] = {k.d0, k.d1, k.d2, 1};
  cuuint64_t strides[3] = {k.s1, k.s2, k.s2 * in.h};
  cuuint32_t box[4] = {k.b0, k.b1, k.b2, 1};
  cuuint32_t es[4] = {1, k.e1, k.e2, 1};
  CUresult r = enc(out, dt_of(f16), 4, (void *)k.ptr, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)k.swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled(A) failed: %d", (int)r);
  if (g_maps.size() > 8192) g_maps.clear();
  g_maps[k] = *out;
  return DCVC_OK;
}

// epilogue map over an output-side view (out / residual): dims (C, W, H), box (32 channels,
// w_box, 32 / w_box rows) = the 32 pixels one epilogue warp owns, 128B swizzle
// sub > 0: the view is the OUTPUT of a depth-to-space conv and the map addresses sub-pixel
// (dy, dx) = ((sub-1) >> 1, (sub-1) & 1) of every 2x2 block: an (h/2, w/2) tensor with doubled strides
int make_map_epi(const dcvc_view &v, int channels, int w_box, bool f16, CUtensorMap *out, int sub = 0) {
  MapKey k{};
  const uint64_t esz = esz_of(f16);
  k.ptr = (uint64_t)v.data + (uint64_t)v.coff * esz;
  k.d0 = channels; k.d1 = v.w; k.d2 = v.h;
  k.s1 = (uint64_t)v.cstride * esz; k.s2 = (uint64_t)v.cstride * esz * v.w;
  if (sub > 0) {
    const int dy = (sub - 1) >> 1, dx = (sub - 1) & 1;
    k.ptr += ((uint64_t)dy * v.w + dx) * v.cstride * esz;
    k.d1 = v.w / 2; k.d2 = v.h / 2; k.s1 *= 2; k.s2 *= 2;
  }
  k.b0 = 32; k.b1 = w_box; k.b2 = 32 / w_box;
  k.e1 = 1; k.e2 = 1; k.swz = swizzle_of(32 * (int)esz); k.rank = 3 | (f16 ? 0x100 : 0);
  std::lock_guard<std::mutex> g(g_maps_mu);
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return DCVC_OK; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[3] = {k.d0, k.d1, k.d2};
  cuuint64_t strides[2] = {k.s1, k.s2};
  cuuint32_t box[3] = {k.b0, k.b1, k.b2};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = enc(out, dt_of(f16), 3, (void *)k.ptr, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)k.swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled(epilogue) failed: %d", (int)r);
  if (g_maps.size() > 8192) g_maps.clear();
  g_maps[k] = *out;
  return DCVC_OK;
}

// weight map: dims (Ktot, N_pad), box (KC, N_tile)
int make_map_b(const void *w, int ktot, int n_pad, int n_tile, int kc, bool f16, CUtensorMap *out) {
  MapKey k{};
  const uint64_t esz = esz_of(f16);
  k.ptr = (uint64_t)w; k.d0 = ktot; k.d1 = n_pad; k.s1 = (uint64_t)ktot * esz;
  k.b0 = kc; k.b1 = n_tile; k.swz = swizzle_of(kc * (int)esz); k.rank = 2 | (f16 ? 0x100 : 0);
  std::lock_guard<std::mutex> g(g_maps_mu);
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return DCVC_OK; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[2] = {k.d0, k.d1};
  cuuint64_t strides[1] = {k.s1};
  cuuint32_t box[2] = {k.b0, k.b1};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(out, dt_of(f16), 2, (void *)k.ptr, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)k.swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled(B) failed: %d", (int)r);
  g_maps[k] = *out;
  return DCVC_OK;
}

inline int ru(int v, int m) { return (v + m - 1) / m * m; }

// N tiling: all output channels in one CTA when they fit 256 TMEM columns, else equal tiles
inline void n_tiling(int cout_w, int *tiles_n, int *n_tile) {
  const int n16 = ru(cout_w, 16);
  *tiles_n = (n16 + 255) / 256;
  // several N tiles: whole 32-column epilogue chunks per tile (a chunk never straddles tiles)
  *n_tile = *tiles_n == 1 ? n16 : ru((n16 + *tiles_n - 1) / *tiles_n, 32);
}

}  // namespace

// K chunk in {32,16,8} floats
extern "C" int dcvc_conv_tc_geometry(int cout_w, int cin, int *n_pad, int *cin_pad, int *kc) {
  if (cout_w < 1 || cin < 1) return dcvc_fail(DCVC_ERR_ARG, "conv_tc_geometry: bad arguments");
  int best = 32, best_pad = ru(cin, 32);
  for (int c : {16, 8})                   // DCVC_TC_KC=0: least padding, ties -> larger chunk
    if (ru(cin, c) < best_pad) { best = c; best_pad = ru(cin, c); }
  // default: fewest (tap, chunk) units, ties -> smaller chunk; zero padding costs MMAs but a
  // unit's fixed cost is larger (3x3 s2 53->64 at 1080p: 7 chunks of 8 -> 0.275 ms, 2 of 32 -> 0.151)
  static const int force = getenv("DCVC_TC_KC") ? atoi(getenv("DCVC_TC_KC")) : 1;
  if (force == 1) {
    int bu = 1 << 30;
    for (int c : {8, 16, 32})
      if (cdiv(cin, c) < bu) { bu = cdiv(cin, c); best = c; best_pad = ru(cin, c); }
  }
  int tiles_n, n_tile;
  n_tiling(cout_w, &tiles_n, &n_tile);
  if (n_pad) *n_pad = tiles_n * n_tile;
  if (cin_pad) *cin_pad = best_pad;
  if (kc) *kc = best;
  return DCVC_OK;
}

extern "C" int dcvc_conv_tc_geometry16(int cout_w, int cin, int *n_pad, int *cin_pad, int *kc) {
  if (cout_w < 1 || cin < 1) return dcvc_fail(DCVC_ERR_ARG, "conv_tc_geometry16: bad arguments");
  int best = 64, best_pad = ru(cin, 64);
  for (int c : {32, 16})
    if (ru(cin, c) < best_pad) { best = c; best_pad = ru(cin, c); }
  // Fewest (tap, chunk) units wins (ties -> smaller chunk): a unit's fixed cost (two TMA
  // instructions, descriptor set-up, narrow smem rows) outweighs the MMAs spent on zero padding;
  // measured 3x3 48->48 at 1080p: chunk 16 -> 0.50 ms, 32 -> 0.37 ms, 64 (zero padded) -> 0.26 ms
  static const int force = getenv("DCVC_TC_KC16") ? atoi(getenv("DCVC_TC_KC16")) : 1;
  if (force == 1) {
    int bu = 1 << 30;
    for (int c : {16, 32, 64})
      if (cdiv(cin, c) < bu) { bu = cdiv(cin, c); best = c; best_pad = ru(cin, c); }
  }
  if (force == 64 || force == 32 || force == 16) { best = force; best_pad = ru(cin, force); }
  int tiles_n, n_tile;
  n_tiling(cout_w, &tiles_n, &n_tile);
  if (n_pad) *n_pad = tiles_n * n_tile;
  if (cin_pad) *cin_pad = best_pad;
  if (kc) *kc = best;
  return DCVC_OK;
}

// Which lean epilogue (fast_chunk) a descriptor can use: 0 none (generic epilogue), 1 plain,
// 2 gated, 3 depth-to-space with weight_tc packed in (subpixel, channel) column order.
int tc_fast_mode(const dcvc_conv *d) {
  static const int no_fast = getenv("DCVC_TC_NOFAST") ? atoi(getenv("DCVC_TC_NOFAST")) : 0;
  if (no_fast) return 0;
  auto al16 = [](const dcvc_view &v) {
    return !v.data || ((((unsigned)v.cstride | (unsigned)v.coff) & 3u) == 0 &&
                       (reinterpret_cast<uintptr_t>(v.data) & 15u) == 0);
  };
  auto unit = [](float s) { return s >= 0.f && s <= 1.f; };
  auto al16h = [](const dcvc_view &v) {     // binary16 view: 8 elements per 16 bytes
    return !v.data || ((((unsigned)v.cstride | (unsigned)v.coff) & 7u) == 0 &&
                       (reinterpret_cast<uintptr_t>(v.data) & 15u) == 0);
  };
  const bool ok = d->cout_w % 4 == 0 && (d->out_f16 ? al16h(d->out) : al16(d->out)) &&
                  (d->res1_f16 ? al16h(d->res1) : al16(d->res1)) &&
                  (d->res2_f16 && d->res2.data ? al16h(d->res2) && d->cout_w % 8 == 0 : al16(d->res2)) &&
                  (d->out_f16 ? al16h(d->out2) : al16(d->out2)) &&   // out2 has out's element type
                  (reinterpret_cast<uintptr_t>(d->bias) & 15u) == 0 &&
                  (reinterpret_cast<uintptr_t>(d->scale) & 15u) == 0 && unit(d->out_slope) &&
                  unit(d->gate_slope) && unit(d->out2_slope);
  if (!ok) return 0;
  if (d->shuffle) return d->shuffle == 2 && d->cout_w % 16 == 0 && !d->gate ? 3 : 0;
  return d->gate ? 2 : 1;
}

// out2 = lrelu(out) through the TMA-store epilogue: plain convs without residuals
static bool out2_tma_ok(const dcvc_conv *d) {
  static const int on = getenv("DCVC_TC_OUT2_TMA") ? atoi(getenv("DCVC_TC_OUT2_TMA")) : 1;
  return on && d->out2.data && !d->res1.data && !d->res2.data && tc_fast_mode(d) == 1;
}

bool conv_tc_ok(const dcvc_conv *d) {
  if (d->in_slope != 1.0f) return false;                       // no input prologue on the TMA path
  if (d->stride != 1 && d->stride != 2) return false;
  const int amask = d->in_f16 ? 7 : 3;
  if ((d->in.cstride & amask) || (d->in.coff & amask)) return false;   // TMA: 16-byte strides/base
  if (reinterpret_cast<uintptr_t>(d->in.data) & 15) return false;
  if (reinterpret_cast<uintptr_t>(d->weight_tc) & 15) return false;
  if (d->in_f16 && (!d->weight_tc16 || (reinterpret_cast<uintptr_t>(d->weight_tc16) & 15)))
    return false;
  // binary16 output / residual go through the TMA-store epilogue only
  // (or, output only, through the planar depth-to-space epilogue without residuals)
  const bool shuf16 = d->out_f16 && d->shuffle == 2 && tc_fast_mode(d) == 3 && !d->res1.data &&
                      !d->res2.data && !d->out2.data;
  if ((d->out_f16 || (d->res1_f16 && d->res1.data) || (d->res2_f16 && d->res2.data)) && !shuf16 &&
      (tc_fast_mode(d) != 1 || (d->out2.data && !out2_tma_ok(d))))
    return false;
  if (d->res2_f16 && d->res2.data && !d->res1.data) return false;   // f16 res2: TMA-store epilogue only
  if (d->shuffle == 2 && tc_fast_mode(d) != 3) return false;   // planar packing needs the lean path
  const int ho = (d->in.h + 2 * d->pad - d->ksize) / d->stride + 1;
  const int wo = (d->in.w + 2 * d->pad - d->ksize) / d->stride + 1;
  if ((int64_t)ho * wo < 2 * TILE_M) return false;             // tiny layers stay on CUDA cores
  return true;
}

int conv_tc_launch(const dcvc_conv *d, int hout, int wout, cudaStream_t s) {
  TcArgs a{};
  a.out = d->out; a.res1 = d->res1; a.res2 = d->res2; a.out2 = d->out2;
  a.bias = d->bias; a.scale = d->scale;
  a.cout_w = d->cout_w; a.gate = d->gate; a.shuffle = d->shuffle; a.clamp01 = d->clamp01;
  a.out_slope = d->out_slope; a.gate_slope = d->gate_slope; a.out2_slope = d->out2_slope;
  a.hout = hout; a.wout = wout;
  a.fast = tc_fast_mode(d);
  a.ksize = d->ksize; a.stride = d->stride; a.pad = d->pad;
  int cin_pad, n_pad, tiles_n;
  const bool f16 = d->in_f16 != 0;
  a.f16_in = f16; a.f16_out = d->out_f16 != 0; a.f16_res = d->res1.data && d->res1_f16;
  a.f16_res2 = d->res2.data && d->res2_f16;
  static const int res_direct_on = getenv("DCVC_TC_RES_DIRECT") ? atoi(getenv("DCVC_TC_RES_DIRECT")) : 1;
  // measured: helps the full-resolution 3x3 convs with up to 48 filters (0.24 -> 0.21 ms), costs a
  // few % on wide 1x1 tiles whose TMA residual is already hidden; res_direct_on == 2 forces it
  a.res_direct = res_direct_on && a.f16_res && (res_direct_on == 2 || (d->ksize >= 3 && d->cout_w <= 48));
  if (f16) dcvc_conv_tc_geometry16(d->cout_w, d->in.c, &n_pad, &cin_pad, &a.kc);
  else dcvc_conv_tc_geometry(d->cout_w, d->in.c, &n_pad, &cin_pad, &a.kc);
  a.kcb = a.kc * (f16 ? 2 : 4);
  {
    const int per_mma = f16 ? 16 : 8;                 // K elements one instruction multiplies
    const int tail = d->in.c - (cin_pad / a.kc - 1) * a.kc;   // valid channels in the last chunk
    a.kk_last = cdiv(tail, per_mma);
  }
  n_tiling(d->cout_w, &tiles_n, &a.n_tile);
  a.chunks_per_tap = cin_pad / a.kc;
  a.units = a.ksize * a.ksize * a.chunks_per_tap;
  // tile shape: the (TH, TW) with TH*TW = 128 that covers the output with least padding
  int best_area = 1 << 30;
  for (int tw : {16, 8, 32, 4, 64, 128, 2}) {
    const int th = TILE_M / tw;
    if (d->stride == 2 && (tw * 2 > 256 || th * 2 > 256)) continue;
    const int area = ru(wout, tw) * ru(hout, th);
    if (area < best_area) { best_area = area; a.tw = tw; a.th = th; }
  }
  // halo mode: stride-1 k x k convs load the (16+k-1) x 16 pixel input window of a 16x8 tile
  // once per channel chunk and feed all k*k taps from it through shifted smem descriptors
  // DCVC_TC_HALO: 0 off, 1 only 3x3 with 64-byte channel chunks or wide N (the selection made
  // when the MMA issue loop was the bound), 2 (default) every stride-1 k >= 3 convolution: with
  // the uniform issue loops the taps' L2 -> smem traffic is what is left, 7x7 64->32 f16 at 1080p
  // 0.86 -> 0.68 ms, 3x3 48->48 fp32 0.39 -> 0.32 ms
  static const int halo_on = getenv("DCVC_TC_HALO") ? atoi(getenv("DCVC_TC_HALO")) : 2;
  a.halo = halo_on && d->stride == 1 && d->ksize >= 3 && d->ksize <= 9 &&
           a.kcb >= (f16 ? 32 : 64) && d->pad == d->ksize / 2;
  if (a.halo && halo_on == 1)
    a.halo = d->ksize == 3 && !d->shuffle && (a.kcb == 64 || a.n_tile >= 96);
  // resident weights: one channel chunk, one N tile and all taps' B tiles within 72 KB.  One CTA
  // per SM then issues every MMA back to back (~130 cycles each for M = 128 whatever N <= 192):
  // 3x3 48->48 f16 at 1080p 0.254 -> 0.206 ms = 27 MMAs x 110 tiles per SM, the issue floor of
  // this GEMM orientation; 7x7 only where a tap is a single MMA (16->2: 0.525 -> 0.40 ms)
  static const int bres_on = getenv("DCVC_TC_BRES") ? atoi(getenv("DCVC_TC_BRES")) : 1;
  const uint32_t bres_bytes = (uint32_t)(d->ksize * d->ksize * a.chunks_per_tap) *
                              (uint32_t)ru(a.n_tile * a.kcb, 1024);
  // one channel chunk: up to 72 KB of weights (two CTAs per SM stay possible); 3x3 with 2-3
  // chunks: up to 112 KB, one CTA per SM (96->48 f16 0.32 -> , 48->3 fp32 0.40 -> ms at 1080p)
  static const uint32_t bres_mc_kb = getenv("DCVC_TC_BRES_MC_KB") ? atoi(getenv("DCVC_TC_BRES_MC_KB")) : 112;
  const bool bres_one = a.chunks_per_tap == 1 && bres_bytes <= 72u * 1024u &&
                        (d->ksize == 3 || a.kk_last == 1) &&   // shapes mma_bres<> is instantiated for
                        (bres_on == 2 || (d->ksize == 3 && !d->shuffle) || (a.kk_last == 1 && a.n_tile <= 32));
  const bool bres_mc = a.chunks_per_tap > 1 && a.chunks_per_tap <= 3 && d->ksize == 3 && !d->shuffle &&
                       bres_bytes <= bres_mc_kb * 1024u;
  a.bres = bres_on && halo_on && d->stride == 1 && d->ksize >= 3 && d->ksize <= 7 &&
           d->pad == d->ksize / 2 && a.kcb >= (f16 ? 32 : 64) && tiles_n == 1 &&
           (!d->out2.data || out2_tma_ok(d)) &&
           (bres_one || bres_mc);
  if (a.bres) a.halo = 1;
  const int halo_ph = TILE_M / 8 + d->ksize - 1;
  static const int halo_pw16 = getenv("DCVC_TC_HALO_PW16") ? atoi(getenv("DCVC_TC_HALO_PW16")) : 0;
  if (a.halo) { a.tw = 8; a.th = TILE_M / 8; a.halo_pw = halo_pw16 ? 16 : 8 + d->ksize - 1; }
  a.tiles_x = cdiv(wout, a.tw);
  const int tiles_y = cdiv(hout, a.th);
  a.unit_a = (uint32_t)ru(TILE_M * a.kcb, 1024);
  a.unit_b = (uint32_t)ru(a.n_tile * a.kcb, 1024);
  a.unit_tx = (uint32_t)(TILE_M * a.kcb + a.n_tile * a.kcb);
  if (a.halo) {
    a.unit_tx = (uint32_t)(halo_ph * a.halo_pw * a.kcb);
    a.unit_a = (uint32_t)ru((int)a.unit_tx, 1024);
    a.halo_sbo = (uint32_t)(a.halo_pw * a.kcb);
  }
  const uint32_t unit = a.unit_a + a.unit_b;
  static const int dbg = getenv("DCVC_TC_DEBUG") ? atoi(getenv("DCVC_TC_DEBUG")) : 0;
  static const int dbg_stages = getenv("DCVC_TC_STAGES") ? atoi(getenv("DCVC_TC_STAGES")) : 0;
  static const int dbg_group = getenv("DCVC_TC_GROUP") ? atoi(getenv("DCVC_TC_GROUP")) : 0;
  static const int stage_kb = getenv("DCVC_TC_STAGE_KB") ? atoi(getenv("DCVC_TC_STAGE_KB")) : 36;
  static const int dbg_ctas = getenv("DCVC_TC_CTAS") ? atoi(getenv("DCVC_TC_CTAS")) : 3;
  static const int dbg_epi = getenv("DCVC_TC_EPI") ? atoi(getenv("DCVC_TC_EPI")) : 0;
  static const int tma_epi_on = getenv("DCVC_TC_TMA_EPI") ? atoi(getenv("DCVC_TC_TMA_EPI")) : 1;
  // TMA-store epilogue: a tile with a trailing 16-column chunk (48, 80 ... columns) gets one warp
  // quad per chunk (8 warps) even when several CTAs share the SM -- measured 64->48 1x1 + residual
  // 0.179 -> 0.143 ms, 48->48 1x1 0.091 -> 0.076 ms; tiles of whole chunks are faster with 4
  static const int epi8_odd = getenv("DCVC_TC_EPI8_ODD") ? atoi(getenv("DCVC_TC_EPI8_ODD")) : 1;
  int num_sms = 0;
  { int rc_ = dcvc_device_info(nullptr, &num_sms); if (rc_) return rc_; }
  a.debug = dbg;
  a.tiles_n = tiles_n;
  a.total_tiles = a.tiles_x * tiles_y * tiles_n;
  a.rcp_tiles_n = ((1ull << 40) + (uint64_t)tiles_n - 1) / (uint64_t)tiles_n;
  a.rcp_tiles_x = ((1ull << 40) + (uint64_t)a.tiles_x - 1) / (uint64_t)a.tiles_x;
  // two TMEM accumulators (epilogue of tile i overlaps the MMAs of tile i+1)
  // (splitting K over two partial accumulators so that consecutive MMAs are independent was
  // measured: no change on any shape -- the MMA queue is not serialised by the accumulator)
  a.acc_stride = (uint32_t)a.n_tile;
  a.tmem_cols = 32;
  while ((int)a.tmem_cols < 2 * a.n_tile) a.tmem_cols *= 2;
  // one pipeline stage = `group` units (~36 KB): amortises the barrier/commit round trip
  int group = (int)((uint32_t)stage_kb * 1024u / unit);
  if (group < 1) group = 1;
  if (dbg_group) group = dbg_group;
  if (group > a.units) group = a.units;
  a.group = group;
  const uint32_t stage = (uint32_t)group * unit;
  const int groups = cdiv(a.units, group);
  // co-resident CTAs per SM: bounded by TMEM columns, then by what the smem ring needs
  int ctas = (int)(512u / a.tmem_cols);
  if (ctas > dbg_ctas) ctas = dbg_ctas;
  if (ctas < 1) ctas = 1;
  static const int halo_ctas = getenv("DCVC_TC_HALO_CTAS") ? atoi(getenv("DCVC_TC_HALO_CTAS")) : 2;
  if (a.halo && ctas > halo_ctas) ctas = halo_ctas;
  if (a.bres) {   // weights + two windows + ~40 KB of epilogue/barrier space per CTA
    const int fit = (int)(227u * 1024u / (bres_bytes + 2u * a.unit_a + 40u * 1024u));
    ctas = fit < 1 ? 1 : (fit < ctas ? fit : ctas);
  }
  // epilogue warps: 4 per 32-column chunk group; a lone CTA on the SM gets 16 so that enough
  // loads/stores are in flight, and the register file (96 regs/thread) caps threads per SM
  const int nchunks = cdiv(a.n_tile, 32);
  // TMA-store epilogue: plain convs with at most one residual, views TMA-addressable
  a.out2_tma = out2_tma_ok(d);
  a.tma_epi = (tma_epi_on || a.f16_out || a.f16_res || a.f16_res2) && a.fast == 1 &&
              (!d->out2.data || a.out2_tma) &&
              (!d->res2.data || d->res1.data);      // a second residual rides on direct loads
  if (!a.tma_epi) a.out2_tma = 0;
  // planar depth-to-space without residuals whose sub-pixel channel count is whole 32-column
  // chunks: the same epilogue, storing through one strided tensor map per sub-pixel
  static const int shuf_tma_on = getenv("DCVC_TC_SHUF_TMA") ? atoi(getenv("DCVC_TC_SHUF_TMA")) : 1;
  a.shuf_tma = shuf_tma_on && a.fast == 3 && (d->cout_w >> 2) % 32 == 0 && !d->res1.data &&
               !d->res2.data && !d->out2.data && (d->out.w & 1) == 0 && (d->out.h & 1) == 0;
  if (a.shuf_tma) a.tma_epi = 1;
  int epi = dbg_epi == 4 || dbg_epi == 8 || dbg_epi == 16 ? dbg_epi
            : a.tma_epi ? (ctas == 1 || (epi8_odd && a.n_tile % 32 == 16 && nchunks >= 2) ? 8 : 4)
                        : (ctas == 1 ? 16 : 8);
  while (epi > 4 && epi / 4 > nchunks) epi >>= 1;
  while (ctas > 1 && ctas * (64 + 32 * epi) > 680) --ctas;
  // two epilogue sets (see the kernel): a lone CTA whose tiles need at most 8 warps, binary16
  // rows everywhere (2 KB tile buffers, so that 16 warps take the shared memory 8 took)
  static const int sets_on = getenv("DCVC_TC_EPI_SETS") ? atoi(getenv("DCVC_TC_EPI_SETS")) : 2;
  a.epi_sets = 1;
  a.epi_buf = 4096u;
  if (sets_on == 2 && a.tma_epi && ctas == 1 && epi <= 8 && a.f16_out && !a.shuf_tma &&
      (!d->res1.data || a.f16_res) && a.total_tiles >= 4 * num_sms) {
    a.epi_sets = 2;
    a.epi_buf = 2048u;
    epi *= 2;
  }
  a.epi_warps = epi;
  const uint32_t epi_bytes = (uint32_t)epi * (a.tma_epi ? 2u * a.epi_buf : STG_WARP_FLOATS * 4u) +
                             (a.tma_epi ? 8u * (uint32_t)(tiles_n * ru(a.n_tile, 32)) : 0u);
  const uint32_t fixed = 1024u /*align*/ + 1024u /*barriers*/ + epi_bytes;
  while (ctas > 1 && (227u * 1024u / ctas) - 1024u < fixed + 2u * stage) --ctas;
  const uint32_t budget = (227u * 1024u / ctas) - 1024u - fixed;
  int stages = (int)(budget / stage);
  if (stages < 2) stages = 2;
  if (stages > 6) stages = 6;
  if (dbg_stages) stages = dbg_stages;
  if (stages > groups && a.total_tiles <= ctas * num_sms) stages = groups < 2 ? 2 : groups;
  a.stages = stages;
  a.sbo = 8u * (uint32_t)a.kcb;
  a.layout_type = a.kcb == 128 ? 2u : a.kcb == 64 ? 4u : 6u;
  // smem: [stages x stage][barriers (1 KB)][epi_warps epilogue tiles][bias, scale vectors]
  a.stage_off = (uint32_t)stages * stage + 1024u;
  if (a.bres) {
    // [SA x window][all taps' weights][barriers]...
    const uint32_t room = (227u * 1024u / ctas) - 1024u - fixed;
    if (room < bres_bytes + 2u * a.unit_a)
      return dcvc_fail(DCVC_ERR_UNSUPPORTED, "conv_tc: resident weights do not fit shared memory");
    int sa = (int)((room - bres_bytes) / a.unit_a);
    a.halo_sa = sa > 6 ? 6 : sa;
    a.halo_sb = 1;
    a.halo_bg = a.ksize * a.ksize;
    a.stages = a.halo_sa + 1;
    a.stage_off = (uint32_t)a.halo_sa * a.unit_a + bres_bytes + 1024u;
  } else if (a.halo) {
    // [SA x window][SB x (bg weight taps)][barriers]...: a B stage carries a group of taps (about
    // 24 KB, at most one kernel row) so that one barrier round trip covers several MMAs
    static const int bg_kb = getenv("DCVC_TC_HALO_BKB") ? atoi(getenv("DCVC_TC_HALO_BKB")) : 24;
    int bg = (int)((uint32_t)bg_kb * 1024u / a.unit_b);
    if (bg > a.ksize) bg = a.ksize;
    if (bg < 1) bg = 1;
    a.halo_bg = bg;
    const uint32_t stage_b = (uint32_t)bg * a.unit_b;
    while (ctas > 1 && (227u * 1024u / ctas) - 1024u < fixed + 2u * a.unit_a + 2u * stage_b) --ctas;
    const uint32_t room = (227u * 1024u / ctas) - 1024u - fixed;
    a.halo_sa = room >= 3u * a.unit_a + 3u * stage_b ? 3 : 2;
    if (room < 2u * a.unit_a + 2u * stage_b)
      return dcvc_fail(DCVC_ERR_UNSUPPORTED, "conv_tc: halo window does not fit shared memory");
    int sb = (int)((room - (uint32_t)a.halo_sa * a.unit_a) / stage_b);
    static const int sb_cap = getenv("DCVC_TC_HALO_SB") ? atoi(getenv("DCVC_TC_HALO_SB")) : 8;
    a.halo_sb = sb > sb_cap ? sb_cap : sb;
    a.stages = a.halo_sa + a.halo_sb;
    a.stage_off = (uint32_t)a.halo_sa * a.unit_a + (uint32_t)a.halo_sb * stage_b + 1024u;
  }
  a.vec_off = a.stage_off + (uint32_t)epi * (a.tma_epi ? 2u * a.epi_buf : 8192u);
  const size_t smem = 1024 + (size_t)a.stage_off + epi_bytes;
  if (smem > 226 * 1024) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "conv_tc: smem %zu too large", smem);

  CUtensorMap ma, mb, mo, mr, mo2, mo3;
  int rc = a.halo ? make_map_a(d->in, a.kc, halo_ph, a.halo_pw, 1, f16, &ma)
                  : make_map_a(d->in, a.kc, a.th, a.tw, d->stride, f16, &ma);
  if (rc) return rc;
  if (a.tma_epi) {
    const int w_box = a.tw < 32 ? a.tw : 32;
    if (a.shuf_tma) {
      rc = make_map_epi(d->out, d->cout_w >> 2, w_box, a.f16_out != 0, &mo, 1);
      if (!rc) rc = make_map_epi(d->out, d->cout_w >> 2, w_box, a.f16_out != 0, &mr, 2);
      if (!rc) rc = make_map_epi(d->out, d->cout_w >> 2, w_box, a.f16_out != 0, &mo2, 3);
      if (!rc) rc = make_map_epi(d->out, d->cout_w >> 2, w_box, a.f16_out != 0, &mo3, 4);
      if (rc) return rc;
    } else {
    rc = make_map_epi(d->out, d->cout_w, w_box, a.f16_out != 0, &mo);
    if (rc) return rc;
    }
    if (a.out2_tma) {
      rc = make_map_epi(d->out2, d->cout_w, w_box, a.f16_out != 0, &mo2);
      if (rc) return rc;
    }
    if (a.shuf_tma) {
    } else if (d->res1.data) {
      rc = make_map_epi(d->res1, d->cout_w, w_box, a.f16_res != 0, &mr);
      if (rc) return rc;
    } else {
      mr = mo;
    }
  } else {
    mo = ma; mr = ma;   // unused by the kernel
  }
  rc = make_map_b(f16 ? d->weight_tc16 : (const void *)d->weight_tc, a.ksize * a.ksize * cin_pad,
                  n_pad, a.n_tile, a.kc, f16, &mb);
  if (rc) return rc;
  rc = dcvc_once_per_device(DCVC_ONCE_CONV_TC, [] {
    return cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
  });
  if (rc) return rc;
  const int grid = a.total_tiles < ctas * num_sms ? a.total_tiles : ctas * num_sms;
  if (!a.shuf_tma) { if (!a.out2_tma) mo2 = mo; mo3 = mo; }
  // programmatic dependent launch (DCVC_TC_PDL=1): measured within noise once the bias / scale
  // vector fill sits behind griddepcontrol.wait, as it must; off by default
  static const int pdl_on = getenv("DCVC_TC_PDL") ? atoi(getenv("DCVC_TC_PDL")) : 0;
  a.pdl = pdl_on;
  if (pdl_on) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)(64 + 32 * epi));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DCVC_CUDA(cudaLaunchKernelEx(&cfg, conv_tc_kernel, ma, mb, mo, mr, mo2, mo3, a));
  } else {
    conv_tc_kernel<<<grid, 64 + 32 * epi, smem, s>>>(ma, mb, mo, mr, mo2, mo3, a);
  }
  DCVC_LAUNCHED();
  return DCVC_OK;
}

#ifdef DCVC_TC_TIMELINE
extern "C" int dcvc_debug_timeline(unsigned long long *host, int bytes) {
  return (int)cudaMemcpyFromSymbol(host, g_tl, bytes);
}
#endif
