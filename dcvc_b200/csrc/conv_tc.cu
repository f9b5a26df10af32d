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

#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace {

constexpr int TILE_M = 128;
constexpr int NTHREADS = 192;
constexpr long long SPIN_CYCLES = 4000000000ll;  // ~2 s: a protocol bug faults instead of hanging

struct TcArgs {
  dcvc_view out, res1, res2, out2;
  const float *bias, *scale;
  int cout_w, n_pad, gate, shuffle, clamp01;
  float out_slope, gate_slope, out2_slope;
  int hout, wout, th, tw, tiles_x;
  int ksize, stride, pad, kc, chunks_per_tap, stages;
  uint32_t a_bytes, b_bytes, tx_bytes, tmem_cols, sbo, layout_type;
};

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
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
// bounded wait: a protocol bug must fault, not hang the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity))
    if (clock64() - t0 > SPIN_CYCLES) __trap();
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

// ------------------------------------------------------------------ epilogue store helpers
__device__ __forceinline__ float finish(const TcArgs &a, float v, size_t o1, size_t o2, int c,
                                        bool has1, bool has2) {
  v = lrelu(v, a.out_slope);
  if (has1) v += a.res1.data[o1 + c];
  if (has2) v += a.res2.data[o2 + c];
  if (a.scale) v *= a.scale[c];
  if (a.clamp01) v = fminf(fmaxf(v, 0.f), 1.f);
  return v;
}

// ------------------------------------------------------------------ the kernel
__global__ void __launch_bounds__(NTHREADS)
conv_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the swizzled tiles
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_bytes = a.a_bytes + a.b_bytes;
  const uint32_t bars = base + a.stages * stage_bytes;       // full[s], empty[s], acc_full
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_x = blockIdx.x % a.tiles_x, tile_y = blockIdx.x / a.tiles_x;
  const int ox0 = tile_x * a.tw, oy0 = tile_y * a.th;
  const int iters = a.ksize * a.ksize * a.chunks_per_tap;

  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (a.stages + s); };
  const uint32_t acc_bar = bars + 8u * (2 * a.stages);

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(acc_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(&tmem_base_slot)), "r"(a.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int it = 0;
      for (int tap = 0; tap < a.ksize * a.ksize; ++tap) {
        const int dy = tap / a.ksize, dx = tap - dy * a.ksize;
        const int ix = ox0 * a.stride + dx - a.pad, iy = oy0 * a.stride + dy - a.pad;
        for (int ch = 0; ch < a.chunks_per_tap; ++ch, ++it) {
          const int s = it % a.stages;
          const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
          mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t sa = base + s * stage_bytes, sb = sa + a.a_bytes;
          mbar_expect_tx(full_bar(s), a.tx_bytes);
          tma_load_4d(sa, &map_a, full_bar(s), ch * a.kc, ix, iy, 0);
          tma_load_2d(sb, &map_b, full_bar(s), (tap * a.chunks_per_tap + ch) * a.kc, 0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) |
                             ((uint32_t)(a.n_pad >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
      const int kk_n = a.kc / 8;
      for (int it = 0; it < iters; ++it) {
        const int s = it % a.stages;
        const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t sa = base + s * stage_bytes, sb = sa + a.a_bytes;
        const uint64_t da = make_smem_desc(sa, a.sbo, a.layout_type);
        const uint64_t db = make_smem_desc(sb, a.sbo, a.layout_type);
        for (int kk = 0; kk < kk_n; ++kk)      // +32 bytes along K per MMA: +2 in the >>4 field
          umma_tf32(tmem_base, da + 2u * kk, db + 2u * kk, idesc, (it | kk) ? 1u : 0u);
        umma_commit(empty_bar(s));             // frees the smem slot when these MMAs retire
      }
      umma_commit(acc_bar);                    // accumulator complete
    }
  } else {
    // ===== epilogue: warp w owns TMEM lanes 32*(w%4) .. +31 =====
    const int q = warp & 3;
    const int row = q * 32 + lane;             // pixel of the tile
    const int oy = oy0 + row / a.tw, ox = ox0 + row % a.tw;
    const bool valid = oy < a.hout && ox < a.wout;
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
    const bool has1 = a.res1.data != nullptr, has2 = a.res2.data != nullptr;
    for (int n0 = 0; n0 < a.n_pad; n0 += 16) {
      float v[16];
      tmem_ld16(trow + (uint32_t)n0, v);       // warp-collective: every lane takes part
      if (!valid || n0 >= a.cout_w) continue;
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] += (a.bias && n0 + j < a.cout_w) ? a.bias[n0 + j] : 0.f;
      if (a.shuffle) {
        // raw channel n -> pixel (2oy + ((n>>1)&1), 2ox + (n&1)), channel n>>2
#pragma unroll
        for (int sp = 0; sp < 4; ++sp) {
          const int y = 2 * oy + (sp >> 1), x = 2 * ox + (sp & 1);
          const size_t oo = vidx(a.out, y, x);
          const size_t o1 = has1 ? vidx(a.res1, y, x) : 0, o2 = has2 ? vidx(a.res2, y, x) : 0;
          const size_t oo2 = a.out2.data ? vidx(a.out2, y, x) : 0;
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) {
            const int n = n0 + 4 * cc + sp, c = n >> 2;
            if (n < a.cout_w) {
              const float r = finish(a, v[4 * cc + sp], o1, o2, c, has1, has2);
              a.out.data[oo + c] = r;
              if (a.out2.data) a.out2.data[oo2 + c] = lrelu(r, a.out2_slope);
            }
          }
        }
      } else {
        const size_t oo = vidx(a.out, oy, ox);
        const size_t o1 = has1 ? vidx(a.res1, oy, ox) : 0, o2 = has2 ? vidx(a.res2, oy, ox) : 0;
        const size_t oo2 = a.out2.data ? vidx(a.out2, oy, ox) : 0;
        if (a.gate) {
#pragma unroll
          for (int j = 0; j < 16; j += 2) {
            const int c = (n0 + j) >> 1;
            if (n0 + j + 1 < a.cout_w) {
              const float r = finish(a, v[j] * lrelu(v[j + 1], a.gate_slope), o1, o2, c, has1, has2);
              a.out.data[oo + c] = r;
              if (a.out2.data) a.out2.data[oo2 + c] = lrelu(r, a.out2_slope);
            }
          }
        } else {
          float r[16];
#pragma unroll
          for (int j = 0; j < 16; ++j)
            r[j] = (n0 + j < a.cout_w) ? finish(a, v[j], o1, o2, n0 + j, has1, has2) : 0.f;
          float *op = a.out.data + oo + n0;
          if (n0 + 16 <= a.cout_w && ((reinterpret_cast<uintptr_t>(op) & 15) == 0)) {
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              *reinterpret_cast<float4 *>(op + j) = make_float4(r[j], r[j + 1], r[j + 2], r[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (n0 + j < a.cout_w) op[j] = r[j];
          }
          if (a.out2.data) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (n0 + j < a.cout_w) a.out2.data[oo2 + n0 + j] = lrelu(r[j], a.out2_slope);
          }
        }
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

CUtensorMapSwizzle swizzle_of(int kc) {
  return kc == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : kc == 16 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                          : CU_TENSOR_MAP_SWIZZLE_32B;
}

// activation map: dims (C, W, H, 1), box (KC, TW', TH', 1), traversal strides (1, s, s, 1)
int make_map_a(const dcvc_view &in, int kc, int th, int tw, int stride, CUtensorMap *out) {
  MapKey k{};
  k.ptr = (uint64_t)(in.data + in.coff);
  k.d0 = in.c; k.d1 = in.w; k.d2 = in.h;
  k.s1 = (uint64_t)in.cstride * 4; k.s2 = (uint64_t)in.cstride * 4 * in.w;
  k.b0 = kc; k.b1 = stride == 1 ? tw : tw * 2 - 1; k.b2 = stride == 1 ? th : th * 2 - 1;
  k.e1 = stride; k.e2 = stride; k.swz = swizzle_of(kc); k.rank = 4;
  std::lock_guard<std::mutex> g(g_maps_mu);
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return DCVC_OK; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[4] = {k.d0, k.d1, k.d2, 1};
  cuuint64_t strides[3] = {k.s1, k.s2, k.s2 * in.h};
  cuuint32_t box[4] = {k.b0, k.b1, k.b2, 1};
  cuuint32_t es[4] = {1, k.e1, k.e2, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void *)k.ptr, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)k.swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled(A) failed: %d", (int)r);
  if (g_maps.size() > 8192) g_maps.clear();
  g_maps[k] = *out;
  return DCVC_OK;
}

// weight map: dims (Ktot, N_pad), box (KC, N_pad)
int make_map_b(const float *w, int ktot, int n_pad, int kc, CUtensorMap *out) {
  MapKey k{};
  k.ptr = (uint64_t)w; k.d0 = ktot; k.d1 = n_pad; k.s1 = (uint64_t)ktot * 4;
  k.b0 = kc; k.b1 = n_pad; k.swz = swizzle_of(kc); k.rank = 2;
  std::lock_guard<std::mutex> g(g_maps_mu);
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return DCVC_OK; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[2] = {k.d0, k.d1};
  cuuint64_t strides[1] = {k.s1};
  cuuint32_t box[2] = {k.b0, k.b1};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)k.ptr, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)k.swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return dcvc_fail(DCVC_ERR_CUDA, "cuTensorMapEncodeTiled(B) failed: %d", (int)r);
  g_maps[k] = *out;
  return DCVC_OK;
}

inline int ru(int v, int m) { return (v + m - 1) / m * m; }

}  // namespace

// K chunk: the candidate in {32,16,8} that pads Cin least (ties -> larger chunk)
extern "C" int dcvc_conv_tc_geometry(int cout_w, int cin, int *n_pad, int *cin_pad, int *kc) {
  if (cout_w < 1 || cin < 1) return dcvc_fail(DCVC_ERR_ARG, "conv_tc_geometry: bad arguments");
  int best = 32, best_pad = ru(cin, 32);
  for (int c : {16, 8})
    if (ru(cin, c) < best_pad) { best = c; best_pad = ru(cin, c); }
  if (n_pad) *n_pad = ru(cout_w, 16);
  if (cin_pad) *cin_pad = best_pad;
  if (kc) *kc = best;
  return DCVC_OK;
}

bool conv_tc_ok(const dcvc_conv *d) {
  if (d->in_slope != 1.0f) return false;                       // no input prologue on the TMA path
  if (d->stride != 1 && d->stride != 2) return false;
  if (d->cout_w > 256) return false;
  if ((d->in.cstride & 3) || (d->in.coff & 3)) return false;   // TMA: 16-byte strides/base
  if (reinterpret_cast<uintptr_t>(d->in.data) & 15) return false;
  if (reinterpret_cast<uintptr_t>(d->weight_tc) & 15) return false;
  const int ho = (d->in.h + 2 * d->pad - d->ksize) / d->stride + 1;
  const int wo = (d->in.w + 2 * d->pad - d->ksize) / d->stride + 1;
  if ((int64_t)ho * wo < 4 * TILE_M) return false;             // tiny layers stay on CUDA cores
  return true;
}

int conv_tc_launch(const dcvc_conv *d, int hout, int wout, cudaStream_t s) {
  TcArgs a{};
  a.out = d->out; a.res1 = d->res1; a.res2 = d->res2; a.out2 = d->out2;
  a.bias = d->bias; a.scale = d->scale;
  a.cout_w = d->cout_w; a.gate = d->gate; a.shuffle = d->shuffle; a.clamp01 = d->clamp01;
  a.out_slope = d->out_slope; a.gate_slope = d->gate_slope; a.out2_slope = d->out2_slope;
  a.hout = hout; a.wout = wout;
  a.ksize = d->ksize; a.stride = d->stride; a.pad = d->pad;
  int cin_pad;
  dcvc_conv_tc_geometry(d->cout_w, d->in.c, &a.n_pad, &cin_pad, &a.kc);
  a.chunks_per_tap = cin_pad / a.kc;
  // tile shape: the (TH, TW) with TH*TW = 128 that covers the output with least padding
  int best_area = 1 << 30;
  for (int tw : {16, 8, 32, 4, 64, 128, 2}) {
    const int th = TILE_M / tw;
    if (d->stride == 2 && (tw * 2 > 256 || th * 2 > 256)) continue;
    const int area = ru(wout, tw) * ru(hout, th);
    if (area < best_area) { best_area = area; a.tw = tw; a.th = th; }
  }
  a.tiles_x = cdiv(wout, a.tw);
  const int tiles_y = cdiv(hout, a.th);
  a.a_bytes = (uint32_t)ru(TILE_M * a.kc * 4, 1024);
  a.b_bytes = (uint32_t)ru(a.n_pad * a.kc * 4, 1024);
  a.tx_bytes = (uint32_t)(TILE_M * a.kc * 4 + a.n_pad * a.kc * 4);
  const uint32_t stage = a.a_bytes + a.b_bytes;
  int stages = (int)(98304u / stage);
  if (stages < 4) stages = 4;
  if (stages > 8) stages = 8;
  const int iters = a.ksize * a.ksize * a.chunks_per_tap;
  if (stages > iters) stages = iters < 2 ? 2 : iters;
  a.stages = stages;
  a.tmem_cols = 32;
  while ((int)a.tmem_cols < a.n_pad) a.tmem_cols *= 2;
  a.sbo = 8u * a.kc * 4u;
  a.layout_type = a.kc == 32 ? 2u : a.kc == 16 ? 4u : 6u;
  const size_t smem = 1024 + (size_t)stages * stage + 8 * (2 * stages + 1) + 16;
  if (smem > 226 * 1024) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "conv_tc: smem %zu too large", smem);

  CUtensorMap ma, mb;
  int rc = make_map_a(d->in, a.kc, a.th, a.tw, d->stride, &ma);
  if (rc) return rc;
  rc = make_map_b(d->weight_tc, a.ksize * a.ksize * cin_pad, a.n_pad, a.kc, &mb);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    DCVC_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   226 * 1024));
    attr_set = true;
  }
  conv_tc_kernel<<<a.tiles_x * tiles_y, NTHREADS, smem, s>>>(ma, mb, a);
  DCVC_LAUNCHED();
  return DCVC_OK;
}
