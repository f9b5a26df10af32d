// conv_direct.cu -- fp32 implicit-GEMM convolution on CUDA cores ("exact mode") with the fused
// prologue/epilogue described in include/dcvc_b200.h, plus the depthwise 3x3.
//
// Replaces nn.Conv2d + activation + residual + PixelShuffle + `* q` chains of the reference
// (DCVC-DC/src/models/layers.py:18-222, video_net.py:58-251).  This kernel takes every shape;
// conv_tc.cu takes the tensor-core eligible ones in mode 1.
//
// GEMM view: M = output pixels (row-major over (oy, ox)), N = raw output channels,
// K = (tap, cin).  CTA tile 128 x BN, K chunk 16 (never crossing a tap), 256 threads,
// thread tile TM x 4.
#include "common.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 16;
constexpr int NT = 256;

struct ConvArgs {
  dcvc_view in, out, res1, res2, out2;
  const float *weight, *bias, *scale;
  int ksize, stride, pad, cout_w, gate, shuffle, clamp01;
  float in_slope, out_slope, gate_slope, out2_slope;
  int hout, wout;  // raw conv output grid (before shuffle)
};

// store one finished value (after gate) for raw-channel-domain index `n` at raw pixel (oy, ox)
__device__ __forceinline__ void epilogue_store(const ConvArgs &a, int oy, int ox, int n, float v) {
  v = lrelu(v, a.out_slope);
  int y = oy, x = ox, c = n;
  if (a.shuffle) {
    y = oy * 2 + ((n >> 1) & 1);
    x = ox * 2 + (n & 1);
    c = n >> 2;
  }
  if (a.res1.data) v += a.res1.data[vidx(a.res1, y, x) + c];
  if (a.res2.data) v += a.res2.data[vidx(a.res2, y, x) + c];
  if (a.scale) v *= a.scale[c];
  if (a.clamp01) v = fminf(fmaxf(v, 0.f), 1.f);
  a.out.data[vidx(a.out, y, x) + c] = v;
  if (a.out2.data) a.out2.data[vidx(a.out2, y, x) + c] = lrelu(v, a.out2_slope);
}

template <int BN, int TM>
__global__ void __launch_bounds__(NT) conv_direct_kernel(const ConvArgs a) {
  constexpr int TN = 4;
  constexpr int NTN = BN / TN;        // threads along n
  static_assert(NTN * (BM / TM) == NT, "tile/thread mismatch");
  __shared__ __align__(16) float As[BK][BM];
  __shared__ __align__(16) float Bs[BK][BN];

  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int M = a.hout * a.wout;
  const int cin = a.in.c;

  // ---- A loader role: pixel lp = tid % 128, channel half lh = tid / 128 (8 channels)
  const int lp = tid & (BM - 1);
  const int lh = tid >> 7;
  const int lm = m0 + lp;
  const bool lvalid = lm < M;
  const int loy = lvalid ? lm / a.wout : 0;
  const int lox = lvalid ? lm - loy * a.wout : 0;
  const int iy0 = loy * a.stride - a.pad;
  const int ix0 = lox * a.stride - a.pad;
  const bool vec_in = ((a.in.cstride | a.in.coff) & 3) == 0 &&
                      ((reinterpret_cast<uintptr_t>(a.in.data) & 15) == 0);
  // ---- B loader role: row bk = tid / (BN/4) ... each thread loads 4 consecutive n
  constexpr int BROW_T = BN / 4;            // threads per B row
  constexpr int BROWS_PER_PASS = NT / BROW_T;
  const bool vec_w = (a.cout_w & 3) == 0 && ((reinterpret_cast<uintptr_t>(a.weight) & 15) == 0);

  // ---- compute role
  const int tn = tid % NTN;
  const int tm = tid / NTN;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int taps = a.ksize * a.ksize;
  for (int tap = 0; tap < taps; ++tap) {
    const int dy = tap / a.ksize, dx = tap - dy * a.ksize;
    const int iy = iy0 + dy, ix = ix0 + dx;
    const bool pv = lvalid && iy >= 0 && iy < a.in.h && ix >= 0 && ix < a.in.w;
    const float *ip = pv ? a.in.data + vidx(a.in, iy, ix) : a.in.data;
    for (int c0 = 0; c0 < cin; c0 += BK) {
      // A chunk: channels c0 + lh*8 .. +7 of pixel lp
      {
        float v[8];
        const int cb = c0 + lh * 8;
        if (pv && vec_in && cb + 8 <= cin) {
          const float4 p0 = *reinterpret_cast<const float4 *>(ip + cb);
          const float4 p1 = *reinterpret_cast<const float4 *>(ip + cb + 4);
          v[0] = p0.x; v[1] = p0.y; v[2] = p0.z; v[3] = p0.w;
          v[4] = p1.x; v[5] = p1.y; v[6] = p1.z; v[7] = p1.w;
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = (pv && cb + j < cin) ? ip[cb + j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) As[lh * 8 + j][lp] = lrelu(v[j], a.in_slope);
      }
      // B chunk: rows tap*cin + c0 .. +15, columns n0 .. n0+BN-1
#pragma unroll
      for (int r = tid / BROW_T; r < BK; r += BROWS_PER_PASS) {
        const int c = c0 + r;
        const int n = n0 + (tid % BROW_T) * 4;
        float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < cin) {
          const float *wp = a.weight + ((size_t)tap * cin + c) * a.cout_w + n;
          if (vec_w && n + 4 <= a.cout_w) {
            wv = *reinterpret_cast<const float4 *>(wp);
          } else {
            if (n + 0 < a.cout_w) wv.x = wp[0];
            if (n + 1 < a.cout_w) wv.y = wp[1];
            if (n + 2 < a.cout_w) wv.z = wp[2];
            if (n + 3 < a.cout_w) wv.w = wp[3];
          }
        }
        *reinterpret_cast<float4 *>(&Bs[r][(tid % BROW_T) * 4]) = wv;
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        float av[TM];
#pragma unroll
        for (int i = 0; i < TM; i += 4) {
          if constexpr (TM >= 4) {
            const float4 t = *reinterpret_cast<const float4 *>(&As[k][tm * TM + i]);
            av[i] = t.x; av[i + 1] = t.y; av[i + 2] = t.z; av[i + 3] = t.w;
          }
        }
        if constexpr (TM == 2) {
          const float2 t = *reinterpret_cast<const float2 *>(&As[k][tm * TM]);
          av[0] = t.x; av[1] = t.y;
        }
        const float4 bv = *reinterpret_cast<const float4 *>(&Bs[k][tn * TN]);
#pragma unroll
        for (int i = 0; i < TM; ++i) {
          acc[i][0] = fmaf(av[i], bv.x, acc[i][0]);
          acc[i][1] = fmaf(av[i], bv.y, acc[i][1]);
          acc[i][2] = fmaf(av[i], bv.z, acc[i][2]);
          acc[i][3] = fmaf(av[i], bv.w, acc[i][3]);
        }
      }
      __syncthreads();
    }
  }

  // ---- epilogue
  const int nb = n0 + tn * TN;
  float bias[TN];
#pragma unroll
  for (int j = 0; j < TN; ++j) bias[j] = (a.bias && nb + j < a.cout_w) ? a.bias[nb + j] : 0.f;
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + tm * TM + i;
    if (m >= M) continue;
    const int oy = m / a.wout, ox = m - oy * a.wout;
    if (a.gate) {
#pragma unroll
      for (int j = 0; j < TN; j += 2) {
        if (nb + j + 1 < a.cout_w) {
          const float x1 = acc[i][j] + bias[j], x2 = acc[i][j + 1] + bias[j + 1];
          epilogue_store(a, oy, ox, (nb + j) >> 1, x1 * lrelu(x2, a.gate_slope));
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < TN; ++j)
        if (nb + j < a.cout_w) epilogue_store(a, oy, ox, nb + j, acc[i][j] + bias[j]);
    }
  }
}

// depthwise 3x3 pad 1, scalar fallback: one thread per (pixel, up to 4 channels)
__global__ void dwconv3x3_kernel(dcvc_view in, dcvc_view out, const float *__restrict__ w,
                                 const float *__restrict__ b, int c4n) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)in.h * in.w * c4n;
  if (t >= total) return;
  const int cq = (int)(t % c4n);
  const int64_t p = t / c4n;
  const int x = (int)(p % in.w), y = (int)(p / in.w);
  const int c = cq * 4;
  const int nc = min(4, in.c - c);
  float acc[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) acc[j] = (b && j < nc) ? b[c + j] : 0.f;
#pragma unroll
  for (int dy = 0; dy < 3; ++dy) {
    const int iy = y + dy - 1;
    if (iy < 0 || iy >= in.h) continue;
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
      const int ix = x + dx - 1;
      if (ix < 0 || ix >= in.w) continue;
      const float *ip = in.data + vidx(in, iy, ix) + c;
      const float *wp = w + (dy * 3 + dx) * in.c + c;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < nc) acc[j] = fmaf(ip[j], wp[j], acc[j]);
    }
  }
  float *op = out.data + vidx(out, y, x) + c;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (j < nc) op[j] = acc[j];
}

// float4 path: one thread per (DW_TY x DW_TX output pixels, 4 channels).  The (TY+2) x (TX+2)
// input window and the nine weight vectors live in registers, so an output costs
// (TY+2)(TX+2)/(TY*TX) = 3 sixteen-byte loads instead of 18 and the kernel is HBM-bound rather
// than L1-bound.  Taps are accumulated in the same (dy, dx) order as the scalar kernel.
constexpr int DW_TX = 4, DW_TY = 2;

// four consecutive channels of an fp32 (16 bytes) or binary16 (8 bytes) view; `off` in elements
template <bool H>
__device__ __forceinline__ float4 ld4(const float *data, size_t off) {
  if (!H) return *reinterpret_cast<const float4 *>(data + off);
  const uint2 u = *reinterpret_cast<const uint2 *>(reinterpret_cast<const uint16_t *>(data) + off);
  float4 r;
  asm("{\n\t.reg .f16 lo, hi;\n\tmov.b32 {lo, hi}, %2;\n\tcvt.f32.f16 %0, lo;\n\tcvt.f32.f16 %1, hi;\n\t}"
      : "=f"(r.x), "=f"(r.y) : "r"(u.x));
  asm("{\n\t.reg .f16 lo, hi;\n\tmov.b32 {lo, hi}, %2;\n\tcvt.f32.f16 %0, lo;\n\tcvt.f32.f16 %1, hi;\n\t}"
      : "=f"(r.z), "=f"(r.w) : "r"(u.y));
  return r;
}
template <bool H>
__device__ __forceinline__ void st4(float *data, size_t off, float4 v) {
  if (!H) { *reinterpret_cast<float4 *>(data + off) = v; return; }
  uint2 u;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u.x) : "f"(v.y), "f"(v.x));
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u.y) : "f"(v.w), "f"(v.z));
  *reinterpret_cast<uint2 *>(reinterpret_cast<uint16_t *>(data) + off) = u;
}

// H: both views hold binary16 elements (dcvc_dwconv3x3_f16); arithmetic is fp32 either way.
// EDGE = false is the interior instantiation: the whole (TY+2) x (TX+2) window and all outputs
// are inside the image, so the body is straight-line loads / FMAs / stores with 32-bit offsets
// (the bounds-checked form spends ~8 instructions per FMA on predicates and is issue-bound).
template <bool H, bool EDGE>
__device__ __forceinline__ void dwconv_tile(const dcvc_view &in, const dcvc_view &out,
                                            const float4 (&wv)[9], const float4 bv, int x0, int y0,
                                            int c) {
  float4 acc[DW_TY][DW_TX];
#pragma unroll
  for (int i = 0; i < DW_TY; ++i)
#pragma unroll
    for (int j = 0; j < DW_TX; ++j) acc[i][j] = bv;
  const size_t ibase = (size_t)in.coff + c;
  // interior: element offset of window pixel (0, 0), then constant strides
  const size_t iorg = EDGE ? 0 : ibase + ((size_t)(y0 - 1) * in.w + (x0 - 1)) * in.cstride;
  const uint32_t irow = (uint32_t)in.w * (uint32_t)in.cstride, ipix = (uint32_t)in.cstride;
#pragma unroll
  for (int r = 0; r < DW_TY + 2; ++r) {
    const int iy = y0 + r - 1;
    if (EDGE && (iy < 0 || iy >= in.h)) continue;     // zero padding: the row contributes nothing
    float4 row[DW_TX + 2];
#pragma unroll
    for (int q = 0; q < DW_TX + 2; ++q) {
      if (EDGE) {
        const int ix = x0 + q - 1;
        row[q] = (ix >= 0 && ix < in.w)
                     ? ld4<H>(in.data, ibase + ((size_t)iy * in.w + ix) * in.cstride)
                     : make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        row[q] = ld4<H>(in.data, iorg + (uint32_t)r * irow + (uint32_t)q * ipix);
      }
    }
#pragma unroll
    for (int i = 0; i < DW_TY; ++i) {
      const int dy = r - i;                             // input row r feeds output row i as tap dy
      if (dy < 0 || dy > 2) continue;
#pragma unroll
      for (int j = 0; j < DW_TX; ++j)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          const float4 iv = row[j + dx], k = wv[dy * 3 + dx];
          acc[i][j].x = fmaf(iv.x, k.x, acc[i][j].x); acc[i][j].y = fmaf(iv.y, k.y, acc[i][j].y);
          acc[i][j].z = fmaf(iv.z, k.z, acc[i][j].z); acc[i][j].w = fmaf(iv.w, k.w, acc[i][j].w);
        }
    }
  }
  const size_t oorg = (size_t)out.coff + c + ((size_t)y0 * out.w + x0) * out.cstride;
  const uint32_t orow = (uint32_t)out.w * (uint32_t)out.cstride, opix = (uint32_t)out.cstride;
#pragma unroll
  for (int i = 0; i < DW_TY; ++i) {
    if (EDGE && y0 + i >= out.h) continue;
#pragma unroll
    for (int j = 0; j < DW_TX; ++j)
      if (!EDGE || x0 + j < out.w)
        st4<H>(out.data, oorg + (uint32_t)i * orow + (uint32_t)j * opix, acc[i][j]);
  }
}

template <bool H>
__global__ void __launch_bounds__(128)
dwconv3x3_v4_kernel(dcvc_view in, dcvc_view out, const float *__restrict__ w,
                    const float *__restrict__ b, uint32_t c4n, uint32_t bx_n, uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const uint32_t blk = t / c4n, cq = t - blk * c4n;
  const uint32_t by = blk / bx_n, bx = blk - by * bx_n;
  const int x0 = (int)bx * DW_TX, y0 = (int)by * DW_TY, c = (int)cq * 4;
  float4 wv[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) wv[k] = __ldg(reinterpret_cast<const float4 *>(w + k * in.c + c));
  const float4 bv = b ? __ldg(reinterpret_cast<const float4 *>(b + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
  const bool interior = x0 >= 1 && y0 >= 1 && x0 + DW_TX + 1 <= in.w && y0 + DW_TY + 1 <= in.h;
  if (interior) dwconv_tile<H, false>(in, out, wv, bv, x0, y0, c);
  else dwconv_tile<H, true>(in, out, wv, bv, x0, y0, c);
}

}  // namespace

// implemented in conv_tc.cu
bool conv_tc_ok(const dcvc_conv *d);
int conv_tc_launch(const dcvc_conv *d, int hout, int wout, cudaStream_t s);

static int conv_check(const dcvc_conv *d, int *hout, int *wout) {
  if (!d) return dcvc_fail(DCVC_ERR_ARG, "conv: null descriptor");
  if (!view_ok(&d->in) || !view_ok(&d->out)) return dcvc_fail(DCVC_ERR_ARG, "conv: bad in/out view");
  if (!d->weight) return dcvc_fail(DCVC_ERR_ARG, "conv: null weight");
  if (d->ksize < 1 || d->ksize > 7 || d->stride < 1 || d->stride > 2 || d->pad < 0)
    return dcvc_fail(DCVC_ERR_ARG, "conv: bad geometry k=%d s=%d p=%d", d->ksize, d->stride, d->pad);
  const int ho = (d->in.h + 2 * d->pad - d->ksize) / d->stride + 1;
  const int wo = (d->in.w + 2 * d->pad - d->ksize) / d->stride + 1;
  if (ho < 1 || wo < 1) return dcvc_fail(DCVC_ERR_ARG, "conv: empty output");
  if (d->gate && d->shuffle) return dcvc_fail(DCVC_ERR_ARG, "conv: gate+shuffle unsupported");
  int eh = ho, ew = wo, ec = d->cout_w;
  if (d->shuffle) {
    if (d->cout_w % 4) return dcvc_fail(DCVC_ERR_ARG, "conv: shuffle needs cout_w %% 4 == 0");
    eh *= 2; ew *= 2; ec /= 4;
  }
  if (d->gate) {
    if (d->cout_w % 2) return dcvc_fail(DCVC_ERR_ARG, "conv: gate needs even cout_w");
    ec /= 2;
  }
  if (d->out.h != eh || d->out.w != ew || d->out.c != ec)
    return dcvc_fail(DCVC_ERR_ARG, "conv: out view is %dx%dx%d, expected %dx%dx%d", d->out.h,
                     d->out.w, d->out.c, eh, ew, ec);
  const dcvc_view *opt[3] = {&d->res1, &d->res2, &d->out2};
  for (int i = 0; i < 3; ++i) {
    if (!opt[i]->data) continue;
    if (!view_ok(opt[i]) || opt[i]->h != eh || opt[i]->w != ew || opt[i]->c != ec)
      return dcvc_fail(DCVC_ERR_ARG, "conv: optional view %d has wrong extent", i);
  }
  *hout = ho;
  *wout = wo;
  return DCVC_OK;
}

static int conv_direct_launch(const dcvc_conv *d, int hout, int wout, cudaStream_t s) {
  ConvArgs a;
  a.in = d->in; a.out = d->out; a.res1 = d->res1; a.res2 = d->res2; a.out2 = d->out2;
  a.weight = d->weight; a.bias = d->bias; a.scale = d->scale;
  a.ksize = d->ksize; a.stride = d->stride; a.pad = d->pad; a.cout_w = d->cout_w;
  a.gate = d->gate; a.shuffle = d->shuffle; a.clamp01 = d->clamp01;
  a.in_slope = d->in_slope; a.out_slope = d->out_slope; a.gate_slope = d->gate_slope;
  a.out2_slope = d->out2_slope;
  a.hout = hout; a.wout = wout;
  const int M = hout * wout;
  const int gm = cdiv(M, BM);
  if (d->cout_w > 32) {
    dim3 grid(gm, cdiv(d->cout_w, 64));
    conv_direct_kernel<64, 8><<<grid, NT, 0, s>>>(a);
  } else if (d->cout_w > 16) {
    dim3 grid(gm, 1);
    conv_direct_kernel<32, 4><<<grid, NT, 0, s>>>(a);
  } else {
    dim3 grid(gm, 1);
    conv_direct_kernel<16, 2><<<grid, NT, 0, s>>>(a);
  }
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_conv_tc_eligible(const dcvc_conv *d) {
  int ho, wo;
  if (conv_check(d, &ho, &wo) != DCVC_OK) return 0;
  return (d->mode == 1 && d->weight_tc && conv_tc_ok(d)) ? 1 : 0;
}

extern "C" int dcvc_conv2d(const dcvc_conv *d, void *stream) {
  int ho, wo;
  int rc = conv_check(d, &ho, &wo);
  if (rc != DCVC_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (d->mode == 1 && d->weight_tc && conv_tc_ok(d)) return conv_tc_launch(d, ho, wo, s);
  if (d->in_f16 || d->out_f16 || (d->res1_f16 && d->res1.data) || (d->res2_f16 && d->res2.data))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED,
                     "conv: binary16 views need the tensor-core path, which this descriptor "
                     "does not qualify for (mode, shape, alignment or epilogue)");
  return conv_direct_launch(d, ho, wo, s);
}

extern "C" int dcvc_dwconv3x3(const dcvc_view *in, const dcvc_view *out, const float *weight,
                              const float *bias, void *stream) {
  if (!view_ok(in) || !view_ok(out) || !weight) return dcvc_fail(DCVC_ERR_ARG, "dwconv: bad args");
  if (in->h != out->h || in->w != out->w || in->c != out->c)
    return dcvc_fail(DCVC_ERR_ARG, "dwconv: extent mismatch");
  const int c4n = (in->c + 3) / 4;
  const int64_t total = (int64_t)in->h * in->w * c4n;
  const bool vec = (in->c & 3) == 0 && ((in->cstride | in->coff | out->cstride | out->coff) & 3) == 0 &&
                   ((reinterpret_cast<uintptr_t>(in->data) | reinterpret_cast<uintptr_t>(out->data) |
                     reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(bias)) & 15) == 0;
  const int64_t bx_n = cdiv(in->w, DW_TX), by_n = cdiv(in->h, DW_TY);
  const int64_t total_v = bx_n * by_n * c4n;
  if (vec && total_v < (1ll << 31))
    dwconv3x3_v4_kernel<false><<<(unsigned)cdiv(total_v, 128), 128, 0, (cudaStream_t)stream>>>(
        *in, *out, weight, bias, (uint32_t)c4n, (uint32_t)bx_n, (uint32_t)total_v);
  else
    dwconv3x3_kernel<<<cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(*in, *out, weight, bias, c4n);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_dwconv3x3_f16(const dcvc_view *in, const dcvc_view *out, const float *weight,
                                  const float *bias, void *stream) {
  if (!view_ok(in) || !view_ok(out) || !weight) return dcvc_fail(DCVC_ERR_ARG, "dwconv_f16: bad args");
  if (in->h != out->h || in->w != out->w || in->c != out->c)
    return dcvc_fail(DCVC_ERR_ARG, "dwconv_f16: extent mismatch");
  const int c4n = in->c / 4;
  const int64_t bx_n = cdiv(in->w, DW_TX), by_n = cdiv(in->h, DW_TY);
  const int64_t total_v = bx_n * by_n * c4n;
  // 8-byte channel groups of binary16; weights/bias are fp32 float4 groups
  const bool vec = (in->c & 3) == 0 && ((in->cstride | in->coff | out->cstride | out->coff) & 3) == 0 &&
                   ((reinterpret_cast<uintptr_t>(in->data) | reinterpret_cast<uintptr_t>(out->data)) & 7) == 0 &&
                   ((reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(bias)) & 15) == 0;
  if (!vec || total_v >= (1ll << 31))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "dwconv_f16: views must hold whole 4-channel groups");
  dwconv3x3_v4_kernel<true><<<(unsigned)cdiv(total_v, 128), 128, 0, (cudaStream_t)stream>>>(
      *in, *out, weight, bias, (uint32_t)c4n, (uint32_t)bx_n, (uint32_t)total_v);
  DCVC_LAUNCHED();
  return DCVC_OK;
}
