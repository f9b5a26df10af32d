// warp.cu -- motion / resampling kernels (HBM-bound stencil and gather kernels, NHWC fp32).
//
// Replaces, on the hot path of the reference (DCVC-DC/src/models):
//   flow_warp / torch_warp (F.grid_sample bilinear, border, align_corners=True)  video_net.py:8-38
//   bilinearupsacling / bilineardownsacling (F.interpolate, align_corners=False) video_net.py:41-55
//   F.avg_pool2d(2,2) pyramid                                                   video_net.py:111-113
//   nn.MaxPool2d(2,2) in UNet                                                   video_net.py:154,157
//   OffsetDiversity from the offset tensor onward                               video_model.py:46-63
//   torch.cat / F.pad(replicate) / negative F.pad                               common_model.py:70-86
// Every kernel maps consecutive threads to consecutive channels (then pixels), i.e. coalesced
// NHWC access; all are one read + one write of their tensors.
#include <cuda_fp16.h>

#include "common.cuh"

namespace {

constexpr int TPB = 256;

__global__ void nchw_to_nhwc_kernel(const float *__restrict__ src, dcvc_view dst) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t hw = (int64_t)dst.h * dst.w;
  if (t >= hw * dst.c) return;
  const int c = (int)(t / hw);
  const int64_t p = t - (int64_t)c * hw;  // coalesced read of a plane
  dst.data[p * dst.cstride + dst.coff + c] = src[t];
}

__global__ void nhwc_to_nchw_kernel(dcvc_view src, float *__restrict__ dst, int clamp01) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t hw = (int64_t)src.h * src.w;
  if (t >= hw * src.c) return;
  const int c = (int)(t / hw);
  const int64_t p = t - (int64_t)c * hw;
  float v = src.data[p * src.cstride + src.coff + c];
  if (clamp01) v = fminf(fmaxf(v, 0.f), 1.f);
  dst[t] = v;
}

__global__ void copy_kernel(dcvc_view src, dcvc_view dst, float mul, float slope) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)dst.h * dst.w * dst.c) return;
  const int c = (int)(t % dst.c);
  const int64_t p = t / dst.c;
  const int x = (int)(p % dst.w), y = (int)(p / dst.w);
  const int sy = min(y, src.h - 1), sx = min(x, src.w - 1);
  dst.data[vidx(dst, y, x) + c] = mul * lrelu(src.data[vidx(src, sy, sx) + c], slope);
}

// float4 path of copy: one thread per (pixel, 4 channels), 32-bit index math.  SAME: source and
// destination have equal sizes (no replicate pad / crop), so the pixel index is shared.
template <bool SAME>
__global__ void copy_v4_kernel(dcvc_view src, dcvc_view dst, float mul, float slope, uint32_t c4n,
                               uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const uint32_t p = t / c4n, cq = t - p * c4n;
  size_t so;
  if (SAME) {
    so = (size_t)p * src.cstride;
  } else {
    const uint32_t y = p / (uint32_t)dst.w, x = p - y * (uint32_t)dst.w;
    so = ((size_t)min((int)y, src.h - 1) * src.w + min((int)x, src.w - 1)) * src.cstride;
  }
  float4 v = *reinterpret_cast<const float4 *>(src.data + so + src.coff + 4 * cq);
  v.x = mul * lrelu(v.x, slope); v.y = mul * lrelu(v.y, slope);
  v.z = mul * lrelu(v.z, slope); v.w = mul * lrelu(v.w, slope);
  *reinterpret_cast<float4 *>(dst.data + (size_t)p * dst.cstride + dst.coff + 4 * cq) = v;
}

// fp32 (or binary16, HS) -> binary16 converting copy: one thread per (pixel, 8 channels): two
// float4 loads (one 16-byte load of eight binary16), one 16-byte store of four half2 (round to
// nearest even)
template <bool HS>
__global__ void copy_f16_kernel(dcvc_view src, dcvc_view dst, float mul, float slope, uint32_t c8n,
                                uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const uint32_t p = t / c8n, cq = t - p * c8n;
  float4 u, v;
  if (HS) {
    const uint4 h = *reinterpret_cast<const uint4 *>(
        reinterpret_cast<const uint16_t *>(src.data) + (size_t)p * src.cstride + src.coff + 8 * cq);
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&h.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&h.y));
    const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&h.z));
    const float2 d = __half22float2(*reinterpret_cast<const __half2 *>(&h.w));
    u = make_float4(a.x, a.y, b.x, b.y);
    v = make_float4(c.x, c.y, d.x, d.y);
  } else {
    const float4 *sp = reinterpret_cast<const float4 *>(src.data + (size_t)p * src.cstride + src.coff + 8 * cq);
    u = sp[0]; v = sp[1];
  }
  auto f = [&](float x) { return mul * lrelu(x, slope); };
  auto pk = [](float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
  };
  uint4 o;
  o.x = pk(f(u.x), f(u.y)); o.y = pk(f(u.z), f(u.w));
  o.z = pk(f(v.x), f(v.y)); o.w = pk(f(v.z), f(v.w));
  uint16_t *dp = reinterpret_cast<uint16_t *>(dst.data) + (size_t)p * dst.cstride + dst.coff + 8 * cq;
  *reinterpret_cast<uint4 *>(dp) = o;
}

// fewer than eight fp32 channels into one 16-byte binary16 group, zero-padded (the image next to
// a binary16 context in a concatenation buffer)
__global__ void copy_f16_narrow_kernel(dcvc_view src, dcvc_view dst, uint32_t total) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= total) return;
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j)
    v[j] = j < src.c ? src.data[(size_t)p * src.cstride + src.coff + j] : 0.f;
  auto pk = [](float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
  };
  *reinterpret_cast<uint4 *>(reinterpret_cast<uint16_t *>(dst.data) + (size_t)p * dst.cstride + dst.coff) =
      make_uint4(pk(v[0], v[1]), pk(v[2], v[3]), pk(v[4], v[5]), pk(v[6], v[7]));
}

__global__ void fill_kernel(dcvc_view dst, float value) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)dst.h * dst.w * dst.c) return;
  const int c = (int)(t % dst.c);
  const int64_t p = t / dst.c;
  dst.data[p * dst.cstride + dst.coff + c] = value;
}

__global__ void pool2_kernel(dcvc_view src, dcvc_view dst, int mode) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)dst.h * dst.w * dst.c) return;
  const int c = (int)(t % dst.c);
  const int64_t p = t / dst.c;
  const int x = (int)(p % dst.w), y = (int)(p / dst.w);
  const float a = src.data[vidx(src, 2 * y, 2 * x) + c];
  const float b = src.data[vidx(src, 2 * y, 2 * x + 1) + c];
  const float d = src.data[vidx(src, 2 * y + 1, 2 * x) + c];
  const float e = src.data[vidx(src, 2 * y + 1, 2 * x + 1) + c];
  float v;
  if (mode == 0) v = (a + b + d + e) * 0.25f;   // ATen avg_pool2d: sum / pool_size
  else v = fmaxf(fmaxf(a, b), fmaxf(d, e));
  dst.data[vidx(dst, y, x) + c] = v;
}

// float4 path of pool2: one thread per (output pixel, 4 channels)
__global__ void pool2_v4_kernel(dcvc_view src, dcvc_view dst, int mode, uint32_t c4n, uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const uint32_t p = t / c4n, cq = t - p * c4n;
  const uint32_t y = p / (uint32_t)dst.w, x = p - y * (uint32_t)dst.w;
  const float *sp = src.data + ((size_t)(2 * y) * src.w + 2 * x) * src.cstride + src.coff + 4 * cq;
  const size_t row = (size_t)src.w * src.cstride;
  const float4 a = *reinterpret_cast<const float4 *>(sp);
  const float4 b = *reinterpret_cast<const float4 *>(sp + src.cstride);
  const float4 d = *reinterpret_cast<const float4 *>(sp + row);
  const float4 e = *reinterpret_cast<const float4 *>(sp + row + src.cstride);
  float4 v;
  if (mode == 0) {
    v = make_float4((a.x + b.x + d.x + e.x) * 0.25f, (a.y + b.y + d.y + e.y) * 0.25f,
                    (a.z + b.z + d.z + e.z) * 0.25f, (a.w + b.w + d.w + e.w) * 0.25f);
  } else {
    v = make_float4(fmaxf(fmaxf(a.x, b.x), fmaxf(d.x, e.x)), fmaxf(fmaxf(a.y, b.y), fmaxf(d.y, e.y)),
                    fmaxf(fmaxf(a.z, b.z), fmaxf(d.z, e.z)), fmaxf(fmaxf(a.w, b.w), fmaxf(d.w, e.w)));
  }
  *reinterpret_cast<float4 *>(dst.data + (size_t)p * dst.cstride + dst.coff + 4 * cq) = v;
}

// binary16 views on both sides: one thread per (output pixel, 8 channels); max is exact, the
// average is summed in fp32 and rounded once
__global__ void pool2_f16_kernel(dcvc_view src, dcvc_view dst, int mode, uint32_t c8n, uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const uint32_t p = t / c8n, cq = t - p * c8n;
  const uint32_t y = p / (uint32_t)dst.w, x = p - y * (uint32_t)dst.w;
  const uint16_t *sp = reinterpret_cast<const uint16_t *>(src.data) +
                       ((size_t)(2 * y) * src.w + 2 * x) * src.cstride + src.coff + 8 * cq;
  const size_t row = (size_t)src.w * src.cstride;
  const uint4 q[4] = {*reinterpret_cast<const uint4 *>(sp), *reinterpret_cast<const uint4 *>(sp + src.cstride),
                      *reinterpret_cast<const uint4 *>(sp + row),
                      *reinterpret_cast<const uint4 *>(sp + row + src.cstride)};
  uint32_t o[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float2 f[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      f[k] = __half22float2(*reinterpret_cast<const __half2 *>(reinterpret_cast<const uint32_t *>(&q[k]) + j));
    float2 v;
    if (mode == 0) {
      v = make_float2((f[0].x + f[1].x + f[2].x + f[3].x) * 0.25f, (f[0].y + f[1].y + f[2].y + f[3].y) * 0.25f);
    } else {
      v = make_float2(fmaxf(fmaxf(f[0].x, f[1].x), fmaxf(f[2].x, f[3].x)),
                      fmaxf(fmaxf(f[0].y, f[1].y), fmaxf(f[2].y, f[3].y)));
    }
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(o[j]) : "f"(v.y), "f"(v.x));
  }
  *reinterpret_cast<uint4 *>(reinterpret_cast<uint16_t *>(dst.data) + (size_t)p * dst.cstride + dst.coff + 8 * cq) =
      make_uint4(o[0], o[1], o[2], o[3]);
}

// ATen upsample_bilinear2d, align_corners=False: src = scale*(dst+0.5)-0.5 clamped at 0
__device__ __forceinline__ void src_index(float scale, int dst, int in_size, int &i0, int &i1,
                                          float &l0, float &l1) {
  float s = scale * (dst + 0.5f) - 0.5f;
  if (s < 0.f) s = 0.f;
  i0 = min((int)s, in_size - 1);
  i1 = min(i0 + 1, in_size - 1);
  l1 = s - (float)i0;
  l0 = 1.f - l1;
}

__device__ __forceinline__ float bilerp_resize(const dcvc_view &src, int c, int y0, int y1,
                                               int x0, int x1, float hl0, float hl1, float wl0,
                                               float wl1) {
  const float v00 = src.data[vidx(src, y0, x0) + c], v01 = src.data[vidx(src, y0, x1) + c];
  const float v10 = src.data[vidx(src, y1, x0) + c], v11 = src.data[vidx(src, y1, x1) + c];
  return hl0 * (wl0 * v00 + wl1 * v01) + hl1 * (wl0 * v10 + wl1 * v11);
}

__global__ void resize_bilinear_kernel(dcvc_view src, dcvc_view dst, float sh, float sw, float mul) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)dst.h * dst.w * dst.c) return;
  const int c = (int)(t % dst.c);
  const int64_t p = t / dst.c;
  const int x = (int)(p % dst.w), y = (int)(p / dst.w);
  int y0, y1, x0, x1;
  float hl0, hl1, wl0, wl1;
  src_index(sh, y, src.h, y0, y1, hl0, hl1);
  src_index(sw, x, src.w, x0, x1, wl0, wl1);
  dst.data[vidx(dst, y, x) + c] = mul * bilerp_resize(src, c, y0, y1, x0, x1, hl0, hl1, wl0, wl1);
}

// torch.linspace(-1, 1, n)[i] in fp32 (symmetric evaluation, ATen RangeFactories)
__device__ __forceinline__ float linspace_pm1(int i, int n) {
  const float step = 2.0f / (float)(n - 1);
  return i < n / 2 ? -1.0f + step * (float)i : 1.0f - step * (float)(n - 1 - i);
}

struct Bilin {
  int x0, y0, x1, y1;
  float nw, ne, sw, se;
  bool vx1, vy1;
};

// reference arithmetic: grid = linspace + flow/((size-1)/2); unnormalise ((g+1)*((size-1)/2));
// clip to [0, size-1]; bilinear with out-of-range corners dropped.
__device__ __forceinline__ Bilin warp_coords(int x, int y, float fx, float fy, int w, int h) {
  const float half_w = (float)(w - 1) / 2.0f, half_h = (float)(h - 1) / 2.0f;
  float gx = linspace_pm1(x, w) + fx / half_w;
  float gy = linspace_pm1(y, h) + fy / half_h;
  if (w == 1) gx = -1.0f + fx / half_w;
  if (h == 1) gy = -1.0f + fy / half_h;
  float sx = (gx + 1.0f) * half_w;
  float sy = (gy + 1.0f) * half_h;
  sx = fminf((float)(w - 1), fmaxf(sx, 0.f));
  sy = fminf((float)(h - 1), fmaxf(sy, 0.f));
  Bilin b;
  const float xf = floorf(sx), yf = floorf(sy);
  b.x0 = (int)xf; b.y0 = (int)yf;
  b.x1 = b.x0 + 1; b.y1 = b.y0 + 1;
  const float e = sx - xf, wv = 1.0f - e, s = sy - yf, n = 1.0f - s;
  b.nw = n * wv; b.ne = n * e; b.sw = s * wv; b.se = s * e;
  b.vx1 = b.x1 < w; b.vy1 = b.y1 < h;
  if (!b.vx1) b.x1 = b.x0;
  if (!b.vy1) b.y1 = b.y0;
  return b;
}

__device__ __forceinline__ float warp_sample(const dcvc_view &f, const Bilin &b, int c) {
  float v = b.nw * f.data[vidx(f, b.y0, b.x0) + c];
  if (b.vx1) v += b.ne * f.data[vidx(f, b.y0, b.x1) + c];
  if (b.vy1) v += b.sw * f.data[vidx(f, b.y1, b.x0) + c];
  if (b.vx1 && b.vy1) v += b.se * f.data[vidx(f, b.y1, b.x1) + c];
  return v;
}

__global__ void flow_warp_kernel(dcvc_view feat, dcvc_view flow, dcvc_view out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)out.h * out.w * out.c) return;
  const int c = (int)(t % out.c);
  const int64_t p = t / out.c;
  const int x = (int)(p % out.w), y = (int)(p / out.w);
  const float *fl = flow.data + vidx(flow, y, x);
  const Bilin b = warp_coords(x, y, fl[0], fl[1], feat.w, feat.h);
  out.data[vidx(out, y, x) + c] = warp_sample(feat, b, c);
}

// float4 path: one thread per (pixel, 4 channels); the four bilinear taps are 16-byte gathers
__device__ __forceinline__ float4 ld4h(const float *data, size_t off_elems) {   // 4 binary16 -> fp32
  const uint2 u = *reinterpret_cast<const uint2 *>(reinterpret_cast<const uint16_t *>(data) + off_elems);
  const float2 lo = __half22float2(*reinterpret_cast<const __half2 *>(&u.x));
  const float2 hi = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
  return make_float4(lo.x, lo.y, hi.x, hi.y);
}

template <bool HO, bool HS>   // HO / HS: out / feat hold binary16 elements
__global__ void flow_warp_v4_kernel(dcvc_view feat, dcvc_view flow, dcvc_view out, uint32_t c4n,
                                    uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const uint32_t p = t / c4n, cq = t - p * c4n;
  const uint32_t y = p / (uint32_t)out.w, x = p - y * (uint32_t)out.w;
  const float2 fl = *reinterpret_cast<const float2 *>(flow.data + (size_t)p * flow.cstride + flow.coff);
  const Bilin b = warp_coords((int)x, (int)y, fl.x, fl.y, feat.w, feat.h);
  const size_t cbase = (size_t)feat.coff + 4 * cq;
  const size_t r0 = (size_t)b.y0 * feat.w, r1 = (size_t)b.y1 * feat.w;
  auto tap = [&](size_t pix) {
    const size_t off = cbase + pix * feat.cstride;
    return HS ? ld4h(feat.data, off) : *reinterpret_cast<const float4 *>(feat.data + off);
  };
  // same accumulation order as warp_sample (dropped corners contribute nothing)
  float4 a = tap(r0 + b.x0);
  float4 v = make_float4(b.nw * a.x, b.nw * a.y, b.nw * a.z, b.nw * a.w);
  if (b.vx1) {
    a = tap(r0 + b.x1);
    v.x += b.ne * a.x; v.y += b.ne * a.y; v.z += b.ne * a.z; v.w += b.ne * a.w;
  }
  if (b.vy1) {
    a = tap(r1 + b.x0);
    v.x += b.sw * a.x; v.y += b.sw * a.y; v.z += b.sw * a.z; v.w += b.sw * a.w;
  }
  if (b.vx1 && b.vy1) {
    a = tap(r1 + b.x1);
    v.x += b.se * a.x; v.y += b.se * a.y; v.z += b.se * a.z; v.w += b.se * a.w;
  }
  if (HO) {
    const __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<const uint32_t *>(&lo);
    u.y = *reinterpret_cast<const uint32_t *>(&hi);
    *reinterpret_cast<uint2 *>(reinterpret_cast<__half *>(out.data) + (size_t)p * out.cstride +
                               out.coff + 4 * cq) = u;
  } else {
    *reinterpret_cast<float4 *>(out.data + (size_t)p * out.cstride + out.coff + 4 * cq) = v;
  }
}

// binary16 -> binary16, one thread per (pixel, 8 channels): 16-byte gathers; same arithmetic and
// accumulation order per channel as flow_warp_v4_kernel
__global__ void flow_warp_v8h_kernel(dcvc_view feat, dcvc_view flow, dcvc_view out, uint32_t c8n,
                                     uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const uint32_t p = t / c8n, cq = t - p * c8n;
  const uint32_t y = p / (uint32_t)out.w, x = p - y * (uint32_t)out.w;
  const float2 fl = *reinterpret_cast<const float2 *>(flow.data + (size_t)p * flow.cstride + flow.coff);
  const Bilin b = warp_coords((int)x, (int)y, fl.x, fl.y, feat.w, feat.h);
  const uint16_t *base = reinterpret_cast<const uint16_t *>(feat.data) + feat.coff + 8 * cq;
  const size_t r0 = (size_t)b.y0 * feat.w, r1 = (size_t)b.y1 * feat.w;
  float v[8];
  auto tap = [&](size_t pix, float wgt, bool first) {
    const uint4 u = *reinterpret_cast<const uint4 *>(base + pix * feat.cstride);
    const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&w4[j]));
      if (first) { v[2 * j] = wgt * f.x; v[2 * j + 1] = wgt * f.y; }
      else { v[2 * j] += wgt * f.x; v[2 * j + 1] += wgt * f.y; }
    }
  };
  tap(r0 + b.x0, b.nw, true);
  if (b.vx1) tap(r0 + b.x1, b.ne, false);
  if (b.vy1) tap(r1 + b.x0, b.sw, false);
  if (b.vx1 && b.vy1) tap(r1 + b.x1, b.se, false);
  uint4 o;
  uint32_t *ow = reinterpret_cast<uint32_t *>(&o);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const __half2 h = __floats2half2_rn(v[2 * j], v[2 * j + 1]);
    ow[j] = *reinterpret_cast<const uint32_t *>(&h);
  }
  *reinterpret_cast<uint4 *>(reinterpret_cast<uint16_t *>(out.data) + (size_t)p * out.cstride + out.coff + 8 * cq) = o;
}

// one thread per (pixel, group q of 16): samples n = 2q, 2q+1
__global__ void offset_diversity_kernel(dcvc_view x, dcvc_view offs, dcvc_view flow,
                                        const float *__restrict__ fw, const float *__restrict__ fb,
                                        dcvc_view out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)out.h * out.w * 16) return;
  const int q = (int)(t & 15);
  const int64_t p = t >> 4;
  const int px = (int)(p % out.w), py = (int)(p / out.w);
  // bilinear x2 source coordinates in the half-resolution offset tensor
  int y0, y1, x0, x1;
  float hl0, hl1, wl0, wl1;
  src_index((float)offs.h / (float)out.h, py, offs.h, y0, y1, hl0, hl1);
  src_index((float)offs.w / (float)out.w, px, offs.w, x0, x1, wl0, wl1);
  const float *fl = flow.data + vidx(flow, py, px);
  const float fx = fl[0], fy = fl[1];
  float v[6];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int n = 2 * q + k;
    const float ox = bilerp_resize(offs, 2 * n, y0, y1, x0, x1, hl0, hl1, wl0, wl1);
    const float oy = bilerp_resize(offs, 2 * n + 1, y0, y1, x0, x1, hl0, hl1, wl0, wl1);
    const float mk = bilerp_resize(offs, 64 + n, y0, y1, x0, x1, hl0, hl1, wl0, wl1);
    const float mask = 1.0f / (1.0f + expf(-mk));
    const Bilin b = warp_coords(px, py, 40.0f * tanhf(ox) + fx, 40.0f * tanhf(oy) + fy, x.w, x.h);
    const int g = n & 15;
#pragma unroll
    for (int j = 0; j < 3; ++j) v[k * 3 + j] = warp_sample(x, b, g * 3 + j) * mask;
  }
  float *op = out.data + vidx(out, py, px) + 3 * q;
#pragma unroll
  for (int o = 0; o < 3; ++o) {
    const float *wr = fw + (3 * q + o) * 6;
    float acc = fb ? fb[3 * q + o] : 0.f;
#pragma unroll
    for (int r = 0; r < 6; ++r) acc = fmaf(v[r], wr[r], acc);
    op[o] = acc;
  }
}

// the same with the offset / mask channels of a sample pair fetched as one float4 / float2 per
// corner of the x2 resize (offs view 16-byte aligned), corner addresses computed once, 32-bit
// indices.  Every value is produced by the same expression as above: results are bit-identical.
template <bool HO, bool X2, bool HS = false>   // HO: binary16 out; X2: 8-byte loads from x; HS: binary16 x
__global__ void __launch_bounds__(256)
offset_diversity_v4_kernel(dcvc_view x, dcvc_view offs, dcvc_view flow,
                           const float *__restrict__ fw, const float *__restrict__ fb,
                           dcvc_view out, uint32_t total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int q = (int)(t & 15u);
  const uint32_t p = t >> 4;
  const int py = (int)(p / (uint32_t)out.w), px = (int)(p - (uint32_t)py * (uint32_t)out.w);
  int y0, y1, x0, x1;
  float hl0, hl1, wl0, wl1;
  src_index((float)offs.h / (float)out.h, py, offs.h, y0, y1, hl0, hl1);
  src_index((float)offs.w / (float)out.w, px, offs.w, x0, x1, wl0, wl1);
  const float *o00 = offs.data + vidx(offs, y0, x0), *o01 = offs.data + vidx(offs, y0, x1);
  const float *o10 = offs.data + vidx(offs, y1, x0), *o11 = offs.data + vidx(offs, y1, x1);
  const float4 a00 = __ldg(reinterpret_cast<const float4 *>(o00 + 4 * q));
  const float4 a01 = __ldg(reinterpret_cast<const float4 *>(o01 + 4 * q));
  const float4 a10 = __ldg(reinterpret_cast<const float4 *>(o10 + 4 * q));
  const float4 a11 = __ldg(reinterpret_cast<const float4 *>(o11 + 4 * q));
  const float2 m00 = __ldg(reinterpret_cast<const float2 *>(o00 + 64 + 2 * q));
  const float2 m01 = __ldg(reinterpret_cast<const float2 *>(o01 + 64 + 2 * q));
  const float2 m10 = __ldg(reinterpret_cast<const float2 *>(o10 + 64 + 2 * q));
  const float2 m11 = __ldg(reinterpret_cast<const float2 *>(o11 + 64 + 2 * q));
  const float2 fl = __ldg(reinterpret_cast<const float2 *>(flow.data + vidx(flow, py, px)));
  auto bil = [&](float v00, float v01, float v10, float v11) {
    return hl0 * (wl0 * v00 + wl1 * v01) + hl1 * (wl0 * v10 + wl1 * v11);
  };
  const float oxs[2] = {bil(a00.x, a01.x, a10.x, a11.x), bil(a00.z, a01.z, a10.z, a11.z)};
  const float oys[2] = {bil(a00.y, a01.y, a10.y, a11.y), bil(a00.w, a01.w, a10.w, a11.w)};
  const float mks[2] = {bil(m00.x, m01.x, m10.x, m11.x), bil(m00.y, m01.y, m10.y, m11.y)};
  float v[6];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int n = 2 * q + k;
    const float mask = 1.0f / (1.0f + expf(-mks[k]));
    const Bilin b = warp_coords(px, py, 40.0f * tanhf(oxs[k]) + fl.x, 40.0f * tanhf(oys[k]) + fl.y,
                                x.w, x.h);
    const int c = (n & 15) * 3;
    // the group's 3 channels (12 bytes at a 4-byte-aligned offset) lie inside the 16-byte window
    // that starts at the 8-byte boundary below them: two 8-byte loads per corner instead of three
    // 4-byte ones (the kernel is bound by L1 sector requests); X2: the view allows it
    const int c2 = c & ~1, odd = c & 1;
    auto corner = [&](int yy, int xx, float (&o)[3]) {
      if (HS) {
        // 3 binary16 channels (6 bytes at a 2-byte-aligned offset) inside two aligned 4-byte words
        const uint16_t *hp = reinterpret_cast<const uint16_t *>(x.data) + vidx(x, yy, xx);
        const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t *>(hp + c2));
        const uint32_t w1 = __ldg(reinterpret_cast<const uint32_t *>(hp + c2 + 2));
        const float2 lo = __half22float2(*reinterpret_cast<const __half2 *>(&w0));
        const float2 hi = __half22float2(*reinterpret_cast<const __half2 *>(&w1));
        o[0] = odd ? lo.y : lo.x; o[1] = odd ? hi.x : lo.y; o[2] = odd ? hi.y : hi.x;
        return;
      }
      const float *cp = x.data + vidx(x, yy, xx);
      if (X2) {
        const float2 lo = __ldg(reinterpret_cast<const float2 *>(cp + c2));
        const float2 hi = __ldg(reinterpret_cast<const float2 *>(cp + c2 + 2));
        o[0] = odd ? lo.y : lo.x; o[1] = odd ? hi.x : lo.y; o[2] = odd ? hi.y : hi.x;
      } else {
        o[0] = __ldg(cp + c); o[1] = __ldg(cp + c + 1); o[2] = __ldg(cp + c + 2);
      }
    };
    float q00[3], q01[3] = {0.f, 0.f, 0.f}, q10[3] = {0.f, 0.f, 0.f}, q11[3] = {0.f, 0.f, 0.f};
    corner(b.y0, b.x0, q00);
    if (b.vx1) corner(b.y0, b.x1, q01);
    if (b.vy1) corner(b.y1, b.x0, q10);
    if (b.vx1 && b.vy1) corner(b.y1, b.x1, q11);
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      float s = b.nw * q00[j];
      if (b.vx1) s += b.ne * q01[j];
      if (b.vy1) s += b.sw * q10[j];
      if (b.vx1 && b.vy1) s += b.se * q11[j];
      v[k * 3 + j] = s * mask;
    }
  }
  float *op = out.data + vidx(out, py, px) + 3 * q;
#pragma unroll
  for (int o = 0; o < 3; ++o) {
    const float *wr = fw + (3 * q + o) * 6;
    float acc = fb ? fb[3 * q + o] : 0.f;
#pragma unroll
    for (int r = 0; r < 6; ++r) acc = fmaf(v[r], wr[r], acc);
    if (HO) reinterpret_cast<__half *>(out.data)[vidx(out, py, px) + 3 * q + o] = __float2half_rn(acc);
    else op[o] = acc;
  }
}

}  // namespace

#define LAUNCH_1D(kernel, total, stream, ...)                                              \
  do {                                                                                     \
    kernel<<<cdiv((total), TPB), TPB, 0, (cudaStream_t)(stream)>>>(__VA_ARGS__);           \
    DCVC_LAUNCHED();                                                                       \
  } while (0)

extern "C" int dcvc_nchw_to_nhwc(const float *src, const dcvc_view *dst, void *stream) {
  if (!src || !view_ok(dst)) return dcvc_fail(DCVC_ERR_ARG, "nchw_to_nhwc: bad args");
  LAUNCH_1D(nchw_to_nhwc_kernel, (int64_t)dst->h * dst->w * dst->c, stream, src, *dst);
  return DCVC_OK;
}

extern "C" int dcvc_nhwc_to_nchw(const dcvc_view *src, float *dst, int clamp01, void *stream) {
  if (!dst || !view_ok(src)) return dcvc_fail(DCVC_ERR_ARG, "nhwc_to_nchw: bad args");
  LAUNCH_1D(nhwc_to_nchw_kernel, (int64_t)src->h * src->w * src->c, stream, *src, dst, clamp01);
  return DCVC_OK;
}

extern "C" int dcvc_copy(const dcvc_view *src, const dcvc_view *dst, float mul, float slope,
                         void *stream) {
  if (!view_ok(src) || !view_ok(dst) || src->c != dst->c)
    return dcvc_fail(DCVC_ERR_ARG, "copy: bad args");
  const int64_t total4 = (int64_t)dst->h * dst->w * (dst->c / 4);
  if (view_al16(src) && view_al16(dst) && total4 < (1ll << 31)) {
    if (src->h == dst->h && src->w == dst->w)
      LAUNCH_1D(copy_v4_kernel<true>, total4, stream, *src, *dst, mul, slope,
                (uint32_t)(dst->c / 4), (uint32_t)total4);
    else
      LAUNCH_1D(copy_v4_kernel<false>, total4, stream, *src, *dst, mul, slope,
                (uint32_t)(dst->c / 4), (uint32_t)total4);
    return DCVC_OK;
  }
  LAUNCH_1D(copy_kernel, (int64_t)dst->h * dst->w * dst->c, stream, *src, *dst, mul, slope);
  return DCVC_OK;
}

extern "C" int dcvc_copy_f16(const dcvc_view *src, const dcvc_view *dst, float mul, float slope,
                             void *stream) {
  if (!view_ok(src) || !view_ok(dst) || src->c != dst->c || src->h != dst->h || src->w != dst->w)
    return dcvc_fail(DCVC_ERR_ARG, "copy_f16: bad args");
  const int64_t total8 = (int64_t)dst->h * dst->w * (dst->c / 8);
  if (!view_al16(src) || ((dst->cstride | dst->coff | dst->c) & 7) ||
      (reinterpret_cast<uintptr_t>(dst->data) & 15) || total8 >= (1ll << 31))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "copy_f16: views must hold whole 16-byte channel groups");
  LAUNCH_1D(copy_f16_kernel<false>, total8, stream, *src, *dst, mul, slope, (uint32_t)(dst->c / 8),
            (uint32_t)total8);
  return DCVC_OK;
}

extern "C" int dcvc_copy_f16_narrow(const dcvc_view *src, const dcvc_view *dst, void *stream) {
  if (!view_ok(src) || !view_ok(dst) || src->c != dst->c || src->h != dst->h || src->w != dst->w ||
      src->c >= 8)
    return dcvc_fail(DCVC_ERR_ARG, "copy_f16_narrow: bad args");
  const int64_t total = (int64_t)dst->h * dst->w;
  if (((dst->cstride | dst->coff) & 7) || dst->coff + 8 > dst->cstride ||
      (reinterpret_cast<uintptr_t>(dst->data) & 15) || total >= (1ll << 31))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "copy_f16_narrow: the destination must own a whole 16-byte group");
  LAUNCH_1D(copy_f16_narrow_kernel, total, stream, *src, *dst, (uint32_t)total);
  return DCVC_OK;
}

extern "C" int dcvc_copy_f16_f16(const dcvc_view *src, const dcvc_view *dst, float mul, float slope,
                                 void *stream) {
  if (!view_ok(src) || !view_ok(dst) || src->c != dst->c || src->h != dst->h || src->w != dst->w)
    return dcvc_fail(DCVC_ERR_ARG, "copy_f16_f16: bad args");
  const int64_t total8 = (int64_t)dst->h * dst->w * (dst->c / 8);
  if (((src->cstride | src->coff | dst->cstride | dst->coff | dst->c) & 7) ||
      ((reinterpret_cast<uintptr_t>(src->data) | reinterpret_cast<uintptr_t>(dst->data)) & 15) ||
      total8 >= (1ll << 31))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "copy_f16_f16: views must hold whole 16-byte channel groups");
  LAUNCH_1D(copy_f16_kernel<true>, total8, stream, *src, *dst, mul, slope, (uint32_t)(dst->c / 8),
            (uint32_t)total8);
  return DCVC_OK;
}

extern "C" int dcvc_fill(const dcvc_view *dst, float value, void *stream) {
  if (!view_ok(dst)) return dcvc_fail(DCVC_ERR_ARG, "fill: bad args");
  LAUNCH_1D(fill_kernel, (int64_t)dst->h * dst->w * dst->c, stream, *dst, value);
  return DCVC_OK;
}

extern "C" int dcvc_pool2(const dcvc_view *src, const dcvc_view *dst, int mode, void *stream) {
  if (!view_ok(src) || !view_ok(dst) || src->c != dst->c || dst->h != src->h / 2 ||
      dst->w != src->w / 2)
    return dcvc_fail(DCVC_ERR_ARG, "pool2: bad args");
  const int64_t total4 = (int64_t)dst->h * dst->w * (dst->c / 4);
  if (view_al16(src) && view_al16(dst) && total4 < (1ll << 31)) {
    LAUNCH_1D(pool2_v4_kernel, total4, stream, *src, *dst, mode, (uint32_t)(dst->c / 4),
              (uint32_t)total4);
    return DCVC_OK;
  }
  LAUNCH_1D(pool2_kernel, (int64_t)dst->h * dst->w * dst->c, stream, *src, *dst, mode);
  return DCVC_OK;
}

extern "C" int dcvc_pool2_f16(const dcvc_view *src, const dcvc_view *dst, int mode, void *stream) {
  if (!view_ok(src) || !view_ok(dst) || src->c != dst->c || dst->h != src->h / 2 ||
      dst->w != src->w / 2)
    return dcvc_fail(DCVC_ERR_ARG, "pool2_f16: bad args");
  const int64_t total8 = (int64_t)dst->h * dst->w * (dst->c / 8);
  if (((src->cstride | src->coff | dst->cstride | dst->coff | dst->c) & 7) ||
      ((reinterpret_cast<uintptr_t>(src->data) | reinterpret_cast<uintptr_t>(dst->data)) & 15) ||
      total8 >= (1ll << 31))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "pool2_f16: views must hold whole 16-byte channel groups");
  LAUNCH_1D(pool2_f16_kernel, total8, stream, *src, *dst, mode, (uint32_t)(dst->c / 8), (uint32_t)total8);
  return DCVC_OK;
}

extern "C" int dcvc_resize_bilinear(const dcvc_view *src, const dcvc_view *dst, float mul,
                                    void *stream) {
  if (!view_ok(src) || !view_ok(dst) || src->c != dst->c)
    return dcvc_fail(DCVC_ERR_ARG, "resize: bad args");
  const float sh = (float)src->h / (float)dst->h, sw = (float)src->w / (float)dst->w;
  LAUNCH_1D(resize_bilinear_kernel, (int64_t)dst->h * dst->w * dst->c, stream, *src, *dst, sh, sw,
            mul);
  return DCVC_OK;
}

extern "C" int dcvc_flow_warp(const dcvc_view *feat, const dcvc_view *flow, const dcvc_view *out,
                              void *stream) {
  if (!view_ok(feat) || !view_ok(flow) || !view_ok(out) || flow->c != 2 || out->c != feat->c ||
      out->h != flow->h || out->w != flow->w)
    return dcvc_fail(DCVC_ERR_ARG, "flow_warp: bad args");
  const int64_t total4 = (int64_t)out->h * out->w * (out->c / 4);
  if (view_al16(feat) && view_al16(out) && ((flow->cstride | flow->coff) & 1) == 0 &&
      (reinterpret_cast<uintptr_t>(flow->data) & 7) == 0 && total4 < (1ll << 31)) {
    LAUNCH_1D((flow_warp_v4_kernel<false, false>), total4, stream, *feat, *flow, *out, (uint32_t)(out->c / 4),
              (uint32_t)total4);
    return DCVC_OK;
  }
  LAUNCH_1D(flow_warp_kernel, (int64_t)out->h * out->w * out->c, stream, *feat, *flow, *out);
  return DCVC_OK;
}

extern "C" int dcvc_offset_diversity(const dcvc_view *x, const dcvc_view *offs_half,
                                     const dcvc_view *flow, const float *fusion_w,
                                     const float *fusion_b, const dcvc_view *out, void *stream) {
  if (!view_ok(x) || !view_ok(offs_half) || !view_ok(flow) || !view_ok(out) || !fusion_w ||
      x->c != 48 || offs_half->c != 96 || flow->c != 2 || out->c != 48 || out->h != x->h ||
      out->w != x->w || flow->h != x->h || flow->w != x->w)
    return dcvc_fail(DCVC_ERR_ARG, "offset_diversity: bad args");
  const int64_t total = (int64_t)out->h * out->w * 16;
  const bool vec = ((offs_half->cstride | offs_half->coff) & 3) == 0 &&
                   (reinterpret_cast<uintptr_t>(offs_half->data) & 15) == 0 &&
                   ((flow->cstride | flow->coff) & 1) == 0 &&
                   (reinterpret_cast<uintptr_t>(flow->data) & 7) == 0 && total < (1ll << 31);
  if (vec) {
    const bool x2 = ((x->cstride | x->coff) & 1) == 0 && (reinterpret_cast<uintptr_t>(x->data) & 7) == 0;
    if (x2)
      offset_diversity_v4_kernel<false, true><<<(unsigned)cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(
          *x, *offs_half, *flow, fusion_w, fusion_b, *out, (uint32_t)total);
    else
      offset_diversity_v4_kernel<false, false><<<(unsigned)cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(
          *x, *offs_half, *flow, fusion_w, fusion_b, *out, (uint32_t)total);
    DCVC_LAUNCHED();
    return DCVC_OK;
  }
  LAUNCH_1D(offset_diversity_kernel, total, stream, *x, *offs_half, *flow, fusion_w, fusion_b, *out);
  return DCVC_OK;
}

extern "C" int dcvc_flow_warp_f16(const dcvc_view *feat, const dcvc_view *flow, const dcvc_view *out,
                                  void *stream) {
  if (!view_ok(feat) || !view_ok(flow) || !view_ok(out) || flow->c != 2 || out->c != feat->c ||
      out->h != flow->h || out->w != flow->w)
    return dcvc_fail(DCVC_ERR_ARG, "flow_warp_f16: bad args");
  const int64_t total4 = (int64_t)out->h * out->w * (out->c / 4);
  const bool ok = view_al16(feat) && ((out->cstride | out->coff | out->c) & 3) == 0 &&
                  (reinterpret_cast<uintptr_t>(out->data) & 7) == 0 &&
                  ((flow->cstride | flow->coff) & 1) == 0 &&
                  (reinterpret_cast<uintptr_t>(flow->data) & 7) == 0 && total4 < (1ll << 31);
  if (!ok) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "flow_warp_f16: views must hold whole 4-channel groups");
  LAUNCH_1D((flow_warp_v4_kernel<true, false>), total4, stream, *feat, *flow, *out, (uint32_t)(out->c / 4),
            (uint32_t)total4);
  return DCVC_OK;
}

extern "C" int dcvc_flow_warp_f16_f16(const dcvc_view *feat, const dcvc_view *flow, const dcvc_view *out,
                                      void *stream) {
  if (!view_ok(feat) || !view_ok(flow) || !view_ok(out) || flow->c != 2 || out->c != feat->c ||
      out->h != flow->h || out->w != flow->w)
    return dcvc_fail(DCVC_ERR_ARG, "flow_warp_f16_f16: bad args");
  const int64_t total4 = (int64_t)out->h * out->w * (out->c / 4);
  const bool ok = ((feat->cstride | feat->coff | out->cstride | out->coff | out->c) & 3) == 0 &&
                  ((reinterpret_cast<uintptr_t>(feat->data) | reinterpret_cast<uintptr_t>(out->data)) & 7) == 0 &&
                  ((flow->cstride | flow->coff) & 1) == 0 &&
                  (reinterpret_cast<uintptr_t>(flow->data) & 7) == 0 && total4 < (1ll << 31);
  if (!ok) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "flow_warp_f16_f16: views must hold whole 4-channel groups");
  if (((feat->cstride | feat->coff | out->cstride | out->coff | out->c) & 7) == 0 &&
      ((reinterpret_cast<uintptr_t>(feat->data) | reinterpret_cast<uintptr_t>(out->data)) & 15) == 0) {
    LAUNCH_1D(flow_warp_v8h_kernel, total4 / 2, stream, *feat, *flow, *out, (uint32_t)(out->c / 8),
              (uint32_t)(total4 / 2));
    return DCVC_OK;
  }
  LAUNCH_1D((flow_warp_v4_kernel<true, true>), total4, stream, *feat, *flow, *out, (uint32_t)(out->c / 4),
            (uint32_t)total4);
  return DCVC_OK;
}

extern "C" int dcvc_offset_diversity_f16_f16(const dcvc_view *x, const dcvc_view *offs_half,
                                             const dcvc_view *flow, const float *fusion_w,
                                             const float *fusion_b, const dcvc_view *out, void *stream) {
  if (!view_ok(x) || !view_ok(offs_half) || !view_ok(flow) || !view_ok(out) || !fusion_w ||
      x->c != 48 || offs_half->c != 96 || flow->c != 2 || out->c != 48 || out->h != x->h ||
      out->w != x->w || flow->h != x->h || flow->w != x->w)
    return dcvc_fail(DCVC_ERR_ARG, "offset_diversity_f16_f16: bad args");
  const int64_t total = (int64_t)out->h * out->w * 16;
  const bool vec = ((offs_half->cstride | offs_half->coff) & 3) == 0 &&
                   (reinterpret_cast<uintptr_t>(offs_half->data) & 15) == 0 &&
                   ((flow->cstride | flow->coff) & 1) == 0 &&
                   (reinterpret_cast<uintptr_t>(flow->data) & 7) == 0 && total < (1ll << 31) &&
                   ((x->cstride | x->coff) & 1) == 0 && (reinterpret_cast<uintptr_t>(x->data) & 3) == 0;
  if (!vec) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "offset_diversity_f16_f16: unaligned views");
  offset_diversity_v4_kernel<true, true, true><<<(unsigned)cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(
      *x, *offs_half, *flow, fusion_w, fusion_b, *out, (uint32_t)total);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_offset_diversity_f16(const dcvc_view *x, const dcvc_view *offs_half,
                                         const dcvc_view *flow, const float *fusion_w,
                                         const float *fusion_b, const dcvc_view *out, void *stream) {
  if (!view_ok(x) || !view_ok(offs_half) || !view_ok(flow) || !view_ok(out) || !fusion_w ||
      x->c != 48 || offs_half->c != 96 || flow->c != 2 || out->c != 48 || out->h != x->h ||
      out->w != x->w || flow->h != x->h || flow->w != x->w)
    return dcvc_fail(DCVC_ERR_ARG, "offset_diversity_f16: bad args");
  const int64_t total = (int64_t)out->h * out->w * 16;
  const bool vec = ((offs_half->cstride | offs_half->coff) & 3) == 0 &&
                   (reinterpret_cast<uintptr_t>(offs_half->data) & 15) == 0 &&
                   ((flow->cstride | flow->coff) & 1) == 0 &&
                   (reinterpret_cast<uintptr_t>(flow->data) & 7) == 0 && total < (1ll << 31);
  if (!vec) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "offset_diversity_f16: unaligned offset/flow views");
  const bool x2 = ((x->cstride | x->coff) & 1) == 0 && (reinterpret_cast<uintptr_t>(x->data) & 7) == 0;
  if (x2)
    offset_diversity_v4_kernel<true, true><<<(unsigned)cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(
        *x, *offs_half, *flow, fusion_w, fusion_b, *out, (uint32_t)total);
  else
    offset_diversity_v4_kernel<true, false><<<(unsigned)cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(
        *x, *offs_half, *flow, fusion_w, fusion_b, *out, (uint32_t)total);
  DCVC_LAUNCHED();
  return DCVC_OK;
}
