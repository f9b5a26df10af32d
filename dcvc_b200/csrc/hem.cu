// hem.cu -- the pieces DCVC-HEM needs beyond the DCVC-DC kernel set: the two-step checkerboard
// ("dual") spatial prior, per-channel scaling / division, and the squeeze-and-excitation gate.
//
//   dual prior : DCVC-HEM/src/models/common_model.py:84-188 (get_mask, process_with_mask,
//                forward_dual_prior, decompress_dual_prior) + GaussianEncoder.build_indexes
//                (src/entropy_models/entropy_models.py:264-268) + the bit estimates (:53-75)
//   scaling    : y / curr_q, y_hat * curr_q with q = LowerBound(q_basic, 0.5) * q_scale
//                (src/models/video_model.py:263-271, image_model.py:61-63)
//   SE layer   : SELayer (src/models/video_net.py:149-163): mean over H x W, two bias-free
//                linear layers (ReLU, sigmoid), per-channel gate
//
// Dual-prior algebra.  Channels split in halves (half 0 = first C/2 channels); the checkerboard
// mask_0 is 1 where (y + x) is even.  Step 0 codes half 0 at mask_0 positions and half 1 at
// mask_1 positions with (scales, means) from the prior fusion; step 1 codes the complement with
// (scales, means) from y_spatial_prior(cat(y_hat_0_0, y_hat_1_1, means, scales, quant_step)).
// The coder's symbol tensor of a step is [1, C/2, h, w]: entry (jj, y, x) holds channel
// jj + (coded half) * C/2, the one channel of the pair that the step codes at that position.
#include <cuda_fp16.h>

#include "common.cuh"

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ int scale_index(const float *__restrict__ edges, float s) {
  s = fmaxf(s, 1e-5f);
  int lo = 0, hi = 256;
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(edges + mid) <= s) lo = mid + 1; else hi = mid;
  }
  if (lo < 256 && __ldg(edges + lo) <= s) ++lo;
  return max(lo - 1, 0);
}
__device__ __forceinline__ float laplace_cdf(float x, float b) {
  const float sg = (x > 0.f) - (x < 0.f);
  return 0.5f - 0.5f * sg * expm1f(-fabsf(x) / b);
}
__device__ __forceinline__ float normal_cdf(float x, float s) {
  return 0.5f * (1.0f + erff(x * (1.0f / s) / 1.41421356237309504880f));
}
__device__ __forceinline__ float probs_to_bits(float p) {
  return fmaxf(-1.0f * logf(p + 1e-5f) / 0.69314718055994530942f, 0.f);
}
__device__ __forceinline__ float block_sum(float v) {
  __shared__ float sm[TPB / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x == 0)
    for (int i = 0; i < TPB / 32; ++i) r += sm[i];
  return r;
}
__global__ void reduce_partials_kernel(const float *__restrict__ part, int n, float *bits) {
  __shared__ float sm[TPB];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += TPB) acc += part[i];
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int o = TPB / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *bits += sm[0];
}

struct DualGeom {
  int jj, x, y, C, H2, j, jo;   // j = coded channel of the pair, jo = the other one
  size_t sym_pos;
};

__device__ __forceinline__ bool dual_geom(const dcvc_dual &q, int64_t t, DualGeom &g) {
  const int C = q.cat.c / 4, H2 = C / 2;
  const int64_t total = (int64_t)q.cat.h * q.cat.w * H2;
  if (t >= total) return false;
  g.C = C; g.H2 = H2;
  g.jj = (int)(t % H2);
  const int64_t p = t / H2;
  g.x = (int)(p % q.cat.w);
  g.y = (int)(p / q.cat.w);
  const bool mask0 = ((g.x + g.y) & 1) == 0;
  const bool half1 = (q.step == 0) ? !mask0 : mask0;    // which half the step codes here
  g.j = g.jj + (half1 ? H2 : 0);
  g.jo = g.jj + (half1 ? 0 : H2);
  g.sym_pos = ((size_t)g.jj * q.cat.h + g.y) * q.cat.w + g.x;
  return true;
}

// (scale, mean, lower-bounded quant step) of channel j at the step's position
__device__ __forceinline__ void dual_params(const dcvc_dual &q, const DualGeom &g, int j, float &scale,
                                            float &mean, float &qs) {
  if (q.step == 0) {
    const float *f = q.fused.data + vidx(q.fused, g.y, g.x);     // [q_step | scales | means]
    qs = fmaxf(f[j], 0.5f);
    scale = f[g.C + j];
    mean = f[2 * g.C + j];
  } else {
    const float *c = q.cat.data + vidx(q.cat, g.y, g.x);         // [.. | means | scales | q_step]
    const float *p = q.prior.data + vidx(q.prior, g.y, g.x);     // [sc_0 | mean_0 | sc_1 | mean_1]
    const int half1 = j >= g.H2;
    qs = c[3 * g.C + j];
    scale = p[(half1 ? 2 : 0) * g.H2 + g.jj];
    mean = p[(half1 ? 3 : 1) * g.H2 + g.jj];
  }
}

// step 0 also builds the spatial-prior conv input: cat = [y_hat_0_0 | y_hat_1_1 | means | scales |
// max(q_step, .5)], zeros where the pair's other channel is not coded yet
__device__ __forceinline__ void dual_fill_cat(const dcvc_dual &q, const DualGeom &g) {
  const float *f = q.fused.data + vidx(q.fused, g.y, g.x);
  float *c = q.cat.data + vidx(q.cat, g.y, g.x);
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int j = k ? g.jo : g.j;
    c[g.C + j] = f[2 * g.C + j];
    c[2 * g.C + j] = f[g.C + j];
    c[3 * g.C + j] = fmaxf(f[j], 0.5f);
  }
  c[g.jo] = 0.f;
}

__global__ void dual_step_enc_kernel(dcvc_dual q, int16_t *__restrict__ symbols,
                                     uint8_t *__restrict__ indexes, float *__restrict__ partials) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  DualGeom g;
  float bits = 0.f;
  if (dual_geom(q, t, g)) {
    float scale, mean, qs;
    dual_params(q, g, g.j, scale, mean, qs);
    if (q.step == 0) dual_fill_cat(q, g);
    const float yv = q.y.data[vidx(q.y, g.y, g.x) + g.j] / qs;
    const float yq = rintf(yv - mean);
    const float yh = yq + mean;
    if (q.step == 0) q.cat.data[vidx(q.cat, g.y, g.x) + g.j] = yh;
    float out = yh * qs;
    if (q.post_scale) out *= q.post_scale[g.j];
    q.y_out.data[vidx(q.y_out, g.y, g.x) + g.j] = out;
    if (symbols) {
      symbols[g.sym_pos] = (int16_t)fminf(fmaxf(yq, -30000.f), 30000.f);
      indexes[g.sym_pos] = (uint8_t)scale_index(q.edges, scale);
    }
    if (partials) {
      const float sg = q.distribution == 0 ? fminf(fmaxf(scale, 1e-5f), 1e10f)
                                           : fminf(fmaxf(scale, 0.11f), 1e10f);
      const float p = q.distribution == 0
                          ? laplace_cdf(yq + 0.5f, sg) - laplace_cdf(yq - 0.5f, sg)
                          : normal_cdf(yq + 0.5f, sg) - normal_cdf(yq - 0.5f, sg);
      bits = probs_to_bits(p);
    }
  }
  if (partials) {
    const float s = block_sum(bits);
    if (threadIdx.x == 0) partials[blockIdx.x] = s;
  }
}

__global__ void dual_step_indexes_kernel(dcvc_dual q, uint8_t *__restrict__ indexes) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  DualGeom g;
  if (!dual_geom(q, t, g)) return;
  float scale, mean, qs;
  dual_params(q, g, g.j, scale, mean, qs);
  indexes[g.sym_pos] = (uint8_t)scale_index(q.edges, scale);
}

__global__ void dual_step_dec_kernel(dcvc_dual q, const int16_t *__restrict__ symbols) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  DualGeom g;
  if (!dual_geom(q, t, g)) return;
  float scale, mean, qs;
  dual_params(q, g, g.j, scale, mean, qs);
  if (q.step == 0) dual_fill_cat(q, g);
  const float yh = (float)symbols[g.sym_pos] + mean;
  if (q.step == 0) q.cat.data[vidx(q.cat, g.y, g.x) + g.j] = yh;
  float out = yh * qs;
  if (q.post_scale) out *= q.post_scale[g.j];
  q.y_out.data[vidx(q.y_out, g.y, g.x) + g.j] = out;
}

int dual_check(const dcvc_dual *q, bool need_y) {
  if (!q || !view_ok(&q->cat) || q->cat.c % 8 || !q->edges || q->step < 0 || q->step > 1)
    return dcvc_fail(DCVC_ERR_ARG, "dual: bad cat buffer / step");
  const int C = q->cat.c / 4;
  const int h = q->cat.h, w = q->cat.w;
  if (!view_ok(&q->y_out) || q->y_out.c != C || q->y_out.h != h || q->y_out.w != w)
    return dcvc_fail(DCVC_ERR_ARG, "dual: y_out extent mismatch");
  if (need_y && (!view_ok(&q->y) || q->y.c != C || q->y.h != h || q->y.w != w))
    return dcvc_fail(DCVC_ERR_ARG, "dual: y extent mismatch");
  if (q->step == 0 && (!view_ok(&q->fused) || q->fused.c != 3 * C || q->fused.h != h || q->fused.w != w))
    return dcvc_fail(DCVC_ERR_ARG, "dual: step 0 needs the prior-fusion output (3C channels)");
  if (q->step == 1 && (!view_ok(&q->prior) || q->prior.c != 2 * C || q->prior.h != h || q->prior.w != w))
    return dcvc_fail(DCVC_ERR_ARG, "dual: step 1 needs the spatial-prior output (2C channels)");
  return DCVC_OK;
}

// ---------------------------------------------------------------- per-channel scale / divide
__global__ void channel_scale_kernel(dcvc_view src, dcvc_view dst, const float *__restrict__ v, int divide) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)src.h * src.w * src.c;
  if (t >= total) return;
  const int c = (int)(t % src.c);
  const int64_t p = t / src.c;
  const int x = (int)(p % src.w), y = (int)(p / src.w);
  const float a = src.data[vidx(src, y, x) + c];
  dst.data[vidx(dst, y, x) + c] = divide ? a / v[c] : a * v[c];
}
// the same, one thread per 4 channels (views whose pixels are whole 16-byte groups)
__global__ void channel_scale4_kernel(dcvc_view src, dcvc_view dst, const float *__restrict__ v, int divide) {
  const int c4n = src.c >> 2;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)src.h * src.w * c4n) return;
  const int c = (int)(t % c4n) * 4;
  const int64_t p = t / c4n;
  const float4 a = *reinterpret_cast<const float4 *>(src.data + (size_t)p * src.cstride + src.coff + c);
  const float4 q = __ldg(reinterpret_cast<const float4 *>(v + c));
  float4 r;
  if (divide) { r.x = a.x / q.x; r.y = a.y / q.y; r.z = a.z / q.z; r.w = a.w / q.w; }
  else { r.x = a.x * q.x; r.y = a.y * q.y; r.z = a.z * q.z; r.w = a.w * q.w; }
  *reinterpret_cast<float4 *>(dst.data + (size_t)p * dst.cstride + dst.coff + c) = r;
}

// binary16 views (the SE branch of a ConvBlockResidual in tensor-core mode): 8 channels per thread,
// fp32 arithmetic, result rounded to nearest even
__global__ void channel_scale8h_kernel(dcvc_view src, dcvc_view dst, const float *__restrict__ v, int divide) {
  const int c8n = src.c >> 3;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)src.h * src.w * c8n) return;
  const int c = (int)(t % c8n) * 8;
  const int64_t p = t / c8n;
  const uint4 a = *reinterpret_cast<const uint4 *>(reinterpret_cast<const __half *>(src.data) +
                                                   (size_t)p * src.cstride + src.coff + c);
  const float4 q0 = __ldg(reinterpret_cast<const float4 *>(v + c)), q1 = __ldg(reinterpret_cast<const float4 *>(v + c + 4));
  const float q[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
  const uint32_t w[4] = {a.x, a.y, a.z, a.w};
  uint32_t o[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&w[j]));
    const float lo = divide ? f.x / q[2 * j] : f.x * q[2 * j];
    const float hi = divide ? f.y / q[2 * j + 1] : f.y * q[2 * j + 1];
    const __half2 h = __floats2half2_rn(lo, hi);
    o[j] = *reinterpret_cast<const uint32_t *>(&h);
  }
  *reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(dst.data) + (size_t)p * dst.cstride + dst.coff + c) =
      make_uint4(o[0], o[1], o[2], o[3]);
}

// ---------------------------------------------------------------- squeeze-and-excitation gate
// stage 1: each block sums SE_PIX pixels of every channel (float4 loads, coalesced along the
// channel axis; fp32 runs of 32 pixels folded into float64); stage 2: fixed-order sum over the
// blocks, the two bias-free linear layers, sigmoid.  Deterministic: no atomics.
constexpr int SE_PIX = 4096;
constexpr int SE_TPB = 256;
constexpr int SE_FIN_TPB = 512;       // threads of the final kernel (>= c)
template <bool H16>
__global__ void __launch_bounds__(SE_TPB) se_partial_kernel(dcvc_view x, double *__restrict__ part) {
  __shared__ double sm[SE_TPB * 4];
  const int nvec = x.c >> 2;                 // 4-channel groups per pixel (c % 4 == 0, nvec <= 64)
  const int lanes = SE_TPB / nvec;           // pixels in flight per iteration
  const int vec = threadIdx.x % nvec, lane = threadIdx.x / nvec;
  const int64_t npix = (int64_t)x.h * x.w;
  const int64_t p0 = (int64_t)blockIdx.x * SE_PIX;
  const int64_t p1 = p0 + SE_PIX < npix ? p0 + SE_PIX : npix;
  double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
  auto load = [&](int64_t p) -> float4 {
    if (H16) {
      const uint2 u = *reinterpret_cast<const uint2 *>(reinterpret_cast<const __half *>(x.data) +
                                                       (size_t)p * x.cstride + x.coff + 4 * vec);
      const float2 lo = __half22float2(*reinterpret_cast<const __half2 *>(&u.x));
      const float2 hi = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
      return make_float4(lo.x, lo.y, hi.x, hi.y);
    }
    return *reinterpret_cast<const float4 *>(x.data + (size_t)p * x.cstride + x.coff + 4 * vec);
  };
  if (lane < lanes) {
    // the same summation order as ever (runs of 32 pixels in fp32, folded into float64), with four
    // loads in flight per thread
    float4 run = make_float4(0.f, 0.f, 0.f, 0.f);
    int cnt = 0;
    int64_t p = p0 + lane;
    for (; p + 3 * (int64_t)lanes < p1; p += 4 * (int64_t)lanes) {
      const float4 v0 = load(p), v1 = load(p + lanes), v2 = load(p + 2 * (int64_t)lanes),
                   v3 = load(p + 3 * (int64_t)lanes);
      const float4 vs[4] = {v0, v1, v2, v3};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        run.x += vs[j].x; run.y += vs[j].y; run.z += vs[j].z; run.w += vs[j].w;
        if (++cnt == 32) {
          a0 += run.x; a1 += run.y; a2 += run.z; a3 += run.w;
          run = make_float4(0.f, 0.f, 0.f, 0.f);
          cnt = 0;
        }
      }
    }
    for (; p < p1; p += lanes) {
      const float4 v = load(p);
      run.x += v.x; run.y += v.y; run.z += v.z; run.w += v.w;
      if (++cnt == 32) {
        a0 += run.x; a1 += run.y; a2 += run.z; a3 += run.w;
        run = make_float4(0.f, 0.f, 0.f, 0.f);
        cnt = 0;
      }
    }
    a0 += run.x; a1 += run.y; a2 += run.z; a3 += run.w;
  }
  sm[threadIdx.x * 4 + 0] = a0; sm[threadIdx.x * 4 + 1] = a1;
  sm[threadIdx.x * 4 + 2] = a2; sm[threadIdx.x * 4 + 3] = a3;
  __syncthreads();
  if (threadIdx.x < x.c) {                   // thread = channel: sum its column over the lanes
    const int c = threadIdx.x, v = c >> 2, k = c & 3;
    double s = 0.0;
    for (int l = 0; l < lanes; ++l) s += sm[(l * nvec + v) * 4 + k];
    part[(size_t)blockIdx.x * x.c + c] = s;
  }
}
__global__ void se_gate_kernel(const double *__restrict__ part, int nblocks, int c, int64_t pixels,
                               const float *__restrict__ w1, const float *__restrict__ w2, int hidden,
                               float *__restrict__ gate) {
  extern __shared__ float sm[];      // mean[c], hid[hidden]
  __shared__ double psum[SE_FIN_TPB];
  float *mean = sm, *hid = sm + c;
  // the blocks' partial sums, in a fixed order: thread (g, i) adds the blocks b = g, g + G, ...
  // of channel i (G = blockDim / c groups, so that the loads of one channel are spread over G
  // threads instead of one dependent chain), then channel i's G sums are added in order of g
  const int G = blockDim.x / c;
  {
    const int g = threadIdx.x / c, i = threadIdx.x - g * c;
    double acc = 0.0;
    if (g < G)
      for (int b = g; b < nblocks; b += G) acc += part[(size_t)b * c + i];
    psum[threadIdx.x] = acc;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < c; i += blockDim.x) {
    double acc = 0.0;
    for (int g = 0; g < G; ++g) acc += psum[g * c + i];
    mean[i] = (float)(acc / (double)pixels);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < hidden; k += blockDim.x) {
    float acc = 0.f;
    for (int i = 0; i < c; ++i) acc += mean[i] * w1[(size_t)k * c + i];     // Linear(c, c/r).weight [c/r][c]
    hid[k] = fmaxf(acc, 0.f);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < c; i += blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < hidden; ++k) acc += hid[k] * w2[(size_t)i * hidden + k];   // [c][c/r]
    gate[i] = 1.0f / (1.0f + expf(-acc));
  }
}

}  // namespace

extern "C" int dcvc_dual_step_enc(const dcvc_dual *q, int16_t *symbols, uint8_t *indexes, float *bits,
                                  void *bits_ws, void *stream) {
  int rc = dual_check(q, true);
  if (rc) return rc;
  if ((symbols == nullptr) != (indexes == nullptr) || (bits && !bits_ws))
    return dcvc_fail(DCVC_ERR_ARG, "dual_step_enc: symbols/indexes/bits arguments inconsistent");
  const int64_t n = (int64_t)q->cat.h * q->cat.w * (q->cat.c / 8);
  const int blocks = cdiv(n, TPB);
  cudaStream_t s = (cudaStream_t)stream;
  dual_step_enc_kernel<<<blocks, TPB, 0, s>>>(*q, symbols, indexes, bits ? (float *)bits_ws : nullptr);
  DCVC_LAUNCHED();
  if (bits) {
    reduce_partials_kernel<<<1, TPB, 0, s>>>((const float *)bits_ws, blocks, bits);
    DCVC_LAUNCHED();
  }
  return DCVC_OK;
}

extern "C" int dcvc_dual_step_indexes(const dcvc_dual *q, uint8_t *indexes, void *stream) {
  int rc = dual_check(q, false);
  if (rc) return rc;
  if (!indexes) return dcvc_fail(DCVC_ERR_ARG, "dual_step_indexes: null output");
  const int64_t n = (int64_t)q->cat.h * q->cat.w * (q->cat.c / 8);
  dual_step_indexes_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(*q, indexes);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_dual_step_dec(const dcvc_dual *q, const int16_t *symbols, void *stream) {
  int rc = dual_check(q, false);
  if (rc) return rc;
  if (!symbols) return dcvc_fail(DCVC_ERR_ARG, "dual_step_dec: null symbols");
  const int64_t n = (int64_t)q->cat.h * q->cat.w * (q->cat.c / 8);
  dual_step_dec_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(*q, symbols);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_channel_scale(const dcvc_view *src, const dcvc_view *dst, const float *vec, int divide,
                                  void *stream) {
  if (!view_ok(src) || !view_ok(dst) || !vec || src->h != dst->h || src->w != dst->w || src->c != dst->c)
    return dcvc_fail(DCVC_ERR_ARG, "channel_scale: bad arguments");
  const int64_t n = (int64_t)src->h * src->w * src->c;
  if (view_al16(src) && view_al16(dst) && (reinterpret_cast<uintptr_t>(vec) & 15) == 0)
    channel_scale4_kernel<<<cdiv(n / 4, TPB), TPB, 0, (cudaStream_t)stream>>>(*src, *dst, vec, divide);
  else
    channel_scale_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(*src, *dst, vec, divide);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_channel_scale_f16(const dcvc_view *src, const dcvc_view *dst, const float *vec,
                                      int divide, void *stream) {
  if (!view_ok(src) || !view_ok(dst) || !vec || src->h != dst->h || src->w != dst->w || src->c != dst->c)
    return dcvc_fail(DCVC_ERR_ARG, "channel_scale_f16: bad arguments");
  auto al = [](const dcvc_view *v) {
    return ((v->cstride | v->coff | v->c) & 7) == 0 && (reinterpret_cast<uintptr_t>(v->data) & 15) == 0;
  };
  if (!al(src) || !al(dst) || (reinterpret_cast<uintptr_t>(vec) & 15))
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "channel_scale_f16: views must hold whole 8-channel groups");
  const int64_t n = (int64_t)src->h * src->w * (src->c / 8);
  channel_scale8h_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(*src, *dst, vec, divide);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" size_t dcvc_se_ws_bytes(int h, int w, int c) {
  return (size_t)cdiv((int64_t)h * w, SE_PIX) * (size_t)c * sizeof(double);
}

static int se_gate_impl(const dcvc_view *x, bool f16, const float *w1, const float *w2, int hidden,
                        float *gate, void *ws, void *stream) {
  if (!view_ok(x) || !w1 || !w2 || hidden < 1 || !gate || !ws)
    return dcvc_fail(DCVC_ERR_ARG, "se_gate: bad arguments");
  const bool al = f16 ? ((x->cstride | x->coff | x->c) & 3) == 0 && (reinterpret_cast<uintptr_t>(x->data) & 7) == 0
                      : view_al16(x);
  if (!al || x->c > SE_TPB)
    return dcvc_fail(DCVC_ERR_UNSUPPORTED, "se_gate: needs an aligned view of whole 4-channel groups, at most %d channels", SE_TPB);
  const int nblocks = cdiv((int64_t)x->h * x->w, SE_PIX);
  cudaStream_t s = (cudaStream_t)stream;
  if (f16) se_partial_kernel<true><<<nblocks, SE_TPB, 0, s>>>(*x, (double *)ws);
  else se_partial_kernel<false><<<nblocks, SE_TPB, 0, s>>>(*x, (double *)ws);
  DCVC_LAUNCHED();
  se_gate_kernel<<<1, SE_FIN_TPB, (size_t)(x->c + hidden) * sizeof(float), s>>>(
      (const double *)ws, nblocks, x->c, (int64_t)x->h * x->w, w1, w2, hidden, gate);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_se_gate(const dcvc_view *x, const float *w1, const float *w2, int hidden, float *gate,
                            void *ws, void *stream) {
  return se_gate_impl(x, false, w1, w2, hidden, gate, ws, stream);
}

extern "C" int dcvc_se_gate_f16(const dcvc_view *x, const float *w1, const float *w2, int hidden,
                                float *gate, void *ws, void *stream) {
  return se_gate_impl(x, true, w1, w2, hidden, gate, ws, stream);
}
