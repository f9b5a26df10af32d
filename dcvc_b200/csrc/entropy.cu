// entropy.cu -- quad-part spatial-temporal entropy model kernels (fp32, NHWC latents).
//
// Replaces the ~150 tiny ATen launches per latent of the reference
// (DCVC-DC/src/models/common_model.py:92-321: process_with_mask, combine_four_parts, masks,
// round/clamp/chunk), GaussianEncoder.build_indexes (entropy_models.py:269-273), the bit
// estimates (common_model.py:39-61) and BitEstimator.get_cdf (entropy_models.py:58-122) with one
// fused kernel per coding step.  HBM-bound, tiny: each element is read once and written once.
//
// Mask algebra: position id m = (y&1)*2 + (x&1); step s codes group g = m ^ XOR_OF[s]
// (the reference's table [[0,1,2,3],[3,2,1,0],[2,3,0,1],[1,0,3,2]], common_model.py:168-220).
#include "common.cuh"

#include <mutex>

namespace {

constexpr int TPB = 256;
__device__ __constant__ int XOR_OF[4] = {0, 3, 2, 1};

// number of edges <= s, minus 1  ==  GaussianEncoder.build_indexes (edges built by the host)
__device__ __forceinline__ int scale_index(const float *__restrict__ edges, float s) {
  s = fmaxf(s, 1e-5f);
  int lo = 0, hi = 256;  // first position with edges[pos] > s
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(edges + mid) <= s) lo = mid + 1; else hi = mid;
  }
  if (lo < 256 && __ldg(edges + lo) <= s) ++lo;
  return max(lo - 1, 0);
}

__device__ __forceinline__ float laplace_cdf(float x, float b) {
  // torch.distributions.Laplace.cdf: 0.5 - 0.5 * sign(z) * expm1(-|z| / b)
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
  return r;  // valid on thread 0
}

__global__ void reduce_partials_kernel(const float *__restrict__ part, int n, float *bits) {
  // single block, fixed order => run-to-run deterministic
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

struct StepGeom {
  int j, x, y, ch, C;
  size_t sym_pos;
};

__device__ __forceinline__ bool step_geom(const dcvc_view &params, int step, int64_t t, StepGeom &g) {
  const int C = params.c / 3, C4 = C / 4;
  const int64_t total = (int64_t)params.h * params.w * C4;
  if (t >= total) return false;
  g.C = C;
  g.j = (int)(t % C4);
  const int64_t p = t / C4;
  g.x = (int)(p % params.w);
  g.y = (int)(p / params.w);
  const int m = (g.y & 1) * 2 + (g.x & 1);
  g.ch = (m ^ XOR_OF[step]) * C4 + g.j;
  g.sym_pos = ((size_t)g.j * params.h + g.y) * params.w + g.x;
  return true;
}

__device__ __forceinline__ void step_scale_mean(const dcvc_quad &q, const StepGeom &g, float &scale,
                                                float &mean) {
  if (q.prior.data) {
    const float *pp = q.prior.data + vidx(q.prior, g.y, g.x);
    scale = pp[g.ch];
    mean = pp[g.C + g.ch];
  } else {
    const float *pp = q.params.data + vidx(q.params, g.y, g.x);
    scale = pp[g.C + g.ch];
    mean = pp[2 * g.C + g.ch];
  }
}

__global__ void quad_step_enc_kernel(dcvc_quad q, int16_t *__restrict__ symbols,
                                     uint8_t *__restrict__ indexes, float *__restrict__ partials) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  StepGeom g;
  float bits = 0.f;
  if (step_geom(q.params, q.step, t, g)) {
    float scale, mean;
    step_scale_mean(q, g, scale, mean);
    const float qs = fmaxf(q.params.data[vidx(q.params, g.y, g.x) + g.ch], 0.5f);
    const float yv = q.y.data[vidx(q.y, g.y, g.x) + g.ch] / qs;
    const float yq = rintf(yv - mean);          // torch.round: half to even
    const float yh = yq + mean;
    q.y_hat.data[vidx(q.y_hat, g.y, g.x) + g.ch] = yh;
    q.y_out.data[vidx(q.y_out, g.y, g.x) + g.ch] = yh * qs;
    if (symbols) {
      symbols[g.sym_pos] = (int16_t)fminf(fmaxf(yq, -30000.f), 30000.f);
      indexes[g.sym_pos] = (uint8_t)scale_index(q.edges, scale);
    }
    if (partials) {
      const float sg = fminf(fmaxf(scale, 1e-5f), 1e10f);
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

__global__ void quad_step_indexes_kernel(dcvc_quad q, uint8_t *__restrict__ indexes) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  StepGeom g;
  if (!step_geom(q.params, q.step, t, g)) return;
  float scale, mean;
  step_scale_mean(q, g, scale, mean);
  indexes[g.sym_pos] = (uint8_t)scale_index(q.edges, scale);
}

__global__ void quad_step_dec_kernel(dcvc_quad q, const int16_t *__restrict__ symbols) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  StepGeom g;
  if (!step_geom(q.params, q.step, t, g)) return;
  float scale, mean;
  step_scale_mean(q, g, scale, mean);
  const float qs = fmaxf(q.params.data[vidx(q.params, g.y, g.x) + g.ch], 0.5f);
  const float yh = (float)symbols[g.sym_pos] + mean;
  q.y_hat.data[vidx(q.y_hat, g.y, g.x) + g.ch] = yh;
  q.y_out.data[vidx(q.y_out, g.y, g.x) + g.ch] = yh * qs;
}

__device__ __forceinline__ float softplusf(float x) { return x > 20.f ? x : log1pf(expf(x)); }

__device__ __forceinline__ float factorized_cdf(const float *__restrict__ fp, int C, int c, float x) {
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    x = x * softplusf(fp[(3 * k) * C + c]) + fp[(3 * k + 1) * C + c];
    x = x + tanhf(x) * tanhf(fp[(3 * k + 2) * C + c]);
  }
  x = x * softplusf(fp[9 * C + c]) + fp[10 * C + c];
  return 1.0f / (1.0f + expf(-x));
}

__global__ void z_quant_kernel(dcvc_view z, dcvc_view zh, int16_t *__restrict__ symbols,
                               const float *__restrict__ fparams, float *__restrict__ partials) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)z.h * z.w * z.c;
  float bits = 0.f;
  if (t < total) {
    const int c = (int)(t % z.c);
    const int64_t p = t / z.c;
    const int x = (int)(p % z.w), y = (int)(p / z.w);
    const float v = rintf(z.data[vidx(z, y, x) + c]);
    zh.data[vidx(zh, y, x) + c] = v;
    if (symbols)
      symbols[((size_t)c * z.h + y) * z.w + x] = (int16_t)fminf(fmaxf(v, -30000.f), 30000.f);
    if (partials && fparams) {
      const float pr = factorized_cdf(fparams, z.c, c, v + 0.5f) - factorized_cdf(fparams, z.c, c, v - 0.5f);
      bits = probs_to_bits(pr);
    }
  }
  if (partials) {
    const float s = block_sum(bits);
    if (threadIdx.x == 0) partials[blockIdx.x] = s;
  }
}

__global__ void z_dequant_kernel(const int16_t *__restrict__ symbols, dcvc_view zh) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)zh.h * zh.w * zh.c;
  if (t >= total) return;
  const int c = (int)(t % zh.c);
  const int64_t p = t / zh.c;
  const int x = (int)(p % zh.w), y = (int)(p / zh.w);
  zh.data[vidx(zh, y, x) + c] = (float)symbols[((size_t)c * zh.h + y) * zh.w + x];
}

}  // namespace

static int quad_check(const dcvc_quad *q, bool need_y) {
  if (!q || !view_ok(&q->params) || q->params.c % 12 || !q->edges || q->step < 0 || q->step > 3)
    return dcvc_fail(DCVC_ERR_ARG, "quad: bad params/step");
  const int C = q->params.c / 3;
  const dcvc_view *vs[3] = {&q->y_hat, &q->y_out, &q->y};
  for (int i = 0; i < (need_y ? 3 : 2); ++i)
    if (!view_ok(vs[i]) || vs[i]->c != C || vs[i]->h != q->params.h || vs[i]->w != q->params.w)
      return dcvc_fail(DCVC_ERR_ARG, "quad: view %d extent mismatch", i);
  if (q->prior.data && (!view_ok(&q->prior) || q->prior.c != 2 * C || q->prior.h != q->params.h ||
                        q->prior.w != q->params.w))
    return dcvc_fail(DCVC_ERR_ARG, "quad: prior extent mismatch");
  if (q->step > 0 && !q->prior.data) return dcvc_fail(DCVC_ERR_ARG, "quad: step>0 needs prior");
  return DCVC_OK;
}

extern "C" size_t dcvc_quad_bits_ws_bytes(int64_t n) { return (size_t)cdiv(n, TPB) * sizeof(float); }

extern "C" int dcvc_quad_step_enc(const dcvc_quad *q, int16_t *symbols, uint8_t *indexes,
                                  float *bits, void *bits_ws, void *stream) {
  int rc = quad_check(q, true);
  if (rc) return rc;
  if ((symbols == nullptr) != (indexes == nullptr) || (bits && !bits_ws))
    return dcvc_fail(DCVC_ERR_ARG, "quad_step_enc: symbols/indexes/bits arguments inconsistent");
  const int64_t n = (int64_t)q->params.h * q->params.w * (q->params.c / 12);
  const int blocks = cdiv(n, TPB);
  cudaStream_t s = (cudaStream_t)stream;
  quad_step_enc_kernel<<<blocks, TPB, 0, s>>>(*q, symbols, indexes, bits ? (float *)bits_ws : nullptr);
  DCVC_LAUNCHED();
  if (bits) {
    reduce_partials_kernel<<<1, TPB, 0, s>>>((const float *)bits_ws, blocks, bits);
    DCVC_LAUNCHED();
  }
  return DCVC_OK;
}

extern "C" int dcvc_quad_step_indexes(const dcvc_quad *q, uint8_t *indexes, void *stream) {
  int rc = quad_check(q, false);
  if (rc) return rc;
  if (!indexes) return dcvc_fail(DCVC_ERR_ARG, "quad_step_indexes: null output");
  const int64_t n = (int64_t)q->params.h * q->params.w * (q->params.c / 12);
  quad_step_indexes_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(*q, indexes);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_quad_step_dec(const dcvc_quad *q, const int16_t *symbols, void *stream) {
  int rc = quad_check(q, false);
  if (rc) return rc;
  if (!symbols) return dcvc_fail(DCVC_ERR_ARG, "quad_step_dec: null symbols");
  const int64_t n = (int64_t)q->params.h * q->params.w * (q->params.c / 12);
  quad_step_dec_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(*q, symbols);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" int dcvc_z_quant(const dcvc_view *z, const dcvc_view *z_hat, int16_t *symbols,
                            const float *fparams, float *bits, void *bits_ws, void *stream) {
  if (!view_ok(z) || !view_ok(z_hat) || z->h != z_hat->h || z->w != z_hat->w || z->c != z_hat->c)
    return dcvc_fail(DCVC_ERR_ARG, "z_quant: bad views");
  if (bits && (!bits_ws || !fparams)) return dcvc_fail(DCVC_ERR_ARG, "z_quant: bits needs ws+params");
  const int64_t n = (int64_t)z->h * z->w * z->c;
  const int blocks = cdiv(n, TPB);
  cudaStream_t s = (cudaStream_t)stream;
  z_quant_kernel<<<blocks, TPB, 0, s>>>(*z, *z_hat, symbols, fparams, bits ? (float *)bits_ws : nullptr);
  DCVC_LAUNCHED();
  if (bits) {
    reduce_partials_kernel<<<1, TPB, 0, s>>>((const float *)bits_ws, blocks, bits);
    DCVC_LAUNCHED();
  }
  return DCVC_OK;
}

extern "C" int dcvc_z_dequant(const int16_t *symbols, const dcvc_view *z_hat, void *stream) {
  if (!symbols || !view_ok(z_hat)) return dcvc_fail(DCVC_ERR_ARG, "z_dequant: bad args");
  const int64_t n = (int64_t)z_hat->h * z_hat->w * z_hat->c;
  z_dequant_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(symbols, *z_hat);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

// ---------------------------------------------------------------------------------------------
// dcvc_quad_step_enc_static: dcvc_quad_step_enc without a caller workspace and without any
// allocation.  The per-block bit partials live in a file-scope device array that every GPU
// loading the library holds until exit (64 KiB).  Every call shares it: a call overwrites the
// first cdiv(n, TPB) entries before reading them, so nothing needs clearing, but two calls that
// overlap on the GPU (different streams) race on it.
#define DCVC_STATIC_MAX_SYMBOLS (1 << 22)  /* positions of ONE step (C/4 * h * w) */
static __device__ __align__(256) float g_bits_partials[DCVC_STATIC_MAX_SYMBOLS / 256];

namespace {
static_assert(TPB == 256, "g_bits_partials is sized for 256-thread blocks");
constexpr int64_t STATIC_MAX_SYMBOLS = DCVC_STATIC_MAX_SYMBOLS;

constexpr int MAX_DEVICES = 64;
float *g_partials_addr[MAX_DEVICES];  // device address of g_bits_partials per device (host table)
std::mutex g_partials_mu;

// In host code `g_bits_partials` is only a stand-in symbol; the device address must be asked for,
// once per device, with that device current.
int static_partials(float **out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return dcvc_fail(DCVC_ERR_CUDA, "quad_step_enc_static: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= MAX_DEVICES) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "quad_step_enc_static: device %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_partials_mu);
  if (!g_partials_addr[dev]) {
    void *p = nullptr;
    e = cudaGetSymbolAddress(&p, g_bits_partials);
    if (e != cudaSuccess || !p)
      return dcvc_fail(DCVC_ERR_CUDA, "quad_step_enc_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    g_partials_addr[dev] = (float *)p;
  }
  *out = g_partials_addr[dev];
  return DCVC_OK;
}
}  // namespace

extern "C" int64_t dcvc_quad_step_enc_static_max(void) { return STATIC_MAX_SYMBOLS; }

extern "C" int dcvc_quad_step_enc_static(const dcvc_quad *q, int16_t *symbols, uint8_t *indexes,
                                         float *bits, void *stream) {
  int rc = quad_check(q, true);
  if (rc) return rc;
  if ((symbols == nullptr) != (indexes == nullptr))
    return dcvc_fail(DCVC_ERR_ARG, "quad_step_enc_static: symbols/indexes arguments inconsistent");
  const int64_t n = (int64_t)q->params.h * q->params.w * (q->params.c / 12);
  if (n > STATIC_MAX_SYMBOLS)
    return dcvc_fail(DCVC_ERR_ARG, "quad_step_enc_static: %lld positions per step exceed the static scratch limit of %lld; "
                     "use dcvc_quad_step_enc with a caller workspace", (long long)n, (long long)STATIC_MAX_SYMBOLS);
  float *partials = nullptr;
  if (bits && (rc = static_partials(&partials))) return rc;
  const int blocks = cdiv(n, TPB);
  cudaStream_t s = (cudaStream_t)stream;
  quad_step_enc_kernel<<<blocks, TPB, 0, s>>>(*q, symbols, indexes, partials);
  DCVC_LAUNCHED();
  if (bits) {
    reduce_partials_kernel<<<1, TPB, 0, s>>>(partials, blocks, bits);
    DCVC_LAUNCHED();
  }
  return DCVC_OK;
}
