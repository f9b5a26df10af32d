// frame_io.cu -- frame ingest / egress on the device: raw 8-bit planes
// become the padded 4:4:4 fp32 tensor the models read, and the decoded tensor becomes 4:2:0 (or
// RGB) planes plus per-plane sums of squared error, without leaving the GPU.
//
// Replaces, per frame, in the reference harness: np_image_to_tensor / .astype(float32)/255
// (DCVC-DC/src/utils/video_reader.py:45-161), ycbcr420_to_444(order=0) = scipy.ndimage.zoom nearest
// (src/transforms/functional.py:61-71), F.pad(mode="replicate") to a multiple of 16
// (test_video.py:130-135), ycbcr444_to_420 2x2 chroma mean + clip (functional.py:74-92), and the
// squared-error part of calc_psnr in float64 (src/utils/metrics.py:81-92, test_video.py:169-195),
// which the reference does on the host after a device -> numpy copy of every decoded frame.
// HBM-bound, tiny: a 1080p frame is 3.1 MB in, 25 MB out.
#include "common.cuh"

namespace {

constexpr int TPB = 256;

// dst(y, x, c) = plane_c(min(y, h-1), min(x, w-1)) / 255 ; chroma planes of 4:2:0 are read at
// (y >> 1, x >> 1): zoom(order=0) by exactly 2 repeats every sample 2x2
__global__ void u8_to_view_kernel(const uint8_t *__restrict__ p, int fmt, int h, int w, dcvc_view dst) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)dst.h * dst.w) return;
  const int x = (int)(t % dst.w), y = (int)(t / dst.w);
  const int sx = min(x, w - 1), sy = min(y, h - 1);
  float v[3];
  if (fmt == DCVC_FMT_RGB444) {
    const size_t plane = (size_t)h * w, o = (size_t)sy * w + sx;
    v[0] = (float)p[o] / 255.0f;
    v[1] = (float)p[plane + o] / 255.0f;
    v[2] = (float)p[2 * plane + o] / 255.0f;
  } else {
    const size_t plane = (size_t)h * w, cw = (size_t)(w >> 1), cplane = (size_t)(h >> 1) * cw;
    const size_t oc = (size_t)(sy >> 1) * cw + (sx >> 1);
    v[0] = (float)p[(size_t)sy * w + sx] / 255.0f;
    v[1] = (float)p[plane + oc] / 255.0f;
    v[2] = (float)p[plane + cplane + oc] / 255.0f;
  }
  float *d = dst.data + vidx(dst, y, x);
  d[0] = v[0]; d[1] = v[1]; d[2] = v[2];
}

__device__ __forceinline__ float clip01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

// one thread per 2x2 block of luma positions; partial sums in double, block-reduced in a fixed order
__global__ void plane_sse_kernel(dcvc_view rec, int fmt, int h, int w, const uint8_t *__restrict__ src,
                                 uint8_t *__restrict__ out, double *__restrict__ partials) {
  const int bw = (w + 1) >> 1, bh = (h + 1) >> 1;      // odd sizes: RGB only (4:2:0 needs even)
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double acc[3] = {0.0, 0.0, 0.0};
  if (t < (int64_t)bw * bh) {
    const int bx = (int)(t % bw), by = (int)(t / bw);
    const size_t plane = (size_t)h * w;
    float r[2][2][3];
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const float *q = rec.data + vidx(rec, min(2 * by + dy, h - 1), min(2 * bx + dx, w - 1));
        r[dy][dx][0] = q[0]; r[dy][dx][1] = q[1]; r[dy][dx][2] = q[2];
      }
    if (fmt == DCVC_FMT_RGB444) {
#pragma unroll
      for (int dy = 0; dy < 2; ++dy)
#pragma unroll
        for (int dx = 0; dx < 2; ++dx) {
          if (2 * by + dy >= h || 2 * bx + dx >= w) continue;
          const size_t o = (size_t)(2 * by + dy) * w + 2 * bx + dx;
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const float v = clip01(r[dy][dx][c]);
            const double d = (double)v - (double)((float)src[c * plane + o] / 255.0f);
            acc[c] += d * d;
            if (out) out[c * plane + o] = (uint8_t)rintf(v * 255.0f);
          }
        }
    } else {
      const size_t cw = (size_t)bw, cplane = (size_t)bh * cw, oc = (size_t)by * cw + bx;
#pragma unroll
      for (int dy = 0; dy < 2; ++dy)
#pragma unroll
        for (int dx = 0; dx < 2; ++dx) {
          const size_t o = (size_t)(2 * by + dy) * w + 2 * bx + dx;
          const float v = clip01(r[dy][dx][0]);
          const double d = (double)v - (double)((float)src[o] / 255.0f);
          acc[0] += d * d;
          if (out) out[o] = (uint8_t)rintf(v * 255.0f);
        }
#pragma unroll
      for (int c = 1; c < 3; ++c) {
        // np.mean over the 2x2 block in float32, then clip (functional.py:84-90)
        const float m = clip01(((r[0][0][c] + r[0][1][c]) + (r[1][0][c] + r[1][1][c])) * 0.25f);
        const size_t o = plane + (c - 1) * cplane + oc;
        const double d = (double)m - (double)((float)src[o] / 255.0f);
        acc[c] += d * d;
        if (out) out[o] = (uint8_t)rintf(m * 255.0f);
      }
    }
  }
  __shared__ double sm[3][TPB / 32];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    double v = acc[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sm[c][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double s = 0.0;
    for (int i = 0; i < TPB / 32; ++i) s += sm[threadIdx.x][i];
    partials[(size_t)blockIdx.x * 3 + threadIdx.x] = s;
  }
}

__global__ void sse_final_kernel(const double *__restrict__ partials, int nblocks, double *__restrict__ sse3) {
  __shared__ double sm[TPB];
  for (int c = 0; c < 3; ++c) {
    double acc = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += TPB) acc += partials[(size_t)i * 3 + c];
    sm[threadIdx.x] = acc;
    __syncthreads();
    for (int o = TPB / 2; o > 0; o >>= 1) {
      if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) sse3[c] = sm[0];
    __syncthreads();
  }
}

bool fmt_ok(int fmt, int h, int w) {
  if (fmt != DCVC_FMT_RGB444 && fmt != DCVC_FMT_YUV420) return false;
  return h > 0 && w > 0 && (fmt == DCVC_FMT_RGB444 || ((h & 1) == 0 && (w & 1) == 0));
}

}  // namespace

extern "C" size_t dcvc_frame_bytes(int fmt, int h, int w) {
  if (!fmt_ok(fmt, h, w)) return 0;
  return fmt == DCVC_FMT_RGB444 ? (size_t)3 * h * w : (size_t)h * w * 3 / 2;
}

extern "C" int dcvc_frame_u8_to_view(const uint8_t *planes, int fmt, int h, int w, const dcvc_view *dst,
                                     void *stream) {
  if (!planes || !fmt_ok(fmt, h, w) || !view_ok(dst) || dst->c != 3 || dst->h < h || dst->w < w)
    return dcvc_fail(DCVC_ERR_ARG, "frame_u8_to_view: bad arguments (4:2:0 needs even h, w; 3-channel view at least h x w)");
  const int64_t n = (int64_t)dst->h * dst->w;
  u8_to_view_kernel<<<cdiv(n, TPB), TPB, 0, (cudaStream_t)stream>>>(planes, fmt, h, w, *dst);
  DCVC_LAUNCHED();
  return DCVC_OK;
}

extern "C" size_t dcvc_plane_sse_ws_bytes(int h, int w) {
  return (size_t)cdiv((int64_t)((h + 1) >> 1) * ((w + 1) >> 1), TPB) * 3 * sizeof(double);
}

extern "C" int dcvc_view_plane_sse(const dcvc_view *rec, int fmt, int h, int w, const uint8_t *src_planes,
                                   double *sse3, uint8_t *rec_planes_out, void *ws, void *stream) {
  if (!view_ok(rec) || rec->c != 3 || rec->h < h || rec->w < w || !fmt_ok(fmt, h, w) || !src_planes ||
      !sse3 || !ws)
    return dcvc_fail(DCVC_ERR_ARG, "view_plane_sse: bad arguments");
  const int blocks = cdiv((int64_t)((h + 1) >> 1) * ((w + 1) >> 1), TPB);
  cudaStream_t s = (cudaStream_t)stream;
  plane_sse_kernel<<<blocks, TPB, 0, s>>>(*rec, fmt, h, w, src_planes, rec_planes_out, (double *)ws);
  DCVC_LAUNCHED();
  sse_final_kernel<<<1, TPB, 0, s>>>((const double *)ws, blocks, sse3);
  DCVC_LAUNCHED();
  return DCVC_OK;
}
