// common.cuh -- shared helpers of libdcvc_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <atomic>

#include "../../include/dcvc_b200.h"

extern thread_local char g_dcvc_err[512];
extern std::atomic<int64_t> g_dcvc_launches;

static inline int dcvc_fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_dcvc_err, sizeof(g_dcvc_err), fmt, ap);
  va_end(ap);
  return code;
}

#define DCVC_CUDA(expr)                                                                   \
  do {                                                                                    \
    cudaError_t e_ = (expr);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return dcvc_fail(DCVC_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #expr,          \
                       cudaGetErrorString(e_));                                           \
  } while (0)

// call after every kernel launch
#define DCVC_LAUNCHED()                                                                   \
  do {                                                                                    \
    g_dcvc_launches.fetch_add(1, std::memory_order_relaxed);                              \
    cudaError_t e_ = cudaGetLastError();                                                  \
    if (e_ != cudaSuccess)                                                                \
      return dcvc_fail(DCVC_ERR_CUDA, "%s:%d launch: %s", __FILE__, __LINE__,             \
                       cudaGetErrorString(e_));                                           \
  } while (0)

static inline bool view_ok(const dcvc_view *v) {
  return v && v->data && v->h > 0 && v->w > 0 && v->c > 0 && v->coff >= 0 &&
         v->cstride >= v->coff + v->c;
}

// every pixel's channel slice starts 16-byte aligned and holds whole float4 groups
static inline bool view_al16(const dcvc_view *v) {
  return ((v->cstride | v->coff | v->c) & 3) == 0 && (reinterpret_cast<uintptr_t>(v->data) & 15) == 0;
}

__device__ __forceinline__ float lrelu(float v, float slope) { return v >= 0.f ? v : v * slope; }

__device__ __forceinline__ size_t vidx(const dcvc_view &v, int y, int x) {
  return ((size_t)y * v.w + x) * v.cstride + v.coff;
}

static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
