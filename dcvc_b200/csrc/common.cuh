// common.cuh -- shared helpers of libdcvc_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <atomic>
#include <mutex>

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


// ---- per-device one-time set-up (cudaFuncSetAttribute is per device; one process may drive
// several GPUs, reference test_video.py:282-294 maps workers to cuda:i).  Thread-safe.
constexpr int DCVC_MAX_DEVICES = 64;
enum { DCVC_ONCE_CONV_TC = 0, DCVC_ONCE_FFN_TC = 1, DCVC_ONCE_RANS_DEC = 2, DCVC_ONCE_FFN128 = 3,
       DCVC_ONCE_DCB = 4, DCVC_ONCE_RANS_DEC2 = 5, DCVC_ONCE_FFN2_TC = 6 };
extern std::atomic<uint32_t> g_dcvc_once[DCVC_MAX_DEVICES];
extern int g_dcvc_num_sms[DCVC_MAX_DEVICES];
extern std::mutex g_dcvc_once_mu;

// current device and its SM count (cached per device)
static inline int dcvc_device_info(int *dev_out, int *num_sms) {
  int dev = 0;
  DCVC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= DCVC_MAX_DEVICES) return dcvc_fail(DCVC_ERR_UNSUPPORTED, "device %d out of range", dev);
  if (!g_dcvc_num_sms[dev]) {
    std::lock_guard<std::mutex> lk(g_dcvc_once_mu);
    int n = 0;
    DCVC_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    g_dcvc_num_sms[dev] = n;
  }
  if (dev_out) *dev_out = dev;
  if (num_sms) *num_sms = g_dcvc_num_sms[dev];
  return DCVC_OK;
}

// runs f() (returning a cudaError_t) once per (device, slot)
template <class F>
static inline int dcvc_once_per_device(int slot, F &&f) {
  int dev = 0;
  int rc = dcvc_device_info(&dev, nullptr);
  if (rc) return rc;
  const uint32_t bit = 1u << slot;
  if (g_dcvc_once[dev].load(std::memory_order_acquire) & bit) return DCVC_OK;
  std::lock_guard<std::mutex> lk(g_dcvc_once_mu);
  if (g_dcvc_once[dev].load(std::memory_order_relaxed) & bit) return DCVC_OK;
  DCVC_CUDA(f());
  g_dcvc_once[dev].fetch_or(bit, std::memory_order_release);
  return DCVC_OK;
}
