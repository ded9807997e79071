// stm_common.cuh -- shared host/device helpers of libstm_b200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/stm_abi.h"

namespace stm {

// ---- error reporting (thread-local message, C status codes) -------------------------
char *last_error_buffer();            // 512 bytes, thread local (abi.cu)
extern std::atomic<unsigned long long> g_launches;

inline int fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

// Checks the launch that was just enqueued (no device sync; the reference's
// cudaDeviceSynchronize after every op, include/stereomatch/kernel.hpp:50,59, is dropped).
inline int check_launch(const char *what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return fail(STM_ECUDA, "Cuda check error %s: %s", what, cudaGetErrorString(e));
  return STM_OK;
}

#define STM_CUDA_TRY(expr)                                                              \
  do {                                                                                  \
    cudaError_t e__ = (expr);                                                           \
    if (e__ != cudaSuccess)                                                             \
      return ::stm::fail(STM_ECUDA, "Cuda call error %s: %s", #expr,                    \
                         cudaGetErrorString(e__));                                      \
  } while (0)

inline int check_shape(const char *op, int H, int W, int D) {
  if (H <= 0 || W <= 0 || D <= 0)
    return fail(STM_EINVAL, "%s: invalid shape H=%d W=%d D=%d", op, H, W, D);
  if ((long long)H * W >= (1ll << 31))
    return fail(STM_EINVAL, "%s: H*W must be < 2^31", op);
  return STM_OK;
}

// Everything cached from a device query or set by cudaFuncSetAttribute is PER DEVICE: one process
// may drive several GPUs from several host threads.  Caches are arrays indexed by the device
// ordinal; slots hold value + 1 so that zero-initialised storage means "not yet queried".  Two
// threads racing on a slot compute the same value (plain int stores; benign).
constexpr int kMaxDevices = 64;
inline int device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return -1;
  return dev < kMaxDevices ? dev : -1;  // -1: no slot, the value is recomputed on every call
}
struct PerDeviceInt {
  int slot[kMaxDevices] = {0};
  bool get(int dev, int *v) const {
    if (dev < 0 || slot[dev] == 0) return false;
    *v = slot[dev] - 1;
    return true;
  }
  void set(int dev, int v) {
    if (dev >= 0) slot[dev] = v + 1;
  }
};

inline int num_sms() {
  static PerDeviceInt cache;
  const int dev = device_slot();
  int n = 0;
  if (!cache.get(dev, &n)) {
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, cur) != cudaSuccess || n <= 0)
      n = 148;
    cache.set(dev, n);
  }
  return n;
}

#ifdef __CUDACC__
constexpr unsigned FULL = 0xffffffffu;

// Order-preserving float -> uint key (total order on non-NaN floats), and back.
__device__ __forceinline__ unsigned f2key(float v) {
  unsigned u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(unsigned k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// Warp-wide float minimum in ONE redux.sync (integer min on order-preserving keys).
__device__ __forceinline__ float warp_min(float v) {
  return key2f(__reduce_min_sync(FULL, f2key(v)));
}

// Streaming (read-once) loads / write-once stores: keep them out of L1.
__device__ __forceinline__ float4 ld_stream4(const float *p) {
  return __ldcs(reinterpret_cast<const float4 *>(p));
}
__device__ __forceinline__ float2 ld_stream2(const float *p) {
  return __ldcs(reinterpret_cast<const float2 *>(p));
}
__device__ __forceinline__ void st_stream4(float *p, float4 v) {
  __stcs(reinterpret_cast<float4 *>(p), v);
}
__device__ __forceinline__ void st_stream2(float *p, float2 v) {
  __stcs(reinterpret_cast<float2 *>(p), v);
}
#endif

}  // namespace stm
