// abi.cu -- the extern "C" surface of libstm_b200.so (include/stm_abi.h).
#include <cstring>

#include "stm_common.cuh"

namespace stm {

std::atomic<unsigned long long> g_launches{0};

char *last_error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

// kernel launchers (one translation unit each)
int launch_wta(const float *, int32_t *, int, int, int, cudaStream_t);
int launch_ssd(const float *, const float *, float *, int, int, int, int, cudaStream_t);
int launch_birchfield(const float *, const float *, float *, int, int, int, cudaStream_t);
int launch_birchfield_wta(const float *, const float *, int32_t *, int, int, int, cudaStream_t);
int launch_ssd_wta(const float *, const float *, int32_t *, unsigned long long *, int, int, int, int, cudaStream_t);
size_t sgm_workspace_bytes(int, int, int);
size_t sgm_workspace_bytes_flags(int, int, int, unsigned);
int sgm_sweep_aborted(unsigned *);  // sgm_vsweep.cu
int sgm_pass_timing(bool);
int sgm_last_pass_ms(float *, int *, int);
int launch_sgm(const float *, const float *, float *, int, int, int, float, float, unsigned, void *,
               size_t, cudaStream_t, int32_t *disp = nullptr);
int launch_dp(const float *, int8_t *, float *, int32_t *, int, int, int, unsigned, cudaStream_t);

static int require_device() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(STM_ECUDA, "no usable CUDA device (%s); libstm_b200 has no CPU fallback",
                cudaGetErrorString(e));
  }
  return STM_OK;
}

// Per-thread device scratch for the host-buffer entry point.
struct HostPipelineScratch {
  int device = -1;
  size_t cap[8] = {0};
  void *ptr[8] = {nullptr};
  int ensure(int slot, size_t bytes) {
    if (cap[slot] >= bytes) return STM_OK;
    if (ptr[slot]) cudaFree(ptr[slot]);
    ptr[slot] = nullptr;
    cap[slot] = 0;
    STM_CUDA_TRY(cudaMalloc(&ptr[slot], bytes));
    cap[slot] = bytes;
    return STM_OK;
  }
  void reset() {
    for (int i = 0; i < 8; ++i) {
      if (ptr[i]) cudaFree(ptr[i]);
      ptr[i] = nullptr;
      cap[i] = 0;
    }
  }
  // thread exit: give the scratch back (errors ignored: at process exit the context may be gone)
  ~HostPipelineScratch() {
    if (device < 0) return;
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess) { cudaGetLastError(); return; }
    if (cur != device && cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return; }
    reset();
    if (cur != device) cudaSetDevice(cur);
    cudaGetLastError();
  }
};

}  // namespace stm

using namespace stm;

#define STM_REQUIRE(cond, ...)                         \
  do {                                                 \
    if (!(cond)) return fail(STM_EINVAL, __VA_ARGS__); \
  } while (0)

extern "C" {

const char *stm_version(void) { return "stm_b200 0.1 sm_100a"; }
const char *stm_last_error(void) { return last_error_buffer(); }
unsigned long long stm_launch_count(void) { return g_launches.load(); }

int stm_ssd_f32(const float *left, const float *right, float *cost, int H, int W, int D,
                int kernel_size, void *stream) {
  int rc = check_shape("ssd", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(left && right && cost, "ssd: null pointer");
  STM_REQUIRE(kernel_size >= 0, "ssd: kernel_size must be >= 0");
  if ((rc = require_device())) return rc;
  return launch_ssd(left, right, cost, H, W, D, kernel_size, (cudaStream_t)stream);
}

int stm_birchfield_f32(const float *left, const float *right, float *cost, int H, int W, int D,
                       void *stream) {
  int rc = check_shape("birchfield", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(left && right && cost, "birchfield: null pointer");
  if ((rc = require_device())) return rc;
  return launch_birchfield(left, right, cost, H, W, D, (cudaStream_t)stream);
}

int stm_birchfield_wta_f32(const float *left, const float *right, int32_t *disparity, int H, int W, int D,
                           void *stream) {
  int rc = check_shape("birchfield_wta", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(left && right && disparity, "birchfield_wta: null pointer");
  if ((rc = require_device())) return rc;
  rc = launch_birchfield_wta(left, right, disparity, H, W, D, (cudaStream_t)stream);
  if (rc == 1)
    return fail(STM_EUNSUPPORTED, "birchfield_wta: needs D %% 32 == 0 (64..256) and 8 <= W <= ~6000 (got W=%d D=%d); "
                "use stm_birchfield_f32 + stm_wta_f32", W, D);
  return rc;
}

size_t stm_ssd_wta_scratch_bytes(int H, int W) {
  if (H <= 0 || W <= 0) return 0;
  return (size_t)H * W * sizeof(unsigned long long);
}

int stm_ssd_wta_f32(const float *left, const float *right, int32_t *disparity, int H, int W, int D,
                    int kernel_size, void *scratch, size_t scratch_bytes, void *stream) {
  int rc = check_shape("ssd_wta", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(left && right && disparity && scratch, "ssd_wta: null pointer");
  STM_REQUIRE(scratch_bytes >= stm_ssd_wta_scratch_bytes(H, W), "ssd_wta: scratch too small");
  STM_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 7) == 0, "ssd_wta: scratch must be 8-byte aligned");
  if ((rc = require_device())) return rc;
  rc = launch_ssd_wta(left, right, disparity, static_cast<unsigned long long *>(scratch), H, W, D, kernel_size,
                      (cudaStream_t)stream);
  if (rc == 1)
    return fail(STM_EUNSUPPORTED, "ssd_wta: needs 1 <= kernel_size <= 10 and D %% 4 == 0 (got k=%d D=%d); "
                "use stm_ssd_f32 + stm_wta_f32", kernel_size, D);
  return rc;
}

size_t stm_sgm_workspace_bytes(int H, int W, int D) {
  if (H <= 0 || W <= 0 || D <= 0) return 0;
  return sgm_workspace_bytes(H, W, D);
}

size_t stm_sgm_workspace_bytes_flags(int H, int W, int D, unsigned flags) {
  if (H <= 0 || W <= 0 || D <= 0) return 0;
  return sgm_workspace_bytes_flags(H, W, D, flags);
}

int stm_sgm_f32(const float *cost, const float *left, float *out, int H, int W, int D,
                float penalty1, float penalty2, unsigned flags, void *workspace,
                size_t workspace_bytes, void *stream) {
  int rc = check_shape("sgm", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(cost && left && out, "sgm: null pointer");
  STM_REQUIRE(cost != out, "sgm: output must not alias the cost volume");
  STM_REQUIRE((flags & ~(STM_SGM_MIN_ALL_D | STM_SGM_OVERWRITE | STM_SGM_FUSED_SWEEPS | STM_SGM_SEQUENTIAL | STM_SGM_OVERLAP)) == 0, "sgm: unknown flags 0x%x",
              flags);
  if ((rc = require_device())) return rc;
  return launch_sgm(cost, left, out, H, W, D, penalty1, penalty2, flags, workspace,
                    workspace_bytes, (cudaStream_t)stream);
}

int stm_sgm_wta_f32(const float *cost, const float *left, float *out, int32_t *disparity, int H,
                    int W, int D, float penalty1, float penalty2, unsigned flags, void *workspace,
                    size_t workspace_bytes, void *stream) {
  int rc = check_shape("sgm_wta", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(cost && left && out && disparity, "sgm_wta: null pointer");
  STM_REQUIRE(cost != out, "sgm_wta: output must not alias the cost volume");
  STM_REQUIRE((flags & ~(STM_SGM_MIN_ALL_D | STM_SGM_OVERWRITE | STM_SGM_DISPARITY_ONLY | STM_SGM_SEQUENTIAL |
                         STM_SGM_FUSED_SWEEPS | STM_SGM_OVERLAP)) == 0,
              "sgm_wta: unknown flags 0x%x", flags);
  if ((rc = require_device())) return rc;
  return launch_sgm(cost, left, out, H, W, D, penalty1, penalty2, flags, workspace, workspace_bytes,
                    (cudaStream_t)stream, disparity);
}

int stm_sgm_pass_timing(int enable) { return sgm_pass_timing(enable != 0); }

int stm_sgm_sweep_aborted(unsigned *info3) {
  const int rc = require_device();
  if (rc) return rc;
  return sgm_sweep_aborted(info3);
}

int stm_sgm_last_pass_ms(float *ms, int *directions, int capacity) {
  STM_REQUIRE(ms && directions, "sgm_last_pass_ms: null pointer");
  return sgm_last_pass_ms(ms, directions, capacity);
}

int stm_wta_f32(const float *cost, int32_t *disparity, int H, int W, int D, void *stream) {
  int rc = check_shape("wta", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(cost && disparity, "wta: null pointer");
  if ((rc = require_device())) return rc;
  return launch_wta(cost, disparity, H, W, D, (cudaStream_t)stream);
}

int stm_dp_f32(const float *cost, int8_t *path, float *row_final, int32_t *disparity, int H, int W,
               int D, unsigned flags, void *stream) {
  int rc = check_shape("dp", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(cost && path && row_final && disparity, "dp: null pointer");
  STM_REQUIRE((flags & ~STM_DP_CPU_EXACT) == 0, "dp: unknown flags 0x%x", flags);
  if ((rc = require_device())) return rc;
  return launch_dp(cost, path, row_final, disparity, H, W, D, flags, (cudaStream_t)stream);
}

int stm_pipeline_host_f32(const float *left, const float *right, int32_t *disparity, int H, int W,
                          int D, int cost_kind, int kernel_size, int use_sgm, float penalty1,
                          float penalty2, unsigned sgm_flags, int reducer, unsigned dp_flags) {
  int rc = check_shape("pipeline", H, W, D);
  if (rc) return rc;
  STM_REQUIRE(left && right && disparity, "pipeline: null pointer");
  STM_REQUIRE(cost_kind == 0 || cost_kind == 1, "pipeline: cost_kind must be 0 (SSD) or 1 (Birchfield)");
  STM_REQUIRE(reducer == 0 || reducer == 1, "pipeline: reducer must be 0 (WTA) or 1 (DP)");
  if ((rc = require_device())) return rc;

  static thread_local HostPipelineScratch s;
  int dev = 0;
  STM_CUDA_TRY(cudaGetDevice(&dev));
  if (s.device != dev) {
    s.reset();
    s.device = dev;
  }
  const size_t P = (size_t)H * W, N = P * D;
  enum { kL, kR, kC, kS, kWs, kDisp, kPath, kRowFinal };
  if ((rc = s.ensure(kL, P * 4)) || (rc = s.ensure(kR, P * 4)) || (rc = s.ensure(kC, N * 4)) ||
      (rc = s.ensure(kDisp, P * 4)))
    return rc;
  float *dl = (float *)s.ptr[kL], *dr = (float *)s.ptr[kR], *dc = (float *)s.ptr[kC];
  cudaStream_t st = nullptr;  // legacy default stream, like the reference
  STM_CUDA_TRY(cudaMemcpyAsync(dl, left, P * 4, cudaMemcpyHostToDevice, st));
  STM_CUDA_TRY(cudaMemcpyAsync(dr, right, P * 4, cudaMemcpyHostToDevice, st));
  rc = cost_kind == 0 ? launch_ssd(dl, dr, dc, H, W, D, kernel_size, st)
                      : launch_birchfield(dl, dr, dc, H, W, D, st);
  if (rc) return rc;
  float *vol = dc;
  if (use_sgm) {
    const size_t wsb = sgm_workspace_bytes_flags(H, W, D, sgm_flags);
    if ((rc = s.ensure(kS, N * 4)) || (rc = s.ensure(kWs, wsb ? wsb : 16))) return rc;
    const bool fuse_wta = reducer == 0;
    rc = launch_sgm(dc, dl, (float *)s.ptr[kS], H, W, D, penalty1, penalty2,
                    sgm_flags | STM_SGM_OVERWRITE | (fuse_wta ? STM_SGM_DISPARITY_ONLY : 0u),
                    s.ptr[kWs], s.cap[kWs], st, fuse_wta ? (int32_t *)s.ptr[kDisp] : nullptr);
    if (rc) return rc;
    vol = (float *)s.ptr[kS];
    if (fuse_wta) reducer = -1;  // the last aggregation pass already wrote the disparity map
  }
  if (reducer < 0) {
  } else if (reducer == 0) {
    rc = launch_wta(vol, (int32_t *)s.ptr[kDisp], H, W, D, st);
  } else {
    if ((rc = s.ensure(kPath, N)) || (rc = s.ensure(kRowFinal, (size_t)H * D * 4))) return rc;
    if (dp_flags & STM_DP_CPU_EXACT) STM_CUDA_TRY(cudaMemsetAsync(s.ptr[kPath], 0, N, st));
    rc = launch_dp(vol, (int8_t *)s.ptr[kPath], (float *)s.ptr[kRowFinal], (int32_t *)s.ptr[kDisp],
                   H, W, D, dp_flags, st);
  }
  if (rc) return rc;
  STM_CUDA_TRY(cudaMemcpyAsync(disparity, s.ptr[kDisp], P * 4, cudaMemcpyDeviceToHost, st));
  STM_CUDA_TRY(cudaStreamSynchronize(st));
  return STM_OK;
}

}  // extern "C"
