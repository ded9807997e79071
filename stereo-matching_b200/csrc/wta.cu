// wta.cu -- winner-takes-all argmin over the disparity axis (HBM-bound streaming reduction).
//
// Replaces DisparityReduceOps::RunWinnersTakeAll (reference src/winners_take_all.cu:89-114).
// Contract = the reference CPU functor (src/winners_take_all.cu:21-40): strict `<` scan
// from d = 0, i.e. lowest d among equal minima, all-+inf -> 0, NaN never wins, NaN at
// d = 0 -> 0.  (The reference CUDA functor's tree reduction breaks ties towards higher d
// and needs power-of-two D; BASELINE.json pins the lowest-d rule.)
//
// Layout: 8 lanes own one pixel; each lane streams float4 chunks d = 32*j + 4*sub .. +3,
// so every warp load instruction covers 4 pixels x 128 contiguous bytes.  The 8 partial
// (value, index) pairs meet in three xor-shuffles.  Algorithmic bytes: 4*D read +
// 4 written per pixel.
#include "stm_common.cuh"

namespace stm {

constexpr int kWtaSentinel = 0x7fffffff;

__device__ __forceinline__ void wta_take(float &best, int &idx, float v, int d) {
  if (v < best) {  // strict: first minimum wins, NaN never does
    best = v;
    idx = d;
  }
}

__device__ __forceinline__ void wta_merge(float &best, int &idx, float ov, int oi) {
  if (ov < best || (ov == best && oi < idx)) {
    best = ov;
    idx = oi;
  }
}

// D % 4 == 0, cost 16-byte aligned.
__global__ void __launch_bounds__(256) wta_vec_kernel(const float *__restrict__ cost,
                                                      int32_t *__restrict__ disp,
                                                      long long num_pixels, int D) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & 7;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long num_warps = (long long)gridDim.x * (blockDim.x >> 5);
  const long long groups = (num_pixels + 3) >> 2;

  for (long long g = warp; g < groups; g += num_warps) {
    const long long p = g * 4 + (lane >> 3);
    const bool live = p < num_pixels;
    const float *src = cost + (live ? p : 0) * D;
    float best = CUDART_INF_F;
    int idx = kWtaSentinel;
    bool nan0 = false;
#pragma unroll 4
    for (int d = sub * 4; d < D; d += 32) {
      const float4 v = ld_stream4(src + d);
      if (d == 0) nan0 = (v.x != v.x);
      wta_take(best, idx, v.x, d);
      wta_take(best, idx, v.y, d + 1);
      wta_take(best, idx, v.z, d + 2);
      wta_take(best, idx, v.w, d + 3);
    }
#pragma unroll
    for (int m = 1; m < 8; m <<= 1) {
      const float ov = __shfl_xor_sync(FULL, best, m);
      const int oi = __shfl_xor_sync(FULL, idx, m);
      wta_merge(best, idx, ov, oi);
    }
    if (sub == 0 && live) disp[p] = (nan0 || idx == kWtaSentinel) ? 0 : idx;
  }
}

// Any D: one warp per pixel, scalar loads.
__global__ void __launch_bounds__(256) wta_generic_kernel(const float *__restrict__ cost,
                                                          int32_t *__restrict__ disp,
                                                          long long num_pixels, int D) {
  const int lane = threadIdx.x & 31;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long num_warps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long p = warp; p < num_pixels; p += num_warps) {
    const float *src = cost + p * D;
    float best = CUDART_INF_F;
    int idx = kWtaSentinel;
    bool nan0 = false;
    for (int d = lane; d < D; d += 32) {
      const float v = __ldcs(src + d);
      if (d == 0) nan0 = (v != v);
      wta_take(best, idx, v, d);
    }
    nan0 = __shfl_sync(FULL, nan0 ? 1 : 0, 0) != 0;
#pragma unroll
    for (int m = 1; m < 32; m <<= 1) {
      const float ov = __shfl_xor_sync(FULL, best, m);
      const int oi = __shfl_xor_sync(FULL, idx, m);
      wta_merge(best, idx, ov, oi);
    }
    if (lane == 0) disp[p] = (nan0 || idx == kWtaSentinel) ? 0 : idx;
  }
}

int launch_wta(const float *cost, int32_t *disp, int H, int W, int D, cudaStream_t stream) {
  const long long P = (long long)H * W;
  const int warps_per_block = 8;
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(cost) & 15) == 0);
  const long long units = vec ? (P + 3) / 4 : P;
  long long blocks = (units + warps_per_block - 1) / warps_per_block;
  const long long cap = (long long)num_sms() * 8 * 4;  // 8 resident CTAs/SM, 4 rounds
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (vec)
    wta_vec_kernel<<<(unsigned)blocks, 256, 0, stream>>>(cost, disp, P, D);
  else
    wta_generic_kernel<<<(unsigned)blocks, 256, 0, stream>>>(cost, disp, P, D);
  return check_launch("wta");
}

}  // namespace stm
