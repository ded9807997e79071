// lanevec.cuh -- a pixel's D-vector spread over the 32 lanes of a warp (see sgm_paths.cuh).
#pragma once

#include "stm_common.cuh"

namespace stm {

template <int NV, int VW>
struct LaneVec {
  float v[NV][VW];
};

template <int NV, int VW>
__device__ __forceinline__ int lv_d(int lane, int k, int j) {
  return k * 32 * VW + lane * VW + j;
}

// Load one pixel's vector.  `stream` = read-once data (bypass L1 allocation).
template <int NV, int VW, bool STREAM>
__device__ __forceinline__ void lv_load(LaneVec<NV, VW> &r, const float *__restrict__ p,
                                        int lane, int D) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const float *q = p + k * 32 * VW + lane * VW;
    if constexpr (VW == 4) {
      const float4 t = STREAM ? ld_stream4(q) : *reinterpret_cast<const float4 *>(q);
      r.v[k][0] = t.x; r.v[k][1] = t.y; r.v[k][2] = t.z; r.v[k][3] = t.w;
    } else if constexpr (VW == 2) {
      const float2 t = STREAM ? ld_stream2(q) : *reinterpret_cast<const float2 *>(q);
      r.v[k][0] = t.x; r.v[k][1] = t.y;
    } else {
      const int d = k * 32 + lane;
      r.v[k][0] = d < D ? (STREAM ? __ldcs(q) : *q) : CUDART_INF_F;
    }
  }
}

template <int NV, int VW, bool STREAM>
__device__ __forceinline__ void lv_store(const LaneVec<NV, VW> &r, float *__restrict__ p,
                                         int lane, int D) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    float *q = p + k * 32 * VW + lane * VW;
    if constexpr (VW == 4) {
      const float4 t = make_float4(r.v[k][0], r.v[k][1], r.v[k][2], r.v[k][3]);
      if (STREAM) st_stream4(q, t); else *reinterpret_cast<float4 *>(q) = t;
    } else if constexpr (VW == 2) {
      const float2 t = make_float2(r.v[k][0], r.v[k][1]);
      if (STREAM) st_stream2(q, t); else *reinterpret_cast<float2 *>(q) = t;
    } else {
      const int d = k * 32 + lane;
      if (d < D) { if (STREAM) __stcs(q, r.v[k][0]); else *q = r.v[k][0]; }
    }
  }
}

// lo[k][j] = value at d-1, hi[k][j] = value at d+1 (+inf outside [0, Dpad)).
template <int NV, int VW>
__device__ __forceinline__ void lv_neighbours(const LaneVec<NV, VW> &a, LaneVec<NV, VW> &lo,
                                              LaneVec<NV, VW> &hi, int lane) {
  float from_left[NV], from_right[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    from_left[k] = __shfl_sync(FULL, a.v[k][VW - 1], (lane + 31) & 31);  // lane-1 (rotating)
    from_right[k] = __shfl_sync(FULL, a.v[k][0], (lane + 1) & 31);       // lane+1 (rotating)
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
#pragma unroll
    for (int j = 0; j < VW; ++j) {
      if (j > 0) lo.v[k][j] = a.v[k][j - 1];
      else lo.v[k][j] = lane > 0 ? from_left[k] : (k > 0 ? from_left[k > 0 ? k - 1 : 0] : CUDART_INF_F);
      if (j < VW - 1) hi.v[k][j] = a.v[k][j + 1];
      else hi.v[k][j] = lane < 31 ? from_right[k]
                                  : (k < NV - 1 ? from_right[k < NV - 1 ? k + 1 : k] : CUDART_INF_F);
    }
  }
}

// Minimum over d in [0, limit) of the vector (limit = D or D-1).
template <int NV, int VW>
__device__ __forceinline__ float lv_min(const LaneVec<NV, VW> &a, int lane, int limit) {
  float m = CUDART_INF_F;
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j) {
      const float x = a.v[k][j];
      if (lv_d<NV, VW>(lane, k, j) < limit) m = fminf(m, x);
    }
  return warp_min(m);
}

}  // namespace stm
