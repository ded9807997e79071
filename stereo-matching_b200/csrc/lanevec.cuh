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

// Per-lane constants of the lane<->disparity map, computed once per kernel.
struct LaneCtx {
  int src_lo, src_hi;  // rotating shuffle sources: lane-1, lane+1
  bool first, last;    // lane 0 / lane 31
  bool excl_last;      // this lane's last slot [NV-1][VW-1] holds a disparity >= min limit
};

template <int NV, int VW>
__device__ __forceinline__ LaneCtx lv_ctx(int lane, int min_limit) {
  LaneCtx c;
  c.src_lo = (lane + 31) & 31;
  c.src_hi = (lane + 1) & 31;
  c.first = lane == 0;
  c.last = lane == 31;
  // The only disparity ever excluded from the minimum is D-1 (and the +inf padding above
  // it), and it always sits in the last slot of its lane (see sgm_paths.cuh).
  c.excl_last = lv_d<NV, VW>(lane, NV - 1, VW - 1) >= min_limit;
  return c;
}

// Warp-wide float minimum: one redux.sync.min.f32 (CREDUX.MIN.F32 on sm_100a).
__device__ __forceinline__ float warp_min_f32(float v) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}

// lo[k][j] = value at d-1, hi[k][j] = value at d+1 (+inf outside [0, Dpad)).
template <int NV, int VW>
__device__ __forceinline__ void lv_neighbours(const LaneVec<NV, VW> &a, LaneVec<NV, VW> &lo,
                                              LaneVec<NV, VW> &hi, const LaneCtx &cx) {
  float from_left[NV], from_right[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    from_left[k] = __shfl_sync(FULL, a.v[k][VW - 1], cx.src_lo);
    from_right[k] = __shfl_sync(FULL, a.v[k][0], cx.src_hi);
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
#pragma unroll
    for (int j = 0; j < VW; ++j) {
      if (j > 0) lo.v[k][j] = a.v[k][j - 1];
      else lo.v[k][j] = cx.first ? (k > 0 ? from_left[k > 0 ? k - 1 : 0] : CUDART_INF_F) : from_left[k];
      if (j < VW - 1) hi.v[k][j] = a.v[k][j + 1];
      else hi.v[k][j] = cx.last ? (k < NV - 1 ? from_right[k < NV - 1 ? k + 1 : k] : CUDART_INF_F)
                                : from_right[k];
    }
  }
}

// Minimum over the disparities below the limit the context was built with.
template <int NV, int VW>
__device__ __forceinline__ float lv_min(const LaneVec<NV, VW> &a, const LaneCtx &cx) {
  float m = cx.excl_last ? CUDART_INF_F : a.v[NV - 1][VW - 1];
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j)
      if (!(k == NV - 1 && j == VW - 1)) m = fminf(m, a.v[k][j]);
  return warp_min_f32(m);
}

// One step of the SGM recurrence (reference src/semiglobal.cpp:145-156):
//   L[d] = (c[d] + min(L[d], L[d-1]+P1, L[d+1]+P1, m+P2')) - m
template <int NV, int VW>
__device__ __forceinline__ void lv_sgm_step(LaneVec<NV, VW> &L, const LaneVec<NV, VW> &c, float p1,
                                            float p2_adj, const LaneCtx &cx) {
  const float m = lv_min<NV, VW>(L, cx);
  LaneVec<NV, VW> lo, hi;
  lv_neighbours<NV, VW>(L, lo, hi, cx);
  const float far = __fadd_rn(m, p2_adj);
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j) {
      const float a = __fadd_rn(lo.v[k][j], p1);
      const float b = __fadd_rn(hi.v[k][j], p1);
      const float best = fminf(fminf(L.v[k][j], a), fminf(b, far));
      L.v[k][j] = __fsub_rn(__fadd_rn(c.v[k][j], best), m);
    }
}

// Packed fp32 pair arithmetic (FADD2 on sm_100a: add.rn.f32x2, same rounding as two FADDs).
__device__ __forceinline__ void f2_add(float &d0, float &d1, float a0, float a1, float b0, float b1) {
  unsigned long long ra, rb, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b0), "f"(b1));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(rd));
}
__device__ __forceinline__ void f2_sub(float &d0, float &d1, float a0, float a1, float b0, float b1) {
  unsigned long long ra, rb, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b0), "f"(b1));
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(rd));
}

// lv_sgm_step with the additions issued as packed pairs (VW even): identical results, 7 FADD2
// instead of 14 FADD per 4 values.  For kernels that are bound by issue slots (sgm_vsweep.cu).
template <int NV, int VW>
__device__ __forceinline__ void lv_sgm_step_packed(LaneVec<NV, VW> &L, const LaneVec<NV, VW> &c, float p1,
                                                   float p2_adj, const LaneCtx &cx) {
  static_assert(VW % 2 == 0, "packed step needs an even vector width");
  const float m = lv_min<NV, VW>(L, cx);
  float from_left[NV], from_right[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    from_left[k] = __shfl_sync(FULL, L.v[k][VW - 1], cx.src_lo);
    from_right[k] = __shfl_sync(FULL, L.v[k][0], cx.src_hi);
  }
  const float far = __fadd_rn(m, p2_adj);
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    // values at d-1 of the first element and d+1 of the last one (+inf outside [0, Dpad))
    const float lo0 = cx.first ? (k > 0 ? from_left[k > 0 ? k - 1 : 0] : CUDART_INF_F) : from_left[k];
    const float hiN = cx.last ? (k < NV - 1 ? from_right[k < NV - 1 ? k + 1 : k] : CUDART_INF_F) : from_right[k];
    float Lp[VW];  // L[j] + p1: the left neighbour term of j+1 and the right neighbour term of j-1
#pragma unroll
    for (int j = 0; j < VW; j += 2) f2_add(Lp[j], Lp[j + 1], L.v[k][j], L.v[k][j + 1], p1, p1);
    float e0, e1;  // the two terms that come from the neighbouring lanes
    f2_add(e0, e1, lo0, hiN, p1, p1);
    float best[VW];
#pragma unroll
    for (int j = 0; j < VW; ++j) {
      const float a = j > 0 ? Lp[j - 1] : e0;
      const float b = j < VW - 1 ? Lp[j + 1] : e1;
      best[j] = fminf(fminf(L.v[k][j], a), fminf(b, far));
    }
#pragma unroll
    for (int j = 0; j < VW; j += 2) {
      float s0, s1;
      f2_add(s0, s1, c.v[k][j], c.v[k][j + 1], best[j], best[j + 1]);
      f2_sub(L.v[k][j], L.v[k][j + 1], s0, s1, m, m);
    }
  }
}

// max(P1, P2/|dI|) with std::max semantics (x/0 = inf, 0/0 = NaN -> P1), IEEE division.
__device__ __forceinline__ float sgm_p2_adjusted(float p1, float p2, float intensity, float prev) {
  const float q = __fdiv_rn(p2, fabsf(__fsub_rn(intensity, prev)));
  return (p1 < q) ? q : p1;
}

// Winner-takes-all of one pixel's vector held across the warp, with the rule of wta.cu /
// reference src/winners_take_all.cu:21-40: first (lowest-d) minimum, all-inf -> 0, NaN never
// wins, NaN at d = 0 -> 0.  Used when the LAST aggregation pass also emits the disparity map.
template <int NV, int VW>
__device__ __forceinline__ void lv_argmin_store(const LaneVec<NV, VW> &o, int lane, int D,
                                                int32_t *dst) {
  // two warp reductions instead of a (value, index) shuffle butterfly, which costs this
  // latency-sensitive kernel as much as the separate WTA pass it replaces (measured):
  // m = minimum over the non-NaN entries (redux.sync.min.f32 returns the non-NaN operand),
  // then the lowest d whose entry equals m (integer redux).  All-inf: m = inf matches d = 0.
  float m = CUDART_INF_F;
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j)
      if (lv_d<NV, VW>(lane, k, j) < D) m = fminf(m, o.v[k][j]);
  m = warp_min_f32(m);
  unsigned idx = 0x7fffffffu;
#pragma unroll
  for (int k = NV - 1; k >= 0; --k)
#pragma unroll
    for (int j = VW - 1; j >= 0; --j) {
      const int d = lv_d<NV, VW>(lane, k, j);
      if (d < D && o.v[k][j] == m) idx = (unsigned)d;  // descending d: the lowest one sticks
    }
  idx = __reduce_min_sync(FULL, idx);
  // NaN at d = 0 -> 0: d = 0 is element (0, 0) of lane 0, the lane that stores: no broadcast needed
  if (lane == 0) *dst = (o.v[0][0] != o.v[0][0] || idx == 0x7fffffffu) ? 0 : (int)idx;
}

}  // namespace stm
