// sgm.cu -- semiglobal path aggregation over the reference path set.
//
// Replaces AggregationOps::RunSemiglobal (reference src/semiglobal.cpp:167-197; recurrence
// SGMCostOperator::operator(), :116-159; CUDA variant src/semiglobal_gpu.cu:43-106).
//
// Per directed path p0..p(n-1):  L0 = C[p0];  for i >= 1
//     m   = min_{k in K} L(i-1)[k]              K = [0, D-2] (CPU reference) or [0, D-1]
//     P2' = max(P1, P2 / |I[pi] - I[p(i-1)]|)   std::max semantics (x/0 = inf, 0/0 -> P1)
//     Li[d] = (C[pi,d] + min(L[d], L[d-1]+P1, L[d+1]+P1, m+P2')) - m
//   out[pi,:] += Li for every i including 0.
//
// Round-1 kernel: one warp per directed path, one launch per direction in the
// reference's accumulation order (-> <- v ^ \v ^\ /v ^/), so that every out[] element sees
// exactly the reference's sequence of fp32 additions (bit-exact for W >= H).  The
// disparity vector of the previous pixel lives in registers spread over the lanes; L[d+-1]
// come from two rotating shuffles per vector, min_k from one redux.sync on
// order-preserving integer keys.  The next pixel's cost vector is prefetched while the
// current one is being reduced.
#include <mutex>

#include "lanevec.cuh"
#include <cstdlib>
#include <cstring>

#include "sgm_paths.cuh"

namespace stm {

enum : unsigned { kSgmAccumulate = 1u, kSgmStoreL = 2u, kSgmNoOutput = 4u, kSgmEmitWta = 8u,
                  kSgmSkipVolumeStore = 16u /* with kSgmEmitWta: the sum is reduced, not written */,
                  kSgmReverseOrder = 32u /* warps take the paths of the direction last to first */ };

// Prefetch depth per lane-vector size.  PFSEL 0 = default; 1 = deep (twice as many pixels in
// flight, for directions with so few paths that they all fit on the GPU anyway); 2 = shallow:
// at most 128 registers, so 4 instead of 3 CTAs (16 instead of 12 warps) are resident per SM
// and all 1920 vertical paths of a 1080p frame run in ONE wave instead of 1776 + a 144-warp
// tail that costs a full chain of 1080 dependent steps at next to no bandwidth.  (CTAs stay
// at 4 warps: registers are allocated per 4 warps, a 2-warp CTA pays for 4.)
template <int NV, int VW, int PFSEL>
struct SgmPf {
  static constexpr int kBase = (NV * VW <= 4) ? 8 : ((NV * VW <= 6) ? 5 : 4);
  static constexpr bool kHasDeep = NV * VW <= 4;     // larger vectors would spill at twice the depth
  static constexpr bool kHasShallow = NV * VW <= 8;  // larger vectors cannot get under 128 registers
  static constexpr int kShallow = (NV * VW <= 4) ? 5 : ((NV * VW <= 6) ? 3 : 2);
  static constexpr int value = (PFSEL == 1 && kHasDeep) ? 2 * kBase
                               : ((PFSEL == 2 && kHasShallow) ? kShallow : kBase);
  static constexpr int kThreads = 128;
  static constexpr int kMinBlocks = (PFSEL == 2 && kHasShallow) ? 4 : 0;  // 0 = unspecified
};

template <int NV, int VW, bool ACC, bool WTA, int PFSEL>
__global__ void __launch_bounds__(SgmPf<NV, VW, PFSEL>::kThreads, SgmPf<NV, VW, PFSEL>::kMinBlocks)
sgm_path_kernel(const float *__restrict__ cost, const float *__restrict__ image,
                float *__restrict__ out, int H, int W, int D, float p1, float p2, int dir,
                int path_begin, int path_count, int min_limit, unsigned mode,
                float *__restrict__ l_store, int l_store_path, int32_t *__restrict__ disp) {
  const int lane = threadIdx.x & 31;
  const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= path_count) return;
  const int path = path_begin + ((mode & kSgmReverseOrder) ? path_count - 1 - w : w);
  const PathGeom g = sgm_path_geom(dir, path, H, W);
  if (g.mult == 0) return;
  const bool write_out = !(mode & kSgmNoOutput);
  const bool store_l = (mode & kSgmStoreL) && path == l_store_path;
  const bool store_vol = !(WTA && (mode & kSgmSkipVolumeStore));
  constexpr int dpad = NV * 32 * VW;

  // Pixel i of the path sits at pix0 + i*step: addresses are known up front, so the cost
  // vector, the output vector (for the read-modify-write) and the intensity of the next
  // PF pixels are kept in flight in registers while the recurrence runs.  The loop runs
  // whole groups of PF pixels: prefetches past the end are clamped to the last pixel and
  // the surplus steps are computed but not stored (branch-free steady state).
  constexpr int PF = SgmPf<NV, VW, PFSEL>::value;
  constexpr int RB = (PFSEL == 1 && SgmPf<NV, VW, PFSEL>::kHasDeep) ? 4 : 1;  // refill batch
  const LaneCtx cx = lv_ctx<NV, VW>(lane, min_limit);
  const long long pix0 = (long long)g.sy * W + g.sx;
  const long long step = (long long)g.dy * W + g.dx;
  const long long step_d = step * D;
  const int n = g.n;
  const int lane_off = lane * VW;

  // all pointers are pre-offset by the lane's first disparity
  const float *cfetch = cost + pix0 * D + lane_off;
  const float *sfetch = out + pix0 * D + lane_off;
  const float *ifetch = image + pix0;
  float *optr = out + pix0 * D + lane_off;
  float *lptr = l_store + lane_off;
  const int dl = D - lane_off;  // lv_load/lv_store bound relative to the lane offset

  // step 0: L = C[p0]
  LaneVec<NV, VW> L;
  lv_load<NV, VW, true>(L, cfetch, 0, dl);
  float prev_intensity = __ldg(ifetch);
  if (write_out) {
    if (ACC) {
      LaneVec<NV, VW> o;
      lv_load<NV, VW, true>(o, sfetch, 0, dl);
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int j = 0; j < VW; ++j) o.v[k][j] = __fadd_rn(o.v[k][j], L.v[k][j]);
      if (store_vol) lv_store<NV, VW, true>(o, optr, 0, dl);
      if (WTA) lv_argmin_store<NV, VW>(o, lane, D, disp + pix0);
    } else {
      lv_store<NV, VW, true>(L, optr, 0, dl);
    }
  }
  if (store_l) lv_store<NV, VW, false>(L, lptr, 0, dpad);
  if (n == 1) return;

  // prologue: pixels 1..PF in flight
  LaneVec<NV, VW> cbuf[PF], sbuf[PF];
  float ibuf[PF];
  int left = n - 1;  // pixels not yet fetched (beyond the current fetch pointer)
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    if (left > 0) { cfetch += step_d; sfetch += step_d; ifetch += step; --left; }
    lv_load<NV, VW, true>(cbuf[u], cfetch, 0, dl);
    if (ACC) lv_load<NV, VW, true>(sbuf[u], sfetch, 0, dl);
    ibuf[u] = __ldg(ifetch);
  }

  for (int i = 1; i < n; i += PF) {
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const LaneVec<NV, VW> c = cbuf[u];
      LaneVec<NV, VW> o;
      if (ACC) o = sbuf[u];
      const float intensity = ibuf[u];
      // refill with pixel i+u+PF (clamped): slot u at once, or RB consumed slots together so
      // that one warp asks for RB consecutive pixels (RB*D*4 contiguous bytes) back to back
      if ((u + 1) % RB == 0) {
#pragma unroll
        for (int b = u + 1 - RB; b <= u; ++b) {
          if (left > 0) { cfetch += step_d; sfetch += step_d; ifetch += step; --left; }
          lv_load<NV, VW, true>(cbuf[b], cfetch, 0, dl);
          if (ACC) lv_load<NV, VW, true>(sbuf[b], sfetch, 0, dl);
          ibuf[b] = __ldg(ifetch);
        }
      }

      lv_sgm_step<NV, VW>(L, c, p1, sgm_p2_adjusted(p1, p2, intensity, prev_intensity), cx);
      prev_intensity = intensity;
      const bool live = i + u < n;
      optr += step_d;
      lptr += dpad;
      if (live && write_out) {
        if (ACC) {
#pragma unroll
          for (int k = 0; k < NV; ++k)
#pragma unroll
            for (int j = 0; j < VW; ++j) o.v[k][j] = __fadd_rn(o.v[k][j], L.v[k][j]);
          if (store_vol) lv_store<NV, VW, true>(o, optr, 0, dl);
          if (WTA) lv_argmin_store<NV, VW>(o, lane, D, disp + pix0 + (long long)(i + u) * step);
        } else {
          lv_store<NV, VW, true>(L, optr, 0, dl);
        }
      }
      if (live && store_l) lv_store<NV, VW, false>(L, lptr, 0, dpad);
    }
  }
}

// Diagonal directions (4..7) of a landscape or square image as a ROW-SYNCHRONOUS sweep.
// Warp w is a column slot: at step i it is at row y = (dy > 0 ? i : H-1-i), column
// (w + dx*i) mod W.  A slot that leaves the image on one side re-enters on the other, where a
// new path of the same direction begins (L = C there); with W >= H that happens at most once
// per slot.  Same paths, same arithmetic and same per-pixel accumulation as sgm_path_kernel --
// only the assignment of paths to warps differs:
//  * every active warp does exactly H steps (one-path-per-warp gives lengths 1..H, and the
//    pass ends with a few long chains running alone at next to no bandwidth);
//  * all warps are on the same image row at the same time, so together they read whole
//    contiguous rows (the pattern that streams at ~6.6 TB/s) instead of 983-KB-strided pieces;
//  * for the defective up-diagonals (5, 7) a slot is either active in both of
//    its paths or in neither, so the skipped band costs nothing: idle slots exit at once.
template <int NV, int VW, int PFSEL>
__global__ void __launch_bounds__(SgmPf<NV, VW, PFSEL>::kThreads, SgmPf<NV, VW, PFSEL>::kMinBlocks)
sgm_diag_rows_kernel(const float *__restrict__ cost, const float *__restrict__ image,
                     float *__restrict__ out, int H, int W, int D, float p1, float p2, int dir,
                     int min_limit, unsigned mode, float *__restrict__ l_store, int l_store_path) {
  const int lane = threadIdx.x & 31;
  const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= W) return;
  const int dx = (dir == 4 || dir == 7) ? 1 : -1;
  const int dy = (dir == 4 || dir == 6) ? 1 : -1;
  // the path this slot is on at step 0 (it starts on the top or bottom row at column w)
  const int idx0 = (dir == 4 || dir == 7) ? w + H - 1 : w;
  if (sgm_path_geom(dir, idx0, H, W).mult == 0) return;  // slot inside the skipped band
  const bool store_l = (mode & kSgmStoreL) && idx0 == l_store_path;
  constexpr int dpad = NV * 32 * VW;
  constexpr int PF = SgmPf<NV, VW, PFSEL>::value;
  const LaneCtx cx = lv_ctx<NV, VW>(lane, min_limit);

  const int n = H;
  const int iw = dx > 0 ? W - w : w + 1;  // first step on the slot's second path (>= H: none)
  const long long step = (long long)dy * W + dx;
  const long long wrap = -(long long)dx * W;  // added to `step` when the column wraps
  const long long pix0 = (long long)(dy > 0 ? 0 : H - 1) * W + w;
  const int lane_off = lane * VW;
  const float *cfetch = cost + pix0 * D + lane_off;
  const float *sfetch = out + pix0 * D + lane_off;
  const float *ifetch = image + pix0;
  float *optr = out + pix0 * D + lane_off;
  float *lptr = l_store + lane_off;
  const int dl = D - lane_off;

  // step 0: L = C
  LaneVec<NV, VW> L;
  lv_load<NV, VW, true>(L, cfetch, 0, dl);
  float prev_intensity = __ldg(ifetch);
  {
    LaneVec<NV, VW> o;
    lv_load<NV, VW, true>(o, sfetch, 0, dl);
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int j = 0; j < VW; ++j) o.v[k][j] = __fadd_rn(o.v[k][j], L.v[k][j]);
    lv_store<NV, VW, true>(o, optr, 0, dl);
  }
  if (store_l) lv_store<NV, VW, false>(L, lptr, 0, dpad);
  if (n == 1) return;

  LaneVec<NV, VW> cbuf[PF], sbuf[PF];
  float ibuf[PF];
  int left = n - 1;    // pixels not yet fetched
  int fnext = 1;       // step index the fetch pointers move to next
  auto advance_fetch = [&]() {  // pointer arithmetic only: no load sits under this branch
    if (left > 0) {
      const long long adv = step + (fnext == iw ? wrap : 0);
      cfetch += adv * D; sfetch += adv * D; ifetch += adv;
      ++fnext; --left;
    }
  };
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    advance_fetch();
    lv_load<NV, VW, true>(cbuf[u], cfetch, 0, dl);
    lv_load<NV, VW, true>(sbuf[u], sfetch, 0, dl);
    ibuf[u] = __ldg(ifetch);
  }

  for (int i = 1; i < n; i += PF) {
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const LaneVec<NV, VW> c = cbuf[u];
      LaneVec<NV, VW> o = sbuf[u];
      const float intensity = ibuf[u];
      advance_fetch();
      lv_load<NV, VW, true>(cbuf[u], cfetch, 0, dl);
      lv_load<NV, VW, true>(sbuf[u], sfetch, 0, dl);
      ibuf[u] = __ldg(ifetch);

      const int cur = i + u;
      if (cur == iw) {
        L = c;  // the slot re-entered the image: first pixel of its second path
      } else {
        lv_sgm_step<NV, VW>(L, c, p1, sgm_p2_adjusted(p1, p2, intensity, prev_intensity), cx);
      }
      prev_intensity = intensity;
      optr += (step + (cur == iw ? wrap : 0)) * D;
      lptr += dpad;
      if (cur < n) {
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int j = 0; j < VW; ++j) o.v[k][j] = __fadd_rn(o.v[k][j], L.v[k][j]);
        lv_store<NV, VW, true>(o, optr, 0, dl);
        if (store_l) lv_store<NV, VW, false>(L, lptr, 0, dpad);
      }
    }
  }
}

// Adds the stored path vector `times` more times, one fp32 addition at a time exactly as
// the reference does when it re-runs the corner diagonal.
__global__ void sgm_repeat_add_kernel(float *__restrict__ out, const float *__restrict__ l_store,
                                      int W, int D, int dpad, int sx, int sy, int dx, int dy,
                                      int n, int times) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * D) return;
  const int i = t / D, d = t - i * D;
  const size_t pix = (size_t)(sy + i * dy) * W + (sx + i * dx);
  float s = out[pix * D + d];
  const float l = l_store[(size_t)i * dpad + d];
  for (int r = 0; r < times; ++r) s = __fadd_rn(s, l);
  out[pix * D + d] = s;
}

// Warps of one kernel variant that can be resident on the whole GPU at once (occupancy API).
template <int NV, int VW, bool ACC, bool WTA, int PFSEL>
static int resident_warps() {
  static PerDeviceInt cache;  // per instantiation and device; a race recomputes the same value
  const int dev = device_slot();
  int cached = 0;
  if (!cache.get(dev, &cached)) {
    int nb = 0;
    const int threads = SgmPf<NV, VW, PFSEL>::kThreads;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
            &nb, sgm_path_kernel<NV, VW, ACC, WTA, PFSEL>, threads, 0) != cudaSuccess) {
      cudaGetLastError();
      nb = 0;
    }
    cached = nb * (threads / 32) * num_sms();
    cache.set(dev, cached);
  }
  return cached;
}

template <int NV, int VW, bool ACC, bool WTA, int PFSEL, typename... Args>
static void launch_path_kernel(int path_count, cudaStream_t stream, Args... args) {
  const int warps = SgmPf<NV, VW, PFSEL>::kThreads / 32;
  const int blocks = (path_count + warps - 1) / warps;
  sgm_path_kernel<NV, VW, ACC, WTA, PFSEL><<<blocks, warps * 32, 0, stream>>>(args...);
}

// One wave if possible: the deepest prefetch whose resident-warp capacity still holds every
// path of the direction; the default when nothing fits (diagonals: many short paths, the
// block scheduler balances them).  The result does not depend on the choice.
template <int NV, int VW, bool ACC>
static int pick_pfsel(int path_count) {
  if (SgmPf<NV, VW, 1>::kHasDeep && path_count <= resident_warps<NV, VW, ACC, false, 1>()) return 1;
  if (path_count <= resident_warps<NV, VW, ACC, false, 0>()) return 0;
  if (SgmPf<NV, VW, 2>::kHasShallow && path_count <= resident_warps<NV, VW, ACC, false, 2>()) return 2;
  return 0;
}

// the pass that also emits the disparity map: default or 128-register variant, same rule
template <int NV, int VW>
static int pick_pfsel_wta(int path_count) {
  if (path_count <= resident_warps<NV, VW, true, true, 0>()) return 0;
  if (SgmPf<NV, VW, 2>::kHasShallow && path_count <= resident_warps<NV, VW, true, true, 2>()) return 2;
  return 0;
}

// sgm_row.cu
int launch_sgm_row(const float *, const float *, float *, int, int, int, float, float, int, int,
                   bool, cudaStream_t);

// Developer knob: STM_DEBUG_SGM_PFSEL="01220000" forces the variant per direction ('-' = auto).
static int pfsel_override(int dir) {
  static int tbl[8];
  static bool init = false;
  if (!init) {
    const char *e = getenv("STM_DEBUG_SGM_PFSEL");
    const bool ok = e && strlen(e) == 8;
    for (int i = 0; i < 8; ++i) tbl[i] = (ok && e[i] >= '0' && e[i] <= '2') ? e[i] - '0' : -1;
    init = true;
  }
  return tbl[dir & 7];
}

// Row-synchronous diagonal pass (sgm_diag_rows_kernel): W slots; `busy` of them do H steps.
template <int NV, int VW, int PFSEL>
static int diag_rows_resident() {
  static PerDeviceInt cache;
  const int dev = device_slot();
  int cached = 0;
  if (!cache.get(dev, &cached)) {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, sgm_diag_rows_kernel<NV, VW, PFSEL>,
                                                      SgmPf<NV, VW, PFSEL>::kThreads, 0) != cudaSuccess) {
      cudaGetLastError();
      nb = 0;
    }
    cached = nb * (SgmPf<NV, VW, PFSEL>::kThreads / 32) * num_sms();
    cache.set(dev, cached);
  }
  return cached;
}

// Returns 1 when the pass is better left to sgm_path_kernel.  Measured at 1080p D=128
// (profiles/r01_v13_sgm_diag_rows.txt): with one slot per column the full diagonals (4, 6) are
// no faster than one warp per path (503 vs 495 us), and the defective ones (5, 7; H busy slots)
// only win with 16 pixels in flight per warp (355 vs 365 us; 410 us at depth 8) -- with so few
// warps the bytes in flight per warp decide, not the balance.  So: directions 5 and 7, and only
// when the deep variant exists for the lane map and all busy slots are resident at once.
template <int NV, int VW>
static int run_diag_rows(const float *cost, const float *image, float *out, int H, int W, int D,
                         float p1, float p2, int dir, int min_limit, unsigned mode, float *l_store,
                         int l_store_path, cudaStream_t stream) {
  if (dir != 5 && dir != 7) return 1;
  if (!SgmPf<NV, VW, 1>::kHasDeep) return 1;
  const int busy = H + 1;  // the other W-H-1 slots lie in the skipped band and exit at once
  if (busy > diag_rows_resident<NV, VW, 1>()) return 1;
  const int warps = 4, blocks = (W + warps - 1) / warps;
  sgm_diag_rows_kernel<NV, VW, 1><<<blocks, warps * 32, 0, stream>>>(
      cost, image, out, H, W, D, p1, p2, dir, min_limit, mode, l_store, l_store_path);
  return check_launch("sgm_diag_rows_kernel");
}

template <int NV, int VW>
static int run_direction(const float *cost, const float *image, float *out, int H, int W, int D,
                         float p1, float p2, int dir, int path_begin, int path_count,
                         int min_limit, unsigned mode, float *l_store, int l_store_path,
                         cudaStream_t stream, int32_t *disp = nullptr) {
#define STM_SGM_ARGS cost, image, out, H, W, D, p1, p2, dir, path_begin, path_count, min_limit, mode, \
                     l_store, l_store_path
  // horizontal directions: tiles staged through shared memory by bulk copies (sgm_row.cu)
  if (dir <= 1 && path_begin == 0 && !(mode & (kSgmStoreL | kSgmNoOutput | kSgmEmitWta))) {
    const int rc = launch_sgm_row(cost, image, out, H, W, D, p1, p2, dir, min_limit,
                                  (mode & kSgmAccumulate) != 0, stream);
    if (rc != 1) return rc;
  }
  // diagonal directions of landscape / square images: row-synchronous slots
  if (dir >= 4 && W >= H && H >= 2 && path_begin == 0 && (mode & kSgmAccumulate) &&
      !(mode & (kSgmNoOutput | kSgmEmitWta))) {
    static const char *off = getenv("STM_DEBUG_SGM_DIAG_ROWS");  // developer knob: "0" disables
    if (!(off && off[0] == '0')) {
      const int rc = run_diag_rows<NV, VW>(cost, image, out, H, W, D, p1, p2, dir, min_limit, mode,
                                           l_store, l_store_path, stream);
      if (rc != 1) return rc;
    }
  }
  const int ov = pfsel_override(dir);
  {
    // Launch order of a direction's paths (the result does not depend on it).  The paths of a
    // diagonal that start on the top/bottom edge advance through the image rows together, so
    // the resident warps read contiguous memory; those that start on the left/right edge are
    // each in a row of their own.  Starting with the row-synchronous set is ~80 us faster at
    // 1080p (measured, profiles/r01_v10_sgm_order_sweep.txt): direction 4 enumerates its
    // left-edge paths first, so it is taken last to first; 5, 6, 7 are fine as they are.
    static const char *flip = getenv("STM_DEBUG_SGM_FLIP");  // developer knob, 8 x '0'/'1'
    const bool dflt = dir == 4;
    const bool rev = (flip && strlen(flip) == 8) ? flip[dir & 7] == '1' : dflt;
    if (rev) mode |= kSgmReverseOrder;
  }
  if ((mode & kSgmEmitWta) && disp) {  // last pass of the fused aggregate+reduce call
    if ((ov >= 0 ? ov : pick_pfsel_wta<NV, VW>(path_count)) == 2)
      launch_path_kernel<NV, VW, true, true, 2>(path_count, stream, STM_SGM_ARGS, disp);
    else launch_path_kernel<NV, VW, true, true, 0>(path_count, stream, STM_SGM_ARGS, disp);
  } else if (mode & kSgmAccumulate) {
    switch (ov >= 0 ? ov : pick_pfsel<NV, VW, true>(path_count)) {
      case 1: launch_path_kernel<NV, VW, true, false, 1>(path_count, stream, STM_SGM_ARGS, nullptr); break;
      case 2: launch_path_kernel<NV, VW, true, false, 2>(path_count, stream, STM_SGM_ARGS, nullptr); break;
      default: launch_path_kernel<NV, VW, true, false, 0>(path_count, stream, STM_SGM_ARGS, nullptr);
    }
  } else {
    switch (ov >= 0 ? ov : pick_pfsel<NV, VW, false>(path_count)) {
      case 1: launch_path_kernel<NV, VW, false, false, 1>(path_count, stream, STM_SGM_ARGS, nullptr); break;
      case 2: launch_path_kernel<NV, VW, false, false, 2>(path_count, stream, STM_SGM_ARGS, nullptr); break;
      default: launch_path_kernel<NV, VW, false, false, 0>(path_count, stream, STM_SGM_ARGS, nullptr);
    }
  }
#undef STM_SGM_ARGS
  return check_launch("sgm_path_kernel");
}

int sgm_repeat_add(float *out, float *l_store, int H, int W, int D, int dpad, int dir, int path,
                   int times, cudaStream_t stream) {
  const PathGeom g = sgm_path_geom(dir, path, H, W);
  const int total = g.n * D;
  sgm_repeat_add_kernel<<<(total + 255) / 256, 256, 0, stream>>>(out, l_store, W, D, dpad, g.sx,
                                                                g.sy, g.dx, g.dy, g.n, times);
  return check_launch("sgm_repeat_add_kernel");
}

// ---- per-pass timing (measurement API, include/stm_abi.h: stm_sgm_pass_timing) -------------
// CUDA events on the launching stream around each direction's launches (its corner-diagonal
// replay included), so bench.py can report every pass against the roofline without a profiler.
static bool g_pass_timing = false;
static cudaEvent_t g_pass_ev[9];
static bool g_pass_ev_ready = false;
static int g_pass_dirs[8];
static int g_pass_count = 0;

int sgm_pass_timing(bool enable) {
  if (enable && !g_pass_ev_ready) {
    for (int i = 0; i < 9; ++i)
      if (cudaEventCreate(&g_pass_ev[i]) != cudaSuccess) return fail(STM_ECUDA, "sgm: cudaEventCreate failed");
    g_pass_ev_ready = true;
  }
  g_pass_timing = enable;
  g_pass_count = 0;
  return STM_OK;
}

int sgm_last_pass_ms(float *ms, int *dirs, int capacity) {
  if (g_pass_count == 0) return 0;
  if (capacity < g_pass_count) return fail(STM_EINVAL, "sgm: pass buffer too small");
  if (cudaEventSynchronize(g_pass_ev[g_pass_count]) != cudaSuccess)
    return fail(STM_ECUDA, "sgm: cudaEventSynchronize failed");
  for (int i = 0; i < g_pass_count; ++i) {
    if (cudaEventElapsedTime(&ms[i], g_pass_ev[i], g_pass_ev[i + 1]) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: cudaEventElapsedTime failed");
    dirs[i] = g_pass_dirs[i];
  }
  return g_pass_count;
}

static inline void pass_mark(int pass, int dir, cudaStream_t stream) {
  if (!g_pass_timing) return;
  cudaEventRecord(g_pass_ev[pass], stream);
  if (pass < 8) g_pass_dirs[pass] = dir;
  g_pass_count = pass;  // number of complete intervals so far
}

template <int NV, int VW>
static int sgm_by_direction(const float *cost, const float *image, float *out, int H, int W,
                            int D, float p1, float p2, unsigned flags, float *ws,
                            cudaStream_t stream, int32_t *disp) {
  const int min_limit = (flags & STM_SGM_MIN_ALL_D) ? D : D - 1;
  const int dpad = NV * 32 * VW;
  int rc;
  // Reference order 0..7.  When the disparity map is emitted by the last pass, that pass must
  // visit every pixel exactly once after everything else (the corner-diagonal replays
  // included); direction 7 skips a band of pixels on landscape images, so the
  // last two directions swap and direction 6 runs last.
  const int plain[8] = {0, 1, 2, 3, 4, 5, 6, 7};
  // (a vertical direction last, one wave at 16 warps/SM, was measured: 60 us slower per frame)
  const int fused[8] = {0, 1, 2, 3, 4, 5, 7, 6};
  const int *order = disp ? fused : plain;
  for (int pass = 0; pass < 8; ++pass) {
    const int dir = order[pass];
    pass_mark(pass, dir, stream);
    const int np = sgm_num_paths(dir, H, W);
    unsigned mode = (pass == 0 && (flags & STM_SGM_OVERWRITE)) ? 0u : kSgmAccumulate;
    if (disp && pass == 7)
      mode |= kSgmEmitWta | ((flags & STM_SGM_DISPARITY_ONLY) ? kSgmSkipVolumeStore : 0u);
    const int multi = sgm_multi_path_index(dir, H, W);
    if (multi < 0) {
      rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, dir, 0, np, min_limit, mode,
                                 nullptr, -1, stream, disp);
      if (rc) return rc;
      continue;
    }
    // landscape: the extra runs of the corner up-diagonal follow that diagonal's own
    // forward+inverse pass in the reference's list order -> run the direction, then add
    // the stored vector (mult-1) more times.
    const int times = sgm_path_geom(dir, multi, H, W).mult - 1;
    rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, dir, 0, np, min_limit,
                               mode | kSgmStoreL, ws, multi, stream);
    if (rc) return rc;
    if (times > 0) {
      rc = sgm_repeat_add(out, ws, H, W, D, dpad, dir, multi, times, stream);
      if (rc) return rc;
    }
  }
  pass_mark(8, -1, stream);
  return STM_OK;
}

// Portrait images (H > W): the defective descriptors precede the corner diagonal's own
// descriptor in the reference's list, so the extra (mult-1) additions of the dir-5 corner
// path land BEFORE that diagonal's dir-4 and dir-5 contributions.  Same kernels, different
// launch order: compute the corner path only (no output), add it mult-1 times, then run
// dir 4 and dir 5 normally.
template <int NV, int VW>
static int sgm_by_direction_portrait(const float *cost, const float *image, float *out, int H,
                                     int W, int D, float p1, float p2, unsigned flags, float *ws,
                                     cudaStream_t stream, int32_t *disp) {
  const int min_limit = (flags & STM_SGM_MIN_ALL_D) ? D : D - 1;
  const int dpad = NV * 32 * VW;
  int rc;
  for (int dir = 0; dir < 8; ++dir) {
    pass_mark(dir, dir, stream);
    const int np = sgm_num_paths(dir, H, W);
    unsigned mode = (dir == 0 && (flags & STM_SGM_OVERWRITE)) ? 0u : kSgmAccumulate;
    if (disp && dir == 7)  // portrait: direction 7 visits every pixel once
      mode |= kSgmEmitWta | ((flags & STM_SGM_DISPARITY_ONLY) ? kSgmSkipVolumeStore : 0u);
    if (dir == 4) {
      const int multi = sgm_multi_path_index(5, H, W);
      const int times = sgm_path_geom(5, multi, H, W).mult - 1;
      if (times > 0) {
        rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, 5, multi, 1, min_limit,
                                   kSgmNoOutput | kSgmStoreL, ws, multi, stream);
        if (rc) return rc;
        rc = sgm_repeat_add(out, ws, H, W, D, dpad, 5, multi, times, stream);
        if (rc) return rc;
      }
    }
    rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, dir, 0, np, min_limit, mode,
                               nullptr, -1, stream, disp);
    if (rc) return rc;
  }
  pass_mark(8, -1, stream);
  return STM_OK;
}

// sgm_vsweep.cu
size_t sgm_vsweep_workspace_bytes(int H, int W, int D);
bool sgm_vsweep_can_run(int H, int W, int D, int variant_group);  // 0: the plain pair; 1: + overwrite / 3 streams; 2: + 4 streams
int launch_sgm_vsweep_prepare(const float *, int, int, int, float, float, void *, cudaStream_t);
int launch_sgm_vsweep(const float *cost, const float *image, float *out, const float *extra, const float *extra2,
                      int32_t *disp, int H, int W, int D, float p1, float p2, int min_limit, int up,
                      bool store_vol, int volume_streams, void *workspace, size_t workspace_bytes,
                      cudaStream_t stream);

// STM_SGM_FUSED_SWEEPS: the two horizontal directions as passes of their own, then the six
// directions with a vertical component as two fused sweeps (sgm_vsweep.cu): four passes over
// the volumes instead of eight.  The last sweep can emit the disparity map.  Returns 1 when
// the shape is not handled by the sweeps (nothing has been launched then).
template <int NV, int VW>
static int sgm_fused(const float *cost, const float *image, float *out, int H, int W, int D,
                     float p1, float p2, unsigned flags, void *ws, size_t ws_bytes,
                     cudaStream_t stream, int32_t *disp) {
  const int min_limit = (flags & STM_SGM_MIN_ALL_D) ? D : D - 1;
  if (ws == nullptr || ws_bytes < sgm_vsweep_workspace_bytes(H, W, D) || !sgm_vsweep_can_run(H, W, D, 0))
    return 1;
  if ((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(out) |
       reinterpret_cast<uintptr_t>(ws)) & 15)
    return 1;
  int rc;
  for (int dir = 0; dir < 2; ++dir) {
    pass_mark(dir, dir, stream);
    const unsigned mode = (dir == 0 && (flags & STM_SGM_OVERWRITE)) ? 0u : kSgmAccumulate;
    rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, dir, 0, H, min_limit, mode,
                               nullptr, -1, stream);
    if (rc) return rc;
  }
  pass_mark(2, 2, stream);
  if ((rc = launch_sgm_vsweep_prepare(image, H, W, D, p1, p2, ws, stream))) return rc;
  rc = launch_sgm_vsweep(cost, image, out, nullptr, nullptr, nullptr, H, W, D, p1, p2, min_limit, 0, true, 2, ws,
                         ws_bytes, stream);
  if (rc == 1) return fail(STM_EUNSUPPORTED, "sgm: fused sweeps refused a shape they accepted");
  if (rc) return rc;
  pass_mark(3, 3, stream);
  rc = launch_sgm_vsweep(cost, image, out, nullptr, nullptr, disp, H, W, D, p1, p2, min_limit, 1,
                         !(disp && (flags & STM_SGM_DISPARITY_ONLY)), 2, ws, ws_bytes, stream);
  if (rc == 1) return fail(STM_EUNSUPPORTED, "sgm: fused sweeps refused a shape they accepted");
  if (rc) return rc;
  pass_mark(4, -1, stream);
  return STM_OK;
}

static size_t sgm_direction_workspace_bytes(int H, int W, int D) {
  const LaneMap m = pick_lane_map(D);
  const size_t dpad = (size_t)m.NV * 32 * m.VW;
  return (size_t)(H < W ? H : W) * dpad * sizeof(float);
}

// Small, L2-resident volumes (BASELINE.json config 2: 450x375x64 = 43 MB) are not bandwidth-
// but latency-bound: eight passes of <= 450 dependent steps each, one after the other.  For
// those the eight directions run CONCURRENTLY on library-owned streams, each into its own
// buffer, and one kernel then adds the buffers in the reference's order (corner-diagonal
// replays included), so the result is bit-identical to the sequential passes.  It costs
// 8 volume-sized buffers of workspace and ~10 % more traffic.  Up to 48 MB the buffers are part of
// the workspace every caller provides (stm_sgm_workspace_bytes).  MID-SIZE volumes (up to 384 MB)
// that do not take the fused sweeps -- 375x450x128, the reference's own SGM benchmark shape; VGA;
// 720p at D=64 -- are latency-bound as eight passes one after the other too (a pass is W or H
// dependent steps of ~0.17 us, its bytes need less): they run concurrently when the caller provides
// the buffers (stm_sgm_workspace_bytes_flags asks for them; measured 0.68 -> 0.45 ms at 375x450x128,
// 0.83 -> 0.45 at 480x640x64, 1.54 -> 1.14 at 720x1280x64, 0.92 -> 0.76 at 480x640x128); above that
// a pass is bound by its bytes and the sequential schedule moves fewer of them.
constexpr size_t kConcurrentMaxVolumeBytes = (size_t)48 << 20;
constexpr size_t kConcurrentMidVolumeBytes = (size_t)384 << 20;
// (portrait images too since r02 v12: the sum kernel takes the multiplicities per pixel and puts the
// corner diagonal's extra runs where the reference's list order puts them for H > W)
static bool sgm_concurrent_eligible(int H, int W, int D) {
  return H >= 2 && W >= 2 && (size_t)H * W * D * sizeof(float) <= kConcurrentMaxVolumeBytes;
}
static bool sgm_concurrent_mid_size(int H, int W, int D) {
  return H >= 2 && W >= 2 && (size_t)H * W * D * sizeof(float) <= kConcurrentMidVolumeBytes;
}
static size_t sgm_base_workspace_bytes(int H, int W, int D) {
  const size_t a = sgm_direction_workspace_bytes(H, W, D), b = sgm_vsweep_workspace_bytes(H, W, D);
  return a > b ? a : b;
}
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

size_t sgm_workspace_bytes(int H, int W, int D) {
  const size_t base = sgm_base_workspace_bytes(H, W, D);
  if (!sgm_concurrent_eligible(H, W, D)) return base;
  return align256(base) + (size_t)8 * H * W * D * sizeof(float);
}
static bool sgm_overlap_shape(int H, int W, int D);
static bool sgm_four_stream_shape(int D);
// with STM_SGM_FUSED_SWEEPS | STM_SGM_OVERLAP: one more volume after the base workspace, and a
// second one with STM_SGM_DISPARITY_ONLY (the four-stream schedule, see sgm_fused_overlap)
size_t sgm_workspace_bytes_flags(int H, int W, int D, unsigned flags) {
  size_t need = sgm_workspace_bytes(H, W, D);
  // mid-size volumes without the fused sweeps: the eight direction buffers of the concurrent schedule
  const bool sweeps = (flags & STM_SGM_FUSED_SWEEPS) && sgm_vsweep_workspace_bytes(H, W, D) != 0;
  if (!(flags & STM_SGM_SEQUENTIAL) && !sweeps && sgm_concurrent_mid_size(H, W, D)) {
    const size_t c = align256(sgm_base_workspace_bytes(H, W, D)) + (size_t)8 * H * W * D * sizeof(float);
    if (c > need) need = c;
  }
  if ((flags & (STM_SGM_FUSED_SWEEPS | STM_SGM_OVERLAP)) == (STM_SGM_FUSED_SWEEPS | STM_SGM_OVERLAP) &&
      sgm_overlap_shape(H, W, D)) {
    const size_t volumes = ((flags & STM_SGM_DISPARITY_ONLY) && sgm_four_stream_shape(D)) ? 2 : 1;
    const size_t o = align256(sgm_base_workspace_bytes(H, W, D)) + volumes * H * W * D * sizeof(float);
    if (o > need) need = o;
  }
  return need;
}

// out = [out +] b0 + b1 + ... + b7, one fp32 addition at a time in the reference's order.  The
// corner up-diagonals (multiplicity m > 1: direction 5 where x - y == W - H, direction 7 where
// x + y == H - 1 on landscape images) get their vector m - 1 more times where the sequential
// schedules put them: right after that direction on landscape images (sgm_by_direction), BEFORE
// direction 4 on portrait ones (sgm_by_direction_portrait; the defective descriptors precede the
// corner diagonal's own in the reference's list).  Pixels a defective direction never visits hold 0
// in its buffer (memset by the caller).
// GROUP > 0 (VEC = 4, D = 4 * GROUP in {32, 64, 128}): the GROUP consecutive threads that hold a pixel's
// sums also take its winner -- first minimum, NaN never wins, NaN at d = 0 -> 0, all +inf -> 0: the rule
// of wta.cu / reference src/winners_take_all.cu:21-40 -- by a butterfly under the full mask (xor
// distances below GROUP stay inside the pixel's lanes), so the separate WTA pass over the volume is
// not needed; with store_out = 0 (STM_SGM_DISPARITY_ONLY) the sums are not written either.
template <int VEC, int GROUP>
__global__ void sgm_sum8_kernel(float *__restrict__ out, const float *__restrict__ bufs,
                                size_t n_vec, size_t stride, int H, int W, int D, int accumulate,
                                int32_t *__restrict__ disp, int store_out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (GROUP == 0 && i >= n_vec) return;
  const bool live = i < n_vec;      // GROUP > 0: n_vec is a multiple of GROUP, so a pixel's threads agree
  if (!live) i = n_vec - 1;
  const size_t e = i * VEC;           // first element of this thread
  const size_t pixel = e / D;
  const int x = (int)(pixel % W), y = (int)(pixel / W);
  const int extra5 = max(sgm_mult5(x, y, H, W) - 1, 0), extra7 = max(sgm_mult7(x, y, H, W) - 1, 0);
  const bool portrait = H > W;
  float s[VEC], b[VEC];
  auto load = [&](const float *p) {
    if constexpr (VEC == 4) {
      const float4 t = *reinterpret_cast<const float4 *>(p + e);
      b[0] = t.x; b[1] = t.y; b[2] = t.z; b[3] = t.w;
    } else {
      b[0] = p[e];
    }
  };
  auto add_b = [&](int times) {
    for (int r = 0; r < times; ++r)
#pragma unroll
      for (int k = 0; k < VEC; ++k) s[k] = __fadd_rn(s[k], b[k]);
  };
  load(bufs);
  if (accumulate) {
#pragma unroll
    for (int k = 0; k < VEC; ++k) s[k] = b[k];
    load(out);
#pragma unroll
    for (int k = 0; k < VEC; ++k) s[k] = __fadd_rn(b[k], s[k]);  // out_old + L0
  } else {
#pragma unroll
    for (int k = 0; k < VEC; ++k) s[k] = b[k];
  }
  for (int dir = 1; dir < 8; ++dir) {
    if (dir == 4 && portrait && extra5 > 0) {
      load(bufs + (size_t)5 * stride);
      add_b(extra5);
    }
    load(bufs + (size_t)dir * stride);
    add_b(1);
    if (dir == 5 && !portrait) add_b(extra5);
    if (dir == 7) add_b(extra7);
  }
  if (live && (GROUP == 0 || store_out)) {
    if constexpr (VEC == 4) {
      *reinterpret_cast<float4 *>(out + e) = make_float4(s[0], s[1], s[2], s[3]);
    } else {
      out[e] = s[0];
    }
  }
  if constexpr (GROUP > 0) {
    const int d0 = (int)(e - pixel * D);   // this thread's first disparity
    float best = CUDART_INF_F;
    int bd = 0x7fffffff;
#pragma unroll
    for (int k = 0; k < VEC; ++k)
      if (s[k] < best) { best = s[k]; bd = d0 + k; }   // strict: the lowest d keeps a tie; NaN never passes
    int nan0 = (d0 == 0 && s[0] != s[0]) ? 1 : 0;
#pragma unroll
    for (int w = 1; w < GROUP; w <<= 1) {
      const float ob = __shfl_xor_sync(FULL, best, w);
      const int od = __shfl_xor_sync(FULL, bd, w);
      nan0 |= __shfl_xor_sync(FULL, nan0, w);
      if (ob < best || (ob == best && od < bd)) { best = ob; bd = od; }
    }
    if (live && d0 == 0) disp[pixel] = (nan0 || bd == 0x7fffffff) ? 0 : bd;
  }
}

// Helper streams and events, one set per device ordinal (a process normally drives one GPU,
// but streams and events belong to the device that was current when they were created).
// Threading: the fork event, the seven helper streams and the join events of a device are shared by
// every caller on that device, so the whole fork ... join ENQUEUE sequence of one call runs under
// the set's mutex (another host thread could otherwise re-record `fork` between this call's record
// and its helper streams' waits).  cudaStreamWaitEvent captures the event's state when it is
// called, so once the sequence has been enqueued the next caller may reuse the events.  Calls from
// several threads / streams on one device are therefore correct and share the helper streams.
struct AuxStreams {
  std::mutex mu;
  cudaStream_t stream[7];
  cudaEvent_t fork, join[7];
  bool ready = false;
};
static AuxStreams g_aux[kMaxDevices];
// Returns the current device's set; the caller must hold a->mu while it uses it.
static int aux_streams(AuxStreams **out) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices)
    return fail(STM_ECUDA, "sgm: unsupported device ordinal");
  AuxStreams &a = g_aux[dev];
  std::lock_guard<std::mutex> init_lock(a.mu);
  if (!a.ready) {
    if (cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: cudaEventCreate failed");
    for (int i = 0; i < 7; ++i) {
      if (cudaStreamCreateWithFlags(&a.stream[i], cudaStreamNonBlocking) != cudaSuccess ||
          cudaEventCreateWithFlags(&a.join[i], cudaEventDisableTiming) != cudaSuccess)
        return fail(STM_ECUDA, "sgm: cudaStreamCreate failed");
    }
    a.ready = true;
  }
  *out = &a;
  return STM_OK;
}

// STM_SGM_OVERLAP (with STM_SGM_FUSED_SWEEPS): the same four passes, but the horizontal passes run
// NEXT TO the down sweep instead of before it.  In the plain fused schedule everything chains
// through `out` (-> writes it, <- and both sweeps read-modify-write it), so 0.98 ms of
// bandwidth-bound horizontal passes (0.82-0.85 of the copy rate) sit in front of 1.44 ms of sweeps
// that are bound by hand-off latency and issue slots and use about half the bandwidth.  With one
// more volume T in the workspace nothing has to chain:
//     stream:       penalties, down sweep  T   = L2 + L4 + L6        (reads C, writes T: 8 B/pix-disp)
//     helper:       ->  out  = [out +] L0 ;  <-  out += L1           (8 + 12 B, concurrently)
//     stream:       up sweep   out = (out + T) + L3 + L5 + L7, argmin  (reads C, out, T)
// Same bytes as before (40-44 B/pix-disp); the sweep CTA (one per SM, 94 KB of shared memory in
// its overwrite form) leaves room for a 96 KB CTA of sgm_row_kernel on every SM.  The sum is
// associated ((L0 + L1) + (L2+L4+L6)) + L3 + L5 + L7: within the 1e-4 bar like the plain fused
// schedule, not bit-identical to it.  Developer knob STM_DEBUG_SGM_OVERLAP=x selects the other
// variant that was measured (-> and <- concurrently with each other, <- into T, then both sweeps).
static size_t sgm_base_workspace_bytes(int H, int W, int D);
static size_t align256(size_t x);
static size_t sgm_overlap_offset(int H, int W, int D) { return align256(sgm_base_workspace_bytes(H, W, D)); }
static bool sgm_overlap_shape(int H, int W, int D) {  // pure function of the shape (workspace query)
  if (H < 2 || W < 2 || sgm_vsweep_workspace_bytes(H, W, D) == 0) return false;
  const LaneMap m = pick_lane_map(D);
  // measured at 1080p: D=128 2.44 -> 2.23 ms; D=192 3.36 -> 3.65 ms (the three-stream up sweep has to
  // give up a row of prefetch to fit in shared memory), so only up to 128 padded disparities
  return m.NV * 32 * m.VW <= 128;
}

// The four-stream form of the schedule (disparity only, see below) pays at 128 padded disparities only:
// measured at 1080p, D=128 2.287 -> 2.248 ms; D=64 1.95 ms against 1.60 ms for the three-stream form that
// even stores the volume (two row kernels next to the sweep slow it down more than the 4 B/pix-disp save).
static bool sgm_four_stream_shape(int D) {
  const LaneMap m = pick_lane_map(D);
  return m.NV * 32 * m.VW == 128;
}

template <int NV, int VW>
static int sgm_fused_overlap(const float *cost, const float *image, float *out, int H, int W, int D,
                             float p1, float p2, unsigned flags, void *ws, size_t ws_bytes,
                             cudaStream_t stream, int32_t *disp) {
  const int min_limit = (flags & STM_SGM_MIN_ALL_D) ? D : D - 1;
  const size_t N = (size_t)H * W * D, off = sgm_overlap_offset(H, W, D);
  if (!sgm_overlap_shape(H, W, D) || ws == nullptr || ws_bytes < off + N * sizeof(float) ||
      !sgm_vsweep_can_run(H, W, D, 1))
    return 1;
  if ((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(ws)) & 15)
    return 1;
  float *T = reinterpret_cast<float *>(static_cast<char *>(ws) + off);
  static const char *knob = getenv("STM_DEBUG_SGM_OVERLAP");
  const bool variant_x = knob && (knob[0] == 'x' || knob[0] == 'X');
  AuxStreams *aux = nullptr;
  int rc = aux_streams(&aux);
  if (rc) return rc;
  std::lock_guard<std::mutex> enqueue_lock(aux->mu);  // fork ... join is one critical section
  // callers on different streams (two frames in flight, stereomatch.streaming) get different helpers
  const int hi = (int)((reinterpret_cast<uintptr_t>(stream) >> 6) % 7u);
  cudaStream_t helper = aux->stream[hi];
  const unsigned mode0 = (flags & STM_SGM_OVERWRITE) ? 0u : kSgmAccumulate;
  const bool store_up = !(disp && (flags & STM_SGM_DISPARITY_ONLY));
  // Disparity only: nothing has to end up in `out`, so <- does not have to chain through it either.
  // With a second workspace volume U the three producers are independent and all of them only write:
  //     stream:    down sweep  T = L2 + L4 + L6          helper a:  ->  out = [out +] L0
  //     helper b:  <-  U = L1                            (8 + 8 + 8 B/pix-disp instead of 8 + 8 + 12)
  //     stream:    up sweep    argmin(((out + T) + U) + L3 + L5 + L7), nothing stored (16 B)
  // The up sweep is bound by its instruction stream, not by bandwidth, so the fourth stream it
  // reads costs less (0.81 -> 0.85 ms) than the read-modify-write it replaces (phase one 1.51 -> 1.39 ms).
  // Developer knob STM_DEBUG_SGM_OVERLAP=3 keeps the three-stream form.
  const bool four = !store_up && !variant_x && !(knob && knob[0] == '3') && sgm_four_stream_shape(D) &&
                    ws_bytes >= off + 2 * N * sizeof(float) && sgm_vsweep_can_run(H, W, D, 2);
  if (four) {
    float *U = T + N;
    const int hb = (hi + 3) % 7;
    cudaStream_t helper_b = aux->stream[hb];
    pass_mark(0, 8, stream);
    if (cudaEventRecord(aux->fork, stream) != cudaSuccess || cudaStreamWaitEvent(helper, aux->fork, 0) != cudaSuccess ||
        cudaStreamWaitEvent(helper_b, aux->fork, 0) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: stream fork failed");
    if ((rc = launch_sgm_vsweep_prepare(image, H, W, D, p1, p2, ws, stream))) return rc;
    rc = launch_sgm_vsweep(cost, image, T, nullptr, nullptr, nullptr, H, W, D, p1, p2, min_limit, 0, true, 1, ws, ws_bytes, stream);
    if (rc == 1) return fail(STM_EUNSUPPORTED, "sgm: fused sweeps refused a shape they accepted");
    if (rc) return rc;
    if ((rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, 0, 0, H, min_limit, mode0, nullptr, -1, helper)))
      return rc;
    if ((rc = run_direction<NV, VW>(cost, image, U, H, W, D, p1, p2, 1, 0, H, min_limit, 0u, nullptr, -1, helper_b)))
      return rc;
    if (cudaEventRecord(aux->join[hi], helper) != cudaSuccess || cudaStreamWaitEvent(stream, aux->join[hi], 0) != cudaSuccess ||
        cudaEventRecord(aux->join[hb], helper_b) != cudaSuccess || cudaStreamWaitEvent(stream, aux->join[hb], 0) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: stream join failed");
    pass_mark(1, 3, stream);
    rc = launch_sgm_vsweep(cost, image, out, T, U, disp, H, W, D, p1, p2, min_limit, 1, false, 4, ws, ws_bytes, stream);
    if (rc == 1) return fail(STM_EUNSUPPORTED, "sgm: fused sweeps refused a shape they accepted");
    if (rc) return rc;
    pass_mark(2, -1, stream);
    return STM_OK;
  }
  pass_mark(0, 8, stream);
  if (cudaEventRecord(aux->fork, stream) != cudaSuccess || cudaStreamWaitEvent(helper, aux->fork, 0) != cudaSuccess)
    return fail(STM_ECUDA, "sgm: stream fork failed");
  if ((rc = launch_sgm_vsweep_prepare(image, H, W, D, p1, p2, ws, stream))) return rc;
  if (!variant_x) {
    // the cooperative sweep first: it needs a slot on every SM; the row CTAs then fill what is left
    rc = launch_sgm_vsweep(cost, image, T, nullptr, nullptr, nullptr, H, W, D, p1, p2, min_limit, 0, true, 1, ws, ws_bytes, stream);
    if (rc == 1) return fail(STM_EUNSUPPORTED, "sgm: fused sweeps refused a shape they accepted");
    if (rc) return rc;
    if ((rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, 0, 0, H, min_limit, mode0, nullptr, -1, helper)))
      return rc;
    if ((rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, 1, 0, H, min_limit, kSgmAccumulate, nullptr, -1, helper)))
      return rc;
    if (cudaEventRecord(aux->join[hi], helper) != cudaSuccess || cudaStreamWaitEvent(stream, aux->join[hi], 0) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: stream join failed");
  } else {
    if ((rc = run_direction<NV, VW>(cost, image, T, H, W, D, p1, p2, 1, 0, H, min_limit, 0u, nullptr, -1, helper)))
      return rc;
    if ((rc = run_direction<NV, VW>(cost, image, out, H, W, D, p1, p2, 0, 0, H, min_limit, mode0, nullptr, -1, stream)))
      return rc;
    rc = launch_sgm_vsweep(cost, image, out, nullptr, nullptr, nullptr, H, W, D, p1, p2, min_limit, 0, true, 2, ws, ws_bytes, stream);
    if (rc == 1) return fail(STM_EUNSUPPORTED, "sgm: fused sweeps refused a shape they accepted");
    if (rc) return rc;
    if (cudaEventRecord(aux->join[hi], helper) != cudaSuccess || cudaStreamWaitEvent(stream, aux->join[hi], 0) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: stream join failed");
  }
  pass_mark(1, 3, stream);
  rc = launch_sgm_vsweep(cost, image, out, T, nullptr, disp, H, W, D, p1, p2, min_limit, 1, store_up, 3, ws, ws_bytes, stream);
  if (rc == 1) return fail(STM_EUNSUPPORTED, "sgm: fused sweeps refused a shape they accepted");
  if (rc) return rc;
  pass_mark(2, -1, stream);
  return STM_OK;
}

// ---- D = 64, concurrent schedule: TWO paths per warp --------------------------------------------
// With 64 disparities a warp of sgm_path_kernel holds two per lane: every instruction of the step does
// half the work it does at D = 128, and eight directions at once are bound by instruction issue
// (450x375x64: ~0.2 ms for the directions where their bytes need a third).  Here a warp runs two
// adjacent paths of the direction, one per 16 lanes, four disparities per lane: the step is the same
// arithmetic in the same order (bit-identical), d-1 / d+1 come from shuffles that stay inside the half
// (full mask, per-lane source), and the minimum over a path's 64 values is a four-step butterfly in the
// half instead of one redux over the warp.  Only what the concurrent schedule needs: L written to the
// direction's buffer, no accumulation, no corner replays (the sum kernel does both).
struct HalfCtx {
  int src_lo, src_hi;
  bool first, last, excl_last;
};
__device__ __forceinline__ float half_min(const float (&v)[4], const HalfCtx &cx) {
  float m = cx.excl_last ? CUDART_INF_F : v[3];
  m = fminf(m, fminf(v[0], fminf(v[1], v[2])));
#pragma unroll
  for (int w = 1; w < 16; w <<= 1) m = fminf(m, __shfl_xor_sync(FULL, m, w));
  return m;
}
// lv_sgm_step<1, 4> for one half: L[d] = (c[d] + min(L[d], L[d-1]+P1, L[d+1]+P1, m+P2')) - m
__device__ __forceinline__ void half_sgm_step(float (&L)[4], const float (&c)[4], float p1, float p2_adj,
                                              const HalfCtx &cx) {
  const float m = half_min(L, cx);
  const float from_left = __shfl_sync(FULL, L[3], cx.src_lo);
  const float from_right = __shfl_sync(FULL, L[0], cx.src_hi);
  const float far = __fadd_rn(m, p2_adj);
  float lo[4], hi[4];
  lo[0] = cx.first ? CUDART_INF_F : from_left;
  lo[1] = L[0]; lo[2] = L[1]; lo[3] = L[2];
  hi[0] = L[1]; hi[1] = L[2]; hi[2] = L[3];
  hi[3] = cx.last ? CUDART_INF_F : from_right;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float a = __fadd_rn(lo[j], p1);
    const float b = __fadd_rn(hi[j], p1);
    const float best = fminf(fminf(L[j], a), fminf(b, far));
    L[j] = __fsub_rn(__fadd_rn(c[j], best), m);
  }
}

__global__ void __launch_bounds__(128)
sgm_path_half64_kernel(const float *__restrict__ cost, const float *__restrict__ image, float *__restrict__ out,
                       int H, int W, float p1, float p2, int dir, int path_count, int min_limit) {
  constexpr int D = 64, PF = 8;
  const int lane = threadIdx.x & 31, l16 = lane & 15, half = lane >> 4;
  const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (2 * w >= path_count) return;  // warp-uniform
  int path = 2 * w + half;
  const bool have = path < path_count;   // an odd count leaves the last warp's second half idle:
  if (!have) path = path_count - 1;      // it shadows the first half and stores nothing
  const PathGeom g = sgm_path_geom(dir, path, H, W);
  const bool writes = have && g.mult != 0;   // paths the reference never runs: the buffer holds 0 (memset)
  HalfCtx cx;
  cx.src_lo = (lane & 16) | ((l16 + 15) & 15);
  cx.src_hi = (lane & 16) | ((l16 + 1) & 15);
  cx.first = l16 == 0;
  cx.last = l16 == 15;
  cx.excl_last = l16 * 4 + 3 >= min_limit;
  const long long pix0 = (long long)g.sy * W + g.sx;
  const long long step = (long long)g.dy * W + g.dx;
  const long long step_d = step * D;
  const int n = g.n;
  const int n_max = max(n, __shfl_xor_sync(FULL, n, 16));
  const float *cfetch = cost + pix0 * D + l16 * 4;
  const float *ifetch = image + pix0;
  float *optr = out + pix0 * D + l16 * 4;
  auto load4 = [](float (&r)[4], const float *q) {
    const float4 t = ld_stream4(q);
    r[0] = t.x; r[1] = t.y; r[2] = t.z; r[3] = t.w;
  };
  float L[4];
  load4(L, cfetch);
  if (writes) st_stream4(optr, make_float4(L[0], L[1], L[2], L[3]));
  if (n_max == 1) return;
  // The adaptive penalty max(P1, P2/|dI|) (an IEEE division, ~15 dependent instructions) depends on the
  // intensities only.  Lane u of a half computes it ONCE for pixel i + u of the current group of PF
  // steps; a shuffle per step hands it out.  Evaluated by every lane at every step it is part of the
  // in-order instruction stream of a latency-bound chain (sgm_row.cu did the same for the same reason).
  const float *ibase = image + pix0;
  auto intensity_of = [&](int i) { return __ldg(ibase + (long long)min(i, n - 1) * step); };  // lane u: pixel i + u
  const int hbase = lane & 16;
  float iprev_grp = intensity_of(0);              // every lane: pixel 0 (the carry into group 1 .. PF)
  float icur = intensity_of(1 + (l16 & (PF - 1)));
  float cbuf[PF][4];
  int left = n - 1;
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    if (left > 0) { cfetch += step_d; --left; }
    load4(cbuf[u], cfetch);
  }
  bool first_group = true;
  for (int i = 1; i < n_max; i += PF) {
    const float inext = intensity_of(i + PF + (l16 & (PF - 1)));
    const float left_px = __shfl_sync(FULL, icur, hbase | ((l16 + 15) & 15));
    const float carry = first_group ? iprev_grp : __shfl_sync(FULL, iprev_grp, hbase | (PF - 1));
    const float p2_grp = sgm_p2_adjusted(p1, p2, icur, (l16 & (PF - 1)) == 0 ? carry : left_px);
    first_group = false;
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      float c[4] = {cbuf[u][0], cbuf[u][1], cbuf[u][2], cbuf[u][3]};
      if (left > 0) { cfetch += step_d; --left; }
      load4(cbuf[u], cfetch);
      half_sgm_step(L, c, p1, __shfl_sync(FULL, p2_grp, hbase | u), cx);
      optr += step_d;
      if (writes && i + u < n) st_stream4(optr, make_float4(L[0], L[1], L[2], L[3]));
    }
    iprev_grp = icur;
    icur = inext;
  }
}

static bool sgm_half64_enabled() {
  static const char *off = getenv("STM_DEBUG_SGM_HALF");  // developer knob: "0" disables
  return !(off && off[0] == '0');
}
static int launch_sgm_half64(const float *cost, const float *image, float *out, int H, int W, float p1, float p2,
                             int dir, int min_limit, cudaStream_t stream) {
  const int np = sgm_num_paths(dir, H, W);
  const int warps = (np + 1) / 2;
  sgm_path_half64_kernel<<<(warps + 3) / 4, 128, 0, stream>>>(cost, image, out, H, W, p1, p2, dir, np, min_limit);
  return check_launch("sgm_path_half64_kernel");
}

int launch_wta(const float *, int32_t *, int, int, int, cudaStream_t);  // wta.cu

template <int NV, int VW>
static int sgm_concurrent(const float *cost, const float *image, float *out, int H, int W, int D,
                          float p1, float p2, unsigned flags, float *bufs, cudaStream_t stream,
                          int32_t *disp) {
  AuxStreams *aux = nullptr;
  int rc = aux_streams(&aux);
  if (rc) return rc;
  std::lock_guard<std::mutex> enqueue_lock(aux->mu);  // fork ... join is one critical section
  const int min_limit = (flags & STM_SGM_MIN_ALL_D) ? D : D - 1;
  const size_t N = (size_t)H * W * D;
  if (cudaEventRecord(aux->fork, stream) != cudaSuccess)
    return fail(STM_ECUDA, "sgm: cudaEventRecord failed");
  for (int dir = 0; dir < 8; ++dir) {
    cudaStream_t s = dir == 0 ? stream : aux->stream[dir - 1];
    float *b = bufs + (size_t)dir * N;
    if (dir > 0 && cudaStreamWaitEvent(s, aux->fork, 0) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: cudaStreamWaitEvent failed");
    // the defective up-diagonals skip a band of the image (direction 5 whenever W != H, direction 7
    // on landscape images): those pixels contribute 0
    if (((dir == 5 && W != H) || (dir == 7 && W > H)) &&
        cudaMemsetAsync(b, 0, N * sizeof(float), s) != cudaSuccess)
      return fail(STM_ECUDA, "sgm: cudaMemsetAsync failed");
    if (D == 64 && sgm_half64_enabled() && !((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(b)) & 15))
      rc = launch_sgm_half64(cost, image, b, H, W, p1, p2, dir, min_limit, s);   // two paths per warp
    else
      rc = run_direction<NV, VW>(cost, image, b, H, W, D, p1, p2, dir, 0, sgm_num_paths(dir, H, W),
                                 min_limit, 0u, nullptr, -1, s);
    if (rc) return rc;
    if (dir > 0) {
      if (cudaEventRecord(aux->join[dir - 1], s) != cudaSuccess ||
          cudaStreamWaitEvent(stream, aux->join[dir - 1], 0) != cudaSuccess)
        return fail(STM_ECUDA, "sgm: stream join failed");
    }
  }
  const int accumulate = (flags & STM_SGM_OVERWRITE) ? 0 : 1;
  const size_t n_vec = N / 4;
  const unsigned blocks4 = (unsigned)((n_vec + 255) / 256);
  const int store_out = (flags & STM_SGM_DISPARITY_ONLY) ? 0 : 1;
  bool emitted = false;
  if (disp && (D == 32 || D == 64 || D == 128)) {
    // the sum kernel takes the winner itself: no pass over the volume for the argmin, no store of the
    // sums when only the map is wanted (375x450x128: 0.47 -> ~0.43 ms)
    if (D == 128) sgm_sum8_kernel<4, 32><<<blocks4, 256, 0, stream>>>(out, bufs, n_vec, N, H, W, D, accumulate, disp, store_out);
    else if (D == 64) sgm_sum8_kernel<4, 16><<<blocks4, 256, 0, stream>>>(out, bufs, n_vec, N, H, W, D, accumulate, disp, store_out);
    else sgm_sum8_kernel<4, 8><<<blocks4, 256, 0, stream>>>(out, bufs, n_vec, N, H, W, D, accumulate, disp, store_out);
    emitted = true;
  } else if (D % 4 == 0) {
    sgm_sum8_kernel<4, 0><<<blocks4, 256, 0, stream>>>(out, bufs, n_vec, N, H, W, D, accumulate, nullptr, 1);
  } else {
    sgm_sum8_kernel<1, 0><<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(out, bufs, N, N, H, W, D, accumulate,
                                                                           nullptr, 1);
  }
  rc = check_launch("sgm_sum8_kernel");
  if (rc) return rc;
  if (disp && !emitted) return launch_wta(out, disp, H, W, D, stream);
  return STM_OK;
}

// Developer hook (tools/sgm_plan.py): resident-warp capacity of the D=128 variants and the
// variant chosen for `paths` paths.  Not part of the ABI.
extern "C" __attribute__((visibility("default"))) void stm_debug_sgm_plan(int paths, int *out6) {
  out6[0] = resident_warps<1, 4, true, false, 0>();
  out6[1] = resident_warps<1, 4, true, false, 1>();
  out6[2] = resident_warps<1, 4, true, false, 2>();
  out6[3] = pick_pfsel<1, 4, true>(paths);
  out6[4] = pick_pfsel<1, 4, false>(paths);
  out6[5] = num_sms();
}

int launch_sgm(const float *cost, const float *image, float *out, int H, int W, int D, float p1,
               float p2, unsigned flags, void *workspace, size_t workspace_bytes,
               cudaStream_t stream, int32_t *disp) {
  if (D < 2) return fail(STM_EINVAL, "sgm: D must be >= 2");
  if (H > 32767 || W > 32767)  // the reference stores coordinates as int16 (semiglobal.hpp:34-36)
    return fail(STM_EINVAL, "sgm: H and W must be <= 32767");
  if (workspace_bytes < sgm_workspace_bytes(H, W, D) || (workspace == nullptr && H != W))
    return fail(STM_EINVAL, "sgm: workspace too small (%zu < %zu)", workspace_bytes,
                sgm_workspace_bytes(H, W, D));
  const LaneMap m = pick_lane_map(D);
  if (flags & STM_SGM_FUSED_SWEEPS) {
    // opt-in: horizontal passes + two fused vertical sweeps (sgm_vsweep.cu); 1 = not handled
    int rc = 1;
    if (flags & STM_SGM_OVERLAP) {
#define STM_SGM_FO(NV_, VW_)        \
  if (m.NV == NV_ && m.VW == VW_)   \
    rc = sgm_fused_overlap<NV_, VW_>(cost, image, out, H, W, D, p1, p2, flags, workspace, workspace_bytes, stream, disp);
      STM_SGM_FO(1, 4) STM_SGM_FO(1, 2) STM_SGM_FO(3, 2) STM_SGM_FO(1, 1) STM_SGM_FO(3, 1) STM_SGM_FO(5, 1)
      STM_SGM_FO(2, 1) STM_SGM_FO(4, 1)   // ranges that are not a multiple of 32: the sweeps mask a vector map
#undef STM_SGM_FO
      if (rc != 1) return rc;
    }
#define STM_SGM_FS(NV_, VW_)        \
  if (m.NV == NV_ && m.VW == VW_)   \
    rc = sgm_fused<NV_, VW_>(cost, image, out, H, W, D, p1, p2, flags, workspace, workspace_bytes, stream, disp);
    STM_SGM_FS(1, 4) STM_SGM_FS(2, 4) STM_SGM_FS(1, 2) STM_SGM_FS(3, 2) STM_SGM_FS(1, 1) STM_SGM_FS(3, 1) STM_SGM_FS(5, 1)
    STM_SGM_FS(2, 1) STM_SGM_FS(4, 1)
#undef STM_SGM_FS
    if (rc != 1) return rc;
  }
  float *ws = static_cast<float *>(workspace);
  {
    static const char *off = getenv("STM_DEBUG_SGM_CONCURRENT");  // developer knob: "0" disables
    // (per-pass timing measures the sequential passes, so it selects them)
    const size_t with_buffers = align256(sgm_base_workspace_bytes(H, W, D)) + (size_t)8 * H * W * D * sizeof(float);
    const bool buffers = sgm_concurrent_eligible(H, W, D) ||
                         (sgm_concurrent_mid_size(H, W, D) && workspace != nullptr && workspace_bytes >= with_buffers);
    if (buffers && !(flags & STM_SGM_SEQUENTIAL) && !g_pass_timing && !(off && off[0] == '0')) {
      float *bufs = reinterpret_cast<float *>(static_cast<char *>(workspace) +
                                              align256(sgm_base_workspace_bytes(H, W, D)));
#define STM_SGM_CC(NV_, VW_)          \
  if (m.NV == NV_ && m.VW == VW_)     \
    return sgm_concurrent<NV_, VW_>(cost, image, out, H, W, D, p1, p2, flags, bufs, stream, disp);
      STM_SGM_CC(1, 4) STM_SGM_CC(2, 4) STM_SGM_CC(3, 4) STM_SGM_CC(4, 4)
      STM_SGM_CC(1, 2) STM_SGM_CC(3, 2) STM_SGM_CC(5, 2) STM_SGM_CC(7, 2)
      STM_SGM_CC(1, 1) STM_SGM_CC(2, 1) STM_SGM_CC(3, 1) STM_SGM_CC(4, 1)
      STM_SGM_CC(5, 1) STM_SGM_CC(6, 1) STM_SGM_CC(7, 1) STM_SGM_CC(8, 1)
#undef STM_SGM_CC
    }
  }
#define STM_SGM_CASE(NV_, VW_)                                                            \
  if (m.NV == NV_ && m.VW == VW_)                                                         \
    return H > W ? sgm_by_direction_portrait<NV_, VW_>(cost, image, out, H, W, D, p1, p2,  \
                                                       flags, ws, stream, disp)           \
                 : sgm_by_direction<NV_, VW_>(cost, image, out, H, W, D, p1, p2, flags,    \
                                              ws, stream, disp);
  STM_SGM_CASE(1, 4) STM_SGM_CASE(2, 4) STM_SGM_CASE(3, 4) STM_SGM_CASE(4, 4)
  STM_SGM_CASE(1, 2) STM_SGM_CASE(3, 2) STM_SGM_CASE(5, 2) STM_SGM_CASE(7, 2)
  STM_SGM_CASE(1, 1) STM_SGM_CASE(2, 1) STM_SGM_CASE(3, 1) STM_SGM_CASE(4, 1)
  STM_SGM_CASE(5, 1) STM_SGM_CASE(6, 1) STM_SGM_CASE(7, 1) STM_SGM_CASE(8, 1)
#undef STM_SGM_CASE
  return fail(STM_EUNSUPPORTED, "sgm: D=%d not supported (D <= 512 if D %% 128 == 0, <= 448 if "
              "D %% 64 == 0, else <= 256)", D);
}

// ---- stm_sgm_static_f32: aggregation without any allocation by the caller or the library ------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The
// per-direction schedule's only scratch -- the L vectors of the corner up-diagonal that the
// reference re-runs W-H times (min(H, W) * Dpad floats) -- lives in this file-scope
// device array instead of a caller-provided workspace.  2 MiB per GPU that loads the library, held
// until exit.  Every call on a device shares it and it keeps the last call's contents: nothing in
// it has to start at zero (the store pass writes every element the replay reads, in stream order),
// but two calls that OVERLAP on one GPU (different streams / host threads) race on it.
constexpr size_t kSgmStaticScratchFloats = STM_SGM_STATIC_MAX_FLOATS;  // 2 MiB (include/stm_abi.h)
static __device__ __align__(256) float g_sgm_static_scratch[kSgmStaticScratchFloats];

// In host code `g_sgm_static_scratch` is a host-side stand-in, not a device address: the address
// is taken once per device, with that device current, and cached per ordinal.
static int sgm_static_scratch(float **out) {
  static std::mutex mu;
  static float *ptr[kMaxDevices] = {nullptr};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    cudaGetLastError();
    return fail(STM_ECUDA, "sgm_static: no usable CUDA device (libstm_b200 has no CPU fallback)");
  }
  std::lock_guard<std::mutex> lock(mu);
  if (ptr[dev] == nullptr) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_sgm_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      cudaGetLastError();
      return fail(STM_ECUDA, "Cuda call error cudaGetSymbolAddress(g_sgm_static_scratch): %s",
                  cudaGetErrorString(e));
    }
    ptr[dev] = static_cast<float *>(p);
  }
  *out = ptr[dev];
  return STM_OK;
}

}  // namespace stm

extern "C" int stm_sgm_static_f32(const float *cost, const float *left, float *out, int H, int W, int D,
                                  float penalty1, float penalty2, unsigned flags, void *stream) {
  using namespace stm;
  int rc = check_shape("sgm_static", H, W, D);
  if (rc) return rc;
  if (!cost || !left || !out) return fail(STM_EINVAL, "sgm_static: null pointer");
  if (cost == out) return fail(STM_EINVAL, "sgm_static: output must not alias the cost volume");
  if (flags & ~(STM_SGM_MIN_ALL_D | STM_SGM_OVERWRITE))
    return fail(STM_EINVAL, "sgm_static: unknown flags 0x%x (STM_SGM_MIN_ALL_D | STM_SGM_OVERWRITE only)", flags);
  if (D < 2) return fail(STM_EINVAL, "sgm_static: D must be >= 2");
  if (H > 32767 || W > 32767) return fail(STM_EINVAL, "sgm_static: H and W must be <= 32767");
  const LaneMap m = pick_lane_map(D);
  const size_t need = (size_t)(H < W ? H : W) * m.NV * 32 * m.VW;
  if (need > kSgmStaticScratchFloats)
    return fail(STM_EINVAL, "sgm_static: min(H, W) * padded D = %zu floats exceeds the built-in scratch of %zu "
                "floats (2 MiB); use stm_sgm_f32 with a workspace", need, kSgmStaticScratchFloats);
  float *ws = nullptr;
  if ((rc = sgm_static_scratch(&ws))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // the eight direction passes one after the other, in the reference's order: exactly what
  // stm_sgm_f32 runs with STM_SGM_SEQUENTIAL
#define STM_SGM_ST(NV_, VW_)                                                                                  \
  if (m.NV == NV_ && m.VW == VW_)                                                                             \
    return H > W ? sgm_by_direction_portrait<NV_, VW_>(cost, left, out, H, W, D, penalty1, penalty2, flags,   \
                                                       ws, st, nullptr)                                       \
                 : sgm_by_direction<NV_, VW_>(cost, left, out, H, W, D, penalty1, penalty2, flags, ws, st,    \
                                              nullptr);
  STM_SGM_ST(1, 4) STM_SGM_ST(2, 4) STM_SGM_ST(3, 4) STM_SGM_ST(4, 4)
  STM_SGM_ST(1, 2) STM_SGM_ST(3, 2) STM_SGM_ST(5, 2) STM_SGM_ST(7, 2)
  STM_SGM_ST(1, 1) STM_SGM_ST(2, 1) STM_SGM_ST(3, 1) STM_SGM_ST(4, 1)
  STM_SGM_ST(5, 1) STM_SGM_ST(6, 1) STM_SGM_ST(7, 1) STM_SGM_ST(8, 1)
#undef STM_SGM_ST
  return fail(STM_EUNSUPPORTED, "sgm_static: D=%d not supported", D);
}
