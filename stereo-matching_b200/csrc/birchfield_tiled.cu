// birchfield_tiled.cu -- Birchfield-Tomasi cost volume, one CTA per scanline (sm_100a).
//
// Replaces CostOps::ComputeBirchfield (reference src/birchfield_cost.cu:153-181; semantics of
// its CUDA functor, :83-139: scanlines zero-padded by one pixel, 8-tap 1-D window):
//   I-  = 0.5 (I[p] + I[p-1]),  I+ = 0.5 (I[p] + I[p+1]),  Imin/Imax = min/max(I-, I+, I)
//   e(p,d) = min( max(0, L[p]-Rmax[q], Rmin[q]-L[p]),  max(0, R[q]-Lmax[p], Lmin[p]-R[q]) ), q = p-d
//   C[y,x,d] = sum_{p = max(d, x-4)}^{min(W, x+4)-1} e(p,d)   (left to right)   for d <= x, else +inf
//
// The straightforward kernel (cost_generic.cu) re-evaluates every e(p,d) for each of the 8
// windows it belongs to, with ~48 global loads per output.  Here
//  * the six per-pixel arrays L, Lmin, Lmax, R, Rmin, Rmax of the scanline are computed once
//    into shared memory (the reference CUDA kernel stages the raw scanlines the same way, but
//    with blockDim.x = W it cannot run for W > 1024);
//  * a warp walks a run of consecutive x; lane l owns the disparities d = l + 32 j and keeps
//    the last 8 values e(x-4..x+3, d) in a rotating register window (static indices: the x
//    loop is unrolled by 8), so each e is evaluated ONCE: 3 broadcast + 3 conflict-free
//    shared loads and ~10 ALU operations, then 7 additions per output;
//  * each output row C[y,x,32j..32j+31] is one fully coalesced 128-byte streaming store.
// Taps outside [d, W) contribute an exact +0 in the same left-to-right order, so the result
// is bit-identical to the straightforward kernel (and to the oracle's restatement).
// HBM traffic = the 4 B/pix-disp store of C; the scanlines come from L2.
#include <cmath>
#include <cstdlib>
#include <type_traits>

#include "stm_common.cuh"

namespace stm {

template <int NJ>
__global__ void __launch_bounds__(256)
birchfield_row_kernel(const float *__restrict__ left, const float *__restrict__ right,
                      float *__restrict__ cost, int H, int W, int D) {
  extern __shared__ float sm[];
  float *sL = sm, *sLmi = sL + W, *sLma = sLmi + W;
  float *sR = sLma + W, *sRmi = sR + W, *sRma = sRmi + W;
  const int y = blockIdx.x;
  const float *lrow = left + (size_t)y * W, *rrow = right + (size_t)y * W;

  for (int p = threadIdx.x; p < W; p += blockDim.x) {
    const float li = __ldg(lrow + p), ri = __ldg(rrow + p);
    const float lm = p > 0 ? __ldg(lrow + p - 1) : 0.f, lp = p + 1 < W ? __ldg(lrow + p + 1) : 0.f;
    const float rm = p > 0 ? __ldg(rrow + p - 1) : 0.f, rp = p + 1 < W ? __ldg(rrow + p + 1) : 0.f;
    const float la = 0.5f * (li + lm), lb = 0.5f * (li + lp);
    const float ra = 0.5f * (ri + rm), rb = 0.5f * (ri + rp);
    sL[p] = li;
    sLmi[p] = fminf(la, fminf(lb, li));
    sLma[p] = fmaxf(la, fmaxf(lb, li));
    sR[p] = ri;
    sRmi[p] = fminf(ra, fminf(rb, ri));
    sRma[p] = fmaxf(ra, fmaxf(rb, ri));
  }
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int chunk = (((W + nw - 1) / nw) + 7) & ~7;  // multiple of 8: window slots stay static
  const int xa = warp * chunk;
  const int xb = min(W, xa + chunk);
  if (xa >= W) return;

  // e(p, d_j) for this lane's disparities; exact 0 for taps outside [d, W)
  auto eval = [&](int p, float (&out)[NJ]) {
    const bool pin = p >= 0 && p < W;  // warp-uniform
    const int pc = pin ? p : 0;
    const float li = sL[pc], lmi = sLmi[pc], lma = sLma[pc];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int q = p - (lane + 32 * j);
      const bool ok = pin && q >= 0;
      const int qc = ok ? q : 0;
      const float ri = sR[qc], rmi = sRmi[qc], rma = sRma[qc];
      const float a = fmaxf(0.f, fmaxf(li - rma, rmi - li));
      const float b = fmaxf(0.f, fmaxf(ri - lma, lmi - ri));
      out[j] = ok ? fminf(a, b) : 0.f;
    }
  };

  float win[8][NJ];  // win[p & 7][j] = e(p, d_j)
#pragma unroll
  for (int k = 0; k < 8; ++k)
#pragma unroll
    for (int j = 0; j < NJ; ++j) win[k][j] = 0.f;
  // taps xa-4 .. xa+2 (xa is a multiple of 8: p & 7 == (k + 4) & 7 for p = xa - 4 + k)
#pragma unroll
  for (int k = 0; k < 7; ++k) eval(xa - 4 + k, win[(k + 4) & 7]);

  float *orow = cost + ((size_t)y * W + xa) * D + lane;
  for (int x0 = xa; x0 < xb; x0 += 8) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int x = x0 + u;
      if (x < xb) {  // warp-uniform; no loads depend on it
        eval(x + 3, win[(u + 3) & 7]);
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          float s = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) s = __fadd_rn(s, win[(u + 4 + i) & 7][j]);  // p = x-4 .. x+3
          const int d = lane + 32 * j;
          if (d < D) __stcs(orow + (size_t)u * D + 32 * j, d > x ? CUDART_INF_F : s);
        }
      }
    }
    orow += (size_t)8 * D;
  }
}

// ---- packed variant (D a multiple of 64): same values, a third of the instructions ---------------
// ncu of the kernel above at 1080p D=128: 39 warp instructions per 128-byte store, issue slots 86 %
// busy -- it is bound by instruction issue, not by the store stream (0.36 ms where the stores alone
// need ~0.17).  Three changes, none of which touches a value:
//  * no predicates: both scanlines are padded in shared memory with SENTINEL pixels
//    (I = 0, Imin = -inf, -Imax = -inf).  A tap whose left pixel p lies outside [0, W) or whose
//    right pixel q = p - d is < 0 then evaluates to max(0, -inf, -inf) = +0 by itself, which is
//    exactly the "+0 for taps outside [d, W)" of the kernel above;
//  * one 16-byte shared load per tap: pixel records are float4 (I, I, -Imax, Imin), and the four
//    differences of a tap are two packed FMAs (I,I) * (1,-1) + (-Imax', Imin') = (I - Imax', Imin' - I)
//    (a multiplication by +-1 is exact, so each lane rounds once, like the subtraction it replaces);
//  * the window sums of two disparity groups (j, j+1) run as packed additions (add.rn.f32x2: the
//    rounding of two FADDs), in the same left-to-right order, starting from the first tap instead
//    of 0 + tap (identical: every tap is >= +0).
// 11 instead of 39 instructions per store.  d > x (+inf) only exists for x < D: those 8-column
// blocks take a variant with the select, the others do not test anything.
namespace bt2 {
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk(u64 r, float &a, float &b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r));
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
  u64 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
}  // namespace bt2

// WTA = true: the volume is never written.  The warp that produced a pixel's D costs holds them
// (4 per lane at D = 128), so it takes the argmin itself -- first (lowest-d) minimum, NaN never
// wins, NaN at d = 0 -> 0, the rule of wta.cu / reference src/winners_take_all.cu:21-40 -- and writes
// 4 bytes per pixel instead of 4 bytes per pixel-disparity (cost + reduce of a
// pipeline without aggregation in one kernel).
// ODD: the range is 32 * (NJ - 1) -- 96, 160, 224: the pairs are formed as for 32 * NJ and the last
// group of 32 disparities is computed but neither stored nor offered to the argmin (a quarter more
// arithmetic at D = 96 than needed, still a third faster than the unpacked kernel those ranges took).
template <int NJ, bool WTA, bool ODD = false>
__global__ void __launch_bounds__(256)
birchfield_row_packed_kernel(const float *__restrict__ left, const float *__restrict__ right,
                             float *__restrict__ cost, int32_t *__restrict__ disp, int H, int W, int part_w) {
  using namespace bt2;
  static_assert(NJ % 2 == 0, "disparity groups are processed in pairs");
  constexpr int D = 32 * NJ, NP = NJ / 2;
  constexpr int DR = ODD ? D - 32 : D;   // the range in memory
  extern __shared__ float4 smq[];
  float4 *sLq = smq;               // [W + 8]:      pixel p at index p + 4,      p in [-4, W + 3]
  float4 *sRq = smq + (W + 8);     // [W + D + 8]:  pixel q at index q + D + 4,  q in [-D - 4, W + 3]
  // A CTA owns the columns [X0, X1) of scanline y (part_w = W: the whole line).  Several parts per
  // line make the work units shorter: 1080 lines on 148 SMs are 7.3 lines per SM, i.e. 8 on the
  // busiest one (measured: 1184 lines take the time of 1080).
  const int y = blockIdx.x, X0 = blockIdx.y * part_w, X1 = min(W, X0 + part_w);
  const float *lrow = left + (size_t)y * W, *rrow = right + (size_t)y * W;
  const float4 sentinel = make_float4(0.f, 0.f, -CUDART_INF_F, -CUDART_INF_F);

  auto record = [&](const float *row, int p) {
    if (p < 0 || p >= W) return sentinel;
    const float v = __ldg(row + p);
    const float m = p > 0 ? __ldg(row + p - 1) : 0.f, n = p + 1 < W ? __ldg(row + p + 1) : 0.f;
    const float a = 0.5f * (v + m), b = 0.5f * (v + n);
    return make_float4(v, v, -fmaxf(a, fmaxf(b, v)), fminf(a, fminf(b, v)));
  };
  // only what the part reads: left pixels X0-4 .. X1+3, right pixels X0-D-4 .. X1+3 (same indices as
  // for a whole line)
  for (int i = X0 + threadIdx.x; i < X1 + 8; i += blockDim.x) sLq[i] = record(lrow, i - 4);
  for (int i = X0 + threadIdx.x; i < X1 + D + 8; i += blockDim.x) sRq[i] = record(rrow, i - D - 4);
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int chunk = (((X1 - X0 + nw - 1) / nw) + 7) & ~7;  // multiple of 8 (and so is X0): window slots stay static
  const int xa = X0 + warp * chunk;
  const int xb = min(X1, xa + chunk);
  if (xa >= X1) return;
  const u64 one_neg = pk(1.f, -1.f);
  const float4 *rbase = sRq + (D + 4 - lane);  // + p - 32 j

  // e(p, d) for d = lane + 32 j, pairs (j, j + 1) packed
  auto eval = [&](int p, u64 (&out)[NP]) {
    const float4 l = sLq[p + 4];
    const u64 ll = pk(l.x, l.y), lm = pk(l.z, l.w);
    const float4 *rp = rbase + p;
    float e[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const float4 r = rp[-32 * j];
      float t0, t1, u0, u1;
      upk(fma2(ll, one_neg, pk(r.z, r.w)), t0, t1);            // L - Rmax,  Rmin - L
      upk(fma2(pk(r.x, r.y), one_neg, lm), u0, u1);            // R - Lmax,  Lmin - R
      e[j] = fminf(fmaxf(0.f, fmaxf(t0, t1)), fmaxf(0.f, fmaxf(u0, u1)));
    }
#pragma unroll
    for (int q = 0; q < NP; ++q) out[q] = pk(e[2 * q], e[2 * q + 1]);
  };

  u64 win[8][NP];  // win[p & 7][j / 2] = e(p, d_j), e(p, d_j+1)
  // taps xa-4 .. xa+2 (xa is a multiple of 8: p & 7 == (k + 4) & 7 for p = xa - 4 + k)
#pragma unroll
  for (int k = 0; k < 7; ++k) eval(xa - 4 + k, win[(k + 4) & 7]);

  float *orow = WTA ? nullptr : cost + ((size_t)y * W + xa) * DR + lane;
  int32_t *drow = WTA ? disp + (size_t)y * W : nullptr;
  auto block8 = [&](auto edge_tag, auto tail_tag, int x0) {
    constexpr bool EDGE = decltype(edge_tag)::value, TAIL = decltype(tail_tag)::value;
    int mine = 0;  // WTA: lane u keeps the disparity of column x0 + u
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int x = x0 + u;
      if (!TAIL || x < xb) {  // warp-uniform
        eval(x + 3, win[(u + 3) & 7]);
        float best = CUDART_INF_F;  // WTA: this lane's minimum over its NJ disparities ...
        int best_d = 0x7fffffff;    // ... and the lowest d that attains it
        bool nan_at_0 = false;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
          u64 s = win[(u + 4) & 7][q];  // p = x-4, then x-3 .. x+3 left to right
#pragma unroll
          for (int i = 1; i < 8; ++i) s = add2(s, win[(u + 4 + i) & 7][q]);
          float s0, s1;
          upk(s, s0, s1);
          if (EDGE) {
            if (lane + 64 * q > x) s0 = CUDART_INF_F;
            if (lane + 64 * q + 32 > x) s1 = CUDART_INF_F;
          }
          if constexpr (!WTA) {
            __stcs(orow + u * DR + 64 * q, s0);
            if (!(ODD && q == NP - 1)) __stcs(orow + u * DR + 64 * q + 32, s1);
          } else {
            // ascending d within the lane, strict <: the lowest d of equal values sticks; NaN never passes
            if (q == 0) nan_at_0 = s0 != s0;
            if (s0 < best) { best = s0; best_d = lane + 64 * q; }
            if (!(ODD && q == NP - 1) && s1 < best) { best = s1; best_d = lane + 64 * q + 32; }
          }
        }
        if constexpr (WTA) {
          float m;  // warp minimum (redux.sync.min.f32 returns the non-NaN operand), then the lowest d equal to it
          asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(m) : "f"(best));
          unsigned idx = (best == m) ? (unsigned)best_d : 0x7fffffffu;
          idx = __reduce_min_sync(FULL, idx);
          const bool nan0 = __shfl_sync(FULL, nan_at_0 ? 1 : 0, 0) != 0;
          const int w = (nan0 || idx == 0x7fffffffu) ? 0 : (int)idx;
          if (lane == u) mine = w;
        }
      }
    }
    if constexpr (WTA) {
      if (lane < 8 && x0 + lane < xb) drow[x0 + lane] = mine;
    } else {
      orow += 8 * DR;
    }
  };
  int x0 = xa;
  for (; x0 + 8 <= xb && x0 < D; x0 += 8) block8(std::true_type{}, std::false_type{}, x0);
  for (; x0 + 8 <= xb; x0 += 8) block8(std::false_type{}, std::false_type{}, x0);
  if (x0 < xb) block8(std::true_type{}, std::true_type{}, x0);
}

// Parts per scanline.  The SMs take lines as they become free, so a launch ends when the busiest SM
// has done ceil(H / SMs) of them: the idle share is 1 - (H / SMs) / ceil(H / SMs) -- 9 % at 1080
// lines on 148 SMs, 3 % at 720 or 2160.  Two half lines per line halve the unit at the price of
// D + 8 more staged pixels per part.  Measured at 1920 columns (tools/stage_bench.py with
// STM_DEBUG_BT_PARTS = 1 / 2 / 3 / 4): 1080 lines D=128 0.228 / 0.223 / 0.223 / 0.229 ms, D=192
// 0.296 / 0.280 / 0.280 / 0.290; 720 lines 0.105 / 0.108; 2160 x 3840 0.671 / 0.716.  So: two parts
// where whole lines idle more than 6 % of the machine and half lines do better, else one.
static int bt_parts(int H, int W) {
  if (W < 256) return 1;
  const double sms = (double)num_sms();
  auto idle = [&](int units) { const double per = units / sms; return 1.0 - per / ceil(per); };
  return (idle(H) > 0.06 && idle(2 * H) < idle(H) - 0.03) ? 2 : 1;
}

template <int NJ, bool WTA, bool ODD = false>
static int run_birchfield_packed(const float *left, const float *right, float *cost, int32_t *disp, int H,
                                 int W, cudaStream_t stream) {
  const size_t smem = (size_t)(2 * W + 32 * NJ + 16) * sizeof(float4);
  auto kernel = birchfield_row_packed_kernel<NJ, WTA, ODD>;
  if (smem > 48 * 1024)  // per device and cheap: set on every call (stm_common.cuh, per-device state)
    STM_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int parts = bt_parts(H, W);
  {
    static const char *force = getenv("STM_DEBUG_BT_PARTS");  // developer knob
    if (force && atoi(force) > 0) parts = atoi(force);
  }
  int part_w = (((W + parts - 1) / parts) + 63) & ~63;   // 8 warps x 8 columns
  if (part_w > W) part_w = (W + 7) & ~7;
  const int ny = (W + part_w - 1) / part_w;
  if (ny > 65535) return 1;
  kernel<<<dim3(H, ny), 256, smem, stream>>>(left, right, cost, disp, H, W, part_w);
  return check_launch("birchfield_row_packed_kernel");
}

// Birchfield cost + winner-takes-all without the volume.  Returns 1 when the shape is not handled.
int launch_birchfield_wta(const float *left, const float *right, int32_t *disp, int H, int W, int D,
                          cudaStream_t stream) {
  if (D % 32 != 0 || W < 8 || (size_t)(2 * W + D + 48) * sizeof(float4) > 200 * 1024) return 1;
  switch (D / 32) {
    case 2: return run_birchfield_packed<2, true>(left, right, nullptr, disp, H, W, stream);
    case 3: return run_birchfield_packed<4, true, true>(left, right, nullptr, disp, H, W, stream);
    case 4: return run_birchfield_packed<4, true>(left, right, nullptr, disp, H, W, stream);
    case 5: return run_birchfield_packed<6, true, true>(left, right, nullptr, disp, H, W, stream);
    case 6: return run_birchfield_packed<6, true>(left, right, nullptr, disp, H, W, stream);
    case 7: return run_birchfield_packed<8, true, true>(left, right, nullptr, disp, H, W, stream);
    case 8: return run_birchfield_packed<8, true>(left, right, nullptr, disp, H, W, stream);
    default: return 1;
  }
}

template <int NJ>
static int run_birchfield_tiled(const float *left, const float *right, float *cost, int H, int W,
                                int D, cudaStream_t stream) {
  const size_t smem = (size_t)6 * W * sizeof(float);
  auto kernel = birchfield_row_kernel<NJ>;
  if (smem > 48 * 1024)
    STM_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kernel<<<H, 256, smem, stream>>>(left, right, cost, H, W, D);
  return check_launch("birchfield_row_kernel");
}

// Returns 1 when this shape is not handled here (caller runs the straightforward kernel).
int launch_birchfield_tiled(const float *left, const float *right, float *cost, int H, int W, int D,
                            cudaStream_t stream) {
  if ((size_t)6 * W * sizeof(float) > 200 * 1024 || W < 8) return 1;
  if (D % 32 == 0 && (size_t)(2 * W + D + 48) * sizeof(float4) <= 200 * 1024 &&
      !((reinterpret_cast<uintptr_t>(cost)) & 3)) {
    static const char *off = getenv("STM_DEBUG_BT_PACKED");  // developer knob: "0" = kernel above
    if (!(off && off[0] == '0')) switch (D / 32) {
      case 2: return run_birchfield_packed<2, false>(left, right, cost, nullptr, H, W, stream);
      case 3: return run_birchfield_packed<4, false, true>(left, right, cost, nullptr, H, W, stream);
      case 4: return run_birchfield_packed<4, false>(left, right, cost, nullptr, H, W, stream);
      case 5: return run_birchfield_packed<6, false, true>(left, right, cost, nullptr, H, W, stream);
      case 6: return run_birchfield_packed<6, false>(left, right, cost, nullptr, H, W, stream);
      case 7: return run_birchfield_packed<8, false, true>(left, right, cost, nullptr, H, W, stream);
      case 8: return run_birchfield_packed<8, false>(left, right, cost, nullptr, H, W, stream);
      default: break;
    }
  }
  switch ((D + 31) / 32) {
    case 1: return run_birchfield_tiled<1>(left, right, cost, H, W, D, stream);
    case 2: return run_birchfield_tiled<2>(left, right, cost, H, W, D, stream);
    case 3: return run_birchfield_tiled<3>(left, right, cost, H, W, D, stream);
    case 4: return run_birchfield_tiled<4>(left, right, cost, H, W, D, stream);
    case 5: return run_birchfield_tiled<5>(left, right, cost, H, W, D, stream);
    case 6: return run_birchfield_tiled<6>(left, right, cost, H, W, D, stream);
    case 7: return run_birchfield_tiled<7>(left, right, cost, H, W, D, stream);
    case 8: return run_birchfield_tiled<8>(left, right, cost, H, W, D, stream);
    default: return 1;
  }
}

}  // namespace stm
