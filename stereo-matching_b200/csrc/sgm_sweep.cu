// sgm_sweep.cu -- semiglobal aggregation as TWO fused raster sweeps (sm_100a).
//
// Replaces AggregationOps::RunSemiglobal (reference src/semiglobal.cpp:167-197; recurrence
// SGMCostOperator::operator(), :116-159) when stm_sgm_f32 is given STM_SGM_FUSED_SWEEPS.
// EXPERIMENTAL: correct (same parity tests as the default path) and it does cut HBM traffic to
// the designed 19.7 B/pix-disp (ncu), but with one warp per image row the wavefront is
// latency-bound: ~420 ns per pixel step and ~19 pixels of lag per row at 1080p, 20.6 ms per
// frame against 4.3 ms for the per-direction kernels (profiles/r01_v3_fused_sweeps.md).  The
// per-direction kernels in sgm.cu move ~92 B per pixel-disparity (8 x read C, read S,
// write S); here the forward sweep computes -> v \v /v for every pixel in one pass and the
// backward sweep <- ^ ^\ ^/ in a second one: read C twice, write S, read+write S = 20 B.
//
// Schedule.  In raster order pixel (y,x) depends on (y,x-1), (y-1,x-1), (y-1,x), (y-1,x+1):
// any schedule in which row y trails row y-1 by >= 2 pixels is valid.  One warp walks one
// image row left to right (disparities across lanes, lanevec.cuh); a CTA holds R
// consecutive rows.  Row y publishes its three downward-looking L vectors per pixel into
// a small shared-memory ring that row y+1 consumes (acquire/release flags in shared
// memory, ~2-3 pixels of lag).  Between CTAs the same vectors travel through an
// L2-resident ring in global memory moved by two helper warps per CTA (exporter:
// smem -> global + release; importer: cp.async global -> smem), so that every compute warp
// sees only shared memory and all rows advance at the same pace.  The backward sweep is
// the same kernel on mirrored coordinates (y -> H-1-y, x -> W-1-x).
//
// All CTAs must be co-resident (a ring producer waits for its consumer): the launch is
// cooperative and the host picks R so that the grid fits.  Every spin loop is bounded and
// traps instead of hanging.
//
// Reference path set: the two up-diagonal directions of the backward sweep
// carry the multiplicity rule of sgm_paths.cuh -- multiplicity 0 contributes nothing, and
// for the one diagonal with multiplicity m > 1 the L vectors are also stored so that
// sgm_repeat_add_kernel (sgm.cu) can add them the remaining m-1 times.
//
// Summation order differs from the reference's (it adds direction after direction); all
// terms are non-negative, so the result stays within a few ulp (bar: 1e-4 relative).
// The default mode of stm_sgm_f32 is the per-direction kernels of sgm.cu.
#include <cooperative_groups.h>

#include "lanevec.cuh"
#include "sgm_paths.cuh"

namespace stm {

constexpr int kRingK = 8;     // slots of a compute-warp -> compute-warp ring
constexpr int kRingK0 = 16;   // slots of the importer -> first-row ring (hides L2 latency)
constexpr int kGlobalK = 32;  // slots of a CTA -> CTA ring in global memory
constexpr unsigned kSpinCap = 1u << 22;

struct SweepParams {
  const float *cost;
  const float *image;
  float *out;
  int H, W, D;
  float p1, p2;
  int min_limit;
  unsigned *gprog;  // [grid] pixels exported by CTA j (read by CTA j+1)
  unsigned *gcons;  // [grid] pixels of CTA j's ring already imported by CTA j+1
  float *gring;     // [grid][kGlobalK][3][dpad]
  // backward sweep only: storage of the multiplicity > 1 diagonals (or null)
  float *lstore_c;  // dir 5 (-1,-1)
  float *lstore_e;  // dir 7 (+1,-1)
  int c_sx, e_sx;   // x of the first pixel of those paths
};

// ---- progress flags ---------------------------------------------------------------------
// Shared-memory flags (row -> row inside a CTA) are STRONG RELAXED accesses, i.e. plain
// LDS/STS with no fence.  st.release.cta would compile to MEMBAR.ALL.CTA + STS, and that
// barrier waits for every earlier memory operation of the thread -- including the global
// prefetch loads a compute warp keeps in flight -- which costs one DRAM round trip per
// pixel (measured: 28 ms instead of ~4 ms per frame).  Ordering of "data, then flag" relies
// on (a) __syncwarp() between the lanes' data stores and lane 0's flag store, (b) the "memory"
// clobbers (compiler order), and (c) the SM performing one warp's shared-memory requests in
// issue order through its single shared-memory pipeline: a reader that sees the new flag
// value therefore sees the data stored before it.  (c) is a property of the hardware, not
// of the PTX memory model; the GPU parity tests (bit-identical reruns included) guard it.
// Global-memory flags (CTA -> CTA, helper warps only) use real acquire/release at gpu scope.
__device__ __forceinline__ unsigned ld_flag_smem(const unsigned *p) {
  unsigned v;
  asm volatile("ld.relaxed.cta.shared.u32 %0, [%1];"
               : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_flag_smem(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.cta.shared.u32 [%0], %1;"
               :: "r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_gmem(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gmem(unsigned *p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// Waits until *flag >= need; returns the value seen.  Bounded: a protocol bug traps
// (loud launch failure) instead of hanging the GPU.  Short waits (the steady state of the
// wavefront) poll back to back; only long ones (a row waiting for the wavefront to arrive)
// sleep.
__device__ __noinline__ unsigned wait_smem_slow(const unsigned *flag, unsigned need) {
  unsigned v = ld_flag_smem(flag);
  unsigned tries = 0, ns = 32;
  while (v < need) {
    if (++tries > 64) {
      __nanosleep(ns);
      if (ns < 512) ns <<= 1;
      if (tries > kSpinCap) __trap();
    }
    v = ld_flag_smem(flag);
  }
  return v;
}
__device__ __forceinline__ unsigned wait_smem(const unsigned *flag, unsigned need) {
  const unsigned v = ld_flag_smem(flag);
  return v >= need ? v : wait_smem_slow(flag, need);
}
__device__ __forceinline__ unsigned wait_gmem(const unsigned *flag, unsigned need) {
  unsigned v = ld_acquire_gmem(flag);
  unsigned tries = 0, ns = 64;
  while (v < need) {
    __nanosleep(ns);
    if (ns < 512) ns <<= 1;
    v = ld_acquire_gmem(flag);
    if (++tries > kSpinCap) __trap();
  }
  return v;
}

// ---- shared-memory layout --------------------------------------------------------------
// ring i feeds compute warp i (i = 0: written by the importer, i >= 1: by warp i-1);
// ring R is written by warp R-1 and drained by the exporter.  A slot holds the three
// vectors {vertical, diagonal-from-left, diagonal-from-right} of one pixel, 3*dpad floats.
template <int DPAD, int R>
struct SweepSmem {
  static constexpr int kSlot = 3 * DPAD;
  static constexpr int kRingFloats = (kRingK0 + R * kRingK) * kSlot;
  static constexpr size_t kBytes = (size_t)kRingFloats * 4 + 2 * (R + 1) * sizeof(unsigned);
  __device__ static float *ring(float *base, int i) {
    return base + (i == 0 ? 0 : (kRingK0 + (i - 1) * kRingK) * kSlot);
  }
  __device__ static int mask(int i) { return i == 0 ? kRingK0 - 1 : kRingK - 1; }
  __device__ static unsigned *prog(float *base) {
    return reinterpret_cast<unsigned *>(base + kRingFloats);
  }
  __device__ static unsigned *cons(float *base) { return prog(base) + (R + 1); }
};

__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
               :: "r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// Importer: global ring of CTA j-1  ->  shared ring 0 (consumed by this CTA's first row).
template <int DPAD, int R>
__device__ void sweep_importer(float *smem, const SweepParams &P, int cta, int lane) {
  using S = SweepSmem<DPAD, R>;
  constexpr int n4 = S::kSlot / 4;  // float4 per slot
  float *ring0 = S::ring(smem, 0);
  unsigned *prog0 = S::prog(smem), *cons0 = S::cons(smem);
  const unsigned *gprog = P.gprog + (cta - 1);
  unsigned *gcons = P.gcons + (cta - 1);
  const float *gring = P.gring + (size_t)(cta - 1) * kGlobalK * S::kSlot;
  const unsigned W = P.W;
  unsigned done = 0;
  while (done < W) {
    unsigned avail = wait_gmem(gprog, done + 1);
    // slot of pixel q is free once the reader has finished step q - K0 (cons >= q-K0+1)
    const unsigned c = wait_smem(cons0, done + 1 > (unsigned)kRingK0 ? done + 1 - kRingK0 : 0);
    unsigned limit = avail < W ? avail : W;
    if (limit > c + kRingK0) limit = c + kRingK0;
    for (unsigned q = done; q < limit; ++q) {
      const float *src = gring + (size_t)(q & (kGlobalK - 1)) * S::kSlot;
      float *dst = ring0 + (q & (kRingK0 - 1)) * S::kSlot;
      for (int i = lane; i < n4; i += 32) cp_async16(dst + 4 * i, src + 4 * i);
    }
    cp_async_wait_all();
    __syncwarp();
    if (lane == 0) {
      st_flag_smem(prog0, limit);
      st_release_gmem(gcons, limit);
    }
    done = limit;
  }
}

// Exporter: shared ring R (written by this CTA's last row)  ->  global ring of CTA j.
template <int DPAD, int R>
__device__ void sweep_exporter(float *smem, const SweepParams &P, int cta, int lane) {
  using S = SweepSmem<DPAD, R>;
  constexpr int n4 = S::kSlot / 4;
  const float *ringR = S::ring(smem, R);
  const unsigned *progR = S::prog(smem) + R;
  unsigned *consR = S::cons(smem) + R;
  unsigned *gprog = P.gprog + cta;
  const unsigned *gcons = P.gcons + cta;
  float *gring = P.gring + (size_t)cta * kGlobalK * S::kSlot;
  const unsigned W = P.W;
  unsigned done = 0, gc = 0;
  while (done < W) {
    const unsigned avail = wait_smem(progR, done + 1);
    if (done + 1 > gc + kGlobalK) gc = wait_gmem(gcons, done + 1 - kGlobalK);
    unsigned limit = avail < W ? avail : W;
    if (limit > gc + kGlobalK) limit = gc + kGlobalK;
    for (unsigned q = done; q < limit; ++q) {
      const float4 *src =
          reinterpret_cast<const float4 *>(ringR + (q & (kRingK - 1)) * S::kSlot);
      float4 *dst = reinterpret_cast<float4 *>(gring + (size_t)(q & (kGlobalK - 1)) * S::kSlot);
      for (int i = lane; i < n4; i += 32) __stcg(dst + i, src[i]);
    }
    __syncwarp();
    if (lane == 0) {
      st_flag_smem(consR, limit);     // the shared slots may be overwritten
      st_release_gmem(gprog, limit);  // ... and the global ones may be read
    }
    done = limit;
  }
}

// ---- reference path-set multiplicities of the two up-diagonal directions (sgm_paths.cuh) --
__device__ __forceinline__ int sweep_mult_dir5(int x, int y, int H, int W) {
  const int c = x - y, cs = W - H;
  if (W > H) {
    if (c >= 0 && c < cs) return 0;
    if (c == cs) return cs + 1;
  } else if (H > W) {
    if (c > cs && c <= -1) return 0;
    if (c == cs) return -cs;
  }
  return 1;
}
__device__ __forceinline__ int sweep_mult_dir7(int x, int y, int H, int W) {
  if (W > H) {
    const int s = x + y;
    if (s >= H && s <= W - 1) return 0;
    if (s == H - 1) return W - H + 1;
  }
  return 1;
}

template <int NV, int VW>
__device__ __forceinline__ void lv_add(LaneVec<NV, VW> &a, const LaneVec<NV, VW> &b) {
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j) a.v[k][j] = __fadd_rn(a.v[k][j], b.v[k][j]);
}

// One sweep.  Logical coordinates (ly, lx) run in raster order; BACKWARD mirrors them onto
// the image (y = H-1-ly, x = W-1-lx).  Per pixel, in logical terms:
//   A: from (ly, lx-1)     forward ->  (dir 0)   backward <-  (dir 1)
//   B: from (ly-1, lx)     forward v   (dir 2)   backward ^   (dir 3)
//   C: from (ly-1, lx-1)   forward \v  (dir 4)   backward ^\  (dir 5)
//   E: from (ly-1, lx+1)   forward /v  (dir 6)   backward ^/  (dir 7)
template <int NV, int VW, int R, bool BACKWARD, bool ACC>
__global__ void __launch_bounds__((R + 2) * 32)
sgm_sweep_kernel(const SweepParams P) {
  constexpr int DPAD = NV * 32 * VW;
  using S = SweepSmem<DPAD, R>;
  using Vec = LaneVec<NV, VW>;
  extern __shared__ __align__(16) float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, cta = blockIdx.x;
  if (threadIdx.x < 2 * (R + 1)) S::prog(smem)[threadIdx.x] = 0;
  __syncthreads();
  const int H = P.H, W = P.W, D = P.D;
  if (warp == R) {
    if (cta > 0) sweep_importer<DPAD, R>(smem, P, cta, lane);
    return;
  }
  if (warp == R + 1) {
    if ((cta + 1) * R < H) sweep_exporter<DPAD, R>(smem, P, cta, lane);
    return;
  }
  const int ly = cta * R + warp;
  if (ly >= H) return;

  const bool has_above = ly > 0, has_below = ly + 1 < H;
  const int y = BACKWARD ? H - 1 - ly : ly;
  const int ya = has_above ? (BACKWARD ? y + 1 : y - 1) : y;
  const LaneCtx cx = lv_ctx<NV, VW>(lane, P.min_limit);
  const int lane_off = lane * VW;
  const int dl = D - lane_off;
  const float p1 = P.p1, p2 = P.p2;
  const long long stepd = BACKWARD ? -(long long)D : (long long)D;
  const long long base = ((long long)y * W + (BACKWARD ? W - 1 : 0)) * D + lane_off;
  const float *crow = P.cost + base;
  float *orow = P.out + base;
  const float *irow = P.image + (long long)y * W;
  const float *irow_a = P.image + (long long)ya * W;
  auto phys = [&](int l) { return BACKWARD ? W - 1 - l : l; };

  float *ring_in = S::ring(smem, warp) + lane_off;
  const int mask_in = S::mask(warp);
  float *ring_out = S::ring(smem, warp + 1) + lane_off;
  const unsigned *prog_in = S::prog(smem) + warp;
  unsigned *cons_in = S::cons(smem) + warp;
  unsigned *prog_out = S::prog(smem) + warp + 1;
  const unsigned *cons_out = S::cons(smem) + warp + 1;

  // C (and S for the read-modify-write) of the next PF pixels stay in flight in registers
  constexpr int PF = (NV * VW <= 4) ? 8 : 4;
  Vec cbuf[PF], sbuf[PF];
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    const long long off = (long long)(u < W ? u : W - 1) * stepd;
    lv_load<NV, VW, true>(cbuf[u], crow + off, 0, dl);
    if (ACC) lv_load<NV, VW, true>(sbuf[u], orow + off, 0, dl);
  }

  Vec LA = cbuf[0], aB = cbuf[0], aC = cbuf[0], aE = cbuf[0];
  unsigned seen_prog = 0, seen_cons = 0;
  if (has_above) {  // vectors of the row above for lx = 0: pixels 0 and 1
    seen_prog = wait_smem(prog_in, W < 2 ? W : 2);
    lv_load<NV, VW, false>(aB, ring_in, 0, DPAD);
    lv_load<NV, VW, false>(aE, ring_in + ((W > 1 ? 1 : 0) & mask_in) * S::kSlot + 2 * DPAD, 0, DPAD);
  }
  float bA = 0.f, bB = 0.f, bC = 0.f, bE = 0.f;  // lane l: adaptive P2 of pixel (lx & ~31) + l

  // adaptive penalties of 32 consecutive pixels, one pixel per lane (4 divisions per lane per
  // 32 pixels instead of 4 per pixel); broadcast by shuffle in step()
  auto penalties = [&](int lx0) {
    int q = lx0 + lane;
    if (q > W - 1) q = W - 1;
    const int ql = q > 0 ? q - 1 : 0, qr = q < W - 1 ? q + 1 : W - 1;
    const float i0 = __ldg(irow + phys(q));
    bA = sgm_p2_adjusted(p1, p2, i0, __ldg(irow + phys(ql)));
    bB = sgm_p2_adjusted(p1, p2, i0, __ldg(irow_a + phys(q)));
    bC = sgm_p2_adjusted(p1, p2, i0, __ldg(irow_a + phys(ql)));
    bE = sgm_p2_adjusted(p1, p2, i0, __ldg(irow_a + phys(qr)));
  };

  // One pixel.  Every load whose latency matters is issued UNCONDITIONALLY (clamped indices):
  // a load under a branch makes the compiler load into a temporary and copy at the join,
  // i.e. wait for it at once (that cost 5x in the first version of this kernel).
  auto step = [&](const int lx, const Vec &c, Vec acc) {
    // vectors of the row above for step lx+1: pixels lx, lx+1, lx+2 (row 0 reads unused smem)
    Vec nB, nC, nE;
    {
      const int l1 = lx + 1 < W ? lx + 1 : W - 1, l2 = lx + 2 < W ? lx + 2 : W - 1;
      const unsigned need = has_above ? (unsigned)(lx + 3 < W ? lx + 3 : W) : 0u;
      if (seen_prog < need) seen_prog = wait_smem(prog_in, need);
      lv_load<NV, VW, false>(nB, ring_in + (l1 & mask_in) * S::kSlot, 0, DPAD);
      lv_load<NV, VW, false>(nC, ring_in + (lx & mask_in) * S::kSlot + DPAD, 0, DPAD);
      lv_load<NV, VW, false>(nE, ring_in + (l2 & mask_in) * S::kSlot + 2 * DPAD, 0, DPAD);
    }
    const int bl = lx & 31;
    const float pA = __shfl_sync(FULL, bA, bl), pB = __shfl_sync(FULL, bB, bl);
    const float pC = __shfl_sync(FULL, bC, bl), pE = __shfl_sync(FULL, bE, bl);

    // four recurrences; a path that STARTS at this pixel takes L = C (select, no branch)
    Vec LB = aB, LC = aC, LE = aE;
    lv_sgm_step<NV, VW>(LA, c, p1, pA, cx);
    lv_sgm_step<NV, VW>(LB, c, p1, pB, cx);
    lv_sgm_step<NV, VW>(LC, c, p1, pC, cx);
    lv_sgm_step<NV, VW>(LE, c, p1, pE, cx);
    const bool sA = lx == 0, sB = !has_above, sC = !has_above || lx == 0,
               sE = !has_above || lx == W - 1;
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int j = 0; j < VW; ++j) {
        LA.v[k][j] = sA ? c.v[k][j] : LA.v[k][j];
        LB.v[k][j] = sB ? c.v[k][j] : LB.v[k][j];
        LC.v[k][j] = sC ? c.v[k][j] : LC.v[k][j];
        LE.v[k][j] = sE ? c.v[k][j] : LE.v[k][j];
      }

    // out (+)= A + B + C + E, one fp32 addition per direction like the reference
    if (ACC) lv_add<NV, VW>(acc, LA); else acc = LA;
    lv_add<NV, VW>(acc, LB);
    if (BACKWARD) {
      const int x = W - 1 - lx;
      const int mc = sweep_mult_dir5(x, y, H, W), me = sweep_mult_dir7(x, y, H, W);
      if (mc) lv_add<NV, VW>(acc, LC);
      if (me) lv_add<NV, VW>(acc, LE);
      if (mc > 1 && P.lstore_c)
        lv_store<NV, VW, false>(LC, P.lstore_c + (size_t)(P.c_sx - x) * DPAD + lane_off, 0, DPAD);
      if (me > 1 && P.lstore_e)
        lv_store<NV, VW, false>(LE, P.lstore_e + (size_t)(x - P.e_sx) * DPAD + lane_off, 0, DPAD);
    } else {
      lv_add<NV, VW>(acc, LC);
      lv_add<NV, VW>(acc, LE);
    }
    lv_store<NV, VW, true>(acc, orow + (long long)lx * stepd, 0, dl);

    if (has_below) {  // hand this pixel's downward-looking vectors to the next row
      const unsigned need = lx >= kRingK ? (unsigned)(lx - kRingK + 1) : 0u;
      if (seen_cons < need) seen_cons = wait_smem(cons_out, need);
      float *slot = ring_out + (lx & (kRingK - 1)) * S::kSlot;
      lv_store<NV, VW, false>(LB, slot, 0, DPAD);
      lv_store<NV, VW, false>(LC, slot + DPAD, 0, DPAD);
      lv_store<NV, VW, false>(LE, slot + 2 * DPAD, 0, DPAD);
    }
    __syncwarp();
    if (lane == 0) {
      if (has_below) st_flag_smem(prog_out, lx + 1);
      if (has_above) st_flag_smem(cons_in, lx + 1);
    }
    aB = nB; aC = nC; aE = nE;
  };

  // full groups of PF pixels: branch-free, the refill loads run PF pixels ahead
  int lx = 0;
  for (int g = W / PF; g > 0; --g) {
    if ((lx & 31) == 0) penalties(lx);
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const Vec c = cbuf[u];
      Vec acc = c;
      if (ACC) acc = sbuf[u];
      const long long off = (long long)(lx + PF < W ? lx + PF : W - 1) * stepd;
      lv_load<NV, VW, true>(cbuf[u], crow + off, 0, dl);
      if (ACC) lv_load<NV, VW, true>(sbuf[u], orow + off, 0, dl);
      step(lx, c, acc);
      ++lx;
    }
  }
  // tail: fewer than PF pixels, already sitting in cbuf / sbuf
  if (lx < W && (lx & 31) == 0) penalties(lx);
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    if (lx < W) {
      Vec acc = cbuf[u];
      if (ACC) acc = sbuf[u];
      step(lx, cbuf[u], acc);
      ++lx;
    }
  }
}

// ---- host side ---------------------------------------------------------------------------
int sgm_repeat_add(float *out, float *l_store, int H, int W, int D, int dpad, int dir, int path,
                   int times, cudaStream_t stream);  // sgm.cu

static size_t align16(size_t n) { return (n + 15) & ~(size_t)15; }

// Workspace: [lstore dir5][lstore dir7][global rings][flags: 2 sweeps x {gprog, gcons}],
// sized for the smallest R (most CTAs).  The first region doubles as the per-direction
// mode's l_store (sgm.cu).
struct SweepWorkspace {
  size_t lstore_bytes, ring_bytes, flag_bytes, total;
  int gmax;
};
static SweepWorkspace sweep_workspace(int H, int W, int D) {
  const LaneMap m = pick_lane_map(D);
  const size_t dpad = (size_t)m.NV * 32 * m.VW;
  SweepWorkspace w;
  w.gmax = (H + 3) / 4;
  w.lstore_bytes = align16((size_t)(H < W ? H : W) * dpad * sizeof(float));
  w.ring_bytes = align16((size_t)w.gmax * kGlobalK * 3 * dpad * sizeof(float));
  w.flag_bytes = align16((size_t)4 * w.gmax * sizeof(unsigned));
  w.total = 2 * w.lstore_bytes + w.ring_bytes + w.flag_bytes;
  return w;
}
size_t sgm_sweep_workspace_bytes(int H, int W, int D) { return sweep_workspace(H, W, D).total; }

template <int NV, int VW, int R, bool BACKWARD, bool ACC>
static int launch_one_sweep(const SweepParams &P, int grid, cudaStream_t stream, bool dry_run,
                            bool *fits) {
  constexpr int DPAD = NV * 32 * VW;
  const size_t smem = SweepSmem<DPAD, R>::kBytes;
  auto kernel = sgm_sweep_kernel<NV, VW, R, BACKWARD, ACC>;
  STM_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (dry_run) {
    int per_sm = 0;
    STM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, (R + 2) * 32, smem));
    *fits = (long long)per_sm * num_sms() >= grid;
    return STM_OK;
  }
  SweepParams p = P;
  void *args[] = {&p};
  STM_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)kernel, dim3(grid), dim3((R + 2) * 32), args,
                                           smem, stream));
  return check_launch("sgm_sweep_kernel");
}

// Returns 1 when this configuration cannot run with all CTAs co-resident.
template <int NV, int VW, int R>
static int run_sweeps(const float *cost, const float *image, float *out, int H, int W, int D,
                      float p1, float p2, unsigned flags, char *ws, cudaStream_t stream) {
  constexpr int DPAD = NV * 32 * VW;
  if (SweepSmem<DPAD, R>::kBytes > 227 * 1024) return 1;
  const int grid = (H + R - 1) / R;
  const SweepWorkspace w = sweep_workspace(H, W, D);
  SweepParams P;
  P.cost = cost; P.image = image; P.out = out;
  P.H = H; P.W = W; P.D = D; P.p1 = p1; P.p2 = p2;
  P.min_limit = (flags & STM_SGM_MIN_ALL_D) ? D : D - 1;
  P.gring = reinterpret_cast<float *>(ws + 2 * w.lstore_bytes);
  unsigned *flagbase = reinterpret_cast<unsigned *>(ws + 2 * w.lstore_bytes + w.ring_bytes);
  P.lstore_c = P.lstore_e = nullptr;
  P.c_sx = P.e_sx = 0;
  const bool overwrite = (flags & STM_SGM_OVERWRITE) != 0;

  bool fits = false;
  int rc;
  P.gprog = flagbase; P.gcons = flagbase + w.gmax;
  // co-residency check on the heavier (read-modify-write) variant
  if ((rc = launch_one_sweep<NV, VW, R, true, true>(P, grid, stream, true, &fits))) return rc;
  if (!fits) return 1;
  if (!overwrite) {
    if ((rc = launch_one_sweep<NV, VW, R, false, true>(P, grid, stream, true, &fits))) return rc;
    if (!fits) return 1;
  }

  STM_CUDA_TRY(cudaMemsetAsync(flagbase, 0, w.flag_bytes, stream));
  rc = overwrite ? launch_one_sweep<NV, VW, R, false, false>(P, grid, stream, false, nullptr)
                 : launch_one_sweep<NV, VW, R, false, true>(P, grid, stream, false, nullptr);
  if (rc) return rc;

  P.gprog = flagbase + 2 * w.gmax; P.gcons = flagbase + 3 * w.gmax;
  const int m5 = sgm_multi_path_index(5, H, W), m7 = sgm_multi_path_index(7, H, W);
  PathGeom g5{}, g7{};
  if (m5 >= 0) { g5 = sgm_path_geom(5, m5, H, W); if (g5.mult > 1) { P.lstore_c = reinterpret_cast<float *>(ws); P.c_sx = g5.sx; } }
  if (m7 >= 0) { g7 = sgm_path_geom(7, m7, H, W); if (g7.mult > 1) { P.lstore_e = reinterpret_cast<float *>(ws + w.lstore_bytes); P.e_sx = g7.sx; } }
  if ((rc = launch_one_sweep<NV, VW, R, true, true>(P, grid, stream, false, nullptr))) return rc;
  if (P.lstore_c && (rc = sgm_repeat_add(out, P.lstore_c, H, W, D, DPAD, 5, m5, g5.mult - 1, stream))) return rc;
  if (P.lstore_e && (rc = sgm_repeat_add(out, P.lstore_e, H, W, D, DPAD, 7, m7, g7.mult - 1, stream))) return rc;
  return STM_OK;
}

template <int NV, int VW>
static int run_sweeps_pick_rows(const float *cost, const float *image, float *out, int H, int W,
                                int D, float p1, float p2, unsigned flags, char *ws,
                                cudaStream_t stream) {
  // few rows per CTA spread the active band of the wavefront over more SMs; more rows per CTA
  // are needed when the image has more rows than co-resident CTAs x 4
  int rc = run_sweeps<NV, VW, 4>(cost, image, out, H, W, D, p1, p2, flags, ws, stream);
  if (rc != 1) return rc;
  return run_sweeps<NV, VW, 8>(cost, image, out, H, W, D, p1, p2, flags, ws, stream);
}

// Returns 1 when the fused sweeps do not handle this shape (caller falls back to the
// per-direction kernels).
int launch_sgm_sweeps(const float *cost, const float *image, float *out, int H, int W, int D,
                      float p1, float p2, unsigned flags, void *workspace, size_t workspace_bytes,
                      cudaStream_t stream) {
  if (workspace == nullptr || workspace_bytes < sgm_sweep_workspace_bytes(H, W, D))
    return fail(STM_EINVAL, "sgm: workspace too small (%zu < %zu)", workspace_bytes,
                sgm_sweep_workspace_bytes(H, W, D));
  if ((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(out) |
       reinterpret_cast<uintptr_t>(workspace)) & 15)
    return 1;
  const LaneMap m = pick_lane_map(D);
  char *ws = static_cast<char *>(workspace);
#define STM_SWEEP_CASE(NV_, VW_) \
  if (m.NV == NV_ && m.VW == VW_) \
    return run_sweeps_pick_rows<NV_, VW_>(cost, image, out, H, W, D, p1, p2, flags, ws, stream);
  STM_SWEEP_CASE(1, 4) STM_SWEEP_CASE(2, 4) STM_SWEEP_CASE(1, 2) STM_SWEEP_CASE(3, 2)
  STM_SWEEP_CASE(1, 1)
#undef STM_SWEEP_CASE
  return 1;
}

}  // namespace stm
