// sgm_vsweep.cu -- three SGM directions per pass: the fused vertical sweeps.
//
// The six directions with a vertical component (reference src/semiglobal.cpp:18-91: v ^ \v ^\
// /v ^/) are two groups of three that advance through the image rows together.  A pass per
// direction reads the cost volume and read-modify-writes the sum volume once per direction
// (12 B/pix-disp each); a sweep does it once for three directions.
//
// Decomposition (what the first two fused designs, profiles/r01_v3_fused_sweeps.md, lacked:
// thousands of independent chains AND on-chip summation).  Sweep step t is image row
// y = t (down: directions 2, 4, 6) or y = H-1-t (up: 3, 5, 7).  Warp `u` is a SKEWED column
// slot: at step t it owns pixel x = (u + t) mod W, so
//   * the diagonal whose predecessor is (x-1, previous row)  [4 down, 7 up]  stays in the
//     warp's registers,
//   * the vertical direction's predecessor (x) belongs to slot u+1, the other diagonal's
//     (x+1) [6 down, 5 up] to slot u+2 -- both at the PREVIOUS step.
// Every hand-off therefore points the same way (u <- u+1, u+2): a pipeline, not a lock-step
// wavefront; slots slip against each other by the hand-off latency instead of waiting for
// it every row, and there is no chain along a row.  A slot that leaves the image on the right
// re-enters on the left, where the register-held diagonal starts a new path (L = C); the
// u+2 diagonal starts a new path at x = W-1.  All W slots do exactly H steps.
//
// CTA c owns `ncols` (2..13) consecutive slots, one warp each, plus an importer and an exporter
// warp.  Inside a CTA a slot publishes its vertical / u+2-diagonal vectors to 4-row rings in
// shared memory behind a (full, empty) mbarrier pair per slot and ring row; a waiting warp is
// suspended by try_wait (spinning neighbours took the issue slots of the warp everybody waits
// for).  Between CTAs the exporter warp writes slots 0 and 1's vectors as (value, row tag) pairs
// in single 8-byte words into an L2-resident ring (the tag travels with the data: no fence, no
// separate flag) and the previous CTA's importer warp polls them a few rows ahead and feeds two
// virtual slots of the shared rings, so column warps see one interface.  Credits flow back
// through the `empty` barriers and a per-CTA imported-rows counter that the importer mirrors
// into shared memory.  Importer and exporter are warps of their own because a .gpu-scope load or
// store issued by a warp with cp.async prefetches in flight was measured at 4-6 us instead of
// 0.35 us.  All CTAs must be co-resident: cooperative launch.
//
// Cost and sum vectors of the next rows are in flight as per-lane cp.async copies into a
// private shared-memory ring (a lane only reads back bytes it copied itself: no barrier).  The
// adaptive penalties max(P1, P2/|dI|) of both sweeps come from a pre-kernel (one float4 per
// pixel and sweep) and are fetched with the row.
//
// Values: same recurrences and same per-direction vectors as sgm.cu; the sum is associated
// (((S + v) + d1) + d2) per sweep instead of the reference's 0..7 order, and the corner
// up-diagonals add mult*L in one fused multiply-add instead of mult additions
// -- within the 1e-4 bar (measured 1.3e-5), not bit-identical: callers that want the reference
// order bit for bit leave STM_SGM_FUSED_SWEEPS off.  Measurements and the designs that were
// tried and removed: profiles/r01_v26_fused_vertical_sweeps.md, DESIGN.md 3.4.
#include "lanevec.cuh"
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <type_traits>

#include "sgm_paths.cuh"

namespace stm {
namespace vsweep {

constexpr int kCols = 13;      // column slots (compute warps) per CTA; 13 x 148 >= 1920
constexpr int kR = 4;          // rows in the shared hand-off rings
constexpr int kRX = 16;        // rows in the CTA->CTA ring
#ifndef STM_VSWEEP_PF
#define STM_VSWEEP_PF 5
#endif
constexpr int kPfSmall = STM_VSWEEP_PF;
#ifndef STM_VSWEEP_LOOK
#define STM_VSWEEP_LOOK 1
#endif
constexpr int kImporterLook = STM_VSWEEP_LOOK;  // 1: measured, see importer_core
constexpr int kLookMax = 4;       // rows of that ring the importer keeps loads in flight for
constexpr int kFlagStride = 32;  // unsigned per counter: counters of different CTAs never share a line
constexpr int kMaxCtas = 320;  // workspace is sized for this many (pure function of the shape): two per SM on 160 SMs
constexpr int kThreads = (kCols + 2) * 32;  // column slots + importer + exporter
constexpr int kExporterBox = kCols + 1;      // barrier index of the exporter's inbox (never a consumer slot)
// A loop that polls GLOBAL memory (the importer on its neighbour's ring, the exporter on its credit)
// and makes no progress for `Params::spin_ns` (60 s) of wall clock GIVES UP instead of hanging the
// GPU (a bug, not a slow neighbour: profilers and time-slicing can stall a CTA for a long time, so
// the guard is on %globaltimer, looked at every 4096 polls, not on an iteration count).  Giving up
// is cooperative, not a __trap() (which would take the host application's whole CUDA context with
// it): the warp raises an abort word in the workspace (cleared by every launch), records where it
// was in a host-mapped status block, and from then on DRAINS -- the importer stops looking at the
// ring and feeds its CTA whatever is there, the exporter stops waiting for credit.  Every other
// global polling loop of the launch sees the word and drains too.  Nothing inside a CTA depends on
// another CTA any more, the column warps (whose waits are untouched: they are the hot path) run
// their remaining rows at full speed on garbage, and the kernel ENDS with an unspecified result
// that the host is told about (stm_sgm_sweep_aborted; the next sweep launch on the device fails
// with the record).
constexpr unsigned long long kSpinSeconds = 60;
constexpr int kAbortWord = 1;  // index into Params::xcons: a spare word of CTA 0's counter line
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#ifdef STM_VSWEEP_DEBUG
constexpr bool kDebug = true;   // STM_DEBUG_VSWEEP knobs, statistics and trace compiled in
#else
constexpr bool kDebug = false;
#endif

// NS = volume vectors prefetched per pixel: 1 = cost only (the sweep OVERWRITES `out` with its three
// directions: no sum is read), 2 = cost + sum (read-modify-write of `out`), 3 = cost + sum + one more
// volume `extra` that is added too (the other sweep's result, when the two sweeps did not chain
// through `out`; see sgm.cu, STM_SGM_OVERLAP), 4 = and a second one, `extra2` (the <- pass's own volume).
template <int DPAD, int NS = 2>
struct Cfg {
  static constexpr int PF = (DPAD <= 128 || (DPAD <= 192 && NS <= 2)) ? kPfSmall : 4;  // prefetch ring entries per slot
  static constexpr int kBarOff = 0;                                    // (full, empty)[kCols + 2][kR]
  static constexpr int kCreditOff = 1008;                              // unsigned: rows the previous CTA has imported
  static constexpr int kRingBOff = 1024;                               // [kCols+1][kR][DPAD]
  static constexpr int kRingCOff = kRingBOff + (kCols + 1) * kR * DPAD * 4;  // [kCols+2][kR][DPAD]
  static constexpr int kPfOff = kRingCOff + (kCols + 2) * kR * DPAD * 4;     // [kCols][PF][NS][DPAD]
  static constexpr int kPfiOff = kPfOff + kCols * PF * NS * DPAD * 4;        // [kCols][PF] float4 penalties
  static constexpr int kBytes = kPfiOff + kCols * PF * 16;
  static constexpr bool kFits = kBytes <= 227 * 1024;
};

struct Params {
  const float *cost;
  const float4 *pen;          // [H][W] adaptive penalties (A, B, C, -) of this sweep, see sgm_vsweep_penalties
  float *out;
  const float *extra;         // NS >= 3: a second sum volume that is added (read only)
  const float *extra2;        // NS == 4: a third one
  int32_t *disp;
  unsigned long long *xring;  // [G][kRX][3][DPAD] (value, tag) words
  unsigned *xcons;            // [G][kFlagStride] rows imported so far by CTA c (one 128-byte line each)
  int H, W, D, G, q, r;       // W = q*G + r: CTAs [0, r) own q+1 slots, the others q
  int up;                     // 0: directions 2,4,6 top to bottom; 1: 3,5,7 bottom to top
  int store_vol, emit_wta, defect, min_limit;
  int dbg;  // developer knob STM_DEBUG_VSWEEP: 1 = importer does not poll, 2 = slots do not wait (wrong results)
  float p1, p2;
  unsigned long long spin_ns;      // a polling loop without progress for this long gives up
  volatile unsigned *host_status;  // host-mapped {gave up, CTA, row, site}, or nullptr
  int fault_cta, fault_row;        // stm_debug_vsweep_fault: that CTA's exporter stops at that row (-1: none)
};

// Developer statistics (STM_DEBUG_VSWEEP & 8), read by stm_debug_vsweep_stats: clock64 cycles.
__device__ unsigned long long g_stats[64];
// Trace of one CTA (STM_DEBUG_VSWEEP & 16): clock64 of [warp][row - kTraceRow0][event].
constexpr int kTraceRow0 = 500, kTraceRows = 24, kTraceCta = 5;
__device__ long long g_trace[16][kTraceRows][4];
__device__ __forceinline__ void stat_add(int i, unsigned long long v) { atomicAdd(&g_stats[i], v); }

// ---- flags ----------------------------------------------------------------------------------
// Hand-offs inside a CTA: one (full, empty) mbarrier pair per slot and ring row.  A waiting warp
// is suspended by the hardware (try_wait) instead of spinning: spinning neighbours were measured
// to take the issue slots of the one warp everybody is waiting for.
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}" :: "r"(bar), "r"(parity) : "memory");
}
// One bounded try (the hardware gives up after a while): for waits that must do something else
// while they wait.
__device__ __forceinline__ bool mbar_try(unsigned bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
// barrier addresses for slot s (0..kCols+1) and ring row e.  `full` is the INBOX of consumer s:
// it completes when both of its inputs for that row are there (the vertical vector of slot s+1
// and the diagonal of slot s+2: 2 x 32 arrivals, from one or two producers), so a consumer waits
// once per row; index kExporterBox is the exporter's inbox (slots 0 and 1).  `empty` belongs to
// producer s: its readers are done with ring row e.
template <int R = kR>
__device__ __forceinline__ unsigned bar_full(unsigned base, int s, int e) { return base + (unsigned)((s * R + e) * 16); }
template <int R = kR>
__device__ __forceinline__ unsigned bar_empty(unsigned base, int s, int e) { return base + (unsigned)((s * R + e) * 16 + 8); }
// readers of slot s's ring rows: slot s-1 takes the vertical vector, slot s-2 the diagonal
// and the exporter warp takes both from slots 0 and 1
__device__ __forceinline__ int slot_readers(int s, int ncols) {
  return ((s - 1 >= 0 && s - 1 < ncols) ? 1 : 0) + ((s - 2 >= 0 && s - 2 < ncols) ? 1 : 0) + (s < 2 ? 1 : 0);
}
__device__ __forceinline__ unsigned ld_flag_s(const unsigned *p) {
  unsigned v;
  asm volatile("ld.relaxed.cta.shared.u32 %0, [%1];"
               : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_flag_s(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.cta.shared.u32 [%0], %1;"
               :: "r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_g(const unsigned *p) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_g(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// Guard of a loop that polls GLOBAL memory.  poll() is true when the loop must stop waiting: somebody
// else has given up, or this loop has seen no progress for P.spin_ns (then it raises the abort
// word; the first to do so also writes the host-visible record).
// (kept inline: a __noinline__ slow part measured 3 % slower on the Pipeline path, by-value 6 %; these
// kernels are sensitive to where cold code sits)
struct SpinGuard {
  unsigned polls = 0;
  unsigned long long t0 = 0;
  __device__ __forceinline__ bool poll(const Params &P, int site, int row) {
    if ((++polls & 4095u) != 0) return false;
    if (ld_relaxed_g(P.xcons + kAbortWord) != 0) return true;
    const unsigned long long now = global_ns();
    if (t0 == 0) { t0 = now; return false; }
    if (now - t0 <= P.spin_ns) return false;
    st_relaxed_g(P.xcons + kAbortWord, 1u);
    if (P.host_status != nullptr && (threadIdx.x & 31) == 0 && atomicExch(P.xcons + kAbortWord + 1, 1u) == 0u) {
      P.host_status[1] = blockIdx.x; P.host_status[2] = (unsigned)row; P.host_status[3] = (unsigned)site;
      __threadfence_system();
      P.host_status[0] = 1u;
      __threadfence_system();
    }
    return true;
  }
};
// The recurrence with packed additions where the lane map allows it (this kernel is bound by
// issue slots; the per-direction kernels are not and keep lv_sgm_step).
template <int NV, int VW>
__device__ __forceinline__ void sgm_step(LaneVec<NV, VW> &L, const LaneVec<NV, VW> &c, float p1, float p2_adj,
                                         const LaneCtx &cx) {
  if constexpr (VW % 2 == 0) lv_sgm_step_packed<NV, VW>(L, c, p1, p2_adj, cx);
  else lv_sgm_step<NV, VW>(L, c, p1, p2_adj, cx);
}

// ---- (value, tag) words ---------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ll_pack(float v, unsigned tag) {
  return ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
}
template <int NV, int VW>
__device__ __forceinline__ void lv_st_ll(const LaneVec<NV, VW> &a, unsigned long long *p, int lane,
                                         unsigned tag) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    unsigned long long *q = p + k * 32 * VW + lane * VW;
    if constexpr (VW % 2 == 0) {
#pragma unroll
      for (int j = 0; j < VW; j += 2)
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};"
                     :: "l"(q + j), "l"(ll_pack(a.v[k][j], tag)), "l"(ll_pack(a.v[k][j + 1], tag))
                     : "memory");
    } else {
      asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(q), "l"(ll_pack(a.v[k][0], tag))
                   : "memory");
    }
  }
}
template <int NV, int VW>
__device__ __forceinline__ void lv_ld_ll(unsigned long long (&w)[NV * VW], const unsigned long long *p,
                                         int lane) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const unsigned long long *q = p + k * 32 * VW + lane * VW;
    if constexpr (VW % 2 == 0) {
#pragma unroll
      for (int j = 0; j < VW; j += 2)
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];"
                     : "=l"(w[k * VW + j]), "=l"(w[k * VW + j + 1]) : "l"(q + j) : "memory");
    } else {
      asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w[k * VW]) : "l"(q) : "memory");
    }
  }
}

// ---- shared-memory vectors ------------------------------------------------------------------
template <int NV, int VW>
__device__ __forceinline__ void lv_lds(LaneVec<NV, VW> &r, const float *p, int lane) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const float *q = p + k * 32 * VW + lane * VW;
    if constexpr (VW == 4) {
      const float4 t = *reinterpret_cast<const float4 *>(q);
      r.v[k][0] = t.x; r.v[k][1] = t.y; r.v[k][2] = t.z; r.v[k][3] = t.w;
    } else if constexpr (VW == 2) {
      const float2 t = *reinterpret_cast<const float2 *>(q);
      r.v[k][0] = t.x; r.v[k][1] = t.y;
    } else {
      r.v[k][0] = *q;
    }
  }
}
template <int NV, int VW>
__device__ __forceinline__ void lv_sts(const LaneVec<NV, VW> &r, float *p, int lane) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    float *q = p + k * 32 * VW + lane * VW;
    if constexpr (VW == 4) *reinterpret_cast<float4 *>(q) = make_float4(r.v[k][0], r.v[k][1], r.v[k][2], r.v[k][3]);
    else if constexpr (VW == 2) *reinterpret_cast<float2 *>(q) = make_float2(r.v[k][0], r.v[k][1]);
    else *q = r.v[k][0];
  }
}

template <int BYTES>
__device__ __forceinline__ void cp_async(float *smem_dst, const float *gmem_src) {
  if constexpr (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                 :: "r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;"
                 :: "r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src), "n"(BYTES)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory");
}

// ---- importer: next CTA's slots 0 and 1 -> virtual slots ncols, ncols+1 ----------------------
template <int NV, int VW, int R, int LOOK>
__device__ void importer_core(const unsigned bars, float *ringB, float *ringC, unsigned *xcred_dst,
                              const Params &P, int ncols, int lane) {
  constexpr int DPAD = NV * 32 * VW, NE = NV * VW;
  const int nrows = P.H - 1;  // the last row's vectors have no consumer
  if (nrows <= 0) return;
  const int c = blockIdx.x;
  const int nb = c + 1 == P.G ? 0 : c + 1;
  const unsigned long long *src = P.xring + (size_t)nb * kRX * 3 * DPAD;
  float *dst[3] = {ringB + (size_t)ncols * R * DPAD, ringC + (size_t)ncols * R * DPAD,
                   ringC + (size_t)(ncols + 1) * R * DPAD};

  // rows of (value, tag) words in flight, bounded by registers (3 vectors x NE words per lane each)
  constexpr int kLook = LOOK;
  unsigned long long buf[kLook][3][NE];
  auto load = [&](unsigned long long (&w)[3][NE], int row) {
    const unsigned long long *s = src + (size_t)(row & (kRX - 1)) * 3 * DPAD;
#pragma unroll
    for (int v = 0; v < 3; ++v) lv_ld_ll<NV, VW>(w[v], s + v * DPAD, lane);
  };
  auto ready = [&](const unsigned long long (&w)[3][NE], int row) {
    bool ok = true;
#pragma unroll
    for (int v = 0; v < 3; ++v)
#pragma unroll
      for (int e = 0; e < NE; ++e) ok = ok && (unsigned)(w[v][e] >> 32) == (unsigned)(row + 1);
    return __all_sync(FULL, ok) != 0;
  };

  const bool poll = !kDebug || !(P.dbg & 1), stats = kDebug && (P.dbg & 8) != 0;
  bool drain = false;  // the launch is giving up (SpinGuard)
  const bool tracing = kDebug && (P.dbg & 16) != 0;
  unsigned long long st_polls = 0, st_ready = 0, st_credit = 0, st_rtt = 0;
  // kLook rows of (value, tag) words in flight: the L2 round trip is longer than a row
#pragma unroll
  for (int i = 0; i < kLook; ++i) load(buf[i], i < nrows ? i : nrows - 1);
#pragma unroll 1
  for (int r0 = 0; r0 < nrows; r0 += kLook) {
#pragma unroll
    for (int i = 0; i < kLook; ++i) {
      const int r = r0 + i;
      if (r < nrows) {
        const int rr = r & (R - 1);
        long long c0 = (stats || tracing) ? clock64() : 0;
        if (r >= R && poll) {  // the readers of ring row r-R are done with it
          const unsigned par = (unsigned)(((r / R) - 1) & 1);
          mbar_wait(bar_empty<R>(bars, ncols, rr), par);
          mbar_wait(bar_empty<R>(bars, ncols + 1, rr), par);
        }
        long long c1 = (stats || tracing) ? clock64() : 0;
        unsigned tries = 0;
        SpinGuard guard;
        while (poll && !drain && !ready(buf[i], r)) {
          ++tries;
          if (guard.poll(P, 0, r)) { drain = true; break; }  // give up: feed the CTA what is there
          const long long r0 = stats ? clock64() : 0;
          load(buf[i], r);
          if (stats) { (void)ready(buf[i], r); st_rtt += clock64() - r0; }
        }
        if (stats) { st_credit += c1 - c0; st_ready += clock64() - c1; st_polls += tries; }
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          LaneVec<NV, VW> t;
#pragma unroll
          for (int k = 0; k < NV; ++k)
#pragma unroll
            for (int j = 0; j < VW; ++j) t.v[k][j] = __uint_as_float((unsigned)buf[i][v][k * VW + j]);
          lv_sts<NV, VW>(t, dst[v] + rr * DPAD, lane);
        }
        // vertical of slot ncols and diagonal of slot ncols+1 go to slot ncols-1, the diagonal
        // of slot ncols to slot ncols-2
        mbar_arrive(bar_full<R>(bars, ncols - 1, rr));
        mbar_arrive(bar_full<R>(bars, ncols - 1, rr));
        mbar_arrive(bar_full<R>(bars, ncols - 2, rr));
        if (tracing && c == kTraceCta && lane == 0 && r >= kTraceRow0 && r < kTraceRow0 + kTraceRows) {
          g_trace[15][r - kTraceRow0][0] = c1; g_trace[15][r - kTraceRow0][1] = clock64();
        }
        __syncwarp();
        if (lane == 0) st_relaxed_g(&P.xcons[(size_t)c * kFlagStride], (unsigned)(r + 1));
        const int rn = r + kLook;
        load(buf[i], rn < nrows ? rn : nrows - 1);
      }
    }
  }
  if (stats && lane == 0) {
    stat_add(0, (unsigned long long)nrows); stat_add(1, st_polls); stat_add(2, st_ready);
    stat_add(3, st_credit); stat_add(4, st_rtt);
  }
}

template <int NV, int VW>
__device__ void importer_warp(unsigned char *smem, const Params &P, int ncols, int lane) {
  constexpr int DPAD = NV * 32 * VW, NE = NV * VW;
  using C = Cfg<DPAD>;
  // rows of (value, tag) words in flight, bounded by registers (3 vectors x NE words per lane each)
  constexpr int LOOK = kImporterLook;
  importer_core<NV, VW, kR, LOOK>((unsigned)__cvta_generic_to_shared(smem + C::kBarOff),
                                  reinterpret_cast<float *>(smem + C::kRingBOff),
                                  reinterpret_cast<float *>(smem + C::kRingCOff),
                                  reinterpret_cast<unsigned *>(smem + C::kCreditOff), P, ncols, lane);
}

// ---- exporter: slots 0 and 1 of this CTA -> the previous CTA ---------------------------------
// A warp of its own, with no prefetch in flight: the .gpu-scope stores (and loads) of a column
// warp were measured to take ~5 us behind its outstanding cp.async traffic, the same accesses
// from a warp without any ~0.4 us -- and that latency is paid once per CTA on the ring.
template <int NV, int VW>
__device__ void exporter_warp(unsigned char *smem, const Params &P, int lane) {
  constexpr int DPAD = NV * 32 * VW;
  using C = Cfg<DPAD>;
  const int nrows = P.H - 1;  // the last row's vectors have no consumer
  const unsigned bars = (unsigned)__cvta_generic_to_shared(smem + C::kBarOff);
  // rows the previous CTA's importer has taken out of this CTA's ring (the exporter has no
  // other global traffic, so the read costs one L2 round trip, every ~14 rows)
  const unsigned *xcredit = &P.xcons[(size_t)(blockIdx.x == 0 ? P.G - 1 : blockIdx.x - 1) * kFlagStride];
  const float *ringB = reinterpret_cast<const float *>(smem + C::kRingBOff);
  const float *ringC = reinterpret_cast<const float *>(smem + C::kRingCOff);
  unsigned long long *xdst = P.xring + (size_t)blockIdx.x * kRX * 3 * DPAD;
  unsigned xc_seen = 0;
  bool drain = false;  // the launch is giving up (SpinGuard): no more waiting for credit
  const int fault_row = (int)blockIdx.x == P.fault_cta ? P.fault_row : 0x7fffffff;  // stm_debug_vsweep_fault
#pragma unroll 1
  for (int t = 0; t < nrows; ++t) {
    const int e = t & (kR - 1);
    const unsigned par = (unsigned)((t / kR) & 1);
    mbar_wait(bar_full(bars, kExporterBox, e), par);  // slots 0 and 1 have published row t
    LaneVec<NV, VW> B0, C0, C1;
    lv_lds<NV, VW>(B0, ringB + e * DPAD, lane);
    lv_lds<NV, VW>(C0, ringC + e * DPAD, lane);
    lv_lds<NV, VW>(C1, ringC + (kR + e) * DPAD, lane);
    mbar_arrive(bar_empty(bars, 0, e));
    mbar_arrive(bar_empty(bars, 1, e));
    if (t >= kRX) {  // the previous CTA has imported row t-kRX
      const unsigned need_x = (unsigned)(t - kRX + 1);
      for (SpinGuard guard; xc_seen < need_x && !drain; drain = guard.poll(P, 1, t)) xc_seen = ld_relaxed_g(xcredit);
    }
    if (t >= fault_row) continue;  // injected fault: this CTA stops exporting (its neighbour starves)
    unsigned long long *ex = xdst + (size_t)(t & (kRX - 1)) * 3 * DPAD;
    lv_st_ll<NV, VW>(B0, ex, lane, (unsigned)(t + 1));
    lv_st_ll<NV, VW>(C0, ex + DPAD, lane, (unsigned)(t + 1));
    lv_st_ll<NV, VW>(C1, ex + 2 * DPAD, lane, (unsigned)(t + 1));
  }
}

// ---- one column slot ------------------------------------------------------------------------
template <int NV, int VW>
__device__ __forceinline__ void add_weighted(LaneVec<NV, VW> &o, const LaneVec<NV, VW> &L, int mult) {
  if (mult == 1) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      if constexpr (VW % 2 == 0) {
#pragma unroll
        for (int j = 0; j < VW; j += 2) f2_add(o.v[k][j], o.v[k][j + 1], o.v[k][j], o.v[k][j + 1], L.v[k][j], L.v[k][j + 1]);
      } else {
#pragma unroll
        for (int j = 0; j < VW; ++j) o.v[k][j] = __fadd_rn(o.v[k][j], L.v[k][j]);
      }
    }
  } else if (mult != 0) {
    const float w = (float)mult;
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int j = 0; j < VW; ++j) o.v[k][j] = __fmaf_rn(w, L.v[k][j], o.v[k][j]);
  }
}

// o += w * L for the defective up-diagonals, w in {0, 1, W-H+1} warp-uniform: w = 0 must leave o alone
// even where L is +inf (0 * inf), hence the select; fma(1, L, o) is o + L with one rounding, exactly the
// packed addition, and fma(w, L, o) is what the branchy version computed for the corner diagonal.
template <int NV, int VW>
__device__ __forceinline__ void add_scaled(LaneVec<NV, VW> &o, const LaneVec<NV, VW> &L, float w) {
  const bool skip = w == 0.f;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    if constexpr (VW % 2 == 0) {
#pragma unroll
      for (int j = 0; j < VW; j += 2) {  // fma.rn.f32x2: the rounding of two FFMAs
        unsigned long long a, b, c, d;
        asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(w), "f"(w));
        asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(skip ? 0.f : L.v[k][j]), "f"(skip ? 0.f : L.v[k][j + 1]));
        asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(o.v[k][j]), "f"(o.v[k][j + 1]));
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(o.v[k][j]), "=f"(o.v[k][j + 1]) : "l"(d));
      }
    } else {
#pragma unroll
      for (int j = 0; j < VW; ++j) o.v[k][j] = __fmaf_rn(w, skip ? 0.f : L.v[k][j], o.v[k][j]);
    }
  }
}

// MASKED: D < DPAD with whole lane vectors beyond D (D % VW == 0): those lanes never load or store and
// hold +inf costs -- exactly the reference's out-of-range neighbour value -- so a range such as
// D = 96 or 100 runs the 4-wide lane map of D = 128 instead of scalar lanes.  VW == 1 maps always
// mask (any D <= DPAD).
template <int NV, int VW, bool UP, bool WTA, int NS, bool MASKED>
__device__ void column_warp(unsigned char *smem, const Params &P, int ncols, int base, int j, int lane) {
  constexpr int DPAD = NV * 32 * VW;
  constexpr bool kMask = MASKED || VW == 1;
  using C = Cfg<DPAD, NS>;
  constexpr int PF = C::PF;
  const int H = P.H, W = P.W, D = P.D;
  constexpr bool up = UP;
  const float p1 = P.p1;
  const LaneCtx cx = lv_ctx<NV, VW>(lane, P.min_limit);
  const int lane_off = lane * VW;
  // shared memory, all pre-offset by this slot and lane
  const unsigned bars = (unsigned)__cvta_generic_to_shared(smem + C::kBarOff);
  float *ringB = reinterpret_cast<float *>(smem + C::kRingBOff) + (size_t)j * kR * DPAD + lane_off;
  float *ringC = reinterpret_cast<float *>(smem + C::kRingCOff) + (size_t)j * kR * DPAD + lane_off;
  float *pf = reinterpret_cast<float *>(smem + C::kPfOff) + (size_t)j * PF * NS * DPAD + lane_off;
  float4 *pfi = reinterpret_cast<float4 *>(smem + C::kPfiOff) + j * PF;  // lane 0 copies, all read

  const long long rowstep = up ? -(long long)W : (long long)W;
  int x = base + j, y = up ? H - 1 : 0;
  // fetch side: row pointers (lane offset included) + column
  const float *crow = P.cost + (long long)y * W * D + lane_off;
  const float *srow = P.out + (long long)y * W * D + lane_off;
  const float *xrow = (NS >= 3 ? P.extra : P.out) + (long long)y * W * D + lane_off;
  const float *zrow = (NS == 4 ? P.extra2 : P.out) + (long long)y * W * D + lane_off;
  const float4 *prow = P.pen + (long long)y * W;
  const long long rowstep_d = rowstep * D;
  int fx = x, ft = 0;

  // row ft: cost, sum and penalties into ring entry `e`; always exactly one commit group.
  // Past the last row the last row is fetched again (never used).
  auto issue = [&](int e) {
    const float *cs = crow + fx * D, *ss = srow + fx * D, *xs = xrow + fx * D, *zs = zrow + fx * D;
    float *dc = pf + e * NS * DPAD;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int off = k * 32 * VW;
      if ((!kMask || off + lane * VW < D) && !(kDebug && (P.dbg & 32))) {
        cp_async<VW * 4>(dc + off, cs + off);
        if constexpr (NS >= 2) cp_async<VW * 4>(dc + DPAD + off, ss + off);
        if constexpr (NS >= 3) cp_async<VW * 4>(dc + 2 * DPAD + off, xs + off);
        if constexpr (NS == 4) cp_async<VW * 4>(dc + 3 * DPAD + off, zs + off);
      }
    }
    if (lane == 0) cp_async<16>(reinterpret_cast<float *>(pfi + e), reinterpret_cast<const float *>(prow + fx));
    cp_async_commit();
    if (ft + 1 < H) {
      ++ft;
      prow += rowstep; crow += rowstep_d; srow += rowstep_d; xrow += rowstep_d; zrow += rowstep_d;
      fx = fx + 1 == W ? 0 : fx + 1;
    }
  };
#pragma unroll 1
  for (int i = 0; i < PF; ++i) issue(i);

  LaneVec<NV, VW> LA;
  float *orow = P.out + (long long)y * W * D + lane_off;
  int32_t *drow = P.disp + (long long)y * W;
  const bool sync = !kDebug || !(P.dbg & 2), stats = kDebug && (P.dbg & 8) != 0;
  unsigned long long st_dep = 0, st_cred = 0, st_g = 0, st_gn = 0, st_mem = 0;
  const long long st_begin = kDebug ? clock64() : 0;
  const bool store_vol = !WTA || P.store_vol != 0, defect = UP && P.defect != 0;
  constexpr bool emit_wta = WTA;
  int e = 0;

  // penalties of the row about to be processed, computed one row ahead (off the hand-off path);
  // pw: the up-diagonal multiplicities of that row's pixel, (mult5 << 16) | mult7 (first row: closed form)
  float pa = 0.f, pb = 0.f, pc = 0.f;
  unsigned pw = UP ? (((unsigned)sgm_mult5(x, y, H, W) << 16) | (unsigned)sgm_mult7(x, y, H, W)) : 0u;

  const bool trace = kDebug && (P.dbg & 16) && blockIdx.x == kTraceCta && lane == 0;
  auto mark = [&](int t, int ev) {
    if (trace && t >= kTraceRow0 && t < kTraceRow0 + kTraceRows) g_trace[j][t - kTraceRow0][ev] = clock64();
  };

  auto row = [&](auto first_tag, int t) {
    constexpr bool FIRST = decltype(first_tag)::value;
    {
      const long long c0 = stats ? clock64() : 0;
      cp_async_wait<PF - 2>();  // rows t and t+1 have landed
      if (stats) st_mem += clock64() - c0;
    }
    LaneVec<NV, VW> c, o;
    lv_lds<NV, VW>(c, pf + e * NS * DPAD, 0);
    if constexpr (NS >= 2) {
      lv_lds<NV, VW>(o, pf + e * NS * DPAD + DPAD, 0);
    } else {
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int q = 0; q < VW; ++q) o.v[k][q] = 0.f;  // this sweep starts the sum (0 + L is exact)
    }
    if constexpr (NS >= 3) {
#pragma unroll
      for (int extra_i = 2; extra_i < NS; ++extra_i) {
        LaneVec<NV, VW> x2;
        lv_lds<NV, VW>(x2, pf + e * NS * DPAD + extra_i * DPAD, 0);
        if constexpr (kMask) {
#pragma unroll
          for (int k = 0; k < NV; ++k)
            if (k * 32 * VW + lane * VW >= D) {
#pragma unroll
              for (int q = 0; q < VW; ++q) x2.v[k][q] = 0.f;
            }
        }
        add_weighted<NV, VW>(o, x2, 1);
      }
    }
    if constexpr (kMask) {
#pragma unroll
      for (int k = 0; k < NV; ++k)
        if (k * 32 * VW + lane * VW >= D) {
#pragma unroll
          for (int q = 0; q < VW; ++q) { c.v[k][q] = CUDART_INF_F; o.v[k][q] = 0.f; }
        }
    }
    const int en = e + 1 == PF ? 0 : e + 1;
    __syncwarp();                  // lane 0's penalty copy of row t+1 is visible to the warp
    const float4 pn = pfi[en];     // penalties of row t+1
    const int e_refill = e;
    e = en;

    // ring row t of this slot must be free (its readers consumed row t-kR): checked here, off
    // the hand-off path -- it has normally been true for a while
    {
      const int ew = t & (kR - 1);
      if (t >= kR && sync) {  // ring row t-kR has been read by slots j-1, j-2
        const long long c0 = stats ? clock64() : 0;
        mbar_wait(bar_empty(bars, j, ew), (unsigned)(((t >> 2) - 1) & 1));
        if (stats) st_cred += clock64() - c0;
      }
    }
    LaneVec<NV, VW> LB, LC;
    if constexpr (FIRST) {
      LB = c; LC = c; LA = c;
    } else {
      const int er = (t - 1) & (kR - 1);
      const unsigned parr = (unsigned)(((t - 1) >> 2) & 1);
      mark(t, 0);
      if (sync) {
        const long long c0 = stats ? clock64() : 0;
        mbar_wait(bar_full(bars, j, er), parr);  // both inputs of row t are there
        if (stats) st_dep += clock64() - c0;
      }
      mark(t, 1);
      lv_lds<NV, VW>(LB, ringB + kR * DPAD + er * DPAD, 0);
      lv_lds<NV, VW>(LC, ringC + 2 * kR * DPAD + er * DPAD, 0);
      mbar_arrive(bar_empty(bars, j + 1, er));
      mbar_arrive(bar_empty(bars, j + 2, er));
      // the two recurrences the neighbours wait for, interleaved; a path start is a select
      sgm_step<NV, VW>(LB, c, p1, pb, cx);
      sgm_step<NV, VW>(LC, c, p1, pc, cx);
      const bool startC = x == W - 1;
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int q = 0; q < VW; ++q) LC.v[k][q] = startC ? c.v[k][q] : LC.v[k][q];
    }

    {  // hand the vertical and the u+2 diagonal to slots j-1, j-2 (the last row's go nowhere)
      const int ew = t & (kR - 1);
      lv_sts<NV, VW>(LB, ringB + ew * DPAD, 0);
      lv_sts<NV, VW>(LC, ringC + ew * DPAD, 0);
      // into the inbox of each reader: slot j-1 (vertical), slot j-2 (diagonal), the exporter
      if (j >= 1) mbar_arrive(bar_full(bars, j - 1, ew));
      if (j >= 2) mbar_arrive(bar_full(bars, j - 2, ew));
      if (j < 2) mbar_arrive(bar_full(bars, kExporterBox, ew));
      mark(t, 2);
    }

    // Off the hand-off path: the register-held diagonal (nobody's input), the refill of the
    // prefetch entry just consumed with row t + PF, then this row's sum / store / argmin.
    if constexpr (!FIRST) {
      sgm_step<NV, VW>(LA, c, p1, pa, cx);
      const bool startA = x == 0;
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int q = 0; q < VW; ++q) LA.v[k][q] = startA ? c.v[k][q] : LA.v[k][q];
    }
    asm volatile("" ::: "memory");
    issue(e_refill);
    pa = pn.x; pb = pn.y; pc = pn.z;
    const unsigned w_now = pw;
    pw = __float_as_uint(pn.w);

    // sum in the reference's direction order: down 2, 4, 6 = B, A, C; up 3, 5, 7 = B, C, A
    add_weighted<NV, VW>(o, LB, 1);
    if (!up) {
      add_weighted<NV, VW>(o, LA, 1);
      add_weighted<NV, VW>(o, LC, 1);
    } else if (!defect) {
      add_weighted<NV, VW>(o, LC, 1);
      add_weighted<NV, VW>(o, LA, 1);
    } else {
      add_scaled<NV, VW>(o, LC, (float)(w_now >> 16));
      add_scaled<NV, VW>(o, LA, (float)(w_now & 0xffffu));
    }
    static_assert(!MASKED || NV == 1, "masked vector maps have one vector per lane");
    if (store_vol && (!MASKED || lane_off < D) && !(kDebug && (P.dbg & 32)))
      lv_store<NV, VW, true>(o, orow + x * D, 0, D - lane_off);
    if (emit_wta) lv_argmin_store<NV, VW>(o, lane, D, drow + x);

    mark(t, 3);
    x = x + 1 == W ? 0 : x + 1;
    y += up ? -1 : 1;
    orow += rowstep_d; drow += rowstep;
  };

  row(std::true_type{}, 0);
#pragma unroll 1
  for (int t = 1; t < H; ++t) row(std::false_type{}, t);
  cp_async_wait<0>();
  if (stats && lane == 0) {
    stat_add(8 + j, st_dep); stat_add(24 + j, st_cred); stat_add(40, st_g); stat_add(41, st_gn);
    stat_add(42, (unsigned long long)(clock64() - st_begin)); stat_add(43, 1ull); stat_add(44, st_mem);
  }
}

// MINB = 2: the same kernel compiled for two CTAs per SM (64 registers), for images wider than
// kCols columns per SM; only for the lane maps with at most two disparities per lane (D <= 64).
template <int NV, int VW, bool UP, bool WTA, int NS, int MINB = 1, bool MASKED = false>
__global__ void __launch_bounds__(kThreads, MINB) sgm_vsweep_kernel(const Params P) {
  extern __shared__ __align__(128) unsigned char smem[];
  using C = Cfg<NV * 32 * VW, NS>;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int c = blockIdx.x;
  const int ncols = P.q + (c < P.r ? 1 : 0);
  const int base = c * P.q + (c < P.r ? c : P.r);
  if (threadIdx.x < (kCols + 2) * kR) {
    const unsigned bars = (unsigned)__cvta_generic_to_shared(smem + C::kBarOff);
    const int sl = threadIdx.x / kR, e = threadIdx.x % kR;
    const int readers = slot_readers(sl, ncols);
    mbar_init(bar_full(bars, sl, e), 64u);                             // two inputs x every lane of their writer
    mbar_init(bar_empty(bars, sl, e), readers ? 32u * readers : 1u);   // every lane of each reader
  }
  if (threadIdx.x == 0) {
    *reinterpret_cast<unsigned *>(smem + C::kCreditOff) = 0u;
  }
  __syncthreads();
  if (wid == kCols) importer_warp<NV, VW>(smem, P, ncols, lane);
  else if (wid == kCols + 1) exporter_warp<NV, VW>(smem, P, lane);
  else if (wid < ncols) column_warp<NV, VW, UP, WTA, NS, MASKED>(smem, P, ncols, base, wid, lane);
}

static size_t align256(size_t n) { return (n + 255) & ~(size_t)255; }

// Adaptive penalties max(P1, P2/|dI|) (reference src/semiglobal.cpp:134-138) of both sweeps, once
// per call instead of a division and three shuffles per pixel inside the sweep: pen[up][y][x] =
// (A, B, C, 0) for the pixel behind (x, y) at x-1, x, x+1 in row y-1 (down) / y+1 (up).  Entries
// whose predecessor lies outside the image are never used (the path starts there).
__global__ void sgm_vsweep_penalties(const float *__restrict__ image, float4 *__restrict__ pen,
                                     int H, int W, float p1, float p2) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= H * W) return;
  const int y = i / W, x = i - y * W;
  const float v = image[i];
  const int xl = x > 0 ? x - 1 : x, xr = x < W - 1 ? x + 1 : x;
#pragma unroll
  for (int up = 0; up < 2; ++up) {
    const int yp = up ? y + 1 : y - 1;
    float4 r = make_float4(p1, p1, p1, 0.f);
    if (yp >= 0 && yp < H) {
      const float *q = image + (size_t)yp * W;
      r.x = sgm_p2_adjusted(p1, p2, v, q[xl]);
      r.y = sgm_p2_adjusted(p1, p2, v, q[x]);
      r.z = sgm_p2_adjusted(p1, p2, v, q[xr]);
    }
    // up sweep: how often the reference runs the up-diagonals through this pixel (0, 1 or W-H+1,
    // SURVEY.md F5), packed as (mult5 << 16) | mult7 -- both <= 32767 -- so that the sweep does
    // not evaluate the closed forms per pixel
    if (up) r.w = __uint_as_float(((unsigned)sgm_mult5(x, y, H, W) << 16) | (unsigned)sgm_mult7(x, y, H, W));
    pen[(size_t)up * H * W + i] = r;
  }
}
static size_t flags_and_rings_bytes(size_t dpad) {
  return align256(kMaxCtas * kFlagStride * sizeof(unsigned)) +
         align256((size_t)kMaxCtas * kRX * 3 * dpad * sizeof(unsigned long long));
}

// The six variants of one lane map: (up?, argmin?, volume streams).  nullptr = does not fit in
// shared memory for this lane map (three streams at D = 256).
template <int NV, int VW, bool UP, bool WTA, int NS, bool MASKED>
static const void *variant(size_t *smem_bytes, int minb = 1) {
  using C = Cfg<NV * 32 * VW, NS>;
  if constexpr (C::kFits) {
    *smem_bytes = (size_t)C::kBytes;
    if (minb == 2) {
      if constexpr (NV * VW <= 2 && 2 * C::kBytes + 2048 <= 227 * 1024)
        return (const void *)sgm_vsweep_kernel<NV, VW, UP, WTA, NS, 2, MASKED>;
      else
        return nullptr;
    }
    return (const void *)sgm_vsweep_kernel<NV, VW, UP, WTA, NS, 1, MASKED>;
  } else {
    *smem_bytes = 0;
    return nullptr;
  }
}
template <int NV, int VW, bool M = false>
static const void *sweep_kernel(bool up, bool wta, int ns, size_t *smem_bytes, int minb = 1) {
  size_t *b = smem_bytes;
  if (!up) return ns == 1 ? variant<NV, VW, false, false, 1, M>(b, minb) : variant<NV, VW, false, false, 2, M>(b, minb);
  if (ns == 4) return wta ? variant<NV, VW, true, true, 4, M>(b, minb) : nullptr;  // Pipeline path only
  if (ns == 3) return wta ? variant<NV, VW, true, true, 3, M>(b, minb) : variant<NV, VW, true, false, 3, M>(b, minb);
  return wta ? variant<NV, VW, true, true, 2, M>(b, minb) : variant<NV, VW, true, false, 2, M>(b, minb);
}

// CTAs for an image W columns wide (all co-resident, 2..kCols slots each), or 0.  group 0: the plain
// pair of sweeps (NS = 2); 1: the overwriting down sweep and the three-stream up sweep as well; 2: and
// the four-stream one.  *minb = CTAs per SM the kernels must be compiled for: 1, or 2 when the image
// is wider than kCols columns per SM and the lane map has a two-CTA build (D <= 64).
template <int NV, int VW, bool M = false>
static int ctas(int W, int group, int *minb_out) {
  static PerDeviceInt cache[3][2];  // per instantiation and device (attribute opt-in + occupancy)
  const int dev = device_slot();
  for (int minb = 1; minb <= 2; ++minb) {
    int cap = 0;
    if (!cache[group][minb - 1].get(dev, &cap)) {
      int per_sm = 1 << 30;
      for (int v = 0; v < 7; ++v) {
        const int ns = v < 3 ? 2 : (v == 3 ? 1 : (v == 6 ? 4 : 3));
        if (group < 1 && ns != 2) continue;
        if (ns == 4 && group < 2) continue;
        size_t bytes = 0;
        const void *k = sweep_kernel<NV, VW, M>(v != 0 && v != 3, v == 2 || v == 5 || v == 6, ns, &bytes, minb);
        int n = 0;
        if (k == nullptr) { per_sm = 0; break; }
        if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess ||
            (minb == 2 && cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout,
                                               cudaSharedmemCarveoutMaxShared) != cudaSuccess) ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k, kThreads, bytes) != cudaSuccess) {
          cudaGetLastError();
          return 0;
        }
        per_sm = n < per_sm ? n : per_sm;
      }
      per_sm = per_sm >= minb ? minb : 0;  // exactly minb CTAs per SM: the rings are sized for it
      cap = per_sm * num_sms();
      cache[group][minb - 1].set(dev, cap);
    }
    int G = cap < kMaxCtas ? cap : kMaxCtas;
    if (G > W / 2) G = W / 2;
    if (G >= 1 && (W + G - 1) / G <= kCols) {
      *minb_out = minb;
      return G;
    }
  }
  return 0;
}

// ---- give-up status (SpinGuard): one host-mapped block of four words per device ----------------
struct SweepStatus {
  std::mutex mu;
  unsigned *host = nullptr;  // {gave up, CTA, row, site}; pinned and mapped, written by the kernel
  unsigned *dev = nullptr;
  bool tried = false;
};
static SweepStatus g_status[kMaxDevices];
static int g_fault_cta = -1, g_fault_row = 0;             // stm_debug_vsweep_fault
static unsigned long long g_spin_ns = kSpinSeconds * 1000000000ull;

// The current device's block (allocated on first use; without it the kernel still gives up, but the
// host is not told), or nullptr.
static SweepStatus *sweep_status() {
  const int dev = device_slot();
  if (dev < 0) return nullptr;
  SweepStatus &st = g_status[dev];
  std::lock_guard<std::mutex> lock(st.mu);
  if (!st.tried) {
    st.tried = true;
    void *h = nullptr, *d = nullptr;
    if (cudaHostAlloc(&h, 64, cudaHostAllocMapped) == cudaSuccess &&
        cudaHostGetDevicePointer(&d, h, 0) == cudaSuccess) {
      st.host = static_cast<unsigned *>(h);
      st.dev = static_cast<unsigned *>(d);
      for (int i = 0; i < 16; ++i) st.host[i] = 0u;
    } else {
      cudaGetLastError();
    }
  }
  return st.host ? &st : nullptr;
}
// 1 and the record (CTA, row, site) if a sweep on the current device has given up since the last
// call; the record is cleared.
static int take_sweep_abort(unsigned *info3) {
  SweepStatus *st = sweep_status();
  if (st == nullptr) return 0;
  std::lock_guard<std::mutex> lock(st->mu);
  volatile unsigned *h = st->host;
  if (h[0] == 0u) return 0;
  if (info3) { info3[0] = h[1]; info3[1] = h[2]; info3[2] = h[3]; }
  h[0] = 0u;
  return 1;
}

template <int NV, int VW, bool M = false>
static int run(const float *cost, const float *image, float *out, const float *extra, const float *extra2,
               int32_t *disp, int H, int W, int D, float p1, float p2, int min_limit, int up, bool store_vol,
               int ns, char *ws, cudaStream_t stream) {
  constexpr int DPAD = NV * 32 * VW;
  size_t smem_bytes = 0;
  int minb = 1;
  const int G = ctas<NV, VW, M>(W, ns == 2 ? 0 : (ns == 4 ? 2 : 1), &minb);
  if (G < 1) return 1;
  const void *kernel = sweep_kernel<NV, VW, M>(up != 0, disp != nullptr, ns, &smem_bytes, minb);
  if (kernel == nullptr) return 1;
  Params P;
  P.cost = cost; P.out = out; P.extra = extra; P.extra2 = extra2; P.disp = disp;
  P.pen = reinterpret_cast<const float4 *>(ws + flags_and_rings_bytes(DPAD)) + (up ? (size_t)H * W : 0);
  P.xcons = reinterpret_cast<unsigned *>(ws);
  P.xring = reinterpret_cast<unsigned long long *>(ws + align256(kMaxCtas * kFlagStride * sizeof(unsigned)));
  P.H = H; P.W = W; P.D = D; P.G = G; P.q = W / G; P.r = W % G;
  P.up = up; P.store_vol = store_vol ? 1 : 0; P.emit_wta = disp ? 1 : 0;
  P.defect = (up && W != H) ? 1 : 0;
  P.min_limit = min_limit; P.p1 = p1; P.p2 = p2;
  static const char *dbg = getenv("STM_DEBUG_VSWEEP");
  P.dbg = dbg ? atoi(dbg) : 0;
  SweepStatus *status = sweep_status();
  P.host_status = status ? status->dev : nullptr;
  P.spin_ns = g_spin_ns; P.fault_cta = g_fault_cta; P.fault_row = g_fault_row;
  {
    unsigned info[3];
    if (take_sweep_abort(info)) {
      char msg[200];
      snprintf(msg, sizeof msg,
               "sgm: a fused sweep launched earlier on this device made no progress for %.0f s and gave up "
               "(CTA %u, row %u, %s); that call's output is invalid",
               (double)g_spin_ns * 1e-9, info[0], info[1], info[2] ? "exporter waiting for credit" : "importer waiting for its neighbour");
      return fail(STM_ECUDA, "%s", msg);
    }
  }
  // tags are row numbers: clear the counters and the rings this launch will use
  STM_CUDA_TRY(cudaMemsetAsync(ws, 0, align256(kMaxCtas * kFlagStride * sizeof(unsigned)) +
                                          (size_t)G * kRX * 3 * DPAD * sizeof(unsigned long long), stream));
  void *args[] = {&P};
  STM_CUDA_TRY(cudaLaunchCooperativeKernel(kernel, dim3(G), dim3(kThreads), args, smem_bytes, stream));
  return check_launch("sgm_vsweep_kernel");
}

}  // namespace vsweep

// Developer hook (tests/test_gpu_sweep_abort.py): CTA `cta`'s exporter stops at row `row` (cta < 0: no
// fault) and polling loops give up after `timeout_ms` (0: the default 60 s).  Process-wide.
extern "C" __attribute__((visibility("default"))) int stm_debug_vsweep_fault(int cta, int row, unsigned timeout_ms) {
  vsweep::g_fault_cta = cta; vsweep::g_fault_row = row;
  vsweep::g_spin_ns = timeout_ms ? (unsigned long long)timeout_ms * 1000000ull : vsweep::kSpinSeconds * 1000000000ull;
  return 0;
}

extern "C" __attribute__((visibility("default"))) int stm_debug_vsweep_trace(long long *out) {
  return cudaMemcpyFromSymbol(out, vsweep::g_trace, sizeof(long long) * 16 * vsweep::kTraceRows * 4) == cudaSuccess ? 0 : -1;
}

// Developer hook (tools/vsweep_probe.py): copies and clears the statistics counters.
extern "C" __attribute__((visibility("default"))) int stm_debug_vsweep_stats(unsigned long long *out64) {
  if (cudaMemcpyFromSymbol(out64, vsweep::g_stats, sizeof(unsigned long long) * 64) != cudaSuccess) return -1;
  static unsigned long long zero[64];
  return cudaMemcpyToSymbol(vsweep::g_stats, zero, sizeof(zero)) == cudaSuccess ? 0 : -1;
}

// The lane map the sweeps run a disparity range with (not always pick_lane_map's: a range that is a
// multiple of 4 but not of 128 takes the 4-wide map of the next 128 with its last lanes masked --
// D = 96 or 100 run as D = 128 does, where three or four scalar vectors per lane cost 10-100 % more).
struct VsMap { int NV, VW; bool masked, ok; };
static VsMap vsweep_map(int D) {
  VsMap m{0, 0, false, true};
  if (D >= 1 && D <= 32) { m.NV = 1; m.VW = 1; }
  else if (D == 64) { m.NV = 1; m.VW = 2; }
  else if (D < 64 && D % 2 == 0) { m.NV = 1; m.VW = 2; m.masked = true; }
  else if (D == 128 || D == 256) { m.NV = D / 128; m.VW = 4; }
  else if (D < 128 && D % 4 == 0) { m.NV = 1; m.VW = 4; m.masked = true; }
  else if (D == 192) { m.NV = 3; m.VW = 2; }
  else if (D > 64 && D <= 96) { m.NV = 3; m.VW = 1; }
  else if (D > 128 && D <= 160) { m.NV = 5; m.VW = 1; }
  else m.ok = false;
  return m;
}
static bool vsweep_lane_map(int D) { return vsweep_map(D).ok; }
static size_t vsweep_dpad(int D) { const VsMap m = vsweep_map(D); return (size_t)m.NV * 32 * m.VW; }

// [imported-row counters][CTA->CTA rings][penalties of both sweeps]
size_t sgm_vsweep_workspace_bytes(int H, int W, int D) {
  if (H < 2 || W < 2 || !vsweep_lane_map(D)) return 0;
  return vsweep::flags_and_rings_bytes(vsweep_dpad(D)) + (size_t)2 * H * W * sizeof(float4);
}

// Once per aggregation, before the two sweeps: the adaptive penalties of both (stream order).
int launch_sgm_vsweep_prepare(const float *image, int H, int W, int D, float p1, float p2,
                              void *workspace, cudaStream_t stream) {
  float4 *pen = reinterpret_cast<float4 *>(static_cast<char *>(workspace) +
                                           vsweep::flags_and_rings_bytes(vsweep_dpad(D)));
  const int n = H * W;
  vsweep::sgm_vsweep_penalties<<<(n + 255) / 256, 256, 0, stream>>>(image, pen, H, W, p1, p2);
  return check_launch("sgm_vsweep_penalties");
}

// stm_sgm_sweep_aborted (include/stm_abi.h)
int sgm_sweep_aborted(unsigned *info3) { return vsweep::take_sweep_abort(info3); }

// Whether launch_sgm_vsweep will take this shape on the current device.
bool sgm_vsweep_can_run(int H, int W, int D, int variant_group) {
  if (H < 2 || W < 2 || !vsweep_lane_map(D)) return false;
  const VsMap m = vsweep_map(D);
  int minb = 1;
#define STM_VSWEEP_CASE(NV_, VW_, M_) \
  if (m.NV == NV_ && m.VW == VW_ && m.masked == M_) return vsweep::ctas<NV_, VW_, M_>(W, variant_group, &minb) > 0;
  STM_VSWEEP_CASE(1, 4, false) STM_VSWEEP_CASE(2, 4, false) STM_VSWEEP_CASE(1, 2, false) STM_VSWEEP_CASE(3, 2, false)
  STM_VSWEEP_CASE(1, 1, false) STM_VSWEEP_CASE(3, 1, false) STM_VSWEEP_CASE(5, 1, false)
  STM_VSWEEP_CASE(1, 4, true) STM_VSWEEP_CASE(1, 2, true)
#undef STM_VSWEEP_CASE
  return false;
}

// One fused sweep: out += L2+L4+L6 (up = 0) or L3+L5+L7 (up = 1, reference multiplicities),
// optionally without writing the sums back (store_vol = false) and/or emitting the disparity
// map.  Returns 1 when the shape is not handled here (the caller runs the directions one by one).
int launch_sgm_vsweep(const float *cost, const float *image, float *out, const float *extra, const float *extra2,
                      int32_t *disp, int H, int W, int D, float p1, float p2, int min_limit, int up,
                      bool store_vol, int volume_streams, void *workspace, size_t workspace_bytes,
                      cudaStream_t stream) {
  if (H < 2 || W < 2 || !vsweep_lane_map(D)) return 1;
  if (workspace == nullptr || workspace_bytes < sgm_vsweep_workspace_bytes(H, W, D)) return 1;
  if ((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(out) |
       reinterpret_cast<uintptr_t>(workspace) | reinterpret_cast<uintptr_t>(extra) |
       reinterpret_cast<uintptr_t>(extra2)) & 15)
    return 1;
  if (volume_streams < 1 || volume_streams > 4 || (volume_streams >= 3) != (extra != nullptr) ||
      (volume_streams == 4) != (extra2 != nullptr))
    return 1;
  const VsMap m = vsweep_map(D);
  char *ws = static_cast<char *>(workspace);
#define STM_VSWEEP_CASE(NV_, VW_, M_)                             \
  if (m.NV == NV_ && m.VW == VW_ && m.masked == M_)               \
    return vsweep::run<NV_, VW_, M_>(cost, image, out, extra, extra2, disp, H, W, D, p1, p2, min_limit, up, store_vol, \
                                     volume_streams, ws, stream);
  STM_VSWEEP_CASE(1, 4, false) STM_VSWEEP_CASE(2, 4, false) STM_VSWEEP_CASE(1, 2, false) STM_VSWEEP_CASE(3, 2, false)
  STM_VSWEEP_CASE(1, 1, false) STM_VSWEEP_CASE(3, 1, false) STM_VSWEEP_CASE(5, 1, false)
  STM_VSWEEP_CASE(1, 4, true) STM_VSWEEP_CASE(1, 2, true)
#undef STM_VSWEEP_CASE
  return 1;
}

}  // namespace stm
