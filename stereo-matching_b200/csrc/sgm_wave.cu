// sgm_wave.cu -- semiglobal aggregation as two raster sweeps, DIRECTION-SPLIT warps (sm_100a).
//
// EXPERIMENTAL, opt-in (STM_SGM_FUSED_SWEEPS): passes the same parity tests as the default
// per-direction kernels and moves ~20 instead of ~92 B/pix-disp, but is ~5x SLOWER (20.3 ms vs
// 4.33 ms at 1080p D=128): a wavefront is serial along each row and a dependent step costs
// ~300 ns here, see profiles/r01_v3_fused_sweeps.md.  Second design; the first one (one warp per
// row, removed; same file in git history as sgm_sweep.cu) met the traffic goal (19.7 B/pix-disp instead of ~92) but
// put all four recurrences of a pixel into ONE warp: ~250 instructions per pixel on the
// critical path of a wavefront that keeps only ~one warp per scheduler busy -> 20.6 ms.
//
// Here a row is served by FIVE warps:
//   A  horizontal recurrence          (-> forward, <- backward)   depends on nothing above
//   B  vertical                       (v / ^)    needs B of the row above at the same pixel
//   C  diagonal from the left         (\v / ^\)  needs C of the row above at pixel lx-1
//   E  diagonal from the right        (/v / ^/)  needs E of the row above at pixel lx+1
//   S  sum: out (+)= A + B + C + E (reference path-set multiplicities applied here)
// The four recurrences are independent chains (each direction only ever reads its OWN
// direction's vector from the row above), so every warp has a ~50-instruction step, rows are
// coupled only through B/C/E with a lag of 1-2 pixels, and A is not on the wavefront at all.
// Vectors travel through shared-memory rings (one per direction and row boundary, plus a
// 4-vector "sum ring" per row); progress counters are CUMULATIVE pixel counts (iteration *
// W + lx), so persistent CTAs can stride over groups of kR rows without ever resetting a flag.
// Groups are chained through an L2-resident ring in global memory by an exporter and an
// importer warp per CTA, as in the first design.
//
// Flags in shared memory are strong relaxed loads/stores (no fence: st.release.cta compiles to
// MEMBAR.ALL.CTA + STS, which waits for the warp's in-flight global loads) ordered by __syncwarp(), "memory" clobbers and
// the in-order shared-memory pipeline; flags in global memory are acquire/release at gpu
// scope.  Every wait is bounded and traps.  The launch is cooperative: a ring producer waits
// for its consumer, so all CTAs must be co-resident.
#include "lanevec.cuh"
#include "sgm_paths.cuh"

namespace stm {
namespace wave {

constexpr int kR = 4;     // rows per group (and per CTA iteration)
constexpr int kK = 8;     // slots of a row -> row ring
constexpr int kK0 = 16;   // slots of the importer -> first-row ring
constexpr int kKs = 8;    // slots of a sum ring
constexpr int kGK = 32;   // slots of a group -> group ring in global memory
constexpr int kWarps = 5 * kR + 2;
constexpr unsigned kSpinCap = 1u << 22;
enum { kRoleA = 0, kRoleB = 1, kRoleC = 2, kRoleE = 3, kRoleS = 4 };

struct Params {
  const float *cost;
  const float *image;
  float *out;
  int H, W, D, groups;
  float p1, p2;
  int min_limit;
  unsigned *gprog;  // [groups] pixels of group g's last row exported
  unsigned *gcons;  // [groups] ... and imported by group g+1
  float *gring;     // [groups][kGK][3][dpad]
  float *lstore_c, *lstore_e;  // backward sweep: multiplicity > 1 diagonals (or null)
  int c_sx, e_sx;
};

__device__ __forceinline__ unsigned ld_flag(const unsigned *p) {
  unsigned v;
  asm volatile("ld.relaxed.cta.shared.u32 %0, [%1];"
               : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_flag(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.cta.shared.u32 [%0], %1;"
               :: "r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_g(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_g(unsigned *p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

__device__ __noinline__ unsigned wait_flag_slow(const unsigned *flag, unsigned need) {
  unsigned v = ld_flag(flag);
  unsigned tries = 0, ns = 32;
  while (v < need) {
    if (++tries > 64) {
      __nanosleep(ns);
      if (ns < 256) ns <<= 1;
      if (tries > kSpinCap) __trap();
    }
    v = ld_flag(flag);
  }
  return v;
}
// Waits until *flag >= need (cumulative pixel counters never decrease).
__device__ __forceinline__ unsigned wait_flag(const unsigned *flag, unsigned need) {
  const unsigned v = ld_flag(flag);
  return v >= need ? v : wait_flag_slow(flag, need);
}
__device__ __forceinline__ unsigned wait_gflag(const unsigned *flag, unsigned need) {
  unsigned v = ld_acquire_g(flag);
  unsigned tries = 0, ns = 64;
  while (v < need) {
    __nanosleep(ns);
    if (ns < 512) ns <<= 1;
    v = ld_acquire_g(flag);
    if (++tries > kSpinCap) __trap();
  }
  return v;
}

// ---- shared memory -------------------------------------------------------------------------
// vertical rings: boundary b in [0, kR] (b = 0 fed by the importer, b = r by row r-1, b = kR
// drained by the exporter) x direction v in {B, C, E}; sum rings: row r x direction k in
// {A, B, C, E}.  One slot = one pixel's vector of one direction (DPAD floats).
template <int DPAD>
struct Smem {
  static constexpr int kVSlots = kK0 + kR * kK;  // per direction
  static constexpr int kVFloats = 3 * kVSlots * DPAD;
  static constexpr int kSFloats = kR * 4 * kKs * DPAD;
  static constexpr int kFlags = 2 * 3 * (kR + 1) + 4 * kR + kR;
  static constexpr size_t kBytes = (size_t)(kVFloats + kSFloats) * 4 + kFlags * sizeof(unsigned);
  __device__ static float *vring(float *base, int b, int v) {
    return base + (size_t)(v * kVSlots + (b == 0 ? 0 : kK0 + (b - 1) * kK)) * DPAD;
  }
  __device__ static unsigned vmask(int b) { return b == 0 ? kK0 - 1 : kK - 1; }
  __device__ static float *sring(float *base, int r, int k) {
    return base + kVFloats + (size_t)((r * 4 + k) * kKs) * DPAD;
  }
  __device__ static unsigned *flags(float *base) {
    return reinterpret_cast<unsigned *>(base + kVFloats + kSFloats);
  }
  __device__ static unsigned *vprog(float *base, int b, int v) { return flags(base) + b * 3 + v; }
  __device__ static unsigned *vcons(float *base, int b, int v) {
    return flags(base) + 3 * (kR + 1) + b * 3 + v;
  }
  __device__ static unsigned *sprog(float *base, int r, int k) {
    return flags(base) + 6 * (kR + 1) + r * 4 + k;
  }
  __device__ static unsigned *scons(float *base, int r) {
    return flags(base) + 6 * (kR + 1) + 4 * kR + r;
  }
};

// ---- one direction of one row ---------------------------------------------------------------
// ROLE A: previous pixel of the path = (ly, lx-1), state stays in registers.
// ROLE B / C / E: previous pixel = (ly-1, lx + OFF), OFF = 0 / -1 / +1, read from the ring
// this direction's warp of the row above fills.  A path STARTS (L = C) where that pixel is
// outside the image.  All loads are unconditional (clamped): a load under a branch is waited
// for at the join (profiles/r01_v3_fused_sweeps.md).
template <int NV, int VW, bool BACKWARD, int ROLE>
__device__ void dir_warp(float *smem, const Params &P, const int r, const int lane) {
  constexpr int DPAD = NV * 32 * VW;
  using S = Smem<DPAD>;
  using Vec = LaneVec<NV, VW>;
  constexpr int V = ROLE == kRoleA ? 0 : ROLE - 1;  // vertical ring of this direction
  constexpr int OFF = ROLE == kRoleC ? -1 : (ROLE == kRoleE ? 1 : 0);
  constexpr int POFF = ROLE == kRoleA ? -1 : OFF;   // column offset of the path's previous pixel
  constexpr int PF = (NV * VW <= 4) ? 4 : 2;
  const int H = P.H, W = P.W, D = P.D;
  const LaneCtx cx = lv_ctx<NV, VW>(lane, P.min_limit);
  const int lane_off = lane * VW, dl = D - lane_off;
  const float p1 = P.p1, p2 = P.p2;
  const long long stepd = BACKWARD ? -(long long)D : (long long)D;
  auto phys = [&](int l) { return BACKWARD ? W - 1 - l : l; };

  float *sum_ring = S::sring(smem, r, ROLE) + lane_off;
  unsigned *sum_prog = S::sprog(smem, r, ROLE);
  const unsigned *sum_cons = S::scons(smem, r);
  const float *ring_in = S::vring(smem, r, V) + lane_off;
  const unsigned mask_in = S::vmask(r);
  float *ring_out = S::vring(smem, r + 1, V) + lane_off;
  const unsigned *prog_in = S::vprog(smem, r, V);
  unsigned *cons_in = S::vcons(smem, r, V);
  unsigned *prog_out = S::vprog(smem, r + 1, V);
  const unsigned *cons_out = S::vcons(smem, r + 1, V);

  unsigned seen_in = 0, seen_out = 0, seen_sum = 0;
  unsigned base = 0;  // cumulative pixel count of this CTA's earlier iterations
  for (int g = blockIdx.x; g < P.groups; g += gridDim.x, base += W) {
    const int ly = g * kR + r;
    if (ly >= H) break;
    const bool has_above = ROLE != kRoleA && ly > 0;
    const bool has_below = ROLE != kRoleA && ly + 1 < H;
    const int y = BACKWARD ? H - 1 - ly : ly;
    const int ya = (ROLE != kRoleA && ly > 0) ? (BACKWARD ? y + 1 : y - 1) : y;
    const float *crow = P.cost + ((long long)y * W + (BACKWARD ? W - 1 : 0)) * D + lane_off;
    const float *irow = P.image + (long long)y * W;
    const float *iprev = P.image + (long long)ya * W;

    Vec cbuf[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u)
      lv_load<NV, VW, true>(cbuf[u], crow + (long long)(u < W ? u : W - 1) * stepd, 0, dl);
    Vec L = cbuf[0];
    float pblk = 0.f;  // lane l: adaptive P2 of pixel (lx & ~31) + l for this direction
    auto penalties = [&](int lx0) {
      int q = lx0 + lane;
      if (q > W - 1) q = W - 1;
      int qp = q + POFF;
      qp = qp < 0 ? 0 : (qp > W - 1 ? W - 1 : qp);
      pblk = sgm_p2_adjusted(p1, p2, __ldg(irow + phys(q)), __ldg(iprev + phys(qp)));
    };

    auto step = [&](const int lx, const Vec &c) {
      const unsigned n = base + lx;
      bool start;
      int consumed = 0;
      if constexpr (ROLE == kRoleA) {
        start = lx == 0;
      } else {
        const int m = lx + OFF;
        start = !has_above || m < 0 || m >= W;
        const int mc = m < 0 ? 0 : (m >= W ? W - 1 : m);
        consumed = m + 1 < 0 ? 0 : (m + 1 > W ? W : m + 1);
        const unsigned need = has_above ? base + mc + 1 : 0u;
        if (seen_in < need) seen_in = wait_flag(prog_in, need);
        lv_load<NV, VW, false>(L, ring_in + (size_t)((base + mc) & mask_in) * DPAD, 0, DPAD);
      }
      const float p = __shfl_sync(FULL, pblk, lx & 31);
      lv_sgm_step<NV, VW>(L, c, p1, p, cx);
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int j = 0; j < VW; ++j) L.v[k][j] = start ? c.v[k][j] : L.v[k][j];

      if (has_below) {  // next row, same direction
        const unsigned need = n >= (unsigned)kK ? n - kK + 1 : 0u;
        if (seen_out < need) seen_out = wait_flag(cons_out, need);
        lv_store<NV, VW, false>(L, ring_out + (size_t)(n & (kK - 1)) * DPAD, 0, DPAD);
      }
      {  // this row's sum warp
        const unsigned need = n >= (unsigned)kKs ? n - kKs + 1 : 0u;
        if (seen_sum < need) seen_sum = wait_flag(sum_cons, need);
        lv_store<NV, VW, false>(L, sum_ring + (size_t)(n & (kKs - 1)) * DPAD, 0, DPAD);
      }
      __syncwarp();
      if (lane == 0) {
        if (has_below) st_flag(prog_out, n + 1);
        st_flag(sum_prog, n + 1);
        if (has_above) st_flag(cons_in, base + consumed);
      }
    };

    int lx = 0;
    for (int grp = W / PF; grp > 0; --grp) {
      if ((lx & 31) == 0) penalties(lx);
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const Vec c = cbuf[u];
        lv_load<NV, VW, true>(cbuf[u], crow + (long long)(lx + PF < W ? lx + PF : W - 1) * stepd, 0, dl);
        step(lx, c);
        ++lx;
      }
    }
    if (lx < W && (lx & 31) == 0) penalties(lx);
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      if (lx < W) {
        step(lx, cbuf[u]);
        ++lx;
      }
    }
    if (ROLE != kRoleA) {  // everything of the row above is dead (keeps the counters cumulative)
      __syncwarp();
      if (lane == 0) st_flag(cons_in, base + W);
    }
  }
}

// ---- reference path-set multiplicities of the two up-diagonal directions (sgm_paths.cuh) ----
__device__ __forceinline__ int mult_dir5(int x, int y, int H, int W) {
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
__device__ __forceinline__ int mult_dir7(int x, int y, int H, int W) {
  if (W > H) {
    const int s = x + y;
    if (s >= H && s <= W - 1) return 0;
    if (s == H - 1) return W - H + 1;
  }
  return 1;
}

template <int NV, int VW>
__device__ __forceinline__ void vec_add(LaneVec<NV, VW> &a, const LaneVec<NV, VW> &b) {
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j) a.v[k][j] = __fadd_rn(a.v[k][j], b.v[k][j]);
}

// ---- sum warp of one row: out (+)= A + B + C + E --------------------------------------------
template <int NV, int VW, bool BACKWARD, bool ACC>
__device__ void sum_warp(float *smem, const Params &P, const int r, const int lane) {
  constexpr int DPAD = NV * 32 * VW;
  using S = Smem<DPAD>;
  using Vec = LaneVec<NV, VW>;
  constexpr int PF = (NV * VW <= 4) ? 4 : 2;
  const int H = P.H, W = P.W, D = P.D;
  const int lane_off = lane * VW, dl = D - lane_off;
  const long long stepd = BACKWARD ? -(long long)D : (long long)D;
  const float *ringA = S::sring(smem, r, 0) + lane_off;
  const float *ringB = S::sring(smem, r, 1) + lane_off;
  const float *ringC = S::sring(smem, r, 2) + lane_off;
  const float *ringE = S::sring(smem, r, 3) + lane_off;
  const unsigned *prog = S::sprog(smem, r, 0);
  unsigned *cons = S::scons(smem, r);
  unsigned seen0 = 0, seen1 = 0, seen2 = 0, seen3 = 0;
  unsigned base = 0;
  for (int g = blockIdx.x; g < P.groups; g += gridDim.x, base += W) {
    const int ly = g * kR + r;
    if (ly >= H) break;
    const int y = BACKWARD ? H - 1 - ly : ly;
    float *orow = P.out + ((long long)y * W + (BACKWARD ? W - 1 : 0)) * D + lane_off;

    Vec sbuf[PF];
    if (ACC) {
#pragma unroll
      for (int u = 0; u < PF; ++u)
        lv_load<NV, VW, true>(sbuf[u], orow + (long long)(u < W ? u : W - 1) * stepd, 0, dl);
    }

    auto step = [&](const int lx, const Vec &old) {
      const unsigned n = base + lx;
      if (seen0 < n + 1) seen0 = wait_flag(prog + 0, n + 1);
      if (seen1 < n + 1) seen1 = wait_flag(prog + 1, n + 1);
      if (seen2 < n + 1) seen2 = wait_flag(prog + 2, n + 1);
      if (seen3 < n + 1) seen3 = wait_flag(prog + 3, n + 1);
      const size_t slot = (size_t)(n & (kKs - 1)) * DPAD;
      Vec a, b, c, e;
      lv_load<NV, VW, false>(a, ringA + slot, 0, DPAD);
      lv_load<NV, VW, false>(b, ringB + slot, 0, DPAD);
      lv_load<NV, VW, false>(c, ringC + slot, 0, DPAD);
      lv_load<NV, VW, false>(e, ringE + slot, 0, DPAD);
      // one fp32 addition per direction, like the reference
      Vec acc = a;
      if (ACC) { acc = old; vec_add<NV, VW>(acc, a); }
      vec_add<NV, VW>(acc, b);
      if (BACKWARD) {
        const int x = W - 1 - lx;
        const int mc = mult_dir5(x, y, H, W), me = mult_dir7(x, y, H, W);
        if (mc) vec_add<NV, VW>(acc, c);
        if (me) vec_add<NV, VW>(acc, e);
        if (mc > 1 && P.lstore_c)
          lv_store<NV, VW, false>(c, P.lstore_c + (size_t)(P.c_sx - x) * DPAD + lane_off, 0, DPAD);
        if (me > 1 && P.lstore_e)
          lv_store<NV, VW, false>(e, P.lstore_e + (size_t)(x - P.e_sx) * DPAD + lane_off, 0, DPAD);
      } else {
        vec_add<NV, VW>(acc, c);
        vec_add<NV, VW>(acc, e);
      }
      lv_store<NV, VW, true>(acc, orow + (long long)lx * stepd, 0, dl);
      __syncwarp();
      if (lane == 0) st_flag(cons, n + 1);
    };

    int lx = 0;
    for (int grp = W / PF; grp > 0; --grp) {
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        Vec old;
        if (ACC) {
          old = sbuf[u];
          lv_load<NV, VW, true>(sbuf[u], orow + (long long)(lx + PF < W ? lx + PF : W - 1) * stepd, 0, dl);
        }
        step(lx, old);
        ++lx;
      }
    }
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      if (lx < W) {
        Vec old;
        if (ACC) old = sbuf[u];
        step(lx, old);
        ++lx;
      }
    }
  }
}

// ---- group -> group transport through global memory ----------------------------------------
__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
               :: "r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// Importer: global ring of group g-1 -> the three b = 0 rings (first row of this group).
template <int DPAD>
__device__ void importer_warp(float *smem, const Params &P, const int lane) {
  using S = Smem<DPAD>;
  constexpr int n4 = DPAD / 4;
  const unsigned W = P.W;
  unsigned base = 0;
  for (int g = blockIdx.x; g < P.groups; g += gridDim.x, base += W) {
    if (g == 0) continue;
    const unsigned *gprog = P.gprog + (g - 1);
    unsigned *gcons = P.gcons + (g - 1);
    const float *gring = P.gring + (size_t)(g - 1) * kGK * 3 * DPAD;
    unsigned done = 0;
    while (done < W) {
      const unsigned avail = wait_gflag(gprog, done + 1);
      const unsigned n0 = base + done;
      const unsigned need = n0 + 1 > (unsigned)kK0 ? n0 + 1 - kK0 : 0u;
      unsigned c = wait_flag(S::vcons(smem, 0, 0), need);
      const unsigned c1 = wait_flag(S::vcons(smem, 0, 1), need);
      const unsigned c2 = wait_flag(S::vcons(smem, 0, 2), need);
      c = c < c1 ? c : c1;
      c = c < c2 ? c : c2;
      unsigned limit = base + (avail < W ? avail : W);  // cumulative
      if (limit > c + kK0) limit = c + kK0;
      for (unsigned n = n0; n < limit; ++n) {
        const float *src = gring + (size_t)((n - base) & (kGK - 1)) * 3 * DPAD;
        const size_t slot = (size_t)(n & (kK0 - 1)) * DPAD;
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          float *dst = S::vring(smem, 0, v) + slot;
          for (int i = lane; i < n4; i += 32) cp_async16(dst + 4 * i, src + v * DPAD + 4 * i);
        }
      }
      cp_async_wait_all();
      __syncwarp();
      if (lane == 0) {
        st_flag(S::vprog(smem, 0, 0), limit);
        st_flag(S::vprog(smem, 0, 1), limit);
        st_flag(S::vprog(smem, 0, 2), limit);
        st_release_g(gcons, limit - base);
      }
      done = limit - base;
    }
  }
}

// Exporter: the three b = kR rings (last row of this group) -> global ring of group g.
template <int DPAD>
__device__ void exporter_warp(float *smem, const Params &P, const int lane) {
  using S = Smem<DPAD>;
  constexpr int n4 = DPAD / 4;
  const unsigned W = P.W;
  unsigned base = 0;
  for (int g = blockIdx.x; g < P.groups; g += gridDim.x, base += W) {
    if ((g + 1) * kR >= P.H) break;  // no next group
    unsigned *gprog = P.gprog + g;
    const unsigned *gcons = P.gcons + g;
    float *gring = P.gring + (size_t)g * kGK * 3 * DPAD;
    unsigned done = 0, gc = 0;
    while (done < W) {
      const unsigned need = base + done + 1;
      unsigned a = wait_flag(S::vprog(smem, kR, 0), need);
      const unsigned a1 = wait_flag(S::vprog(smem, kR, 1), need);
      const unsigned a2 = wait_flag(S::vprog(smem, kR, 2), need);
      a = a < a1 ? a : a1;
      a = a < a2 ? a : a2;
      if (done + 1 > gc + kGK) gc = wait_gflag(gcons, done + 1 - kGK);
      unsigned limit = a - base;
      if (limit > W) limit = W;
      if (limit > gc + kGK) limit = gc + kGK;
      for (unsigned q = done; q < limit; ++q) {
        const size_t slot = (size_t)((base + q) & (kK - 1)) * DPAD;
        float4 *dst = reinterpret_cast<float4 *>(gring + (size_t)(q & (kGK - 1)) * 3 * DPAD);
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          const float4 *src = reinterpret_cast<const float4 *>(S::vring(smem, kR, v) + slot);
          for (int i = lane; i < n4; i += 32) __stcg(dst + v * n4 + i, src[i]);
        }
      }
      __syncwarp();
      if (lane == 0) {
        st_flag(S::vcons(smem, kR, 0), base + limit);
        st_flag(S::vcons(smem, kR, 1), base + limit);
        st_flag(S::vcons(smem, kR, 2), base + limit);
        st_release_g(gprog, limit);
      }
      done = limit;
    }
  }
}

// ---- kernel: 5 warps per row x kR rows + importer + exporter --------------------------------
template <int NV, int VW, bool BACKWARD, bool ACC>
__global__ void __launch_bounds__(kWarps * 32, 1) sgm_wave_kernel(const Params P) {
  constexpr int DPAD = NV * 32 * VW;
  using S = Smem<DPAD>;
  extern __shared__ __align__(16) float smem[];
  for (int i = threadIdx.x; i < S::kFlags; i += blockDim.x) S::flags(smem)[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp == 5 * kR) {
    importer_warp<DPAD>(smem, P, lane);
  } else if (warp == 5 * kR + 1) {
    exporter_warp<DPAD>(smem, P, lane);
  } else {
    const int r = warp / 5;
    switch (warp - 5 * r) {
      case kRoleA: dir_warp<NV, VW, BACKWARD, kRoleA>(smem, P, r, lane); break;
      case kRoleB: dir_warp<NV, VW, BACKWARD, kRoleB>(smem, P, r, lane); break;
      case kRoleC: dir_warp<NV, VW, BACKWARD, kRoleC>(smem, P, r, lane); break;
      case kRoleE: dir_warp<NV, VW, BACKWARD, kRoleE>(smem, P, r, lane); break;
      default: sum_warp<NV, VW, BACKWARD, ACC>(smem, P, r, lane); break;
    }
  }
}

}  // namespace wave

// ---- host side --------------------------------------------------------------------------------
int sgm_repeat_add(float *out, float *l_store, int H, int W, int D, int dpad, int dir, int path,
                   int times, cudaStream_t stream);  // sgm.cu

namespace wave {

static size_t align16(size_t n) { return (n + 15) & ~(size_t)15; }

// Layout: [lstore dir5][lstore dir7][group rings][flags: 2 sweeps x
// {gprog, gcons}]; groups = ceil(H / 4).
struct Workspace {
  size_t lstore_bytes, ring_bytes, flag_bytes, total;
  int groups;
};
static Workspace workspace(int H, int W, int D) {
  const LaneMap m = pick_lane_map(D);
  const size_t dpad = (size_t)m.NV * 32 * m.VW;
  Workspace w;
  w.groups = (H + kR - 1) / kR;
  w.lstore_bytes = align16((size_t)(H < W ? H : W) * dpad * sizeof(float));
  w.ring_bytes = align16((size_t)w.groups * kGK * 3 * dpad * sizeof(float));
  w.flag_bytes = align16((size_t)4 * w.groups * sizeof(unsigned));
  w.total = 2 * w.lstore_bytes + w.ring_bytes + w.flag_bytes;
  return w;
}

template <int NV, int VW, bool BACKWARD, bool ACC>
static int launch_one(const Params &P, int grid, cudaStream_t stream, int *capacity) {
  constexpr int DPAD = NV * 32 * VW;
  const size_t smem = Smem<DPAD>::kBytes;
  auto kernel = sgm_wave_kernel<NV, VW, BACKWARD, ACC>;
  STM_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (capacity) {  // query only
    int per_sm = 0;
    STM_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kWarps * 32, smem));
    *capacity = per_sm * num_sms();
    return STM_OK;
  }
  Params p = P;
  void *args[] = {&p};
  STM_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)kernel, dim3(grid), dim3(kWarps * 32), args,
                                           smem, stream));
  return check_launch("sgm_wave_kernel");
}

// Returns 1 when this configuration cannot run.
template <int NV, int VW>
static int run(const float *cost, const float *image, float *out, int H, int W, int D, float p1,
               float p2, unsigned flags, char *ws, cudaStream_t stream) {
  constexpr int DPAD = NV * 32 * VW;
  if (Smem<DPAD>::kBytes > 227 * 1024) return 1;
  const Workspace w = workspace(H, W, D);
  Params P;
  P.cost = cost; P.image = image; P.out = out;
  P.H = H; P.W = W; P.D = D; P.groups = w.groups; P.p1 = p1; P.p2 = p2;
  P.min_limit = (flags & STM_SGM_MIN_ALL_D) ? D : D - 1;
  P.gring = reinterpret_cast<float *>(ws + 2 * w.lstore_bytes);
  unsigned *flagbase = reinterpret_cast<unsigned *>(ws + 2 * w.lstore_bytes + w.ring_bytes);
  P.lstore_c = P.lstore_e = nullptr;
  P.c_sx = P.e_sx = 0;
  P.gprog = flagbase; P.gcons = flagbase + w.groups;
  const bool overwrite = (flags & STM_SGM_OVERWRITE) != 0;

  int rc, cap = 0;
  if ((rc = launch_one<NV, VW, true, true>(P, 0, stream, &cap))) return rc;
  if (cap < 1) return 1;
  // persistent CTAs stride over the groups; never more CTAs than can be co-resident
  const int grid = w.groups < cap ? w.groups : cap;

  STM_CUDA_TRY(cudaMemsetAsync(flagbase, 0, w.flag_bytes, stream));
  rc = overwrite ? launch_one<NV, VW, false, false>(P, grid, stream, nullptr)
                 : launch_one<NV, VW, false, true>(P, grid, stream, nullptr);
  if (rc) return rc;

  P.gprog = flagbase + 2 * w.groups; P.gcons = flagbase + 3 * w.groups;
  const int m5 = sgm_multi_path_index(5, H, W), m7 = sgm_multi_path_index(7, H, W);
  PathGeom g5{}, g7{};
  if (m5 >= 0) {
    g5 = sgm_path_geom(5, m5, H, W);
    if (g5.mult > 1) { P.lstore_c = reinterpret_cast<float *>(ws); P.c_sx = g5.sx; }
  }
  if (m7 >= 0) {
    g7 = sgm_path_geom(7, m7, H, W);
    if (g7.mult > 1) { P.lstore_e = reinterpret_cast<float *>(ws + w.lstore_bytes); P.e_sx = g7.sx; }
  }
  if ((rc = launch_one<NV, VW, true, true>(P, grid, stream, nullptr))) return rc;
  if (P.lstore_c && (rc = sgm_repeat_add(out, P.lstore_c, H, W, D, DPAD, 5, m5, g5.mult - 1, stream))) return rc;
  if (P.lstore_e && (rc = sgm_repeat_add(out, P.lstore_e, H, W, D, DPAD, 7, m7, g7.mult - 1, stream))) return rc;
  return STM_OK;
}

}  // namespace wave

size_t sgm_wave_workspace_bytes(int H, int W, int D) { return wave::workspace(H, W, D).total; }

// Returns 1 when the shape is not handled here (caller falls back).
int launch_sgm_wave(const float *cost, const float *image, float *out, int H, int W, int D, float p1,
                    float p2, unsigned flags, void *workspace, size_t workspace_bytes,
                    cudaStream_t stream) {
  if (workspace == nullptr || workspace_bytes < sgm_wave_workspace_bytes(H, W, D))
    return fail(STM_EINVAL, "sgm: workspace too small (%zu < %zu)", workspace_bytes,
                sgm_wave_workspace_bytes(H, W, D));
  if ((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(out) |
       reinterpret_cast<uintptr_t>(workspace)) & 15)
    return 1;
  if ((long long)W * ((H + wave::kR - 1) / wave::kR) >= (1ll << 31)) return 1;  // cumulative counters
  const LaneMap m = pick_lane_map(D);
  char *ws = static_cast<char *>(workspace);
#define STM_WAVE_CASE(NV_, VW_) \
  if (m.NV == NV_ && m.VW == VW_) \
    return wave::run<NV_, VW_>(cost, image, out, H, W, D, p1, p2, flags, ws, stream);
  STM_WAVE_CASE(1, 4) STM_WAVE_CASE(2, 4) STM_WAVE_CASE(1, 2) STM_WAVE_CASE(3, 2) STM_WAVE_CASE(1, 1)
#undef STM_WAVE_CASE
  return 1;
}

}  // namespace stm
