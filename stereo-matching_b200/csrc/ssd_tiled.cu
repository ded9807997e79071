// ssd_tiled.cu -- SSD cost volume as a separable, add-only box filter (sm_100a).
//
// Replaces CostOps::ComputeSSD (reference src/ssd.cu:64-81, functor :31-61):
//   C[y,x,d] = sum_{r in [y-K, y+K)} sum_{c in [max(x-K,d), min(x+K,W))} (L[r,c]-R[r,c-d])^2
// for d <= x, +inf for d > x (window rows/cols clipped to the image, never renormalised).
//
// The reference spends 2*(2K)^2 loads and (2K)^2 FMAs per output.  Here every squared
// difference e(r,c,d) is evaluated once, summed vertically and then horizontally with
// van Herk / Gil-Werman prefix+suffix blocks, ~3 additions per element and direction,
// and NO subtraction: a running sum that subtracts the leaving row would turn the exact
// zeros of a perfect match into rounding noise (unbounded relative error), whereas an
// add-only sum of non-negative terms stays within ~2K*2K ulp of the reference for any
// input and is bit-exact for integer-valued images (every partial sum is an exact
// integer < 2^24).
//
// CTA = one x-strip of TX = RUNS*2K output columns  x  one chunk of 32 disparities,
// walking down the rows of its y-range in HALF-BLOCKS of K rows.  Two warp roles meet at one
// barrier per half-block (it was one per row: each warp then had ~55 instructions of work
// between barriers and the kernel ran at 40 % issue utilisation, latency-bound):
//  * vertical warps: thread = (column c, 4 consecutive d).  Per input row it squares 4
//    differences (image rows staged in shared memory, conflict-free reads) and advances a
//    2K-deep van Herk ring held in REGISTERS (fully unrolled, static indices); the
//    vertical window sum of each of the K rows goes to the half-block's tile in shared
//    memory as one float4 (a warp writes 512 contiguous bytes).
//  * horizontal warps: lane = disparity, warp = a run of 2K output columns.  For each of the
//    K rows of the PREVIOUS half-block it reads the 4K-1 vertical sums it needs (128-byte
//    rows, conflict-free), does the prefix/suffix blocks in registers and streams 2K
//    fully-coalesced 128-byte stores into C.
// The image rows of the next half-block are requested with cp.async at the start of a
// half-block and waited for at its end (double-buffered), so their latency is covered by
// K rows of arithmetic.  HBM traffic = the 4 B/pix-disp store of C plus the images.
#include <cstdlib>

#include "stm_common.cuh"

namespace stm {

template <int K, int RUNS>
struct SsdCfg {
  static constexpr int WIN = 2 * K;
  static constexpr int HALF = K;                // rows per barrier interval
  static constexpr int TX = RUNS * WIN;         // output columns per CTA
  static constexpr int TXH = TX + WIN - 1;      // columns with halo
  static constexpr int DC = 32;                 // disparities per CTA
  static constexpr int VTHREADS = ((TXH * 8 + 31) / 32) * 32;
  static constexpr int HWARPS = RUNS;
  static constexpr int THREADS = VTHREADS + HWARPS * 32;
  static constexpr int RROW = TXH + DC - 1;     // staged right-image span
  static constexpr int RPITCH = RROW + 1;
  // dynamic shared memory, in floats: vbuf[2][HALF][TXH][DC], lrow[2][HALF][TXH],
  // rrow[2][HALF][RPITCH]
  static constexpr int VBUF = 2 * HALF * TXH * DC;
  static constexpr int LROW = 2 * HALF * TXH;
  static constexpr int RROWS = 2 * HALF * RPITCH;
  static constexpr size_t SMEM = (size_t)(VBUF + LROW + RROWS) * sizeof(float);
};

__device__ __forceinline__ void named_barrier(int threads) {
  asm volatile("bar.sync 1, %0;" ::"r"(threads) : "memory");
}

template <int K, int RUNS>
__global__ void __launch_bounds__(SsdCfg<K, RUNS>::THREADS, 1)
ssd_tiled_kernel(const float *__restrict__ left, const float *__restrict__ right,
                 float *__restrict__ cost, int H, int W, int D, int rows_per_chunk) {
  using Cfg = SsdCfg<K, RUNS>;
  constexpr int WIN = Cfg::WIN, HALF = Cfg::HALF, TXH = Cfg::TXH, DC = Cfg::DC;

  extern __shared__ __align__(16) float smem[];
  float *vbuf = smem;                       // [(b*HALF + i)*TXH + col][DC]
  float *lrow = smem + Cfg::VBUF;           // [(b*HALF + i)][TXH]
  float *rrow = lrow + Cfg::LROW;           // [(b*HALF + i)][RPITCH]

  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * Cfg::TX;
  const int d0 = blockIdx.y * DC;
  const int ya = blockIdx.z * rows_per_chunk;
  const int yb = min(H, ya + rows_per_chunk);
  const int T = (yb - ya) + WIN - 1;        // input rows ya-K .. yb+K-2
  const int n_half = (T + HALF - 1) / HALF; // barrier intervals
  const int cbase = x0 - K;                 // image column of staged column 0
  const int ubase = cbase - d0 - (DC - 1);  // right-image column of rrow[.][0]

  // Cooperative staging of the image rows of one half-block (zeros outside the image):
  // cp.async straight from global to shared memory, one group per half-block.  Threads
  // [0,TXH) stage the left rows, the next RROW threads the right rows; out-of-image pixels
  // are zero-filled by a source size of 0 (the address is clamped to stay valid).
  static_assert(Cfg::VTHREADS >= TXH + Cfg::RROW, "not enough loader threads");
  const bool st_left = tid < TXH;
  const bool st_any = tid < TXH + Cfg::RROW;
  const int st_col = st_left ? cbase + tid : ubase + (tid - TXH);
  const bool st_col_ok = st_any && st_col >= 0 && st_col < W;
  const float *st_src = (st_left ? left : right) + min(max(st_col, 0), W - 1);
  const uint32_t st_dst0 = (uint32_t)__cvta_generic_to_shared(
      st_left ? lrow + tid : rrow + (st_any ? tid - TXH : 0));
  const uint32_t st_pitch = (st_left ? TXH : Cfg::RPITCH) * sizeof(float);  // bytes per row slot
  auto stage_half = [&](int h) {  // request the rows of half-block h (any h >= 0)
    if (st_any) {
#pragma unroll
      for (int i = 0; i < HALF; ++i) {
        const int r = ya - K + h * HALF + i;
        const int ok_bytes = (st_col_ok && r >= 0 && r < H) ? 4 : 0;
        const float *src = st_src + (size_t)min(max(r, 0), H - 1) * W;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(
                         st_dst0 + (uint32_t)((h & 1) * HALF + i) * st_pitch),
                     "l"(src), "r"(ok_bytes)
                     : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  auto staged_landed = [&]() { asm volatile("cp.async.wait_group 0;" ::: "memory"); };

  if (tid < Cfg::VTHREADS) {
    stage_half(0);
    staged_landed();
  }
  __syncthreads();

  if (tid < Cfg::VTHREADS) {
    // ------------------------------ vertical role ------------------------------------
    const int cl = tid >> 3;   // staged column
    const int dq = tid & 7;    // which float4 of the 32-disparity chunk
    const bool active = cl < TXH;
    const int c = cbase + cl;
    const int rbase = cl - 4 * dq + (DC - 1);  // rrow index for j = 0
    bool valid[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      valid[j] = active && c >= 0 && c < W && (c - (d0 + 4 * dq + j)) >= 0;

    float ring[WIN][4];
    float pre[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < WIN; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) ring[i][j] = 0.f;

    // one input row of the current block: jj (0..WIN-1) is a compile-time constant after
    // unrolling, so the ring indices are static and so are the buffer (jj / HALF -- a block is
    // two half-blocks, hence block-relative half == global half & 1) and the row slot
    auto row_step = [&](int jj) {
      const int slot = jj;  // (jj / HALF) * HALF + jj % HALF
      if (active) {
        const float l = lrow[slot * TXH + cl];
        const float *rr = rrow + slot * Cfg::RPITCH + rbase;
        float out[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float diff = __fsub_rn(l, rr[-j]);
          const float e = valid[j] ? __fmul_rn(diff, diff) : 0.f;
          pre[j] = __fadd_rn(pre[j], e);
          out[j] = (jj < WIN - 1) ? __fadd_rn(ring[(jj + 1) % WIN][j], pre[j]) : pre[j];
          ring[jj][j] = e;
        }
        *reinterpret_cast<float4 *>(vbuf + ((size_t)slot * TXH + cl) * DC + dq * 4) =
            make_float4(out[0], out[1], out[2], out[3]);
      }
    };
    // one half-block: request the next half-block's image rows, sum HALF rows, meet the
    // horizontal warps.  `b` is 0 or 1 at every call site (constant after inlining).
    auto half_step = [&](int tb, int b, bool guarded) {
      stage_half(tb / HALF + b + 1);
#pragma unroll
      for (int i = 0; i < HALF; ++i) {
        const int jj = b * HALF + i;
        if (!guarded || tb + jj < T) row_step(jj);  // block-uniform
      }
      staged_landed();
      named_barrier(Cfg::THREADS);
    };
    // turn the finished block of raw values into suffix sums; restart the prefix
    auto close_block = [&]() {
#pragma unroll
      for (int i = WIN - 2; i >= 1; --i)
#pragma unroll
        for (int j = 0; j < 4; ++j) ring[i][j] = __fadd_rn(ring[i][j], ring[i + 1][j]);
#pragma unroll
      for (int j = 0; j < 4; ++j) pre[j] = 0.f;
    };

    int tb = 0;
    for (; tb + WIN <= T; tb += WIN) {
      half_step(tb, 0, false);
      half_step(tb, 1, false);
      close_block();
    }
    if (tb < T) {  // last, partial block: one barrier per half-block that has a row
      half_step(tb, 0, true);
      if (tb + HALF < T) half_step(tb, 1, true);
    }
  } else {
    // ----------------------------- horizontal role -----------------------------------
    const int q = (tid - Cfg::VTHREADS) >> 5;  // run index
    const int lane = tid & 31;
    const int d = d0 + lane;
    const bool d_ok = d < D;
    const int xr = x0 + q * WIN;
    const bool interior = (xr + WIN <= W) && (xr >= d0 + DC - 1) && (d0 + DC <= D);  // warp-uniform
    auto emit_row = [&](int slot, int y) {
      const float *vb = vbuf + ((size_t)slot * TXH + q * WIN) * DC + lane;
      float v[2 * WIN - 1];
#pragma unroll
      for (int i = 0; i < 2 * WIN - 1; ++i) v[i] = vb[i * DC];
      // suffix sums of block A = v[0..WIN), prefix sums of block B = v[WIN..2WIN-1)
#pragma unroll
      for (int i = WIN - 2; i >= 0; --i) v[i] = __fadd_rn(v[i], v[i + 1]);
#pragma unroll
      for (int i = WIN + 1; i < 2 * WIN - 1; ++i) v[i] = __fadd_rn(v[i], v[i - 1]);
      float *orow = cost + ((size_t)y * W + xr) * D + d;
      if (interior) {
        // the whole run lies inside the image, right of every disparity of the chunk, and the
        // chunk inside [0,D): nothing to test per store
        asm volatile("" : "+l"(orow));  // opaque base: keeps the index arithmetic out of the loop
#pragma unroll
        for (int j = 0; j < WIN; ++j)  // 32-bit offsets: one IMAD.WIDE per address
          __stcs(orow + j * D, (j == 0) ? v[0] : __fadd_rn(v[j], v[WIN + j - 1]));
      } else {
#pragma unroll
        for (int j = 0; j < WIN; ++j) {
          const int x = xr + j;
          const float s = (j == 0) ? v[0] : __fadd_rn(v[j], v[WIN + j - 1]);
          if (d_ok && x < W) __stcs(orow + (size_t)j * D, d > x ? CUDART_INF_F : s);
        }
      }
    };
    for (int g = 0; g < n_half; ++g) {
      named_barrier(Cfg::THREADS);  // half-block g sits in tile buffer g & 1
      const int b = g & 1;
#pragma unroll
      for (int i = 0; i < HALF; ++i) {
        const int t = g * HALF + i;
        const int y = ya - WIN + 1 + t;  // output row completed by input row t
        if (t < T && y >= ya) emit_row(b * HALF + i, y);
      }
    }
  }
}

template <int K, int RUNS>
static int run_ssd_tiled(const float *left, const float *right, float *cost, int H, int W, int D,
                         cudaStream_t stream) {
  using Cfg = SsdCfg<K, RUNS>;
  static PerDeviceInt configured;  // the > 48 KB opt-in is per function AND per device
  const int dev = device_slot();
  int done = 0;
  if (!configured.get(dev, &done)) {
    if (cudaFuncSetAttribute(ssd_tiled_kernel<K, RUNS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)Cfg::SMEM) != cudaSuccess) {
      cudaGetLastError();
      return 1;  // fall back to the generic kernel
    }
    configured.set(dev, 1);
  }
  const int nx = (W + Cfg::TX - 1) / Cfg::TX;
  const int nd = (D + Cfg::DC - 1) / Cfg::DC;
  // Rows are split into chunks of ~128 output rows (each chunk pays WIN-1 warm-up rows).  Many
  // short CTAs beat one full-height CTA per SM by a wide margin -- 1080p D=128: 0.507 ms with
  // 1 chunk, 0.420 with 2, 0.383 with 4, 0.373 with 8 or 10 (profiles/r01_v18_ssd.txt): the
  // block scheduler balances them over all SMs, and CTAs working on different image rows
  // spread the store stream over the DRAM instead of all writing the same ~1 MB at once.
  // Small images: at least one CTA per SM as long as a chunk keeps >= 32 rows.
  int ny = (H + 64) / 128;
  const int fill = (num_sms() + nx * nd - 1) / (nx * nd);
  if (ny < fill) ny = fill;
  if (ny > (H + 31) / 32) ny = (H + 31) / 32;
  if (ny < 1) ny = 1;
  {
    static const char *force = getenv("STM_DEBUG_SSD_NY");  // developer knob
    if (force && atoi(force) > 0) ny = atoi(force);
  }
  const int rows = (H + ny - 1) / ny;
  ny = (H + rows - 1) / rows;
  if (nd > 65535 || ny > 65535) return fail(STM_EUNSUPPORTED, "ssd: shape too large");
  ssd_tiled_kernel<K, RUNS><<<dim3(nx, nd, ny), Cfg::THREADS, Cfg::SMEM, stream>>>(left, right, cost,
                                                                                  H, W, D, rows);
  return check_launch("ssd_tiled_kernel");
}

int launch_ssd_tiled(const float *left, const float *right, float *cost, int H, int W, int D,
                     int k, cudaStream_t stream) {
  switch (k) {
    case 7: return run_ssd_tiled<7, 4>(left, right, cost, H, W, D, stream);
    case 3: return run_ssd_tiled<3, 8>(left, right, cost, H, W, D, stream);
    default: return 1;  // not handled here
  }
}

}  // namespace stm
