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
// walking down the rows of its y-range.  Two warp roles meet at one barrier per row:
//  * vertical warps: thread = (column c, 4 consecutive d).  Per input row it squares 4
//    differences (image rows staged in shared memory, conflict-free reads) and advances a
//    2K-deep van Herk ring held in REGISTERS (fully unrolled, static indices); the
//    vertical window sum of the row that just completed goes to shared memory as one
//    float4 (a warp writes 512 contiguous bytes).
//  * horizontal warps: lane = disparity, warp = a run of 2K output columns.  Reads the
//    4K-1 vertical sums it needs (128-byte rows, conflict-free), does the prefix/suffix
//    blocks in registers and streams 2K fully-coalesced 128-byte stores into C.
// The two roles work on consecutive rows concurrently (double-buffered tile).
// HBM traffic = the 4 B/pix-disp store of C plus the (L2-resident) images: the kernel is
// bound by the store stream, shared-memory wavefronts (~4.8 words/output) and FP32 issue
// in roughly equal parts -- see DESIGN.md.
#include "stm_common.cuh"

namespace stm {

template <int K, int RUNS>
struct SsdCfg {
  static constexpr int WIN = 2 * K;
  static constexpr int TX = RUNS * WIN;         // output columns per CTA
  static constexpr int TXH = TX + WIN - 1;      // columns with halo
  static constexpr int DC = 32;                 // disparities per CTA
  static constexpr int VTHREADS = ((TXH * 8 + 31) / 32) * 32;
  static constexpr int HWARPS = RUNS;
  static constexpr int THREADS = VTHREADS + HWARPS * 32;
  static constexpr int RROW = TXH + DC - 1;     // staged right-image span
};

__device__ __forceinline__ void named_barrier(int threads) {
  asm volatile("bar.sync 1, %0;" ::"r"(threads) : "memory");
}

template <int K, int RUNS>
__global__ void __launch_bounds__(SsdCfg<K, RUNS>::THREADS, 1)
ssd_tiled_kernel(const float *__restrict__ left, const float *__restrict__ right,
                 float *__restrict__ cost, int H, int W, int D, int rows_per_chunk) {
  using Cfg = SsdCfg<K, RUNS>;
  constexpr int WIN = Cfg::WIN, TXH = Cfg::TXH, DC = Cfg::DC;

  __shared__ __align__(16) float vbuf[2][TXH][DC];
  __shared__ float lrow[2][TXH];
  __shared__ float rrow[2][Cfg::RROW + 1];

  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * Cfg::TX;
  const int d0 = blockIdx.y * DC;
  const int ya = blockIdx.z * rows_per_chunk;
  const int yb = min(H, ya + rows_per_chunk);
  const int T = (yb - ya) + WIN - 1;   // input rows ya-K .. yb+K-2
  const int cbase = x0 - K;            // image column of staged column 0
  const int ubase = cbase - d0 - (DC - 1);  // right-image column of rrow[.][0]

  // Cooperative staging of the image rows (zeros outside the image), software-pipelined through
  // one register: a row is LOADED two iterations before it is used and STORED to shared memory
  // one iteration before, so no global-load latency sits between two row barriers.  The load is
  // unconditional (clamped address, value masked when it is stored): a load under a branch
  // would be waited for at the join.  Threads [0,TXH) stage the left row, the next RROW the
  // right row; the other vertical threads load a clamped address and store nothing.
  static_assert(Cfg::VTHREADS >= TXH + Cfg::RROW, "not enough loader threads");
  const bool st_left = tid < TXH;
  const bool st_any = tid < TXH + Cfg::RROW;
  const int st_col = st_left ? cbase + tid : ubase + (tid - TXH);
  const bool st_col_ok = st_any && st_col >= 0 && st_col < W;
  const float *st_src = (st_left ? left : right) + min(max(st_col, 0), W - 1);
  float *st_dst0 = st_left ? &lrow[0][tid] : &rrow[0][st_any ? tid - TXH : 0];
  const int st_stride = st_left ? TXH : Cfg::RROW + 1;  // floats between the two buffers
  auto fetch = [&](int t) -> float {  // raw value of staged row t (any t)
    const int r = min(max(ya - K + t, 0), H - 1);
    return __ldg(st_src + (size_t)r * W);
  };
  auto put = [&](int t, float raw) {  // store staged row t into buffer t&1
    const int r = ya - K + t;
    const bool ok = st_col_ok && r >= 0 && r < H;
    if (st_any) st_dst0[(t & 1) * st_stride] = ok ? raw : 0.f;
  };

  float staged = 0.f;  // raw value of row t+1 while iteration t runs (vertical role only)
  if (tid < Cfg::VTHREADS) {
    put(0, fetch(0));
    staged = fetch(1);
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

    // one input row: jj is a compile-time constant after unrolling (static ring indices)
    auto row_step = [&](int t, int jj) {
      const int buf = t & 1;
      // row t+1 (loaded one iteration ago) goes to shared memory, row t+2 is requested.  Past
      // the last row both are harmless: the address is clamped and nobody reads the buffer.
      put(t + 1, staged);
      staged = fetch(t + 2);
      float out[4];
      if (active) {
        const float l = lrow[buf][cl];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float diff = __fsub_rn(l, rrow[buf][rbase - j]);
          const float e = valid[j] ? __fmul_rn(diff, diff) : 0.f;
          pre[j] = __fadd_rn(pre[j], e);
          out[j] = (jj < WIN - 1) ? __fadd_rn(ring[(jj + 1) % WIN][j], pre[j]) : pre[j];
          ring[jj][j] = e;
        }
        *reinterpret_cast<float4 *>(&vbuf[buf][cl][dq * 4]) =
            make_float4(out[0], out[1], out[2], out[3]);
      }
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

    // Full blocks run without any `t < T` test: a load whose result is merged at the join of
    // a branch is waited for at that join, which would put the global latency of `fetch`
    // back between two barriers.  Only the last, partial block is guarded.
    int tb = 0;
    for (; tb + WIN <= T; tb += WIN) {
#pragma unroll
      for (int jj = 0; jj < WIN; ++jj) row_step(tb + jj, jj);
      close_block();
    }
    if (tb < T) {
#pragma unroll
      for (int jj = 0; jj < WIN; ++jj)
        if (tb + jj < T) row_step(tb + jj, jj);  // block-uniform
    }
  } else {
    // ----------------------------- horizontal role -----------------------------------
    const int q = (tid - Cfg::VTHREADS) >> 5;  // run index
    const int lane = tid & 31;
    const int d = d0 + lane;
    const bool d_ok = d < D;
    const int xr = x0 + q * WIN;
    const bool interior = (xr + WIN <= W) && (xr >= d0 + DC - 1) && (d0 + DC <= D);  // warp-uniform
    auto emit = [&](int t) {
      // row produced by the vertical warps in iteration t
      const int y = ya - WIN + 1 + t;
      if (y < ya) return;
      const int buf = t & 1;
      float v[2 * WIN - 1];
#pragma unroll
      for (int i = 0; i < 2 * WIN - 1; ++i) v[i] = vbuf[buf][q * WIN + i][lane];
      // suffix sums of block A = v[0..WIN), prefix sums of block B = v[WIN..2WIN-1)
#pragma unroll
      for (int i = WIN - 2; i >= 0; --i) v[i] = __fadd_rn(v[i], v[i + 1]);
#pragma unroll
      for (int i = WIN + 1; i < 2 * WIN - 1; ++i) v[i] = __fadd_rn(v[i], v[i - 1]);
      float *orow = cost + ((size_t)y * W + xr) * D + d;
      if (interior) {
        // the whole run lies inside the image, right of every disparity of the chunk, and the
        // chunk inside [0,D): nothing to test per store (these 4 warps are the serial part of
        // a row, ~200 instructions each with the tests, ~110 without)
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
    for (int t = 0; t < T; ++t) {
      if (t > 0) emit(t - 1);
      named_barrier(Cfg::THREADS);
    }
    emit(T - 1);
  }
}

template <int K, int RUNS>
static int run_ssd_tiled(const float *left, const float *right, float *cost, int H, int W, int D,
                         cudaStream_t stream) {
  using Cfg = SsdCfg<K, RUNS>;
  const int nx = (W + Cfg::TX - 1) / Cfg::TX;
  const int nd = (D + Cfg::DC - 1) / Cfg::DC;
  // One CTA per SM, so the launch runs in ceil(CTAs / SMs) waves of equal cost.  Split the rows
  // into ny chunks (each pays WIN-1 warm-up rows) so that waves x rows-per-chunk is smallest:
  // D=128 at 1080p is 140 CTAs = one wave (ny = 1); D=192 is 210 CTAs = two waves of full
  // height, but 420 half-height CTAs = three waves, 1.5 instead of 2 units.
  const int sms = num_sms();
  int ny = 1;
  long best = -1;
  for (int cand = 1; cand <= 8; ++cand) {
    const int rows_c = (H + cand - 1) / cand;
    if (cand > 1 && rows_c < 64) break;  // >= 64 rows per chunk amortises the warm-up rows
    const long waves = ((long)nx * nd * cand + sms - 1) / sms;
    const long cost = waves * (rows_c + Cfg::WIN - 1);
    if (best < 0 || cost < best) { best = cost; ny = cand; }
  }
  const int rows = (H + ny - 1) / ny;
  ny = (H + rows - 1) / rows;
  if (nd > 65535 || ny > 65535) return fail(STM_EUNSUPPORTED, "ssd: shape too large");
  ssd_tiled_kernel<K, RUNS><<<dim3(nx, nd, ny), Cfg::THREADS, 0, stream>>>(left, right, cost, H, W,
                                                                          D, rows);
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
