// ssd_rows.cu -- SSD cost volume, horizontal sums first, packed arithmetic (sm_100a).
//
// Same function as ssd_tiled.cu (reference src/ssd.cu:31-61: C[y,x,d] = sum over the 2K x 2K window of
// (L[r,c] - R[r,c-d])^2, clipped to the image, +inf for d > x) and the same add-only van Herk /
// Gil-Werman sums (no subtraction: bit-exact for integer-valued images, a few ulp otherwise).
// ncu of ssd_tiled_kernel at 1080p D=128 (profiles/r02_v2_ncu_full_summary.txt): 22 warp
// instructions per 128-byte store, 57 % of the issue slots busy, a third of all warp samples at the
// barrier between its two warp roles -- 0.36 ms where the store stream alone needs ~0.17.  Doubling
// the horizontal warps bought 4 %: the kernel has too many instructions, not too few warps.
//
// Here the order of the two 1-D sums is swapped and everything is packed (add/sub/mul.rn.f32x2):
//  * H warps: lane = disparity, a work item = (a PAIR of image rows, a run of 2K output columns).
//    The two rows are staged INTERLEAVED in shared memory (float2 per column), so one 8-byte load
//    fetches L (broadcast) or R (one per lane) of both rows and every operation of the item is a
//    packed pair: 27 squared differences, the prefix/suffix blocks, 2K window sums per row.
//  * V threads: thread = (output column, 4 consecutive disparities).  It reads the horizontal sums
//    of a row as one float4, advances the vertical van Herk ring held in registers (a group of rows
//    IS a van Herk block of 2K rows, so all ring indices are static) and writes 16 bytes of C with a
//    streaming store; 8 threads write one pixel's 128-byte chunk.
// The two roles work on consecutive groups through a double-buffered tile and meet at one barrier
// per group; image rows are requested with cp.async two groups ahead.  HBM traffic = the 4 B/pix-disp
// store of C plus the images.
//
// What was measured (1080p D=128, profiles/r02_v6_ssd_investigation.md).  First version: 0.350 ms
// against 0.371 for ssd_tiled_kernel.  A store-only kernel with exactly this write pattern
// (128-byte chunks of each pixel's vector from four CTAs) runs in 0.164 ms, so the pattern is not the
// limit.  r02 v11 (ncu per-line counts, profiles/r02_v11_ssd_ncu.md): two thirds of the V role's
// instruction stream were a 64-bit y*W*D product and four range tests per row, and the H warps
// loaded every van Herk block twice (once per adjacent run): branch-free row loop with a pointer
// bump, staging with one add per element, and STREAMING row-pair items (ssd_rows_pair: each block
// loaded once, square and sum as one packed FMA, left values as 16-byte broadcast loads)
// -> 0.349 -> 0.282 ms (D=192: 0.499 -> 0.388).  The kernel is now bound by the shared-memory /
// LSU data path: 73 % of its wavefront slots are busy (5.1 wavefronts per 128 bytes of C: one for
// the global store, two for the H -> V hand-off, two for the image values), 18 % of the warp
// samples wait at the role barrier.  Variants that were measured and are worse: plain instead of
// streaming stores (0.43), 11 H warps (0.46, spills), V warps helping with H items (0.61, spills),
// 8 / 9 H warps (0.328 / 0.362), staging dealt out statically per V warp (6 registers of state
// instead of 18, no spills, but 0.310: eleven unrolled variants of the staging code).
#include <algorithm>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <vector>

#include "stm_common.cuh"

#ifndef STM_SSD_NH7
#define STM_SSD_NH7 7   /* 11 H warps: 80-register cap, spills, 0.36 -> 0.46 ms (measured) */
#endif

namespace stm {
namespace ssdr {

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk(u64 r, float &a, float &b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r));
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
  u64 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ u64 sub2(u64 a, u64 b) {
  u64 d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
  u64 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

template <int K, int RUNS, int NH>
struct Cfg {
  static constexpr int WIN = 2 * K;            // window, van Herk block, rows per group
  static constexpr int PAIRS = WIN / 2;        // row pairs per group
  static constexpr int TX = RUNS * WIN;        // output columns per CTA
  static constexpr int TXH = TX + WIN - 1;     // staged left columns (halo)
  static constexpr int DC = 32;                // disparities per CTA
  static constexpr int RROW = TXH + DC - 1;    // staged right columns
  static constexpr int VTHREADS = ((TX * 8 + 31) / 32) * 32;
  static constexpr int THREADS = VTHREADS + NH * 32;
  static constexpr int ITEMS = PAIRS * RUNS;   // item-pairs per group
  // The V warps have ~1/3 of an H warp's instructions per group: each of the first V_ITEMS of them
  // also takes one item, so that both roles reach the barrier at about the same time.
  // (measured: with V_ITEMS = ITEMS / 3 the V role spills and becomes the critical path, 0.36 -> 0.61 ms)
  static constexpr int V_ITEMS = 0;
  static constexpr int H_ITEMS = ITEMS - V_ITEMS;
  // one streaming item (a whole row pair) per H warp and group where that keeps every H warp busy
#ifndef STM_SSD_STREAM_PAIRS
#define STM_SSD_STREAM_PAIRS 1
#endif
  static constexpr bool STREAM_PAIRS = STM_SSD_STREAM_PAIRS && V_ITEMS == 0 && PAIRS >= NH;
  // shared memory, in floats: hbuf[2][WIN][TX][DC], rows[3][PAIRS][RSTRIDE] float2
  static constexpr int HBUF = WIN * TX * DC;
  static constexpr int RSTRIDE = (TXH + RROW + 1) & ~1;   // float2 per staged row pair; even: 16-byte loads of two columns
  static constexpr int ROWBUF = PAIRS * RSTRIDE * 2;
  static constexpr size_t SMEM = (size_t)(2 * HBUF + 3 * ROWBUF) * sizeof(float);
};

__device__ __forceinline__ void named_barrier(int threads) {
  asm volatile("bar.sync 1, %0;" ::"r"(threads) : "memory");
}

}  // namespace ssdr

// One horizontal work item: row pair p, run q of group buffer `rb` -> horizontal sums in `hb`.
template <int K, int RUNS, int NH>
__device__ __forceinline__ void ssd_rows_item(const float2 *rb, float *hb, int item, int lane, int x0, int d0, int W) {
  using namespace ssdr;
  using C = Cfg<K, RUNS, NH>;
  constexpr int WIN = C::WIN, TX = C::TX, TXH = C::TXH, DC = C::DC;
  const int dglob = d0 + lane;
  const int p = item / RUNS, q = item - p * RUNS;
  const float2 *lp = rb + p * C::RSTRIDE + q * WIN;                            // + k
  const float2 *rp = rb + p * C::RSTRIDE + TXH + q * WIN + (DC - 1) - lane;    // + k
  const int c0 = x0 + q * WIN - K;  // image column of k = 0
  // warp-uniform: every column of the item lies in [0, W) and right of every disparity
  const bool interior = c0 >= d0 + DC - 1 && c0 + 2 * WIN - 2 < W;
  u64 v[2 * WIN - 1];  // (row 2p, row 2p+1) squared differences of column k
#pragma unroll
  for (int k = 0; k < 2 * WIN - 1; ++k) {
    const u64 l2 = *reinterpret_cast<const u64 *>(lp + k);
    const u64 r2 = *reinterpret_cast<const u64 *>(rp + k);
    const u64 df = sub2(l2, r2);
    v[k] = mul2(df, df);
  }
  if (!interior) {
#pragma unroll
    for (int k = 0; k < 2 * WIN - 1; ++k) {
      const int c = c0 + k;
      if (!(c < W && c - dglob >= 0)) v[k] = 0ull;  // window clipped to [d, W) (reference :40-42)
    }
  }
  // suffix sums of block A = v[0..WIN), prefix sums of block B = v[WIN..2WIN-1)
#pragma unroll
  for (int i = WIN - 2; i >= 0; --i) v[i] = add2(v[i], v[i + 1]);
#pragma unroll
  for (int i = WIN + 1; i < 2 * WIN - 1; ++i) v[i] = add2(v[i], v[i - 1]);
  float *o0 = hb + ((size_t)(2 * p) * TX + q * WIN) * DC + lane;
  float *o1 = o0 + (size_t)TX * DC;
#pragma unroll
  for (int j = 0; j < WIN; ++j) {
    float s0, s1;
    upk(j == 0 ? v[0] : add2(v[j], v[WIN + j - 1]), s0, s1);
    o0[j * DC] = s0;
    o1[j * DC] = s1;
  }
}

// One horizontal work item of the streaming form: row pair p, ALL runs of group buffer `rb`.
// ssd_rows_item loads 2*WIN-1 columns per run of WIN outputs: consecutive runs share a van Herk block
// and load it twice (81 instead of 55 column loads per row pair at K = 7, RUNS = 3).  Here the warp
// walks the blocks of the row pair once: the finished block's suffix sums stay in registers, the next
// block's values are added into a running prefix as they are combined.  Same additions in the same
// order as ssd_rows_item (bit-identical), 17 % fewer instructions and 23 % fewer shared-memory
// wavefronts per group in the H role; one item per H warp and group when PAIRS == NH.
template <int K, int RUNS, int NH>
__device__ __forceinline__ void ssd_rows_pair(const float2 *rb, float *hb, int p, int lane, int x0, int d0, int W) {
  using namespace ssdr;
  using C = Cfg<K, RUNS, NH>;
  constexpr int WIN = C::WIN, TX = C::TX, TXH = C::TXH, DC = C::DC;
  const int dglob = d0 + lane;
  const float2 *lp = rb + p * C::RSTRIDE;                   // + k: image column x0 - K + k
  const float2 *rp = lp + TXH + (DC - 1) - lane;            // + k: the same column, d0 + lane to the left
  const int c0 = x0 - K;
  // s: suffix sums of the finished block; f: the DIFFERENCES of the next block.  Their squares are
  // never formed on their own: each one enters a sum, so the square and the addition are one packed
  // FMA (exact, hence bit-identical to mul + add, for integer-valued images; one rounding instead of
  // two -- a last-bit difference -- otherwise).
  u64 s[WIN], f[WIN];
  auto load_block = [&](int b, u64 (&v)[WIN]) {             // block RUNS (the last) has WIN - 1 columns
    static_assert(WIN % 2 == 0, "two columns per 16-byte load");
#pragma unroll
    for (int k = 0; k < WIN; k += 2) {
      // the left values are the same for every lane: one 16-byte broadcast load fetches two columns
      // (both rows each); the last pair of the last block reads one staged column past the left ones
      const ulonglong2 l4 = *reinterpret_cast<const ulonglong2 *>(lp + b * WIN + k);
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        if (b == RUNS && k + kk == WIN - 1) { v[k + kk] = 0ull; continue; }
        const u64 r2 = *reinterpret_cast<const u64 *>(rp + b * WIN + k + kk);
        v[k + kk] = sub2(kk ? l4.y : l4.x, r2);
      }
    }
    const int cb = c0 + b * WIN;
    // warp-uniform: every column of the block lies in [0, W) and right of every disparity
    const bool interior = cb >= d0 + DC - 1 && cb + WIN - 1 < W;
    if (!interior) {
#pragma unroll
      for (int k = 0; k < WIN; ++k) {
        const int c = cb + k;
        if (!(c < W && c - dglob >= 0)) v[k] = 0ull;        // window clipped to [d, W) (reference :40-42)
      }
    }
  };
  load_block(0, f);
  s[WIN - 1] = mul2(f[WIN - 1], f[WIN - 1]);
#pragma unroll
  for (int i = WIN - 2; i >= 0; --i) s[i] = fma2(f[i], f[i], s[i + 1]);
  float *o0 = hb + ((size_t)(2 * p) * TX) * DC + lane;
  float *o1 = o0 + (size_t)TX * DC;
#pragma unroll
  for (int q = 0; q < RUNS; ++q) {
    load_block(q + 1, f);
    u64 run = 0ull;
#pragma unroll
    for (int j = 0; j < WIN; ++j) {
      float s0, s1;
      upk(j == 0 ? s[0] : add2(s[j], run), s0, s1);
      o0[(q * WIN + j) * DC] = s0;
      o1[(q * WIN + j) * DC] = s1;
      run = j == 0 ? mul2(f[0], f[0]) : fma2(f[j], f[j], run);   // the prefix of the next block
    }
    if (q + 1 < RUNS) {
      s[WIN - 1] = mul2(f[WIN - 1], f[WIN - 1]);
#pragma unroll
      for (int i = WIN - 2; i >= 0; --i) s[i] = fma2(f[i], f[i], s[i + 1]);
    }
  }
}

// WTA = true (stm_ssd_wta_f32): the V threads do not store C; they take the first minimum of their
// four entries, combine the eight threads of a pixel by shuffles and send one 64-bit key
// (cost bits << 32 | disparity) per pixel, row and disparity chunk to `keys` with an atomic minimum:
// costs are >= +0, so the unsigned order of the keys is (cost, then disparity) -- the first-minimum
// rule of wta.cu / reference src/winners_take_all.cu:21-40 across the chunks of all CTAs.
template <int K, int RUNS, int NH, bool WTA>
__global__ void __launch_bounds__(ssdr::Cfg<K, RUNS, NH>::THREADS, 1)
ssd_rows_kernel(const float *__restrict__ left, const float *__restrict__ right, float *__restrict__ cost,
                unsigned long long *__restrict__ keys, int H, int W, int D, int rows_per_chunk, int d_fastest) {
  using namespace ssdr;
  using C = Cfg<K, RUNS, NH>;
  constexpr int WIN = C::WIN, PAIRS = C::PAIRS, TX = C::TX, TXH = C::TXH, DC = C::DC, RROW = C::RROW;
  extern __shared__ __align__(16) float smem[];
  float *hbuf = smem;                 // [(b*WIN + row)*TX + x][DC]
  float *rows = smem + 2 * C::HBUF;   // [b3][pair][RSTRIDE] float2: left columns, then right columns

  const int tid = threadIdx.x;
  const int x0 = (d_fastest ? blockIdx.y : blockIdx.x) * TX;
  const int d0 = (d_fastest ? blockIdx.x : blockIdx.y) * DC;
  const int ya = blockIdx.z * rows_per_chunk;
  const int yb = min(H, ya + rows_per_chunk);
  const int T = (yb - ya) + WIN - 1;          // input rows ya-K .. yb+K-2
  const int n_groups = (T + WIN - 1) / WIN;
  const int cbase = x0 - K;                   // image column of staged left column 0
  const int ubase = cbase - d0 - (DC - 1);    // image column of staged right column 0

  if (tid < C::VTHREADS) {
    // ------------------------------------------------------------------ V role (+ staging)
    // Rows of group g: image rows ya - K + g*WIN + i, i = 0..WIN-1, interleaved in pairs; pixels
    // outside the image are zero-filled by a source size of 0 (the address is clamped to stay valid).
    // A thread copies the same (row-in-group, column) elements in every group: their source column
    // pointer, destination offset and column validity are worked out once (the first version did
    // the index arithmetic per element and group: 47 instructions per 4 bytes, a quarter of the
    // kernel's instructions).
    constexpr int ROWLEN = TXH + RROW, NE = (WIN * ROWLEN + C::VTHREADS - 1) / C::VTHREADS;
    uintptr_t st_addr[NE];     // address of the element in group 0 (row ya - K + i; formed as an integer, only
                               // used for groups whose rows all lie inside the image)
    uint32_t st_dst[NE];       // byte offset in a rows buffer; bit 31: column outside the image (zero-fill);
                               // 0xffffffff: no element (tail)
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      const int idx = tid + e * C::VTHREADS;
      const int i = idx / ROWLEN, col = idx - i * ROWLEN;
      const bool is_left = col < TXH;
      const int c = is_left ? cbase + col : ubase + (col - TXH);
      // the column is clamped so that a zero-filling copy (source size 0) still names a valid address
      st_addr[e] = reinterpret_cast<uintptr_t>((is_left ? left : right) + min(max(c, 0), W - 1)) +
                   (uintptr_t)((long long)(ya - K + i) * W * (long long)sizeof(float));
      st_dst[e] = idx < WIN * ROWLEN ? (uint32_t)((((i >> 1) * C::RSTRIDE + col) * 2 + (i & 1)) * sizeof(float)) |
                                           ((c >= 0 && c < W) ? 0u : 0x80000000u)
                                     : 0xffffffffu;
    }
    const uint32_t rows_s = (uint32_t)__cvta_generic_to_shared(rows);
    auto stage_group = [&](int g) {
      const uint32_t dst0 = rows_s + (uint32_t)((g % 3) * C::ROWBUF * sizeof(float));
      const int r0 = ya - K + g * WIN;
      if (r0 >= 0 && r0 + WIN <= H) {
        // every row of the group is an image row (block-uniform): address = group-0 address + g groups,
        // 5 instructions per element (the general body below: ~20, a quarter of the V role's stream)
        const uintptr_t goff = (uintptr_t)((long long)g * WIN * W * (long long)sizeof(float));
#pragma unroll
        for (int e = 0; e < NE; ++e) {
          if (st_dst[e] != 0xffffffffu)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst0 + (st_dst[e] & 0x7fffffffu)),
                         "l"(st_addr[e] + goff), "r"((int)st_dst[e] >= 0 ? 4 : 0) : "memory");
        }
      } else {
#pragma unroll
        for (int e = 0; e < NE; ++e) {
          if (st_dst[e] != 0xffffffffu) {
            const int i = (tid + e * C::VTHREADS) / ROWLEN;
            const int r = r0 + i;
            const int ok_bytes = ((int)st_dst[e] >= 0 && r >= 0 && r < H) ? 4 : 0;
            // same column, row clamped into the image
            const uintptr_t src = st_addr[e] + (uintptr_t)((long long)(min(max(r, 0), H - 1) - (ya - K + i)) * W * (long long)sizeof(float));
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst0 + (st_dst[e] & 0x7fffffffu)), "l"(src), "r"(ok_bytes) : "memory");
          }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    stage_group(0);
    stage_group(1);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    named_barrier(C::THREADS);  // group 0's rows are in shared memory

    const int xl = tid >> 3, dq = tid & 7;
    const int x = x0 + xl;
    const bool active = xl < TX && x < W;
    const int dfirst = d0 + 4 * dq;
    const bool plain = active && dfirst + 3 < D && dfirst + 3 <= x;   // nothing to test per element
    // WTA: the threads of a warp shuffle with the FULL mask, so a warp branches as one: its four pixels
    // x_w .. x_w + 3 all exist and need no test.  (Shuffles under the 8-lane mask of a pixel split the
    // warp into four groups that then run the row loop one after the other: 0.49 ms instead of 0.3.)
    const int xw = x0 + (tid >> 5) * 4;
    const bool warp_has_pixels = (tid >> 5) * 4 < TX && xw < W;
    const bool warp_plain = (tid >> 5) * 4 + 3 < TX && xw + 3 < W && d0 + DC - 1 < D && d0 + DC - 1 <= xw;
    float ring[WIN][4];
    float pre[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < WIN; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) ring[i][j] = 0.f;
    float *ocol = cost + ((size_t)x * D + dfirst);   // + y * W * D
    const size_t rowstride = (size_t)W * D;
    // WTA: one key per pixel, row and disparity chunk.  `bounded` = entries d > x (+inf) or d >= D
    // (absent) have to be skipped.
    auto emit_key = [&](int y, const float (&out)[4], bool bounded) {
      float best = CUDART_INF_F;
      unsigned bd = (unsigned)dfirst;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int d = dfirst + j;
        if ((!bounded || (d < D && d <= x)) && out[j] < best) {  // strict: the lowest d keeps a tie; NaN never wins
          best = out[j];
          bd = (unsigned)d;
        }
      }
      unsigned hi = __float_as_uint(best), lo = bd;
      if (dfirst == 0 && out[0] != out[0]) hi = 0u;              // NaN at d = 0: the reference answers 0
      if (dfirst >= D || !active) { hi = 0xffffffffu; lo = 0xffffffffu; }   // padding of the last chunk / no pixel
      // The minimum cost of the pixel's eight threads by three shuffles of the cost bits; every thread that
      // holds it sends its key (normally one; on a tie the atomic keeps the lowest disparity).  Measured
      // at 1080p D=128: a butterfly over (cost, d) pairs -- six shuffles per row through the already
      // saturated LSU/MIO path -- 0.516 ms (0.354 without any shuffle, 0.485 without the atomics);
      // electing the owner with a vote, 0.75 ms, and redux.sync.min.u32, 0.82 ms: both are slow with
      // the four 8-lane masks of a warp.
      unsigned m = hi;
#pragma unroll
      for (int w = 1; w < 8; w <<= 1) m = min(m, __shfl_xor_sync(FULL, m, w));   // stays inside the pixel's 8 lanes
      if (hi == m && active) atomicMin(keys + ((size_t)y * W + x), ((unsigned long long)hi << 32) | lo);
    };

    for (int g = 0; g <= n_groups; ++g) {
      stage_group(g + 2);
      if (g >= 1 && (WTA ? warp_has_pixels : active)) {
        // (WTA: a thread without a pixel reads the tile's last column and sends nothing)
        const float *hb = hbuf + (size_t)(((g - 1) & 1) * C::HBUF) + (size_t)(WTA ? min(xl, TX - 1) : xl) * DC + 4 * dq;
        const int ybase = ya - WIN + 1 + (g - 1) * WIN;  // output row completed by row 0 of the group
        // One row: read the horizontal sums, advance prefix and ring, form the window sum.
        auto row_sum = [&](int i, float (&out)[4]) {
          const float4 h = *reinterpret_cast<const float4 *>(hb + (size_t)i * TX * DC);
          const float hv[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
          for (int j = 0; j < 4; j += 2) {
            u64 p2 = add2(pk(pre[j], pre[j + 1]), pk(hv[j], hv[j + 1]));
            upk(p2, pre[j], pre[j + 1]);
            if (i < WIN - 1) p2 = add2(p2, pk(ring[i + 1][j], ring[i + 1][j + 1]));
            upk(p2, out[j], out[j + 1]);
            ring[i][j] = hv[j];
            ring[i][j + 1] = hv[j + 1];
          }
        };
        if ((WTA ? warp_plain : plain) && ybase >= ya && ybase + WIN <= yb) {
          // Every row of the group is an output row of this chunk and none of the four entries needs
          // a test: no per-row branch and one pointer bump per row.  (The general body below re-did a
          // 64-bit y*W*D product and four range tests per row: 26 instructions per row on this path
          // where 9 do the work -- two thirds of the V role's instruction stream.)
          float *o = WTA ? nullptr : ocol + (size_t)ybase * rowstride;
#pragma unroll
          for (int i = 0; i < WIN; ++i) {
            float out[4];
            row_sum(i, out);
            if constexpr (WTA) {
              emit_key(ybase + i, out, false);
            } else {
              // streaming (evict-first) store: plain and write-through stores measured 0.43 vs 0.385 ms
              __stcs(reinterpret_cast<float4 *>(o), make_float4(out[0], out[1], out[2], out[3]));
              o += rowstride;
            }
          }
        } else if (!WTA && dfirst + 3 < D && ybase >= ya && ybase + WIN <= yb) {
          // Left image border (some d > x): the same branch-free loop, the entries beyond x replaced by
          // +inf with selects and still ONE 16-byte store.  Four scalar stores with a test each made the
          // CTAs of the first column tiles several times slower than the others -- the critical path of
          // a launch with few waves (1080p D=64 with long row chunks).
          float *o = ocol + (size_t)ybase * rowstride;
#pragma unroll
          for (int i = 0; i < WIN; ++i) {
            float out[4];
            row_sum(i, out);
#pragma unroll
            for (int j = 0; j < 4; ++j) out[j] = dfirst + j > x ? CUDART_INF_F : out[j];
            __stcs(reinterpret_cast<float4 *>(o), make_float4(out[0], out[1], out[2], out[3]));
            o += rowstride;
          }
        } else {
#pragma unroll
          for (int i = 0; i < WIN; ++i) {
            float out[4];
            row_sum(i, out);
            const int y = ybase + i;
            if (y >= ya && y < yb) {  // block-uniform
              if constexpr (WTA) {
                emit_key(y, out, true);
                continue;
              }
              float *o = ocol + (size_t)y * rowstride;
              if (plain) {
                __stcs(reinterpret_cast<float4 *>(o), make_float4(out[0], out[1], out[2], out[3]));
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (dfirst + j < D) __stcs(o + j, dfirst + j > x ? CUDART_INF_F : out[j]);
              }
            }
          }
        }
        // the finished block of raw values becomes suffix sums; the prefix restarts
#pragma unroll
        for (int i = WIN - 2; i >= 1; --i)
#pragma unroll
          for (int j = 0; j < 4; j += 2) {
            const u64 s2 = add2(pk(ring[i][j], ring[i][j + 1]), pk(ring[i + 1][j], ring[i + 1][j + 1]));
            upk(s2, ring[i][j], ring[i][j + 1]);
          }
#pragma unroll
        for (int j = 0; j < 4; ++j) pre[j] = 0.f;
      }
      if (C::V_ITEMS > 0 && g < n_groups && (tid >> 5) < C::V_ITEMS)
        ssd_rows_item<K, RUNS, NH>(reinterpret_cast<const float2 *>(rows + (g % 3) * C::ROWBUF),
                                   hbuf + (size_t)((g & 1) * C::HBUF), C::H_ITEMS + (tid >> 5), tid & 31, x0, d0, W);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      named_barrier(C::THREADS);
    }
  } else {
    // ------------------------------------------------------------------ H role
    const int hw = (tid - C::VTHREADS) >> 5, lane = tid & 31;
    named_barrier(C::THREADS);  // group 0's rows are in shared memory
    for (int g = 0; g <= n_groups; ++g) {
      if (g < n_groups) {
        const float2 *rb = reinterpret_cast<const float2 *>(rows + (g % 3) * C::ROWBUF);
        float *hb = hbuf + (size_t)((g & 1) * C::HBUF);
        if constexpr (C::STREAM_PAIRS) {
          for (int pr = hw; pr < PAIRS; pr += NH) ssd_rows_pair<K, RUNS, NH>(rb, hb, pr, lane, x0, d0, W);
        } else {
          // items [0, H_ITEMS) belong to the H warps; the V warps take the rest after their own rows
          for (int item = hw; item < C::H_ITEMS; item += NH)
            ssd_rows_item<K, RUNS, NH>(rb, hb, item, lane, x0, d0, W);
        }
      }
      named_barrier(C::THREADS);
    }
  }
}

// Groups of WIN rows per row chunk (a chunk of m groups writes m*WIN - (WIN-1) output rows).  A CTA costs
// one period per group plus ~1.5 for fill and drain, and the CTAs of one launch -- per_z of them per
// chunk, the last chunk shorter -- are handed to the SMs in launch order: the makespan of that greedy
// schedule is evaluated for every m and the best one kept (ties: the larger m).  The fixed m = 20 left
// the last wave half empty where few CTAs exist per chunk (1080p D=32: 230 CTAs on 148 SMs, two waves
// for 1.2 waves of work).
static int ssd_pick_groups(int H, int per_z, int win, int sms) {
  struct Entry { int H, per_z, win, sms, m; };
  static std::mutex mu;
  static Entry cache[8];
  static int used = 0, next = 0;
  {
    std::lock_guard<std::mutex> lock(mu);
    for (int i = 0; i < used; ++i)
      if (cache[i].H == H && cache[i].per_z == per_z && cache[i].win == win && cache[i].sms == sms) return cache[i].m;
  }
  auto cost = [&](int rows) { return (double)((rows + win - 1 + win - 1) / win) + 1.5; };  // groups + 1 periods + fill
  double best = 1e300;
  int best_m = 2;
  std::vector<double> sm((size_t)sms);
  for (int m = 2; m <= 48; ++m) {
    const int rows = m * win - (win - 1);
    const int ny = (H + rows - 1) / rows;
    const int last = H - (ny - 1) * rows;
    // full chunks: uniform jobs in launch order -> SM i gets q or q + 1 of them
    const long long F = (long long)per_z * (ny - 1);
    const double cf = cost(rows), cp = cost(last);
    const long long q = F / sms, r = F % sms;
    for (int i = 0; i < sms; ++i) sm[(size_t)i] = cf * (double)(q + (i < r ? 1 : 0));
    // the last chunk's CTAs go to whichever SM is free first
    std::make_heap(sm.begin(), sm.end(), std::greater<double>());
    for (int j = 0; j < per_z; ++j) {
      std::pop_heap(sm.begin(), sm.end(), std::greater<double>());
      sm.back() += cp;
      std::push_heap(sm.begin(), sm.end(), std::greater<double>());
    }
    const double span = *std::max_element(sm.begin(), sm.end());
    if (span <= best * (1.0 + 1e-9)) { best = span; best_m = m; }
    if (ny == 1) break;  // larger chunks change nothing
  }
  std::lock_guard<std::mutex> lock(mu);
  cache[next] = Entry{H, per_z, win, sms, best_m};
  next = (next + 1) % 8;
  if (used < 8) ++used;
  return best_m;
}

template <int K, int RUNS, int NH, bool WTA>
static int run_ssd_rows(const float *left, const float *right, float *cost, unsigned long long *keys, int H, int W,
                        int D, cudaStream_t stream) {
  using C = ssdr::Cfg<K, RUNS, NH>;
  static PerDeviceInt configured;  // the > 48 KB opt-in is per function AND per device
  const int dev = device_slot();
  int done = 0;
  if (!configured.get(dev, &done)) {
    if (cudaFuncSetAttribute(ssd_rows_kernel<K, RUNS, NH, WTA>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)C::SMEM) != cudaSuccess) {
      cudaGetLastError();
      return 1;  // the caller runs ssd_tiled_kernel
    }
    configured.set(dev, 1);
  }
  const int nx = (W + C::TX - 1) / C::TX;
  const int nd = (D + C::DC - 1) / C::DC;
  // Row chunks of m*WIN - (WIN-1) output rows: such a chunk reads exactly m groups of WIN input
  // rows (1080p D=128, K=7: 0.423 ms with 5 groups per chunk, 0.360 with 10, 0.350 with 20, measured).
  int m = ssd_pick_groups(H, nx * nd, C::WIN, num_sms());
  {
    static const char *force = getenv("STM_DEBUG_SSD_GROUPS");  // developer knob: groups per chunk; 0: the old fixed 20
    if (force && atoi(force) > 0) m = atoi(force);
    else if (force && force[0] == '0') {
      m = 20;
      while (m > 2 && (long long)nx * nd * ((H + (m * C::WIN - C::WIN + 1) - 1) / (m * C::WIN - C::WIN + 1)) < num_sms()) --m;
    }
  }
  const int rows = m * C::WIN - (C::WIN - 1);
  const int ny = (H + rows - 1) / rows;
  if (nd > 65535 || ny > 65535) return fail(STM_EUNSUPPORTED, "ssd: shape too large");
  static const char *order = getenv("STM_DEBUG_SSD_ORDER");  // developer knob: "x" = x strips fastest
  const int d_fastest = !(order && order[0] == 'x');
  if (nx > 65535) return fail(STM_EUNSUPPORTED, "ssd: shape too large");
  const dim3 grid = d_fastest ? dim3(nd, nx, ny) : dim3(nx, nd, ny);
  ssd_rows_kernel<K, RUNS, NH, WTA><<<grid, C::THREADS, C::SMEM, stream>>>(left, right, cost, keys, H, W, D, rows, d_fastest);
  return check_launch("ssd_rows_kernel");
}

// Window half-sizes 1..10 (windows of 2x2 .. 20x20; the reference's default is 7; 9 and 10 spill ~45 registers).  <K, RUNS, NH>: a tile
// of RUNS * 2K columns, NH horizontal warps.  K >= 5: one streaming row-pair item per H warp (NH = K);
// smaller windows have fewer row pairs per group than warps worth running, so their H warps take
// (row pair, run) items.  Tile sizes keep the double-buffered hand-off tile within shared memory and
// the vertical ring (8K registers) within the register cap of the CTA.
template <bool WTA>
static int dispatch_ssd_rows(const float *left, const float *right, float *cost, unsigned long long *keys, int H,
                             int W, int D, int k, cudaStream_t stream) {
  switch (k) {
    case 1: return run_ssd_rows<1, 24, 8, WTA>(left, right, cost, keys, H, W, D, stream);
    case 2: return run_ssd_rows<2, 12, 8, WTA>(left, right, cost, keys, H, W, D, stream);
    case 3: return run_ssd_rows<3, 8, 8, WTA>(left, right, cost, keys, H, W, D, stream);
    case 4: return run_ssd_rows<4, 6, 8, WTA>(left, right, cost, keys, H, W, D, stream);
    case 5: return run_ssd_rows<5, 4, 5, WTA>(left, right, cost, keys, H, W, D, stream);
    case 6: return run_ssd_rows<6, 4, 6, WTA>(left, right, cost, keys, H, W, D, stream);
    case 7: return run_ssd_rows<7, 3, STM_SSD_NH7, WTA>(left, right, cost, keys, H, W, D, stream);
    case 8: return run_ssd_rows<8, 2, 8, WTA>(left, right, cost, keys, H, W, D, stream);
    case 9: return run_ssd_rows<9, 2, 9, WTA>(left, right, cost, keys, H, W, D, stream);
    case 10: return run_ssd_rows<10, 1, 10, WTA>(left, right, cost, keys, H, W, D, stream);
    default: return 1;
  }
}

// Returns 0 on success, 1 when the shape is not handled here (the caller runs ssd_tiled_kernel).
int launch_ssd_rows(const float *left, const float *right, float *cost, int H, int W, int D, int k,
                    cudaStream_t stream) {
  if (D % 4 != 0 || (reinterpret_cast<uintptr_t>(cost) & 15)) return 1;  // 16-byte stores
  {
    static const char *off = getenv("STM_DEBUG_SSD_ROWS");  // developer knob: "0" disables
    if (off && off[0] == '0') return 1;
  }
  return dispatch_ssd_rows<false>(left, right, cost, nullptr, H, W, D, k, stream);
}

__global__ void ssd_wta_extract_kernel(const unsigned long long *__restrict__ keys, int32_t *__restrict__ disp, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) disp[i] = (int32_t)(unsigned)(keys[i] & 0xffffffffull);
}

// SSD cost + winner-takes-all without the volume (stm_ssd_wta_f32).  `keys` = H*W 64-bit words of scratch.
// Returns 0 on success, 1 when the shape is not handled (k not in 1..10, D % 4 != 0).
int launch_ssd_wta(const float *left, const float *right, int32_t *disp, unsigned long long *keys, int H, int W,
                   int D, int k, cudaStream_t stream) {
  if (D % 4 != 0 || k < 1 || k > 10) return 1;
  const size_t n = (size_t)H * W;
  STM_CUDA_TRY(cudaMemsetAsync(keys, 0xff, n * sizeof(unsigned long long), stream));
  const int rc = dispatch_ssd_rows<true>(left, right, nullptr, keys, H, W, D, k, stream);
  if (rc) return rc;
  ssd_wta_extract_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(keys, disp, n);
  return check_launch("ssd_wta_extract_kernel");
}

}  // namespace stm
