// dp_bulk.cu -- scanline dynamic programming, default ("cuda", Jacobi) mode, for D % 64 == 0.
//
// Same recurrence, back-pointers, final costs and disparity map as dp_row_kernel (dp.cu; reference
// src/dynamic_programming.cu:83-130 forward, :166-181 backtrack): one warp per image row, the D
// running costs spread over the lanes, neighbours by shuffle.  What differs is how bytes move; ncu
// of dp_row_kernel at 1080p D=128 (profiles/r02_v2_ncu_full_summary.txt, 433 us) showed
//   * forward pass (59 % of the warp samples): 30 % of them on the first use of a prefetched
//     register.  24 column vectors were in flight per warp, but a warp has six scoreboards, so a
//     wait for the load issued 24 columns ago also waits for the ones issued 6, 12, 18 columns ago:
//     the effective lookahead was ~6 columns, about one DRAM round trip;
//   * backtrack (40 %): e(col) = clamp(e(col+1) + path[col][e(col+1)]) was resolved 16 columns per
//     DRAM round trip (a +-16 window around e), 120 dependent round trips per row.
// Here
//   * a cost row is contiguous in [H,W,D], so lane 0 streams it in tiles of a few columns with
//     cp.async.bulk (global -> shared, completion on the stage's mbarrier), four tiles ahead; the
//     warp reads the vectors back from shared memory.  No prefetch registers, no scoreboards;
//   * the backtrack streams the row's path bytes back through the same stages (tiles of up to 32
//     columns, whole vectors: those copies do not depend on e).  e is warp-uniform, so every lane
//     reads the byte tile[col][e] itself: no shuffle, two instructions on the chain per column.
// Each warp owns its stages and barriers: no CTA-wide barrier, a warp past the last row exits.
#include <cstdlib>
#include <type_traits>

#include "lanevec.cuh"
#include "sgm_paths.cuh"

namespace stm {
namespace dpb {

constexpr int kWarps = 4;
constexpr int kStages = 4;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (int spin = 0; spin < (1 << 24); ++spin) {  // try_wait sleeps in hardware; bounded anyway
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return;
  }
  __trap();  // a bulk copy never completed: fail loudly instead of hanging the GPU
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

template <int NV, int VW>
struct Cfg {
  static constexpr int kD = NV * 32 * VW;
  static constexpr int kTile = kD <= 128 ? 8 : (kD <= 256 ? 4 : 2);  // columns per tile (<= 4 KB)
  static constexpr int kTileFloats = kTile * kD;
  static constexpr size_t kWarpBytes = (size_t)kStages * kTileFloats * sizeof(float);
  static constexpr size_t kSmemBytes = kWarps * kWarpBytes + kWarps * kStages * sizeof(uint64_t);
};

}  // namespace dpb

template <int NV, int VW>
__global__ void __launch_bounds__(dpb::kWarps * 32)
dp_row_bulk_kernel(const float *__restrict__ cost, int8_t *__restrict__ path, float *__restrict__ row_final,
                   int32_t *__restrict__ disp, int H, int W) {
  using namespace dpb;
  using C = Cfg<NV, VW>;
  constexpr int D = C::kD, TILE = C::kTile, TF = C::kTileFloats;
  static_assert(VW == 4 || VW == 2 || VW == 1, "lane maps of sgm_paths.cuh");  // VW == 1: D = 32 NV exactly (host)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
  const int row = blockIdx.x * kWarps + wi;
  if (row >= H) return;  // no CTA-wide barrier anywhere below

  float *stages = reinterpret_cast<float *>(smem_raw + (size_t)wi * C::kWarpBytes);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + kWarps * C::kWarpBytes) + wi * kStages;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(smem_u32(bars + s), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();

  const size_t row_off = (size_t)row * W;
  const float *crow = cost + row_off * D;
  const int ntiles = (W + TILE - 1) / TILE;
  auto issue = [&](int t) {  // lane 0: arm the stage's barrier and request tile t
    const int s = t % kStages;
    const int np = min(TILE, W - t * TILE);
    const uint32_t bytes = (uint32_t)np * D * sizeof(float);
    const uint32_t bar = smem_u32(bars + s);
    mbar_expect_tx(bar, bytes);
    bulk_g2s(smem_u32(stages + s * TF), crow + (size_t)t * TILE * D, bytes, bar);
  };
  if (lane == 0)
    for (int t = 0; t < kStages && t < ntiles; ++t) issue(t);

  const LaneCtx cx = lv_ctx<NV, VW>(lane, D);
  LaneVec<NV, VW> acc;
  int8_t *prow_w = path + row_off * D;

  // one column: three-way decision of the reference (c1 < c2 && c1 < c3 / c2 < c3 / else), all
  // three neighbours from the previous column, back-pointers packed VW bytes per lane
  auto step = [&](const int col, const LaneVec<NV, VW> &cur) {
    LaneVec<NV, VW> lo, hi;
    lv_neighbours<NV, VW>(acc, lo, hi, cx);
    int8_t *pp = prow_w + (size_t)col * D;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      int8_t dirs[VW];
#pragma unroll
      for (int j = 0; j < VW; ++j) {
        const float c1 = lo.v[k][j], c2 = acc.v[k][j], c3 = hi.v[k][j];
        float m;
        int8_t dir;
        if (c1 < c2 && c1 < c3) { m = c1; dir = -1; }
        else if (c2 < c3) { m = c2; dir = 0; }
        else { m = c3; dir = 1; }
        dirs[j] = dir;
        acc.v[k][j] = __fadd_rn(cur.v[k][j], m);
      }
      const int d0 = lv_d<NV, VW>(lane, k, 0);
      if constexpr (VW == 4) *reinterpret_cast<char4 *>(pp + d0) = make_char4(dirs[0], dirs[1], dirs[2], dirs[3]);
      else if constexpr (VW == 2) *reinterpret_cast<char2 *>(pp + d0) = make_char2(dirs[0], dirs[1]);
      else pp[d0] = dirs[0];
    }
  };

  // ---- forward pass over the tiles
  for (int t = 0; t < ntiles; ++t) {
    const int s = t % kStages;
    mbar_wait(smem_u32(bars + s), (uint32_t)(t / kStages) & 1u);
    const float *tb = stages + s * TF;
    const int np = min(TILE, W - t * TILE);
    auto column = [&](int u) {
      LaneVec<NV, VW> c;
      lv_load<NV, VW, false>(c, tb + u * D, lane, D);
      const int col = t * TILE + u;
      if (col == 0) {
        acc = c;  // column 0: the running costs start as the cost vector; its path bytes are 0
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int d0 = lv_d<NV, VW>(lane, k, 0);
          if constexpr (VW == 4) *reinterpret_cast<char4 *>(prow_w + d0) = make_char4(0, 0, 0, 0);
          else if constexpr (VW == 2) *reinterpret_cast<char2 *>(prow_w + d0) = make_char2(0, 0);
          else prow_w[d0] = 0;
        }
      } else {
        step(col, c);
      }
    };
    if (np == TILE) {
#pragma unroll
      for (int u = 0; u < TILE; ++u) column(u);
    } else {
      for (int u = 0; u < np; ++u) column(u);
    }
    __syncwarp();  // every lane has read the tile: its stage can be refilled
    if (lane == 0 && t + kStages < ntiles) issue(t + kStages);
  }

  lv_store<NV, VW, false>(acc, row_final + (size_t)row * D, lane, D);

  // first minimum of the final costs (torch::argmin semantics, reference :207)
  float best = CUDART_INF_F;
  int idx = 0x7fffffff;
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j) {
      const int d = lv_d<NV, VW>(lane, k, j);
      const float v = acc.v[k][j];
      if (v < best || (v == best && d < idx)) { best = v; idx = d; }
    }
#pragma unroll
  for (int m = 1; m < 32; m <<= 1) {
    const float ov = __shfl_xor_sync(FULL, best, m);
    const int oi = __shfl_xor_sync(FULL, idx, m);
    if (ov < best || (ov == best && oi < idx)) { best = ov; idx = oi; }
  }
  int e = idx == 0x7fffffff ? 0 : idx;

  // ---- backtrack: e(col) = clamp(e(col+1) + path[row][col][e(col+1)], 0, D-1), col = W-2 .. 0.
  // The row's path bytes are contiguous as well, so they come back the way the costs came in: tiles
  // of BTC columns (BTC * D bytes, at most one stage) by cp.async.bulk into the SAME stages, kStages
  // tiles ahead.  With a column's whole vector in shared memory no shuffle is needed: e is
  // warp-uniform, every lane reads the one byte tile[col][e] (a broadcast) and keeps the e of the
  // column whose index equals its lane; a tile ends with one coalesced store of BTC disparities.
  // The chain per column is LDS.S8 + add-and-clamp.
  // The bytes were written by this warp's generic-proxy stores and are read by the async proxy:
  // __syncwarp orders the lanes' stores, fence.proxy.async orders them before the bulk copies.
  __syncwarp();
  asm volatile("fence.proxy.async;" ::: "memory");
  constexpr int BTC = (int)((C::kTileFloats * sizeof(float)) / D) < 32 ? (int)((C::kTileFloats * sizeof(float)) / D) : 32;
  static_assert(BTC >= 1 && BTC <= 32, "backtrack tile");
  int32_t *drow = disp + row_off;
  const signed char *prow_r = reinterpret_cast<const signed char *>(prow_w);
  if (lane == 0) drow[W - 1] = e;
  const int nbt = (W - 1 + BTC - 1) / BTC;  // tile b: columns hi = W-2-b*BTC down to lo = max(hi-BTC+1, 0)
  auto bt_issue = [&](int b) {              // lane 0; sequence number ntiles + b continues the stage/phase order
    const int g = ntiles + b, st = g % kStages;
    const int hi = W - 2 - b * BTC, lo = hi - BTC + 1 < 0 ? 0 : hi - BTC + 1;
    const uint32_t bytes = (uint32_t)(hi - lo + 1) * D;
    const uint32_t bar = smem_u32(bars + st);
    mbar_expect_tx(bar, bytes);
    bulk_g2s(smem_u32(stages + st * TF), prow_r + (size_t)lo * D, bytes, bar);
  };
  if (lane == 0)
    for (int b = 0; b < kStages && b < nbt; ++b) bt_issue(b);

  auto bt_resolve = [&](auto full_tag, const signed char *tile, int ncols_tile, int hi) {
    constexpr bool FULL_TILE = decltype(full_tag)::value;  // all BTC columns exist: no test per step
    // column hi - j sits at tile row (ncols_tile - 1 - j)
    const signed char *top = tile + (size_t)(ncols_tile - 1) * D;
    int mine = 0;
#pragma unroll
    for (int j = 0; j < BTC; ++j) {
      if (FULL_TILE || j < ncols_tile) {
        const int v = top[e - j * D];
        int en = e + v;
        en = en < 0 ? 0 : (en > D - 1 ? D - 1 : en);
        e = en;
      }
      if (lane == j) mine = e;
    }
    if (lane < ncols_tile) drow[hi - lane] = mine;
  };

  for (int b = 0; b < nbt; ++b) {
    const int g = ntiles + b, st = g % kStages;
    mbar_wait(smem_u32(bars + st), (uint32_t)(g / kStages) & 1u);
    const int hi = W - 2 - b * BTC, lo = hi - BTC + 1 < 0 ? 0 : hi - BTC + 1;
    const signed char *tile = reinterpret_cast<const signed char *>(stages + st * TF);
    if (hi - lo + 1 == BTC) bt_resolve(std::true_type{}, tile, BTC, hi);
    else bt_resolve(std::false_type{}, tile, hi - lo + 1, hi);
    __syncwarp();  // every lane has read the tile: its stage can be refilled
    if (lane == 0 && b + kStages < nbt) bt_issue(b + kStages);
  }
}

template <int NV, int VW>
static int run_dp_bulk(const float *cost, int8_t *path, float *row_final, int32_t *disp, int H, int W,
                       cudaStream_t stream) {
  using C = dpb::Cfg<NV, VW>;
  static PerDeviceInt configured;  // the > 48 KB opt-in is per function AND per device
  const int dev = device_slot();
  int done = 0;
  if (!configured.get(dev, &done)) {
    if (cudaFuncSetAttribute(dp_row_bulk_kernel<NV, VW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)C::kSmemBytes) != cudaSuccess) {
      cudaGetLastError();
      return 1;
    }
    configured.set(dev, 1);
  }
  const int blocks = (H + dpb::kWarps - 1) / dpb::kWarps;
  dp_row_bulk_kernel<NV, VW><<<blocks, dpb::kWarps * 32, C::kSmemBytes, stream>>>(cost, path, row_final, disp, H, W);
  return check_launch("dp_row_bulk_kernel");
}

// Returns 0 on success, 1 if the shape is not handled here (the caller runs dp_row_kernel), < 0 on error.
int launch_dp_bulk(const float *cost, int8_t *path, float *row_final, int32_t *disp, int H, int W, int D,
                   cudaStream_t stream) {
  if (D % 32 != 0 || W < 2) return 1;   // whole lane vectors: the tiles are copied as they lie in memory
  if ((reinterpret_cast<uintptr_t>(cost) & 15) || (reinterpret_cast<uintptr_t>(path) & 15) ||
      (reinterpret_cast<uintptr_t>(row_final) & 15))
    return 1;
  {
    static const char *off = getenv("STM_DEBUG_DP_BULK");  // developer knob: "0" disables
    if (off && off[0] == '0') return 1;
  }
  const LaneMap m = pick_lane_map(D);
#define STM_DPB_CASE(NV_, VW_) \
  if (m.NV == NV_ && m.VW == VW_) return run_dp_bulk<NV_, VW_>(cost, path, row_final, disp, H, W, stream);
  STM_DPB_CASE(1, 4) STM_DPB_CASE(2, 4) STM_DPB_CASE(3, 4) STM_DPB_CASE(4, 4)
  STM_DPB_CASE(1, 2) STM_DPB_CASE(3, 2) STM_DPB_CASE(5, 2) STM_DPB_CASE(7, 2)
  STM_DPB_CASE(1, 1) STM_DPB_CASE(3, 1) STM_DPB_CASE(5, 1) STM_DPB_CASE(7, 1)
#undef STM_DPB_CASE
  return 1;
}

}  // namespace stm
