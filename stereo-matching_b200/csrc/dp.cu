// dp.cu -- scanline dynamic-programming disparity reducer.
//
// Replaces DisparityReduceOps::RunDynamicProgramming (reference
// src/dynamic_programming.cu:184-225): forward min-plus pass with 3-neighbour transitions
// and int8 back-pointers, first-min of the final row costs (torch::argmin, :207), and the
// per-row backtrack (ReducePathsKernel, :166-181).
//
// Default mode = the reference CUDA functor (ExpandPathsKernelCUDA, :83-130) with its
// shared-memory race resolved as read-before-write: all three neighbours come from the
// previous column.  One warp per image row, the D running costs spread over the lanes
// (lanevec.cuh), neighbours by shuffle, back-pointers stored as packed bytes, the next 8
// cost vectors prefetched in registers; forward pass, argmin and backtrack are fused in one
// kernel; the backtrack fetches a +-15 disparity window for 16 columns at a time.
//
// STM_DP_CPU_EXACT = the reference CPU functor (ExpandPathsKernelCPU, :24-68) literally,
// including its transposed column-0 read, its in-place (Gauss-Seidel) scan over d and the
// path[d][row][0] = 0 clobber; serial in d by construction, kept for literal parity only.
#include "lanevec.cuh"
#include "sgm_paths.cuh"

namespace stm {

template <int NV, int VW, bool DEEP>
__global__ void __launch_bounds__(128)
dp_row_kernel(const float *__restrict__ cost, int8_t *__restrict__ path,
              float *__restrict__ row_final, int32_t *__restrict__ disp, int H, int W, int D) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= H) return;
  const size_t row_off = (size_t)row * W;
  const LaneCtx cx = lv_ctx<NV, VW>(lane, D);
  const float *crow = cost + row_off * D;

  LaneVec<NV, VW> acc;
  lv_load<NV, VW, true>(acc, crow, lane, D);
  // column 0 of the path volume: zero (never read by the backtrack)
  {
    int8_t *p0 = path + row_off * D;
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int j = 0; j < VW; ++j) {
        const int d = lv_d<NV, VW>(lane, k, j);
        if (d < D) p0[d] = 0;
      }
  }

  // ---- forward pass.  The cost vectors of the next PF columns stay in flight in registers;
  // every load is unconditional (clamped to the row): a load under a branch is waited for at
  // the join and the pass then runs at one DRAM round trip per column (it did: 1.75 ms).
  // Only H warps exist (7 per SM at 1080p), so the bytes each warp keeps in flight decide the
  // bandwidth: 8 columns (4 KB at D=128) gave 0.51 ms, 24 columns 0.455 ms.  The deep variant
  // needs ~240 registers (8 warps per SM); the host uses it only while all H rows still fit on
  // the GPU at once, otherwise the 8-column one (20 warps per SM).
  constexpr int PF = (NV * VW <= 4) ? (DEEP ? 24 : 8) : 4;
  LaneVec<NV, VW> cbuf[PF];
#pragma unroll
  for (int u = 0; u < PF; ++u)
    lv_load<NV, VW, true>(cbuf[u], crow + (size_t)(1 + u < W ? 1 + u : W - 1) * D, lane, D);

  auto step = [&](const int col, const LaneVec<NV, VW> &cur) {
    LaneVec<NV, VW> lo, hi;
    lv_neighbours<NV, VW>(acc, lo, hi, cx);
    int8_t *pp = path + (row_off + col) * D;
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
      if constexpr (VW == 4) {
        *reinterpret_cast<char4 *>(pp + d0) = make_char4(dirs[0], dirs[1], dirs[2], dirs[3]);
      } else if constexpr (VW == 2) {
        *reinterpret_cast<char2 *>(pp + d0) = make_char2(dirs[0], dirs[1]);
      } else {
        if (d0 < D) pp[d0] = dirs[0];
      }
    }
  };

  int col = 1;
  for (int g = (W - 1) / PF; g > 0; --g) {
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const LaneVec<NV, VW> cur = cbuf[u];
      lv_load<NV, VW, true>(cbuf[u], crow + (size_t)(col + PF < W ? col + PF : W - 1) * D, lane, D);
      step(col, cur);
      ++col;
    }
  }
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    if (col < W) {  // tail: already in cbuf, no loads here
      step(col, cbuf[u]);
      ++col;
    }
  }

  lv_store<NV, VW, false>(acc, row_final + (size_t)row * D, lane, D);

  // first minimum of the final costs (torch::argmin semantics)
  float best = CUDART_INF_F;
  int idx = 0x7fffffff;
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int j = 0; j < VW; ++j) {
      const int d = lv_d<NV, VW>(lane, k, j);
      const float v = acc.v[k][j];
      if (d < D && (v < best || (v == best && d < idx))) { best = v; idx = d; }
    }
#pragma unroll
  for (int m = 1; m < 32; m <<= 1) {
    const float ov = __shfl_xor_sync(FULL, best, m);
    const int oi = __shfl_xor_sync(FULL, idx, m);
    if (ov < best || (ov == best && oi < idx)) { best = ov; idx = oi; }
  }
  int e = idx == 0x7fffffff ? 0 : idx;

  // ---- backtrack (reference ReducePathsKernel, src/dynamic_programming.cu:166-181):
  //   e(col) = clamp(e(col+1) + path[row][col][e(col+1)], 0, D-1).
  // Literally that is W dependent one-byte loads.  The disparity moves by at most one per
  // column, so the bytes the next 16 columns can ask for lie within e +- 15: lane l fetches
  // path[col][e - 16 + l] for 16 columns at once (16 independent loads in flight), then the 16
  // steps are resolved with shuffles.  The row's path bytes were written by this warp.
  __syncwarp();
  int32_t *drow = disp + row_off;
  const signed char *prow = reinterpret_cast<const signed char *>(path + row_off * D);
  if (lane == 0) drow[W - 1] = e;
  int col_hi = W - 1;  // column whose disparity `e` is known
  while (col_hi > 0) {
    const int nb = col_hi < 16 ? col_hi : 16;  // columns col_hi-1 .. col_hi-nb
    int dl = e - 16 + lane;
    dl = dl < 0 ? 0 : (dl > D - 1 ? D - 1 : dl);
    int w[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int c = col_hi - 1 - j;
      w[j] = __ldcg(prow + (size_t)(c < 0 ? 0 : c) * D + dl);
    }
    const int e0 = e;
    int mine = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int v = __shfl_sync(FULL, w[j], e - e0 + 16);
      int en = e + v;
      en = en < 0 ? 0 : (en > D - 1 ? D - 1 : en);
      if (j < nb) e = en;
      if (lane == j) mine = e;
    }
    if (lane < nb) drow[col_hi - 1 - lane] = mine;
    col_hi -= nb;
  }
}

// ---- literal CPU-functor variant -----------------------------------------------------
__global__ void dp_cpu_exact_forward_kernel(const float *__restrict__ cost, int8_t *__restrict__ path,
                                            float *row_final, int H, int W, int D) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= H) return;
  float *prev = row_final + (size_t)row * D;  // running costs live in row_final[row]
  for (int d = 0; d < D; ++d) prev[d] = cost[((size_t)d * W + row) * D];  // cost[d][row][0] (sic)
  for (int col = 1; col < W; ++col) {
    const float *c = cost + ((size_t)row * W + col) * D;
    int8_t *p = path + ((size_t)row * W + col) * D;
    float left = CUDART_INF_F;  // prev[d-1], already updated for this column
    for (int d = 0; d < D; ++d) {
      const float c1 = left, c2 = prev[d], c3 = d < D - 1 ? prev[d + 1] : CUDART_INF_F;
      float m;
      int8_t dir;
      if (c1 < c2 && c1 < c3) { m = c1; dir = -1; }
      else if (c2 < c3) { m = c2; dir = 0; }
      else { m = c3; dir = 1; }
      p[d] = dir;
      left = __fadd_rn(c[d], m);
      prev[d] = left;
    }
  }
}

// Net effect of the reference's `path[disp][row][0] = 0` executed in increasing row
// order (dynamic_programming.cu:38): path[y][x][0] = 0 for y < D, y < x < H, and column 0.
__global__ void dp_cpu_exact_clobber_kernel(int8_t *__restrict__ path, int H, int W, int D) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;  // = reference `row`
  const int y = blockIdx.y;                             // = reference `disp`
  if (x >= H || y >= D) return;
  if (x == 0 || y < x) path[((size_t)y * W + x) * D] = 0;
}

__global__ void dp_backtrack_kernel(const int8_t *__restrict__ path, const float *__restrict__ row_final,
                                    int32_t *__restrict__ disp, int H, int W, int D) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= H) return;
  const float *f = row_final + (size_t)row * D;
  int e = 0;
  for (int d = 1; d < D; ++d)
    if (f[d] < f[e]) e = d;
  int32_t *drow = disp + (size_t)row * W;
  drow[W - 1] = e;
  const int8_t *prow = path + (size_t)row * W * D;
  for (int col = W - 2; col >= 0; --col) {
    e += prow[(size_t)col * D + e];
    e = e < 0 ? 0 : (e > D - 1 ? D - 1 : e);
    drow[col] = e;
  }
}

template <int NV, int VW>
static int run_dp(const float *cost, int8_t *path, float *row_final, int32_t *disp, int H, int W,
                  int D, cudaStream_t stream) {
  const int warps = 4;
  const int blocks = (H + warps - 1) / warps;
  bool deep = false;
  if (NV * VW <= 4) {
    static PerDeviceInt cache;  // per instantiation and device
    const int dev = device_slot();
    int resident = 0;
    if (!cache.get(dev, &resident)) {
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, dp_row_kernel<NV, VW, true>, warps * 32,
                                                        0) != cudaSuccess) {
        cudaGetLastError();
        nb = 0;
      }
      resident = nb * warps * num_sms();
      cache.set(dev, resident);
    }
    deep = H <= resident;
  }
  if (deep)
    dp_row_kernel<NV, VW, true><<<blocks, warps * 32, 0, stream>>>(cost, path, row_final, disp, H, W, D);
  else
    dp_row_kernel<NV, VW, false><<<blocks, warps * 32, 0, stream>>>(cost, path, row_final, disp, H, W, D);
  return check_launch("dp_row_kernel");
}

int launch_dp_bulk(const float *, int8_t *, float *, int32_t *, int, int, int, cudaStream_t);  // dp_bulk.cu

int launch_dp(const float *cost, int8_t *path, float *row_final, int32_t *disp, int H, int W,
              int D, unsigned flags, cudaStream_t stream) {
  if (flags & STM_DP_CPU_EXACT) {
    if (D > H || H > W)
      return fail(STM_EINVAL, "dp: STM_DP_CPU_EXACT requires D <= H <= W (the reference CPU "
                  "functor reads cost[d][row][0]); got H=%d W=%d D=%d", H, W, D);
    dp_cpu_exact_forward_kernel<<<(H + 63) / 64, 64, 0, stream>>>(cost, path, row_final, H, W, D);
    int rc = check_launch("dp_cpu_exact_forward_kernel");
    if (rc) return rc;
    dp_cpu_exact_clobber_kernel<<<dim3((H + 127) / 128, D), 128, 0, stream>>>(path, H, W, D);
    rc = check_launch("dp_cpu_exact_clobber_kernel");
    if (rc) return rc;
    dp_backtrack_kernel<<<(H + 63) / 64, 64, 0, stream>>>(path, row_final, disp, H, W, D);
    return check_launch("dp_backtrack_kernel");
  }
  {  // D % 32 == 0: cost tiles through the bulk-copy engine, prefetched backtrack (dp_bulk.cu)
    const int rc = launch_dp_bulk(cost, path, row_final, disp, H, W, D, stream);
    if (rc != 1) return rc;
  }
  const LaneMap m = pick_lane_map(D);
#define STM_DP_CASE(NV_, VW_) \
  if (m.NV == NV_ && m.VW == VW_) return run_dp<NV_, VW_>(cost, path, row_final, disp, H, W, D, stream);
  STM_DP_CASE(1, 4) STM_DP_CASE(2, 4) STM_DP_CASE(3, 4) STM_DP_CASE(4, 4)
  STM_DP_CASE(1, 2) STM_DP_CASE(3, 2) STM_DP_CASE(5, 2) STM_DP_CASE(7, 2)
  STM_DP_CASE(1, 1) STM_DP_CASE(2, 1) STM_DP_CASE(3, 1) STM_DP_CASE(4, 1)
  STM_DP_CASE(5, 1) STM_DP_CASE(6, 1) STM_DP_CASE(7, 1) STM_DP_CASE(8, 1)
#undef STM_DP_CASE
  return fail(STM_EUNSUPPORTED, "dp: D=%d not supported", D);
}

}  // namespace stm
