// sgm_row.cu -- the two horizontal SGM directions (-> and <-) with the volumes staged through
// shared memory by 1-D bulk copies (cp.async.bulk + mbarrier; SASS UBLKCP), sm_100a.
//
// Same arithmetic as sgm_path_kernel (sgm.cu) for dir 0 / dir 1 of the reference path set
// (reference src/semiglobal.cpp:145-156, path list :167-197): one warp owns one image row and
// runs the recurrence along it.  What differs is how the bytes move.  A horizontal path is the
// one direction whose pixels are CONTIGUOUS in the [H,W,D] volume, so instead of one 512-byte
// register prefetch per pixel (1080 warps x 3 streams, each 983 KB from the next: the slowest
// access pattern of the eight passes, ~5.4 TB/s against ~6.6 TB/s for the vertical ones) a warp
// moves whole tiles of kTile pixels:
//   lane 0:  cp.async.bulk  global -> shared   C tile and S tile (kTile*D*4 contiguous bytes
//            each), completing on the stage's mbarrier; three stages per warp
//   warp:    waits on the mbarrier, runs kTile steps reading C (and S) from shared memory,
//            overwrites the S tile with S + L in place
//   lane 0:  cp.async.bulk  shared -> global   of the finished tile (bulk group), and re-arms
//            the stage that was stored one tile earlier
// No registers are spent on prefetch slots, no CTA-wide barrier exists (each warp owns its
// stages and barriers), and a warp whose row is outside the image simply exits.
#include <cstdlib>

#include "lanevec.cuh"
#include "sgm_paths.cuh"

namespace stm {

namespace rowk {

constexpr int kStages = 3;  // 4-pixel tiles x 6 stages (same bytes, longer lead) measured slower
constexpr int kWarps = 4;

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (int spin = 0; spin < (1 << 24); ++spin) {  // try_wait sleeps in hardware; bounded anyway
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) return;
  }
  __trap();  // a bulk copy never completed: fail loudly instead of hanging the GPU
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes,
                                         uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

template <int NV, int VW>
struct Cfg {
  static constexpr int kDpad = NV * 32 * VW;
  static constexpr int kTile = kDpad <= 128 ? 8 : (kDpad <= 256 ? 4 : 2);  // pixels per tile
  static constexpr int kTileFloats = kTile * kDpad;                        // <= 1024 floats
  static constexpr size_t kWarpBytes = (size_t)kStages * 2 * kTileFloats * sizeof(float);
  static constexpr size_t kSmemBytes = kWarps * kWarpBytes + kWarps * kStages * sizeof(uint64_t);
};

}  // namespace rowk

// MASKED (NV == 1): the range D_real is a multiple of 4 below the padded 32 * VW; a pixel's vector is
// D_real floats in memory and in the tiles, the lanes beyond it hold +inf and neither load nor store
// (D = 100 or 112 then take this kernel instead of the scalar-lane path kernel).
template <int NV, int VW, bool ACC, bool BACKWARD, bool MASKED = false>
__global__ void __launch_bounds__(rowk::kWarps * 32)
sgm_row_kernel(const float *__restrict__ cost, const float *__restrict__ image,
               float *__restrict__ out, int H, int W, float p1, float p2, int min_limit, int D_real) {
  using namespace rowk;
  using C = Cfg<NV, VW>;
  static_assert(!MASKED || NV == 1, "masked maps have one vector per lane");
  constexpr int TILE = C::kTile, TF = C::kTileFloats;
  const int D = MASKED ? D_real : C::kDpad;  // not masked: the host guarantees D == dpad
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
  const int y = blockIdx.x * kWarps + wi;
  if (y >= H) return;  // no CTA-wide barrier anywhere below

  // this warp's stages: stage s = [C tile | S tile], and its mbarriers
  float *wbase = reinterpret_cast<float *>(smem_raw + (size_t)wi * C::kWarpBytes);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + kWarps * C::kWarpBytes) + wi * kStages;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(smem_u32(bars + s), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_async_smem();
  }
  __syncwarp();

  const size_t row = (size_t)y * W;
  const float *crow = cost + row * D;
  float *orow = out + row * D;
  const float *irow = image + row;
  const int ntiles = (W + TILE - 1) / TILE;

  // tile t = pixels [t*TILE, t*TILE+np) in TRAVERSAL order; in memory it starts at x0(t)
  auto tile_np = [&](int t) { return min(TILE, W - t * TILE); };
  auto tile_x0 = [&](int t) { return BACKWARD ? W - t * TILE - tile_np(t) : t * TILE; };
  auto issue = [&](int t) {  // lane 0: arm the stage's barrier and request the tile(s)
    const int s = t % kStages;
    const uint32_t bytes = (uint32_t)tile_np(t) * D * sizeof(float);
    const size_t off = (size_t)tile_x0(t) * D;
    const uint32_t bar = smem_u32(bars + s);
    mbar_expect_tx(bar, ACC ? 2 * bytes : bytes);
    bulk_g2s(smem_u32(wbase + (s * 2) * TF), crow + off, bytes, bar);
    if (ACC) bulk_g2s(smem_u32(wbase + (s * 2 + 1) * TF), orow + off, bytes, bar);
  };
  // lane u (mod TILE) holds the intensity of traversal pixel t*TILE+u; always a valid address
  auto tile_intensity = [&](int t) {
    const int i = min(t * TILE + (lane & (TILE - 1)), W - 1);
    return __ldg(irow + (BACKWARD ? W - 1 - i : i));
  };

  if (lane == 0) {
    for (int t = 0; t < kStages - 1 && t < ntiles; ++t) issue(t);
  }
  const LaneCtx cx = lv_ctx<NV, VW>(lane, min_limit);
  LaneVec<NV, VW> L;
  float icur = tile_intensity(0);
  float iprev_tile = icur;  // intensities of the previous tile (its last pixel precedes pixel 0)

  for (int t = 0; t < ntiles; ++t) {
    const int s = t % kStages;
    const float inext = tile_intensity(t + 1);
    mbar_wait(smem_u32(bars + s), (uint32_t)(t / kStages) & 1u);
    const float *cb = wbase + (s * 2) * TF;
    float *sb = wbase + (s * 2 + 1) * TF;
    const int np = tile_np(t);
    // The adaptive penalty max(P1, P2/|dI|) (an IEEE division) depends on the intensities only:
    // lane u computes it once for pixel u of the tile, one shuffle per step hands it out.  It
    // was evaluated by every lane at every step -- a dozen instructions of a serial chain that
    // runs at ~5 cycles per instruction (ncu: 69 instructions per step, 35 % issue utilisation).
    float p2_tile;
    {
      const float left_px = __shfl_up_sync(FULL, icur, 1);
      const float carry = __shfl_sync(FULL, iprev_tile, TILE - 1);
      const float before = ((lane & (TILE - 1)) == 0) ? carry : left_px;
      p2_tile = sgm_p2_adjusted(p1, p2, icur, before);
    }

    auto step = [&](int u, int um) {  // u: traversal index in the tile, um: memory index
      LaneVec<NV, VW> c;
      const bool held = !MASKED || lane * VW < D;   // this lane's vector exists
      if (held) {
        lv_load<NV, VW, false>(c, cb + um * D, lane, C::kDpad);
      } else {
#pragma unroll
        for (int j = 0; j < VW; ++j) c.v[0][j] = CUDART_INF_F;
      }
      if (t == 0 && u == 0) {
        L = c;  // first pixel of the path: L = C
      } else {
        lv_sgm_step<NV, VW>(L, c, p1, __shfl_sync(FULL, p2_tile, u), cx);
      }
      if (ACC) {
        if (held) {
          LaneVec<NV, VW> o;
          lv_load<NV, VW, false>(o, sb + um * D, lane, C::kDpad);
#pragma unroll
          for (int k = 0; k < NV; ++k)
#pragma unroll
            for (int j = 0; j < VW; ++j) o.v[k][j] = __fadd_rn(o.v[k][j], L.v[k][j]);
          lv_store<NV, VW, false>(o, sb + um * D, lane, C::kDpad);
        }
      } else if (held) {
        lv_store<NV, VW, false>(L, sb + um * D, lane, C::kDpad);
      }
    };
    if (np == TILE) {
#pragma unroll
      for (int u = 0; u < TILE; ++u) step(u, BACKWARD ? TILE - 1 - u : u);
    } else {
      for (int u = 0; u < np; ++u) step(u, BACKWARD ? np - 1 - u : u);
    }

    fence_async_smem();  // generic-proxy writes of the tile -> visible to the bulk store
    __syncwarp();
    if (lane == 0) {
      bulk_s2g(orow + (size_t)tile_x0(t) * D, smem_u32(sb), (uint32_t)np * D * sizeof(float));
      bulk_commit();
      if (t + kStages - 1 < ntiles) {
        bulk_wait_read<1>();  // the store of tile t-1 has left its stage: re-arm it
        issue(t + kStages - 1);
      }
    }
    iprev_tile = icur;
    icur = inext;
  }
  if (lane == 0) bulk_wait_all();  // shared memory must outlive the last stores
}

template <int NV, int VW, bool MASKED = false>
static int run_sgm_row(const float *cost, const float *image, float *out, int H, int W, int D,
                       float p1, float p2, int dir, int min_limit, bool acc, cudaStream_t stream) {
  using C = rowk::Cfg<NV, VW>;
  if (MASKED ? (D >= C::kDpad || D % 4 != 0) : C::kDpad != D) return 1;  // other ranges stay on the generic kernel
  static PerDeviceInt configured;  // the > 48 KB opt-in is per function AND per device
  const int dev = device_slot();
  int done = 0;
  if (!configured.get(dev, &done)) {
    cudaError_t e = cudaSuccess;
#define STM_ROW_ATTR(A_, B_)                                                                  \
  if (e == cudaSuccess)                                                                      \
    e = cudaFuncSetAttribute(sgm_row_kernel<NV, VW, A_, B_, MASKED>,                         \
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::kSmemBytes);
    STM_ROW_ATTR(false, false) STM_ROW_ATTR(false, true) STM_ROW_ATTR(true, false)
    STM_ROW_ATTR(true, true)
#undef STM_ROW_ATTR
    if (e != cudaSuccess) {
      cudaGetLastError();
      return 1;
    }
    configured.set(dev, 1);
  }
  const int blocks = (H + rowk::kWarps - 1) / rowk::kWarps;
  const dim3 grid(blocks), block(rowk::kWarps * 32);
  const size_t smem = C::kSmemBytes;
  const bool backward = dir == 1;
  if (acc && backward)
    sgm_row_kernel<NV, VW, true, true, MASKED><<<grid, block, smem, stream>>>(cost, image, out, H, W, p1, p2, min_limit, D);
  else if (acc)
    sgm_row_kernel<NV, VW, true, false, MASKED><<<grid, block, smem, stream>>>(cost, image, out, H, W, p1, p2, min_limit, D);
  else if (backward)
    sgm_row_kernel<NV, VW, false, true, MASKED><<<grid, block, smem, stream>>>(cost, image, out, H, W, p1, p2, min_limit, D);
  else
    sgm_row_kernel<NV, VW, false, false, MASKED><<<grid, block, smem, stream>>>(cost, image, out, H, W, p1, p2, min_limit, D);
  return check_launch("sgm_row_kernel");
}

// dir 0 (->) or 1 (<-).  Returns 0 on success, 1 if the shape is not handled here (the caller
// falls back to sgm_path_kernel), < 0 on error.
int launch_sgm_row(const float *cost, const float *image, float *out, int H, int W, int D,
                   float p1, float p2, int dir, int min_limit, bool acc, cudaStream_t stream) {
  if (dir != 0 && dir != 1) return 1;
  if (W < 2) return 1;
  if (D % 32 != 0) {
    // a multiple of 4 between 32 and 128: the 4-wide map of D = 128, last lanes masked
    if (D % 4 != 0 || D < 36 || D > 124) return 1;
    if ((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(out)) & 15) return 1;
    static const char *off = getenv("STM_DEBUG_SGM_ROW");
    if (off && off[0] == '0') return 1;
    return run_sgm_row<1, 4, true>(cost, image, out, H, W, D, p1, p2, dir, min_limit, acc, stream);
  }
  if ((reinterpret_cast<uintptr_t>(cost) | reinterpret_cast<uintptr_t>(out)) & 15) return 1;  // bulk copies
  {
    static const char *off = getenv("STM_DEBUG_SGM_ROW");  // developer knob: "0" disables
    if (off && off[0] == '0') return 1;
  }
  const LaneMap m = pick_lane_map(D);
#define STM_ROW_CASE(NV_, VW_) \
  if (m.NV == NV_ && m.VW == VW_) \
    return run_sgm_row<NV_, VW_>(cost, image, out, H, W, D, p1, p2, dir, min_limit, acc, stream);
  STM_ROW_CASE(1, 4) STM_ROW_CASE(2, 4) STM_ROW_CASE(3, 4) STM_ROW_CASE(4, 4)
  STM_ROW_CASE(1, 2) STM_ROW_CASE(3, 2) STM_ROW_CASE(5, 2) STM_ROW_CASE(7, 2)
  STM_ROW_CASE(1, 1) STM_ROW_CASE(3, 1) STM_ROW_CASE(5, 1) STM_ROW_CASE(7, 1)
#undef STM_ROW_CASE
  return 1;
}

}  // namespace stm
