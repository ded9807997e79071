// cost_generic.cu -- straightforward cost-volume builders (any kernel_size, any D).
//
// SSD:        reference SSDKernel::operator(), src/ssd.cu:31-61
// Birchfield: reference BirchfieldKernelCUDA::operator(), src/birchfield_cost.cu:83-139
//
// These are the fall-back kernels (arbitrary kernel_size / shapes) and the in-repo
// cross-check for the tiled kernels in ssd_tiled.cu.  One thread produces one (y, x, d)
// with the lanes of a warp along d, so the [H,W,D] stores are coalesced and the right
// image reads are a contiguous reversed segment; summation order is the reference's
// (row-major window, left-to-right taps), hence results are bit-identical to the
// reference for every input, not only integer-valued ones.
#include "stm_common.cuh"

namespace stm {

__global__ void __launch_bounds__(256)
ssd_generic_kernel(const float *__restrict__ left, const float *__restrict__ right,
                   float *__restrict__ cost, int H, int W, int D, int k) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  const int x = blockIdx.y;
  const int y = blockIdx.z;
  if (d >= D) return;
  float *out = cost + ((size_t)y * W + x) * D + d;
  if (d > x) {
    *out = CUDART_INF_F;
    return;
  }
  const int r0 = max(y - k, 0), r1 = min(y + k, H);
  const int c0 = max(x - k, d), c1 = min(x + k, W);
  float acc = 0.f;
  for (int r = r0; r < r1; ++r) {
    const float *lrow = left + (size_t)r * W;
    const float *rrow = right + (size_t)r * W - d;
    for (int c = c0; c < c1; ++c) {
      const float diff = __fsub_rn(__ldg(lrow + c), __ldg(rrow + c));
      acc = __fadd_rn(acc, __fmul_rn(diff, diff));  // no FMA contraction: match the CPU op
    }
  }
  *out = acc;
}

__device__ __forceinline__ float bt_pad(const float *__restrict__ row, int i, int W) {
  // zero-padded scanline: index i in [-1, W]
  return (i >= 0 && i < W) ? __ldg(row + i) : 0.f;
}

__global__ void __launch_bounds__(256)
birchfield_generic_kernel(const float *__restrict__ left, const float *__restrict__ right,
                          float *__restrict__ cost, int H, int W, int D) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  const int x = blockIdx.y;
  const int y = blockIdx.z;
  if (d >= D) return;
  float *out = cost + ((size_t)y * W + x) * D + d;
  if (d > x) {
    *out = CUDART_INF_F;
    return;
  }
  const float *lrow = left + (size_t)y * W;
  const float *rrow = right + (size_t)y * W;
  const int start = max(0, x - d - 4) + d;
  const int end = min(W, x + 4);
  float acc = 0.f;
  for (int p = start; p < end; ++p) {
    const int q = p - d;
    const float li = bt_pad(lrow, p, W), ri = bt_pad(rrow, q, W);
    const float la = 0.5f * (li + bt_pad(lrow, p - 1, W));
    const float lb = 0.5f * (li + bt_pad(lrow, p + 1, W));
    const float ra = 0.5f * (ri + bt_pad(rrow, q - 1, W));
    const float rb = 0.5f * (ri + bt_pad(rrow, q + 1, W));
    const float lmi = fminf(la, fminf(lb, li)), lma = fmaxf(la, fmaxf(lb, li));
    const float rmi = fminf(ra, fminf(rb, ri)), rma = fmaxf(ra, fmaxf(rb, ri));
    const float a = fmaxf(0.f, fmaxf(li - rma, rmi - li));
    const float b = fmaxf(0.f, fmaxf(ri - lma, lmi - ri));
    acc = __fadd_rn(acc, fminf(a, b));
  }
  *out = acc;
}

static dim3 cost_grid(int H, int W, int D, int threads) {
  return dim3((D + threads - 1) / threads, W, H);
}

int launch_ssd_generic(const float *left, const float *right, float *cost, int H, int W, int D,
                       int k, cudaStream_t stream) {
  if (W > 65535 || H > 65535) return fail(STM_EUNSUPPORTED, "ssd: H, W must be <= 65535");
  const int threads = D >= 256 ? 256 : ((D + 31) / 32) * 32;
  ssd_generic_kernel<<<cost_grid(H, W, D, threads), threads, 0, stream>>>(left, right, cost, H, W,
                                                                         D, k);
  return check_launch("ssd_generic_kernel");
}

int launch_birchfield_generic(const float *left, const float *right, float *cost, int H, int W,
                              int D, cudaStream_t stream) {
  if (W > 65535 || H > 65535) return fail(STM_EUNSUPPORTED, "birchfield: H, W must be <= 65535");
  const int threads = D >= 256 ? 256 : ((D + 31) / 32) * 32;
  birchfield_generic_kernel<<<cost_grid(H, W, D, threads), threads, 0, stream>>>(left, right, cost,
                                                                                H, W, D);
  return check_launch("birchfield_generic_kernel");
}

}  // namespace stm
