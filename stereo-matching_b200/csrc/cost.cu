// cost.cu -- dispatch of the cost-volume builders.
#include "stm_common.cuh"

namespace stm {

int launch_ssd_generic(const float *, const float *, float *, int, int, int, int, cudaStream_t);
int launch_birchfield_generic(const float *, const float *, float *, int, int, int, cudaStream_t);
int launch_ssd_tiled(const float *, const float *, float *, int, int, int, int, cudaStream_t);
int launch_ssd_rows(const float *, const float *, float *, int, int, int, int, cudaStream_t);
int launch_birchfield_tiled(const float *, const float *, float *, int, int, int, cudaStream_t);

int launch_ssd(const float *left, const float *right, float *cost, int H, int W, int D, int k,
               cudaStream_t stream) {
  // tiled separable kernel for the window sizes it is instantiated for; returns 1 when it
  // does not handle `k` (then the straightforward kernel runs instead)
  // horizontal-first packed kernel (ssd_rows.cu) for k = 7 / 3 and D % 4 == 0, then the older tiled one
  int rc = launch_ssd_rows(left, right, cost, H, W, D, k, stream);
  if (rc != 1) return rc;
  rc = launch_ssd_tiled(left, right, cost, H, W, D, k, stream);
  if (rc != 1) return rc;
  return launch_ssd_generic(left, right, cost, H, W, D, k, stream);
}

int launch_birchfield(const float *left, const float *right, float *cost, int H, int W, int D,
                      cudaStream_t stream) {
  // one CTA per scanline with the BT terms evaluated once (birchfield_tiled.cu); returns 1 for
  // the shapes it does not cover (then the straightforward kernel runs instead)
  const int rc = launch_birchfield_tiled(left, right, cost, H, W, D, stream);
  if (rc != 1) return rc;
  return launch_birchfield_generic(left, right, cost, H, W, D, stream);
}

}  // namespace stm
