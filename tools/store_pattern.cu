// store_pattern.cu -- how fast can [H,W,D] fp32 be WRITTEN by kernels shaped like the cost builders?
// (developer microbenchmark: no arithmetic, only the store pattern.)  nvcc -O3 -arch=sm_100a.
//   pattern A(DC): CTA = 42 columns x DC disparities x a chunk of rows; thread = (column, 4 disparities),
//                  walks down the rows; DC = 32 is ssd_rows_kernel / ssd_tiled_kernel, 64 / 128 = wider chunks.
//   pattern B:     CTA = one image row, 8 warps x 240 columns, lane writes d = lane + 32 j (birchfield).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

template <int DC>
__global__ void pat_a(float *c, int H, int W, int D, int rows) {
  constexpr int TPX = DC / 4;                 // threads per column
  const int TX = blockDim.x / TPX;
  const int xl = threadIdx.x / TPX, dq = threadIdx.x % TPX;
  const int x = blockIdx.y * TX + xl, d0 = blockIdx.x * DC + 4 * dq;
  const int ya = blockIdx.z * rows, yb = min(H, ya + rows);
  if (x >= W) return;
  float *o = c + ((size_t)ya * W + x) * D + d0;
  for (int y = ya; y < yb; ++y, o += (size_t)W * D) __stcs(reinterpret_cast<float4 *>(o), make_float4(y, x, d0, 1.f));
}
__global__ void pat_b(float *c, int H, int W, int D) {
  const int y = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunk = W / 8;
  float *o = c + ((size_t)y * W + warp * chunk) * D + lane;
  for (int x = 0; x < chunk; ++x, o += D)
    for (int j = 0; j < D / 32; ++j) __stcs(o + 32 * j, (float)x);
}

int main() {
  const int H = 1080, W = 1920, D = 128;
  float *c;
  cudaMalloc(&c, (size_t)H * W * D * 4);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  auto time = [&](const char *name, auto launch) {
    for (int i = 0; i < 3; ++i) launch();
    float best = 1e9f;
    for (int i = 0; i < 10; ++i) {
      cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); best = ms < best ? ms : best;
    }
    printf("{\"pattern\": \"%s\", \"ms\": %.4f, \"GBs\": %.1f, \"err\": \"%s\"}\n", name, best,
           (double)H * W * D * 4 / best / 1e6, cudaGetErrorString(cudaGetLastError()));
  };
  const int rows = 127, ny = (H + rows - 1) / rows;
  time("A DC=32 (ssd): 42 cols x 32 d per CTA", [&] { pat_a<32><<<dim3(D / 32, (W + 41) / 42, ny), 42 * 8>>>(c, H, W, D, rows); });
  time("A DC=64: 42 cols x 64 d per CTA", [&] { pat_a<64><<<dim3(D / 64, (W + 41) / 42, ny), 42 * 16>>>(c, H, W, D, rows); });
  time("A DC=128: 28 cols x 128 d per CTA", [&] { pat_a<128><<<dim3(1, (W + 27) / 28, ny), 28 * 32>>>(c, H, W, D, rows); });
  time("A DC=32, 1080-row chunks", [&] { pat_a<32><<<dim3(D / 32, (W + 41) / 42, 1), 42 * 8>>>(c, H, W, D, H); });
  time("A DC=128, 1080-row chunks", [&] { pat_a<128><<<dim3(1, (W + 27) / 28, 1), 28 * 32>>>(c, H, W, D, H); });
  time("B (birchfield): row per CTA, warp writes 512 B per column", [&] { pat_b<<<H, 256>>>(c, H, W, D); });
  cudaMemset(c, 0, 16);
  time("cudaMemset", [&] { cudaMemsetAsync(c, 0, (size_t)H * W * D * 4); });
  return 0;
}
