// texture.cu -- CUDA texture objects as an INPUT format of the SSD stage.
//
// The reference's native module exports a `CUDATexture` class (reference src/cuda_texture.cpp:23-192,
// include/stereomatch/cuda_texture.hpp: a cudaArray + cudaTextureObject_t built from an [H,W,C]
// tensor, point filtering, clamp addressing, element-type reads) and a
// `compute_ssd(CUDATexture, CUDATexture, cost, k)` overload (src/cost.cpp:10-13) whose kernel does
// 2 (2k)^2 tex2D fetches per output (src/ssd.cu:83-153).  Texture objects are CUDA compute API.
//
// Here a texture is the same object (cudaMallocArray + cudaCreateTextureObject, same descriptor),
// owned by an opaque handle.  The SSD overload does NOT walk the texture per output: one small
// kernel fetches every texel of both textures once (tex2D<float> at integer coordinates -- exactly
// the fetches the reference kernel makes, src/ssd.cu:124-127) into two linear [H,W] rows of a
// caller-provided scratch, and the volume is then produced by the same tiled kernel as the tensor
// overload (ssd_tiled.cu): 2 H W fetches instead of 2 (2k)^2 H W D, identical values.
#include <cstring>

#include "stm_common.cuh"

namespace stm {

int launch_ssd(const float *, const float *, float *, int, int, int, int, cudaStream_t);

}  // namespace stm

struct stm_texture {
  cudaTextureObject_t object = 0;
  cudaArray_t array = nullptr;
  int height = 0, width = 0, channels = 0, dtype = 0, normalized = 0, device = 0;
};

namespace stm {
namespace {

// channel 0 of texel (x, y); the reference's kernels read with unnormalised integer coordinates
template <typename T>
__device__ __forceinline__ T tex_fetch(cudaTextureObject_t t, int x, int y, int channels);
template <>
__device__ __forceinline__ float tex_fetch<float>(cudaTextureObject_t t, int x, int y, int channels) {
  if (channels == 1) return tex2D<float>(t, (float)x, (float)y);
  if (channels == 2) return tex2D<float2>(t, (float)x, (float)y).x;
  return tex2D<float4>(t, (float)x, (float)y).x;
}
template <>
__device__ __forceinline__ unsigned char tex_fetch<unsigned char>(cudaTextureObject_t t, int x, int y,
                                                                   int channels) {
  if (channels == 1) return tex2D<unsigned char>(t, (float)x, (float)y);
  if (channels == 2) return tex2D<uchar2>(t, (float)x, (float)y).x;
  return tex2D<uchar4>(t, (float)x, (float)y).x;
}

// One thread per texel, x fastest: 2-D locality for the texture cache, coalesced stores.
template <typename T>
__global__ void tex_to_linear_kernel(cudaTextureObject_t a, cudaTextureObject_t b, T *__restrict__ oa,
                                     T *__restrict__ ob, int H, int W, int ca, int cb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= W || y >= H) return;
  const size_t i = (size_t)y * W + x;
  oa[i] = tex_fetch<T>(a, x, y, ca);
  if (ob != nullptr) ob[i] = tex_fetch<T>(b, x, y, cb);
}

template <typename T>
int tex_to_linear(const stm_texture *a, const stm_texture *b, T *oa, T *ob, cudaStream_t stream) {
  const dim3 block(32, 8), grid((a->width + 31) / 32, (a->height + 7) / 8);
  tex_to_linear_kernel<T><<<grid, block, 0, stream>>>(a->object, b ? b->object : 0, oa, ob, a->height,
                                                      a->width, a->channels, b ? b->channels : 1);
  return check_launch("tex_to_linear_kernel");
}

int check_texture(const char *op, const stm_texture *t) {
  if (t == nullptr || t->object == 0) return fail(STM_EINVAL, "%s: empty texture", op);
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev != t->device)
    return fail(STM_EINVAL, "%s: the texture lives on device %d, the current device is %d", op, t->device, dev);
  return STM_OK;
}

}  // namespace
}  // namespace stm

using namespace stm;

extern "C" {

int stm_texture_create(const void *data, int data_on_device, int H, int W, int channels, int dtype,
                       int normalized_coords, stm_texture **out) {
  if (out == nullptr) return fail(STM_EINVAL, "texture_create: null output handle");
  *out = nullptr;
  if (data == nullptr || H <= 0 || W <= 0) return fail(STM_EINVAL, "texture_create: invalid tensor");
  // reference src/cuda_texture.cpp:29-50
  if (normalized_coords && ((H & (H - 1)) != 0 || (W & (W - 1)) != 0))
    return fail(STM_EINVAL, "CUDATExture::CopyFromTensor: Only power of 2 texture can use normalized coordinates");
  if (channels == 3)
    return fail(STM_EINVAL, "CUDATExture::CopyFromTensor: Tensor with 3 or more than 4 channels aren't supported");
  if (channels < 1 || channels > 4)
    return fail(STM_EINVAL, "CUDATExture::CopyFromTensor: Tensor with more than 4 channels aren't supported");
  if (dtype != STM_TEX_F32 && dtype != STM_TEX_U8)
    return fail(STM_EINVAL, "CUDATExture::CopyFromTensor: Input tensor's dtype not supported.");
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(STM_ECUDA, "no usable CUDA device (%s); libstm_b200 has no CPU fallback", cudaGetErrorString(e));
  }
  const int bits = dtype == STM_TEX_F32 ? 32 : 8;
  const cudaChannelFormatKind kind = dtype == STM_TEX_F32 ? cudaChannelFormatKindFloat : cudaChannelFormatKindUnsigned;
  const cudaChannelFormatDesc desc = cudaCreateChannelDesc(bits, channels >= 2 ? bits : 0, channels >= 3 ? bits : 0,
                                                           channels >= 4 ? bits : 0, kind);
  stm_texture *t = new stm_texture();
  t->height = H; t->width = W; t->channels = channels; t->dtype = dtype;
  t->normalized = normalized_coords ? 1 : 0; t->device = dev;
  const size_t row_bytes = (size_t)W * channels * (bits / 8);
  e = cudaMallocArray(&t->array, &desc, (size_t)W, (size_t)H, 0);
  if (e == cudaSuccess)
    e = cudaMemcpy2DToArray(t->array, 0, 0, data, row_bytes, row_bytes, (size_t)H,
                            data_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    cudaResourceDesc res;
    memset(&res, 0, sizeof(res));
    res.resType = cudaResourceTypeArray;
    res.res.array.array = t->array;
    cudaTextureDesc td;  // reference src/cuda_texture.cpp:112-118
    memset(&td, 0, sizeof(td));
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint;
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = normalized_coords ? 1 : 0;
    e = cudaCreateTextureObject(&t->object, &res, &td, nullptr);
  }
  if (e != cudaSuccess) {
    if (t->array) cudaFreeArray(t->array);
    delete t;
    cudaGetLastError();
    return fail(STM_ECUDA, "Cuda check error texture_create: %s", cudaGetErrorString(e));
  }
  *out = t;
  return STM_OK;
}

int stm_texture_destroy(stm_texture *t) {
  if (t == nullptr) return STM_OK;
  cudaError_t e = cudaSuccess;
  if (t->object) e = cudaDestroyTextureObject(t->object);
  if (t->array) {
    const cudaError_t e2 = cudaFreeArray(t->array);
    if (e == cudaSuccess) e = e2;
  }
  delete t;
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(STM_ECUDA, "Cuda check error texture_destroy: %s", cudaGetErrorString(e));
  }
  return STM_OK;
}

int stm_texture_info(const stm_texture *t, int *H, int *W, int *channels, int *dtype, int *normalized_coords,
                     unsigned long long *cuda_texture_object) {
  if (t == nullptr) return fail(STM_EINVAL, "texture_info: null texture");
  if (H) *H = t->height;
  if (W) *W = t->width;
  if (channels) *channels = t->channels;
  if (dtype) *dtype = t->dtype;
  if (normalized_coords) *normalized_coords = t->normalized;
  if (cuda_texture_object) *cuda_texture_object = (unsigned long long)t->object;
  return STM_OK;
}

int stm_texture_to_host(const stm_texture *t, void *dst) {
  int rc = check_texture("texture_to_host", t);
  if (rc) return rc;
  if (dst == nullptr) return fail(STM_EINVAL, "texture_to_host: null pointer");
  const size_t row_bytes = (size_t)t->width * t->channels * (t->dtype == STM_TEX_F32 ? 4 : 1);
  STM_CUDA_TRY(cudaMemcpy2DFromArray(dst, row_bytes, t->array, 0, 0, row_bytes, (size_t)t->height,
                                     cudaMemcpyDeviceToHost));
  return STM_OK;
}

int stm_texture_read(const stm_texture *t, void *out, int out_dtype, void *stream) {
  int rc = check_texture("texture_read", t);
  if (rc) return rc;
  if (out == nullptr) return fail(STM_EINVAL, "texture_read: null pointer");
  if (t->normalized) return fail(STM_EUNSUPPORTED, "texture_read: texture uses normalized coordinates");
  if (out_dtype != t->dtype)
    return fail(STM_EINVAL, out_dtype == STM_TEX_F32 ? "Expectating a float scalar" : "texture_read: dtype mismatch");
  if (t->dtype == STM_TEX_F32) return tex_to_linear<float>(t, nullptr, (float *)out, nullptr, (cudaStream_t)stream);
  return tex_to_linear<unsigned char>(t, nullptr, (unsigned char *)out, nullptr, (cudaStream_t)stream);
}

size_t stm_ssd_tex_scratch_bytes(int H, int W) {
  if (H <= 0 || W <= 0) return 0;
  return 2 * (((size_t)H * W * sizeof(float) + 255) & ~(size_t)255);  // two 256-byte aligned images
}

int stm_ssd_tex_f32(const stm_texture *left, const stm_texture *right, float *cost, int D, int kernel_size,
                    void *scratch, size_t scratch_bytes, void *stream) {
  int rc = check_texture("ssd_tex", left);
  if (rc || (rc = check_texture("ssd_tex", right))) return rc;
  if (cost == nullptr) return fail(STM_EINVAL, "ssd_tex: null pointer");
  if (kernel_size < 0) return fail(STM_EINVAL, "ssd_tex: kernel_size must be >= 0");
  // reference include/stereomatch/cuda_texture.hpp:30-37 (accessor<float>() on a byte texture)
  if (left->dtype != STM_TEX_F32 || right->dtype != STM_TEX_F32) return fail(STM_EINVAL, "Expectating a float scalar");
  if (left->normalized || right->normalized)  // reference stereomatch/cost.py:66-68
    return fail(STM_EINVAL, "Texture coordinates can't be normalized for this implementation");
  const int H = left->height, W = left->width;
  if (right->height != H || right->width != W) return fail(STM_EINVAL, "ssd_tex: texture sizes differ");
  if ((rc = check_shape("ssd_tex", H, W, D))) return rc;
  if (scratch == nullptr || scratch_bytes < stm_ssd_tex_scratch_bytes(H, W) ||
      (reinterpret_cast<uintptr_t>(scratch) & 15))
    return fail(STM_EINVAL, "ssd_tex: scratch must hold stm_ssd_tex_scratch_bytes(H, W) bytes, 16-byte aligned");
  float *l = (float *)scratch, *r = l + stm_ssd_tex_scratch_bytes(H, W) / (2 * sizeof(float));
  if ((rc = tex_to_linear<float>(left, right, l, r, (cudaStream_t)stream))) return rc;
  return launch_ssd(l, r, cost, H, W, D, kernel_size, (cudaStream_t)stream);
}

}  // extern "C"
