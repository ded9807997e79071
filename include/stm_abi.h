/*
 * stm_abi.h -- C ABI of libstm_b200.so, the B200-native (sm_100a) dense-stereo hot path.
 *
 * This is the drop-in boundary for the reference's pybind11 module
 * `stereomatch._cstereomatch` (reference src/_cstereomatch.cpp:10-17).  Each entry point
 * below replaces one static method of that module; the reference-side binding a
 * maintainer would add is shown in INTEGRATION.md, and
 * stereo-matching_b200/stereomatch/_cstereomatch.py is that binding (ctypes).
 *
 * Conventions
 *  - `*_f32` entry points take DEVICE pointers and enqueue on `stream` (a cudaStream_t
 *    passed as void*; NULL = legacy default stream) of the CURRENT device; they never
 *    allocate, never synchronise the device, and are CUDA-graph capturable.
 *  - `*_host_f32` entry points take HOST pointers, do the host<->device copies themselves
 *    (synchronous; scratch device memory is cached per thread) and return when the
 *    result is in the host buffer.
 *  - Volumes are contiguous row-major [H, W, D] float32 with D innermost (element (y,x,d)
 *    at (y*W + x)*D + d), images [H, W] float32, disparity maps [H, W] int32 -- the
 *    reference's layout.
 *  - Return value: STM_OK (0) or a negative status; stm_last_error() returns a
 *    thread-local message for the last failure (the binding raises RuntimeError with it,
 *    mirroring the reference's C++ exceptions, include/stereomatch/check.hpp:12-37).
 *  - There is NO CPU fallback: without a CUDA device every compute call fails with
 *    STM_ECUDA.
 */
#ifndef STM_ABI_H_
#define STM_ABI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STM_OK 0
#define STM_EINVAL (-1)       /* bad argument (shape, null pointer, alignment) */
#define STM_ECUDA (-2)        /* CUDA runtime error; message holds cudaGetErrorString */
#define STM_EUNSUPPORTED (-3) /* valid request this build does not implement */

/* flags of stm_sgm_f32 */
#define STM_SGM_MIN_ALL_D 1u /* min_k over all D (reference CUDA kernel, src/semiglobal_gpu.cu:70-81)
                                instead of k in [0, D-2] (compiled CPU reference,
                                src/bordered_buffer.hpp:26-27) */
#define STM_SGM_OVERWRITE 2u /* out = sum (skip the caller's zero fill) instead of out += sum */
#define STM_SGM_FUSED_SWEEPS 4u /* the two horizontal directions as passes of their own, then the six
                                   directions with a vertical component as two fused sweeps
                                   (csrc/sgm_vsweep.cu): four passes over the volumes instead of
                                   eight (11.6 instead of 21.3 GB of DRAM traffic at 1080p D=128).
                                   Same per-direction vectors; the sum is associated per sweep
                                   (0, 1, 2+4+6, 3+5+7) and the corner up-diagonals add mult*L in one
                                   fused multiply-add: within 1e-4 of the reference order, not
                                   bit-identical.  Cooperative launch, one CTA per SM.  Shapes the
                                   sweeps do not take (W > 13 x #SMs, D not in {<= 32, 64, 128, 192,
                                   256}, unaligned pointers) run the per-direction passes. */
#define STM_SGM_DISPARITY_ONLY 8u /* stm_sgm_wta_f32 only: `out` is scratch.  The last pass reduces
                                     the finished sums to the disparity map without writing them
                                     back (4 B/pix-disp less HBM traffic); on return the content
                                     of `out` is UNSPECIFIED (partial sums that depend on the
                                     schedule), only `disparity` is a result. */
#define STM_SGM_SEQUENTIAL 16u /* always run the eight direction passes one after the other.  Without
                                  it, landscape / square volumes of at most 48 MB (latency-bound,
                                  L2-resident: BASELINE.json config 2) run the eight directions
                                  concurrently into per-direction buffers of the workspace and
                                  add them in the reference's order -- bit-identical results. */

#define STM_SGM_OVERLAP 32u /* with STM_SGM_FUSED_SWEEPS: the two horizontal passes run on a library-owned
                               helper stream NEXT TO the down sweep, which writes a volume T of its own
                               in the workspace; the up sweep adds out + T.  Same passes and bytes, less
                               time (the sweeps are latency-bound and leave half the bandwidth idle).
                               Needs stm_sgm_workspace_bytes_flags(H, W, D, flags) bytes of workspace
                               (one more volume); padded D <= 128 (at D = 192 it measured slower).  Falls back to the plain fused schedule
                               when the shape or the workspace does not allow it.  Sum associated
                               ((L0 + L1) + (L2+L4+L6)) + L3 + L5 + L7: within 1e-4, not bit-identical.
                               With STM_SGM_DISPARITY_ONLY at 128 padded disparities the query asks for a
                               second volume U: `<-` writes U = L1 instead of adding into `out`, so the three
                               producers only write and run concurrently, and the up sweep reduces
                               ((out + T) + U) + L3 + L5 + L7 (2.29 -> 2.25 ms at 1080p). */

/* flags of stm_dp_f32.  The DEFAULT mode (no flag) is a defined variant: the reference CUDA functor
 * (src/dynamic_programming.cu:83-130) with its shared-memory race resolved read-before-write.  It
 * equals the reference CUDA kernel bit for bit where that kernel is race-free (D <= 32, tested on
 * the box against the compiled reference) and is otherwise pinned only by the oracle's restatement;
 * it is a different algorithm from the CPU functor (a few % of pixels differ). */
#define STM_DP_CPU_EXACT 1u /* literal reference CPU functor (src/dynamic_programming.cu:24-68),
                               bit-identical to the compiled reference CPU op; requires
                               D <= H <= W and a zeroed path volume; serial in d (slow: ~80 ms
                               at 1080p D=128 against 0.32 ms for the default mode) */

/* Library / build identification, e.g. "stm_b200 0.1 sm_100a". */
const char *stm_version(void);
const char *stm_last_error(void);

/* Replaces CostOps.compute_ssd(left, right, cost_volume, kernel_size)
 * (reference src/cost.cpp:6-9 -> CostOps::ComputeSSD, src/ssd.cu:64-81).
 * cost[y,x,d] = sum over rows [max(y-k,0), min(y+k,H)) x cols [max(x-k,d), min(x+k,W))
 * of (left[r,c] - right[r,c-d])^2 for d <= x, +inf for d > x. */
int stm_ssd_f32(const float *left, const float *right, float *cost, int H, int W, int D,
                int kernel_size, void *stream);

/* Replaces CostOps.compute_birchfield(left, right, cost_volume, kernel_size)
 * (reference src/cost.cpp:14 -> CostOps::ComputeBirchfield, src/birchfield_cost.cu:153-181;
 * kernel size is the reference's template constant 4).  Semantics of the reference CUDA
 * functor (zero-padded scanlines); entries d > x, which the reference leaves unwritten,
 * are set to +inf. */
int stm_birchfield_f32(const float *left, const float *right, float *cost, int H, int W,
                       int D, void *stream);

/* Birchfield cost and winner-takes-all in one kernel, the volume never written: what
 * Pipeline.estimate does for Birchfield + WinnerTakesAll WITHOUT aggregation (reference
 * stereomatch/pipeline.py:80-92 chains compute_birchfield, src/cost.cpp:14, and
 * run_winners_take_all, src/disparity_reduce.cpp:6-9).  The disparity map is exactly stm_wta_f32 of
 * stm_birchfield_f32's volume (tested); HBM traffic is 4 bytes per PIXEL (the map) instead of 8 bytes
 * per pixel-disparity (volume written, then read).  D % 32 == 0 (up to 256) only; other shapes return STM_EUNSUPPORTED
 * (the binding then runs the two stages).  New, SURVEY.md 8(f) N1. */
int stm_birchfield_wta_f32(const float *left, const float *right, int32_t *disparity, int H, int W,
                           int D, void *stream);

/* SSD cost and winner-takes-all in one pass, the volume never written: what Pipeline.estimate does
 * for SSD + WinnerTakesAll WITHOUT aggregation (reference stereomatch/pipeline.py:80-92 chains
 * compute_ssd, src/cost.cpp:9, and run_winners_take_all, src/disparity_reduce.cpp:6-9; BASELINE.json
 * config 1 in its GPU form).  The disparity map is exactly stm_wta_f32 of stm_ssd_f32's volume (tested):
 * the threads that would store a pixel's 128-byte chunk of C send one (cost, disparity) key per pixel and
 * disparity chunk to `scratch` with a 64-bit atomic minimum instead; a second small kernel writes the
 * int32 map.  HBM traffic is 28 bytes per PIXEL instead of 8 bytes per pixel-disparity.
 * `scratch`: stm_ssd_wta_scratch_bytes(H, W) bytes of device memory (H*W 64-bit words; contents
 * irrelevant on entry, garbage on exit).  kernel_size 1..10 and D % 4 == 0 only; other shapes return
 * STM_EUNSUPPORTED (the binding then runs the two stages).  New, SURVEY.md 8(f) N1. */
size_t stm_ssd_wta_scratch_bytes(int H, int W);
int stm_ssd_wta_f32(const float *left, const float *right, int32_t *disparity, int H, int W, int D,
                    int kernel_size, void *scratch, size_t scratch_bytes, void *stream);

/* Replaces AggregationOps.run_semiglobal(cost, left_image, p1, p2, out)
 * (reference src/aggregation.cpp:6 -> AggregationOps::RunSemiglobal, src/semiglobal.cpp:167-197).
 * Aggregates over the reference path set (8 directions, including its stored-end quirk)
 * and ACCUMULATES into `out` unless STM_SGM_OVERWRITE.  `workspace` must hold
 * stm_sgm_workspace_bytes(H, W, D) bytes of device memory (may be NULL if that is 0): one
 * path's worth of vectors for the corner-diagonal replay, the CTA-to-CTA rings and the per-pixel
 * penalties of the fused sweeps and, for volumes of at most 48 MB, eight
 * volume-sized buffers (those
 * shapes run the eight directions concurrently on library-owned streams that are forked from and
 * joined to `stream`; see STM_SGM_SEQUENTIAL).  The query is a pure function of the shape. */
size_t stm_sgm_workspace_bytes(int H, int W, int D);
/* The same query for calls that pass `flags`: what the FASTEST schedule for those flags wants.  More than
 * stm_sgm_workspace_bytes in two cases: STM_SGM_FUSED_SWEEPS | STM_SGM_OVERLAP asks for one more volume (two
 * with STM_SGM_DISPARITY_ONLY at D = 128); and a volume of 48 to 384 MB that does not take
 * the fused sweeps (no STM_SGM_FUSED_SWEEPS, or a D they do not handle) asks for the eight direction buffers of
 * the concurrent schedule -- such volumes are latency-bound one pass after the other (375x450x128: 0.68 ->
 * 0.45 ms).  A workspace of only stm_sgm_workspace_bytes is always accepted; the call then runs the passes
 * one after the other.  New; the reference allocates its scratch itself (thrust::device_vector per call,
 * src/semiglobal_gpu.cu:126). */
size_t stm_sgm_workspace_bytes_flags(int H, int W, int D, unsigned flags);
int stm_sgm_f32(const float *cost, const float *left, float *out, int H, int W, int D,
                float penalty1, float penalty2, unsigned flags, void *workspace,
                size_t workspace_bytes, void *stream);

/* Aggregation and winner-takes-all in one call (what Pipeline.estimate does for
 * Semiglobal + WinnerTakesAll, reference stereomatch/pipeline.py:85-92): same result as
 * stm_sgm_f32 followed by stm_wta_f32 on `out`, but the argmin is taken inside the last
 * aggregation pass, so the aggregated volume is never re-read.  To let that pass visit every
 * pixel exactly once, the last two directions are run in swapped order (6 after 7): `out`
 * equals stm_sgm_f32's result up to fp32 association (a few ulp).  `out` is still written
 * unless STM_SGM_DISPARITY_ONLY is given (what Pipeline.estimate uses: it never hands the
 * aggregated volume to its caller).
 * flags: STM_SGM_MIN_ALL_D | STM_SGM_OVERWRITE | STM_SGM_DISPARITY_ONLY | STM_SGM_SEQUENTIAL |
 * STM_SGM_FUSED_SWEEPS (then the up sweep takes the argmin; no direction swap is needed). */
int stm_sgm_wta_f32(const float *cost, const float *left, float *out, int32_t *disparity, int H,
                    int W, int D, float penalty1, float penalty2, unsigned flags, void *workspace,
                    size_t workspace_bytes, void *stream);

/* Aggregation that allocates NOTHING, in the caller or in the library.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * stm_sgm_f32 with STM_SGM_SEQUENTIAL (the eight direction passes one after the other in the
 * reference's order, accumulating into `out` unless STM_SGM_OVERWRITE), but the only scratch that
 * schedule needs -- the corner up-diagonal's L vectors, min(H, W) * padded-D floats -- is a
 * file-scope `static __device__` array of STM_SGM_STATIC_MAX_FLOATS floats (2 MiB, held by every
 * GPU that loads the library until exit) instead of a workspace argument.  padded D = D rounded
 * up to a multiple of 32.  Shapes with min(H, W) * padded D > STM_SGM_STATIC_MAX_FLOATS are
 * rejected with STM_EINVAL (1080p fits up to D = 480; use stm_sgm_f32 beyond).  Calls no
 * cudaMalloc / cudaFree / allocating library routine; enqueues on `stream` like the other *_f32
 * entry points.  flags: STM_SGM_MIN_ALL_D | STM_SGM_OVERWRITE.
 * SHARED STATE: every call on a device uses the same array, and it keeps the last call's contents.
 * Nothing in it needs to start at zero (each call writes what it later reads, in stream order), so
 * back-to-back calls on one stream are safe; calls that OVERLAP on one GPU (two streams, two host
 * threads) race on it and give wrong sums on the corner diagonal. */
#define STM_SGM_STATIC_MAX_FLOATS 524288
int stm_sgm_static_f32(const float *cost, const float *left, float *out, int H, int W, int D,
                       float penalty1, float penalty2, unsigned flags, void *stream);

/* Measurement API (new; bench.py's per-pass roofline lines).  With timing enabled, every
 * stm_sgm_f32 / stm_sgm_wta_f32 call records CUDA events on its stream around each of the
 * passes (a pass = one direction's kernel plus its corner-diagonal replay, see SURVEY.md F5, or
 * one fused sweep: directions[] then holds 2 for 2+4+6 and 3 for 3+5+7).  stm_sgm_last_pass_ms
 * waits for the last call's final event and returns the number of passes (8, or 4 with
 * STM_SGM_FUSED_SWEEPS; 0 if no call was timed), filling ms[i] and directions[i] in launch
 * order.  Process-wide state, not thread-safe: a measurement switch, not a product feature. */
int stm_sgm_pass_timing(int enable);
int stm_sgm_last_pass_ms(float *ms, int *directions, int capacity);

/* New; no counterpart in the reference (whose kernels cannot wait for each other).  The fused sweeps
 * (STM_SGM_FUSED_SWEEPS) are CTAs that hand vectors to each other; a polling loop that sees no
 * progress for 60 s gives up COOPERATIVELY -- every CTA of that launch leaves, the kernel ends, the
 * CUDA context stays usable -- and the call's output is unspecified.  Calls are asynchronous, so the
 * host learns of it here: returns 1 if a sweep on the current device has given up since the last
 * call of this function (info3, if not NULL, gets {CTA, image row step, 0 = importer waiting for
 * its neighbour / 1 = exporter waiting for credit}) and clears the record, 0 if none, < 0 on error.
 * Meaningful after the stream has been synchronised.  The next fused-sweep call on the device also
 * fails with STM_ECUDA and the record in stm_last_error() if nobody has collected it.  It cannot
 * happen unless a CTA stops (a fault); it replaces round 1's __trap(). */
int stm_sgm_sweep_aborted(unsigned *info3);

/* Replaces DisparityReduceOps.run_winners_take_all(cost, disparity)
 * (reference src/disparity_reduce.cpp:6-9 -> RunWinnersTakeAll, src/winners_take_all.cu:89-114).
 * CPU rule: first (lowest-d) minimum, all-inf -> 0, NaN never wins.  Any D >= 1. */
int stm_wta_f32(const float *cost, int32_t *disparity, int H, int W, int D, void *stream);

/* Replaces DisparityReduceOps.run_dynamic_programming(cost, path, row_final, disparity)
 * (reference src/disparity_reduce.cpp:6-9 -> RunDynamicProgramming,
 * src/dynamic_programming.cu:184-225).  path: int8 [H,W,D]; row_final: float32 [H,D]. */
int stm_dp_f32(const float *cost, int8_t *path, float *row_final, int32_t *disparity, int H,
               int W, int D, unsigned flags, void *stream);

/* Host-buffer form of Pipeline.estimate (reference stereomatch/pipeline.py:69-94):
 * cost (0 = SSD, 1 = Birchfield) -> optional SGM (use_sgm) -> reducer (0 = WTA, 1 = DP).
 * left/right/disparity are HOST pointers; copies are inside the call. */
int stm_pipeline_host_f32(const float *left, const float *right, int32_t *disparity, int H,
                          int W, int D, int cost_kind, int kernel_size, int use_sgm,
                          float penalty1, float penalty2, unsigned sgm_flags, int reducer,
                          unsigned dp_flags);

/* ---- CUDA texture objects as SSD inputs (the reference's CUDATexture class and the
 * compute_ssd(CUDATexture, CUDATexture, cost, k) overload) ------------------ */
#define STM_TEX_F32 0
#define STM_TEX_U8 1
typedef struct stm_texture stm_texture; /* opaque: a cudaArray + cudaTextureObject_t */

/* Replaces CUDATexture.copy_from_tensor(tensor[H,W,C], normalized_coords)
 * (reference src/cuda_texture.cpp:23-131): `data` is a contiguous [H, W, channels] image of
 * float32 (STM_TEX_F32) or uint8 (STM_TEX_U8) in host or device memory (data_on_device); 1, 2 or
 * 4 channels; point filter, clamp addressing, element-type reads, as the reference.  Normalized
 * coordinates need power-of-two sizes.  Synchronous (allocates a cudaArray and copies). */
int stm_texture_create(const void *data, int data_on_device, int H, int W, int channels, int dtype,
                       int normalized_coords, stm_texture **out);
/* Replaces CUDATexture.release() (src/cuda_texture.cpp:12-21).  NULL is a no-op. */
int stm_texture_destroy(stm_texture *texture);
/* The read-only attributes of the reference class (src/cuda_texture.cpp:177-186): height, width,
 * channels, dtype, use_normalized_coords, cuda_texture_object.  Any output pointer may be NULL. */
int stm_texture_info(const stm_texture *texture, int *H, int *W, int *channels, int *dtype,
                     int *normalized_coords, unsigned long long *cuda_texture_object);
/* Replaces CUDATexture.to_tensor() (src/cuda_texture.cpp:133-165): copies the texels to a HOST
 * buffer [H, W, channels] of the texture's dtype (the reference always returns a float32 host
 * tensor, which is wrong for uint8 textures; here the dtype is the texture's).  Synchronous. */
int stm_texture_to_host(const stm_texture *texture, void *dst);
/* Replaces CUDATexture._run_transfer_test_kernel(texture, out) (src/cuda_texture_gpu.cu:9-49):
 * a kernel reads every texel with tex2D at (col, row) and writes channel 0 to the DEVICE image
 * out[H, W] of the texture's dtype (out_dtype must match).  Enqueued on `stream`. */
int stm_texture_read(const stm_texture *texture, void *out, int out_dtype, void *stream);
/* Replaces CostOps.compute_ssd(CUDATexture, CUDATexture, cost_volume, kernel_size)
 * (reference src/cost.cpp:10-13 -> CostOps::ComputeSSD, src/ssd.cu:137-153; kernel :83-135).  Same
 * values as stm_ssd_f32 on the images the textures hold.  Both textures float32, unnormalized,
 * same size; cost is a DEVICE volume [H, W, D].  `scratch`: stm_ssd_tex_scratch_bytes(H, W) bytes of
 * device memory (every texel is fetched once into it, then the tiled SSD kernel runs). */
size_t stm_ssd_tex_scratch_bytes(int H, int W);
int stm_ssd_tex_f32(const stm_texture *left, const stm_texture *right, float *cost, int D,
                    int kernel_size, void *scratch, size_t scratch_bytes, void *stream);

/* Number of kernel launches the library has enqueued in this process (bench "gpu_launches"). */
unsigned long long stm_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* STM_ABI_H_ */
