"""Drop-in for the reference's native module `stereomatch._cstereomatch`.

The reference binds three classes of static methods with pybind11
(reference src/_cstereomatch.cpp:10-17, src/cost.cpp:6-14, src/aggregation.cpp:6,
src/disparity_reduce.cpp:6-9).  Here the same names and argument orders are a ctypes
binding over the C ABI of libstm_b200.so (include/stm_abi.h): torch is used only for
device memory and the current stream.

There is NO CPU fallback: CPU tensors raise RuntimeError, and importing this module
fails loudly if the CUDA library has not been built.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

# STM_B200_LIB: developer knob (tools/variant_libs.py: A/B runs of kernel variants on one box)
_LIB_PATH = Path(os.environ.get("STM_B200_LIB") or Path(__file__).resolve().parent / "libstm_b200.so").resolve()

STM_SGM_MIN_ALL_D = 1
STM_SGM_OVERWRITE = 2
STM_SGM_FUSED_SWEEPS = 4
STM_SGM_DISPARITY_ONLY = 8
STM_SGM_SEQUENTIAL = 16
STM_SGM_OVERLAP = 32
STM_DP_CPU_EXACT = 1
STM_TEX_F32, STM_TEX_U8 = 0, 1

if not _LIB_PATH.exists():
    raise ImportError(
        f"{_LIB_PATH} is missing: build it with `python stereo-matching_b200/build.py` "
        "(stereomatch_b200 has no CPU fallback)")

_lib = ctypes.CDLL(str(_LIB_PATH))
_vp, _i, _f, _u, _sz = (ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_uint,
                        ctypes.c_size_t)
_lib.stm_version.restype = ctypes.c_char_p
_lib.stm_last_error.restype = ctypes.c_char_p
_lib.stm_launch_count.restype = ctypes.c_ulonglong
_lib.stm_ssd_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _i, _vp]
_lib.stm_birchfield_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _vp]
_lib.stm_birchfield_wta_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _vp]
_lib.stm_birchfield_wta_f32.restype = _i
_lib.stm_ssd_wta_scratch_bytes.argtypes = [_i, _i]
_lib.stm_ssd_wta_scratch_bytes.restype = _sz
_lib.stm_ssd_wta_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _sz, _vp]
_lib.stm_ssd_wta_f32.restype = _i
_lib.stm_sgm_workspace_bytes.argtypes = [_i, _i, _i]
_lib.stm_sgm_workspace_bytes.restype = _sz
_lib.stm_sgm_workspace_bytes_flags.argtypes = [_i, _i, _i, _u]
_lib.stm_sgm_workspace_bytes_flags.restype = _sz
_lib.stm_sgm_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _f, _f, _u, _vp, _sz, _vp]
_lib.stm_wta_f32.argtypes = [_vp, _vp, _i, _i, _i, _vp]
_lib.stm_sgm_wta_f32.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _f, _u, _vp, _sz, _vp]
_lib.stm_dp_f32.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _i, _u, _vp]
_lib.stm_sgm_pass_timing.argtypes = [_i]
_lib.stm_sgm_last_pass_ms.argtypes = [_vp, _vp, _i]
_lib.stm_sgm_sweep_aborted.argtypes = [_vp]
_lib.stm_sgm_sweep_aborted.restype = _i
_lib.stm_pipeline_host_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _f, _u, _i, _u]
_pp = ctypes.POINTER
_lib.stm_texture_create.argtypes = [_vp, _i, _i, _i, _i, _i, _i, _pp(_vp)]
_lib.stm_texture_destroy.argtypes = [_vp]
_lib.stm_texture_info.argtypes = [_vp, _pp(_i), _pp(_i), _pp(_i), _pp(_i), _pp(_i), _pp(ctypes.c_ulonglong)]
_lib.stm_texture_to_host.argtypes = [_vp, _vp]
_lib.stm_texture_read.argtypes = [_vp, _vp, _i, _vp]
_lib.stm_ssd_tex_scratch_bytes.argtypes = [_i, _i]
_lib.stm_ssd_tex_scratch_bytes.restype = _sz
_lib.stm_ssd_tex_f32.argtypes = [_vp, _vp, _vp, _i, _i, _vp, _sz, _vp]
for _name in ("stm_ssd_f32", "stm_birchfield_f32", "stm_sgm_f32", "stm_sgm_wta_f32", "stm_wta_f32",
              "stm_dp_f32", "stm_pipeline_host_f32", "stm_texture_create", "stm_texture_destroy",
              "stm_texture_info", "stm_texture_to_host", "stm_texture_read", "stm_ssd_tex_f32"):
    getattr(_lib, _name).restype = _i


def version() -> str:
    return _lib.stm_version().decode()


def launch_count() -> int:
    return int(_lib.stm_launch_count())


def _check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError(_lib.stm_last_error().decode())


def _gpu(name: str, t: torch.Tensor, dtype: torch.dtype, ndim: int, ref: torch.Tensor = None):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name}: expected a torch.Tensor")
    if not t.is_cuda:
        # reference wording: include/stereomatch/check.hpp:27-37
        raise RuntimeError(f"Check failed: {name}: Expected a gpu tensor "
                           "(stereomatch_b200 is CUDA-only, no CPU fallback)")
    if ref is not None and t.device != ref.device:
        raise RuntimeError(f"Check failed: {name}: tensors must be on the same device")
    if t.dtype != dtype:
        raise RuntimeError(f"{name}: type not supported ({t.dtype}; expected {dtype})")
    if t.dim() != ndim:
        raise RuntimeError(f"{name}: expected a {ndim}-d tensor, got shape {tuple(t.shape)}")
    return t


def _out(name: str, t: torch.Tensor) -> torch.Tensor:
    if not t.is_contiguous():
        raise RuntimeError(f"{name}: output tensors must be contiguous")
    return t


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


_workspaces = {}


def _sgm_workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    # one workspace per (device, stream): two pipelines aggregating on different streams at the
    # same time (stereomatch.streaming) must not share the corner-diagonal store
    with torch.cuda.device(device):
        key = (device, _stream())
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


class CUDATexture:
    """reference src/cuda_texture.cpp:167-192 (pybind class `CUDATexture`): a cudaArray + CUDA texture
    object holding an [H, W, C] float32 / uint8 image (C in 1, 2, 4; point filter, clamp).  Same
    methods and read-only attributes; the handle is owned by libstm_b200 (stm_texture_*)."""

    _DTYPES = {torch.float32: STM_TEX_F32, torch.uint8: STM_TEX_U8}

    def __init__(self):
        self._handle = None
        self._device = None
        self.width = 0
        self.height = 0
        self.channels = 0
        self.dtype = torch.float32
        self.cuda_texture_object = 0
        self.use_normalized_coords = False

    def is_empty(self) -> bool:
        return self._handle is None

    def copy_from_tensor(self, tensor: torch.Tensor, normalized_coords: bool = False) -> None:
        """tensor: [H, W, C] float32 / uint8, on the host or on a CUDA device."""
        if not self.is_empty():
            raise RuntimeError("CUDATExture::CopyFromTensor: Loading into a non-empty instance.")
        if tensor.dim() != 3:
            raise RuntimeError("CUDATExture::CopyFromTensor: expected a [H, W, C] tensor")
        if tensor.dtype not in self._DTYPES:
            raise RuntimeError("CUDATExture::CopyFromTensor: Input tensor's dtype not supported.")
        tensor = tensor.contiguous()
        H, W, C = tensor.shape
        # reference src/cuda_texture.cpp:29-50, checked before any CUDA call
        if normalized_coords and ((H & (H - 1)) or (W & (W - 1))):
            raise RuntimeError("CUDATExture::CopyFromTensor: Only power of 2 texture can use "
                               "normalized coordinates")
        if C == 3:
            raise RuntimeError("CUDATExture::CopyFromTensor: Tensor with 3 or more than 4 channels "
                               "aren't supported")
        if C < 1 or C > 4:
            raise RuntimeError("CUDATExture::CopyFromTensor: Tensor with more than 4 channels aren't "
                               "supported")
        if not torch.cuda.is_available():
            raise RuntimeError("CUDATexture needs a CUDA device (stereomatch_b200 has no CPU fallback)")
        device = tensor.device if tensor.is_cuda else torch.device("cuda", torch.cuda.current_device())
        handle = _vp()
        with torch.cuda.device(device):
            if tensor.is_cuda:
                torch.cuda.current_stream().synchronize()   # the copy below is not stream-ordered
            _check(_lib.stm_texture_create(tensor.data_ptr(), int(tensor.is_cuda), H, W, C,
                                           self._DTYPES[tensor.dtype], int(bool(normalized_coords)),
                                           ctypes.byref(handle)))
        self._handle, self._device = handle, device
        self.width, self.height, self.channels, self.dtype = W, H, C, tensor.dtype
        obj = ctypes.c_ulonglong()
        _check(_lib.stm_texture_info(handle, None, None, None, None, None, ctypes.byref(obj)))
        self.cuda_texture_object = int(obj.value)
        self.use_normalized_coords = bool(normalized_coords)

    def to_tensor(self) -> torch.Tensor:
        """[H, W, C] HOST tensor of the texture's dtype (the reference returns a host tensor too)."""
        if self.is_empty():
            raise RuntimeError("CUDATexture::ToTensor: empty texture")
        out = torch.empty(self.height, self.width, self.channels, dtype=self.dtype)
        with torch.cuda.device(self._device):
            _check(_lib.stm_texture_to_host(self._handle, out.data_ptr()))
        return out

    def release(self) -> None:
        if self._handle is not None:
            handle, self._handle = self._handle, None
            self.cuda_texture_object = 0
            _check(_lib.stm_texture_destroy(handle))

    @staticmethod
    def _run_transfer_test_kernel(input_texture: "CUDATexture", output_tensor: torch.Tensor) -> None:
        """reference src/cuda_texture_gpu.cu:29-49: a kernel reads every texel through tex2D and
        writes it to `output_tensor` ([H, W], the texture's dtype, CUDA)."""
        if not output_tensor.is_cuda:
            raise RuntimeError("Cost volume must be a CUDA tensor.")
        if output_tensor.dtype not in CUDATexture._DTYPES:
            raise RuntimeError("Cannot test this type")
        if input_texture.is_empty():
            raise RuntimeError("_run_transfer_test_kernel: empty texture")
        if tuple(output_tensor.shape) != (input_texture.height, input_texture.width):
            raise RuntimeError("_run_transfer_test_kernel: output shape must be [H, W]")
        _out("output_tensor", output_tensor)
        with torch.cuda.device(output_tensor.device):
            _check(_lib.stm_texture_read(input_texture._handle, output_tensor.data_ptr(),
                                         CUDATexture._DTYPES[output_tensor.dtype], _stream()))
            torch.cuda.current_stream().synchronize()     # the reference synchronises the device here

    @staticmethod
    def _run_binding_test_kernel(input_tensor: torch.Tensor, output_tensor: torch.Tensor) -> None:
        """reference src/cuda_texture_gpu.cu:69-110: tensor -> temporary texture -> tensor."""
        tex = CUDATexture()
        tex.copy_from_tensor(input_tensor.reshape(input_tensor.size(0), input_tensor.size(1), 1).float())
        try:
            CUDATexture._run_transfer_test_kernel(tex, output_tensor)
        finally:
            tex.release()


_tex_scratch = {}


def _ssd_tex_scratch(device: torch.device, nbytes: int) -> torch.Tensor:
    with torch.cuda.device(device):
        key = (device, _stream())
    ws = _tex_scratch.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _tex_scratch[key] = ws
    return ws


class CostOps:
    """reference include/stereomatch/cost.hpp:8-20"""

    @staticmethod
    def compute_ssd(left_image, right_image, cost_volume, kernel_size: int) -> None:
        if isinstance(left_image, CUDATexture) or isinstance(right_image, CUDATexture):
            # the reference's second overload (src/cost.cpp:10-13)
            return CostOps._compute_ssd_texture(left_image, right_image, cost_volume, kernel_size)
        left = _gpu("SSDKernel: left_image", left_image, torch.float32, 2)
        right = _gpu("SSDKernel: right_image", right_image, torch.float32, 2, left)
        cost = _out("cost_volume", _gpu("SSDKernel: cost_volume", cost_volume, torch.float32, 3, left))
        H, W = left.shape
        if right.shape != left.shape or cost.shape[:2] != left.shape:
            raise RuntimeError("SSDKernel: image / cost volume shapes do not match")
        left, right = left.contiguous(), right.contiguous()
        with torch.cuda.device(left.device):
            _check(_lib.stm_ssd_f32(left.data_ptr(), right.data_ptr(), cost.data_ptr(), H, W,
                                    cost.shape[2], int(kernel_size), _stream()))

    @staticmethod
    def _compute_ssd_texture(left_image, right_image, cost_volume, kernel_size: int) -> None:
        if not (isinstance(left_image, CUDATexture) and isinstance(right_image, CUDATexture)):
            raise TypeError("compute_ssd: both images must be CUDATexture (or both tensors)")
        if left_image.is_empty() or right_image.is_empty():
            raise RuntimeError("compute_ssd: empty texture")
        if not cost_volume.is_cuda:
            raise RuntimeError("Cost volume must be a CUDA tensor.")  # reference src/ssd.cu:140-142
        cost = _out("cost_volume", _gpu("SSDKernel: cost_volume", cost_volume, torch.float32, 3))
        if cost.device != left_image._device or cost.device != right_image._device:
            raise RuntimeError("Check failed: SSDKernel: textures and cost volume must be on the same device")
        H, W = left_image.height, left_image.width
        if tuple(cost.shape[:2]) != (H, W):
            raise RuntimeError("SSDKernel: texture / cost volume shapes do not match")
        nbytes = int(_lib.stm_ssd_tex_scratch_bytes(H, W))
        ws = _ssd_tex_scratch(cost.device, nbytes)
        with torch.cuda.device(cost.device):
            _check(_lib.stm_ssd_tex_f32(left_image._handle, right_image._handle, cost.data_ptr(),
                                        cost.shape[2], int(kernel_size), ws.data_ptr(), ws.numel(), _stream()))

    @staticmethod
    def compute_birchfield(left_image, right_image, cost_volume, kernel_size: int = 4) -> None:
        left = _gpu("BirchfieldKernel: left_image", left_image, torch.float32, 2)
        right = _gpu("BirchfieldKernel: right_image", right_image, torch.float32, 2, left)
        cost = _out("cost_volume",
                    _gpu("BirchfieldKernel: cost_volume", cost_volume, torch.float32, 3, left))
        H, W = left.shape
        if right.shape != left.shape or cost.shape[:2] != left.shape:
            raise RuntimeError("BirchfieldKernel: image / cost volume shapes do not match")
        left, right = left.contiguous(), right.contiguous()
        with torch.cuda.device(left.device):
            _check(_lib.stm_birchfield_f32(left.data_ptr(), right.data_ptr(), cost.data_ptr(), H, W,
                                           cost.shape[2], _stream()))


    @staticmethod
    def compute_ssd_wta(left_image, right_image, disparity_image, max_disparity: int, kernel_size: int) -> bool:
        """New: compute_ssd + run_winners_take_all in one pass, the volume never written
        (include/stm_abi.h, stm_ssd_wta_f32).  Returns False when the shape is not supported
        (kernel_size not in 1..10, D % 4 != 0): the caller then runs the two stages."""
        left = _gpu("SSDKernel: left_image", left_image, torch.float32, 2)
        right = _gpu("SSDKernel: right_image", right_image, torch.float32, 2, left)
        disp = _out("disparity_image",
                    _gpu("RunWinnersTakeAll: disparity_image", disparity_image, torch.int32, 2, left))
        if right.shape != left.shape or disp.shape != left.shape:
            raise RuntimeError("SSDKernel: image / disparity shapes do not match")
        left, right = left.contiguous(), right.contiguous()
        H, W = left.shape
        nbytes = int(_lib.stm_ssd_wta_scratch_bytes(H, W))
        scratch = _sgm_workspace(left.device, nbytes)   # per (device, stream), shared with the SGM workspace
        with torch.cuda.device(left.device):
            rc = _lib.stm_ssd_wta_f32(left.data_ptr(), right.data_ptr(), disp.data_ptr(), H, W,
                                      int(max_disparity), int(kernel_size), scratch.data_ptr(), scratch.numel(),
                                      _stream())
        if rc == -3:   # STM_EUNSUPPORTED
            return False
        _check(rc)
        return True

    @staticmethod
    def compute_birchfield_wta(left_image, right_image, disparity_image, max_disparity: int) -> bool:
        """New: compute_birchfield + run_winners_take_all in one kernel, the volume never written
        (include/stm_abi.h, stm_birchfield_wta_f32).  Returns False when the shape is not supported
        (D % 32 != 0 ...): the caller then runs the two stages."""
        left = _gpu("BirchfieldKernel: left_image", left_image, torch.float32, 2)
        right = _gpu("BirchfieldKernel: right_image", right_image, torch.float32, 2, left)
        disp = _out("disparity_image",
                    _gpu("RunWinnersTakeAll: disparity_image", disparity_image, torch.int32, 2, left))
        if right.shape != left.shape or disp.shape != left.shape:
            raise RuntimeError("BirchfieldKernel: image / disparity shapes do not match")
        left, right = left.contiguous(), right.contiguous()
        H, W = left.shape
        with torch.cuda.device(left.device):
            rc = _lib.stm_birchfield_wta_f32(left.data_ptr(), right.data_ptr(), disp.data_ptr(), H, W,
                                             int(max_disparity), _stream())
        if rc == -3:   # STM_EUNSUPPORTED
            return False
        _check(rc)
        return True


class AggregationOps:
    """reference include/stereomatch/aggregation.hpp:6-13"""

    @staticmethod
    def run_semiglobal(cost_volume, left_image, penalty1: float, penalty2: float,
                       output_cost_volume, flags: int = 0) -> None:
        cost = _gpu("RunSemiglobal: cost_volume", cost_volume, torch.float32, 3)
        left = _gpu("RunSemiglobal: left_image", left_image, torch.float32, 2, cost)
        out = _out("output_cost_volume",
                   _gpu("RunSemiglobal: output_cost_volume", output_cost_volume, torch.float32, 3, cost))
        if out.shape != cost.shape or tuple(left.shape) != tuple(cost.shape[:2]):
            raise RuntimeError("RunSemiglobal: shapes do not match")
        cost, left = cost.contiguous(), left.contiguous()
        H, W, D = cost.shape
        nbytes = int(_lib.stm_sgm_workspace_bytes_flags(H, W, D, int(flags)))
        ws = _sgm_workspace(cost.device, nbytes)
        with torch.cuda.device(cost.device):
            _check(_lib.stm_sgm_f32(cost.data_ptr(), left.data_ptr(), out.data_ptr(), H, W, D,
                                    float(penalty1), float(penalty2), int(flags), ws.data_ptr(),
                                    ws.numel(), _stream()))


    @staticmethod
    def pass_timing(enable: bool) -> None:
        """New (measurement): record CUDA events around each direction pass of the following
        run_semiglobal / run_semiglobal_wta calls (include/stm_abi.h, stm_sgm_pass_timing)."""
        _check(_lib.stm_sgm_pass_timing(1 if enable else 0))

    @staticmethod
    def sweep_aborted():
        """New: None, or (CTA, row step, site) if a fused sweep on the current device gave up (no
        progress for 60 s: a fault) since the last call; that call's output is invalid.  Ask after
        synchronising the stream (include/stm_abi.h, stm_sgm_sweep_aborted)."""
        info = (ctypes.c_uint * 3)()
        rc = _lib.stm_sgm_sweep_aborted(info)
        if rc < 0:
            _check(rc)
        return (int(info[0]), int(info[1]), int(info[2])) if rc == 1 else None

    @staticmethod
    def last_pass_ms():
        """[(direction, milliseconds)] of the last timed call, in launch order; waits for it."""
        ms = (ctypes.c_float * 8)()
        dirs = (ctypes.c_int * 8)()
        n = _lib.stm_sgm_last_pass_ms(ms, dirs, 8)
        if n < 0:
            _check(n)
        return [(int(dirs[i]), float(ms[i])) for i in range(n)]

    @staticmethod
    def run_semiglobal_wta(cost_volume, left_image, penalty1: float, penalty2: float,
                           output_cost_volume, disparity_image, flags: int = 0) -> None:
        """New: run_semiglobal + run_winners_take_all in one call (include/stm_abi.h,
        stm_sgm_wta_f32); the argmin is taken inside the last aggregation pass."""
        cost = _gpu("RunSemiglobal: cost_volume", cost_volume, torch.float32, 3)
        left = _gpu("RunSemiglobal: left_image", left_image, torch.float32, 2, cost)
        out = _out("output_cost_volume",
                   _gpu("RunSemiglobal: output_cost_volume", output_cost_volume, torch.float32, 3, cost))
        disp = _out("disparity_image",
                    _gpu("RunWinnersTakeAll: disparity_image", disparity_image, torch.int32, 2, cost))
        if out.shape != cost.shape or tuple(left.shape) != tuple(cost.shape[:2]) or \
                tuple(disp.shape) != tuple(cost.shape[:2]):
            raise RuntimeError("RunSemiglobal: shapes do not match")
        cost, left = cost.contiguous(), left.contiguous()
        H, W, D = cost.shape
        ws = _sgm_workspace(cost.device, int(_lib.stm_sgm_workspace_bytes_flags(H, W, D, int(flags))))
        with torch.cuda.device(cost.device):
            _check(_lib.stm_sgm_wta_f32(cost.data_ptr(), left.data_ptr(), out.data_ptr(),
                                        disp.data_ptr(), H, W, D, float(penalty1), float(penalty2),
                                        int(flags), ws.data_ptr(), ws.numel(), _stream()))


class DisparityReduceOps:
    """reference include/stereomatch/disparity_reduce.hpp:6-16"""

    @staticmethod
    def run_winners_take_all(cost_volume, disparity_image) -> None:
        cost = _gpu("RunWinnersTakeAll: cost_volume", cost_volume, torch.float32, 3)
        disp = _out("disparity_image",
                    _gpu("RunWinnersTakeAll: disparity_image", disparity_image, torch.int32, 2, cost))
        if tuple(disp.shape) != tuple(cost.shape[:2]):
            raise RuntimeError("RunWinnersTakeAll: shapes do not match")
        cost = cost.contiguous()
        H, W, D = cost.shape
        with torch.cuda.device(cost.device):
            _check(_lib.stm_wta_f32(cost.data_ptr(), disp.data_ptr(), H, W, D, _stream()))

    @staticmethod
    def run_dynamic_programming(cost_volume, path_volume, row_final_costs, disparity_image,
                                flags: int = 0) -> None:
        cost = _gpu("RunDynamicProgramming: cost_volume", cost_volume, torch.float32, 3)
        path = _out("path_volume",
                    _gpu("RunDynamicProgramming: path_volume", path_volume, torch.int8, 3, cost))
        final = _out("row_final_costs",
                     _gpu("RunDynamicProgramming: row_final_costs", row_final_costs, torch.float32, 2, cost))
        disp = _out("disparity_image",
                    _gpu("RunDynamicProgramming: disparity_image", disparity_image, torch.int32, 2, cost))
        H, W, D = cost.shape
        if path.shape != cost.shape or tuple(final.shape) != (H, D) or tuple(disp.shape) != (H, W):
            raise RuntimeError("RunDynamicProgramming: shapes do not match")
        cost = cost.contiguous()
        with torch.cuda.device(cost.device):
            _check(_lib.stm_dp_f32(cost.data_ptr(), path.data_ptr(), final.data_ptr(),
                                   disp.data_ptr(), H, W, D, int(flags), _stream()))


def pipeline_host(left, right, max_disparity: int, cost_kind: int = 0, kernel_size: int = 7,
                  use_sgm: bool = True, penalty1: float = 0.1, penalty2: float = 0.2,
                  sgm_flags: int = 0, reducer: int = 0, dp_flags: int = 0, out=None):
    """Host-buffer form of Pipeline.estimate: CPU float32 tensors in, CPU int32 map out;
    the host<->device copies happen inside the C call (stm_pipeline_host_f32)."""
    if left.is_cuda or right.is_cuda:
        raise RuntimeError("pipeline_host takes host tensors")
    left = left.contiguous().float()
    right = right.contiguous().float()
    H, W = left.shape
    if out is None:
        out = torch.empty(H, W, dtype=torch.int32)
    _check(_lib.stm_pipeline_host_f32(left.data_ptr(), right.data_ptr(), out.data_ptr(), H, W,
                                      int(max_disparity), int(cost_kind), int(kernel_size),
                                      int(bool(use_sgm)), float(penalty1), float(penalty2),
                                      int(sgm_flags), int(reducer), int(dp_flags)))
    return out
