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
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent / "libstm_b200.so"

STM_SGM_MIN_ALL_D = 1
STM_SGM_OVERWRITE = 2
STM_SGM_FUSED_SWEEPS = 4
STM_SGM_DISPARITY_ONLY = 8
STM_SGM_SEQUENTIAL = 16
STM_DP_CPU_EXACT = 1

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
_lib.stm_sgm_workspace_bytes.argtypes = [_i, _i, _i]
_lib.stm_sgm_workspace_bytes.restype = _sz
_lib.stm_sgm_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _f, _f, _u, _vp, _sz, _vp]
_lib.stm_wta_f32.argtypes = [_vp, _vp, _i, _i, _i, _vp]
_lib.stm_sgm_wta_f32.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _f, _u, _vp, _sz, _vp]
_lib.stm_dp_f32.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _i, _u, _vp]
_lib.stm_sgm_pass_timing.argtypes = [_i]
_lib.stm_sgm_last_pass_ms.argtypes = [_vp, _vp, _i]
_lib.stm_pipeline_host_f32.argtypes = [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _f, _u, _i, _u]
for _name in ("stm_ssd_f32", "stm_birchfield_f32", "stm_sgm_f32", "stm_sgm_wta_f32", "stm_wta_f32",
              "stm_dp_f32", "stm_pipeline_host_f32"):
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


class CostOps:
    """reference include/stereomatch/cost.hpp:8-20"""

    @staticmethod
    def compute_ssd(left_image, right_image, cost_volume, kernel_size: int) -> None:
        if not isinstance(left_image, torch.Tensor):
            raise RuntimeError("compute_ssd(CUDATexture, ...) is not part of the B200 hot path")
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
        nbytes = int(_lib.stm_sgm_workspace_bytes(H, W, D))
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
        ws = _sgm_workspace(cost.device, int(_lib.stm_sgm_workspace_bytes(H, W, D)))
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
