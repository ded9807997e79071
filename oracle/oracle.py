"""numpy front-end of the CPU oracle (oracle/stm_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of stm_oracle.c.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.

Every function takes/returns C-contiguous numpy arrays; float32 volumes are laid out
[H, W, D] with D innermost exactly like the reference's cost volumes.
"""
from __future__ import annotations

import ctypes
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libstm_oracle.so"
SRC_PATH = HERE / "stm_oracle.c"

_lib = None


def build(force: bool = False) -> Path:
    """gcc-compiles the C restatement (a few hundred ms)."""
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < SRC_PATH.stat().st_mtime:
        subprocess.check_call(
            ["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-std=c99",
             "-o", str(LIB_PATH), str(SRC_PATH), "-lm"])
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(str(LIB_PATH))
        _lib.stm_oracle_dp.restype = ctypes.c_int
        _lib.stm_oracle_sgm_num_paths.restype = ctypes.c_int
        _lib.stm_oracle_sgm_paths.restype = ctypes.c_int
    return _lib


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def ssd(left, right, max_disparity: int, kernel_size: int = 7) -> np.ndarray:
    left, right = _f32(left), _f32(right)
    H, W = left.shape
    cost = np.empty((H, W, max_disparity), np.float32)
    lib().stm_oracle_ssd(_p(left), _p(right), _p(cost), H, W, max_disparity, kernel_size)
    return cost


def birchfield(left, right, max_disparity: int, fill_inf: bool = True,
               cpu_functor: bool = False, out: np.ndarray | None = None) -> np.ndarray:
    """Canonical (reference CUDA functor) semantics; cpu_functor=True reproduces the
    reference CPU functor (one-pixel shift, out-of-bounds at the right border)."""
    left, right = _f32(left), _f32(right)
    H, W = left.shape
    cost = out if out is not None else np.full((H, W, max_disparity), np.nan, np.float32)
    lib().stm_oracle_birchfield(_p(left), _p(right), _p(cost), H, W, max_disparity,
                                int(fill_inf), int(cpu_functor))
    return cost


def sgm_paths(H: int, W: int) -> np.ndarray:
    """[n, 7] int32: start.x start.y end.x end.y dir.x dir.y size (reference list order)."""
    n = lib().stm_oracle_sgm_num_paths(H, W)
    desc = np.empty((n, 7), np.int32)
    got = lib().stm_oracle_sgm_paths(H, W, _p(desc))
    assert got == n
    return desc


def sgm_visit_counts(H: int, W: int) -> np.ndarray:
    """[8, H, W] visits per direction: (1,0) (-1,0) (0,1) (0,-1) (1,1) (-1,-1) (-1,1) (1,-1)."""
    count = np.empty((8, H, W), np.int32)
    lib().stm_oracle_sgm_visit_counts(H, W, _p(count))
    return count


def sgm(cost, left, penalty1: float = 0.1, penalty2: float = 0.2,
        min_over_all_d: bool = False, out: np.ndarray | None = None) -> np.ndarray:
    cost, left = _f32(cost), _f32(left)
    H, W, D = cost.shape
    out = np.zeros_like(cost) if out is None else out
    lib().stm_oracle_sgm(_p(cost), _p(left), _p(out), H, W, D,
                         ctypes.c_float(penalty1), ctypes.c_float(penalty2),
                         int(min_over_all_d))
    return out


def wta(cost) -> np.ndarray:
    cost = _f32(cost)
    H, W, D = cost.shape
    disp = np.empty((H, W), np.int32)
    lib().stm_oracle_wta(_p(cost), _p(disp), H, W, D)
    return disp


def dp(cost, mode: str = "cuda", return_aux: bool = False):
    """mode 'cuda' (Jacobi, reference CUDA kernel semantics) or 'cpu_exact' (literal
    reference CPU functor; path volume pre-zeroed)."""
    cost = _f32(cost)
    H, W, D = cost.shape
    path = np.zeros((H, W, D), np.int8)
    row_final = np.empty((H, D), np.float32)
    disp = np.empty((H, W), np.int32)
    rc = lib().stm_oracle_dp(_p(cost), _p(path), _p(row_final), _p(disp), H, W, D,
                             {"cuda": 0, "cpu_exact": 1}[mode])
    if rc != 0:
        raise ValueError("dp cpu_exact requires D <= H <= W (the reference reads out of bounds otherwise)")
    if return_aux:
        return disp, path, row_final
    return disp
