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


def _row_chunks(H: int, threads: int):
    n = max(1, min(int(threads), H))
    edges = [round(i * H / n) for i in range(n + 1)]
    return [(a, b) for a, b in zip(edges[:-1], edges[1:]) if b > a]


def default_threads() -> int:
    import os
    return max(1, min(32, os.cpu_count() or 1))


def ssd_threaded(left, right, max_disparity: int, kernel_size: int = 7,
                 threads: int | None = None) -> np.ndarray:
    """stm_oracle_ssd over row chunks on a thread pool (ctypes releases the GIL).  Output row y
    only reads image rows [y-k, y+k), so each chunk is the literal function on a slice of the
    images: bit-identical to ssd() (tests/test_oracle_golden.py), just faster at 1080p."""
    from concurrent.futures import ThreadPoolExecutor
    left, right = _f32(left), _f32(right)
    H, W = left.shape
    k = int(kernel_size)
    cost = np.empty((H, W, max_disparity), np.float32)

    def work(ab):
        a, b = ab
        lo, hi = max(a - k, 0), min(b + k, H)
        cost[a:b] = ssd(left[lo:hi], right[lo:hi], max_disparity, k)[a - lo:b - lo]

    with ThreadPoolExecutor(threads or default_threads()) as ex:
        list(ex.map(work, _row_chunks(H, (threads or default_threads()) * 4)))
    return cost


def ssd_exact_int(left, right, max_disparity: int, kernel_size: int = 7,
                  threads: int | None = None) -> np.ndarray:
    """SSD of INTEGER-VALUED images (the raw 0..255 scale) by int64 integral images.  Every
    partial sum of the reference's fp32 loop (src/ssd.cu:44-56) is then an exact integer below
    2^24 (196 * 255^2 = 12,744,900), so the reference's result is the exact integer sum whatever
    the order: this returns exactly what ssd() returns (checked in tests/test_oracle_golden.py),
    in seconds instead of a minute at 1080p.  Raises for non-integer images or sums >= 2^24."""
    from concurrent.futures import ThreadPoolExecutor
    left, right = _f32(left), _f32(right)
    L, R = left.astype(np.int64), right.astype(np.int64)
    if not (np.array_equal(L, left) and np.array_equal(R, right)):
        raise ValueError("ssd_exact_int needs integer-valued images")
    H, W = L.shape
    k, D = int(kernel_size), int(max_disparity)
    dmax = int(max(L.max(initial=0) - R.min(initial=0), R.max(initial=0) - L.min(initial=0)))
    if (2 * k) * (2 * k) * dmax * dmax >= (1 << 24):
        raise ValueError("window sums may exceed 2^24: not exact in fp32")
    cost = np.full((H, W, D), np.inf, np.float32)
    y, x = np.arange(H), np.arange(W)
    r0, r1 = np.maximum(y - k, 0), np.maximum(np.minimum(y + k, H), np.maximum(y - k, 0))

    def work(d):
        diff = L[:, d:] - R[:, :W - d]                 # column c = d .. W-1 at index c - d
        I = np.zeros((H + 1, W - d + 1), np.int64)
        np.cumsum(np.cumsum(diff * diff, axis=0), axis=1, out=I[1:, 1:])
        xs = x[d:]
        a = np.maximum(xs - k, d) - d                  # src/ssd.cu:40-41
        b = np.maximum(np.minimum(xs + k, W) - d, a)   # src/ssd.cu:42 (empty window -> 0)
        top, bot = I[r0], I[r1]
        cost[:, d:, d] = (bot[:, b] - bot[:, a]) - (top[:, b] - top[:, a])

    with ThreadPoolExecutor(threads or default_threads()) as ex:
        list(ex.map(work, range(min(D, W))))
    return cost


def birchfield_threaded(left, right, max_disparity: int, threads: int | None = None) -> np.ndarray:
    """Canonical birchfield() over row chunks on a thread pool (scanlines are independent)."""
    from concurrent.futures import ThreadPoolExecutor
    left, right = _f32(left), _f32(right)
    H, W = left.shape
    cost = np.empty((H, W, max_disparity), np.float32)

    def work(ab):
        a, b = ab
        birchfield(left[a:b], right[a:b], max_disparity, out=cost[a:b])

    with ThreadPoolExecutor(threads or default_threads()) as ex:
        list(ex.map(work, _row_chunks(H, (threads or default_threads()) * 4)))
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
        min_over_all_d: bool = False, out: np.ndarray | None = None,
        directions: int = 0xff) -> np.ndarray:
    """directions: bit i keeps the directed paths of direction i (order of sgm_visit_counts);
    0xff is the reference's whole aggregation."""
    cost, left = _f32(cost), _f32(left)
    H, W, D = cost.shape
    out = np.zeros_like(cost) if out is None else out
    lib().stm_oracle_sgm_dirs(_p(cost), _p(left), _p(out), H, W, D,
                              ctypes.c_float(penalty1), ctypes.c_float(penalty2),
                              int(min_over_all_d), ctypes.c_uint(directions))
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
