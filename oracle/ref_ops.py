"""Loader for the UNMODIFIED reference native module built by oracle/build_ref.py.

TEST INFRASTRUCTURE ONLY.  `load()` returns the pybind module whose classes are the
reference's own `CostOps`, `AggregationOps`, `DisparityReduceOps`
(reference src/_cstereomatch.cpp:10-17) or None when oracle/_ref/_cstereomatch.so has not
been built.  The thin functions below call the reference ops on CPU tensors the way the
reference's Python wrappers do (stereomatch/cost.py, aggregation.py, disparity_reduce.py)
-- those wrappers themselves are not copied; they only allocate the outputs.
"""
from __future__ import annotations

import importlib.util
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
SO_PATH = HERE / "_ref" / "_cstereomatch.so"

_mod = None


def available() -> bool:
    return SO_PATH.exists()


def load():
    global _mod
    if _mod is None:
        if not SO_PATH.exists():
            return None
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        spec = importlib.util.spec_from_file_location("_cstereomatch", str(SO_PATH))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _mod = mod
    return _mod


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))


def ssd(left, right, max_disparity, kernel_size=7, device="cpu"):
    import torch
    l, r = _t(left).to(device), _t(right).to(device)
    cost = torch.empty(l.shape[0], l.shape[1], max_disparity, dtype=torch.float32, device=device)
    load().CostOps.compute_ssd(l, r, cost, kernel_size)
    return cost.cpu().numpy()


def birchfield(left, right, max_disparity, device="cpu", prefill=float("nan")):
    import torch
    l, r = _t(left).to(device), _t(right).to(device)
    cost = torch.full((l.shape[0], l.shape[1], max_disparity), prefill, dtype=torch.float32,
                      device=device)
    load().CostOps.compute_birchfield(l, r, cost, 4)
    if device != "cpu":
        torch.cuda.synchronize()
    return cost.cpu().numpy()


def sgm(cost, left, penalty1=0.1, penalty2=0.2, device="cpu"):
    import torch
    c, l = _t(cost).to(device), _t(left).to(device)
    out = torch.zeros_like(c)
    load().AggregationOps.run_semiglobal(c, l, penalty1, penalty2, out)
    if device != "cpu":
        torch.cuda.synchronize()
    return out.cpu().numpy()


def wta(cost, device="cpu"):
    import torch
    c = _t(cost).to(device)
    disp = torch.empty(c.shape[0], c.shape[1], dtype=torch.int32, device=device)
    load().DisparityReduceOps.run_winners_take_all(c, disp)
    if device != "cpu":
        torch.cuda.synchronize()
    return disp.cpu().numpy()


def dp(cost, device="cpu", return_aux=False):
    import torch
    c = _t(cost).to(device)
    H, W, D = c.shape
    disp = torch.empty(H, W, dtype=torch.int32, device=device)
    path = torch.zeros(H, W, D, dtype=torch.int8, device=device)  # pre-zeroed
    row_final = torch.empty(H, D, dtype=torch.float32, device=device)
    load().DisparityReduceOps.run_dynamic_programming(c, path, row_final, disp)
    if return_aux:
        return disp.cpu().numpy(), path.cpu().numpy(), row_final.cpu().numpy()
    return disp.cpu().numpy()
