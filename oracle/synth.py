"""Deterministic synthetic inputs (random-dot stereograms), BASELINE.md section 4.

base = randint(0,256,(H, W+D), seed); gt = D//8 everywhere except the centred half-size
rectangle where gt = D//2; right = base[:, D:]; left[y,x] = base[y, x + D - gt[y,x]]
(so left[y,x] == right[y, x-gt]).  float32, raw (0..255) or unit (/255) scale.
Shared by tests/ and bench.py (inputs only -- not part of the oracle's arithmetic).
"""
from __future__ import annotations

import numpy as np
import torch


def random_dot_pair(H: int, W: int, D: int, seed: int = 1000, scale: str = "raw"):
    """Returns (left, right, gt) as numpy arrays (float32, float32, int32)."""
    g = torch.Generator().manual_seed(int(seed))
    base = torch.randint(0, 256, (H, W + D), generator=g).numpy().astype(np.float32)
    gt = np.full((H, W), D // 8, np.int32)
    gt[H // 4:(3 * H) // 4, W // 4:(3 * W) // 4] = D // 2
    right = np.ascontiguousarray(base[:, D:])
    xs = np.arange(W, dtype=np.int64)[None, :] + D - gt
    left = np.take_along_axis(base, xs, axis=1)
    left = np.ascontiguousarray(left, dtype=np.float32)
    if scale == "unit":
        left = left / np.float32(255.0)
        right = right / np.float32(255.0)
    elif scale != "raw":
        raise ValueError(scale)
    return left, right, gt


def tie_heavy_volume(H: int, W: int, D: int, seed: int = 7) -> np.ndarray:
    """floor(4*rand) volume: many exact ties (WTA / DP tie-break tests)."""
    rng = np.random.default_rng(seed)
    return np.floor(4.0 * rng.random((H, W, D))).astype(np.float32)


def random_volume(H: int, W: int, D: int, seed: int = 11, hi: float = 100.0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return (hi * rng.random((H, W, D))).astype(np.float32)
