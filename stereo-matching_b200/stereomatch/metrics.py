"""Disparity-map quality metrics on the device (the evaluation step after the path, reference
workflows/evaluation/workflow.py:74-105).

The reference calls `stereomideval.eval.Metric.calc_rmse / calc_avgerr / calc_bad_pix_error`
(an un-vendored, unpinned dependency: `stereomideval` in workflows/evaluation; absent from
/root/reference and from this image).  Its published definitions, restated as torch ops that run
where the tensors live (the disparity map never leaves the GPU):

    valid   = ground truth finite (Middlebury PFM stores +inf where the disparity is unknown)
    diff    = |prediction - ground_truth| on valid pixels
    rmse    = sqrt(mean(diff^2))           avgerr = mean(diff)
    bad_pix_error(t) = 100 * count(diff > t) / count(valid)      (t = 2.0 by default)

Everything is accumulated in float64 and returned as Python floats (one device->host read per metric;
`evaluate` reads all three with a single transfer).
"""
from __future__ import annotations

from typing import Dict, Optional

import torch


def _diff(ground_truth: torch.Tensor, prediction: torch.Tensor, mask: Optional[torch.Tensor] = None):
    if ground_truth.shape != prediction.shape:
        raise ValueError(f"metrics: shapes differ {tuple(ground_truth.shape)} vs {tuple(prediction.shape)}")
    gt = ground_truth.to(device=prediction.device, dtype=torch.float64)
    valid = torch.isfinite(gt)
    if mask is not None:
        valid = valid & mask.to(prediction.device, torch.bool)
    diff = (prediction.to(torch.float64) - torch.where(valid, gt, torch.zeros_like(gt))).abs()
    return torch.where(valid, diff, torch.zeros_like(diff)), valid


def _stats(ground_truth, prediction, threshold: float, mask=None) -> torch.Tensor:
    diff, valid = _diff(ground_truth, prediction, mask)
    n = valid.sum().to(torch.float64).clamp_min(1.0)
    return torch.stack([torch.sqrt((diff * diff).sum() / n), diff.sum() / n,
                        100.0 * ((diff > threshold) & valid).sum().to(torch.float64) / n, valid.sum().to(torch.float64)])


def evaluate(ground_truth: torch.Tensor, prediction: torch.Tensor, threshold: float = 2.0,
             mask: Optional[torch.Tensor] = None) -> Dict[str, float]:
    """{'rmse', 'avgerr', 'bad_pix_error', 'valid_pixels'} of one map; `mask` (optional, bool [H,W])
    restricts the valid set further (e.g. to x >= d, where a left-image pixel has a match)."""
    rmse, avgerr, bad, n = _stats(ground_truth, prediction, threshold, mask).tolist()
    return dict(rmse=rmse, avgerr=avgerr, bad_pix_error=bad, valid_pixels=int(n))


def calc_rmse(ground_truth: torch.Tensor, prediction: torch.Tensor) -> float:
    return evaluate(ground_truth, prediction)["rmse"]


def calc_avgerr(ground_truth: torch.Tensor, prediction: torch.Tensor) -> float:
    return evaluate(ground_truth, prediction)["avgerr"]


def calc_bad_pix_error(ground_truth: torch.Tensor, prediction: torch.Tensor, threshold: float = 2.0) -> float:
    return evaluate(ground_truth, prediction, threshold)["bad_pix_error"]
