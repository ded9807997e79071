"""Dataset evaluation loop: predict every scene, score it (the two tasks of reference
workflows/evaluation/workflow.py:22-105, without flytekit / pandas / torch.save round trips).

    from stereomatch.evaluation import evaluate_dataset
    table = evaluate_dataset(MiddleburyDataset(root), create_pipeline("ssd", "wta", "sgm"))
"""
from __future__ import annotations

from typing import Dict, Iterator, List, Tuple

import torch

from . import metrics
from .numeric import next_power_of_2
from .pipeline import Pipeline


def to_grayscale_f32(image: torch.Tensor) -> torch.Tensor:
    """uint8 / float [H,W,3] or [H,W] -> float32 [H,W]; ITU-R 601 luma weights, the formula of
    torchvision's rgb_to_grayscale the reference uses (workflow.py:29-32)."""
    if image.dim() == 2:
        return image.float()
    r, g, b = image[..., 0].float(), image[..., 1].float(), image[..., 2].float()
    return (0.2989 * r + 0.587 * g + 0.114 * b).to(image.dtype if image.is_floating_point() else torch.uint8).float()


def predict_depthmaps(dataset, pipeline: Pipeline, device="cuda") -> Iterator[Tuple[int, str, torch.Tensor]]:
    """Yields (index, scene name, int32 [H,W] disparity on the device) for every scene; the cost
    function's max_disparity follows the scene's ndisp rounded up to a power of two, as the
    reference does (workflow.py:34)."""
    for idx in range(len(dataset)):
        item = dataset.get_stereo_pair(idx)
        left = to_grayscale_f32(item["left"]).to(device)
        right = to_grayscale_f32(item["right"]).to(device)
        cost = getattr(pipeline.cost, "cost_function", pipeline.cost)   # texture wrapper or plain
        cost.max_disparity = next_power_of_2(item["max_disparity"])
        yield idx, item["stereo_name"], pipeline.estimate(left, right)


def evaluate_dataset(dataset, pipeline: Pipeline, device="cuda", threshold: float = 2.0) -> Dict[str, object]:
    """Mean rmse / avgerr / bad_pix_error over the dataset (workflow.py:74-105) and the per-scene
    rows.  Predictions are scored on the device they were computed on."""
    rows: List[Dict[str, object]] = []
    for idx, name, disparity in predict_depthmaps(dataset, pipeline, device):
        gt = dataset.get_ground_truth(idx)["gt_disparity"]
        row = metrics.evaluate(gt.to(disparity.device), disparity, threshold)
        row["stereo_name"] = name
        rows.append(row)
    n = max(len(rows), 1)
    return dict(rmse=sum(r["rmse"] for r in rows) / n, avgerr=sum(r["avgerr"] for r in rows) / n,
                bad_pix_error=sum(r["bad_pix_error"] for r in rows) / n, scenes=rows)
