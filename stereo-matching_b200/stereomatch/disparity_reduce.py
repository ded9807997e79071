"""Cost volume -> disparity map (API of reference stereomatch/disparity_reduce.py)."""
from typing import Optional

import torch

from ._common import empty_tensor
from ._cstereomatch import STM_DP_CPU_EXACT
from ._cstereomatch import DisparityReduceOps as _DisparityReduceOps


class WinnerTakesAll:
    """argmin over d with the reference CPU rule: lowest d among equal minima
    (reference stereomatch/disparity_reduce.py:16-46, src/winners_take_all.cu:21-40).
    The reference's power-of-two guard for CUDA tensors (disparity_reduce.py:40-43) is
    dropped: any D >= 1 works."""

    def __call__(self, cost_volume: torch.Tensor,
                 disparity_img: Optional[torch.Tensor] = None) -> torch.Tensor:
        disparity_img = empty_tensor(cost_volume.size(0), cost_volume.size(1), dtype=torch.int32,
                                     device=cost_volume.device, reuse_tensor=disparity_img)
        _DisparityReduceOps.run_winners_take_all(cost_volume, disparity_img)
        return disparity_img


class DynamicProgramming:
    """Scanline dynamic programming (reference stereomatch/disparity_reduce.py:49-90).

    mode="cuda" (default): the reference CUDA kernel's recurrence (all three neighbours
    from the previous column).  mode="cpu_exact": the reference CPU functor literally
    (needs D <= H <= W).  The int8 path volume and the [H,D] final costs are cached on the
    object between calls (the reference re-allocates them every call)."""

    def __init__(self, mode: str = "cuda"):
        if mode not in ("cuda", "cpu_exact"):
            raise ValueError("mode must be 'cuda' or 'cpu_exact'")
        self.mode = mode
        self._path_volume = None
        self._row_final_costs = None

    def __call__(self, cost_volume: torch.Tensor,
                 disparity_img: Optional[torch.Tensor] = None) -> torch.Tensor:
        height, width, max_disparity = cost_volume.shape
        disparity_img = empty_tensor(height, width, dtype=torch.int32, device=cost_volume.device,
                                     reuse_tensor=disparity_img)
        self._path_volume = empty_tensor(height, width, max_disparity, dtype=torch.int8,
                                         device=cost_volume.device, reuse_tensor=self._path_volume)
        self._row_final_costs = empty_tensor(height, max_disparity, dtype=cost_volume.dtype,
                                             device=cost_volume.device,
                                             reuse_tensor=self._row_final_costs)
        flags = 0
        if self.mode == "cpu_exact":
            flags = STM_DP_CPU_EXACT
            self._path_volume.zero_()
        _DisparityReduceOps.run_dynamic_programming(cost_volume, self._path_volume,
                                                    self._row_final_costs, disparity_img, flags)
        return disparity_img
