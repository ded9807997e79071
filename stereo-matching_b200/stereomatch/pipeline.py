"""cost -> (aggregation) -> disparity reduce (API of reference stereomatch/pipeline.py:36-94)."""
from typing import Callable, Optional

import torch

from .aggregation import Semiglobal
from .cost import SSD, Birchfield, SSDTexture
from .cuda_texture import CUDATexture
from .disparity_reduce import WinnerTakesAll

CostFunction = Callable[[torch.Tensor, torch.Tensor, torch.Tensor], torch.Tensor]
AggregationFunction = Callable[[torch.Tensor, torch.Tensor, torch.Tensor], torch.Tensor]
DisparityReduceFunction = Callable[[torch.Tensor, torch.Tensor], torch.Tensor]


class _TexCostFunctionWrapper:
    """Lets Pipeline.estimate take tensors when the cost function wants textures
    (reference stereomatch/pipeline.py:22-33): both images are loaded into new textures per call."""

    def __init__(self, cost_function):
        self.cost_function = cost_function

    def __call__(self, left_image: torch.Tensor, right_image: torch.Tensor, cost_volume: torch.Tensor):
        return self.cost_function(CUDATexture.from_tensor(left_image.cuda()),
                                  CUDATexture.from_tensor(right_image.cuda()),
                                  cost_volume=cost_volume)


class Pipeline:
    """Composes the three stages and keeps their output buffers between frames, so a
    steady-state video loop allocates nothing (reference pipeline.py:65-67, 80-92)."""

    def __init__(self, cost: CostFunction, disparity_reduce: DisparityReduceFunction,
                 aggregation: Optional[AggregationFunction] = None):
        self.cost = _TexCostFunctionWrapper(cost) if isinstance(cost, SSDTexture) else cost
        self.disparity_reduce = disparity_reduce
        self.aggregation = aggregation
        self._cost_volume = None
        self._aggregation_volume = None
        self._disparity_image = None
        # New: Semiglobal followed by WinnerTakesAll runs as one native call (the argmin is
        # folded into the last aggregation pass).  Set to False for two separate stages.
        self.fuse_aggregation_and_reduce = True
        # New: Birchfield followed directly by WinnerTakesAll (no aggregation) runs as one kernel that
        # never writes the cost volume (`_cost_volume` stays None then).  Set to False for two stages.
        self.fuse_cost_and_reduce = True
        # The aggregated volume never leaves the pipeline, so the fused call may skip writing it
        # (its buffer is scratch then).  Set to True to find the aggregation in
        # `_aggregation_volume` after estimate().
        self.keep_aggregation_volume = False

    def _fold_pays(self, left_image: torch.Tensor) -> bool:
        """Birchfield + WTA in one kernel wins at every size.  SSD + WTA pays for its atomics and
        shuffles: when the volume would go through HBM it takes the time of the two stages
        (1080p D=128: 0.43 ms against 0.42-0.45, measured before and after the SSD kernel got
        faster) and needs no 1 GB volume; while the volume stays in L2 it loses (450x375 D=64:
        0.053 -> 0.064 ms).  So the fold is used above 48 MB -- the threshold of the other size
        rules (aggregation.py) -- for its memory, not its speed.
        fuse_cost_and_reduce = "always" forces it."""
        if type(self.cost) is not SSD or self.fuse_cost_and_reduce == "always":
            return True
        return left_image.numel() * self.cost.max_disparity * 4 > (48 << 20)

    def estimate(self, left_image: torch.Tensor, right_image: torch.Tensor,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[H,W] left/right images (CUDA float32) -> int32 [H,W] disparity map.

        out (new, optional): an int32 [H,W] tensor on the images' device that receives the map
        instead of the pipeline's own reused buffer (callers that keep several frames in flight,
        stereomatch.streaming); it is not retained."""
        if out is not None:
            keep = self._disparity_image
            self._disparity_image = out
            try:
                return self.estimate(left_image, right_image)
            finally:
                self._disparity_image = keep
        if (self.fuse_cost_and_reduce and self.aggregation is None and type(self.cost) in (Birchfield, SSD)
                and type(self.disparity_reduce) is WinnerTakesAll
                and self._fold_pays(left_image)):
            # cost + reduce in one pass, the volume never materialised (stm_birchfield_wta_f32,
            # stm_ssd_wta_f32)
            fused = self.cost.cost_and_reduce(left_image, right_image, self._disparity_image)
            if fused is not None:
                self._disparity_image = fused
                return fused
        self._cost_volume = self.cost(left_image, right_image, cost_volume=self._cost_volume)
        if self.aggregation is not None:
            own = self._aggregation_volume
            if own is self._cost_volume:  # left over from a run without aggregation
                own = None
            if (self.fuse_aggregation_and_reduce and type(self.disparity_reduce) is WinnerTakesAll
                    and type(self.aggregation) is Semiglobal):
                self._aggregation_volume, self._disparity_image = self.aggregation.aggregate_and_reduce(
                    self._cost_volume, left_image.to(self._cost_volume.device), own,
                    self._disparity_image, keep_volume=self.keep_aggregation_volume)
                return self._disparity_image
            self._aggregation_volume = self.aggregation(
                self._cost_volume, left_image.to(self._cost_volume.device), own)
        else:
            self._aggregation_volume = self._cost_volume
        self._disparity_image = self.disparity_reduce(self._aggregation_volume,
                                                      self._disparity_image)
        return self._disparity_image
