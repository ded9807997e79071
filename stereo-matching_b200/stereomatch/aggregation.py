"""Cost-volume aggregation (API of reference stereomatch/aggregation.py)."""
from typing import Optional

import torch

from ._common import empty_tensor
from ._cstereomatch import (STM_SGM_MIN_ALL_D, STM_SGM_OVERWRITE, STM_SGM_FUSED_SWEEPS,
                             STM_SGM_DISPARITY_ONLY, STM_SGM_SEQUENTIAL, STM_SGM_OVERLAP)
from ._cstereomatch import AggregationOps as _AggregationOps


class Semiglobal:
    """Semiglobal matching aggregation over the reference's path set: 8 directions with
    the adaptive second penalty max(P1, P2/|dI|) (reference stereomatch/aggregation.py:13-57,
    src/semiglobal.cpp:116-159).

    penalty1 / penalty2 are public and mutable as in the reference.  Keyword-only extra:
    min_over_all_d=False keeps the compiled CPU reference's min over d in [0, D-2]
    (SURVEY F6); True takes the min over all D like the reference CUDA kernel.
    Unlike the reference's CUDA branch, D does not have to be a power of two.
    Three schedules, same per-direction values:
      * one kernel per direction, the eight directions added into the output in the reference's
        order (bit-identical to the compiled CPU reference for W >= H);
      * concurrent_directions=True (default): volumes that are latency-bound
        one pass after the other (<= 48 MB always; up to 384 MB when the fused sweeps do not
        apply, e.g. 375x450x128, VGA, 720p at D=64) run the eight directions at the same time
        into per-direction buffers that are then added in the reference's order -- bit-identical
        to the sequential passes, which concurrent_directions=False forces.  Costs eight
        volume-sized buffers of workspace;
      * fused_sweeps: the two horizontal passes, then the six directions with a vertical
        component as two fused sweeps (csrc/sgm_vsweep.cu): four passes over the volumes instead
        of eight.  The sum is associated per sweep (0, 1, 2+4+6, 3+5+7) -- within 1e-4 of the
        reference order (measured 1.3e-5 at 1080p), not bit-identical.  fused_sweeps=None
        (default) uses them for volumes above 48 MB where a cost model of both schedules says
        they are faster (about W*D >= 600*128, more for D > 128 and for few rows), True always,
        False never (reference order).
        Shapes the sweeps do not handle (more than 13 columns per SM -- 26 at D <= 64 --; D above
        160 other than 192 and 256, odd D between 33 and 63, D between 97 and 127 that is not a
        multiple of 4) take the per-direction passes.
      * overlap (with the fused sweeps; default on): the two horizontal passes run on a helper
        stream next to the down sweep, which writes a workspace volume of its own that the up
        sweep adds (STM_SGM_OVERLAP).  Costs one more volume of workspace (1.06 GB at 1080p
        D=128); overlap=False keeps the four passes one after the other.
    """

    def __init__(self, penalty1: float = 0.1, penalty2: float = 0.2, *,
                 min_over_all_d: bool = False, fused_sweeps: Optional[bool] = None,
                 concurrent_directions: bool = True, overlap: bool = True):
        self.penalty1 = penalty1
        self.penalty2 = penalty2
        self.min_over_all_d = min_over_all_d
        self.fused_sweeps = fused_sweeps
        self.concurrent_directions = concurrent_directions
        self.overlap = overlap

    def _use_fused_sweeps(self, cost_volume: torch.Tensor) -> bool:
        if self.fused_sweeps is None:
            # auto: a cost model fitted to tools/sgm_schedule_ab.py (ms, sweeps vs directions):
            #   720x1280x64 0.93 / 1.20   480x640x128 0.73 / 0.79   400x700x128 0.65 / 0.67
            #   400x600x128 0.62 / 0.59   500x560x192 1.10 / 1.08   375x450x192 1.04 / 0.68
            #   300x500x256 0.89 / 0.74   1080x1920x32 1.73 / 1.60  375x450x128 0.72 / 0.47
            # The two sweeps cost a row period per image row each -- ~0.68 us up to 128 disparities
            # (hand-off latency), proportionally more above -- plus ~0.1 ms of fixed cost; the eight
            # directions (concurrent up to 384 MB, else one after the other) cost ~19.5 ps per
            # pixel-disparity.  Below 48 MB the concurrent directions win anyway.
            H, W, D = cost_volume.shape
            if cost_volume.numel() * cost_volume.element_size() <= (48 << 20):
                return False
            sweeps_us = 2.0 * H * 0.68 * max(1.0, D / 128.0) + 100.0
            directions_us = 19.5e-6 * H * W * D
            return sweeps_us < directions_us
        return bool(self.fused_sweeps)

    def __call__(self, cost_volume: torch.Tensor, left_image: torch.Tensor,
                 sga_volume: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[H,W,D] cost volume + [H,W] left image -> [H,W,D] aggregated volume.  The result
        equals the reference's zero-fill followed by accumulation; the zero fill itself
        is skipped (the first direction overwrites)."""
        sga_volume = empty_tensor(*cost_volume.size(), dtype=cost_volume.dtype,
                                  device=cost_volume.device, reuse_tensor=sga_volume)
        flags = STM_SGM_OVERWRITE | (STM_SGM_MIN_ALL_D if self.min_over_all_d else 0)
        if not self.concurrent_directions:
            flags |= STM_SGM_SEQUENTIAL
        if self._use_fused_sweeps(cost_volume):
            flags |= STM_SGM_FUSED_SWEEPS | (STM_SGM_OVERLAP if self.overlap else 0)
        _AggregationOps.run_semiglobal(cost_volume, left_image, self.penalty1, self.penalty2,
                                       sga_volume, flags)
        return sga_volume

    def aggregate_and_reduce(self, cost_volume: torch.Tensor, left_image: torch.Tensor,
                             sga_volume: Optional[torch.Tensor] = None,
                             disparity_img: Optional[torch.Tensor] = None, keep_volume: bool = True):
        """New: aggregation followed by winner-takes-all in one native call; returns
        (sga_volume, disparity_img).  The argmin is taken inside the last aggregation pass, so
        the aggregated volume is not re-read; that pass must visit every pixel once, hence the
        last two directions run in swapped order and sga_volume equals __call__'s result up to
        fp32 association.  The disparity map is exactly WinnerTakesAll()(sga_volume).
        keep_volume=False: sga_volume is only scratch (the last pass does not write the sums
        back, 4 B/pix-disp less traffic); the returned volume is then NOT the aggregation."""
        sga_volume = empty_tensor(*cost_volume.size(), dtype=cost_volume.dtype,
                                  device=cost_volume.device, reuse_tensor=sga_volume)
        disparity_img = empty_tensor(cost_volume.size(0), cost_volume.size(1), dtype=torch.int32,
                                     device=cost_volume.device, reuse_tensor=disparity_img)
        flags = STM_SGM_OVERWRITE | (STM_SGM_MIN_ALL_D if self.min_over_all_d else 0)
        if not keep_volume:
            flags |= STM_SGM_DISPARITY_ONLY
        if not self.concurrent_directions:
            flags |= STM_SGM_SEQUENTIAL
        if self._use_fused_sweeps(cost_volume):
            flags |= STM_SGM_FUSED_SWEEPS | (STM_SGM_OVERLAP if self.overlap else 0)
        _AggregationOps.run_semiglobal_wta(cost_volume, left_image, self.penalty1, self.penalty2,
                                           sga_volume, disparity_img, flags)
        return sga_volume, disparity_img
