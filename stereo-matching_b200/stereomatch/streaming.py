"""Host frames in, host disparity maps out, with the PCIe copies overlapped with compute.

New (the reference's video loop, stereomatch/cli_video.py:91-104, uploads, estimates and
downloads one frame after the other on the default stream).  `HostFrameStreamer` wraps a
`Pipeline`: the upload of frame k+1 and the download of frame k-1 run on their own CUDA
streams while frame k is being computed; inputs and results are double-buffered on the
device, so nothing is allocated per frame.  The arithmetic is exactly `Pipeline.estimate`.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch

from .pipeline import Pipeline


class HostFrameStreamer:
    """pipeline: a stereomatch.Pipeline; device: the CUDA device to run on."""

    def __init__(self, pipeline: Pipeline, device="cuda"):
        self.pipeline = pipeline
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("HostFrameStreamer: stereomatch_b200 is CUDA-only")
        self._up = torch.cuda.Stream(self.device)
        self._down = torch.cuda.Stream(self.device)
        self._in: List[Tuple[torch.Tensor, torch.Tensor]] = []
        self._res: List[torch.Tensor] = []
        self._shape = None

    def _buffers(self, left: torch.Tensor) -> None:
        if self._shape == tuple(left.shape):
            return
        self._shape = tuple(left.shape)
        mk = lambda dt: torch.empty(self._shape, dtype=dt, device=self.device)
        self._in = [(mk(torch.float32), mk(torch.float32)) for _ in range(2)]
        self._res = [mk(torch.int32) for _ in range(2)]

    def run(self, pairs: Iterable[Tuple[torch.Tensor, torch.Tensor]],
            outs: Sequence[torch.Tensor]) -> Sequence[torch.Tensor]:
        """pairs: (left, right) float32 HOST tensors [H,W] (pinned memory for real overlap);
        outs[k]: int32 HOST tensor [H,W] receiving frame k's disparity map.  Returns `outs`;
        on return the calling stream has been made to wait for the last download, so a
        `torch.cuda.current_stream().synchronize()` (or a CUDA event recorded after the call)
        covers every result."""
        with torch.cuda.device(self.device):
            compute = torch.cuda.current_stream()
            in_free = [None, None]    # compute finished reading input buffer s
            res_free = [None, None]   # download finished reading result buffer s
            last_down = None
            for k, (left, right) in enumerate(pairs):
                if left.is_cuda or right.is_cuda:
                    raise RuntimeError("HostFrameStreamer.run takes host tensors")
                self._buffers(left)
                s = k & 1
                dl, dr = self._in[s]
                with torch.cuda.stream(self._up):
                    if in_free[s] is not None:
                        self._up.wait_event(in_free[s])
                    dl.copy_(left, non_blocking=True)
                    dr.copy_(right, non_blocking=True)
                    uploaded = torch.cuda.Event()
                    uploaded.record(self._up)
                compute.wait_event(uploaded)
                disp = self.pipeline.estimate(dl, dr)
                in_free[s] = torch.cuda.Event()
                in_free[s].record(compute)
                if res_free[s] is not None:
                    compute.wait_event(res_free[s])
                self._res[s].copy_(disp)  # the pipeline reuses `disp` for the next frame
                computed = torch.cuda.Event()
                computed.record(compute)
                with torch.cuda.stream(self._down):
                    self._down.wait_event(computed)
                    outs[k].copy_(self._res[s], non_blocking=True)
                    res_free[s] = torch.cuda.Event()
                    res_free[s].record(self._down)
                    last_down = res_free[s]
            if last_down is not None:
                compute.wait_event(last_down)
        return outs
