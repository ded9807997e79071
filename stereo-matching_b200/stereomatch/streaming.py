"""Batches of frames: several frames in flight, PCIe copies overlapped with compute.

New (the reference's video loop, stereomatch/cli_video.py:91-104, uploads, estimates and
downloads one frame after the other on the default stream).  Two helpers, both running exactly
`Pipeline.estimate` per frame:

* `BatchEstimator` -- device-resident pairs in, disparity maps out, consecutive frames
  round-robin over several `Pipeline`s, each on its own CUDA stream.  The cost stage is
  issue-bound and the aggregation passes are HBM-bound, so two frames in flight overlap them:
  3.81 instead of 4.00 ms per 1080p D=128 frame on a B200 (a third stream adds nothing).
* `HostFrameStreamer` -- host frames in, host maps out: the upload of later frames and the
  download of earlier ones run on their own streams while the compute streams work; inputs and
  results are buffered on the device, so nothing is allocated per frame.

* `GraphedPipeline` -- one `Pipeline.estimate` of a fixed shape captured into a CUDA graph: a small
  frame costs the host ~25 driver calls (0.116 ms at 450x375 D=64), the replay is one (0.035 ms).

Every pipeline owns its volumes (2 x 1.06 GB at 1080p D=128), so a second one doubles that.
Work is ordered after the caller's current stream, and on return that stream has been made to
wait for everything issued here.
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence, Tuple

import torch

from .pipeline import Pipeline


def _as_list(pipelines) -> List[Pipeline]:
    if isinstance(pipelines, Pipeline):
        return [pipelines]
    out = list(pipelines)
    if not out or not all(isinstance(p, Pipeline) for p in out):
        raise TypeError("expected a stereomatch.Pipeline or a non-empty sequence of them")
    if len(set(id(p) for p in out)) != len(out):
        raise ValueError("the pipelines must be distinct objects (each owns its volumes)")
    return out


class BatchEstimator:
    """pipelines: one Pipeline, or several independent ones (one compute stream each)."""

    def __init__(self, pipelines, device="cuda"):
        self.pipelines = _as_list(pipelines)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("BatchEstimator: stereomatch_b200 is CUDA-only")
        self._streams = [torch.cuda.Stream(self.device) for _ in self.pipelines]

    def estimate_batch(self, pairs: Sequence[Tuple[torch.Tensor, torch.Tensor]],
                       outs: Optional[Sequence[torch.Tensor]] = None) -> List[torch.Tensor]:
        """pairs: (left, right) CUDA float32 [H,W]; outs[k] (optional): CUDA int32 [H,W] that
        receives frame k's map.  Returns the list of maps (outs, or freshly allocated)."""
        n = len(self.pipelines)
        results: List[torch.Tensor] = []
        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream()
            start = torch.cuda.Event()
            start.record(main)
            for s in self._streams:
                s.wait_event(start)
            for k, (left, right) in enumerate(pairs):
                pipe, stream = self.pipelines[k % n], self._streams[k % n]
                with torch.cuda.stream(stream):
                    # frames k and k+n must not share a result buffer: the caller's tensor or a new one
                    out = outs[k] if outs is not None else torch.empty(
                        tuple(left.shape), dtype=torch.int32, device=self.device)
                    results.append(pipe.estimate(left, right, out=out))
            for s in self._streams:
                done = torch.cuda.Event()
                done.record(s)
                main.wait_event(done)
        return results


class GraphedPipeline:
    """`Pipeline.estimate` for ONE image shape, captured once into a CUDA graph and replayed.

    What it buys is HOST time.  At 450x375 D=64 (the size the reference's own tests and benchmarks
    use) `estimate` is ~25 driver calls -- the cost kernel, eight aggregation directions on helper
    streams with their fork / join events, the sum, the argmin -- and costs the calling thread
    0.116 ms per frame; the replay is one call, 0.035 ms (measured, tools/graph_probe.py).  The
    device time of the frame does not change (0.32 ms: the GPU, not the host, is the longer side
    even at this size), so this is for callers whose host thread has other work, or who drive
    several small pipelines from one thread.  At 1080p it is correct and pointless.

    The images are copied into buffers the graph owns (pass `self.left` / `self.right`, filled in
    place, to skip the copies); the map is returned in `self.out` (or copied into `out`).  The
    pipeline's volumes and the stream-keyed workspaces are allocated by two eager warm-up frames on
    the capture stream and must not be used by anything else while the graph is alive.
    """

    def __init__(self, pipeline: Pipeline, height: int, width: int, device="cuda"):
        if not isinstance(pipeline, Pipeline):
            raise TypeError("expected a stereomatch.Pipeline")
        self.pipeline = pipeline
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("GraphedPipeline: stereomatch_b200 is CUDA-only")
        with torch.cuda.device(self.device):
            self.left = torch.zeros(height, width, dtype=torch.float32, device=self.device)
            self.right = torch.zeros(height, width, dtype=torch.float32, device=self.device)
            self.out = torch.empty(height, width, dtype=torch.int32, device=self.device)
            self._stream = torch.cuda.Stream(self.device)
            self._stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._stream):
                for _ in range(2):  # allocations, per-device kernel attributes, lazy module loading
                    pipeline.estimate(self.left, self.right, out=self.out)
            self._stream.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=self._stream):
                pipeline.estimate(self.left, self.right, out=self.out)

    def estimate(self, left_image: torch.Tensor, right_image: torch.Tensor,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Same result as `pipeline.estimate(left_image, right_image)`; enqueued on the current stream."""
        for src, dst in ((left_image, self.left), (right_image, self.right)):
            if src is not dst:
                if tuple(src.shape) != tuple(dst.shape):
                    raise RuntimeError(f"GraphedPipeline: captured for {tuple(dst.shape)}, got {tuple(src.shape)}")
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        if out is not None:
            out.copy_(self.out, non_blocking=True)
            return out
        return self.out


class HostFrameStreamer:
    """pipeline: a stereomatch.Pipeline; extra_pipelines: more independent Pipelines, one more
    compute stream each (one extra is worth ~5 % at 1080p); device: the CUDA device."""

    def __init__(self, pipeline: Pipeline, device="cuda", extra_pipelines: Sequence[Pipeline] = ()):
        self.pipelines = _as_list([pipeline, *extra_pipelines])
        self.pipeline = pipeline
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("HostFrameStreamer: stereomatch_b200 is CUDA-only")
        n = len(self.pipelines)
        self._compute = [torch.cuda.Stream(self.device) for _ in range(n)] if n > 1 else [None]
        self._up = torch.cuda.Stream(self.device)
        self._down = torch.cuda.Stream(self.device)
        self._slots = 2 * n   # device input / result buffers: two per compute stream
        self._in: List[Tuple[torch.Tensor, torch.Tensor]] = []
        self._res: List[torch.Tensor] = []
        self._shape = None

    def _buffers(self, left: torch.Tensor) -> None:
        if self._shape == tuple(left.shape):
            return
        self._shape = tuple(left.shape)
        mk = lambda dt: torch.empty(self._shape, dtype=dt, device=self.device)
        self._in = [(mk(torch.float32), mk(torch.float32)) for _ in range(self._slots)]
        self._res = [mk(torch.int32) for _ in range(self._slots)]

    def run(self, pairs: Iterable[Tuple[torch.Tensor, torch.Tensor]],
            outs: Sequence[torch.Tensor]) -> Sequence[torch.Tensor]:
        """pairs: (left, right) float32 HOST tensors [H,W] (pinned memory for real overlap);
        outs[k]: int32 HOST tensor [H,W] receiving frame k's disparity map.  Returns `outs`;
        on return the calling stream has been made to wait for the last download, so a
        `torch.cuda.current_stream().synchronize()` (or a CUDA event recorded after the call)
        covers every result."""
        n = len(self.pipelines)
        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream()
            compute = [main] if n == 1 else self._compute
            if n > 1:
                start = torch.cuda.Event()
                start.record(main)
                for c in compute:
                    c.wait_event(start)
            in_free = [None] * self._slots    # compute finished reading input slot s
            res_free = [None] * self._slots   # download finished reading result slot s
            downloads = []
            for k, (left, right) in enumerate(pairs):
                if left.is_cuda or right.is_cuda:
                    raise RuntimeError("HostFrameStreamer.run takes host tensors")
                self._buffers(left)
                s = k % self._slots
                pipe, cstream = self.pipelines[k % n], compute[k % n]
                dl, dr = self._in[s]
                with torch.cuda.stream(self._up):
                    if in_free[s] is not None:
                        self._up.wait_event(in_free[s])
                    dl.copy_(left, non_blocking=True)
                    dr.copy_(right, non_blocking=True)
                    uploaded = torch.cuda.Event()
                    uploaded.record(self._up)
                with torch.cuda.stream(cstream):
                    cstream.wait_event(uploaded)
                    if res_free[s] is not None:
                        cstream.wait_event(res_free[s])
                    pipe.estimate(dl, dr, out=self._res[s])   # written in place, no copy
                    in_free[s] = torch.cuda.Event()
                    in_free[s].record(cstream)
                    computed = in_free[s]
                with torch.cuda.stream(self._down):
                    self._down.wait_event(computed)
                    outs[k].copy_(self._res[s], non_blocking=True)
                    res_free[s] = torch.cuda.Event()
                    res_free[s].record(self._down)
                    downloads.append(res_free[s])
            if downloads:
                main.wait_event(downloads[-1])   # the download stream is in order
            if n > 1:
                for c in compute:
                    done = torch.cuda.Event()
                    done.record(c)
                    main.wait_event(done)
        return outs
