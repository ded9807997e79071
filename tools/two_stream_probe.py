"""Does running consecutive frames on two streams (two Pipelines, two sets of volumes) raise
whole-job throughput?  SSD is issue-bound, the SGM passes are HBM-bound.  Developer probe."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch as sm  # noqa: E402
from oracle import synth  # noqa: E402

H, W, D, F = 1080, 1920, 128, 8
frames = []
for f in range(F):
    l, r, _ = synth.random_dot_pair(H, W, D, seed=1000 + f, scale="raw")
    frames.append((torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()))


def mk():
    return sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal())


def run(nstreams, reps=5):
    pipes = [mk() for _ in range(nstreams)]
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    main = torch.cuda.current_stream()

    def step():
        start = torch.cuda.Event(); start.record(main)
        for s in streams:
            s.wait_event(start)
        for f, (l, r) in enumerate(frames):
            k = f % nstreams
            with torch.cuda.stream(streams[k]):
                pipes[k].estimate(l, r)
        for s in streams:
            e = torch.cuda.Event(); e.record(s); main.wait_event(e)
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        step()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps / F


out = {n: round(run(n), 4) for n in (1, 2, 3)}
print(json.dumps({"ms_per_frame_by_streams": out}))
