"""Is the small-volume SGM launch-bound?  Times the stage enqueued call by call against the same
calls replayed from a CUDA graph (torch.cuda.CUDAGraph capture on a side stream) -- developer tool.
    python tools/graph_probe.py [H W D]"""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402
from oracle import synth  # noqa: E402

H, W, D = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (375, 450, 64)
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
ssd, sgm = stereomatch.cost.SSD(D), stereomatch.aggregation.Semiglobal()
cost = ssd(l, r)
agg = torch.empty_like(cost)
disp = torch.empty(H, W, dtype=torch.int32, device="cuda")


def timed(fn, reps=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts, hs = [], []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(); fn(); e1.record()
        hs.append((time.perf_counter() - t0) * 1e3)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort(); hs.sort()
    return round(ts[len(ts) // 2], 4), round(hs[len(hs) // 2], 4)


res = {"shape": [H, W, D]}
for name, fn in (("sgm", lambda: sgm(cost, l, agg)),
                 ("sgm_wta", lambda: sgm.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False)),
                 ("ssd", lambda: ssd(l, r, cost))):
    dev_ms, host_ms = timed(fn)
    res[name] = {"device_ms": dev_ms, "host_enqueue_ms": host_ms}
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn(); fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            fn()
        gd, gh = timed(g.replay)
        res[name].update({"graph_device_ms": gd, "graph_host_ms": gh})
    except Exception as e:  # capture may be refused by a call inside the library
        res[name]["graph_error"] = str(e)[:200]
        torch.cuda.synchronize()
pipe = stereomatch.Pipeline(stereomatch.cost.SSD(D), stereomatch.disparity_reduce.WinnerTakesAll(),
                            stereomatch.aggregation.Semiglobal())
gp = stereomatch.streaming.GraphedPipeline(
    stereomatch.Pipeline(stereomatch.cost.SSD(D), stereomatch.disparity_reduce.WinnerTakesAll(),
                         stereomatch.aggregation.Semiglobal()), H, W)
d1, h1 = timed(lambda: pipe.estimate(l, r))
d2, h2 = timed(lambda: gp.estimate(l, r))
res["frame"] = {"device_ms": d1, "host_enqueue_ms": h1, "graph_device_ms": d2, "graph_host_ms": h2}
print(json.dumps(res))
