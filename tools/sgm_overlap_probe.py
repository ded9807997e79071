"""SGM stage at 1080p: plain fused schedule vs STM_SGM_OVERLAP (developer tool).
    python tools/sgm_overlap_probe.py [D]      (STM_DEBUG_SGM_OVERLAP=x selects the other variant)"""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402
from oracle import synth  # noqa: E402

D = int(sys.argv[1]) if len(sys.argv) > 1 else 128
H, W = 1080, 1920
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
cost = stereomatch.cost.SSD(D)(l, r)
vol = torch.empty_like(cost)
disp = torch.empty(H, W, dtype=torch.int32, device="cuda")


def med(fn, reps=15, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return round(ts[len(ts) // 2], 4)


res = {"D": D, "knob": os.environ.get("STM_DEBUG_SGM_OVERLAP", "")}
maps = {}
for name, ov in (("plain", False), ("overlap", True)):
    sgm = stereomatch.aggregation.Semiglobal(0.1, 0.2, overlap=ov)
    res[name + "_volume_ms"] = med(lambda: sgm(cost, l, vol))
    res[name + "_disparity_only_ms"] = med(lambda: sgm.aggregate_and_reduce(cost, l, vol, disp, keep_volume=False))
    maps[name] = disp.clone()
    from stereomatch._cstereomatch import AggregationOps
    AggregationOps.pass_timing(True)
    sgm(cost, l, vol); sgm(cost, l, vol)
    res[name + "_passes"] = [(d, round(t, 4)) for d, t in AggregationOps.last_pass_ms()]
    sgm.aggregate_and_reduce(cost, l, vol, disp, keep_volume=False); sgm.aggregate_and_reduce(cost, l, vol, disp, keep_volume=False)
    res[name + "_disparity_only_passes"] = [(d, round(t, 4)) for d, t in AggregationOps.last_pass_ms()]
    AggregationOps.pass_timing(False)
res["maps_differ"] = float((maps["plain"] != maps["overlap"]).float().mean())
print(json.dumps(res))
