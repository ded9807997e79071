"""SGM + WTA (no volume) under each schedule for a list of shapes: where do the auto rules sit? -- developer tool.
    python tools/sgm_schedule_ab.py "720 1280 64" "480 640 128" ..."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402
from oracle import synth  # noqa: E402


def med(fn, reps=7, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return round(sorted(ts)[len(ts) // 2], 4)


for arg in sys.argv[1:]:
    H, W, D = (int(v) for v in arg.split())
    left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
    l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    cost = stereomatch.cost.SSD(D)(l, r)
    agg = torch.empty_like(cost)
    disp = torch.empty(H, W, dtype=torch.int32, device="cuda")
    rec = {"shape": [H, W, D], "W*D/128": round(W * D / 128, 1)}
    for name, kw in (("auto", {}), ("sweeps", {"fused_sweeps": True}), ("sweeps_no_overlap", {"fused_sweeps": True, "overlap": False}),
                     ("directions", {"fused_sweeps": False}), ("sequential", {"fused_sweeps": False, "concurrent_directions": False})):
        op = stereomatch.aggregation.Semiglobal(**kw)
        rec[name] = med(lambda: op.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False))
    print(json.dumps(rec), flush=True)
    del cost, agg
    torch.cuda.empty_cache()
