"""Birchfield + WTA (no aggregation) at 1080p: two stages vs the fused kernel (developer tool)."""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch, stereomatch
from oracle import synth
D = int(sys.argv[1]) if len(sys.argv) > 1 else 128
H, W = 1080, 1920
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
def med(fn, reps=15, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort(); return round(ts[len(ts) // 2], 4)
fused = stereomatch.Pipeline(stereomatch.cost.Birchfield(D), stereomatch.disparity_reduce.WinnerTakesAll())
plain = stereomatch.Pipeline(stereomatch.cost.Birchfield(D), stereomatch.disparity_reduce.WinnerTakesAll())
plain.fuse_cost_and_reduce = False
res = {"D": D, "two_stage_ms": med(lambda: plain.estimate(l, r)), "fused_ms": med(lambda: fused.estimate(l, r)),
       "equal": bool(torch.equal(plain.estimate(l, r), fused.estimate(l, r)))}
print(json.dumps(res))
