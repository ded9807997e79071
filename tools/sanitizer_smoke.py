"""Small invocations of every kernel family, for compute-sanitizer (developer tool)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch, stereomatch as sm
from oracle import synth
for (H, W, D) in [(40, 72, 64), (33, 130, 128), (20, 45, 20)]:
    left, right, _ = synth.random_dot_pair(H, W, D, seed=1, scale="unit")
    l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    for cost_fn in (sm.cost.SSD(D), sm.cost.SSD(D, 3), sm.cost.Birchfield(D)):
        c = cost_fn(l, r)
        for kw in ({}, {"concurrent_directions": False}, {"fused_sweeps": True}, {"fused_sweeps": True, "overlap": False}):
            sg = sm.aggregation.Semiglobal(**kw)
            sg(c, l)
            sg.aggregate_and_reduce(c, l, keep_volume=False)
        sm.disparity_reduce.WinnerTakesAll()(c)
        sm.disparity_reduce.DynamicProgramming()(c)
    if D % 64 == 0:
        sm.cost.Birchfield(D).cost_and_reduce(l, r)
    torch.cuda.synchronize()
print("sanitizer smoke done")
