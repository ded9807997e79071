"""Fused sweeps: N back-to-back runs at 1080p must be bit-identical (developer tool)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch, stereomatch
from oracle import synth
H, W, D = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (1080, 1920, 128)
N = int(sys.argv[4]) if len(sys.argv) >= 5 else 200
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="unit")
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
cost = stereomatch.cost.SSD(D)(l, r)
fused = stereomatch.aggregation.Semiglobal(fused_sweeps=True)
ref_vol, ref_disp = fused.aggregate_and_reduce(cost, l)
ref_vol, ref_disp = ref_vol.clone(), ref_disp.clone()
vol, disp = torch.empty_like(ref_vol), torch.empty_like(ref_disp)
bad = 0
for i in range(N):
    fused.aggregate_and_reduce(cost, l, vol, disp)
    if i % 10 == 0:
        bad += int(not torch.equal(vol, ref_vol)) + int(not torch.equal(disp, ref_disp))
torch.cuda.synchronize()
print({"shape": [H, W, D], "runs": N, "mismatching_checks": bad})
