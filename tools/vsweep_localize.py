"""Which fused-sweep call fails (developer tool)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch, stereomatch
from oracle import synth
H, W, D = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (1080, 1920, 128)
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="unit")
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
cost = stereomatch.cost.SSD(D)(l, r)
fused = stereomatch.aggregation.Semiglobal(fused_sweeps=True)
if len(sys.argv) >= 5 and sys.argv[4] == "plainfirst":
    plain = stereomatch.aggregation.Semiglobal(concurrent_directions=False, fused_sweeps=False)
    want = plain(cost, l).clone(); torch.cuda.synchronize(); print("plain per-direction ok", flush=True)
if len(sys.argv) >= 5 and sys.argv[4] == "newout":
    got = fused(cost, l); torch.cuda.synchronize(); print("fused new out ok", flush=True)
agg = torch.empty_like(cost); disp = torch.empty(H, W, dtype=torch.int32, device="cuda")
import time
for name, fn in (("plain", lambda: fused(cost, l, agg)),
                 ("wta_keep", lambda: fused.aggregate_and_reduce(cost, l, agg, disp)),
                 ("wta_nokeep", lambda: fused.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False)),
                 ("plain_again", lambda: fused(cost, l, agg))):
    for i in range(6):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        print(name, i, f"{(time.perf_counter()-t0)*1e3:.2f} ms", flush=True)
print("back to back x30"); 
for i in range(30): fused.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False)
torch.cuda.synchronize(); print("ok", flush=True)
