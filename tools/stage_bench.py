"""Per-stage device times (CUDA events, L2-exceeding volumes) -- developer tool.
    python tools/stage_bench.py [H W D] [--stages ssd,bt,sgm,wta,dp,sgm_wta,ssd_wta,bt_wta] [--reps N]
sgm_wta = aggregation with the argmin folded in and no volume stored: what Pipeline.estimate runs."""
import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402
from oracle import synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("shape", nargs="*", type=int, default=[1080, 1920, 128])
ap.add_argument("--stages", default="ssd,bt,sgm,wta,dp,sgm_wta")
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--warm", type=int, default=3)
a = ap.parse_args()
H, W, D = a.shape
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
N = H * W * D
import json as _json
try:
    peak = float(_json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
except Exception:
    peak = 6573.2
algo = {"ssd": 4 * N + 8 * H * W, "bt": 4 * N + 8 * H * W, "sgm": 8 * N + 4 * H * W,
        "wta": 4 * N + 4 * H * W, "dp": 5 * N + 5 * H * W + 8 * H * D, "sgm_wta": 4 * N + 8 * H * W,
        "ssd_wta": 28 * H * W, "bt_wta": 12 * H * W}
ssd, bt = stereomatch.cost.SSD(D), stereomatch.cost.Birchfield(D)
sgm, wta = stereomatch.aggregation.Semiglobal(), stereomatch.disparity_reduce.WinnerTakesAll()
dp = stereomatch.disparity_reduce.DynamicProgramming()
cost = ssd(l, r)
agg = torch.empty_like(cost)
disp = torch.empty(H, W, dtype=torch.int32, device="cuda")
fns = {"ssd": lambda: ssd(l, r, cost), "bt": lambda: bt(l, r, agg), "sgm": lambda: sgm(cost, l, agg),
       "wta": lambda: wta(agg, disp), "dp": lambda: dp(cost, disp),
       "sgm_wta": lambda: sgm.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False),
       "ssd_wta": lambda: ssd.cost_and_reduce(l, r, disp), "bt_wta": lambda: bt.cost_and_reduce(l, r, disp)}
res = {}
for name in a.stages.split(","):
    fn = fns[name]
    if name == "wta":
        sgm(cost, l, agg)
    for _ in range(a.warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    med = ts[len(ts) // 2]
    res[name] = {"ms_med": round(med, 4), "ms_min": round(ts[0], 4),
                 "algo_GBs": round(algo[name] / med / 1e6, 1), "frac": round(algo[name] / med / 1e6 / peak, 4)}
    if name == "bt":
        ssd(l, r, cost)
print(json.dumps({"shape": [H, W, D], **res}))
