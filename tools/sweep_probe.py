"""Developer probe: time of the fused SGM sweeps vs shape, to separate the per-pixel step time
(one CTA: H <= 4) from the cost of CTA -> CTA hand-offs (H = 8, 16, ...).
    python tools/sweep_probe.py [--shapes H,W,D ...] [--reps N]"""
import argparse, json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shapes", nargs="*", default=["1,30000,128", "4,30000,128", "8,30000,128", "16,30000,128",
                                                "64,30000,128", "256,1920,128", "1080,1920,128"])
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--modes", default="sweep,ref")
a = ap.parse_args()
out = []
for shp in a.shapes:
    H, W, D = map(int, shp.split(","))
    g = torch.Generator(device="cuda").manual_seed(1)
    cost = torch.rand(H, W, D, device="cuda", generator=g) * 100
    img = torch.rand(H, W, device="cuda", generator=g) * 255
    agg = torch.empty_like(cost)
    rec = {"shape": [H, W, D]}
    for mode in a.modes.split(","):
        sgm = stereomatch.aggregation.Semiglobal(fused_sweeps=(mode == "sweep"))
        for _ in range(2):
            sgm(cost, img, agg)
        torch.cuda.synchronize()
        ts = []
        for _ in range(a.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); sgm(cost, img, agg); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = min(ts)
        rec[mode + "_ms"] = round(ms, 4)
        if mode == "sweep":  # two sweeps, each ~ (W + lag*H) steps; lower bound on the step: ms / (2 W)
            rec["ns_per_pixel_step_upper"] = round(ms * 1e6 / (2 * W), 1)
    out.append(rec)
    print(json.dumps(rec), flush=True)
