"""Frame time of Pipeline(SSD, WinnerTakesAll, Semiglobal).estimate and of its stages over a list of
shapes (regime scan: small / mid-size / large volumes, portrait, odd disparity ranges) -- developer tool.
    python tools/shape_scan.py > profiles/<tag>_shape_scan.jsonl"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402
from oracle import synth  # noqa: E402

SHAPES = [(100, 150, 32), (240, 320, 64), (375, 450, 64), (450, 375, 64), (375, 450, 128), (480, 640, 64),
          (640, 480, 64), (480, 640, 128), (720, 1280, 64), (720, 1280, 128), (1080, 1920, 32), (1080, 1920, 64),
          (1080, 1920, 96), (1080, 1920, 128), (1920, 1080, 128), (1080, 1920, 160), (1080, 1920, 192),
          (1080, 1920, 256), (2160, 3840, 64)]


def med(fn, reps=7, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


for H, W, D in SHAPES:
    left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
    l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    ssd, bt = stereomatch.cost.SSD(D), stereomatch.cost.Birchfield(D)
    sgm, wta = stereomatch.aggregation.Semiglobal(), stereomatch.disparity_reduce.WinnerTakesAll()
    dp = stereomatch.disparity_reduce.DynamicProgramming()
    pipe = stereomatch.Pipeline(ssd, wta, sgm)
    cost = ssd(l, r)
    agg = torch.empty_like(cost)
    disp = torch.empty(H, W, dtype=torch.int32, device="cuda")
    n = H * W * D
    rec = {"shape": [H, W, D], "volume_MB": round(n * 4 / 2**20, 1)}
    rec["ssd_ms"] = round(med(lambda: ssd(l, r, cost)), 4)
    rec["bt_ms"] = round(med(lambda: bt(l, r, agg)), 4)
    ssd(l, r, cost)
    rec["sgm_wta_ms"] = round(med(lambda: sgm.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False)), 4)
    rec["dp_ms"] = round(med(lambda: dp(cost, disp)), 4)
    frame = med(lambda: pipe.estimate(l, r))
    rec["frame_ms"] = round(frame, 4)
    rec["frame_Gpix_disp_s"] = round(n / frame / 1e6, 1)
    print(json.dumps(rec), flush=True)
    del cost, agg, pipe
    torch.cuda.empty_cache()
