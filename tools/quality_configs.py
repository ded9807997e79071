"""RMSE / average error / bad-pixel(2.0) of the BASELINE.json configurations on their synthetic
random-dot pairs, scored on the device against the generator's ground truth (developer tool).
No Middlebury data is on this box; these are the numbers the evaluation step
(stereomatch/evaluation.py, reference workflows/evaluation/workflow.py:74-105) would publish."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402
from stereomatch.cli_common import create_pipeline  # noqa: E402
from oracle import synth  # noqa: E402

CONFIGS = [("config 1/2: 450x375 D=64 SSD+WTA", 375, 450, 64, "ssd", "wta", None),
           ("config 2: 450x375 D=64 SSD+SGM+WTA", 375, 450, 64, "ssd", "wta", "sgm"),
           ("config 3: 1920x1080 D=128 Birchfield+SGM+WTA", 1080, 1920, 128, "birchfield", "wta", "sgm"),
           ("config 4: 1920x1080 D=128 SSD+DP", 1080, 1920, 128, "ssd", "dyn", None),
           ("headline: 1920x1080 D=128 SSD+SGM+WTA", 1080, 1920, 128, "ssd", "wta", "sgm"),
           ("config 5 frame: 1920x1080 D=192 SSD+SGM+WTA", 1080, 1920, 192, "ssd", "wta", "sgm")]
for name, H, W, D, cm, dm, am in CONFIGS:
    left, right, gt = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
    l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    disp = create_pipeline(cm, dm, am, D).estimate(l, r)
    gtf = torch.from_numpy(gt).float().cuda()
    # a left pixel has a match only where x >= its disparity; the window adds K = 7 columns
    mask = torch.arange(W, device="cuda")[None, :] >= (torch.from_numpy(gt).cuda() + 7)
    res = stereomatch.metrics.evaluate(gtf, disp, threshold=2.0, mask=mask)
    res["exact"] = float((disp[mask] == torch.from_numpy(gt).cuda()[mask]).float().mean())
    print(json.dumps({"config": name, **{k: (round(v, 4) if isinstance(v, float) else v) for k, v in res.items()}}))
