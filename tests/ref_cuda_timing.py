"""Secondary baseline for BASELINE.json config 2: the reference's own CUDA branch (the sm_100a
cubins inside oracle/_ref/_cstereomatch.so, built unmodified from /root/reference) against this
library on the same GPU and the same inputs -- 450x375, D=64 (power-of-two D and W <= 1024, the
only shapes the reference CUDA kernels accept).  NOT collected by pytest (the reference CUDA
kernels race, SURVEY.md F7, and are not a checker): run by hand on a GPU box,

    python tests/ref_cuda_timing.py > gpurun_out/ref_cuda_timing.json

Test infrastructure: this is one of the places allowed to execute oracle/.
"""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))

import numpy as np  # noqa: E402
import torch  # noqa: E402

import stereomatch as sm  # noqa: E402
from oracle import ref_ops, synth  # noqa: E402


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()   # the reference's Birchfield/SGM/WTA launches are async
        ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort()
    return round(ts[len(ts) // 2], 4)


def main():
    H, W, D = 375, 450, 64
    ref = ref_ops.load()
    if ref is None:
        print(json.dumps({"unavailable": "oracle/_ref/_cstereomatch.so not built"}))
        return 0
    left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
    l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    out = {"shape": [H, W, D], "unit": "ms, wall clock around a synchronised call, median of 20"}

    # --- reference CUDA branch (its Python wrappers only allocate; the ops are called directly)
    rc = torch.empty(H, W, D, device="cuda")
    ra = torch.zeros(H, W, D, device="cuda")
    rd = torch.empty(H, W, dtype=torch.int32, device="cuda")
    ref_t = {}
    ref_t["ssd"] = timed(lambda: ref.CostOps.compute_ssd(l, r, rc, 7))
    ref.CostOps.compute_ssd(l, r, rc, 7)
    def ref_sgm():
        ra.zero_()
        ref.AggregationOps.run_semiglobal(rc, l, 0.1, 0.2, ra)
    ref_t["sgm"] = timed(ref_sgm)
    ref_t["wta"] = timed(lambda: ref.DisparityReduceOps.run_winners_take_all(ra, rd))
    ref_t["ssd+sgm+wta"] = round(sum(ref_t.values()), 4)
    out["reference_cuda"] = ref_t

    # --- this library, same stages through the reference-shaped API
    ssd, sgm, wta = sm.cost.SSD(D), sm.aggregation.Semiglobal(), sm.disparity_reduce.WinnerTakesAll()
    c = ssd(l, r)
    a = sgm(c, l)
    d = wta(a)
    ours = {"ssd": timed(lambda: ssd(l, r, c)), "sgm": timed(lambda: sgm(c, l, a)),
            "wta": timed(lambda: wta(a, d))}
    ours["ssd+sgm+wta"] = round(sum(ours.values()), 4)
    pipe = sm.Pipeline(ssd, wta, sgm)
    ours["Pipeline.estimate"] = timed(lambda: pipe.estimate(l, r))
    out["stereomatch_b200"] = ours
    out["speedup"] = {k: round(ref_t[k] / ours[k], 2) for k in ref_t}

    # --- agreement (informative: the reference CUDA SGM races, its WTA breaks ties differently)
    ssd(l, r, c)
    ref.CostOps.compute_ssd(l, r, rc, 7)
    torch.cuda.synchronize()
    fin = torch.isfinite(rc) & torch.isfinite(c)
    out["ssd_max_rel_diff_vs_reference_cuda"] = float(((rc - c).abs()[fin] / rc.abs()[fin].clamp_min(1e-6)).max())
    out["ssd_same_inf_pattern"] = bool(torch.equal(torch.isinf(rc), torch.isinf(c)))
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
