"""Fused vertical sweeps (csrc/sgm_vsweep.cu) against the per-direction passes: agreement,
stage time and per-pass times -- developer tool.
    python tools/vsweep_probe.py [H W D] [--reps N]"""
import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch  # noqa: E402
from stereomatch._cstereomatch import AggregationOps  # noqa: E402
from oracle import synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("shape", nargs="*", type=int, default=[1080, 1920, 128])
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--scale", default="unit")
a = ap.parse_args()
H, W, D = a.shape
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale=a.scale)
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
cost = stereomatch.cost.SSD(D)(l, r)
plain = stereomatch.aggregation.Semiglobal(concurrent_directions=False, fused_sweeps=False)
fused = stereomatch.aggregation.Semiglobal(fused_sweeps=True)
want = plain(cost, l).clone()
got = fused(cost, l)
torch.cuda.synchronize()
fin = torch.isfinite(want)
same_inf = bool(torch.equal(fin, torch.isfinite(got)))
den = torch.maximum(want[fin].abs(), got[fin].abs()).clamp_min(1e-30)
rel = ((want[fin] - got[fin]).abs() / den).max().item()
wta = stereomatch.disparity_reduce.WinnerTakesAll()
d_plain = wta(want)
_, d_fused = fused.aggregate_and_reduce(cost, l, keep_volume=False)
mism = (d_plain != d_fused).float().mean().item()
res = {"shape": [H, W, D], "same_inf_pattern": same_inf, "max_rel_err": rel,
       "disparity_mismatch_frac": mism}


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return round(ts[len(ts) // 2], 4), round(ts[0], 4)


agg = torch.empty_like(cost)
disp = torch.empty(H, W, dtype=torch.int32, device="cuda")
res["plain_ms"] = timed(lambda: plain(cost, l, agg))
res["fused_ms"] = timed(lambda: fused(cost, l, agg))
res["plain_wta_ms"] = timed(lambda: plain.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False))
res["fused_wta_ms"] = timed(lambda: fused.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False))
AggregationOps.pass_timing(True)
try:
    for name, op in (("plain", plain), ("fused", fused)):
        op.aggregate_and_reduce(cost, l, agg, disp, keep_volume=False)
        res[name + "_passes"] = [(d, round(ms, 4)) for d, ms in AggregationOps.last_pass_ms()]
finally:
    AggregationOps.pass_timing(False)
print(json.dumps(res))

import os  # noqa: E402
if int(os.environ.get("STM_DEBUG_VSWEEP", "0")) & 8:
    import ctypes
    from stereomatch._cstereomatch import _lib
    buf = (ctypes.c_ulonglong * 64)()
    _lib.stm_debug_vsweep_stats(buf)            # clear what the timing loops accumulated
    fused(cost, l, agg)                         # one down + one up sweep
    torch.cuda.synchronize()
    _lib.stm_debug_vsweep_stats(buf)
    v = list(buf)
    ghz = 1.9
    rows = max(v[0], 1)
    print(json.dumps({
        "importer_rows": v[0], "extra_polls_per_row": v[1] / rows,
        "importer_us_per_row_waiting_data": v[2] / rows / ghz / 1e3,
        "importer_us_per_row_waiting_credit": v[3] / rows / ghz / 1e3,
        "reload_rtt_us": (v[4] / max(v[1], 1)) / ghz / 1e3,
        "slot_dep_wait_us_per_row": [round(v[8 + j] / (2 * H) / ghz / 1e3 / max(1, 148), 4) for j in range(13)],
        "slot_credit_wait_us_per_row": [round(v[24 + j] / (2 * H) / ghz / 1e3 / 148, 4) for j in range(13)],
        "xcons_wait_us_per_call": v[40] / max(v[41], 1) / ghz / 1e3, "xcons_calls": v[41],
        "cp_async_wait_us_per_row": v[44] / max(v[43], 1) / H / ghz / 1e3,
        "warp_total_us": v[42] / max(v[43], 1) / ghz / 1e3}))
