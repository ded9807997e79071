"""Random-shape parity of every stage against the oracle (developer tool; the fixed shapes live in tests/).
    python tools/fuzz_parity.py [cases] [seed] [big]
big: mid-size volumes (40 .. 400 MB: the concurrent / fused-sweep / sequential decision of the auto rule).
Small shapes so that the C oracle answers at once; every schedule of the aggregation; prints one line per
failure and a summary, exit code 1 on any failure."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch  # noqa: E402
import stereomatch as sm  # noqa: E402
from oracle import oracle as O, synth  # noqa: E402

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 120
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 7)
DS = [4, 8, 12, 16, 20, 32, 36, 48, 64, 64, 64, 80, 96, 100, 128, 128, 150, 160, 192, 224, 256]
fails = []
BIG = len(sys.argv) > 3 and sys.argv[3] == "big"


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def close(got, want, exact, what, shape):
    got = got.cpu().numpy() if isinstance(got, torch.Tensor) else got
    fin = np.isfinite(want)
    ok = got.shape == want.shape and np.array_equal(np.isfinite(got), fin)
    if ok and exact:
        ok = np.array_equal(got[fin], want[fin])
    elif ok:
        den = np.maximum(np.abs(want[fin]), np.abs(got[fin]))
        err = np.abs(got[fin] - want[fin])
        ok = bool((err <= 1e-4 * den).all())
    if not ok:
        fails.append((what, shape))
        print("FAIL", what, shape, flush=True)


for n in range(cases):
    H, W = int(rng.integers(2, 90)), int(rng.integers(2, 140))
    if n % 7 == 0:
        W = int(rng.integers(140, 700))   # several column tiles / CTAs of the sweeps
    D = int(rng.choice(DS))
    if BIG:
        H, W, D = int(rng.integers(150, 520)), int(rng.integers(250, 800)), int(rng.choice([64, 96, 128, 160, 192]))
        if n % 3 == 0:
            H, W = W, H   # portrait
    k = int(rng.integers(1, 4)) if BIG else int(rng.integers(0, 12))   # (the oracle's SSD is (2k)^2 taps per output)
    scale = "raw" if n % 2 == 0 else "unit"
    shape = (H, W, D, k, scale)
    left, right, _ = synth.random_dot_pair(H, W, D, seed=1000 + n, scale=scale)
    l, r = cuda(left), cuda(right)
    wta, dp = sm.disparity_reduce.WinnerTakesAll(), sm.disparity_reduce.DynamicProgramming()
    # costs
    ossd = O.ssd(left, right, D, k)
    ssd = sm.cost.SSD(D, k)
    vol = ssd(l, r)
    close(vol, ossd, scale == "raw", "ssd", shape)
    obt = O.birchfield(left, right, D)
    close(sm.cost.Birchfield(D)(l, r), obt, True, "birchfield", shape)
    folded = ssd.cost_and_reduce(l, r)
    if folded is not None and not torch.equal(folded, wta(vol)):
        fails.append(("ssd_wta", shape)); print("FAIL ssd_wta", shape, flush=True)
    fb = sm.cost.Birchfield(D).cost_and_reduce(l, r)
    if fb is not None and not np.array_equal(fb.cpu().numpy(), O.wta(obt)):
        fails.append(("bt_wta", shape)); print("FAIL bt_wta", shape, flush=True)
    # reducers on the oracle's volume
    c = cuda(ossd)
    if not np.array_equal(wta(c).cpu().numpy(), O.wta(ossd)):
        fails.append(("wta", shape)); print("FAIL wta", shape, flush=True)
    if not np.array_equal(dp(c).cpu().numpy(), O.dp(ossd, "cuda")):
        fails.append(("dp", shape)); print("FAIL dp", shape, flush=True)
    # aggregation: default (concurrent), sequential, fused sweeps (overlapped and plain), and the folds
    osgm = O.sgm(ossd, left, 0.1, 0.2)
    if BIG:   # what the auto rule picks, volume and map
        auto = sm.aggregation.Semiglobal()
        close(auto(c, l), osgm, False, "sgm_auto", shape)
        v, d_ = auto.aggregate_and_reduce(c, l)
        if not torch.equal(d_, wta(v)):
            fails.append(("sgm_auto_wta", shape)); print("FAIL sgm_auto_wta", shape, flush=True)
        _, d2 = auto.aggregate_and_reduce(c, l, keep_volume=False)
        if float((d2 != d_).float().mean()) > 1e-3:
            fails.append(("sgm_auto_wta_only", shape)); print("FAIL sgm_auto_wta_only", shape, flush=True)
    seq = sm.aggregation.Semiglobal(concurrent_directions=False, fused_sweeps=False)(c, l).clone()
    close(seq, osgm, False, "sgm_sequential", shape)
    conc_op = sm.aggregation.Semiglobal(fused_sweeps=False)
    if not torch.equal(conc_op(c, l), seq):
        fails.append(("sgm_concurrent", shape)); print("FAIL sgm_concurrent", shape, flush=True)
    vol_c, disp_c = conc_op.aggregate_and_reduce(c, l)
    if not (torch.equal(vol_c, seq) and torch.equal(disp_c, wta(seq))):
        fails.append(("sgm_concurrent_wta", shape)); print("FAIL sgm_concurrent_wta", shape, flush=True)
    _, disp_only = conc_op.aggregate_and_reduce(c, l, keep_volume=False)
    if not torch.equal(disp_only, disp_c):
        fails.append(("sgm_concurrent_wta_only", shape)); print("FAIL sgm_concurrent_wta_only", shape, flush=True)
    for name, kw in (("sgm_sweeps", {"fused_sweeps": True}), ("sgm_sweeps_plain", {"fused_sweeps": True, "overlap": False})):
        op = sm.aggregation.Semiglobal(**kw)
        close(op(c, l), osgm, False, name, shape)
        v, d_ = op.aggregate_and_reduce(c, l)
        if not torch.equal(d_, wta(v)):
            fails.append((name + "_wta", shape)); print("FAIL", name + "_wta", shape, flush=True)
    torch.cuda.synchronize()

print(f"{cases} random shapes, {len(fails)} failures")
sys.exit(1 if fails else 0)
