"""Generates tests/golden/*.npz from the UNMODIFIED compiled reference (oracle/_ref).

Run in the build container (needs /root/reference built by oracle/build_ref.py):

    python tests/golden/make_golden.py

Every file stores the inputs and the outputs of the reference's own CPU ops
(CostOps / AggregationOps / DisparityReduceOps on CPU tensors).  The fixtures travel to
the GPU box, where /root/reference does not exist.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import ref_ops as R  # noqa: E402
from oracle import synth  # noqa: E402

OUT = Path(__file__).resolve().parent


def save(name, **arrays):
    np.savez_compressed(OUT / f"{name}.npz", **arrays)
    print(name, {k: v.shape for k, v in arrays.items()})


def main():
    assert R.available(), "build the reference first: python oracle/build_ref.py"

    # --- cost builders: landscape raw scale, portrait unit scale --------------------
    for name, (H, W, D, scale, seed) in {
        "cost_raw_24x40x16": (24, 40, 16, "raw", 1),
        "cost_unit_40x24x16": (40, 24, 16, "unit", 2),
        "cost_unit_20x72x32": (20, 72, 32, "unit", 3),
    }.items():
        left, right, _ = synth.random_dot_pair(H, W, D, seed=seed, scale=scale)
        ssd7 = R.ssd(left, right, D, 7)
        ssd3 = R.ssd(left, right, D, 3)
        bt_cpu = R.birchfield(left, right, D)  # reference CPU functor, NaN = never written
        save(name, left=left, right=right, ssd7=ssd7, ssd3=ssd3, bt_cpu=bt_cpu)

    # --- SGM: landscape / portrait / square, default penalties and a large-penalty run --
    for name, (H, W, D, scale, seed, p1, p2) in {
        "sgm_unit_18x30x16": (18, 30, 16, "unit", 4, 0.1, 0.2),
        "sgm_unit_30x18x16": (30, 18, 16, "unit", 5, 0.1, 0.2),
        "sgm_raw_16x16x32": (16, 16, 32, "raw", 6, 30.0, 120.0),
        "sgm_raw_12x40x64": (12, 40, 64, "raw", 7, 10.0, 50.0),
    }.items():
        left, right, _ = synth.random_dot_pair(H, W, D, seed=seed, scale=scale)
        cost = R.ssd(left, right, D, 3)
        agg = R.sgm(cost, left, p1, p2)
        save(name, left=left, cost=cost, agg=agg, p=np.array([p1, p2], np.float32),
             wta=R.wta(agg))

    # SGM on a dense random volume whose running minimum often sits at d = D-1
    H, W, D = 14, 22, 8
    vol = synth.random_volume(H, W, D, seed=8)
    vol[:, :, D - 1] *= 0.05
    img = synth.random_volume(H, W, 1, seed=9, hi=255.0)[:, :, 0].round()
    save("sgm_minlast_14x22x8", left=img, cost=vol, agg=R.sgm(vol, img, 2.0, 9.0),
         p=np.array([2.0, 9.0], np.float32), wta=R.wta(R.sgm(vol, img, 2.0, 9.0)))

    # --- WTA: ties, inf, NaN -------------------------------------------------------
    vol = synth.tie_heavy_volume(16, 24, 32, seed=10)
    vol[0, 0, :] = np.inf
    vol[0, 1, 0] = np.nan
    vol[0, 2, 5] = np.nan
    vol[0, 3, :] = np.nan
    vol[1, :, :8] = np.inf
    save("wta_ties_16x24x32", cost=vol, wta=R.wta(vol))
    vol = synth.random_volume(9, 13, 192, seed=11)
    save("wta_rand_9x13x192", cost=vol, wta=R.wta(vol))

    # --- DP (reference CPU functor, path volume pre-zeroed, D <= H <= W) -------------
    for name, (H, W, D, seed) in {"dp_raw_32x48x16": (32, 48, 16, 12),
                                  "dp_raw_32x32x32": (32, 32, 32, 13)}.items():
        left, right, _ = synth.random_dot_pair(H, W, D, seed=seed, scale="raw")
        cost = R.ssd(left, right, D, 3)
        disp, path, row_final = R.dp(cost, return_aux=True)
        save(name, cost=cost, disp=disp, path=path, row_final=row_final)
    vol = synth.tie_heavy_volume(16, 20, 8, seed=14)
    disp, path, row_final = R.dp(vol, return_aux=True)
    save("dp_ties_16x20x8", cost=vol, disp=disp, path=path, row_final=row_final)


if __name__ == "__main__":
    main()
