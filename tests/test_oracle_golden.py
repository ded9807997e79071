"""The oracle (oracle/stm_oracle.c) against the committed golden vectors, which are outputs
of the UNMODIFIED compiled reference CPU ops (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle import oracle as O

COST_CASES = ["cost_raw_24x40x16", "cost_unit_40x24x16", "cost_unit_20x72x32"]
SGM_CASES = ["sgm_unit_18x30x16", "sgm_unit_30x18x16", "sgm_raw_16x16x32",
             "sgm_raw_12x40x64", "sgm_minlast_14x22x8"]
DP_CASES = ["dp_raw_32x48x16", "dp_raw_32x32x32", "dp_ties_16x20x8"]


@pytest.mark.parametrize("case", COST_CASES)
def test_ssd_bit_exact(golden, case):
    g = golden(case)
    D = g["ssd7"].shape[2]
    assert np.array_equal(O.ssd(g["left"], g["right"], D, 7), g["ssd7"])
    assert np.array_equal(O.ssd(g["left"], g["right"], D, 3), g["ssd3"])


@pytest.mark.parametrize("case", COST_CASES)
def test_birchfield_cpu_functor_and_shift_relation(golden, case):
    g = golden(case)
    ref = g["bt_cpu"]  # NaN where the reference never wrote (d > x)
    H, W, D = ref.shape
    mine = O.birchfield(g["left"], g["right"], D, fill_inf=False, cpu_functor=True)
    assert np.array_equal(np.isnan(mine), np.isnan(ref))
    # the reference CPU functor reads out of bounds on the last row / right border
    m = ~np.isnan(ref)
    m[-1] = False
    m[:, W - 6:] = False
    assert np.array_equal(mine[m], ref[m])
    # canonical (reference CUDA functor) semantics == CPU functor shifted by one pixel,
    # wherever both windows are untruncated on the left: d <= x-4, x+5 <= W-1
    canon = O.birchfield(g["left"], g["right"], D)
    for x in range(4, W - 6):
        dm = min(D, x - 3)
        assert np.array_equal(ref[:-1, x, :dm], canon[:-1, x + 1, :dm])
    # BT-1: d > x is +inf
    for x in range(min(W, D - 1)):
        assert np.all(np.isposinf(canon[:, x, x + 1:]))


@pytest.mark.parametrize("case", SGM_CASES)
def test_sgm_bit_exact(golden, case):
    g = golden(case)
    agg = O.sgm(g["cost"], g["left"], float(g["p"][0]), float(g["p"][1]))
    assert np.array_equal(agg, g["agg"])
    assert np.array_equal(O.wta(agg), g["wta"])


def test_sgm_min_quirk_matters(golden):
    """K={0..D-2} (compiled CPU reference) differs from K={0..D-1} on this volume."""
    g = golden("sgm_minlast_14x22x8")
    a = O.sgm(g["cost"], g["left"], float(g["p"][0]), float(g["p"][1]), min_over_all_d=True)
    assert not np.array_equal(a, g["agg"])


@pytest.mark.parametrize("case", ["wta_ties_16x24x32", "wta_rand_9x13x192"])
def test_wta_bit_exact(golden, case):
    g = golden(case)
    assert np.array_equal(O.wta(g["cost"]), g["wta"])


@pytest.mark.parametrize("case", DP_CASES)
def test_dp_cpu_exact(golden, case):
    g = golden(case)
    disp, path, row_final = O.dp(g["cost"], "cpu_exact", return_aux=True)
    assert np.array_equal(disp, g["disp"])
    assert np.array_equal(path, g["path"])
    assert np.array_equal(row_final, g["row_final"])


def test_dp_cpu_exact_rejects_out_of_bounds_shapes():
    with pytest.raises(ValueError):
        O.dp(np.zeros((4, 8, 16), np.float32), "cpu_exact")


def test_path_set_multiplicities():
    """Closed form of the reference path set: landscape, portrait, square."""
    for H, W in [(5, 8), (8, 5), (6, 6), (3, 11), (1, 4), (4, 1)]:
        cnt = O.sgm_visit_counts(H, W)
        assert np.all(cnt[[0, 1, 2, 3, 4, 6]] == 1)  # -> <- v ^ \v /v : every pixel once
        yy, xx = np.mgrid[0:H, 0:W]
        c = xx - yy
        s = xx + yy
        cs = W - H
        exp5 = np.ones((H, W), int)
        exp7 = np.ones((H, W), int)
        if W > H:
            exp5[(c >= 0) & (c < cs)] = 0
            exp5[c == cs] = cs + 1
            exp7[(s >= H) & (s <= W - 1)] = 0
            exp7[s == H - 1] = cs + 1
        elif H > W:
            exp5[(c > cs) & (c <= -1)] = 0
            exp5[c == cs] = -cs
        assert np.array_equal(cnt[5], exp5), (H, W)
        assert np.array_equal(cnt[7], exp7), (H, W)


@pytest.mark.parametrize("H,W,D,k", [(37, 53, 32, 7), (20, 45, 64, 3), (9, 11, 5, 2), (16, 33, 32, 0),
                                      (30, 20, 40, 7), (5, 3, 8, 7)])
def test_fast_oracle_forms_equal_the_literal_restatement(H, W, D, k):
    """The forms the full-size GPU tests use (tests/test_gpu_named_shapes.py) are the pinned
    restatement, only faster: row chunks on threads (bit-identical by construction: output
    row y reads image rows [y-k, y+k) only) and, for integer-valued images, int64 integral
    images (bit-identical because every fp32 partial sum is an exact integer < 2^24)."""
    from oracle import synth
    l, r, _ = synth.random_dot_pair(H, W, D, seed=H + W, scale="raw")
    want = O.ssd(l, r, D, k)
    assert np.array_equal(O.ssd_exact_int(l, r, D, k, threads=3), want)
    assert np.array_equal(O.ssd_threaded(l, r, D, k, threads=3), want)
    lu, ru, _ = synth.random_dot_pair(H, W, D, seed=H + W, scale="unit")
    assert np.array_equal(O.ssd_threaded(lu, ru, D, k, threads=5), O.ssd(lu, ru, D, k))
    assert np.array_equal(O.birchfield_threaded(lu, ru, D, threads=5), O.birchfield(lu, ru, D))
    with pytest.raises(ValueError):
        O.ssd_exact_int(lu, ru, D, k)           # not integer-valued


def test_fast_ssd_equals_the_golden_reference_output(golden):
    g = golden("cost_raw_24x40x16")
    assert np.array_equal(O.ssd_exact_int(g["left"], g["right"], 16, 7), g["ssd7"])
    assert np.array_equal(O.ssd_exact_int(g["left"], g["right"], 16, 3), g["ssd3"])
