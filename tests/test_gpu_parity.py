"""Parity of the CUDA path (through the C ABI / the reference-shaped Python API) against
the oracle and the committed golden vectors.  Run on the B200 box: pytest -m gpu.

Bars (BASELINE.json): WTA / DP maps bit-exact given identical volumes; cost volumes and
SGM sums within 1e-4 relative on finite entries with an identical inf pattern
(RTOL below); end-to-end maps differ at <= 0.1 % of pixels.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as O, synth

pytestmark = pytest.mark.gpu

RTOL = 1e-4  # north_star: "float cost volumes and SGM aggregated costs within 1e-4 relative"


@pytest.fixture(scope="module")
def sm():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import stereomatch
    return stereomatch


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def assert_volume_close(got, want, exact=False):
    got = got.cpu().numpy() if isinstance(got, torch.Tensor) else got
    assert got.shape == want.shape
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin), "inf pattern differs"
    assert np.array_equal(got[~fin], want[~fin]) or np.array_equal(
        np.isposinf(got[~fin]), np.isposinf(want[~fin]))
    if exact:
        assert np.array_equal(got[fin], want[fin])
        return 0.0
    denom = np.maximum(np.abs(want[fin]), np.abs(got[fin]))
    err = np.abs(got[fin] - want[fin])
    rel = np.where(denom > 0, err / np.where(denom > 0, denom, 1), 0.0)
    assert rel.max(initial=0.0) <= RTOL, f"max rel err {rel.max()}"
    return float(rel.max(initial=0.0))


# ------------------------------------------------------------------------------ costs
COST_CASES = ["cost_raw_24x40x16", "cost_unit_40x24x16", "cost_unit_20x72x32"]


@pytest.mark.parametrize("case", COST_CASES)
def test_ssd_golden(sm, golden, case):
    g = golden(case)
    D = g["ssd7"].shape[2]
    l, r = cuda(g["left"]), cuda(g["right"])
    assert_volume_close(sm.cost.SSD(D, 7)(l, r), g["ssd7"], exact="raw" in case)
    assert_volume_close(sm.cost.SSD(D, 3)(l, r), g["ssd3"], exact="raw" in case)


@pytest.mark.parametrize("H,W,D,k,scale", [
    (37, 53, 32, 7, "raw"), (37, 53, 32, 7, "unit"), (19, 140, 64, 7, "unit"),
    (50, 31, 128, 7, "raw"), (23, 70, 192, 7, "unit"), (9, 11, 5, 2, "unit"),
    (16, 33, 32, 0, "raw"), (40, 200, 128, 7, "unit"),
    # every window half-size the packed kernel is built for (1..10), tile edges in x, d and y;
    # 11 and above take the straightforward kernel
    (45, 130, 64, 1, "raw"), (9, 50, 32, 1, "unit"), (45, 130, 64, 2, "raw"), (61, 97, 128, 2, "unit"),
    (45, 130, 64, 4, "raw"), (33, 49, 32, 4, "unit"), (70, 130, 64, 5, "raw"), (23, 41, 128, 5, "unit"),
    (70, 130, 64, 6, "raw"), (25, 49, 36, 6, "unit"), (70, 130, 64, 8, "raw"), (35, 33, 192, 8, "unit"),
    (300, 90, 32, 5, "raw"), (300, 70, 32, 8, "raw"), (40, 60, 32, 9, "raw"),
    (70, 130, 64, 9, "raw"), (41, 77, 128, 9, "unit"), (70, 130, 64, 10, "raw"), (300, 50, 32, 10, "unit"),
    (30, 60, 32, 11, "raw"),
])
def test_ssd_vs_oracle(sm, H, W, D, k, scale):
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H + W, scale=scale)
    got = sm.cost.SSD(D, k)(cuda(left), cuda(right))
    assert_volume_close(got, O.ssd(left, right, D, k), exact=(scale == "raw"))


@pytest.mark.parametrize("case", COST_CASES)
def test_birchfield_golden_shift_relation(sm, golden, case):
    """Against the compiled reference CPU op under the one-pixel shift relation."""
    g = golden(case)
    ref = g["bt_cpu"]
    H, W, D = ref.shape
    got = sm.cost.Birchfield(D)(cuda(g["left"]), cuda(g["right"])).cpu().numpy()
    for x in range(4, W - 6):
        dm = min(D, x - 3)
        a, b = got[:-1, x + 1, :dm], ref[:-1, x, :dm]
        assert np.allclose(a, b, rtol=RTOL, atol=0)


@pytest.mark.parametrize("H,W,D,scale", [
    (21, 45, 16, "raw"), (21, 45, 32, "unit"), (12, 150, 128, "unit"), (30, 9, 64, "raw"),
    (8, 70, 192, "unit"),
    (5, 7, 4, "unit"),        # W < 8: straightforward kernel
    (6, 61, 160, "unit"), (5, 90, 150, "raw"), (7, 77, 224, "unit"),   # 5 and 7 lane groups
    (4, 50, 300, "unit"),     # 10 lane groups: straightforward kernel
    (4, 203, 256, "unit"),    # 8 lane groups, width not a multiple of 8
    (9, 33, 20, "unit"), (3, 1000, 96, "raw"),
    (4, 1920, 128, "unit"),   # 1080p scanlines: 46 KB of shared memory per CTA
    (2, 2500, 64, "unit"),    # > 48 KB: opt-in shared memory
])
def test_birchfield_vs_oracle(sm, H, W, D, scale):
    """Bit-exact for every input: the kernels keep the oracle's expressions and its
    left-to-right tap order (taps outside [d, W) add an exact +0)."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=3 * H + W, scale=scale)
    got = sm.cost.Birchfield(D)(cuda(left), cuda(right))
    assert_volume_close(got, O.birchfield(left, right, D), exact=True)


# -------------------------------------------------------------------------------- SGM
SGM_CASES = ["sgm_unit_18x30x16", "sgm_unit_30x18x16", "sgm_raw_16x16x32",
             "sgm_raw_12x40x64", "sgm_minlast_14x22x8"]


@pytest.mark.parametrize("case", SGM_CASES)
def test_sgm_golden(sm, golden, case):
    g = golden(case)
    agg = sm.aggregation.Semiglobal(float(g["p"][0]), float(g["p"][1]))(cuda(g["cost"]),
                                                                       cuda(g["left"]))
    assert_volume_close(agg, g["agg"])
    # given the reference's aggregated volume, WTA is bit-exact
    assert np.array_equal(sm.disparity_reduce.WinnerTakesAll()(cuda(g["agg"])).cpu().numpy(),
                          g["wta"])


@pytest.mark.parametrize("H,W,D,scale,p1,p2,all_d", [
    (26, 41, 32, "unit", 0.1, 0.2, False), (41, 26, 32, "unit", 0.1, 0.2, False),
    (20, 20, 64, "raw", 20.0, 90.0, False), (17, 60, 128, "unit", 0.1, 0.2, False),
    (17, 60, 128, "raw", 50.0, 200.0, True), (13, 35, 192, "unit", 0.1, 0.2, False),
    (35, 13, 96, "unit", 0.3, 1.0, True), (11, 29, 20, "raw", 5.0, 9.0, False),
    (1, 9, 32, "unit", 0.1, 0.2, False), (9, 1, 32, "unit", 0.1, 0.2, False),
])
@pytest.mark.parametrize("mode", ["default", "sequential", "fused_sweeps", "fused_sweeps_no_overlap"])
def test_sgm_vs_oracle(sm, H, W, D, scale, p1, p2, all_d, mode):
    """All SGM modes: one kernel per direction -- concurrently for these small volumes
    (default) or one after the other -- and the fused raster sweeps."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H * 7 + W, scale=scale)
    cost = O.ssd(left, right, D, 3)
    want = O.sgm(cost, left, p1, p2, min_over_all_d=all_d)
    got = sm.aggregation.Semiglobal(p1, p2, min_over_all_d=all_d, fused_sweeps=mode.startswith("fused_sweeps"),
                                    overlap=mode != "fused_sweeps_no_overlap",
                                    concurrent_directions=mode != "sequential")(cuda(cost), cuda(left))
    assert_volume_close(got, want)


@pytest.mark.parametrize("H,W,D", [(4, 8, 128), (4, 2, 128), (5, 16, 128), (3, 25, 128), (6, 64, 256),
                                     (5, 19, 512), (7, 24, 320), (3, 17, 448), (9, 31, 160)])
def test_sgm_horizontal_bulk_copy_kernel_tiles(sm, H, W, D):
    """csrc/sgm_row.cu (directions 0 and 1 for D % 32 == 0): every tile size (8, 4, 2 pixels),
    widths that are a multiple of the tile, one more, one tile, less than a tile; rows beyond
    the last CTA's image rows (H % 4 != 0)."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H * 11 + W, scale="unit")
    cost = synth.random_volume(H, W, D, seed=D + W, hi=3.0)
    want = O.sgm(cost, left, 0.1, 0.2)
    c, l = cuda(cost), cuda(left)
    got = sm.aggregation.Semiglobal(concurrent_directions=False)(c, l)   # these kernels run
    assert_volume_close(got, want)
    # the raw op accumulates: direction 0 then runs the read-modify-write variant too
    from stereomatch._cstereomatch import AggregationOps, STM_SGM_SEQUENTIAL
    acc = torch.zeros_like(got)
    AggregationOps.run_semiglobal(c, l, 0.1, 0.2, acc, STM_SGM_SEQUENTIAL)
    assert torch.equal(acc, got)            # 0 + L is exact, the rest is the same sequence


@pytest.mark.parametrize("H,W,D", [(70, 150, 128), (150, 70, 64), (45, 333, 192), (33, 100, 32),
                                     (64, 64, 128), (5, 700, 64), (130, 9, 16), (24, 1900, 64),
                                     (40, 1920, 128), (37, 301, 256), (3, 2, 32), (2, 5, 64),
                                     (20, 3840, 64), (9, 2500, 32), (7, 3848, 48), (6, 3850, 64),
                                     # D = 96 / 160 and ranges below them (three / five scalar vectors per lane)
                                     (60, 200, 96), (150, 70, 96), (40, 1920, 96), (33, 100, 80), (45, 333, 160),
                                     (20, 301, 150), (64, 64, 96),
                                     # masked vector maps: multiples of 4 below 128 (4-wide), even ranges below 64 (2-wide)
                                     (33, 100, 100), (40, 300, 112), (31, 77, 72), (20, 200, 48), (25, 90, 40),
                                     (12, 1920, 100), (90, 60, 120), (30, 50, 36)])
def test_sgm_sweeps_many_ctas_and_ring_wraps(sm, H, W, D):
    """Fused vertical sweeps (csrc/sgm_vsweep.cu): one CTA to 148 CTAs, 2 to 13 column slots
    per CTA, more rows than every hand-off ring (4 rows in shared memory, 16 in global memory),
    slots that wrap around the image edge, landscape (defective up-diagonals),
    portrait and square; widths above 13 columns per SM (two sweep CTAs per SM at D <= 64; 3850 is
    one column too many for 296 CTAs and runs the direction passes)."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H + 3 * W, scale="unit")
    cost = O.ssd(left, right, D, 3)
    want = O.sgm(cost, left, 0.1, 0.2)
    l, c = cuda(left), cuda(cost)
    fused = sm.aggregation.Semiglobal(fused_sweeps=True)
    got = fused(c, l)
    assert_volume_close(got, want)
    # deterministic: a second run is bit-identical
    assert torch.equal(got, fused(c, l))
    # and the modes agree with each other to the same bar
    assert_volume_close(sm.aggregation.Semiglobal()(c, l), got.cpu().numpy())
    plain = sm.aggregation.Semiglobal(fused_sweeps=True, overlap=False)
    assert_volume_close(plain(c, l), want)
    _, d_over = fused.aggregate_and_reduce(c, l, keep_volume=False)
    _, d_plain = plain.aggregate_and_reduce(c, l, keep_volume=False)
    assert float((d_over != d_plain).float().mean()) <= 1e-3
    # the disparity-only forms (four volume streams at D = 128: out, T, U and the cost) pick the
    # oracle's winner wherever the oracle's winner is clear of the 1e-4 bar
    srt = np.sort(want, axis=-1)
    clear = (srt[..., 1] - srt[..., 0]) > 1e-4 * np.abs(srt[..., 0])
    best = want.argmin(-1)
    for d in (d_over, d_plain):
        assert np.array_equal(d.cpu().numpy()[clear], best[clear])
    assert clear.mean() > 0.5


def test_sgm_accumulates_like_the_reference(sm):
    """The raw op ACCUMULATES into the caller's volume (reference aggregation.py:51)."""
    from stereomatch._cstereomatch import AggregationOps
    H, W, D = 12, 20, 32
    left, right, _ = synth.random_dot_pair(H, W, D, seed=5, scale="unit")
    cost = O.ssd(left, right, D, 3)
    base = synth.random_volume(H, W, D, seed=6, hi=3.0)
    out = cuda(base.copy())
    AggregationOps.run_semiglobal(cuda(cost), cuda(left), 0.1, 0.2, out)
    want = O.sgm(cost, left, 0.1, 0.2, out=base.copy())
    assert_volume_close(out, want)


@pytest.mark.parametrize("H,W,D", [(20, 33, 32), (24, 24, 64), (17, 60, 128), (13, 35, 192), (11, 29, 20),
                                   (33, 20, 32), (60, 17, 128), (35, 13, 192), (29, 11, 20),   # portrait
                                   (450, 375, 64), (640, 480, 64),                            # portrait, 43 / 79 MB
                                   (375, 450, 128),    # 86 MB: the reference's own SGM benchmark shape
                                   (480, 640, 64)])    # 79 MB, VGA: mid-size volumes, workspace from the flags query
def test_sgm_concurrent_directions_equal_the_sequential_passes(sm, H, W, D):
    """Small landscape / square volumes run the eight directions concurrently into separate
    buffers and add them in the reference's order: bit-identical to the sequential passes
    (STM_SGM_SEQUENTIAL), for the plain call and the accumulating raw op; with the folded
    winner-takes-all the volume is in reference order (the sequential fold swaps 6 and 7)."""
    from stereomatch._cstereomatch import AggregationOps, STM_SGM_SEQUENTIAL
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H * 3 + W, scale="unit")
    c, l = cuda(O.ssd(left, right, D, 3)), cuda(left)
    conc_op = sm.aggregation.Semiglobal()
    seq_op = sm.aggregation.Semiglobal(concurrent_directions=False)
    base = cuda(synth.random_volume(H, W, D, seed=4, hi=2.0))
    conc, seq = conc_op(c, l).clone(), seq_op(c, l).clone()
    assert torch.equal(conc, seq)
    conc_acc, seq_acc = base.clone(), base.clone()
    AggregationOps.run_semiglobal(c, l, 0.1, 0.2, conc_acc)
    AggregationOps.run_semiglobal(c, l, 0.1, 0.2, seq_acc, STM_SGM_SEQUENTIAL)
    assert torch.equal(conc_acc, seq_acc)
    assert_volume_close(conc, O.sgm(O.ssd(left, right, D, 3), left, 0.1, 0.2))
    vol_c, disp_c = conc_op.aggregate_and_reduce(c, l)
    assert torch.equal(vol_c, seq)
    assert torch.equal(disp_c, sm.disparity_reduce.WinnerTakesAll()(vol_c))
    # per-pass timing measures sequential passes, so it selects them
    AggregationOps.pass_timing(True)
    try:
        assert torch.equal(conc_op(c, l), seq)
        assert len(AggregationOps.last_pass_ms()) == 8
    finally:
        AggregationOps.pass_timing(False)


def test_sgm_pass_timing_api(sm):
    """stm_sgm_pass_timing / stm_sgm_last_pass_ms: eight passes in launch order with positive
    durations (reference order for the plain call, 6 and 7 swapped when WTA is folded in),
    nothing when disabled, and the result is unchanged by timing."""
    from stereomatch._cstereomatch import AggregationOps
    H, W, D = 24, 40, 32
    left, right, _ = synth.random_dot_pair(H, W, D, seed=8, scale="unit")
    c, l = cuda(O.ssd(left, right, D, 3)), cuda(left)
    sgm = sm.aggregation.Semiglobal()
    want = sgm(c, l).clone()
    AggregationOps.pass_timing(True)
    try:
        got = sgm(c, l)
        passes = AggregationOps.last_pass_ms()
        assert [d for d, _ in passes] == [0, 1, 2, 3, 4, 5, 6, 7]
        assert all(t > 0 for _, t in passes)
        assert torch.equal(got, want)
        sgm.aggregate_and_reduce(c, l)
        assert [d for d, _ in AggregationOps.last_pass_ms()] == [0, 1, 2, 3, 4, 5, 7, 6]
    finally:
        AggregationOps.pass_timing(False)
    sgm(c, l)
    assert AggregationOps.last_pass_ms() == []


def test_sgm_reference_arange_recipe(sm):
    """The reference's own equality-test input (tests/test_semiglobalmatch.py:46-47), smaller."""
    H = W = 40
    D = 128
    vol = np.arange(H * W * D, dtype=np.float32).reshape(H, W, D)
    img = np.ones((H, W), np.float32)
    got = sm.aggregation.Semiglobal()(cuda(vol), cuda(img))
    assert_volume_close(got, O.sgm(vol, img, 0.1, 0.2))


# -------------------------------------------------------------------------------- WTA
@pytest.mark.parametrize("case", ["wta_ties_16x24x32", "wta_rand_9x13x192"])
def test_wta_golden(sm, golden, case):
    g = golden(case)
    got = sm.disparity_reduce.WinnerTakesAll()(cuda(g["cost"])).cpu().numpy()
    assert np.array_equal(got, g["wta"])


@pytest.mark.parametrize("H,W,D", [(7, 9, 1), (7, 9, 3), (5, 11, 4), (13, 17, 30), (33, 31, 64),
                                     (10, 23, 128), (6, 7, 192), (3, 5, 1000)])
def test_wta_vs_oracle_adversarial(sm, H, W, D):
    vol = synth.tie_heavy_volume(H, W, D, seed=D)
    vol[0, 0, :] = np.inf
    vol[0, 1, 0] = np.nan
    vol[0, 2, D // 2] = np.nan
    vol[0, 3, :] = np.nan
    vol[1, :, : D // 2] = np.inf
    vol[2, :, :] = -vol[2, :, :]
    got = sm.disparity_reduce.WinnerTakesAll()(cuda(vol)).cpu().numpy()
    assert np.array_equal(got, O.wta(vol))


# --------------------------------------------------------------------------------- DP
@pytest.mark.parametrize("case", ["dp_raw_32x48x16", "dp_raw_32x32x32", "dp_ties_16x20x8"])
def test_dp_cpu_exact_golden(sm, golden, case):
    g = golden(case)
    dp = sm.disparity_reduce.DynamicProgramming(mode="cpu_exact")
    got = dp(cuda(g["cost"])).cpu().numpy()
    assert np.array_equal(got, g["disp"])
    assert np.array_equal(dp._row_final_costs.cpu().numpy(), g["row_final"])
    assert np.array_equal(dp._path_volume.cpu().numpy(), g["path"])


@pytest.mark.parametrize("H,W,D,kind", [(20, 50, 32, "ssd"), (9, 33, 64, "ssd"), (12, 40, 128, "ssd"),
                                          (7, 30, 192, "ssd"), (15, 21, 12, "ties"), (6, 1, 32, "ties"),
                                          (5, 64, 96, "rand"),
                                          # forward prefetch groups of 8: none / exactly one / tail
                                          (4, 2, 32, "rand"), (4, 8, 32, "rand"), (4, 9, 128, "rand"),
                                          # backtrack windows of 16 columns: edges, many windows
                                          (3, 16, 64, "rand"), (3, 17, 64, "rand"), (3, 18, 64, "rand"),
                                          (6, 700, 128, "rand"), (5, 333, 192, "rand"),
                                          # one scalar per lane and vector (D % 32 == 0, not % 64): bulk kernel since r02 v12
                                          (6, 700, 96, "rand"), (5, 333, 160, "ties"), (4, 50, 224, "rand"),
                                          (4, 90, 5, "rand"),    # D smaller than the window
                                          (3, 120, 256, "rand"), (4, 400, 32, "ties"),
                                          # more rows than the 24-column-prefetch variant keeps
                                          # resident: the 8-column variant runs
                                          (3100, 7, 32, "rand"), (2500, 30, 128, "ties")])
def test_dp_cuda_mode_vs_oracle(sm, H, W, D, kind):
    if kind == "ssd":
        left, right, _ = synth.random_dot_pair(H, W, D, seed=H + D, scale="raw")
        vol = O.ssd(left, right, D, 3)
    elif kind == "ties":
        vol = synth.tie_heavy_volume(H, W, D, seed=D)
    else:
        vol = synth.random_volume(H, W, D, seed=D)
    dp = sm.disparity_reduce.DynamicProgramming()
    got = dp(cuda(vol)).cpu().numpy()
    want, wpath, wfinal = O.dp(vol, "cuda", return_aux=True)
    assert np.array_equal(got, want)
    assert np.array_equal(dp._row_final_costs.cpu().numpy(), wfinal)
    assert np.array_equal(dp._path_volume.cpu().numpy()[:, 1:], wpath[:, 1:])


# --------------------------------------------------------------------------- pipeline
@pytest.mark.parametrize("cost_name,agg,red", [("SSD", True, "wta"), ("SSD", False, "wta"),
                                               ("Birchfield", True, "wta"), ("SSD", False, "dp"),
                                               ("Birchfield", False, "dp")])
def test_pipeline_end_to_end(sm, cost_name, agg, red):
    H, W, D = 48, 96, 32
    left, right, gt = synth.random_dot_pair(H, W, D, seed=99, scale="unit")
    cost_fn = sm.cost.SSD(D) if cost_name == "SSD" else sm.cost.Birchfield(D)
    reducer = sm.disparity_reduce.WinnerTakesAll() if red == "wta" else \
        sm.disparity_reduce.DynamicProgramming()
    pipe = sm.Pipeline(cost_fn, reducer, sm.aggregation.Semiglobal() if agg else None)
    got = pipe.estimate(cuda(left), cuda(right))
    first_ptr = got.data_ptr()
    got = pipe.estimate(cuda(left), cuda(right))  # second frame reuses every buffer
    assert got.data_ptr() == first_ptr and got.dtype == torch.int32
    vol = O.ssd(left, right, D, 7) if cost_name == "SSD" else O.birchfield(left, right, D)
    if agg:
        vol = O.sgm(vol, left, 0.1, 0.2)
    want = O.wta(vol) if red == "wta" else O.dp(vol, "cuda")
    mismatch = float((got.cpu().numpy() != want).mean())
    assert mismatch <= 1e-3, mismatch  # north_star: <= 0.1 % of pixels


def test_pipeline_host_entry_point(sm):
    """stm_pipeline_host_f32: host buffers in, host map out, copies inside the call."""
    from stereomatch import _cstereomatch as C
    H, W, D = 40, 72, 64
    left, right, _ = synth.random_dot_pair(H, W, D, seed=123, scale="unit")
    got = C.pipeline_host(torch.from_numpy(left), torch.from_numpy(right), D).numpy()
    want = O.wta(O.sgm(O.ssd(left, right, D, 7), left, 0.1, 0.2))
    assert float((got != want).mean()) <= 1e-3
    got = C.pipeline_host(torch.from_numpy(left), torch.from_numpy(right), D, use_sgm=False,
                          reducer=1).numpy()
    assert np.array_equal(got, O.dp(O.ssd(left, right, D, 7), "cuda"))


@pytest.mark.parametrize("H,W,D,kind", [(26, 41, 32, "ssd"), (41, 26, 32, "ssd"), (20, 20, 64, "ssd"),
                                          (17, 60, 128, "ssd"), (13, 35, 192, "ssd"), (11, 29, 20, "ssd"),
                                          (9, 33, 32, "ties"), (1, 9, 32, "ssd"), (9, 1, 32, "ssd"),
                                          (70, 150, 128, "ssd")])
@pytest.mark.parametrize("concurrent", [False, True])
def test_sgm_with_wta_folded_into_the_last_pass(sm, H, W, D, kind, concurrent):
    """stm_sgm_wta_f32: the disparity map is EXACTLY WinnerTakesAll of the volume it returns
    (landscape, portrait, square, ties, odd D), and that volume is the oracle's SGM to 1e-4 --
    with the argmin folded into the last sequential pass, and with concurrent directions."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H * 5 + W, scale="unit")
    cost = O.ssd(left, right, D, 3) if kind == "ssd" else synth.tie_heavy_volume(H, W, D, seed=D)
    c, l = cuda(cost), cuda(left)
    sgm = sm.aggregation.Semiglobal(concurrent_directions=concurrent)
    vol, disp = sgm.aggregate_and_reduce(c, l)
    assert disp.dtype == torch.int32 and tuple(disp.shape) == (H, W)
    assert_volume_close(vol, O.sgm(cost, left, 0.1, 0.2))
    assert np.array_equal(disp.cpu().numpy(), O.wta(vol.cpu().numpy()))
    assert torch.equal(disp, sm.disparity_reduce.WinnerTakesAll()(vol))
    # buffers are reused when passed back in
    vol2, disp2 = sgm.aggregate_and_reduce(c, l, vol, disp)
    assert vol2.data_ptr() == vol.data_ptr() and disp2.data_ptr() == disp.data_ptr()
    # STM_SGM_DISPARITY_ONLY: the last pass does not write the sums back; same disparity map
    want = disp.clone()
    scratch, disp3 = sgm.aggregate_and_reduce(c, l, keep_volume=False)
    assert torch.equal(disp3, want)
    if not concurrent:
        assert not torch.equal(scratch, vol)  # sequential fold: the volume really was left unfinished


@pytest.mark.parametrize("concurrent", [False, True])
def test_sgm_wta_fold_argmin_rule_with_nan_and_inf(sm, concurrent):
    """Pixels whose whole cost vector is +inf turn into NaN sums (inf - inf) that spread along
    the paths: the folded argmin must follow the same rule as the WTA kernel and the oracle
    (first minimum, NaN never wins, NaN at d = 0 -> 0, all-inf -> 0) on the SAME volume."""
    H, W, D = 12, 20, 32
    vol = synth.random_volume(H, W, D, seed=5, hi=10.0)
    vol[3, 4, :] = np.inf
    vol[7, 11, :] = np.inf
    vol[5, 6, 0] = np.inf
    vol[:, 0, 1:] = np.inf          # like d > x
    img = synth.random_volume(H, W, 1, seed=6, hi=255.0)[:, :, 0]
    agg, disp = sm.aggregation.Semiglobal(concurrent_directions=concurrent).aggregate_and_reduce(
        cuda(vol), cuda(img))
    a = agg.cpu().numpy()
    assert np.isnan(a).any() and np.isinf(a).any()      # the case is what it claims to be
    assert np.array_equal(disp.cpu().numpy(), O.wta(a))
    assert torch.equal(disp, sm.disparity_reduce.WinnerTakesAll()(agg))


def test_pipeline_fused_and_unfused_agree(sm):
    H, W, D = 48, 96, 32
    left, right, _ = synth.random_dot_pair(H, W, D, seed=99, scale="unit")
    l, r = cuda(left), cuda(right)
    mk = lambda: sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal())
    fused, plain = mk(), mk()
    plain.fuse_aggregation_and_reduce = False
    a, b = fused.estimate(l, r), plain.estimate(l, r)
    assert float((a != b).float().mean()) <= 1e-3   # north_star: <= 0.1 % of pixels
    keep = mk()
    keep.keep_aggregation_volume = True             # default False: the volume is scratch
    assert torch.equal(keep.estimate(l, r), a)
    assert torch.equal(sm.disparity_reduce.WinnerTakesAll()(keep._aggregation_volume), a)
    want = O.wta(O.sgm(O.ssd(left, right, D, 7), left, 0.1, 0.2))
    assert float((a.cpu().numpy() != want).mean()) <= 1e-3
    assert float((b.cpu().numpy() != want).mean()) <= 1e-3


def _ssd_sgm_wta(sm, D):
    return sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal())


@pytest.mark.parametrize("H,W,D,n_pipes", [(40, 96, 32, 1), (40, 96, 32, 2), (40, 96, 32, 3),
                                            (300, 400, 128, 2)])   # 61 MB: sequential SGM passes
def test_batch_estimator_matches_per_frame_estimate(sm, H, W, D, n_pipes):
    """Frames round-robin over several Pipelines on their own streams return exactly what one
    Pipeline returns frame by frame (no shared workspace, no shared result buffer), with and
    without caller-provided outputs, and when called twice."""
    n_frames = 5
    frames = [synth.random_dot_pair(H, W, D, seed=700 + f, scale="unit")[:2] for f in range(n_frames)]
    dev = [(cuda(l), cuda(r)) for l, r in frames]
    ref_pipe = _ssd_sgm_wta(sm, D)
    want = [ref_pipe.estimate(l, r).clone() for l, r in dev]
    batch = sm.streaming.BatchEstimator([_ssd_sgm_wta(sm, D) for _ in range(n_pipes)], "cuda")
    got = batch.estimate_batch(dev)
    torch.cuda.current_stream().synchronize()
    assert len(got) == n_frames and all(torch.equal(g, w) for g, w in zip(got, want))
    outs = [torch.full((H, W), -1, dtype=torch.int32, device="cuda") for _ in frames]
    for _ in range(2):
        back = batch.estimate_batch(dev, outs)
        torch.cuda.current_stream().synchronize()
        assert all(b.data_ptr() == o.data_ptr() for b, o in zip(back, outs))
        assert all(torch.equal(o, w) for o, w in zip(outs, want))
    with pytest.raises(ValueError, match="distinct"):
        p = _ssd_sgm_wta(sm, D)
        sm.streaming.BatchEstimator([p, p])


@pytest.mark.parametrize("extra", [0, 1, 2])
def test_host_frame_streamer_matches_per_frame_estimate(sm, extra):
    """Overlapped upload / compute / download returns exactly what Pipeline.estimate returns,
    for more frames than device buffers, with one or several compute streams, called twice."""
    H, W, D = 40, 96, 32
    frames = [synth.random_dot_pair(H, W, D, seed=400 + f, scale="unit")[:2] for f in range(7)]
    pipe = _ssd_sgm_wta(sm, D)
    want = [pipe.estimate(cuda(l), cuda(r)).cpu().clone() for l, r in frames]
    pinned = [(torch.from_numpy(l).pin_memory(), torch.from_numpy(r).pin_memory()) for l, r in frames]
    outs = [torch.full((H, W), -1, dtype=torch.int32).pin_memory() for _ in frames]
    streamer = sm.streaming.HostFrameStreamer(pipe, "cuda",
                                              extra_pipelines=[_ssd_sgm_wta(sm, D) for _ in range(extra)])
    for _ in range(2):
        for o in outs:
            o.fill_(-1)
        streamer.run(pinned, outs)
        torch.cuda.current_stream().synchronize()
        for got, w in zip(outs, want):
            assert torch.equal(got, w)
    with pytest.raises(RuntimeError, match="host tensors"):
        streamer.run([(cuda(frames[0][0]), cuda(frames[0][1]))], outs[:1])


@pytest.mark.parametrize("kind,H,W,D", [("ssd_sgm_wta", 375, 450, 64), ("bt_sgm_wta", 120, 200, 64),
                                         ("ssd_dp", 96, 160, 64), ("ssd_wta", 64, 96, 32),
                                         ("ssd_sgm_wta", 1080, 1920, 128)])
def test_graphed_pipeline_replays_exactly_what_estimate_computes(sm, kind, H, W, D):
    """streaming.GraphedPipeline: Pipeline.estimate of one shape captured into a CUDA graph (helper
    streams of the concurrent / overlapped SGM schedules and the cooperative sweeps included) and
    replayed for new images returns exactly the eager result, frame after frame."""
    def make():
        cost = sm.cost.Birchfield(D) if kind.startswith("bt") else sm.cost.SSD(D)
        if kind.endswith("dp"):
            return sm.Pipeline(cost, sm.disparity_reduce.DynamicProgramming())
        agg = sm.aggregation.Semiglobal() if "sgm" in kind else None
        return sm.Pipeline(cost, sm.disparity_reduce.WinnerTakesAll(), agg)
    frames = [synth.random_dot_pair(H, W, D, seed=900 + f, scale="raw")[:2] for f in range(3)]
    dev = [(cuda(l), cuda(r)) for l, r in frames]
    eager = make()
    want = [eager.estimate(l, r).clone() for l, r in dev]
    graphed = sm.streaming.GraphedPipeline(make(), H, W)
    for _ in range(2):
        for (l, r), w in zip(dev, want):
            got = graphed.estimate(l, r)
            assert got.data_ptr() == graphed.out.data_ptr() and torch.equal(got, w)
    mine = torch.full((H, W), -1, dtype=torch.int32, device="cuda")
    assert graphed.estimate(*dev[1], out=mine) is mine and torch.equal(mine, want[1])
    graphed.left.copy_(dev[2][0]); graphed.right.copy_(dev[2][1])        # in-place inputs: no copies
    assert torch.equal(graphed.estimate(graphed.left, graphed.right), want[2])
    with pytest.raises(RuntimeError, match="captured for"):
        graphed.estimate(dev[0][0][:-1], dev[0][1][:-1])


def test_errors_mirror_the_reference(sm):
    img = torch.zeros(8, 8, device="cuda")
    with pytest.raises(RuntimeError, match="type not supported"):
        sm.cost.SSD(4)(img.to(torch.uint8), img.to(torch.uint8))
    with pytest.raises(RuntimeError, match="gpu tensor"):
        sm.cost.SSD(4)(img, img.cpu())
    with pytest.raises(RuntimeError, match="D <= H <= W"):
        sm.disparity_reduce.DynamicProgramming(mode="cpu_exact")(torch.zeros(4, 8, 16, device="cuda"))


# ------------------------------------------------- full-size, size-independent properties
def test_full_size_1080p_properties(sm):
    """1920x1080 D=128: properties that need no CPU oracle run."""
    H, W, D = 1080, 1920, 128
    left, right, gt = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
    l, r = cuda(left), cuda(right)
    cost = sm.cost.SSD(D)(l, r)
    # (1) d > x is +inf, everything else finite and >= 0
    dd = torch.arange(D, device="cuda")[None, :]
    xx = torch.arange(W, device="cuda")[:, None]
    assert bool(torch.isinf(cost[:, dd.expand(W, D) > xx]).all())
    assert bool((cost[:, dd.expand(W, D) <= xx] >= 0).all())
    # (2) a band of rows against the oracle (bit-exact on integer-valued images)
    band = slice(533, 547)
    want = O.ssd(left[band.start - 7:band.stop + 7], right[band.start - 7:band.stop + 7], D, 7)[7:-7]
    assert np.array_equal(cost[band].cpu().numpy(), want)
    # (3) WTA == torch.argmin on NaN-free volumes, and recovers the ground truth
    wta = sm.disparity_reduce.WinnerTakesAll()(cost)
    assert bool((wta.long() == torch.argmin(cost, dim=2)).all())
    valid = torch.from_numpy(np.arange(W)[None, :] >= gt + 7).cuda()
    acc = float((wta[valid] == cuda(gt)[valid]).float().mean())
    assert acc > 0.99, acc
    # (4) SGM: linear in the cost volume's positive scale when penalties scale too
    sgm = sm.aggregation.Semiglobal(30.0, 120.0)
    agg = sgm(cost, l)
    agg2 = sm.aggregation.Semiglobal(60.0, 240.0)(cost * 2, l)
    fin = torch.isfinite(agg)
    assert bool((torch.isfinite(agg2) == fin).all())
    assert bool(torch.equal(agg2[fin], agg[fin] * 2))  # power-of-two scaling is exact in fp32
    # (4b) at this size the default schedule is the fused sweeps (csrc/sgm_vsweep.cu, 148 CTAs
    #      x 13 column slots): against the reference-order passes within the 1e-4 bar, the
    #      folded argmin equal to WTA of the volume, and two runs bit-identical
    ref_order = sm.aggregation.Semiglobal(30.0, 120.0, fused_sweeps=False)(cost, l)
    assert bool((torch.isfinite(ref_order) == fin).all())
    den = torch.maximum(ref_order[fin].abs(), agg[fin].abs()).clamp_min(1e-30)
    assert float(((ref_order[fin] - agg[fin]).abs() / den).max()) <= RTOL
    del ref_order, den
    vol, disp = sgm.aggregate_and_reduce(cost, l)
    assert bool(torch.equal(vol, agg))
    assert bool(torch.equal(disp, sm.disparity_reduce.WinnerTakesAll()(vol)))
    del vol
    # (5) a path-local check: the first row only sees -> <- and the starts of v \v /v
    #     so a 3-row crop reproduces row 0 of `->`+`<-` ... use the oracle on a narrow crop
    #     of full width is too slow; instead check SGM+WTA recovers the ground truth.
    wta2 = sm.disparity_reduce.WinnerTakesAll()(agg)
    acc2 = float((wta2[valid] == cuda(gt)[valid]).float().mean())
    assert acc2 > 0.99, acc2
    # (6) DP runs at full size and returns in-range disparities; its forward costs are
    #     reproducible (same input -> identical output)
    dp = sm.disparity_reduce.DynamicProgramming()
    d1 = dp(cost).clone()
    d2 = dp(cost)
    assert bool(torch.equal(d1, d2)) and int(d1.min()) >= 0 and int(d1.max()) < D


@pytest.mark.parametrize("H,W,D", [(30, 47, 128), (47, 30, 64), (21, 21, 32), (19, 333, 192), (30, 47, 96),
                                   (19, 333, 160), (25, 90, 70), (25, 90, 100), (47, 30, 48)])
def test_sgm_fused_sweeps_fold_winner_takes_all(sm, H, W, D):
    """stm_sgm_wta_f32 with STM_SGM_FUSED_SWEEPS: the up sweep emits the disparity map; it is
    exactly WinnerTakesAll of the volume the same call returns, the volume meets the oracle
    bar, and STM_SGM_DISPARITY_ONLY (no write-back) gives the same map."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H * 5 + W, scale="unit")
    cost = O.ssd(left, right, D, 3)
    c, l = cuda(cost), cuda(left)
    fused = sm.aggregation.Semiglobal(fused_sweeps=True)
    vol, disp = fused.aggregate_and_reduce(c, l)
    assert_volume_close(vol, O.sgm(cost, left, 0.1, 0.2))
    assert torch.equal(disp, sm.disparity_reduce.WinnerTakesAll()(vol))
    assert torch.equal(vol, fused(c, l))
    _, disp2 = fused.aggregate_and_reduce(c, l, keep_volume=False)
    assert torch.equal(disp2, disp)


@pytest.mark.parametrize("fused_sweeps", [False, True])
def test_full_height_sgm_column_strip_vs_oracle(sm, fused_sweeps):
    """Full-height (1080 rows), narrow strip: exercises long paths and the 1080-fold
    corner-diagonal multiplicity against the oracle (12 CTAs in the fused-sweep mode)."""
    H, W, D = 1080, 24, 32
    left, right, _ = synth.random_dot_pair(H, W, D, seed=77, scale="unit")
    cost = O.ssd(left, right, D, 3)
    sgm = sm.aggregation.Semiglobal(fused_sweeps=fused_sweeps)
    got = sgm(cuda(cost), cuda(left))
    assert_volume_close(got, O.sgm(cost, left, 0.1, 0.2))
    H, W = 24, 1100
    left, right, _ = synth.random_dot_pair(H, W, D, seed=78, scale="unit")
    cost = O.ssd(left, right, D, 3)
    got = sgm(cuda(cost), cuda(left))
    assert_volume_close(got, O.sgm(cost, left, 0.1, 0.2))


@pytest.mark.parametrize("H,W,D,scale", [(21, 45, 64, "unit"), (12, 150, 128, "unit"), (8, 70, 192, "raw"),
                                          (4, 203, 256, "unit"), (30, 9, 64, "raw"), (1080, 1920, 128, "raw")])
def test_birchfield_with_winner_takes_all_folded_in(sm, H, W, D, scale):
    """stm_birchfield_wta_f32: the map is EXACTLY WinnerTakesAll of the Birchfield volume (same
    kernel arithmetic, first-minimum rule, +inf for d > x), without the volume being written;
    Pipeline(Birchfield, WinnerTakesAll) takes that path and leaves no cost volume behind."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=3 * H + W, scale=scale)
    l, r = cuda(left), cuda(right)
    bt = sm.cost.Birchfield(D)
    want = sm.disparity_reduce.WinnerTakesAll()(bt(l, r))
    got = bt.cost_and_reduce(l, r)
    assert got is not None and got.dtype == torch.int32
    assert torch.equal(got, want)
    if H * W * D < 4_000_000:
        assert np.array_equal(got.cpu().numpy(), O.wta(O.birchfield(left, right, D)))
    pipe = sm.Pipeline(sm.cost.Birchfield(D), sm.disparity_reduce.WinnerTakesAll())
    assert torch.equal(pipe.estimate(l, r), want) and pipe._cost_volume is None
    pipe.fuse_cost_and_reduce = False
    assert torch.equal(pipe.estimate(l, r), want) and pipe._cost_volume is not None
    # a tie-heavy pair: constant images make every cost 0 -> lowest d wins everywhere
    z = torch.full((H, W), 7.0, device="cuda")
    assert int(bt.cost_and_reduce(z, z).max()) == 0
    # unsupported D falls back inside Pipeline
    assert sm.cost.Birchfield(48).cost_and_reduce(l, r) is None
    p48 = sm.Pipeline(sm.cost.Birchfield(48), sm.disparity_reduce.WinnerTakesAll())
    assert torch.equal(p48.estimate(l, r), sm.disparity_reduce.WinnerTakesAll()(sm.cost.Birchfield(48)(l, r)))


@pytest.mark.parametrize("H,W,D,k,scale", [(21, 45, 64, 7, "unit"), (33, 150, 128, 7, "raw"), (40, 70, 192, 7, "raw"),
                                            (17, 203, 256, 3, "unit"), (30, 9, 64, 7, "raw"), (12, 50, 36, 7, "raw"),
                                            (64, 96, 4, 3, "raw"), (375, 450, 64, 7, "raw"), (1080, 1920, 128, 7, "raw"),
                                            (40, 90, 64, 1, "raw"), (40, 90, 64, 2, "unit"), (40, 90, 64, 4, "raw"),
                                            (40, 90, 64, 5, "unit"), (40, 90, 64, 6, "raw"), (40, 90, 64, 8, "raw"),
                                            (40, 90, 64, 9, "raw"), (40, 90, 64, 10, "unit")])
def test_ssd_with_winner_takes_all_folded_in(sm, H, W, D, k, scale):
    """stm_ssd_wta_f32: the map is EXACTLY WinnerTakesAll of the SSD volume (same kernel arithmetic,
    first-minimum rule across the disparity chunks of different CTAs, +inf for d > x), without the
    volume being written; Pipeline(SSD, WinnerTakesAll) takes that path and leaves no cost volume."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=5 * H + W, scale=scale)
    l, r = cuda(left), cuda(right)
    ssd = sm.cost.SSD(D, k)
    want = sm.disparity_reduce.WinnerTakesAll()(ssd(l, r))
    got = ssd.cost_and_reduce(l, r)
    assert got is not None and got.dtype == torch.int32
    assert torch.equal(got, want)
    assert torch.equal(ssd.cost_and_reduce(l, r, got), want)  # second call: the scratch keys are re-initialised
    if H * W * D < 12_000_000:
        assert np.array_equal(got.cpu().numpy(), O.wta(O.ssd(left, right, D, k)))
    pipe = sm.Pipeline(sm.cost.SSD(D, k), sm.disparity_reduce.WinnerTakesAll())
    big = H * W * D * 4 > (48 << 20)   # below 48 MB the volume stays in L2 and the two stages are faster
    assert torch.equal(pipe.estimate(l, r), want) and (pipe._cost_volume is None) == big
    pipe = sm.Pipeline(sm.cost.SSD(D, k), sm.disparity_reduce.WinnerTakesAll())
    pipe.fuse_cost_and_reduce = "always"
    assert torch.equal(pipe.estimate(l, r), want) and pipe._cost_volume is None
    pipe.fuse_cost_and_reduce = False
    assert torch.equal(pipe.estimate(l, r), want) and pipe._cost_volume is not None
    # ties: constant images make every cost 0 -> lowest d wins everywhere (also across CTA chunks)
    z = torch.full((H, W), 7.0, device="cuda")
    assert int(ssd.cost_and_reduce(z, z).max()) == 0
    # NaN at d = 0 -> 0 (the reference's `<` never replaces a NaN best); a NaN elsewhere never wins
    ln = l.clone()
    ln[H // 2, W // 2] = float("nan")
    assert torch.equal(ssd.cost_and_reduce(ln, r), sm.disparity_reduce.WinnerTakesAll()(ssd(ln, r)))


def test_ssd_with_winner_takes_all_unsupported_shapes_fall_back(sm):
    left, right, _ = synth.random_dot_pair(24, 60, 32, seed=9, scale="raw")
    l, r = cuda(left), cuda(right)
    wta = sm.disparity_reduce.WinnerTakesAll()
    assert sm.cost.SSD(30).cost_and_reduce(l, r) is None          # D % 4 != 0
    assert sm.cost.SSD(32, 11).cost_and_reduce(l, r) is None       # window the packed kernel is not built for
    for cost in (sm.cost.SSD(30), sm.cost.SSD(32, 11)):
        pipe = sm.Pipeline(cost, wta)
        assert torch.equal(pipe.estimate(l, r), wta(cost(l, r)))


def test_dp_default_mode_stays_close_to_the_pinned_cpu_exact_mode(sm):
    """The default (Jacobi) DP mode is a defined variant, not the reference CPU functor: they are
    different algorithms (SURVEY.md 8.1 measured 4.56 % of pixels different at 375x450x64).  The
    distance is bounded here so that a regression in either kernel shows: on the random-dot SSD
    volume of BASELINE config 2 both recover the ground truth and differ at < 10 % of pixels."""
    H, W, D = 375, 450, 64
    left, right, gt = synth.random_dot_pair(H, W, D, seed=0, scale="raw")
    vol = sm.cost.SSD(D)(cuda(left), cuda(right))
    a = sm.disparity_reduce.DynamicProgramming()(vol).cpu().numpy()
    b = sm.disparity_reduce.DynamicProgramming(mode="cpu_exact")(vol).cpu().numpy()
    diff = float((a != b).mean())
    valid = np.arange(W)[None, :] >= gt + 7
    acc_a, acc_b = float((a[valid] == gt[valid]).mean()), float((b[valid] == gt[valid]).mean())
    print(f"DP default vs cpu_exact at {H}x{W}x{D}: {diff:.4f} of pixels differ; accuracy {acc_a:.4f} / {acc_b:.4f}")
    assert 0.0 < diff < 0.10
    assert acc_a > 0.90 and acc_b > 0.90
