"""Parity ON THE NAMED SHAPES (BASELINE.json configs 2-5, north_star: "identical synthetic
random-dot stereograms of the named shapes"), through the schedule each shape gets BY DEFAULT
(at 1080p: two horizontal passes + two fused vertical sweeps, the argmin folded into the up
sweep, no write-back of the sums), against the CPU oracle run at FULL size.

    config 2   375 x  450, D =  64   SSD / Birchfield, SGM (all schedules), WTA, DP, Pipeline
    config 3  1080 x 1920, D = 128   Birchfield + SGM + WTA
    config 4  1080 x 1920, D = 128   SSD + DP (both modes)
    headline  1080 x 1920, D = 128   SSD + SGM(P1=.1, P2=.2) + WTA -- bench.py's exact workload
    config 5  1080 x 1920, D = 192   SSD + SGM + WTA, seeds 1000 and 1063 (first / last frame)

Bars (BASELINE.json): cost volumes and SGM sums within 1e-4 relative on finite entries with an
identical inf pattern (bit-exact where the arithmetic is exact: SSD on the raw 0..255 scale,
Birchfield everywhere); WTA / DP maps bit-exact given identical volumes; end-to-end disparity
maps differ from the oracle chain at <= 0.1 % of pixels.  The mismatch rates are printed (-s).
The oracle costs ~12 s (SGM 1080p D=128) / ~20 s (D=192) per run; the whole file takes a few
minutes of the box's host cores.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as O, ref_ops, synth

pytestmark = pytest.mark.gpu

RTOL = 1e-4          # north_star: volumes within 1e-4 relative
MAX_MISMATCH = 1e-3  # north_star: end-to-end maps differ at <= 0.1 % of pixels


@pytest.fixture(scope="module")
def sm():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import stereomatch
    return stereomatch


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def volume_close(got: torch.Tensor, want: np.ndarray, exact: bool = False, chunk_rows: int = 135):
    """`got` (CUDA) against the oracle's `want` (host), compared on the device in row chunks:
    identical inf / NaN pattern, then bit-equality or max relative error <= RTOL on the finite
    entries.  Returns the max relative error."""
    assert tuple(got.shape) == tuple(want.shape)
    worst = 0.0
    for a in range(0, want.shape[0], chunk_rows):
        w = torch.from_numpy(want[a:a + chunk_rows]).cuda()
        g = got[a:a + chunk_rows]
        fin = torch.isfinite(w)
        assert bool((torch.isfinite(g) == fin).all()), "inf pattern differs"
        assert bool((torch.isposinf(g) == torch.isposinf(w)).all()), "+inf pattern differs"
        if exact:
            assert bool((g[fin] == w[fin]).all()), f"not bit-exact in rows {a}..{a + chunk_rows}"
            continue
        gf, wf = g[fin], w[fin]
        den = torch.maximum(gf.abs(), wf.abs())
        rel = torch.where(den > 0, (gf - wf).abs() / den.clamp_min(1e-38), torch.zeros_like(den))
        worst = max(worst, float(rel.max()) if rel.numel() else 0.0)
    assert worst <= RTOL, f"max rel err {worst}"
    return worst


def mismatch(got: torch.Tensor, want: np.ndarray) -> float:
    return float((got.cpu().numpy() != want).mean())


def oracle_ssd(left, right, D, scale):
    # raw scale: integer images, every order of summation is exact (oracle.ssd_exact_int is
    # checked equal to the literal restatement on the CPU side); unit scale: the literal loop
    return O.ssd_exact_int(left, right, D, 7) if scale == "raw" else O.ssd_threaded(left, right, D, 7)


# ------------------------------------------------------------------ config 2: 375 x 450, D = 64
@pytest.mark.parametrize("scale", ["raw", "unit"])
def test_config2_teddy_size_every_stage_and_schedule(sm, scale):
    H, W, D = 375, 450, 64
    left, right, gt = synth.random_dot_pair(H, W, D, seed=1000, scale=scale)
    l, r = cuda(left), cuda(right)
    ocost = O.ssd_threaded(left, right, D, 7)          # the literal loop on both scales
    if scale == "raw":
        assert np.array_equal(ocost, O.ssd_exact_int(left, right, D, 7))
    cost = sm.cost.SSD(D)(l, r)
    volume_close(cost, ocost, exact=scale == "raw")
    volume_close(sm.cost.Birchfield(D)(l, r), O.birchfield_threaded(left, right, D), exact=True)

    oagg = O.sgm(ocost, left, 0.1, 0.2)
    owta = O.wta(oagg)
    oc = cuda(ocost)
    for kw in ({}, {"concurrent_directions": False}, {"fused_sweeps": True}):
        agg = sm.aggregation.Semiglobal(0.1, 0.2, **kw)(oc, l)
        err = volume_close(agg, oagg)
        m = mismatch(sm.disparity_reduce.WinnerTakesAll()(agg), owta)
        print(f"config2 {scale} sgm {kw}: max rel err {err:.2e}, WTA mismatch {m:.2e}")
        assert m <= MAX_MISMATCH
        _, disp = sm.aggregation.Semiglobal(0.1, 0.2, **kw).aggregate_and_reduce(oc, l, keep_volume=False)
        assert mismatch(disp, owta) <= MAX_MISMATCH
    # given identical volumes: bit-exact maps
    assert np.array_equal(sm.disparity_reduce.WinnerTakesAll()(cuda(oagg)).cpu().numpy(), owta)
    assert np.array_equal(sm.disparity_reduce.WinnerTakesAll()(oc).cpu().numpy(), O.wta(ocost))
    for mode in ("cuda", "cpu_exact"):
        dp = sm.disparity_reduce.DynamicProgramming(mode=mode)
        odisp, opath, ofinal = O.dp(ocost, mode, return_aux=True)
        assert np.array_equal(dp(oc).cpu().numpy(), odisp)
        assert np.array_equal(dp._row_final_costs.cpu().numpy(), ofinal)
        if mode == "cpu_exact":
            assert np.array_equal(dp._path_volume.cpu().numpy(), opath)
        else:
            assert np.array_equal(dp._path_volume.cpu().numpy()[:, 1:], opath[:, 1:])
    # end to end through the public API
    pipe = sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal())
    m = mismatch(pipe.estimate(l, r), owta)
    print(f"config2 {scale} Pipeline SSD+SGM+WTA vs oracle chain: {m:.2e}")
    assert m <= MAX_MISMATCH
    pipe = sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.DynamicProgramming())
    assert mismatch(pipe.estimate(l, r), O.dp(ocost, "cuda")) <= (0.0 if scale == "raw" else MAX_MISMATCH)


def test_config2_against_the_compiled_reference_itself(sm):
    """Same pair through the UNMODIFIED reference's CPU ops (oracle/_ref, built from
    /root/reference by oracle/build_ref.py; travels to the box) instead of the restatement."""
    if not ref_ops.available():
        pytest.skip("oracle/_ref/_cstereomatch.so not built")
    H, W, D = 375, 450, 64
    left, right, gt = synth.random_dot_pair(H, W, D, seed=1000, scale="raw")
    l, r = cuda(left), cuda(right)
    rcost = ref_ops.ssd(left, right, D, 7)
    volume_close(sm.cost.SSD(D)(l, r), rcost, exact=True)
    ragg = ref_ops.sgm(rcost, left, 0.1, 0.2)
    volume_close(sm.aggregation.Semiglobal()(cuda(rcost), l), ragg)
    rwta = ref_ops.wta(ragg)
    assert np.array_equal(sm.disparity_reduce.WinnerTakesAll()(cuda(ragg)).cpu().numpy(), rwta)
    pipe = sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal())
    assert mismatch(pipe.estimate(l, r), rwta) <= MAX_MISMATCH
    rdp = ref_ops.dp(rcost)     # the reference CPU functor == mode "cpu_exact"
    got = sm.disparity_reduce.DynamicProgramming(mode="cpu_exact")(cuda(rcost))
    assert np.array_equal(got.cpu().numpy(), rdp)
    # survey anchors (seed 0): the chain recovers the ground truth
    valid = np.arange(W)[None, :] >= gt + 7
    assert (rwta[valid] == gt[valid]).mean() > 0.99


# ------------------------------------------- headline + config 4: 1080 x 1920, D = 128, SSD
@pytest.mark.parametrize("scale", ["raw", "unit"])
def test_1080p_d128_ssd_sgm_wta_default_schedule_vs_oracle(sm, scale):
    """bench.py's exact workload (raw scale, P1 = 0.1, P2 = 0.2) and the unit scale."""
    H, W, D = 1080, 1920, 128
    left, right, gt = synth.random_dot_pair(H, W, D, seed=1000, scale=scale)
    l, r = cuda(left), cuda(right)
    ocost = oracle_ssd(left, right, D, scale)
    cost = sm.cost.SSD(D)(l, r)
    err = volume_close(cost, ocost, exact=scale == "raw")   # the WHOLE volume, borders and d > x included
    print(f"1080p D=128 {scale} SSD max rel err {err:.2e}")
    if scale == "raw":   # the fast exact form against the literal loop: first / last rows, a middle band
        for a, b in ((0, 14), (533, 547), (H - 14, H)):
            lo, hi = max(a - 7, 0), min(b + 7, H)
            want = O.ssd_threaded(left[lo:hi], right[lo:hi], D, 7)[a - lo:b - lo]
            assert np.array_equal(ocost[a:b], want)

    oagg = O.sgm(ocost, left, 0.1, 0.2)
    owta = O.wta(oagg)
    oc = cuda(ocost)
    default = sm.aggregation.Semiglobal(0.1, 0.2)
    assert default._use_fused_sweeps(oc)                    # the schedule the bench times
    agg = default(oc, l)
    err = volume_close(agg, oagg)
    m = mismatch(sm.disparity_reduce.WinnerTakesAll()(agg), owta)
    print(f"1080p D=128 {scale} fused sweeps: max rel err {err:.2e}, WTA mismatch {m:.2e}")
    assert m <= MAX_MISMATCH
    agg = sm.aggregation.Semiglobal(0.1, 0.2, fused_sweeps=False)(oc, l, agg)
    err = volume_close(agg, oagg)
    m = mismatch(sm.disparity_reduce.WinnerTakesAll()(agg), owta)
    print(f"1080p D=128 {scale} per-direction passes: max rel err {err:.2e}, WTA mismatch {m:.2e}")
    assert m <= MAX_MISMATCH
    del agg
    assert np.array_equal(sm.disparity_reduce.WinnerTakesAll()(cuda(oagg)).cpu().numpy(), owta)
    del oagg

    # end to end: Pipeline.estimate = SSD kernel + folded argmin, sums not written back
    pipe = sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal(0.1, 0.2))
    m = mismatch(pipe.estimate(l, r), owta)
    print(f"1080p D=128 {scale} Pipeline.estimate vs oracle chain: {m:.2e} of pixels differ")
    assert m <= MAX_MISMATCH
    del pipe

    # config 4: SSD + DP, both modes, maps bit-exact given the identical volume
    for mode in ("cuda", "cpu_exact"):
        dp = sm.disparity_reduce.DynamicProgramming(mode=mode)
        odisp, opath, ofinal = O.dp(ocost, mode, return_aux=True)
        assert np.array_equal(dp(oc).cpu().numpy(), odisp), mode
        assert np.array_equal(dp._row_final_costs.cpu().numpy(), ofinal), mode
        p = dp._path_volume.cpu().numpy()
        assert np.array_equal(p if mode == "cpu_exact" else p[:, 1:], opath if mode == "cpu_exact" else opath[:, 1:])
        del dp, p
    pipe = sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.DynamicProgramming())
    m = mismatch(pipe.estimate(l, r), O.dp(ocost, "cuda"))
    print(f"1080p D=128 {scale} Pipeline SSD+DP vs oracle chain: {m:.2e}")
    assert m <= (0.0 if scale == "raw" else MAX_MISMATCH)


# ------------------------------------------------- config 3: 1080 x 1920, D = 128, Birchfield
@pytest.mark.parametrize("scale", ["unit", "raw"])
def test_1080p_d128_birchfield_sgm_wta_vs_oracle(sm, scale):
    H, W, D = 1080, 1920, 128
    left, right, gt = synth.random_dot_pair(H, W, D, seed=1000, scale=scale)
    l, r = cuda(left), cuda(right)
    ocost = O.birchfield_threaded(left, right, D)
    cost = sm.cost.Birchfield(D)(l, r)
    volume_close(cost, ocost, exact=True)
    p1, p2 = (0.1, 0.2) if scale == "unit" else (10.0, 40.0)
    oagg = O.sgm(ocost, left, p1, p2)
    owta = O.wta(oagg)
    err = volume_close(sm.aggregation.Semiglobal(p1, p2)(cost, l), oagg)
    del oagg
    pipe = sm.Pipeline(sm.cost.Birchfield(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal(p1, p2))
    m = mismatch(pipe.estimate(l, r), owta)
    print(f"1080p D=128 {scale} Birchfield+SGM+WTA: sums max rel err {err:.2e}, map mismatch {m:.2e}")
    assert m <= MAX_MISMATCH


# ------------------------------------------------------ config 5: 1080 x 1920, D = 192 frames
@pytest.mark.parametrize("seed", [1000, 1063])
def test_1080p_d192_frame_of_the_video_batch_vs_oracle(sm, seed):
    H, W, D = 1080, 1920, 192
    left, right, gt = synth.random_dot_pair(H, W, D, seed=seed, scale="raw")
    l, r = cuda(left), cuda(right)
    ocost = O.ssd_exact_int(left, right, D, 7)
    cost = sm.cost.SSD(D)(l, r)
    volume_close(cost, ocost, exact=True)
    oagg = O.sgm(ocost, left, 0.1, 0.2)
    owta = O.wta(oagg)
    default = sm.aggregation.Semiglobal(0.1, 0.2)
    assert default._use_fused_sweeps(cost)
    err = volume_close(default(cost, l), oagg)
    del oagg
    pipe = sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal(0.1, 0.2))
    m = mismatch(pipe.estimate(l, r), owta)
    print(f"1080p D=192 seed {seed}: sums max rel err {err:.2e}, Pipeline map mismatch {m:.2e}")
    assert m <= MAX_MISMATCH
    valid = np.arange(W)[None, :] >= gt + 7
    assert (owta[valid] == gt[valid]).mean() > 0.99


# ------------------------------------------------- beyond the named shapes: 3840 x 2160, D = 64
def test_4k_d64_two_sweep_ctas_per_sm_vs_oracle(sm):
    """3840 columns are more than 13 column slots x 148 SMs: the fused sweeps run their two-CTAs-
    per-SM build (64 registers, lane maps with <= 2 disparities per lane; csrc/sgm_vsweep.cu) instead
    of falling back to eight direction passes.  Same bars as the named shapes; the number of passes
    the library timed says which schedule ran."""
    from stereomatch._cstereomatch import AggregationOps
    H, W, D = 2160, 3840, 64
    left, right, gt = synth.random_dot_pair(H, W, D, seed=4000, scale="raw")
    l, r = cuda(left), cuda(right)
    ocost = O.ssd_exact_int(left, right, D, 7)
    cost = sm.cost.SSD(D)(l, r)
    volume_close(cost, ocost, exact=True)
    oagg = O.sgm(ocost, left, 0.1, 0.2)
    owta = O.wta(oagg)
    del ocost
    default = sm.aggregation.Semiglobal(0.1, 0.2)
    assert default._use_fused_sweeps(cost)
    AggregationOps.pass_timing(True)
    agg = default(cost, l)
    passes = AggregationOps.last_pass_ms()
    plain = sm.aggregation.Semiglobal(0.1, 0.2, overlap=False)
    agg_plain = plain(cost, l)
    passes_plain = AggregationOps.last_pass_ms()
    AggregationOps.pass_timing(False)
    assert len(passes) == 2 and len(passes_plain) == 4, (passes, passes_plain)   # not the 8 direction passes
    err = volume_close(agg, oagg)
    err_plain = volume_close(agg_plain, oagg)
    del oagg, agg, agg_plain
    pipe = sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(), sm.aggregation.Semiglobal(0.1, 0.2))
    m = mismatch(pipe.estimate(l, r), owta)
    print(f"4K D=64: sums max rel err {err:.2e} (overlapped) {err_plain:.2e} (plain fused), Pipeline map "
          f"mismatch {m:.2e}; passes ms {[round(t, 3) for _, t in passes]} / {[round(t, 3) for _, t in passes_plain]}")
    assert m <= MAX_MISMATCH
    valid = np.arange(W)[None, :] >= gt + 7
    assert (owta[valid] == gt[valid]).mean() > 0.99
