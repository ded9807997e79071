"""Pins that only the reference's own CUDA branch can give, taken on the box.

oracle/_ref/_cstereomatch.so is the UNMODIFIED reference compiled from /root/reference by
oracle/build_ref.py; its fatbin holds sm_100a cubins, so on the B200 box its CUDA ops run.  Three
parts of the contract are defined by kernels that have no CPU twin in the reference:

  * DP default mode ("cuda", Jacobi): reference ExpandPathsKernelCUDA, src/dynamic_programming.cu:83-130.
    Its shared-memory recurrence has no barrier between the reads and the write of a column
    (the __syncthreads are commented out, :98,:126), so it is only well defined while one warp
    runs in lock step: D <= 32.  There O.dp(vol, "cuda") must equal it bit for bit.
  * canonical Birchfield (zero-padded scanlines): reference BirchfieldKernelCUDA,
    src/birchfield_cost.cu:83-139, launched with blockDim.x = W, i.e. W <= 1024 only.
  * SGM with min over all D: reference SemiglobalKernel, src/semiglobal_gpu.cu:43-106, for
    power-of-two D.  All directions of one launch accumulate with a non-atomic `+=` from
    different blocks (:62,:97), so updates can be lost: compared with an allowance.

Skipped when oracle/_ref is absent.  Everything else on the path is pinned by the reference's
CPU ops (tests/test_oracle_vs_reference.py, tests/golden/).
"""
import numpy as np
import pytest
import torch

from oracle import oracle as O, ref_ops, synth

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_ops.available(), reason="oracle/_ref/_cstereomatch.so not built")]


@pytest.fixture(scope="module")
def sm():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import stereomatch
    return stereomatch


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ------------------------------------------------------------------------------------ DP
@pytest.mark.parametrize("H,W,D,kind", [(40, 96, 8, "rand"), (48, 128, 16, "ssd"), (64, 200, 32, "ssd"),
                                          (33, 77, 32, "ties"), (130, 333, 32, "rand"), (20, 64, 4, "ties")])
def test_dp_jacobi_equals_reference_cuda_kernel_single_warp(sm, H, W, D, kind):
    """D <= 32: the reference kernel's block is one warp, its smem reads precede the write in
    lock step -- the read-before-write resolution O.dp(.., "cuda") restates.  Forward costs and
    back-pointers (columns >= 1) equal bit for bit; the maps equal on the rows the reference's
    start-up clobber `path[disp][row][0] = 0` (:102, a cross-block race on path[y < D][x < H][0])
    cannot reach, and everywhere when it did not change a back-pointer that is followed."""
    if kind == "ssd":
        left, right, _ = synth.random_dot_pair(H, W, D, seed=H + D, scale="raw")
        vol = O.ssd(left, right, D, 3)
    elif kind == "ties":
        vol = synth.tie_heavy_volume(H, W, D, seed=D)
    else:
        vol = synth.random_volume(H, W, D, seed=D)
    assert D <= H <= W                                  # the clobber stays inside the tensor
    rdisp, rpath, rfinal = ref_ops.dp(vol, device="cuda", return_aux=True)
    odisp, opath, ofinal = O.dp(vol, "cuda", return_aux=True)
    assert np.array_equal(rfinal, ofinal)               # the whole forward recurrence
    clobber = np.zeros((H, W, D), bool)
    clobber[:D, :H, 0] = True                           # path[disp][row][0], disp < D, row < H
    keep = ~clobber
    keep[:, 0, :] = False                               # column 0 is never written by the reference
    assert np.array_equal(rpath[keep], opath[keep])
    assert np.array_equal(rdisp[D:], odisp[D:])
    print(f"DP {H}x{W}x{D} {kind}: rows < D differing from the race-free restatement: "
          f"{(rdisp[:D] != odisp[:D]).any(axis=1).sum()} of {D}")
    # and the product kernel is that restatement (bit-exact), hence the reference kernel
    dp = sm.disparity_reduce.DynamicProgramming()
    got = dp(cuda(vol)).cpu().numpy()
    assert np.array_equal(got, odisp)
    assert np.array_equal(got[D:], rdisp[D:])
    assert np.array_equal(dp._row_final_costs.cpu().numpy(), rfinal)
    assert np.array_equal(dp._path_volume.cpu().numpy()[keep], rpath[keep])


# ---------------------------------------------------------------------------- Birchfield
@pytest.mark.parametrize("H,W,D,scale", [(21, 45, 16, "raw"), (21, 45, 32, "unit"), (12, 150, 128, "unit"),
                                          (375, 450, 64, "unit"), (375, 450, 64, "raw"), (6, 1024, 64, "unit"),
                                          (9, 33, 20, "unit"), (5, 7, 4, "unit")])
def test_birchfield_equals_reference_cuda_kernel(sm, H, W, D, scale):
    """Every entry the reference CUDA kernel writes (d <= x), both borders and the last row
    included, bit for bit; what it leaves unwritten (d > x) is +inf here (decision BT-1)."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=3 * H + W, scale=scale)
    ref = ref_ops.birchfield(left, right, D, device="cuda")     # NaN-prefilled
    written = ~np.isnan(ref)
    dd, xx = np.arange(D)[None, None, :], np.arange(W)[None, :, None]
    assert np.array_equal(written, np.broadcast_to(dd <= xx, ref.shape))
    canon = O.birchfield(left, right, D)
    assert np.array_equal(canon[written], ref[written])
    got = sm.cost.Birchfield(D)(cuda(left), cuda(right)).cpu().numpy()
    assert np.array_equal(got[written], ref[written])
    assert np.all(np.isposinf(got[~written]))


# ------------------------------------------------------------------------------------ SGM
@pytest.mark.parametrize("H,W,D", [(48, 48, 32), (40, 72, 64), (60, 100, 128), (96, 64, 32), (375, 450, 64)])
def test_sgm_min_over_all_d_against_reference_cuda_kernel(sm, H, W, D):
    """STM_SGM_MIN_ALL_D (K = {0..D-1}) is the reference CUDA kernel's rule.  That kernel loses
    `+=` updates when two blocks hit a pixel together (F7), which only ever makes an entry
    SMALLER by one direction's term, so: the bulk of the finite entries agree to 1e-4, no entry
    of the reference exceeds ours beyond the bar, and with K = {0..D-2} (the CPU reference's
    rule) the agreement is measurably worse wherever the minimum sits at d = D-1."""
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H * 7 + W, scale="unit")
    cost = O.ssd(left, right, D, 3)
    ref = ref_ops.sgm(cost, left, 0.1, 0.2, device="cuda")
    want = O.sgm(cost, left, 0.1, 0.2, min_over_all_d=True)
    got = sm.aggregation.Semiglobal(0.1, 0.2, min_over_all_d=True)(cuda(cost), cuda(left)).cpu().numpy()
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin)
    den = np.maximum(np.abs(want[fin]), 1e-30)
    assert (np.abs(got[fin] - want[fin]) / den).max() <= 1e-4        # product == restatement
    assert np.array_equal(np.isfinite(ref), fin)
    rel = (ref[fin] - want[fin]) / den
    agree = float((np.abs(rel) <= 1e-4).mean())
    lost = float((rel < -1e-4).mean())
    above = float((rel > 1e-4).mean())
    print(f"SGM MIN_ALL_D {H}x{W}x{D} vs reference CUDA: agree {agree:.5f}, lost updates {lost:.2e}, "
          f"above {above:.2e}")
    assert above == 0.0          # a lost update can only remove a term
    assert agree >= 0.90         # allowance for the reference's race (measured rate is printed)
