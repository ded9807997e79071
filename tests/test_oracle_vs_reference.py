"""The oracle against the compiled reference itself (only where oracle/_ref was built,
i.e. in the build container and on boxes that received the .so)."""
import numpy as np
import pytest

from oracle import oracle as O, ref_ops as R, synth

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built")


@pytest.mark.parametrize("H,W,D,scale", [(21, 37, 16, "raw"), (37, 21, 16, "unit"),
                                           (32, 44, 32, "unit")])
def test_pipeline_stages_bit_exact(H, W, D, scale):
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H * W, scale=scale)
    cost = O.ssd(left, right, D, 7)
    assert np.array_equal(cost, R.ssd(left, right, D, 7))
    agg = O.sgm(cost, left, 0.1, 0.2)
    assert np.array_equal(agg, R.sgm(cost, left, 0.1, 0.2))
    assert np.array_equal(O.wta(agg), R.wta(agg))
    if D <= H <= W:
        assert np.array_equal(O.dp(cost, "cpu_exact"), R.dp(cost))


def test_wta_adversarial():
    vol = synth.tie_heavy_volume(11, 17, 64, seed=3)
    vol[2, 3, :] = np.inf
    vol[2, 4, 0] = np.nan
    vol[2, 5, 17] = np.nan
    assert np.array_equal(O.wta(vol), R.wta(vol))
