"""The fused sweeps give up cooperatively instead of trapping (csrc/sgm_vsweep.cu, SpinGuard).

A sweep is ~150-300 CTAs that hand vectors to each other through rings; if one of them stopped,
the others would poll forever.  Round 1 ended that with __trap() after 60 s, which takes the host
application's CUDA context with it.  Now a polling loop that sees no progress raises an abort word
that every other wait of the launch observes, the kernel ENDS, the host can ask
(stm_sgm_sweep_aborted), and the next call works.  The fault is injected through the developer hook
stm_debug_vsweep_fault: one CTA's exporter stops at a given row, the time limit is set to 300 ms.
"""
import ctypes
import time

import numpy as np
import pytest
import torch

from oracle import oracle as O, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("H,W,D,cta,row", [(96, 640, 64, 0, 8), (80, 1920, 128, 77, 30), (60, 3840, 64, 200, 3)])
def test_sweep_gives_up_reports_and_the_context_survives(H, W, D, cta, row):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import stereomatch
    from stereomatch._cstereomatch import AggregationOps, _lib
    _lib.stm_debug_vsweep_fault.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint]
    left, right, _ = synth.random_dot_pair(H, W, D, seed=W + D, scale="unit")
    cost = O.ssd(left, right, D, 3)
    want = O.sgm(cost, left, 0.1, 0.2)
    c, l = torch.from_numpy(cost).cuda(), torch.from_numpy(left).cuda()
    sgm = stereomatch.aggregation.Semiglobal(0.1, 0.2, fused_sweeps=True, overlap=False)
    assert AggregationOps.sweep_aborted() is None
    try:
        _lib.stm_debug_vsweep_fault(cta, row, 300)
        t0 = time.time()
        sgm(c, l)                       # down sweep starves, gives up; the up sweep runs on garbage
        torch.cuda.synchronize()        # returns: nothing hangs, nothing traps
        took = time.time() - t0
    finally:
        _lib.stm_debug_vsweep_fault(-1, 0, 0)
    assert took < 20.0
    info = AggregationOps.sweep_aborted()
    assert info is not None, "the host was not told"
    print(f"{H}x{W}x{D}: fault at CTA {cta} row {row} -> gave up after {took:.2f} s, record {info}")
    assert AggregationOps.sweep_aborted() is None          # collected once
    # the context is alive and the same call is right again
    got = sgm(c, l).cpu().numpy()
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin)
    assert (np.abs(got[fin] - want[fin]) <= 1e-4 * np.abs(want[fin])).all()
    assert AggregationOps.sweep_aborted() is None


def test_uncollected_record_fails_the_next_sweep_call():
    import stereomatch
    from stereomatch._cstereomatch import AggregationOps, _lib
    _lib.stm_debug_vsweep_fault.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint]
    H, W, D = 64, 512, 64
    left, right, _ = synth.random_dot_pair(H, W, D, seed=5, scale="unit")
    c, l = torch.from_numpy(O.ssd(left, right, D, 3)).cuda(), torch.from_numpy(left).cuda()
    sgm = stereomatch.aggregation.Semiglobal(0.1, 0.2, fused_sweeps=True, overlap=False)
    try:
        _lib.stm_debug_vsweep_fault(1, 5, 300)
        sgm(c, l)
        torch.cuda.synchronize()
    finally:
        _lib.stm_debug_vsweep_fault(-1, 0, 0)
    with pytest.raises(RuntimeError, match="gave up"):
        sgm(c, l)
    sgm(c, l)                                               # the record was consumed by the error
    torch.cuda.synchronize()
    assert AggregationOps.sweep_aborted() is None
