"""The library's per-device and per-thread state (ADVICE): the > 48 KB
dynamic-shared-memory opt-in, the SM count and the occupancy caches are per DEVICE; the helper
streams of the concurrent SGM schedule are shared per device behind a mutex; error messages and
the host-pipeline scratch are per THREAD.  include/stm_abi.h states the contract."""
import threading

import numpy as np
import pytest
import torch

from oracle import oracle as O, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sm():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import stereomatch
    return stereomatch


def _frame(seed, H=60, W=100, D=64):
    left, right, _ = synth.random_dot_pair(H, W, D, seed=seed, scale="unit")
    return left, right


def _pipelines(sm, D):
    """One pipeline per schedule: concurrent directions (small volume default), sequential
    passes, fused sweeps, and DP."""
    mk = lambda **kw: sm.Pipeline(sm.cost.SSD(D), sm.disparity_reduce.WinnerTakesAll(),
                                  sm.aggregation.Semiglobal(**kw))
    return [mk(), mk(concurrent_directions=False), mk(fused_sweeps=True),
            sm.Pipeline(sm.cost.Birchfield(D), sm.disparity_reduce.DynamicProgramming())]


def test_threads_on_one_device_share_the_library(sm):
    """Four host threads, each with its own stream and pipelines, hammer one device at the same
    time (the concurrent-direction schedule forks onto helper streams shared per device); every
    result equals the single-threaded one."""
    D = 64
    frames = [_frame(300 + i) for i in range(4)]
    want = []
    for l, r in frames:
        lc, rc = torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()
        want.append([p.estimate(lc, rc).cpu().clone() for p in _pipelines(sm, D)])
    errors, results = [], [None] * 4

    def work(i):
        try:
            torch.cuda.set_device(0)
            stream = torch.cuda.Stream()
            l, r = frames[i]
            with torch.cuda.stream(stream):
                lc, rc = torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()
                pipes = _pipelines(sm, D)
                got = None
                for _ in range(25):
                    got = [p.estimate(lc, rc).clone() for p in pipes]
                stream.synchronize()
            results[i] = [g.cpu() for g in got]
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for i in range(4):
        for got, w in zip(results[i], want[i]):
            assert torch.equal(got, w)


def test_error_messages_are_per_thread(sm):
    from stereomatch import _cstereomatch as C
    seen = {}

    def bad(key, H):
        rc = C._lib.stm_wta_f32(None, None, H, 4, 4, None)
        seen[key] = (rc, C._lib.stm_last_error().decode())

    t = threading.Thread(target=bad, args=("other", -7))
    bad("main", -3)
    t.start()
    t.join()
    assert seen["main"][0] == -1 and "H=-3" in seen["main"][1]
    assert seen["other"][0] == -1 and "H=-7" in seen["other"][1]
    assert "H=-3" in C._lib.stm_last_error().decode()        # the other thread did not overwrite it


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices in one process")
def test_two_devices_from_one_process(sm):
    """Every schedule on cuda:0 then cuda:1 then cuda:0 again from one process, and from two threads
    at once: the kernels that need > 48 KB of dynamic shared memory (ssd_tiled, sgm_row, the
    fused sweeps) must be opted in on EACH device."""
    D = 64
    l, r = _frame(77, H=70, W=150, D=D)
    want = None
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        with torch.cuda.device(dev):
            lc, rc = torch.from_numpy(l).to(dev), torch.from_numpy(r).to(dev)
            got = [p.estimate(lc, rc).cpu() for p in _pipelines(sm, D)]
        if want is None:
            want = got
        assert all(torch.equal(a, b) for a, b in zip(got, want)), dev
    ovol = O.sgm(O.ssd(l, r, D, 7), l, 0.1, 0.2)
    assert float((want[0].numpy() != O.wta(ovol)).mean()) <= 1e-3
    errors = []

    def work(dev):
        try:
            torch.cuda.set_device(dev)
            lc, rc = torch.from_numpy(l).to(dev), torch.from_numpy(r).to(dev)
            for _ in range(10):
                got = [p.estimate(lc, rc).cpu() for p in _pipelines(sm, D)]
            assert all(torch.equal(a, b) for a, b in zip(got, want))
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=work, args=(d,)) for d in ("cuda:0", "cuda:1")]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
