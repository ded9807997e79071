"""stm_sgm_static_f32 (include/stm_abi.h): the aggregation entry point that allocates nothing --
its scratch is a file-scope `static __device__` array of the library.  Called straight through the
C ABI (ctypes) and compared with the existing entry point stm_sgm_f32 (STM_SGM_SEQUENTIAL: the
same schedule with a caller-provided workspace) within the suite's tolerance for SGM sums (1e-4
relative, identical inf pattern), on two calls in a row and on a second, different frame (the
array keeps the last call's contents).  Collected last on purpose (the round's end runs -x)."""
import ctypes
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "stereo-matching_b200" / "stereomatch" / "libstm_b200.so"
RTOL = 1e-4
STM_SGM_MIN_ALL_D, STM_SGM_OVERWRITE, STM_SGM_SEQUENTIAL = 1, 2, 16
MAX_FLOATS = 524288        # STM_SGM_STATIC_MAX_FLOATS


@pytest.fixture(scope="module")
def lib():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    lib = ctypes.CDLL(str(LIB))
    vp, i, f, u, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_uint, ctypes.c_size_t
    lib.stm_last_error.restype = ctypes.c_char_p
    lib.stm_sgm_workspace_bytes.restype = sz
    lib.stm_sgm_workspace_bytes.argtypes = [i, i, i]
    lib.stm_sgm_f32.argtypes = [vp, vp, vp, i, i, i, f, f, u, vp, sz, vp]
    lib.stm_sgm_static_f32.argtypes = [vp, vp, vp, i, i, i, f, f, u, vp]
    lib.stm_sgm_f32.restype = lib.stm_sgm_static_f32.restype = i
    return lib


def make_inputs(H, W, D, seed):
    g = torch.Generator().manual_seed(seed)
    cost = (torch.rand(H, W, D, generator=g) * 50.0).cuda()
    dd = torch.arange(D, device="cuda")[None, None, :]
    xx = torch.arange(W, device="cuda")[None, :, None]
    cost = torch.where(dd > xx, torch.full_like(cost, float("inf")), cost).contiguous()   # d > x, like SSD
    image = (torch.randint(0, 256, (H, W), generator=g).float() / 255.0).cuda().contiguous()
    return cost, image


def existing(lib, cost, image, flags):
    H, W, D = cost.shape
    n = int(lib.stm_sgm_workspace_bytes(H, W, D))
    ws = torch.empty(max(n, 16), dtype=torch.uint8, device="cuda")
    out = torch.full_like(cost, 3.0)
    rc = lib.stm_sgm_f32(cost.data_ptr(), image.data_ptr(), out.data_ptr(), H, W, D, 0.1, 0.2,
                         flags | STM_SGM_SEQUENTIAL, ws.data_ptr(), ws.numel(), None)
    assert rc == 0, lib.stm_last_error().decode()
    torch.cuda.synchronize()
    return out


def close(got, want):
    fin = torch.isfinite(want)
    assert bool((torch.isfinite(got) == fin).all()), "inf pattern differs"
    den = torch.maximum(got[fin].abs(), want[fin].abs()).clamp_min(1e-30)
    err = float(((got[fin] - want[fin]).abs() / den).max())
    assert err <= RTOL, err
    return err


@pytest.mark.parametrize("H,W,D,flags", [(60, 100, 64, STM_SGM_OVERWRITE),          # landscape: corner replay of 5 and 7
                                          (100, 60, 32, STM_SGM_OVERWRITE),          # portrait: replay before direction 4
                                          (48, 48, 128, 0),                          # square, accumulating into `out`
                                          (33, 70, 20, STM_SGM_OVERWRITE | STM_SGM_MIN_ALL_D)])   # padded D
def test_static_scratch_entry_point_matches_the_workspace_one(lib, H, W, D, flags):
    for seed in (5, 6):                                   # a second frame over the first one's leftovers
        cost, image = make_inputs(H, W, D, seed)
        want = existing(lib, cost, image, flags)
        for call in range(2):                             # two calls in a row
            out = torch.full_like(cost, 3.0)
            before = torch.cuda.memory_allocated()
            rc = lib.stm_sgm_static_f32(cost.data_ptr(), image.data_ptr(), out.data_ptr(), H, W, D,
                                        0.1, 0.2, flags, None)
            assert rc == 0, lib.stm_last_error().decode()
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before
            err = close(out, want)
            assert torch.equal(out[torch.isfinite(want)], want[torch.isfinite(want)]), \
                f"same schedule, same kernels: expected bit-identical sums (max rel err {err})"


def test_static_scratch_entry_point_rejects_what_does_not_fit(lib):
    H, W, D = 2100, 2100, 256                             # 2100 * 256 = 537,600 floats > 524,288
    assert H * D > MAX_FLOATS
    rc = lib.stm_sgm_static_f32(1 << 20, 1 << 20, 2 << 20, H, W, D, 0.1, 0.2, STM_SGM_OVERWRITE, None)
    assert rc == -1
    msg = lib.stm_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "524288" in msg
    rc = lib.stm_sgm_static_f32(1 << 20, 1 << 20, 2 << 20, 8, 8, 32, 0.1, 0.2, 4, None)   # FUSED_SWEEPS: not here
    assert rc == -1 and "unknown flags" in lib.stm_last_error().decode()
    rc = lib.stm_sgm_static_f32(None, None, None, 8, 8, 32, 0.1, 0.2, 0, None)
    assert rc == -1 and "null pointer" in lib.stm_last_error().decode()
