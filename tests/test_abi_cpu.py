"""CPU-side checks: the C-ABI library loads and exports every symbol include/stm_abi.h
declares, the Python API mirrors the reference's, and nothing computes without a GPU."""
import ctypes
import re
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]
HEADER = ROOT / "include" / "stm_abi.h"
LIB = ROOT / "stereo-matching_b200" / "stereomatch" / "libstm_b200.so"


def declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(stm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    assert LIB.exists(), "run python stereo-matching_b200/build.py"
    lib = ctypes.CDLL(str(LIB))
    names = declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), n


def test_version_and_error_strings():
    lib = ctypes.CDLL(str(LIB))
    lib.stm_version.restype = ctypes.c_char_p
    lib.stm_last_error.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.stm_version()
    # bad shape is rejected before any CUDA call
    rc = lib.stm_wta_f32(None, None, 0, 4, 4, None)
    assert rc == -1 and b"invalid shape" in lib.stm_last_error()


def test_workspace_query_is_pure():
    lib = ctypes.CDLL(str(LIB))
    lib.stm_sgm_workspace_bytes.restype = ctypes.c_size_t

    def expect(H, W, dpad, sweep_dpad=None):
        # max of csrc/sgm.cu (one path's vectors for the corner-diagonal replay) and
        # csrc/sgm_vsweep.cu (320 imported-row counters of 128 B + 320 CTA->CTA rings of
        # 16 rows x 3 vectors of 8-byte (value, tag) words; only for the lane maps it handles).
        # sweep_dpad: the sweeps run ranges that are a multiple of 4 below 128 on the masked 4-wide map
        direction = min(H, W) * dpad * 4
        sd = dpad if sweep_dpad is None else sweep_dpad
        # + the adaptive penalties of both sweeps, one float4 per pixel each
        sweeps = (40960 + 320 * 16 * 3 * sd * 8 + 2 * H * W * 16
                  if (H >= 2 and W >= 2 and sd in (32, 64, 96, 128, 160, 192, 256)) else 0)
        return max(direction, sweeps)

    assert lib.stm_sgm_workspace_bytes(1080, 1920, 128) == expect(1080, 1920, 128)
    assert lib.stm_sgm_workspace_bytes(1920, 1080, 128) == expect(1920, 1080, 128)  # portrait
    assert lib.stm_sgm_workspace_bytes(0, 0, 0) == 0
    # volumes up to 48 MB also get one buffer per direction (csrc/sgm.cu,
    # concurrent directions), after the base workspace rounded up to 256 bytes
    a256 = lambda n: (n + 255) // 256 * 256
    for H, W, D, dpad in ((375, 450, 64, 64), (20, 30, 32, 32), (64, 64, 128, 128), (450, 375, 64, 64),
                          (30, 18, 16, 32)):   # portrait too; D=16 pads to 32 lanes
        assert lib.stm_sgm_workspace_bytes(H, W, D) == a256(expect(H, W, dpad)) + 8 * H * W * D * 4
    assert lib.stm_sgm_workspace_bytes(480, 640, 128) == expect(480, 640, 128)   # 157 MB: the minimum is sequential
    # the overlapped schedule (STM_SGM_FUSED_SWEEPS 4 | STM_SGM_OVERLAP 32): one more volume after the
    # base workspace, two with STM_SGM_DISPARITY_ONLY (8) at 128 padded disparities; none at D > 128
    lib.stm_sgm_workspace_bytes_flags.restype = ctypes.c_size_t
    H, W = 1080, 1920
    for D, dpad, flags, volumes in ((128, 128, 36, 1), (128, 128, 44, 2), (64, 64, 44, 1),
                                    (192, 192, 44, 0), (128, 128, 4, 0), (128, 128, 32, 0)):
        want = expect(H, W, dpad) if volumes == 0 else a256(expect(H, W, dpad)) + volumes * H * W * D * 4
        assert lib.stm_sgm_workspace_bytes_flags(H, W, D, flags) == want, (D, flags)
    # mid-size volumes (48..384 MB) that do not take the fused sweeps: the flags query asks for the eight
    # direction buffers of the concurrent schedule; not with STM_SGM_SEQUENTIAL (16), not when the sweeps run
    for H, W, D in ((375, 450, 128), (480, 640, 64), (720, 1280, 64)):
        base = expect(H, W, D)
        assert lib.stm_sgm_workspace_bytes(H, W, D) == base
        assert lib.stm_sgm_workspace_bytes_flags(H, W, D, 1) == a256(base) + 8 * H * W * D * 4
        assert lib.stm_sgm_workspace_bytes_flags(H, W, D, 1 | 16) == base
        assert lib.stm_sgm_workspace_bytes_flags(H, W, D, 1 | 4) == base
    # D = 224 is not a lane map of the sweeps: the flag does not change what the call needs
    assert lib.stm_sgm_workspace_bytes_flags(400, 900, 224, 1 | 4) == a256(expect(400, 900, 224)) + 8 * 400 * 900 * 224 * 4
    assert lib.stm_sgm_workspace_bytes_flags(540, 960, 96, 1 | 4) == expect(540, 960, 96, 128)   # D = 96: the sweeps run (masked 4-wide map)
    assert lib.stm_sgm_workspace_bytes_flags(540, 960, 100, 1 | 4) == expect(540, 960, 128, 128)
    assert lib.stm_sgm_workspace_bytes_flags(1080, 1920, 64, 1) == expect(1080, 1920, 64)   # 531 MB: sequential


def test_python_api_surface_matches_reference():
    import stereomatch
    from stereomatch import aggregation, cost, disparity_reduce, pipeline
    ssd = cost.SSD(64)
    assert (ssd.max_disparity, ssd.kernel_size, ssd.cost_volume_dtype) == (64, 7, torch.float32)
    ssd.max_disparity = 32  # attributes stay mutable (reference workflow.py:34)
    bt = cost.Birchfield(32)
    assert bt.max_disparity == 32 and bt.kerne_size == 7
    sgm = aggregation.Semiglobal()
    assert (sgm.penalty1, sgm.penalty2) == (0.1, 0.2)
    p = pipeline.Pipeline(ssd, disparity_reduce.WinnerTakesAll(), sgm)
    assert p.cost is ssd and p.aggregation is sgm
    assert hasattr(disparity_reduce.DynamicProgramming(), "__call__")
    assert stereomatch.Pipeline is pipeline.Pipeline


def test_no_cpu_fallback():
    from stereomatch import aggregation, cost, disparity_reduce
    img = torch.zeros(8, 8)
    vol = torch.zeros(8, 8, 4)
    with pytest.raises(RuntimeError, match="gpu tensor"):
        cost.SSD(4)(img, img)
    with pytest.raises(RuntimeError, match="gpu tensor"):
        cost.Birchfield(4)(img, img)
    with pytest.raises(RuntimeError, match="gpu tensor"):
        aggregation.Semiglobal()(vol, img)
    with pytest.raises(RuntimeError, match="gpu tensor"):
        disparity_reduce.WinnerTakesAll()(vol)
    with pytest.raises(RuntimeError, match="gpu tensor"):
        disparity_reduce.DynamicProgramming()(vol)


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without a GPU")
def test_compute_call_fails_loudly_without_gpu():
    lib = ctypes.CDLL(str(LIB))
    lib.stm_last_error.restype = ctypes.c_char_p
    buf = (ctypes.c_float * 64)()
    out = (ctypes.c_int32 * 16)()
    rc = lib.stm_wta_f32(buf, out, 4, 4, 4, None)
    assert rc == -2 and b"no CPU fallback" in lib.stm_last_error()


def test_numeric_known_answers():
    from stereomatch.numeric import is_power_of_two, next_power_of_2
    # reference tests/test_numeric.py:11-17
    for n, e in [(0, 1), (1, 1), (2, 2), (3, 4), (4, 4), (50, 64), (127, 128)]:
        assert next_power_of_2(n) == e
    assert is_power_of_two(64) and not is_power_of_two(0) and not is_power_of_two(192)


def test_buffer_reuse_rules():
    from stereomatch._common import empty_tensor, zeros_tensor_like
    a = torch.ones(3, 4)
    assert empty_tensor(3, 4, reuse_tensor=a) is a
    assert empty_tensor(3, 5, reuse_tensor=a) is not a
    assert empty_tensor(3, 4, dtype=torch.int32, reuse_tensor=a).dtype == torch.int32
    z = zeros_tensor_like(a, reuse_tensor=a)
    assert z is a and float(a.sum()) == 0.0
    assert zeros_tensor_like(a, reuse_tensor=torch.ones(2, 2)).shape == (3, 4)
