"""CUDA textures as SSD inputs: the reference's own texture tests restated
(reference tests/test_texture.py:11-120, tests/test_cost.py:41-57, tests/test_pipeline.py:24)
plus parity of the texture overload of compute_ssd against the oracle."""
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


def test_initialization(sm):
    """reference tests/test_texture.py:11-28"""
    from stereomatch.cuda_texture import CUDATexture
    tex1 = CUDATexture.from_tensor(torch.rand(100, 100, dtype=torch.float32, device="cuda:0"))
    assert (tex1.width, tex1.height, tex1.channels) == (100, 100, 1)
    assert tex1.cuda_texture_object != 0 and not tex1.use_normalized_coords and tex1.dtype == torch.float32
    tex2 = CUDATexture.from_tensor(torch.rand(60, 100, 4, dtype=torch.float32, device="cuda:0"))
    assert (tex2.width, tex2.height, tex2.channels) == (100, 60, 4)
    assert "width=100, height=60, channels=4" in str(tex2)
    tex2.release()
    assert tex2.is_empty() and tex2.cuda_texture_object == 0
    tex2.release()                                    # idempotent


def test_rejected_inputs(sm):
    """reference tests/test_texture.py:31-37 and src/cuda_texture.cpp:24-50"""
    from stereomatch.cuda_texture import CUDATexture
    with pytest.raises(RuntimeError, match="3 or more than 4 channels"):
        CUDATexture.from_tensor(torch.rand(100, 100, 3))
    with pytest.raises(RuntimeError, match="more than 4 channels"):
        CUDATexture.from_tensor(torch.rand(10, 10, 5))
    with pytest.raises(RuntimeError, match="power of 2"):
        CUDATexture.from_tensor(torch.rand(100, 128), normalized_coords=True)
    with pytest.raises(RuntimeError, match="dtype not supported"):
        CUDATexture.from_tensor(torch.zeros(8, 8, dtype=torch.int32))
    tex = CUDATexture.from_tensor(torch.rand(8, 8))
    with pytest.raises(RuntimeError, match="non-empty"):
        tex.copy_from_tensor(torch.rand(8, 8, 1), False)


@pytest.mark.parametrize("shape", [(128, 128, 4), (100, 100), (37, 53, 2), (513, 123, 1)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.uint8])
@pytest.mark.parametrize("device", ["cpu", "cuda:0"])
def test_transfers_round_trip(sm, shape, dtype, device):
    """reference tests/test_texture.py:40-59, 99-120: tensor -> texture -> tensor, host and device
    sources, float and byte, 1 / 2 / 4 channels, odd sizes (array pitch != row bytes)."""
    from stereomatch.cuda_texture import CUDATexture
    t0 = (torch.rand(*shape) * 255.0).to(dtype).to(device)
    tex = CUDATexture.from_tensor(t0)
    assert (tex.height, tex.width) == tuple(shape[:2])
    t1 = tex.to_tensor()
    assert t1.dtype == dtype and not t1.is_cuda
    assert torch.equal(t1.reshape(t0.shape), t0.cpu())
    # through a kernel that reads tex2D at (col, row): channel 0
    out = torch.zeros(shape[0], shape[1], dtype=dtype, device="cuda:0")
    CUDATexture._run_transfer_test_kernel(tex, out)
    assert torch.equal(out.cpu(), t0.cpu().reshape(shape[0], shape[1], -1)[:, :, 0])


def test_bindings(sm):
    """reference tests/test_texture.py:84-96"""
    from stereomatch.cuda_texture import CUDATexture
    tensor = (torch.rand(128, 128, device="cuda:0") * 255.0).float()
    out = torch.zeros_like(tensor)
    CUDATexture._run_binding_test_kernel(tensor, out)
    assert torch.equal(tensor, out)


@pytest.mark.parametrize("H,W,D,k,scale", [(37, 53, 32, 7, "raw"), (37, 53, 32, 7, "unit"), (50, 31, 128, 7, "raw"),
                                            (23, 70, 192, 3, "unit"), (9, 11, 5, 2, "unit"), (375, 450, 64, 7, "raw")])
@pytest.mark.parametrize("source", ["cpu", "cuda"])
def test_ssd_texture_equals_ssd_and_the_oracle(sm, H, W, D, k, scale, source):
    """reference tests/test_cost.py:41-57 (cpu == gpu == gpu-tex), with the oracle as the cpu leg:
    bit-exact on the raw scale, 1e-4 on the unit scale, and bit-identical to the tensor overload."""
    from stereomatch.cuda_texture import CUDATexture
    left, right, _ = synth.random_dot_pair(H, W, D, seed=H + W, scale=scale)
    lt, rt = torch.from_numpy(left).to(source), torch.from_numpy(right).to(source)
    tex_l, tex_r = CUDATexture.from_tensor(lt), CUDATexture.from_tensor(rt)
    vol_tex = sm.cost.SSDTexture(D, k)(tex_l, tex_r)
    vol = sm.cost.SSD(D, k)(lt.cuda(), rt.cuda())
    assert torch.equal(vol_tex, vol)
    want = O.ssd_threaded(left, right, D, k)
    got = vol_tex.cpu().numpy()
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin) and np.all(np.isposinf(got[~fin]))
    if scale == "raw":
        assert np.array_equal(got[fin], want[fin])
    else:
        assert np.allclose(got[fin], want[fin], rtol=1e-4, atol=0)
    # buffer reuse rule (reference _common.py:9-12)
    again = sm.cost.SSDTexture(D, k)(tex_l, tex_r, cost_volume=vol_tex)
    assert again.data_ptr() == vol_tex.data_ptr()


def test_ssd_texture_errors(sm):
    from stereomatch.cuda_texture import CUDATexture
    f = CUDATexture.from_tensor(torch.rand(16, 16, device="cuda"))
    b = CUDATexture.from_tensor((torch.rand(16, 16) * 255).byte())
    n = CUDATexture.from_tensor(torch.rand(16, 16), normalized_coords=True)
    small = CUDATexture.from_tensor(torch.rand(8, 16))
    with pytest.raises(RuntimeError, match="float scalar"):     # cuda_texture.hpp:30-37
        sm.cost.SSDTexture(8)(f, b)
    with pytest.raises(RuntimeError, match="normalized"):       # cost.py:66-68
        sm.cost.SSDTexture(8)(n, n)
    with pytest.raises(RuntimeError, match="sizes differ"):
        sm.cost.SSDTexture(8)(f, small)
    with pytest.raises(RuntimeError, match="CUDA tensor"):      # ssd.cu:140-142
        sm._cstereomatch.CostOps.compute_ssd(f, f, torch.empty(16, 16, 8), 7)


@pytest.mark.parametrize("reduce_name,agg", [("wta", False), ("wta", True), ("dyn", False)])
def test_pipeline_with_texture_cost(sm, reduce_name, agg):
    """reference tests/test_pipeline.py:24-38: Pipeline(SSDTexture(32), ...) takes tensors and
    wraps them into textures per call; same map as the tensor pipeline, <= 0.1 % off the oracle."""
    from stereomatch.cli_common import create_pipeline
    H, W, D = 48, 96, 32
    left, right, _ = synth.random_dot_pair(H, W, D, seed=99, scale="unit")
    l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    tex_pipe = create_pipeline("ssd-texture", reduce_name, "sgm" if agg else None, D)
    pipe = create_pipeline("ssd", reduce_name, "sgm" if agg else None, D)
    got = tex_pipe.estimate(l, r)
    assert torch.equal(got, pipe.estimate(l, r))
    assert torch.equal(tex_pipe.estimate(torch.from_numpy(left), torch.from_numpy(right)), got)  # host tensors too
    vol = O.ssd(left, right, D, 7)
    if agg:
        vol = O.sgm(vol, left, 0.1, 0.2)
    want = O.wta(vol) if reduce_name == "wta" else O.dp(vol, "cuda")
    assert float((got.cpu().numpy() != want).mean()) <= 1e-3
