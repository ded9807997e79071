"""The drop-in claim, run: the REFERENCE's own unmodified Python package (cost.py, aggregation.py,
disparity_reduce.py, cuda_texture.py, pipeline.py, cli_common.py ...) on top of the binding printed
in INTEGRATION.md section 2 and libstm_b200.so.  oracle/build_dropin.py assembles the package under
oracle/_ref/dropin (git-ignored; travels to the GPU box); it is imported in a FRESH interpreter
because the reference uses absolute `stereomatch.` imports that would collide with this
repository's own package of the same name."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from oracle import build_dropin, oracle as O, synth

ROOT = Path(__file__).resolve().parents[1]
REFERENCE = Path("/root/reference")


@pytest.fixture(scope="module")
def dropin(tmp_path_factory):
    if REFERENCE.exists():                      # this container: assemble a fresh one
        out = build_dropin.build(REFERENCE, tmp_path_factory.mktemp("dropin"))
    elif build_dropin.refresh_library():        # the GPU box: the prebuilt one, current library + stub
        out = build_dropin.DEFAULT_OUT
    else:
        pytest.skip("neither /root/reference nor oracle/_ref/dropin is present")
    return out


def run_in_dropin(dropin, code: str, *args) -> str:
    env = dict(os.environ, PYTHONPATH=str(dropin))
    r = subprocess.run([sys.executable, "-c", code, *map(str, args)], cwd=str(dropin), env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    return r.stdout


def test_binding_in_integration_md_is_complete():
    """Every C symbol the documented binding uses is declared in include/stm_abi.h."""
    import re
    stub = build_dropin.integration_stub()
    header = (ROOT / "include" / "stm_abi.h").read_text()
    used = set(re.findall(r"_lib\.(stm_[a-z0-9_]+)", stub))
    assert {"stm_ssd_f32", "stm_sgm_f32", "stm_wta_f32", "stm_dp_f32", "stm_birchfield_f32",
            "stm_ssd_tex_f32", "stm_texture_create"} <= used
    for name in used:
        assert re.search(rf"\b{name}\s*\(", header), name


def test_reference_package_imports_and_rejects_cpu_tensors(dropin):
    out = run_in_dropin(dropin, r'''
import json, sys, torch
import stereomatch
from stereomatch.cost import SSD, SSDTexture, Birchfield
from stereomatch.aggregation import Semiglobal
from stereomatch.disparity_reduce import WinnerTakesAll, DynamicProgramming
from stereomatch.pipeline import Pipeline
from stereomatch.cli_common import create_pipeline
from stereomatch.cuda_texture import CUDATexture
assert stereomatch.__file__.startswith(sys.argv[1]), stereomatch.__file__
assert "ctypes" in open(stereomatch._cstereomatch.__file__).read()      # the stub, not a pybind .so
p = Pipeline(SSD(64), WinnerTakesAll(), Semiglobal())
q = create_pipeline("birchfield", "dyn", "sgm", 32)
assert (p.cost.max_disparity, p.aggregation.penalty1, q.cost.max_disparity) == (64, 0.1, 32)
img = torch.zeros(8, 16)
errors = []
for fn in (lambda: SSD(8)(img, img), lambda: Birchfield(8)(img, img), lambda: p.estimate(img, img),
           lambda: WinnerTakesAll()(torch.zeros(8, 16, 8)), lambda: DynamicProgramming()(torch.zeros(8, 16, 8)),
           lambda: Semiglobal()(torch.zeros(8, 16, 8), img)):
    try:
        fn()
        errors.append(None)
    except RuntimeError as e:
        errors.append(str(e))
print(json.dumps(errors))
''', dropin)
    errors = json.loads(out.strip().splitlines()[-1])
    assert all(e is not None for e in errors), errors      # CUDA only: every CPU call raises
    assert any("gpu tensor" in e for e in errors)


@pytest.mark.gpu
def test_reference_pipeline_over_the_binding_matches_the_oracle(dropin, tmp_path):
    H, W, D = 96, 160, 64
    left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="unit")
    np.save(tmp_path / "l.npy", left)
    np.save(tmp_path / "r.npy", right)
    run_in_dropin(dropin, r'''
import sys, numpy as np, torch
from stereomatch.cost import SSD, SSDTexture, Birchfield
from stereomatch.aggregation import Semiglobal
from stereomatch.disparity_reduce import WinnerTakesAll, DynamicProgramming
from stereomatch.pipeline import Pipeline
d = sys.argv[1]
l = torch.from_numpy(np.load(d + "/l.npy")).cuda()
r = torch.from_numpy(np.load(d + "/r.npy")).cuda()
D = 64
out = {}
out["ssd_sgm_wta"] = Pipeline(SSD(D), WinnerTakesAll(), Semiglobal()).estimate(l, r)
out["ssd_wta"] = Pipeline(SSD(D), WinnerTakesAll()).estimate(l, r)
out["tex_wta"] = Pipeline(SSDTexture(D), WinnerTakesAll()).estimate(l, r)
out["bt_sgm_wta"] = Pipeline(Birchfield(D), WinnerTakesAll(), Semiglobal()).estimate(l, r)
out["ssd_dp"] = Pipeline(SSD(D), DynamicProgramming()).estimate(l, r)
out["ssd_vol"] = SSD(D)(l, r)
out["sgm_vol"] = Semiglobal()(out["ssd_vol"], l)
torch.cuda.synchronize()
np.savez(d + "/out.npz", **{k: v.cpu().numpy() for k, v in out.items()})
''', tmp_path)
    got = np.load(tmp_path / "out.npz")
    ssd = O.ssd(left, right, D, 7)
    agg = O.sgm(ssd, left, 0.1, 0.2)
    fin = np.isfinite(ssd)
    assert np.array_equal(np.isfinite(got["ssd_vol"]), fin)
    assert np.allclose(got["ssd_vol"][fin], ssd[fin], rtol=1e-4, atol=0)
    assert np.allclose(got["sgm_vol"][np.isfinite(agg)], agg[np.isfinite(agg)], rtol=1e-4, atol=0)
    bt = O.birchfield(left, right, D)
    for key, want in (("ssd_sgm_wta", O.wta(agg)), ("ssd_wta", O.wta(ssd)), ("tex_wta", O.wta(ssd)),
                      ("bt_sgm_wta", O.wta(O.sgm(bt, left, 0.1, 0.2))), ("ssd_dp", O.dp(ssd, "cuda"))):
        assert got[key].dtype == np.int32 and got[key].shape == (H, W)
        assert float((got[key] != want).mean()) <= 1e-3, key       # north_star: <= 0.1 % of pixels
