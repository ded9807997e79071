"""Assemble the package a reference maintainer would end up with (test infrastructure only).

    oracle/_ref/dropin/stereomatch/
        __init__.py _common.py numeric.py cost.py aggregation.py disparity_reduce.py
        cuda_texture.py pipeline.py cli_common.py      <- the REFERENCE's own files, unmodified,
                                                          copied from /root/reference/stereomatch/
        _cstereomatch.py                               <- the code block of INTEGRATION.md section 2
        libstm_b200.so                                 <- this repository's library

Everything lands under oracle/_ref/ (git-ignored, so no reference source enters the history;
not gpurun-ignored, so it travels to the GPU box).  tests/test_dropin_reference_package.py imports
the package in a fresh interpreter.  Usage: python oracle/build_dropin.py [--reference DIR] [--out DIR]
"""
from __future__ import annotations

import argparse
import re
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
DEFAULT_OUT = HERE / "_ref" / "dropin"
REFERENCE_FILES = ["__init__.py", "_common.py", "numeric.py", "cost.py", "aggregation.py",
                   "disparity_reduce.py", "cuda_texture.py", "pipeline.py", "cli_common.py"]


def integration_stub() -> str:
    """The python code block of INTEGRATION.md section 2: the file under test IS the document."""
    text = (ROOT / "INTEGRATION.md").read_text()
    m = re.search(r"```python\n(import ctypes, torch\n.*?)```", text, flags=re.S)
    if not m:
        raise RuntimeError("INTEGRATION.md: binding code block not found")
    return m.group(1)


def build(reference: Path, out: Path = DEFAULT_OUT) -> Path | None:
    """Returns the directory to put on sys.path, or None when the reference tree is absent."""
    src = reference / "stereomatch"
    if not (src / "pipeline.py").exists():
        return None
    lib = ROOT / "stereo-matching_b200" / "stereomatch" / "libstm_b200.so"
    if not lib.exists():
        raise RuntimeError(f"{lib} is missing: run python stereo-matching_b200/build.py first")
    pkg = out / "stereomatch"
    if pkg.exists():
        shutil.rmtree(pkg)
    pkg.mkdir(parents=True)
    for name in REFERENCE_FILES:
        shutil.copy2(src / name, pkg / name)
    (pkg / "_cstereomatch.py").write_text(integration_stub())
    shutil.copy2(lib, pkg / "libstm_b200.so")
    return out


def refresh_library(out: Path = DEFAULT_OUT) -> bool:
    """Re-copy libstm_b200.so and the stub into an existing drop-in (no reference tree needed)."""
    pkg = out / "stereomatch"
    if not (pkg / "pipeline.py").exists():
        return False
    shutil.copy2(ROOT / "stereo-matching_b200" / "stereomatch" / "libstm_b200.so", pkg / "libstm_b200.so")
    (pkg / "_cstereomatch.py").write_text(integration_stub())
    return True


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--out", default=str(DEFAULT_OUT))
    a = ap.parse_args()
    r = build(Path(a.reference), Path(a.out))
    if r is None:
        print("reference tree not found; nothing built", file=sys.stderr)
        sys.exit(1)
    print(r)
