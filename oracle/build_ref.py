"""Build the UNMODIFIED reference native module into oracle/_ref/ (test infrastructure only).

The reference (otaviog/stereo-matching) keeps its CPU functors inside .cu files, so the
only way to run "the reference's own CPU implementation" is to compile its 13 native
sources (the list in the reference's setup.py:63-77) with nvcc + libtorch.  This script
compiles them *from where they lie* under /root/reference (nothing is copied into this
repository) with torch.utils.cpp_extension.load and leaves one file,

    oracle/_ref/_cstereomatch.so

which is git-ignored but travels to the GPU box with the gpurun snapshot.  The module
exposes the reference's `CostOps`, `AggregationOps` and `DisparityReduceOps` static
methods; `oracle/ref_ops.py` loads it.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may use it, and only as the checker or
the timed CPU baseline -- never on the product path.

Usage:  python oracle/build_ref.py [--reference /root/reference] [--force]
"""
from __future__ import annotations

import argparse
import os
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT_DIR = HERE / "_ref"
SO_NAME = "_cstereomatch.so"

# setup.py:63-77 of the reference
SOURCES = [
    "src/cuda_utils.cpp",
    "src/cuda_texture.cpp",
    "src/cuda_texture_gpu.cu",
    "src/cost.cpp",
    "src/ssd.cu",
    "src/birchfield_cost.cu",
    "src/winners_take_all.cu",
    "src/dynamic_programming.cu",
    "src/disparity_reduce.cpp",
    "src/semiglobal.cpp",
    "src/semiglobal_gpu.cu",
    "src/aggregation.cpp",
    "src/_cstereomatch.cpp",
]


def build(reference: Path, force: bool = False, verbose: bool = True) -> Path | None:
    """Returns the path of the built module, or None when the reference tree is absent."""
    target = OUT_DIR / SO_NAME
    if target.exists() and not force:
        return target
    if not (reference / "src" / "ssd.cu").exists():
        return None

    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 8))
    from torch.utils.cpp_extension import load

    build_dir = OUT_DIR / "build"
    build_dir.mkdir(parents=True, exist_ok=True)
    # torch's cpp_extension serialises builds with a `lock` file and waits on it forever; a
    # build that was killed leaves it behind and every later build() would hang.  This is the
    # only builder of this directory (tests never build, they only check ref_ops.available()),
    # so a lock found here is stale.
    stale = build_dir / "lock"
    if stale.exists():
        print(f"[build_ref] removing stale {stale} left by an interrupted build", file=sys.stderr)
        stale.unlink()
    load(
        name="_cstereomatch",
        sources=[str(reference / s) for s in SOURCES],
        extra_include_paths=[str(reference / "include"),
                             str(reference / "include" / "stereomatch")],
        extra_cflags=["-std=c++17", "-O2"],
        extra_cuda_cflags=["-std=c++17", "-O2"],
        build_directory=str(build_dir),
        is_python_module=False,
        verbose=verbose,
    )
    shutil.copy2(build_dir / SO_NAME, target)
    # keep only the module itself (objects are large and pointless on the GPU box)
    shutil.rmtree(build_dir, ignore_errors=True)
    return target


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--force", action="store_true")
    args = ap.parse_args()
    out = build(Path(args.reference), force=args.force)
    if out is None:
        print("reference tree not found; nothing built", file=sys.stderr)
        return 1
    print(out)
    return 0


if __name__ == "__main__":
    sys.exit(main())
