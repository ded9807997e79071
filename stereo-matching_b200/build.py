"""Builds libstm_b200.so (hand-written sm_100a CUDA + the C ABI of include/stm_abi.h).

    python stereo-matching_b200/build.py [--force] [--verbose]

nvcc cross-compiles for sm_100a without a GPU.  The library is built IN-TREE next to the
Python package (stereo-matching_b200/stereomatch/libstm_b200.so) so that it travels to
the GPU box with the repository snapshot.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OUT = HERE / "stereomatch" / "libstm_b200.so"
OBJ = HERE / "build"

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr",
]


def sources():
    return sorted(CSRC.glob("*.cu"))


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    headers = list(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "stm_abi.h", Path(__file__)]
    OBJ.mkdir(exist_ok=True)
    jobs = []
    for src in sources():
        obj = OBJ / (src.stem + ".o")
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        # e.g. STM_NVCC_EXTRA="-DSTM_VSWEEP_DEBUG" (with --force) builds the developer knobs in
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("STM_NVCC_EXTRA", "").split(), "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        (OBJ / (src.stem + ".ptxas.log")).write_text(r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        list(ex.map(compile_one, jobs))
    objs = [OBJ / (s.stem + ".o") for s in sources()]
    if force or jobs or _stale(OUT, objs):
        # --no-undefined: a launcher whose signature drifted from its declaration in another
        # translation unit must fail HERE, not when the library is loaded on the GPU box
        cmd = [nvcc, "-shared", "-o", str(OUT), *map(str, objs),
               "-gencode", "arch=compute_100a,code=sm_100a", "-Xlinker", "--no-undefined"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stderr}")
    return OUT


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
