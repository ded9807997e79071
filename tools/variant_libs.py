"""A/B libraries of one translation unit (developer tool): builds tools/bin/libstm_<name>.so for each
variant of csrc/sgm_vsweep.cu below, all other objects from stereo-matching_b200/build/.  Run the
probes on one box with STM_B200_LIB=tools/bin/libstm_<name>.so.
    python tools/variant_libs.py            (after stereo-matching_b200/build.py)"""
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
CSRC = ROOT / "stereo-matching_b200" / "csrc"
OBJ = ROOT / "stereo-matching_b200" / "build"
BIN = ROOT / "tools" / "bin"
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr", f"-I{CSRC}", f"-I{ROOT / 'include'}"]


def sub(s, a, b, cnt=1):
    assert s.count(a) == cnt, (s.count(a), a[:60])
    return s.replace(a, b)


def head(_):  # the committed kernel file under the current host code
    src = subprocess.run(["git", "show", "HEAD:stereo-matching_b200/csrc/sgm_vsweep.cu"], cwd=ROOT, check=True,
                         capture_output=True, text=True).stdout
    return src + "\nnamespace stm { int sgm_sweep_aborted(unsigned *) { return 0; } }\n"


VARIANTS = {

    "head": head,
    "new": lambda s: s,
}


def build(name):
    src = CSRC / f"_variant_{name}.cu"
    src.write_text(VARIANTS[name]((CSRC / "sgm_vsweep.cu").read_text()))
    obj = BIN / f"{name}.o"
    try:
        subprocess.run(["nvcc", *FLAGS, "-c", str(src), "-o", str(obj)], check=True, capture_output=True, text=True)
    finally:
        src.unlink()
    others = [str(o) for o in sorted(OBJ.glob("*.o")) if o.name != "sgm_vsweep.o"]
    subprocess.run(["nvcc", "-shared", "-o", str(BIN / f"libstm_{name}.so"), str(obj), *others, "-gencode",
                    "arch=compute_100a,code=sm_100a", "-Xlinker", "--no-undefined"], check=True)
    obj.unlink()
    return name


if __name__ == "__main__":
    BIN.mkdir(exist_ok=True)
    names = sys.argv[1:] or list(VARIANTS)
    with ThreadPoolExecutor(max_workers=6) as ex:
        print(list(ex.map(build, names)))
