"""A/B libraries of one translation unit built with different -D flags (developer tool).
    python tools/variant_defs.py ssd_rows stream="-DSTM_SSD_STREAM_PAIRS=1" items="-DSTM_SSD_STREAM_PAIRS=0"
builds tools/bin/libstm_<name>.so from csrc/<tu>.cu with the flags and every other object from
stereo-matching_b200/build/ (run stereo-matching_b200/build.py first).  Select one on the box with
STM_B200_LIB=tools/bin/libstm_<name>.so (tools/ab_stage.sh)."""
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
CSRC = ROOT / "stereo-matching_b200" / "csrc"
OBJ = ROOT / "stereo-matching_b200" / "build"
BIN = ROOT / "tools" / "bin"
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr"]


def build(tu, name, defs):
    obj = BIN / f"{name}.o"
    subprocess.run(["nvcc", *FLAGS, *defs.split(), "-c", str(CSRC / f"{tu}.cu"), "-o", str(obj)], check=True)
    others = [str(o) for o in sorted(OBJ.glob("*.o")) if o.name != f"{tu}.o"]
    subprocess.run(["nvcc", "-shared", "-o", str(BIN / f"libstm_{name}.so"), str(obj), *others, "-gencode",
                    "arch=compute_100a,code=sm_100a", "-Xlinker", "--no-undefined"], check=True)
    obj.unlink()
    return name


if __name__ == "__main__":
    BIN.mkdir(exist_ok=True)
    tu = sys.argv[1]
    pairs = [a.split("=", 1) for a in sys.argv[2:]]
    with ThreadPoolExecutor(max_workers=6) as ex:
        print(list(ex.map(lambda nv: build(tu, nv[0], nv[1]), pairs)))
