"""Print the SGM launch plan (resident-warp capacities, chosen prefetch variant) -- developer tool."""
import ctypes, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
import torch
torch.cuda.init()
lib = ctypes.CDLL(str(ROOT / "stereo-matching_b200" / "stereomatch" / "libstm_b200.so"))
for paths in (1080, 1920, 2999):
    out = (ctypes.c_int * 6)()
    lib.stm_debug_sgm_plan(paths, out)
    print(paths, "resident default/deep/shallow", out[0], out[1], out[2], "pick acc/ovw", out[3], out[4], "sms", out[5])
