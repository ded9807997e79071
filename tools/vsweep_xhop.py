"""CTA -> CTA hop of the fused sweep on %globaltimer (STM_DEBUG_VSWEEP=16, debug build):
publish of slots 0/1 of CTA 6 -> exporter's stores -> importer delivery in CTA 5 -> unblock of
CTA 5's slots 12/11 (developer tool)."""
import ctypes, os, sys
from pathlib import Path
os.environ["STM_DEBUG_VSWEEP"] = "16"
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
import torch, stereomatch
from stereomatch._cstereomatch import _lib
from oracle import synth
H, W, D = 1080, 1920, 128
left, right, _ = synth.random_dot_pair(H, W, D, seed=1000, scale="unit")
l, r = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
cost = stereomatch.cost.SSD(D)(l, r)
fused = stereomatch.aggregation.Semiglobal(fused_sweeps=True)
for _ in range(2):
    fused(cost, l)
torch.cuda.synchronize()
R = 24
buf = (ctypes.c_longlong * (32 * R * 4))()
_lib.stm_debug_vsweep_trace(buf)
a = torch.tensor(list(buf), dtype=torch.int64).view(32, R, 4)
t0 = int(a[16, 0, 2])
f = lambda v: int(v) - t0
print("row r: CTA6 slot0 publish, slot1 publish, exporter stores done | CTA5 importer delivered | CTA5 slot12 unblocked (row r+1), slot11 unblocked (row r+1); ns")
for r in range(R - 1):
    print(500 + r, f(a[16, r, 2]), f(a[17, r, 2]), f(a[30, r, 0]), "|", f(a[15, r, 1]), "|", f(a[12, r + 1, 1]), f(a[11, r + 1, 1]),
          "| hop", f(a[15, r, 1]) - f(a[16, r, 2]), "| importer: top", f(a[15, r, 3]), "after ring-space wait", f(a[15, r, 0]), "polls", int(a[15, r, 2]))
