"""Per-row timeline of one CTA of the fused sweep (STM_DEBUG_VSWEEP=16) -- developer tool."""
import ctypes, os, sys
from pathlib import Path
os.environ["STM_DEBUG_VSWEEP"] = os.environ.get("STM_DEBUG_VSWEEP", "16")
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
buf = (ctypes.c_longlong * (16 * R * 4))()
_lib.stm_debug_vsweep_trace(buf)   # the up sweep ran last: rows are sweep steps either way
a = torch.tensor(list(buf), dtype=torch.int64).view(16, R, 4)
t0 = int(a[:13, 0, 0].min())
ns = lambda v: (int(v) - t0) / 1.9  # cycles -> ns at ~1.9 GHz
print("sweep steps 500..; per slot: w = wait for inputs, p = unblocked->published, d = published->row done, @ = start of wait (ns)")
for row in range(R):
    line = [f"t={500+row:4d}"]
    for j in (12, 11, 10, 6, 2, 1, 0):
        e = a[j, row]
        line.append(f"s{j}: w{ns(e[1])-ns(e[0]):6.0f} p{ns(e[2])-ns(e[1]):5.0f} d{ns(e[3])-ns(e[2]):5.0f} @{ns(e[0]):7.0f}")
    imp = a[15, row]
    line.append(f"imp: poll{ns(imp[1])-ns(imp[0]):6.0f} @{ns(imp[1]):7.0f}")
    print(" | ".join(line))
