"""How fast can the [H,W,D] volume be WRITTEN, by access pattern? (developer tool, torch ops only)
Context for the cost builders, which are pure store streams: contiguous fill vs. the pattern of a
kernel whose CTAs own a 32-disparity (128-byte) chunk of every pixel's D-vector."""
import json
import torch

H, W, D = 1080, 1920, 128
vol = torch.empty(H, W, D, device="cuda")
src = torch.rand(H, W, D, device="cuda")


def med(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return round(ts[len(ts) // 2], 4)


res = {"bytes": vol.numel() * 4}
res["fill_contiguous_ms"] = med(lambda: vol.fill_(1.0))
res["zero_memset_ms"] = med(lambda: vol.zero_())
q = vol.view(H * W, D // 32, 32)
res["fill_quarter_128B_stride_512B_ms"] = med(lambda: q[:, 0, :].fill_(1.0))
res["fill_four_quarters_sequential_ms"] = med(lambda: [q[:, k, :].fill_(1.0) for k in range(D // 32)])
res["copy_read_write_ms"] = med(lambda: vol.copy_(src))
rows = vol.view(H, W * D)
res["fill_every_other_row_ms"] = med(lambda: rows[::2].fill_(1.0))
for k, v in list(res.items()):
    if k.endswith("_ms") and "quarter_128B" not in k and "other_row" not in k:
        res[k.replace("_ms", "_GBs")] = round(res["bytes"] * (2 if "copy" in k else 1) / v / 1e6, 1)
print(json.dumps(res))
