"""Pinned host -> device copy rate of the box (what bounds the end-to-end number of bench.py)."""
import json
import torch

x = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
d = torch.empty_like(x, device="cuda")
out = {}
for name, (src, dst) in {"h2d": (x, d), "d2h": (d, x)}.items():
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        dst.copy_(src, non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    out[name + "_GBps"] = 5 * x.numel() / (a.elapsed_time(b) / 1e3) / 1e9
print(json.dumps(out))
