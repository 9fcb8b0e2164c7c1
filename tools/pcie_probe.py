"""Pinned host -> device copy rate of this box (the ceiling of bench.py's e2e leg, which ships 10 GB per step)."""
import json
import torch

n = 2 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    d.copy_(h, non_blocking=True)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 5
print(json.dumps({"probe": "pinned H2D", "bytes": n, "ms": ms, "GBps": n / ms / 1e6, "ms_for_10GB": 1e10 / (n / ms)}))
