#!/usr/bin/env python3
"""Write-only HBM bandwidth of the box (torch fill of a large tensor): the ceiling of a kernel that only writes."""
import json
import torch
n = 40 * (1 << 30)
x = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2):
    x.zero_()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
ev[0].record()
for i in range(5):
    x.zero_()
    ev[i + 1].record()
torch.cuda.synchronize()
ms = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(5))[2]
print(json.dumps({"memset_GBps": n / ms / 1e6, "bytes": n, "ms": ms}))
