#!/usr/bin/env python
"""Pinned host -> device copy bandwidth of the box (the floor of bench.py's e2e: 12.6 GB per layer cross PCIe)."""
import json
import torch
x = torch.empty(16, 2048, 11008, dtype=torch.float16).pin_memory()
d = torch.empty_like(x, device='cuda')
for _ in range(2):
    d.copy_(x, non_blocking=True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    d.copy_(x, non_blocking=True)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
print(json.dumps({'k': 'h2d_pinned', 'MB': x.numel() * 2 / 1e6, 'ms': ms, 'GBps': x.numel() * 2 / ms / 1e6}))
