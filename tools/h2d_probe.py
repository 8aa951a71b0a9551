#!/usr/bin/env python
"""Host->device copy rate of this box for the tip-state matrix shapes (tuning tool): one contiguous 100 MB copy vs 2-D chunked copies."""
import time
import torch

x = torch.empty(100_000_000, dtype=torch.uint8).pin_memory()
d = torch.empty_like(x, device="cuda")
for name, fn in (("contiguous 100 MB", lambda: d.copy_(x, non_blocking=True)),
                 ("8 x 12.5 MB", lambda: [d[i * 12_500_000:(i + 1) * 12_500_000].copy_(x[i * 12_500_000:(i + 1) * 12_500_000], non_blocking=True) for i in range(8)]),
                 ("2-D 100 rows x 8 chunks", lambda: [d.view(100, -1)[:, i * 125_000:(i + 1) * 125_000].copy_(x.view(100, -1)[:, i * 125_000:(i + 1) * 125_000], non_blocking=True) for i in range(8)])):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("%-28s %.3f ms  %.1f GB/s" % (name, ms, 0.1 / ms * 1e3))
y = torch.empty(8_000_000, dtype=torch.uint8).pin_memory()
dy = torch.empty_like(y, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
y.copy_(dy, non_blocking=True); torch.cuda.synchronize()
e0.record(); y.copy_(dy, non_blocking=True); e1.record(); torch.cuda.synchronize()
print("D2H 8 MB %.3f ms" % e0.elapsed_time(e1))
