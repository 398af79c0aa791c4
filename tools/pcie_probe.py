#!/usr/bin/env python
"""Diagnostic: pinned-memory PCIe copy bandwidth of this box (the bound of the host-buffer decode path)."""
import torch

dev = torch.device("cuda:0")
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device=dev)
h2 = torch.empty(n // 8, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n // 8, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for name, fn in (("D2H alone", lambda: h.copy_(d, non_blocking=True)), ("H2D alone", lambda: d.copy_(h, non_blocking=True))):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(name, "%.1f GB/s" % (4 * n / 1e6 / e0.elapsed_time(e1)))
# D2H with a concurrent H2D of 1/8 the size (the decode path's mix)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(s1)
with torch.cuda.stream(s1):
    for _ in range(4):
        h.copy_(d, non_blocking=True)
with torch.cuda.stream(s2):
    for _ in range(4):
        d2.copy_(h2, non_blocking=True)
e1.record(s1)
torch.cuda.synchronize()
print("D2H with concurrent H2D/8", "%.1f GB/s" % (4 * n / 1e6 / e0.elapsed_time(e1)))
# chunked D2H (6 MB pieces, like one image per copy)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
piece = 6220800
e0.record()
k = 0
for off in range(0, n - piece, piece):
    h[off:off + piece].copy_(d[off:off + piece], non_blocking=True)
    k += 1
e1.record()
torch.cuda.synchronize()
print("D2H in %d pieces of 6.2 MB" % k, "%.1f GB/s" % (k * piece / 1e6 / e0.elapsed_time(e1)))
