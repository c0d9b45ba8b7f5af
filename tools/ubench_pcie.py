"""Pinned host<->device copy rates of the box: the floor of bench.py's end-to-end step (200 MB in, 87 MB out).

    python tools/ubench_pcie.py
"""
import torch, time
n = 200_000_000
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(87_000_000, dtype=torch.uint8).pin_memory()
d2 = torch.empty(87_000_000, dtype=torch.uint8, device="cuda")
s2 = torch.cuda.Stream()
for _ in range(3):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    d.copy_(h, non_blocking=True)
b.record(); torch.cuda.synchronize()
print("H2D 200MB single copy: %.3f ms -> %.1f GB/s" % (a.elapsed_time(b)/5, n/ (a.elapsed_time(b)/5) / 1e6))
a.record()
for _ in range(5):
    for k in range(64):
        lo, hi = n*k//64, n*(k+1)//64
        d[lo:hi].copy_(h[lo:hi], non_blocking=True)
b.record(); torch.cuda.synchronize()
print("H2D 200MB in 64 pieces: %.3f ms -> %.1f GB/s" % (a.elapsed_time(b)/5, n/ (a.elapsed_time(b)/5) / 1e6))
# duplex
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 5
print("duplex 200MB in + 87MB out: %.3f ms" % (dt*1e3))
a.record()
for _ in range(5):
    h2.copy_(d2, non_blocking=True)
b.record(); torch.cuda.synchronize()
print("D2H 87MB: %.3f ms -> %.1f GB/s" % (a.elapsed_time(b)/5, 87e6/ (a.elapsed_time(b)/5) / 1e6))
