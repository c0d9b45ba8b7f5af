"""bayes_shrink alone at the C3 cross-section size:  python tools/bench_bayes.py [N] [T]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))
from bmfm import _cabi            # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
T = int(sys.argv[2]) if len(sys.argv) > 2 else 750
ctx = _cabi.default_context()
g = torch.Generator(device="cuda")
g.manual_seed(1)
vol = (0.02 + 0.01 * torch.randn(N * T, dtype=torch.float64, device="cuda", generator=g)).abs()
cap = torch.exp(15.0 + 1.5 * torch.randn(N * T, dtype=torch.float64, device="cuda", generator=g))
off = torch.arange(T + 1, dtype=torch.int64, device="cuda") * N
out = torch.empty_like(vol)
st = torch.empty(T, dtype=torch.int32, device="cuda")
for _ in range(3):
    ctx.bayes_shrink(T, N, off, vol, cap, 10, 1.0, out, st)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    ctx.bayes_shrink(T, N, off, vol, cap, 10, 1.0, out, st)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 5
print(f"bayes_shrink N={N} T={T}: {ms:.3f} ms, {24 * N * T / ms / 1e6:.0f} GB/s ({24 * N * T / ms / 1e6 / 6538:.3f} of HBM), status ok={bool((st == 0).all())}")
