#!/usr/bin/env python
"""Time the Newey-West and vol-regime entry points alone on a synthetic factor-return series (CUDA events).

    python tools/bench_nwvr.py [T] [K] [q]           BENCH_NW_CHUNKS=n sets the nw_chunks context option
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))
from bmfm import _cabi            # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 2500
K = int(sys.argv[2]) if len(sys.argv) > 2 else 42
q = int(sys.argv[3]) if len(sys.argv) > 3 else 2
ctx = _cabi.default_context()
if os.environ.get("BENCH_NW_CHUNKS"):
    ctx.set_option("nw_chunks", int(os.environ["BENCH_NW_CHUNKS"]))
g = torch.Generator(device="cuda")
g.manual_seed(7)
fr = 0.01 * torch.randn((T, K), dtype=torch.float64, device="cuda", generator=g)
cov = torch.empty((T, K, K), dtype=torch.float64, device="cuda")
out = torch.empty_like(cov)
st = torch.empty(T, dtype=torch.int32, device="cuda")
lamb = torch.empty(T, dtype=torch.float64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
nw, vr = [], []
for i in range(10):
    flush.fill_(i)
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    ctx.newey_west_by_time(T, K, q, 252.0, fr, cov, st)
    e[1].record()
    ctx.vol_regime_adj_by_time(T, K, 42.0, fr, cov, st, out, lamb)
    e[2].record()
    torch.cuda.synchronize()
    nw.append(e[0].elapsed_time(e[1]))
    vr.append(e[1].elapsed_time(e[2]))
mb = T * K * K * 8 / 1e6
n, v = float(np.median(nw[3:])), float(np.median(vr[3:]))
print(f"T={T} K={K} q={q}: newey_west {n * 1e3:.1f} us ({mb / n:.0f} GB/s of {mb:.0f} MB written), "
      f"vol_regime {v * 1e3:.1f} us ({2 * mb / v:.0f} GB/s of {2 * mb:.0f} MB), "
      f"cov checksum {float(torch.nan_to_num(cov).sum()):.12e} out checksum {float(torch.nan_to_num(out).sum()):.12e} lamb[-1] {float(lamb[-1]):.12f}")
