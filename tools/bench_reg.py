#!/usr/bin/env python
"""Time the regression entry point alone on a synthetic panel of a given shape (CUDA events).

    python tools/bench_reg.py N T [mode]        mode: the reg_mode context option (0 auto; 3 / 5 / 4 = FP64-MMA kernel with 1 / 2 / 4 warps per date; 6 = 4 warps per date fed by cp.async.bulk;
                                                1 CTA kernel, 2 register-Gram stream kernel: earlier designs)
    BENCH_UNGROUPED=1: rows not grouped by industry; BENCH_GATHER=n: the fused-gather entry point with n local "peers"
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))
from bmfm import _cabi            # noqa: E402
from bmfm.panel import pack_arrays  # noqa: E402
from bmfm.synth import make_panel   # noqa: E402

N, T = int(sys.argv[1]), int(sys.argv[2])
MODE = int(sys.argv[3]) if len(sys.argv) > 3 else 0
P, Q = 31, 10
K = 1 + P + Q
pan = make_panel(T, N, P, Q, seed=3, scale_log10=(-2, 2))
csr = pan.to_csr()
GROUPED = os.environ.get("BENCH_UNGROUPED") != "1"     # rows sorted by industry inside a date (what pack_frame does)
(cap, ret, ind, style), pos = pack_arrays(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], P,
                                          group_by_industry=GROUPED)
d = lambda a: torch.from_numpy(a).cuda()
off, cap, ret, ind, style = d(csr["date_off"]), d(cap), d(ret), d(ind), d(style)
rows = int(csr["date_off"][-1])
fr = torch.empty((T, K), dtype=torch.float64, device="cuda")
spec = torch.empty(rows, dtype=torch.float64, device="cuda")
r2 = torch.empty(T, dtype=torch.float64, device="cuda")
st = torch.empty(T, dtype=torch.int32, device="cuda")
ctx = _cabi.default_context()
ctx.set_option("reg_mode", MODE)
for name in ("reg_smma", "reg_warps", "reg_ctas", "reg_static"):
    if os.environ.get("BENCH_" + name.upper()):
        ctx.set_option(name, int(os.environ["BENCH_" + name.upper()]))
GATHER = int(os.environ.get("BENCH_GATHER", "0"))
fr_all = torch.empty((T, K), dtype=torch.float64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
times = []
for i in range(8):
    flush.fill_(i)        # evict the panel from L2 between iterations
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if GATHER:      # the fused-gather entry point with this GPU's own buffer as the only "peer"
        ctx.reg_by_time_gather(T, P, Q, N, off, cap, ret, ind, style, fr, spec, r2, st, [fr_all.data_ptr()] * GATHER, 0)
    else:
        ctx.reg_by_time(T, P, Q, N, off, cap, ret, ind, style, fr, spec, r2, st)
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
ms = float(np.median(times[3:]))
print(f"{'grouped' if GROUPED else 'ungrouped'} N={N} T={T} mode={MODE} (ran {ctx.get_option('last_reg_mode')}): {ms:.3f} ms, {rows * 108 / ms / 1e6:.0f} GB/s "
      f"({rows * 108 / ms / 1e6 / 6538:.3f} of measured HBM), {T / ms * 1e3:.0f} dates/s, status ok={bool((st == 0).all())}")
