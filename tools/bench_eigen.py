"""Eigen Monte-Carlo loop alone (bmfm_eigen_mc_accumulate), K = 42, window 252:
    python tools/bench_eigen.py [dates] [sims] [modes, e.g. 0,3]
prints ms, sims/s and the fraction of the measured FP64 peak by the algorithmic 15 K^3 flop count."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))
from bmfm import _cabi            # noqa: E402
from bmfm.synth import make_factor_returns   # noqa: E402

T_e = int(sys.argv[1]) if len(sys.argv) > 1 else 256
M = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
modes = [int(x) for x in (sys.argv[3] if len(sys.argv) > 3 else "0,3").split(",")]
K = 42
ctx = _cabi.default_context()
fr = torch.from_numpy(make_factor_returns(T_e + 300, K, seed=5)).cuda()
T = fr.shape[0]
cov = torch.empty((T, K, K), dtype=torch.float64, device="cuda")
st = torch.empty(T, dtype=torch.int32, device="cuda")
ctx.newey_west_by_time(T, K, 2, 252.0, fr, cov, st)
cov_e, st_e = cov[-T_e:].contiguous(), st[-T_e:].contiguous()
D0 = torch.empty((T_e, K), dtype=torch.float64, device="cuda")
U0 = torch.empty((T_e, K, K), dtype=torch.float64, device="cuda")
est = torch.empty(T_e, dtype=torch.int32, device="cuda")
ctx.eigen_decompose(T_e, K, cov_e, st_e, D0, U0, est)
Cm = torch.empty((M, K, K), dtype=torch.float64, device="cuda")
t0 = time.time(); ctx.gauss_cov(0, M, 252, K, Cm); torch.cuda.synchronize(); print(f"gauss_cov M={M}: {1e3 * (time.time() - t0):.1f} ms")
# numpy check of one date: v_j = w_j' D0 w_j / lambda_j of S = D0^1/2 C_m D0^1/2, summed over the simulations
d0 = D0[-1].cpu().numpy()
Cn = Cm.cpu().numpy()
want = np.zeros(K)
for m in range(M):
    S = np.sqrt(d0)[:, None] * Cn[m] * np.sqrt(d0)[None, :]
    lam, W = np.linalg.eigh(S)
    want += ((d0[:, None] * W * W).sum(0) / lam)[::-1]
ref = None
for mode in modes:
    ctx.set_option("eig_mode", mode % 10)
    ctx.set_option("eig_c_blocks", mode // 10)        # e.g. 20 = tile solver with 2 CTAs / SM in the vector kernel
    vsum = torch.zeros((T_e, K), dtype=torch.float64, device="cuda")
    ctx.eigen_mc_accumulate(T_e, K, M, Cm, D0, est, vsum, False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    a.record()
    for _ in range(reps):
        ctx.eigen_mc_accumulate(T_e, K, M, Cm, D0, est, vsum, False)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    v = vsum.cpu().numpy()
    if ref is None:
        ref = v
    print(f"mode {mode}: {T_e} dates x {M} sims: {ms:.2f} ms, {T_e * M / ms / 1e3:.2f} M sims/s, "
          f"{T_e * M * 15 * K ** 3 / ms / 1e9:.2f} TFLOP/s algorithmic, max rel diff vs first mode "
          f"{np.max(np.abs(v - ref) / np.abs(ref)):.2e}, vs numpy on the last date {np.max(np.abs(v[-1] - want) / want):.2e}, "
          f"ok dates {(est == 0).sum().item()}")
