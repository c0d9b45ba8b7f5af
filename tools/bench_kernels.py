#!/usr/bin/env python
"""Per-kernel measurement table for every row of SURVEY.md §8(a): CUDA-event time at the config
size, algorithmic bytes (or flops) per unit, achieved rate against the measured HBM peak, and the
CPU port (oracle/) timed on a bounded sample on the same box.

    python tools/bench_kernels.py [--out profiles/r01_kernel_table.json]

Inputs are resident in HBM; an L2 flush (256 MB fill) separates timed launches.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))
sys.path.insert(0, ROOT)
from bmfm import _cabi                      # noqa: E402
from bmfm.panel import pack_arrays           # noqa: E402
from bmfm.synth import make_panel, make_factor_returns, make_price_panel  # noqa: E402
from oracle import barra_oracle as orc       # noqa: E402

PEAK = 6538.0
if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])

ctx = _cabi.default_context()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
dv = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
rows_out = []


def timeit(fn, reps=7):
    ts = []
    for i in range(reps):
        flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts[2:]))


def cpu_time(fn):
    t0 = time.perf_counter()
    fn()
    return time.perf_counter() - t0


def add(name, config, ms, units, unit_name, bytes_per_unit, cpu_units_per_s, cpu_sample):
    gbs = units * bytes_per_unit / ms / 1e6 if bytes_per_unit else None
    row = {"kernel": name, "config": config, "ms": round(ms, 4), "units_per_s": units / ms * 1e3, "unit": unit_name,
           "algorithmic_bytes_per_unit": bytes_per_unit, "achieved_GBps": gbs,
           "frac_of_measured_hbm": (gbs / PEAK if gbs else None),
           "cpu_port_units_per_s": cpu_units_per_s, "cpu_sample": cpu_sample,
           "speedup_vs_cpu_port": (units / ms * 1e3 / cpu_units_per_s if cpu_units_per_s else None)}
    rows_out.append(row)
    print(f"{name:28s} {config:34s} {ms:9.3f} ms  {units / ms * 1e3:14.0f} {unit_name}/s"
          + (f"  {gbs:7.0f} GB/s ({gbs / PEAK:.3f})" if gbs else "")
          + (f"  cpu {cpu_units_per_s:10.1f}/s  x{units / ms * 1e3 / cpu_units_per_s:.0f}" if cpu_units_per_s else ""), flush=True)


def bench_reg(N, T, tag):
    P, Q = 31, 10
    K = 1 + P + Q
    pan = make_panel(T, N, P, Q, seed=3, scale_log10=(-2, 2))
    csr = pan.to_csr()
    (cap, ret, ind, style), _ = pack_arrays(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], P)
    off, dcap, dret, dind, dsty = dv(csr["date_off"]), dv(cap), dv(ret), dv(ind), dv(style)
    rows = len(cap)
    fr = torch.empty((T, K), dtype=torch.float64, device="cuda")
    spec = torch.empty(rows, dtype=torch.float64, device="cuda")
    r2 = torch.empty(T, dtype=torch.float64, device="cuda")
    st = torch.empty(T, dtype=torch.int32, device="cuda")
    ms = timeit(lambda: ctx.reg_by_time(T, P, Q, N, off, dcap, dret, dind, dsty, fr, spec, r2, st))
    nd = 6
    ar = np.arange(P)
    def cpu():
        for t in range(nd):
            a, b = int(csr["date_off"][t]), int(csr["date_off"][t + 1])
            oh = (csr["ind"][a:b, None] == ar[None, :]).astype(np.int64)
            orc.cross_section_reg(csr["cap"][a:b], csr["ret"][a:b], oh, csr["style"][a:b])
    c = cpu_time(cpu)
    add("reg_by_time", f"{tag}: N={N} T={T} P=31 Q=10", ms, rows, "stock-dates", 108, nd * N / c, f"{nd} dates")
    return fr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    K = 42
    # ---- part 1
    bench_reg(300, 1400, "C1")
    bench_reg(800, 2500, "C2")
    bench_reg(5000, 750, "C3 (1/5 of the dates)")
    # ---- part 3
    for (T, tag) in [(2500, "C2"), (3750, "C3")]:
        frn = make_factor_returns(T, K, seed=4)
        fr = dv(frn)
        cov = torch.empty((T, K, K), dtype=torch.float64, device="cuda")
        st = torch.empty(T, dtype=torch.int32, device="cuda")
        ms = timeit(lambda: ctx.newey_west_by_time(T, K, 2, 252.0, fr, cov, st))
        pick = np.linspace(K + 1, T - 1, 12).astype(int)
        c = cpu_time(lambda: [orc.newey_west(frn[:t + 1], 2, 252) for t in pick])
        add("newey_west_by_time", f"{tag}: T={T} K=42 q=2", ms, T, "dates", 8 * K * K + 8 * K, len(pick) / c, "12 windows")
        out = torch.empty_like(cov)
        lamb = torch.empty(T, dtype=torch.float64, device="cuda")
        ms = timeit(lambda: ctx.vol_regime_adj_by_time(T, K, 42.0, fr, cov, st, out, lamb))
        covs = cov.cpu().numpy()
        stn = st.cpu().numpy()
        Tc = 600
        c = cpu_time(lambda: orc.vol_regime_adj_by_time(frn[:Tc], [covs[t] if stn[t] == 0 else None for t in range(Tc)], 42))
        add("vol_regime_adj_by_time", f"{tag}: T={T} K=42", ms, T, "dates", 16 * K * K, Tc / c / (T / Tc), f"{Tc}-date loop, cost scaled x{T / Tc:.1f}")
    # eigen MC (C4 slice): 96 dates x 1024 sims, T_sim 252
    T_e, M = 96, 1024
    cov_e = cov[-T_e:].contiguous()
    st_e = st[-T_e:].contiguous()
    D0 = torch.empty((T_e, K), dtype=torch.float64, device="cuda")
    U0 = torch.empty((T_e, K, K), dtype=torch.float64, device="cuda")
    est = torch.empty(T_e, dtype=torch.int32, device="cuda")
    vsum = torch.zeros((T_e, K), dtype=torch.float64, device="cuda")
    eo = torch.empty((T_e, K, K), dtype=torch.float64, device="cuda")
    Cm = torch.empty((M, K, K), dtype=torch.float64, device="cuda")
    ms_g = timeit(lambda: ctx.gauss_cov(0, M, 252, K, Cm), reps=3)
    add("gauss_cov (one-off)", f"C4: M={M} T_sim=252 K=42", ms_g, M, "sims", None, None, None)
    def eig():
        ctx.eigen_decompose(T_e, K, cov_e, st_e, D0, U0, est)
        ctx.eigen_mc_accumulate(T_e, K, M, Cm, D0, est, vsum, False)
        ctx.eigen_finalize(T_e, K, M, 1.4, D0, U0, vsum, est, eo)
    ms = timeit(eig, reps=5)
    c = cpu_time(lambda: orc.eigen_risk_adj(cov_e[-1].cpu().numpy(), 252, 16, 1.4))
    add("eigen_risk_adj_by_time", f"C4 slice: {T_e} dates x {M} sims, T_sim=252", ms, T_e * M, "sims", None, 16 / c, "16 sims of one date")
    rows_out[-1]["algorithmic_flops_per_unit"] = 15 * K ** 3
    rows_out[-1]["achieved_algorithmic_TFLOPs"] = T_e * M * 15 * K ** 3 / ms / 1e9
    # bayes shrink (C3 cross-sections)
    N, T = 5000, 750
    rng = np.random.default_rng(0)
    off = (np.arange(T + 1) * N).astype(np.int64)
    vol = np.abs(rng.normal(0.02, 0.01, N * T))
    capn = np.exp(rng.normal(15, 1.5, N * T))
    dvol, dcapn, doff = dv(vol), dv(capn), dv(off)
    o = torch.empty_like(dvol)
    sst = torch.empty(T, dtype=torch.int32, device="cuda")
    ms = timeit(lambda: ctx.bayes_shrink(T, N, doff, dvol, dcapn, 10, 1.0, o, sst))
    c = cpu_time(lambda: [orc.bayes_shrink(vol[i * N:(i + 1) * N], capn[i * N:(i + 1) * N]) for i in range(20)])
    add("bayes_shrink", f"C3: N={N} T={T} (1/5 of the dates)", ms, N * T, "stock-dates", 24, 20 * N / c, "20 dates")
    # ---- part 2: descriptors at C5 scale (5000 stocks x 3750 days)
    S, L = 5000, 3750
    close, idx = make_price_panel(L, S, seed=5, missing=0.005, late=50)
    mkt1 = np.full(L, np.nan)
    mkt1[1:] = idx[1:] / idx[:-1] - 1
    dclose = dv(close)
    dmkt = dv(np.repeat(mkt1[:, None], S, axis=1))
    ret = torch.empty_like(dclose)
    lr = torch.empty_like(dclose)
    ms = timeit(lambda: ctx.prepare_returns(L, S, dclose, ret, lr), reps=5)
    add("prepare_returns", f"C5: S={S} L={L}", ms, S * L, "stock-days", 24, None, None)
    b = torch.empty_like(dclose)
    h = torch.empty_like(dclose)
    ms = timeit(lambda: ctx.beta_hsigma(L, S, ret, dmkt, 252, 63, 42, b, h), reps=5)
    rs = ret[:, 0].cpu().numpy()
    c = cpu_time(lambda: orc.beta_hsigma(rs[:600], mkt1[:600]))
    add("beta_hsigma", f"C5: S={S} L={L} window 252", ms, S * L, "stock-days", 32, 600 / c, "1 stock x 600 days")
    ms = timeit(lambda: ctx.dastd(L, S, ret, dmkt, 252, 42, 42, b), reps=5)
    c = cpu_time(lambda: orc.dastd(rs[:1500] - mkt1[:1500]))
    add("dastd", f"C5: S={S} L={L} window 252", ms, S * L, "stock-days", 24, 1500 / c, "1 stock x 1500 days")
    ms = timeit(lambda: ctx.cmra(L, S, lr, 252, b), reps=5)
    ls = lr[:, 0].cpu().numpy()
    c = cpu_time(lambda: orc.cmra(ls[:1500]))
    add("cmra", f"C5: S={S} L={L} window 252", ms, S * L, "stock-days", 16, 1500 / c, "1 stock x 1500 days")
    ms = timeit(lambda: ctx.rstr(L, S, lr, 504, 21, 126, 42, b), reps=5)
    c = cpu_time(lambda: orc.rstr(ls[:1500]))
    add("rstr", f"C5: S={S} L={L} window 483", ms, S * L, "stock-days", 16, 1500 / c, "1 stock x 1500 days")
    dto = dv(np.exp(rng.normal(0.2, 0.8, size=(L, S))))
    o3 = [torch.empty_like(dto) for _ in range(3)]
    ms = timeit(lambda: ctx.rolling_log_sum3(L, S, dto, 100.0, [21, 63, 252], [15, 42, 126], o3), reps=5)
    tos = dto[:, 0].cpu().numpy()
    c = cpu_time(lambda: orc.liquidity(tos[:1500]))
    add("liquidity (STOM+STOQ+STOA)", f"C5: S={S} L={L}", ms, S * L, "stock-days", 32, 1500 / c, "1 stock x 1500 days")
    # post-processing on the long table grouped by date (T dates x N stocks)
    N, T = 5000, 750
    off = (np.arange(T + 1) * N).astype(np.int64)
    x = rng.normal(size=N * T)
    x[rng.random(N * T) < 0.03] = np.nan
    y2 = rng.normal(size=N * T)
    z2 = rng.normal(size=N * T)
    dx, dy2, dz2, doff = dv(x), dv(y2), dv(z2), dv(off)
    o = torch.empty_like(dx)
    ms = timeit(lambda: ctx.winsorize(T, doff, dx, 2.5, o))
    c = cpu_time(lambda: [orc.winsorize(x[i * N:(i + 1) * N]) for i in range(50)])
    add("winsorize", f"N={N} T={T}, one column", ms, N * T, "rows", 16, 50 * N / c, "50 dates")
    ms = timeit(lambda: ctx.composite(N * T, [dx, dy2, dz2], [0.7, 0.15, 0.15], o))
    c = cpu_time(lambda: orc.composite([x[:N * 50], y2[:N * 50], z2[:N * 50]], [0.7, 0.15, 0.15]))
    add("composite", f"N={N} T={T}, 3 components", ms, N * T, "rows", 32, 50 * N / c, "50 dates")
    ms = timeit(lambda: ctx.orthogonalize(T, doff, dx, [dy2, dz2], o))
    c = cpu_time(lambda: [orc.orthogonalize(x[i * N:(i + 1) * N], np.column_stack([y2[i * N:(i + 1) * N], z2[i * N:(i + 1) * N]])) for i in range(50)])
    add("orthogonalize", f"N={N} T={T}, k=2", ms, N * T, "rows", 32, 50 * N / c, "50 dates")
    sz = rng.normal(15.5, 1.4, size=N * T)
    dsz = dv(sz)
    ms = timeit(lambda: ctx.nlsize(T, doff, dsz, o))
    c = cpu_time(lambda: [orc.nlsize(sz[i * N:(i + 1) * N]) for i in range(50)])
    add("nlsize", f"N={N} T={T}", ms, N * T, "rows", 16, 50 * N / c, "50 dates")
    if args.out:
        json.dump({"peak_hbm_GBps": PEAK, "cpu_cores": os.cpu_count(), "rows": rows_out}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
