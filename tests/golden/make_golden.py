"""Generate tests/golden/*.npz by running the UNMODIFIED reference source.

Run in the build container only (``/root/reference`` mounted):

    python tests/golden/make_golden.py

Every fixture stores the inputs and the reference's outputs, so the tests need
neither the reference nor pandas semantics at run time.  Inputs come from the
seeded generators in ``bmfm/synth.py``; statsmodels/matplotlib are stubbed as
described in ``ref_harness.py``.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))

from ref_harness import load_reference, quiet, patched_eig, canonical_eig  # noqa: E402
from bmfm.synth import make_panel, make_factor_returns, make_price_panel, make_dates, make_financial_inputs  # noqa: E402

from bmfm.handoff import read_results  # noqa: E402

ref_mfm, ref_cs, ref_utils, ref_fc, ref_pp, ref_cfg = load_reference()


def cov_list_to_array(lst, K):
    """list of DataFrames (empty allowed) -> (array [T,K,K] NaN-filled, ok mask)."""
    T = len(lst)
    out = np.full((T, K, K), np.nan)
    ok = np.zeros(T, dtype=bool)
    for t, c in enumerate(lst):
        if c is not None and len(c) > 0:
            out[t] = np.asarray(c.values, dtype=float)
            ok[t] = True
    return out, ok


def run_reg(panel, shuffle_seed=7):
    df = panel.to_frame(shuffle_seed=shuffle_seed)
    with quiet():
        model = ref_mfm.MFM(df, panel.P, panel.Q)
        fr, spec, r2 = model.reg_by_time()
    csr = panel.to_csr()
    # the reference sorts each date by stocknames (MFM.py:59); synth names sort in stock order
    spec_flat = np.concatenate([s.values[0] for s in spec])
    names_flat = np.concatenate([np.asarray(s.columns) for s in spec])
    return model, df, csr, fr.values, spec_flat, r2.values[:, 0], names_flat


def golden_reg():
    out = {}
    cases = {
        "small": dict(T=6, N=60, P=5, Q=4, seed=11, ragged=0.1, scale_log10=(-2, 2)),
        "csi": dict(T=3, N=150, P=31, Q=10, seed=12, ragged=0.05, scale_log10=(-2, 2)),
        "n_lt_k": dict(T=2, N=35, P=31, Q=10, seed=13, ragged=0.0, scale_log10=(-1, 1)),
        "nostyle_scale": dict(T=2, N=80, P=3, Q=1, seed=14, ragged=0.0, scale_log10=(0, 0)),
    }
    for name, kw in cases.items():
        pan = make_panel(**kw)
        if name == "small":
            # empty NON-last industry on date 2: pinv min-norm branch (SURVEY hard part 3)
            pan.present[2, pan.ind == 1] = False
        model, df, csr, fr, spec, r2, names = run_reg(pan)
        for k, v in csr.items():
            out[f"{name}__{k}"] = v
        out[f"{name}__P"] = np.int64(pan.P)
        out[f"{name}__factor_ret"] = fr
        out[f"{name}__specific_ret"] = spec
        out[f"{name}__R2"] = r2
        # one explicit CrossSection with its exposure matrix (CrossSection.py:104)
        d0 = df[df.date == pan.dates[0]].sort_values("stocknames")
        with quiet():
            cs = ref_cs.CrossSection(d0.iloc[:, :4], d0.iloc[:, -pan.Q:], d0.iloc[:, 4:4 + pan.P])
            f, s, expo, r = cs.reg()
        out[f"{name}__exposure0"] = np.asarray(expo, dtype=float)
        out[f"{name}__style_norm0"] = np.asarray(cs.style_factors, dtype=float)
        out[f"{name}__W0"] = np.asarray(cs.W, dtype=float)
    np.savez_compressed(os.path.join(HERE, "reg.npz"), **out)
    print("reg.npz", {k: v.shape for k, v in out.items() if k.endswith("factor_ret")})


def golden_risk():
    out = {}
    # Newey-West on a plain series
    K = 6
    fr = make_factor_returns(70, K, seed=21)
    frdf = pd.DataFrame(fr, columns=[f"f{i}" for i in range(K)])
    for (q, tao) in [(2, 10), (0, 252), (3, 40)]:
        lst = []
        for t in range(1, len(fr) + 1):
            try:
                lst.append(ref_utils.Newey_West(frdf[:t], q, tao))
            except Exception:
                lst.append(pd.DataFrame())
        arr, ok = cov_list_to_array(lst, K)
        out[f"nw_q{q}_tao{tao}"] = arr
        out[f"nw_q{q}_tao{tao}_ok"] = ok
    out["nw_factor_ret"] = fr

    # eigen_risk_adj: literal and canonical on one covariance
    cov = pd.DataFrame(np.cov(fr.T), columns=frdf.columns, index=frdf.columns)
    out["eig_cov_in"] = cov.values
    for (T, M) in [(40, 5), (252, 8)]:
        lit = ref_utils.eigen_risk_adj(cov, T, M, 1.4)
        with patched_eig(ref_utils):
            can = ref_utils.eigen_risk_adj(cov, T, M, 1.4)
        out[f"eig_T{T}_M{M}_literal"] = lit.values
        out[f"eig_T{T}_M{M}_canonical"] = can.values
    # the legacy gaussian stream itself (first 64 normals of seeds 1,2,100)
    for seed in (1, 2, 100):
        np.random.seed(seed)
        out[f"gauss_seed{seed}"] = np.random.standard_normal(64)
    np.random.seed(3)
    out["mvn_seed3"] = np.random.multivariate_normal(mean=3 * [0], cov=np.diag([3.0, 2.0, 0.5]), size=5)

    # bayes_shrink
    rng = np.random.default_rng(31)
    for n in (57, 300):
        vol = np.abs(rng.normal(0.02, 0.01, size=n))
        cap = np.exp(rng.normal(15, 1.5, size=n))
        with quiet():
            sh = ref_utils.bayes_shrink(vol, cap, 10, 1)
            sh2 = ref_utils.bayes_shrink(vol, cap, 5, 0.5)
        out[f"bs{n}_vol"] = vol
        out[f"bs{n}_cap"] = cap
        out[f"bs{n}_out_g10_q1"] = np.asarray(sh, dtype=float)
        out[f"bs{n}_out_g5_q0.5"] = np.asarray(sh2, dtype=float)
        out[f"bs{n}_codes_g10"] = np.asarray(pd.qcut(cap, 10).codes, dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "risk.npz"), **out)
    print("risk.npz", len(out))


def golden_pipeline():
    """demo.py's four calls (demo.py:38-42) on a small panel, canonical eig."""
    out = {}
    pan = make_panel(T=90, N=48, P=4, Q=3, seed=41, ragged=0.05, scale_log10=(-1, 1))
    df = pan.to_frame(shuffle_seed=3)
    K = 1 + pan.P + pan.Q
    with quiet(), patched_eig(ref_utils), patched_eig(ref_mfm):
        # MFM.py imports eigen_risk_adj by name from mfm.utils, so patching utils suffices;
        model = ref_mfm.MFM(df, pan.P, pan.Q)
        fr, spec, r2 = model.reg_by_time()
        nw = model.Newey_West_by_time(q=2, tao=30)
        er = model.eigen_risk_adj_by_time(M=6, scale_coef=1.4)
        vr, lamb = model.vol_regime_adj_by_time(tao=12)
    for k, v in pan.to_csr().items():
        out[k] = v
    out["P"] = np.int64(pan.P)
    out["factor_ret"] = fr.values
    out["specific_ret"] = np.concatenate([s.values[0] for s in spec])
    out["R2"] = r2.values[:, 0]
    out["last_capital"] = np.asarray(model.last_capital, dtype=float)
    out["nw"], out["nw_ok"] = cov_list_to_array(nw, K)
    out["er"], out["er_ok"] = cov_list_to_array(er, K)
    out["vr"], out["vr_ok"] = cov_list_to_array(vr, K)
    out["lamb"] = np.asarray(lamb, dtype=float)
    # literal run (LAPACK order/sign dependent) kept beside it to document the gap
    with quiet():
        model2 = ref_mfm.MFM(df, pan.P, pan.Q)
        model2.reg_by_time()
        model2.Newey_West_by_time(q=2, tao=30)
        er2 = model2.eigen_risk_adj_by_time(M=6, scale_coef=1.4)
    out["er_literal"], out["er_literal_ok"] = cov_list_to_array(er2, K)
    with quiet(), patched_eig(ref_utils):
        out["bias_stat_nw"] = ref_utils.eigenfactor_bias_stat(nw[40:], fr[40:], predlen=5)
    np.savez_compressed(os.path.join(HERE, "pipeline.npz"), **out)
    # the same panel as a CSV pair in demo.py's on-disk schema (demo.py:22-35)
    print("pipeline.npz", len(out), "er ok:", int(out["er_ok"].sum()), "of", len(out["er_ok"]))


def golden_descriptors():
    out = {}
    T, S = 330, 7
    close, idx_close = make_price_panel(T, S, seed=51, missing=0.01, late=2)
    close[:, 6] = np.nan
    close[200:, 6] = 10.0 * np.exp(np.cumsum(np.random.default_rng(5).normal(0, 0.02, T - 200)))
    dates = pd.to_datetime(make_dates(T), format="%Y/%m/%d")
    rows = []
    for s in range(S):
        # stock 5 has ABSENT rows (not NaN rows): the window is over the stock's own rows
        keep = np.ones(T, bool)
        if s == 5:
            keep[50:80] = False
        for t in np.nonzero(keep)[0]:
            rows.append((f"{s:06d}.SZ", dates[t], close[t, s], t, s))
    prices = pd.DataFrame(rows, columns=["ts_code", "trade_date", "close", "t", "s"])
    # columns of the "next" descriptors: turnover in percent with suspension runs of exact
    # zeros and missing days; total market value with a few missing rows
    rng2 = np.random.default_rng(77)
    to = np.exp(rng2.normal(0.3, 0.8, size=len(prices)))
    to[rng2.random(len(prices)) < 0.03] = np.nan
    susp = (prices["s"].values == 2) & (prices["t"].values >= 100) & (prices["t"].values < 170)
    to[susp] = 0.0
    prices["turnover_rate"] = to
    mv = np.exp(rng2.normal(13.0, 1.2, size=S))[prices["s"].values] * np.exp(rng2.normal(0, 0.02, size=len(prices)))
    mv[rng2.random(len(prices)) < 0.02] = np.nan
    prices["total_mv"] = mv
    pb = np.exp(rng2.normal(0.7, 0.6, size=len(prices)))
    pb[rng2.random(len(prices)) < 0.04] *= -1.0          # negative book value: BP undefined (factor_calculator.py:313)
    pb[rng2.random(len(prices)) < 0.01] = 0.0
    pb[rng2.random(len(prices)) < 0.02] = np.nan
    prices["pb"] = pb
    prices = prices.sample(frac=1.0, random_state=1).reset_index(drop=True)   # arbitrary input order
    index_df = pd.DataFrame({"trade_date": dates, "close": idx_close})
    with quiet():
        calc = ref_fc.FactorCalculator(prices.copy(), index_df.copy())
        bh = calc.compute_beta_hsigma()
        da = calc.compute_dastd()
        cm = calc.compute_cmra()
        rs = calc.compute_rstr()
        lq = calc.compute_liquidity()
        nl = calc.compute_nlsize()
        sz = calc.compute_size()
        bp = calc.compute_bp()
    m = calc.master_df
    res = m[["original_index", "t", "s", "close", "ret", "log_ret", "market_ret", "turnover_rate", "total_mv", "pb"]] \
        .merge(sz, on="original_index").merge(bp, on="original_index") \
        .merge(bh, on="original_index").merge(da, on="original_index").merge(cm, on="original_index") \
        .merge(rs, on="original_index").merge(lq, on="original_index").merge(nl, on="original_index")
    res = res.sort_values(["s", "t"])
    for c in ["t", "s"]:
        out[c] = res[c].values.astype(np.int64)
    for c in ["close", "ret", "log_ret", "market_ret", "BETA", "HSIGMA", "DASTD", "CMRA", "turnover_rate", "total_mv",
              "RSTR", "STOM", "STOQ", "STOA", "NLSIZE", "pb", "SIZE", "BP"]:
        out[c] = res[c].values.astype(float)
    out["index_close"] = idx_close

    # a series longer than the RSTR window (483 rows) so that rows leave the window, two stocks
    T2, S2 = 640, 2
    close2, idx2 = make_price_panel(T2, S2, seed=52, missing=0.02, late=1)
    dates2 = pd.to_datetime(make_dates(T2), format="%Y/%m/%d")
    rows2 = [(f"{s:06d}.SZ", dates2[t], close2[t, s], t, s) for s in range(S2) for t in range(T2)]
    prices2 = pd.DataFrame(rows2, columns=["ts_code", "trade_date", "close", "t", "s"])
    rng3 = np.random.default_rng(78)
    to2 = np.exp(rng3.normal(0.0, 0.7, size=len(prices2)))
    to2[rng3.random(len(prices2)) < 0.05] = np.nan
    prices2["turnover_rate"] = to2
    with quiet():
        calc2 = ref_fc.FactorCalculator(prices2.copy(), pd.DataFrame({"trade_date": dates2, "close": idx2}))
        rs2 = calc2.compute_rstr()
        lq2 = calc2.compute_liquidity()
    res2 = calc2.master_df[["original_index", "t", "s", "log_ret", "turnover_rate"]].merge(rs2, on="original_index") \
        .merge(lq2, on="original_index").sort_values(["s", "t"])
    for c in ["t", "s"]:
        out["long_" + c] = res2[c].values.astype(np.int64)
    for c in ["log_ret", "turnover_rate", "RSTR", "STOM", "STOQ", "STOA"]:
        out["long_" + c] = res2[c].values.astype(float)

    # post-processing on a small long table
    rng = np.random.default_rng(61)
    nd, ns = 5, 40
    frame = pd.DataFrame({
        "trade_date": np.repeat([f"2020/01/{d + 1:02d}" for d in range(nd)], ns),
        "ts_code": np.tile([f"{i:06d}.SZ" for i in range(ns)], nd),
    })
    for c in ["DASTD", "CMRA", "HSIGMA", "BETA", "SIZE", "STOM", "STOQ", "STOA"]:
        v = rng.normal(size=nd * ns) * (1 + 3 * (rng.random(nd * ns) < 0.05))
        v[rng.random(nd * ns) < 0.08] = np.nan
        frame[c] = v
    frame.loc[frame.trade_date == "2020/01/05", "BETA"] = np.nan        # a date with no complete rows
    frame.loc[(frame.trade_date == "2020/01/04") & (frame.index % ns > 1), "STOA"] = np.nan   # n<2 stats
    cols = ["DASTD", "CMRA", "HSIGMA", "BETA", "SIZE", "STOM", "STOQ", "STOA"]
    with quiet():
        wz = ref_pp.winsorize_factors(frame, cols, 2.5)
        cfg = {k: ref_cfg.COMPOSITE_CONFIG[k] for k in ("volatility", "liquidity")}
        comp = ref_pp.calculate_composite_factors(wz, cfg)
        orth = ref_pp.orthogonalize_factors(comp, ref_cfg.ORTHO_RULES)
    out["pp_date_idx"] = np.repeat(np.arange(nd), ns).astype(np.int64)
    for c in cols:
        out[f"pp_in_{c}"] = frame[c].values.astype(float)
        out[f"pp_wz_{c}"] = wz[c].values.astype(float)
    for c in ("volatility", "liquidity"):
        out[f"pp_comp_{c}"] = comp[c].values.astype(float)
        out[f"pp_orth_{c}"] = orth[c].values.astype(float)
    np.savez_compressed(os.path.join(HERE, "descriptors.npz"), **out)
    print("descriptors.npz", len(out), "beta valid:", int(np.isfinite(out["BETA"]).sum()))


def golden_financial():
    """Rows f3/f4: FactorCalculator.run(config.FACTORS_TO_RUN) (all eleven entries, incl. EARNINGS / GROWTH /
    LEVERAGE, factor_calculator.py:371-509), then main.py's own sequence with the reference's functions and
    constants — winsorise every column but the keys (main.py:71-72), composites (:79), orthogonalise (:86),
    industry merge and next-day return (:98-99), rename (:107), column order (:111-112), industry table
    (:128-137) — without the MongoDB calls around it."""
    out = {}
    prices, index_df, sw = make_financial_inputs()
    with quiet():
        calc = ref_fc.FactorCalculator(prices.copy(), index_df.copy())
        raw = calc.run(ref_cfg.FACTORS_TO_RUN)
        factor_columns = [c for c in raw.columns if c not in ['ts_code', 'trade_date']]
        wz = ref_pp.winsorize_factors(raw, factor_columns)
        comp = ref_pp.calculate_composite_factors(wz, ref_cfg.COMPOSITE_CONFIG)
        proc = ref_pp.orthogonalize_factors(comp, ref_cfg.ORTHO_RULES)
        barra = pd.merge(proc, sw[['ts_code', 'l1_code']], on='ts_code', how='left')
        barra['ret'] = barra.groupby('ts_code')['ret'].shift(-1)
        barra.rename(columns=ref_cfg.COLUMN_RENAME_MAP, inplace=True)
        final_columns = [c for c in ref_cfg.BARRA_OUTPUT_COLUMNS if c in barra.columns]
        final = barra[final_columns]
        stk = final[['stocknames']].drop_duplicates().rename(columns={'stocknames': 'ts_code'})
        ind_info = pd.merge(stk, sw[['ts_code', 'l1_code', 'l1_name', 'in_date']], on='ts_code', how='left')
        ind_info = ind_info.drop_duplicates(subset=['l1_code', 'l1_name']).rename(
            columns={'l1_code': 'code', 'l1_name': 'industry_names', 'in_date': 'start_date'})[['code', 'industry_names', 'start_date']]
    out["raw_columns"] = np.asarray(raw.columns.astype(str)).astype("U")
    out["raw_keys"] = np.asarray(raw["ts_code"].astype(str) + "|" + raw["trade_date"].astype(str)).astype("U")
    for c in raw.columns:
        if c not in ("ts_code", "trade_date"):
            out["raw__" + c] = raw[c].values.astype(float)
    out["final_columns"] = np.asarray(final.columns.astype(str)).astype("U")
    out["final_keys"] = np.asarray(final["stocknames"].astype(str) + "|" + final["date"].astype(str)).astype("U")
    out["final_industry"] = np.asarray(final["industry"].astype(str)).astype("U")
    for c in final.columns:
        if c not in ("date", "stocknames", "industry"):
            out["final__" + c] = final[c].values.astype(float)
    out["ind_info_code"] = np.asarray(ind_info["code"].astype(str)).astype("U")
    out["ind_info_names"] = np.asarray(ind_info["industry_names"].astype(str)).astype("U")
    np.savez_compressed(os.path.join(HERE, "financial.npz"), **out)
    print("financial.npz", len(out), list(out["final_columns"]),
          {c: int(np.isfinite(out["raw__" + c]).sum()) for c in ("CETOP", "ETOP", "MLEV", "BLEV", "YOYProfit")})


DEMO_CASE = dict(T=1400, N=300, P=28, Q=10, seed=20240001)     # config C1; demo.py:38 hard-codes MFM(data, 28, 10)


def _run_reference_demo(workdir, canonical):
    """Execute the reference's own Barra-master/demo.py, byte-unchanged, from a scratch copy of its directory
    layout (demo.py next to data_new/), with the reference `mfm` first on sys.path."""
    import runpy
    import shutil
    bm = os.path.join(workdir, "Barra-master")
    os.makedirs(bm, exist_ok=True)
    shutil.copyfile("/root/reference/Barra-master/demo.py", os.path.join(bm, "demo.py"))
    sys.path.insert(0, "/root/reference/Barra-master")
    for name in [m for m in sys.modules if m == "mfm" or m.startswith("mfm.")]:
        del sys.modules[name]
    try:
        import mfm.utils as live_utils
        assert live_utils.__file__.startswith("/root/reference"), live_utils.__file__
        import contextlib
        cm = patched_eig(live_utils) if canonical else contextlib.nullcontext()
        with quiet(), cm:
            runpy.run_path(os.path.join(bm, "demo.py"), run_name="__main__")
    finally:
        sys.path.remove("/root/reference/Barra-master")
        for name in [m for m in sys.modules if m == "mfm" or m.startswith("mfm.")]:
            del sys.modules[name]
    return os.path.join(bm, "results")


def golden_demo():
    """Config C1 / row g: the reference's demo.py run UNCHANGED on the stand-in CSV pair (the
    real files are not in the tree), eigen order canonical.  The fixture also carries demo.py's bytes so the
    GPU test can execute the very same file where /root/reference does not exist."""
    import hashlib
    import tempfile
    from bmfm.synth import write_demo_csv_pair
    out = {}
    with tempfile.TemporaryDirectory() as wd:
        dpath, ipath = write_demo_csv_pair(os.path.join(wd, "Barra-master", "data_new"), **DEMO_CASE)
        out["csv_sha256"] = np.array([hashlib.sha256(open(p, "rb").read()).hexdigest() for p in (dpath, ipath)])
        res = read_results(_run_reference_demo(wd, canonical=True))
    src = open("/root/reference/Barra-master/demo.py", "rb").read()
    out["demo_py"] = np.frombuffer(src, dtype=np.uint8)
    out["demo_py_sha256"] = np.array(hashlib.sha256(src).hexdigest())
    for k, v in DEMO_CASE.items():
        out["case_" + k] = np.int64(v)
    # specific returns: every 8th date and the last 40 in full, per-date sums of all dates as a checksum
    full = res.pop("specific_full")
    T = full.shape[0]
    keep = np.unique(np.concatenate([np.arange(0, T, 8), np.arange(max(0, T - 40), T)]))
    out["specific_keep"] = keep.astype(np.int64)
    out["specific_rows"] = full[keep]
    out["specific_sum"] = np.nansum(full, axis=1)
    out["specific_sumsq"] = np.nansum(full * full, axis=1)
    out["specific_count"] = np.isfinite(full).sum(axis=1).astype(np.int64)
    out.update(res)
    np.savez_compressed(os.path.join(HERE, "demo.npz"), **out)
    print("demo.npz", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    which = sys.argv[1:] or ["reg", "risk", "pipeline", "descriptors", "financial", "demo"]
    for w in which:
        globals()[f"golden_{w}"]()
