"""GPU parity of the descriptor kernels (returns, BETA/HSIGMA, DASTD, CMRA), the per-date
post-processing kernels and bayes_shrink — against outputs of the unmodified reference
(tests/golden/descriptors.npz, risk.npz) and the CPU oracle."""
import os

import numpy as np
import pytest

from util import relerr

pytestmark = pytest.mark.gpu
TOL = 1e-9


@pytest.fixture(scope="module")
def ops():
    from bmfm import ops
    return ops


@pytest.fixture(scope="module")
def desc(golden_dir):
    return np.load(os.path.join(golden_dir, "descriptors.npz"))


def _panel(desc, key):
    """golden long table (sorted by stock, then date) -> dense [row-of-stock, stock]"""
    s = desc["s"]
    stocks = np.unique(s)
    rows = np.concatenate([np.arange((s == k).sum()) for k in stocks])
    L = rows.max() + 1
    out = np.full((L, len(stocks)), np.nan)
    out[rows, np.searchsorted(stocks, s)] = desc[key]
    return out, rows, np.searchsorted(stocks, s)


def test_descriptors_golden(ops, desc):
    close, rows, cols = _panel(desc, "close")
    mkt, _, _ = _panel(desc, "market_ret")
    ret, lr = ops.prepare_returns(close)
    assert relerr(ret[rows, cols], desc["ret"]) < 1e-13
    assert relerr(lr[rows, cols], desc["log_ret"]) < 1e-12
    beta, hs = ops.beta_hsigma(ret, mkt)
    assert relerr(beta[rows, cols], desc["BETA"]) < TOL
    assert relerr(hs[rows, cols], desc["HSIGMA"]) < TOL
    assert relerr(ops.dastd(ret, mkt)[rows, cols], desc["DASTD"]) < TOL
    assert relerr(ops.cmra(lr)[rows, cols], desc["CMRA"]) < TOL


def test_descriptors_oracle_long_series(ops):
    """Several time chunks per stock (chunk warm-up), NaNs, late listings."""
    from bmfm.synth import make_price_panel
    from oracle import barra_oracle as orc
    T, S = 900, 9
    close, idx = make_price_panel(T, S, seed=3, missing=0.01, late=3)
    mkt1 = np.full(T, np.nan)
    mkt1[1:] = idx[1:] / idx[:-1] - 1
    mkt1[400] = np.nan
    mkt = np.repeat(mkt1[:, None], S, axis=1)
    ret, lr = ops.prepare_returns(close)
    beta, hs = ops.beta_hsigma(ret, mkt)
    da = ops.dastd(ret, mkt)
    cm = ops.cmra(lr)
    for s in range(S):
        r, l = orc.prepare_returns(close[:, s])
        assert relerr(ret[:, s], r) < 1e-13
        wb, wh = orc.beta_hsigma(r, mkt1)
        assert relerr(beta[:, s], wb) < TOL and relerr(hs[:, s], wh) < TOL
        assert relerr(da[:, s], orc.dastd(r - mkt1)) < TOL
        assert relerr(cm[:, s], orc.cmra(l)) < TOL


def _panel_of(desc, key, pre=""):
    s = desc[pre + "s"]
    stocks = np.unique(s)
    rows = np.concatenate([np.arange((s == k).sum()) for k in stocks])
    out = np.full((rows.max() + 1, len(stocks)), np.nan)
    cols = np.searchsorted(stocks, s)
    out[rows, cols] = desc[pre + key]
    return out, rows, cols


@pytest.mark.parametrize("pre", ["", "long_"])
def test_next_descriptors_golden(ops, desc, pre):
    """RSTR and STOM/STOQ/STOA against the unmodified reference (short panel with a suspension run of exact
    zeros / absent rows / late listing, and a series longer than the 483-row RSTR window)."""
    lr, rows, cols = _panel_of(desc, "log_ret", pre)
    got = ops.rstr(lr)[rows, cols]
    assert np.array_equal(np.isnan(got), np.isnan(desc[pre + "RSTR"]))
    assert relerr(got, desc[pre + "RSTR"]) < TOL
    to, _, _ = _panel_of(desc, "turnover_rate", pre)
    for name, o in zip(("STOM", "STOQ", "STOA"), ops.liquidity(to)):
        g = o[rows, cols]
        assert np.array_equal(np.isnan(g), np.isnan(desc[pre + name])), name
        assert relerr(g, desc[pre + name]) < TOL, name


def test_size_and_bp_golden(ops, desc):
    for got, key in ((ops.size_factor(desc["total_mv"]), "SIZE"), (ops.bp_factor(desc["pb"]), "BP")):
        assert np.array_equal(np.isnan(got), np.isnan(desc[key])), key
        assert relerr(got, desc[key]) < 1e-14, key
    assert np.isnan(ops.bp_factor(np.array([-2.0, 0.0, np.nan, np.inf]))[:3]).all()


def test_nlsize_golden_and_edges(ops, desc):
    from oracle import barra_oracle as orc
    t = desc["t"]
    with np.errstate(invalid="ignore"):
        size = np.log(desc["total_mv"])
    order = np.argsort(t, kind="stable")
    off = np.concatenate([[0], np.cumsum(np.bincount(t))]).astype(np.int64)
    got = np.empty(len(t))
    got[order] = ops.nlsize(off, size[order])
    assert np.array_equal(np.isnan(got), np.isnan(desc["NLSIZE"]))
    assert relerr(got, desc["NLSIZE"], scale=np.nanmax(np.abs(size ** 3))) < TOL
    # dates with 0, 1, 2 valid rows, a constant cross-section, and a full A-share cross-section
    rng = np.random.default_rng(4)
    blocks = [np.array([np.nan, np.nan]), np.array([15.0, np.nan]), np.array([14.0, 16.5, np.nan]),
              np.full(7, 13.25), rng.normal(15.5, 1.4, size=5000)]
    blocks[-1][rng.random(5000) < 0.01] = np.nan
    off2 = np.concatenate([[0], np.cumsum([len(b) for b in blocks])]).astype(np.int64)
    x = np.concatenate(blocks)
    g2 = ops.nlsize(off2, x)
    want = np.concatenate([orc.nlsize(b) for b in blocks])
    assert np.array_equal(np.isnan(g2), np.isnan(want))
    assert relerr(g2, want, scale=np.nanmax(np.abs(x ** 3))) < TOL


def test_next_descriptors_oracle_long_series(ops):
    """Several row chunks per stock (window rebuilt at every chunk), NaN runs, zero runs, late listings."""
    from oracle import barra_oracle as orc
    rng = np.random.default_rng(12)
    L, S = 1300, 6
    lr = rng.normal(0, 0.02, size=(L, S))
    lr[rng.random((L, S)) < 0.03] = np.nan
    lr[:200, 1] = np.nan
    lr[600:700, 2] = np.nan
    to = np.exp(rng.normal(0.2, 0.9, size=(L, S)))
    to[rng.random((L, S)) < 0.04] = np.nan
    to[300:420, 3] = 0.0
    to[:, 4] = 0.0
    got = ops.rstr(lr)
    liq = ops.liquidity(to)
    # the single-window entry point gives the same numbers as the fused three-window pass
    one = ops.liquidity(to, windows=((63, 42),))[0]
    assert np.array_equal(np.isnan(one), np.isnan(liq[1])) and relerr(one, liq[1]) < 1e-14
    for s in range(S):
        want = orc.rstr(lr[:, s])
        assert np.array_equal(np.isnan(got[:, s]), np.isnan(want))
        assert relerr(got[:, s], want) < TOL
        for g, w in zip(liq, orc.liquidity(to[:, s])):
            assert np.array_equal(np.isnan(g[:, s]), np.isnan(w))
            if np.isfinite(w).any():
                assert relerr(g[:, s], w) < TOL


def test_post_processing_golden(ops, desc):
    d = desc["pp_date_idx"]
    off = np.concatenate([[0], np.cumsum(np.bincount(d))]).astype(np.int64)
    cols = ["DASTD", "CMRA", "HSIGMA", "BETA", "SIZE", "STOM", "STOQ", "STOA"]
    wz = {}
    for c in cols:
        wz[c] = ops.winsorize(off, desc[f"pp_in_{c}"], 2.5)
        assert relerr(wz[c], desc[f"pp_wz_{c}"]) < 1e-12, c
    vol = ops.composite([wz["DASTD"], wz["CMRA"], wz["HSIGMA"]], [0.7, 0.15, 0.15])
    liq = ops.composite([wz["STOM"], wz["STOQ"], wz["STOA"]], [0.5, 0.25, 0.25])
    assert relerr(vol, desc["pp_comp_volatility"]) < 1e-13
    assert relerr(liq, desc["pp_comp_liquidity"]) < 1e-13
    ov = ops.orthogonalize(off, desc["pp_comp_volatility"], np.column_stack([desc["pp_wz_BETA"], desc["pp_wz_SIZE"]]))
    ol = ops.orthogonalize(off, desc["pp_comp_liquidity"], desc["pp_wz_SIZE"])
    assert relerr(ov, desc["pp_orth_volatility"]) < TOL
    assert relerr(ol, desc["pp_orth_liquidity"]) < TOL


def test_post_processing_oracle_large(ops):
    from oracle import barra_oracle as orc
    rng = np.random.default_rng(2)
    T, N = 7, 3000
    off = (np.arange(T + 1) * N).astype(np.int64)
    y = rng.normal(size=T * N) + 23.0
    X = np.column_stack([rng.normal(size=T * N) * 2 + 23.0, rng.normal(size=T * N)])
    y[rng.random(T * N) < 0.05] = np.nan
    X[rng.random(T * N) < 0.05, 0] = np.nan
    got = ops.orthogonalize(off, y, X)
    wz = ops.winsorize(off, y, 2.5)
    for t in range(T):
        sl = slice(off[t], off[t + 1])
        assert relerr(got[sl], orc.orthogonalize(y[sl], X[sl])) < TOL
        assert relerr(wz[sl], orc.winsorize(y[sl], 2.5)) < 1e-12


@pytest.mark.parametrize("n", [57, 300])
def test_bayes_shrink_golden(ops, golden_dir, n):
    risk = np.load(os.path.join(golden_dir, "risk.npz"))
    vol, cap = risk[f"bs{n}_vol"], risk[f"bs{n}_cap"]
    off = np.array([0, n], dtype=np.int64)
    out, st = ops.bayes_shrink_by_time(off, vol, cap, 10, 1)
    assert st[0] == 0 and relerr(out, risk[f"bs{n}_out_g10_q1"]) < 1e-12
    out, st = ops.bayes_shrink_by_time(off, vol, cap, 5, 0.5)
    assert st[0] == 0 and relerr(out, risk[f"bs{n}_out_g5_q0.5"]) < 1e-12


def test_bayes_shrink_batched_and_ties(ops):
    """Several dates of different sizes; sizes n = 10 k + 1 put decile edges exactly ON data points,
    where only a bit-exact quantile position keeps pd.qcut's codes (integer parity)."""
    from oracle import barra_oracle as orc
    rng = np.random.default_rng(9)
    sizes = [301, 5000, 11, 64, 1001]
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    vol = np.abs(rng.normal(0.02, 0.01, off[-1]))
    cap = np.exp(rng.normal(15, 1.5, off[-1]))
    out, st = ops.bayes_shrink_by_time(off, vol, cap, 10, 1)
    assert (st == 0).all()
    for t in range(len(sizes)):
        sl = slice(off[t], off[t + 1])
        assert relerr(out[sl], orc.bayes_shrink(vol[sl], cap[sl], 10, 1)) < 1e-12
    # duplicate edges -> ValueError through the drop-in
    from mfm import utils
    with pytest.raises(ValueError):
        utils.bayes_shrink(np.ones(20), np.ones(20), 10, 1)
    assert relerr(utils.bayes_shrink(vol[:301], cap[:301]), orc.bayes_shrink(vol[:301], cap[:301])) < 1e-12


def test_factor_calculator_dropin(desc):
    """FactorCalculator / post_processing drop-ins on a long DataFrame in arbitrary row order."""
    import contextlib, io, sys
    import pandas as pd
    from bmfm.synth import make_dates, make_price_panel
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                       "llm-driven-multi-factor-model_b200", "Barra_factor_cal")
    sys.path.insert(0, pkg)
    try:
        import factor_calculator as fc
        import post_processing as pp
    finally:
        sys.path.remove(pkg)
    from oracle import barra_oracle as orc
    T, S = 300, 5
    close, idx = make_price_panel(T, S, seed=8, missing=0.01, late=1)
    dates = pd.to_datetime(make_dates(T), format="%Y/%m/%d")
    rows = [(f"{s:06d}.SZ", dates[t], close[t, s]) for s in range(S) for t in range(T) if not (s == 2 and 40 <= t < 60)]
    prices = pd.DataFrame(rows, columns=["ts_code", "trade_date", "close"]).sample(frac=1.0, random_state=0)
    prices["circ_mv"] = 1.0e6            # run() carries circ_mv in its base frame (factor_calculator.py:541)
    index_df = pd.DataFrame({"trade_date": dates, "close": idx})
    with contextlib.redirect_stdout(io.StringIO()):
        calc = fc.FactorCalculator(prices.copy(), index_df.copy())
        res = calc.run(["BETA", "DASTD", "CMRA"])
        wz = pp.winsorize_factors(res, ["BETA", "DASTD"], 2.5)
    m = calc.master_df
    for code, grp in m.groupby("ts_code"):
        r, l = orc.prepare_returns(grp["close"].values)
        assert relerr(grp["ret"].values, r) < 1e-13
        sel = (res["ts_code"] == code).values
        wb, wh = orc.beta_hsigma(r, grp["market_ret"].values)
        assert relerr(res.loc[sel, "BETA"].values, wb) < TOL
        assert relerr(res.loc[sel, "HSIGMA"].values, wh) < TOL
        assert relerr(res.loc[sel, "CMRA"].values, orc.cmra(l)) < TOL
    for date, grp in res.groupby("trade_date"):
        assert relerr(wz.loc[grp.index, "BETA"].values, orc.winsorize(grp["BETA"].values)) < 1e-12


def test_factor_calculator_next_descriptors_dropin(desc):
    """FactorCalculator.run(['RSTR', 'LIQUIDITY', 'NLSIZE']) of the drop-in on the golden panel (rows shuffled)
    against what the unmodified reference returned for it."""
    import contextlib, io, sys
    import pandas as pd
    from bmfm.synth import make_dates
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                       "llm-driven-multi-factor-model_b200", "Barra_factor_cal")
    sys.path.insert(0, pkg)
    try:
        import factor_calculator as fc
    finally:
        sys.path.remove(pkg)
    T = len(desc["index_close"])
    dates = pd.to_datetime(make_dates(T), format="%Y/%m/%d")
    prices = pd.DataFrame({"ts_code": [f"{k:06d}.SZ" for k in desc["s"]], "trade_date": dates[desc["t"]],
                           "close": desc["close"], "turnover_rate": desc["turnover_rate"],
                           "total_mv": desc["total_mv"], "circ_mv": 0.5 * desc["total_mv"],
                           "key": np.arange(len(desc["s"]))})
    prices = prices.sample(frac=1.0, random_state=5).reset_index(drop=True)
    index_df = pd.DataFrame({"trade_date": dates, "close": desc["index_close"]})
    with contextlib.redirect_stdout(io.StringIO()):
        calc = fc.FactorCalculator(prices.copy(), index_df.copy())
        res = calc.run(["RSTR", "LIQUIDITY", "NLSIZE", "SIZE"])
    res = res.merge(calc.master_df[["ts_code", "trade_date", "key"]], on=["ts_code", "trade_date"]).sort_values("key")
    for c in ["RSTR", "STOM", "STOQ", "STOA", "SIZE"]:
        got = res[c].values
        assert np.array_equal(np.isnan(got), np.isnan(desc[c])), c
        assert relerr(got, desc[c]) < TOL, c
    with np.errstate(invalid="ignore"):
        scale = np.nanmax(np.abs(np.log(desc["total_mv"]) ** 3))
    assert np.array_equal(np.isnan(res["NLSIZE"].values), np.isnan(desc["NLSIZE"]))
    assert relerr(res["NLSIZE"].values, desc["NLSIZE"], scale=scale) < TOL


def _import_factor_cal():
    import importlib
    import sys
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                       "llm-driven-multi-factor-model_b200", "Barra_factor_cal")
    sys.path.insert(0, pkg)
    try:
        for name in ("config", "factor_calculator", "post_processing", "main"):
            sys.modules.pop(name, None)
        return importlib.import_module("main")
    finally:
        sys.path.remove(pkg)


def test_full_factor_pipeline_against_reference(golden_dir):
    """Rows f3 / f4: FactorCalculator.run(FACTORS_TO_RUN) with all eleven entries (incl. EARNINGS, GROWTH,
    LEVERAGE) and main.py's post-processing / Barra-format sequence (main.py:51-137) through the drop-ins,
    against the tables the unmodified reference produced for the same seeded inputs; then the in-memory
    hand-off to the risk model (demo.py:25-35) and the result writers (demo.py:59-94)."""
    import contextlib
    import io
    import tempfile
    import pandas as pd
    from bmfm.handoff import barra_frame_for_mfm, write_results
    from bmfm.synth import make_financial_inputs
    from mfm.MFM import MFM
    from util import record, relerr_cols
    g = np.load(os.path.join(golden_dir, "financial.npz"))
    main = _import_factor_cal()
    prices, index_df, sw = make_financial_inputs()
    with contextlib.redirect_stdout(io.StringIO()):
        calc = main.FactorCalculator(prices.copy(), index_df.copy())
        raw = calc.run(main.config.FACTORS_TO_RUN)
    assert list(raw.columns) == list(g["raw_columns"])
    key = (raw["ts_code"].astype(str) + "|" + raw["trade_date"].astype(str)).values
    assert np.array_equal(key, g["raw_keys"])                       # same rows in the same order
    for c in raw.columns:
        if c in ("ts_code", "trade_date"):
            continue
        scale = None
        if c == "NLSIZE":
            with np.errstate(invalid="ignore", divide="ignore"):
                scale = np.nanmax(np.abs(g["raw__SIZE"] ** 3))
        assert record(f"factors.raw.{c}", relerr(raw[c].values.astype(float), g["raw__" + c], scale=scale)) < TOL, c
    with contextlib.redirect_stdout(io.StringIO()):
        final, ind_info = main.run_factor_pipeline(prices.copy(), index_df.copy(), sw)
    assert list(final.columns) == list(g["final_columns"])
    fkey = (final["stocknames"].astype(str) + "|" + final["date"].astype(str)).values
    assert np.array_equal(fkey, g["final_keys"])
    assert np.array_equal(final["industry"].astype(str).values, g["final_industry"])
    for c in final.columns:
        if c in ("date", "stocknames", "industry"):
            continue
        got = final[c].values.astype(float)
        want = g["final__" + c]
        # residuals of small per-date regressions: errors scale with the regressand, not the residual
        scale = None
        if c == "non_linear_size":
            with np.errstate(invalid="ignore", divide="ignore"):
                scale = np.nanmax(np.abs(g["raw__SIZE"] ** 3))
        assert record(f"factors.final.{c}", relerr(got, want, scale=scale)) < TOL, c
    assert list(ind_info["code"].astype(str)) == list(g["ind_info_code"])
    assert list(ind_info["industry_names"].astype(str)) == list(g["ind_info_names"])
    # hand-off: the in-memory table goes into the risk model the way demo.py feeds the CSV pair
    frame, P, Q = barra_frame_for_mfm(final, ind_info)
    assert (P, Q) == (len(ind_info), 10) and frame.shape[1] == 4 + P + Q and not frame.isna().any().any()
    assert set(np.unique(frame.iloc[:, 4:4 + P].values)) <= {0, 1}
    # (an empty LAST industry makes the reference's constraint row -cap/0, CrossSection.py:70, and numpy raise:
    #  keep the dates on which every industry of this 8-stock universe is present)
    full = frame.groupby("date")[list(frame.columns[4:4 + P])].sum().min(axis=1) > 0
    frame = frame[frame["date"].map(full)].reset_index(drop=True)
    assert frame["date"].nunique() > 20
    with contextlib.redirect_stdout(io.StringIO()):
        model = MFM(frame, P, Q)
        fr, spec, r2 = model.reg_by_time()
    assert fr.shape == (model.T, 1 + P + Q) and np.isfinite(fr.values).all()
    with tempfile.TemporaryDirectory() as d:
        files = write_results(d, model, fr, spec, r2, [fr.cov()], [1.0] * model.T)
        assert sorted(os.path.basename(f) for f in files) == sorted([
            "factor_returns.csv", "r_squared.csv", "specific_returns.csv",
            "final_vol_regime_adj_covariance.csv", "volatility_multiplier_lambda.csv"])
        back = pd.read_csv(os.path.join(d, "factor_returns.csv"), index_col=0)
        assert relerr_cols(back.values, fr.values) < 1e-13      # pandas' fast float parser is not exactly round-trip


def test_two_column_elementwise_ops(ops):
    """bmfm_elementwise2 / bmfm_rolling_sum / BMFM_OP_PERCENT against the oracle, incl. the branches the
    reference run cannot reach (total_mv == 0 makes its own NLSIZE step fail)."""
    from oracle import barra_oracle as orc
    rng = np.random.default_rng(3)
    a = rng.normal(size=5000) * 10
    b = rng.normal(size=5000) * 3
    a[::17] = np.nan; b[::23] = np.nan; b[::29] = 0.0; a[::31] = 0.0
    for got, want in ((ops.cetop_factor(a, b), orc.cetop_factor(a, b)), (ops.mlev_factor(a, b), orc.mlev_factor(a, b)),
                      (ops.blev_factor(a, b), orc.blev_factor(a, b)), (ops.percent_factor(a), orc.percent_factor(a)),
                      (ops.etop_factor(b), orc.etop_factor(b))):
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert relerr(got, want) < 1e-15
    panel = rng.normal(size=(23, 40))
    panel[rng.random(panel.shape) < 0.1] = np.nan
    got = ops.rolling_sum(panel, 4, 4)
    want = np.column_stack([orc.rolling_sum(panel[:, s], 4, 4) for s in range(panel.shape[1])])
    assert np.array_equal(np.isnan(got), np.isnan(want)) and relerr(got, want) < 1e-15
    got = ops.rolling_sum(panel, 5, 2)
    want = np.column_stack([orc.rolling_sum(panel[:, s], 5, 2) for s in range(panel.shape[1])])
    assert np.array_equal(np.isnan(got), np.isnan(want)) and relerr(got, want) < 1e-14


@pytest.mark.gpu
@pytest.mark.parametrize("world", [1, 3])
def test_stock_sharded_post_processing_equals_whole_table(world):
    """the cross-sectional stages on a table whose STOCKS are split over ranks, with the per-date
    moments summed over the ranks between the passes (here: the shards of one GPU, summed by hand), give what the
    single-table entry points give; NaN runs, a date with too few complete rows, a date with a constant column."""
    import torch
    from bmfm import _cabi
    from bmfm.dist import winsorize_sharded, orthogonalize_sharded
    from util import record
    rng = np.random.default_rng(5)
    T, S = 40, 61
    cols = [rng.normal(3.0 * j, 1.0 + j, size=(T, S)) for j in range(3)]
    for c in cols:
        c[rng.random((T, S)) < 0.06] = np.nan
    cols[0][7, :] = np.nan                     # a date with no complete row
    cols[1][9, 3:] = np.nan                    # a date with three complete rows at most (k + 2 = 4 needed)
    cols[2][11, :] = 2.5                       # a constant regressor: rank-deficient Gram, pinv
    ctx = _cabi.default_context()
    dv = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    off_full = dv(np.arange(T + 1, dtype=np.int64) * S)
    full = [dv(c.reshape(-1)) for c in cols]
    want_w = [torch.empty_like(f) for f in full]
    for f, w in zip(full, want_w):
        ctx.winsorize(T, off_full, f, 2.5, w)
    want_o = torch.empty_like(full[0])
    ctx.orthogonalize(T, off_full, full[0], [full[1], full[2]], want_o)
    # the shards: contiguous blocks of stocks, each a [T][S_r] table of its own
    cuts = [S * r // world for r in range(world + 1)]
    shards = [[dv(c[:, cuts[r]:cuts[r + 1]].reshape(-1)) for c in cols] for r in range(world)]
    offs = [dv(np.arange(T + 1, dtype=np.int64) * (cuts[r + 1] - cuts[r])) for r in range(world)]

    # the all-reduce of `world` ranks that run in lockstep, emulated: every rank's call is recorded and replayed
    # with the sum once all ranks have reached the same pass
    import threading
    barrier = threading.Barrier(world)
    slots, lock = {}, threading.Lock()

    def make_allreduce(rank):
        step = [0]

        def allreduce(t):
            key = step[0]
            step[0] += 1
            with lock:
                slots.setdefault(key, []).append(t)
            barrier.wait()
            if rank == 0:
                tot = torch.stack(slots[key]).sum(0)
                for x in slots[key]:
                    x.copy_(tot)
                torch.cuda.synchronize()
            barrier.wait()
        return allreduce

    got_w = [[torch.empty_like(x) for x in shards[r]] for r in range(world)]
    got_o = [torch.empty_like(shards[r][0]) for r in range(world)]
    errs = []

    ctxs = [_cabi.Context(0) for _ in range(world)]         # one context per rank (host thread)

    def run(rank):
        try:
            ar = make_allreduce(rank)
            c = ctxs[rank]
            winsorize_sharded(c, T, offs[rank], shards[rank], 2.5, got_w[rank], ar)
            orthogonalize_sharded(c, T, offs[rank], shards[rank][0], [shards[rank][1], shards[rank][2]], got_o[rank], ar)
        except Exception as e:      # noqa: BLE001
            errs.append(e)
            barrier.abort()

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    torch.cuda.synchronize()
    for c in ctxs:
        c.close()
    for j in range(3):
        got = np.concatenate([got_w[r][j].cpu().numpy().reshape(T, -1) for r in range(world)], axis=1)
        want = want_w[j].cpu().numpy().reshape(T, S)
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert relerr(got, want) < 1e-13, j
    got = np.concatenate([got_o[r].cpu().numpy().reshape(T, -1) for r in range(world)], axis=1)
    want = want_o.cpu().numpy().reshape(T, S)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.isnan(got[7]).all() and np.isnan(got[9]).all()
    ok = np.isfinite(want)
    assert record(f"orthogonalize.sharded.w{world}", relerr(got[ok], want[ok])) < 1e-9
