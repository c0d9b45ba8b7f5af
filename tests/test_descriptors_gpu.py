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
