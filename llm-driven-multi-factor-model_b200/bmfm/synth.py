"""Seeded synthetic Barra panels.

The reference ships no data (``*.csv`` is git-ignored, ``Barra-master/data_new``
is empty), so every parity test, golden fixture and bench run uses panels made
here.  The schema is exactly what ``Barra-master/demo.py:22-35`` hands to
``MFM(data, P, Q)``: columns ``date`` (str ``%Y/%m/%d``), ``stocknames``,
``capital``, ``ret``, P one-hot industry columns, Q style columns.

Nothing here touches the GPU: it is the workload generator, shared by
``tests/``, ``bench.py`` and ``tests/golden/make_golden.py``.
"""
from __future__ import annotations

import numpy as np


class Panel:
    """A dense [T, N] panel with an optional presence mask (ragged universes).

    Attributes
    ----------
    dates      : list[str]            length T, ``%Y/%m/%d``
    stocknames : list[str]            length N
    cap, ret   : float64 [T, N]
    ind        : int32   [N]          industry index in [0, P) (fixed through time)
    style      : float64 [T, N, Q]
    present    : bool    [T, N]       False rows are dropped from the long table
    P, Q       : ints
    """

    def __init__(self, dates, stocknames, cap, ret, ind, style, present, P, Q,
                 industry_names=None, style_names=None):
        self.dates = dates
        self.stocknames = stocknames
        self.cap = cap
        self.ret = ret
        self.ind = ind
        self.style = style
        self.present = present
        self.P = P
        self.Q = Q
        self.industry_names = industry_names or [f"ind_{p:02d}" for p in range(P)]
        self.style_names = style_names or [f"style_{q:02d}" for q in range(Q)]

    @property
    def T(self):
        return self.cap.shape[0]

    @property
    def N(self):
        return self.cap.shape[1]

    # ------------------------------------------------------------------ long table
    def to_frame(self, shuffle_seed=None):
        """Long DataFrame in the reference's positional schema (MFM.py:19-26)."""
        import pandas as pd

        T, N = self.cap.shape
        tt, nn = np.nonzero(self.present)
        if shuffle_seed is not None:
            perm = np.random.default_rng(shuffle_seed).permutation(len(tt))
            tt, nn = tt[perm], nn[perm]
        cols = {
            "date": np.asarray(self.dates, dtype=object)[tt],
            "stocknames": np.asarray(self.stocknames, dtype=object)[nn],
            "capital": self.cap[tt, nn],
            "ret": self.ret[tt, nn],
        }
        onehot = (self.ind[nn][:, None] == np.arange(self.P)[None, :]).astype(np.int64)
        for p, name in enumerate(self.industry_names):
            cols[name] = onehot[:, p]
        for q, name in enumerate(self.style_names):
            cols[name] = self.style[tt, nn, q]
        return pd.DataFrame(cols)

    # ------------------------------------------------------------------ packed CSR
    def to_csr(self):
        """Date-major CSR arrays in stock order (what the C-ABI consumes).

        Returns dict(date_off int64[T+1], cap, ret float64[rows], ind int32[rows],
        style float64[rows, Q]).
        """
        T, N = self.cap.shape
        counts = self.present.sum(axis=1)
        date_off = np.zeros(T + 1, dtype=np.int64)
        np.cumsum(counts, out=date_off[1:])
        tt, nn = np.nonzero(self.present)
        return {
            "date_off": date_off,
            "cap": np.ascontiguousarray(self.cap[tt, nn]),
            "ret": np.ascontiguousarray(self.ret[tt, nn]),
            "ind": np.ascontiguousarray(self.ind[nn].astype(np.int32)),
            "style": np.ascontiguousarray(self.style[tt, nn, :]),
        }


def make_dates(T, start="2010-01-04"):
    import pandas as pd

    return [d.strftime("%Y/%m/%d") for d in pd.bdate_range(start, periods=T)]


def make_panel(T, N, P=31, Q=10, seed=20240002, ragged=0.0, scale_log10=(-1.0, 1.0),
               dtype=np.float64, dates=True):
    """Synthetic panel following SURVEY.md §8(d).

    * industries: first P stocks one per industry (no empty industry), the rest
      uniform on P, fixed through time;
    * ``capital = exp(mu_i + 0.02 * random walk)``, ``mu_i ~ N(15, 1.5)`` — caps
      MUST move through time or the factor-return series is exactly collinear
      and the PSD gate of ``utils.eigen_risk_adj`` (utils.py:67) is a coin flip
      (SURVEY.md §7 hard part 4);
    * styles: AR(1) (rho = 0.98) unit-variance paths times ``scale_q``,
      ``scale_q = 10**U(scale_log10)`` (un-standardised columns);
    * ``ret = X_std f_t + eps``, ``f_t ~ N(0, 0.01^2)``, ``eps ~ N(0, 0.02^2)``.

    ``ragged`` drops that fraction of (stock, date) cells at random.
    """
    rng = np.random.default_rng(seed)
    K = 1 + P + Q
    ind = np.empty(N, dtype=np.int32)
    npref = min(P, N)
    ind[:npref] = np.arange(npref, dtype=np.int32)
    if N > P and P > 0:
        ind[P:] = rng.integers(0, P, size=N - P, dtype=np.int32)
    if P == 0:
        ind[:] = -1

    mu = rng.normal(15.0, 1.5, size=N)
    walk = np.cumsum(rng.normal(0.0, 0.02, size=(T, N)), axis=0)
    cap = np.exp(mu[None, :] + walk)

    scale = 10.0 ** rng.uniform(scale_log10[0], scale_log10[1], size=Q)
    rho = 0.98
    style = np.empty((T, N, Q), dtype=np.float64)
    cur = rng.normal(size=(N, Q))
    innov_sd = np.sqrt(1.0 - rho * rho)
    for t in range(T):
        style[t] = cur
        cur = rho * cur + innov_sd * rng.normal(size=(N, Q))
    style *= scale[None, None, :]
    # non-zero column means, as real un-standardised descriptors have
    style += (rng.normal(size=Q) * scale)[None, None, :]

    # returns from a factor structure on a cheaply standardised exposure
    sstd = (style - style.mean(axis=1, keepdims=True)) / style.std(axis=(1, 2), keepdims=True)
    f = rng.normal(0.0, 0.01, size=(T, K))
    ret = f[:, :1] + (f[:, 1:1 + P][:, ind] if P > 0 else 0.0) + np.einsum("tnq,tq->tn", sstd, f[:, 1 + P:])
    ret = ret + rng.normal(0.0, 0.02, size=(T, N))

    present = np.ones((T, N), dtype=bool)
    if ragged > 0.0:
        present = rng.random(size=(T, N)) >= ragged
        present[:, :npref] = True  # keep every industry populated

    d = make_dates(T) if dates else [f"d{t:05d}" for t in range(T)]
    names = [f"{i:06d}.SZ" for i in range(N)]
    return Panel(d, names, cap.astype(dtype), ret.astype(dtype), ind, style.astype(dtype),
                 present, P, Q)


def make_factor_returns(T, K, seed=20240004):
    """Factor-return series ``N(0, Sigma)`` with eigenvalues ``exp(U(-9,-5))``
    (SURVEY.md §8(d), config C4 stand-in when no regression panel is run)."""
    rng = np.random.default_rng(seed)
    q, _ = np.linalg.qr(rng.normal(size=(K, K)))
    lam = np.exp(rng.uniform(-9.0, -5.0, size=K))
    a = q * np.sqrt(lam)[None, :]
    return rng.normal(size=(T, K)) @ a.T


def make_price_panel(T, S, seed=20240005, missing=0.005, late=0):
    """Close prices for the descriptor build (config C5).

    Geometric random walk (sigma = 0.02) per stock, an index series whose return
    is the equal-weighted mean return plus noise, ``missing`` fraction of closes
    set to NaN and ``late`` stocks that list after a random date (leading NaNs),
    so that ``dropna`` / min-periods logic (factor_calculator.py:91-93,167-169)
    is exercised.
    Returns (close [T, S] float64 with NaN, index_close [T] float64).
    """
    rng = np.random.default_rng(seed)
    mkt = rng.normal(0.0003, 0.01, size=T)
    beta = rng.uniform(0.5, 1.5, size=S)
    r = beta[None, :] * mkt[:, None] + rng.normal(0.0, 0.02, size=(T, S))
    close = 10.0 * np.exp(np.cumsum(np.log1p(np.clip(r, -0.5, 0.5)), axis=0))
    index_close = 3000.0 * np.exp(np.cumsum(np.log1p(mkt + rng.normal(0, 0.001, size=T))))
    if missing > 0:
        close[rng.random(size=(T, S)) < missing] = np.nan
    if late > 0:
        who = rng.choice(S, size=min(late, S), replace=False)
        for s in who:
            close[: rng.integers(1, max(2, T // 2)), s] = np.nan
    return close, index_close


def write_demo_csv_pair(dirpath, T=1400, N=300, P=28, Q=10, seed=20240001, nan_frac=0.002, ragged=0.03):
    """Stand-ins for the two files ``Barra-master/demo.py:22,31`` reads (the real ones are not in the
    reference tree): ``barra_data_csi.csv`` with columns date, stocknames, capital, ret,
    industry (a code), then Q style columns, in shuffled row order and with a few NaN cells so that
    demo.py's NaN-row drop (:25-27) does something; ``industry_info.csv`` with columns code,
    industry_names, start_date (P rows, the order of the one-hot block, demo.py:33-34).
    Floats are written with ``repr`` precision, so the CSV round trip is exact.
    Returns the two paths."""
    import os
    import pandas as pd

    pan = make_panel(T=T, N=N, P=P, Q=Q, seed=seed, ragged=ragged, scale_log10=(-2.0, 2.0))
    rng = np.random.default_rng(seed + 7)
    tt, nn = np.nonzero(pan.present)
    perm = rng.permutation(len(tt))
    tt, nn = tt[perm], nn[perm]
    codes = np.array([f"8011{p:02d}.SI" for p in range(P)], dtype=object)
    style_names = ["size", "beta", "momentum", "residual_volatility", "non_linear_size",
                   "book_to_price_ratio", "liquidity", "earnings_yield", "growth", "leverage"]
    style_names = (style_names + [f"style_{q:02d}" for q in range(10, Q)])[:Q]
    cols = {
        "date": np.asarray(pan.dates, dtype=object)[tt],
        "stocknames": np.asarray(pan.stocknames, dtype=object)[nn],
        "capital": pan.cap[tt, nn],
        "ret": pan.ret[tt, nn],
        "industry": codes[pan.ind[nn]],
    }
    # NaN cells never fall on stocks 0..P-1, so every industry stays populated on every date after the drop
    lead = nn < P
    for q, name in enumerate(style_names):
        v = pan.style[tt, nn, q].copy()
        if nan_frac > 0:
            v[(rng.random(len(v)) < nan_frac / Q) & ~lead] = np.nan
        cols[name] = v
    df = pd.DataFrame(cols)
    os.makedirs(dirpath, exist_ok=True)
    data_path = os.path.join(dirpath, "barra_data_csi.csv")
    ind_path = os.path.join(dirpath, "industry_info.csv")
    df.to_csv(data_path, index=False, float_format="%.17g")
    pd.DataFrame({"code": codes, "industry_names": [f"SW_{p:02d}" for p in range(P)],
                  "start_date": ["2014/01/01"] * P}).to_csv(ind_path, index=False)
    return data_path, ind_path


def make_financial_inputs(T=330, S=8, seed=91):
    """Long daily table with every column FactorCalculator.run(FACTORS_TO_RUN) touches, an index series and
    a Shenwan industry table (what load_and_prepare_data hands to main.py:47): prices, market values, pb / pe,
    turnover, and report-period fundamentals forward-filled between report dates (quarterly end_date,
    single-quarter cash flow, growth rates in percent, liabilities, book value)."""
    import pandas as pd

    close, idx_close = make_price_panel(T, S, seed=seed, missing=0.01, late=1)
    dates = pd.to_datetime(make_dates(T), format="%Y/%m/%d")
    rng = np.random.default_rng(seed + 1)
    rows = []
    for s in range(S):
        base_mv = np.exp(rng.normal(13.0, 1.0))
        nq = T // 60 + 2
        cash = rng.normal(2e3, 3e3, size=nq)                 # single-quarter operating cash flow, some negative
        cash[rng.random(nq) < 0.1] = np.nan
        prof = rng.normal(15.0, 40.0, size=nq)
        sales = rng.normal(10.0, 25.0, size=nq)
        ncl = np.abs(rng.normal(0.3, 0.2, size=nq)) * base_mv
        book = rng.normal(0.5, 0.35, size=nq) * base_mv      # occasionally <= 0: BLEV undefined
        dta = np.clip(rng.normal(45.0, 20.0, size=nq), 1.0, 99.0)
        for t in range(T):
            if s == 3 and 70 <= t < 95:
                continue                                     # absent rows
            qi = t // 60
            mv = base_mv * np.exp(0.01 * rng.normal()) * (close[t, s] / 10.0 if np.isfinite(close[t, s]) else 1.0)
            rows.append(dict(ts_code=f"{s:06d}.SZ", trade_date=dates[t], close=close[t, s],
                             total_mv=mv, circ_mv=0.6 * mv, pb=np.exp(rng.normal(0.7, 0.6)) * (1 if rng.random() > 0.04 else -1),
                             pe_ttm=rng.normal(25.0, 30.0), turnover_rate=np.exp(rng.normal(0.3, 0.8)),
                             end_date=f"{2009 + (qi + 3) // 4}{[3, 6, 9, 12][(qi + 3) % 4]:02d}{[31, 30, 30, 31][(qi + 3) % 4]}",
                             n_cashflow_act=cash[qi], q_profit_yoy=prof[qi], q_sales_yoy=sales[qi],
                             total_ncl=ncl[qi], total_hldr_eqy_inc_min_int=book[qi], debt_to_assets=dta[qi]))
    prices = pd.DataFrame(rows)
    for c in ("total_mv", "pe_ttm", "turnover_rate", "total_ncl"):
        v = prices[c].values.copy()
        v[rng.random(len(v)) < 0.02] = np.nan
        prices[c] = v
    # (total_mv == 0 would make MLEV inf -> NaN, but the reference's own NLSIZE regression then fails on
    #  ln(0) = -inf, so that branch is exercised against the oracle only)
    prices = prices.sample(frac=1.0, random_state=2).reset_index(drop=True)
    index_df = pd.DataFrame({"trade_date": dates, "close": idx_close})
    sw = pd.DataFrame({"ts_code": [f"{s:06d}.SZ" for s in range(S)],
                       "l1_code": [f"8011{s % 3:02d}.SI" for s in range(S)],
                       "l1_name": [f"SW_{s % 3:02d}" for s in range(S)],
                       "in_date": ["20140101"] * S})
    return prices, index_df, sw
