"""CPU oracle for the Barra risk-model path — TEST INFRASTRUCTURE ONLY.

A plain numpy restatement of the reference's algorithms for the hot path of
SURVEY.md §8(a).  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module; the product
(``llm-driven-multi-factor-model_b200/``) never does and has no CPU fallback.

Parity status: the reference has NO tests, golden vectors or fixtures for this
path.  The oracle is therefore pinned against outputs of
the UNMODIFIED reference source executed in the build container on seeded
synthetic panels: ``tests/golden/make_golden.py`` produced ``tests/golden/*.npz``
and ``tests/test_oracle_golden.py`` checks every function here against them.
For ``eigen_risk_adj`` the pin is the reference source with ``np.linalg.eig``
patched to the canonical order/sign convention (SURVEY.md §8(c) mode (ii)); the
literal (LAPACK-order-dependent) output is recorded beside it.  For the
statsmodels-backed descriptor/post-processing functions the pin is "reference
source + restated statsmodels stub" (statsmodels is not installed and its
version is pinned nowhere in the reference).

Every function cites the reference file:line it follows (paths relative to
/root/reference).
"""
from __future__ import annotations

import numpy as np

# ----------------------------------------------------------------------------
# Part 1 — cross-sectional regression
# ----------------------------------------------------------------------------


def style_factor_norm(factors, capital):
    """Barra-master/mfm/CrossSection.py:12-20.

    Cap-weighted mean per column (statsmodels ``DescrStatsW.mean`` =
    ``data.T @ w / w.sum()``), divided by ONE scalar ``np.std`` over the whole
    N x Q matrix (ddof 0)."""
    factors = np.asarray(factors, dtype=np.float64)
    capital = np.asarray(capital, dtype=np.float64)
    w_mu = np.dot(factors.T, capital) / capital.sum(0)
    w_std = np.std(factors)
    return (factors - w_mu) / w_std


def cross_section_reg(cap, ret, industry_onehot, style_raw):
    """One date.  Barra-master/mfm/CrossSection.py:35-108 (ctor + ``reg``).

    Restated without the dense N x N ``np.diag(W)`` of CrossSection.py:62,76
    (``A.T @ diag(w) @ B == (A * w[:, None]).T @ B``); everything else — the
    python-``sum`` weights (:50), the industry-cap constraint matrix R (:66-71),
    ``np.linalg.pinv`` of the transformed Gram (:76), exposure, specific return
    and the unweighted ddof-0 R2 (:101-106) — is as in the reference.

    Returns (factor_ret[K], specific_ret[N], exposure[K, K], R2).
    """
    cap = np.asarray(cap, dtype=np.float64)
    ret = np.asarray(ret, dtype=np.float64)
    N = cap.shape[0]
    industry = np.asarray(industry_onehot)
    P = industry.shape[1] if industry.ndim == 2 else 0
    style_raw = np.asarray(style_raw, dtype=np.float64)
    Q = style_raw.shape[1] if style_raw.ndim == 2 else 0
    style = style_factor_norm(style_raw, cap) if Q > 0 else np.zeros((N, 0))
    country = np.ones((N, 1))
    sq = np.sqrt(cap)
    W = sq / sum(sq)                                       # CrossSection.py:50 (python sum)

    if P > 0:
        industry_capital = np.array([sum(industry[:, i] * cap) for i in range(P)])   # :66
        K = 1 + P + Q
        R = np.eye(K)
        R[P, 1:(1 + P)] = -industry_capital / industry_capital[-1]                   # :70
        R = np.delete(R, P, axis=1)                                                  # :71
        factors = np.hstack([country, industry.astype(np.float64), style])
        ft = factors @ R
        ftw = ft.T * W[None, :]                                                      # (XR)^T diag(W)
        weight = R @ np.linalg.pinv(ftw @ ft) @ ftw                                  # :76
    else:
        factors = np.hstack([country, style])
        fw = factors.T * W[None, :]
        weight = np.linalg.pinv(fw @ factors) @ fw                                   # :98

    factor_ret = weight @ ret                                                        # :101
    exposure = weight @ factors                                                      # :104
    specific = ret - factors @ factor_ret                                            # :105
    R2 = 1.0 - np.var(specific) / np.var(ret)                                        # :106
    return factor_ret, specific, exposure, R2


def reg_by_time(date_off, cap, ret, ind, style, P):
    """Barra-master/mfm/MFM.py:48-76 on packed arrays (rows already grouped by
    date and sorted by stock name inside a date, as MFM.py:58-59 does).

    ``ind`` is the industry index per row (int, -1 = no industry column set);
    the one-hot block the reference receives is rebuilt from it.
    Returns (factor_ret [T, K], specific_ret [rows], R2 [T])."""
    T = len(date_off) - 1
    Q = style.shape[1]
    K = 1 + P + Q
    fr = np.empty((T, K))
    spec = np.empty(cap.shape[0])
    r2 = np.empty(T)
    ar = np.arange(P)
    for t in range(T):
        a, b = int(date_off[t]), int(date_off[t + 1])
        onehot = (np.asarray(ind[a:b])[:, None] == ar[None, :]).astype(np.int64)
        f, s, _, r = cross_section_reg(cap[a:b], ret[a:b], onehot, style[a:b])
        fr[t] = f
        spec[a:b] = s
        r2[t] = r
    return fr, spec, r2


# ----------------------------------------------------------------------------
# Part 3 — risk-model adjustments
# ----------------------------------------------------------------------------


def newey_west(ret, q=2, tao=252):
    """Barra-master/mfm/utils.py:16-50 on a [T, K] array.  Raises like the
    reference when ``T <= q or T <= K`` (utils.py:29)."""
    ret = np.asarray(ret, dtype=np.float64)
    T, K = ret.shape
    if T <= q or T <= K:
        raise Exception("T <= q or T <= K")
    weights = 0.5 ** (np.arange(T - 1, -1, -1) / tao)
    weights = weights / sum(weights)                       # utils.py:33 (python sum)
    mean = np.dot(ret.T, weights) / weights.sum(0)         # DescrStatsW.mean, utils.py:36-37
    x = ret - mean
    V = (x * weights[:, None]).T @ x                       # Gamma0, utils.py:40-41
    for i in range(1, q + 1):
        # Gamma_i = sum_t w[i+t] x[t]^T x[i+t], utils.py:46
        G = (x[: T - i] * weights[i:, None]).T @ x[i:]
        V = V + (1 - i / (1 + q)) * (G + G.T)              # utils.py:48
    return V


def newey_west_by_time(factor_ret, q=2, tao=252):
    """Barra-master/mfm/MFM.py:80-101: expanding window, ``None`` where the
    reference appends an empty DataFrame (exception swallowed, MFM.py:95-96)."""
    T = factor_ret.shape[0]
    out = []
    for t in range(1, T + 1):
        try:
            out.append(newey_west(factor_ret[:t], q, tao))
        except Exception:
            out.append(None)
    return out


def canonical_eig(a):
    """SURVEY.md §8(c) mode (ii): eigh of the symmetrised matrix, eigenvalues
    descending, eigenvector sign fixed so its largest-|.| component is > 0."""
    a = np.asarray(a, dtype=np.float64)
    w, v = np.linalg.eigh(0.5 * (a + a.T))
    order = np.argsort(-w, kind="stable")
    w = w[order]
    v = v[:, order]
    idx = np.argmax(np.abs(v), axis=0)
    sgn = np.sign(v[idx, np.arange(v.shape[1])])
    sgn[sgn == 0] = 1.0
    return w, v * sgn[None, :]


def eigen_risk_adj(covmat, T=1000, M=100, scale_coef=1.4, eig=canonical_eig, return_sims=False):
    """Barra-master/mfm/utils.py:55-92 with the eigensolver injectable.

    ``eig=np.linalg.eig`` is the literal reference; the default is the
    canonical convention the CUDA path implements.  The simulated returns come
    from the same legacy stream as the reference's ``np.random.seed(m+1)`` +
    ``np.random.multivariate_normal`` (a private ``RandomState(m+1)`` is the
    same MT19937 + polar-gauss stream without clobbering the global state).
    Raises TypeError on a non-PSD input exactly like the reference's
    ``raise('...')`` of a str (utils.py:67-68)."""
    F0 = np.asarray(covmat, dtype=np.float64)
    K = F0.shape[0]
    D0, U0 = eig(F0)
    if not all(D0 >= 0):
        raise TypeError("exceptions must derive from BaseException")
    v = []
    sims = []
    for m in range(M):
        rs = np.random.RandomState(m + 1)
        bm = rs.multivariate_normal(mean=K * [0], cov=np.diag(D0), size=T).T   # utils.py:74
        fm = U0 @ bm
        Fm = np.cov(fm)
        Dm, Um = eig(Fm)
        Dm_hat = Um.T @ F0 @ Um
        v.append(np.diagonal(Dm_hat) / Dm)
        if return_sims:
            sims.append((Dm.copy(), np.diagonal(Dm_hat).copy()))
    vv = np.sqrt(np.mean(np.array(v), axis=0))
    vv = scale_coef * (vv - 1) + 1
    D0_hat = np.diag(vv ** 2) * np.diag(D0)
    F0_hat = U0 @ D0_hat @ U0.T
    if return_sims:
        return F0_hat, vv, sims
    return F0_hat


def gauss_cov(M, T, K):
    """``C_m = np.cov(z_m.T)`` (ddof 1) of ``z_m = RandomState(m+1).standard_normal((T, K))``
    — the sample covariance every date shares because the reference re-seeds
    with ``m+1`` for every date (utils.py:73)."""
    out = np.empty((M, K, K))
    for m in range(M):
        z = np.random.RandomState(m + 1).standard_normal((T, K))
        out[m] = np.cov(z.T)
    return out


def eigen_risk_adj_by_time(nw_cov, T_sim, M=100, scale_coef=1.4, eig=canonical_eig):
    """Barra-master/mfm/MFM.py:105-126: per date, exception -> None."""
    out = []
    for c in nw_cov:
        try:
            if c is None:
                raise ValueError("empty")
            out.append(eigen_risk_adj(c, T_sim, M, scale_coef, eig))
        except Exception:
            out.append(None)
    return out


def vol_regime_adj_by_time(factor_ret, eig_cov, tao=84):
    """Barra-master/mfm/MFM.py:130-167 on arrays; ``eig_cov`` is a list of K x K
    arrays or None.  Returns (list of K x K / None, lamb list of float)."""
    factor_ret = np.asarray(factor_ret, dtype=np.float64)
    T, K = factor_ret.shape
    factor_var = np.empty((T, K))
    for t in range(T):
        factor_var[t] = np.diag(eig_cov[t]) if eig_cov[t] is not None else np.nan
    with np.errstate(invalid="ignore", divide="ignore"):
        ratio = factor_ret ** 2 / factor_var
    # pandas DataFrame.mean(axis=1) skips NaN; an all-NaN row gives NaN (MFM.py:149)
    cnt = np.sum(~np.isnan(ratio), axis=1)
    with np.errstate(invalid="ignore", divide="ignore"):
        B = np.sqrt(np.where(cnt > 0, np.nansum(ratio, axis=1) / np.maximum(cnt, 1), np.nan))
    weights = 0.5 ** (np.arange(T - 1, -1, -1) / tao)
    lamb, out = [], []
    for t in range(1, T + 1):
        okidx = np.isnan(factor_var[:t]).sum(axis=1) == 0
        ws = weights[:t][okidx]
        with np.errstate(invalid="ignore", divide="ignore"):
            okweights = ws / sum(ws)
            fvm = np.sqrt(sum(okweights * B[:t][okidx] ** 2))
        fvm = float(fvm)
        lamb.append(fvm)
        out.append(eig_cov[t - 1] * fvm ** 2 if eig_cov[t - 1] is not None else None)
    return out, lamb


def qcut_codes(x, ngroup):
    """``pd.qcut(x, ngroup).codes`` restated (utils.py:161): edges are linear
    quantiles, bins right-closed with the lowest edge included, duplicate edges
    raise ValueError."""
    x = np.asarray(x, dtype=np.float64)
    # pandas: Series.quantile(linspace) -> np.percentile(values, qs * 100.0) (method 'linear')
    edges = np.percentile(x, np.linspace(0, 1, ngroup + 1) * 100.0)
    if len(np.unique(edges)) < len(edges):
        raise ValueError("Bin edges must be unique")
    codes = np.searchsorted(edges, x, side="left") - 1
    codes[x == edges[0]] = 0
    return codes, edges


def bayes_shrink(volatility, capital, ngroup=10, q=1):
    """Barra-master/mfm/utils.py:133-168."""
    vol = np.asarray(volatility, dtype=np.float64)
    cap = np.asarray(capital, dtype=np.float64)
    group, _ = qcut_codes(cap, ngroup)
    out = np.empty_like(vol)
    for g in range(ngroup):
        sel = group == g
        if not sel.any():
            continue
        m = sum(vol[sel] * cap[sel]) / sum(cap[sel])                 # utils.py:137
        s = np.sqrt(np.mean((vol[sel] - m) ** 2))                    # :138
        a = q * np.abs(vol[sel] - m)                                 # :146
        with np.errstate(invalid="ignore", divide="ignore"):
            v = a / (a + s)                                          # :148
        out[sel] = v * m + (1 - v) * np.abs(vol[sel])                # :149
    return out


def eigenfactor_bias_stat(cov, ret, predlen=1, eig=canonical_eig):
    """Barra-master/mfm/utils.py:97-117 without the plot."""
    b = []
    for i in range(len(cov) - predlen):
        try:
            D, U = eig(cov[i])
            U = U / np.sum(U, axis=0)
            sigma = np.sqrt(predlen * np.diag(U.T @ cov[i] @ U))
            retlen = (ret[(i + 1):(i + predlen + 1)] + 1).prod(axis=0) - 1
            r = U.T @ retlen
            b.append(r / sigma)
        except Exception:
            pass
    b = np.array(b)
    return np.std(b, axis=0)


# ----------------------------------------------------------------------------
# Part 2 — descriptors (per stock: 1-D arrays over that stock's own rows)
# ----------------------------------------------------------------------------


def _wls_line(y, x, w):
    """statsmodels ``WLS(y, add_constant(x), weights=w).fit()`` (default
    method='pinv'): returns (slope or nan, scale).  ``add_constant`` with
    has_constant='skip' returns X unchanged when x is already a non-zero
    constant column — then there is no 'market_ret'-only design and the slope is
    the single coefficient (factor_calculator.py:96,104)."""
    n = len(y)
    const_x = (np.ptp(x) == 0) and np.all(x != 0)
    X = x[:, None] if const_x else np.column_stack([np.ones(n), x])
    sw = np.sqrt(w)
    wx = X * sw[:, None]
    wy = y * sw
    beta = np.linalg.pinv(wx) @ wy
    rank = np.linalg.matrix_rank(np.diag(np.linalg.svd(wx, compute_uv=False)))
    wres = wy - wx @ beta
    scale = float(wres @ wres) / (n - rank)
    return beta[-1], scale


def beta_hsigma(ret, mkt, window=252, half_life=63, min_periods=42):
    """Barra_factor_cal/factor_calculator.py:79-125 for ONE stock.

    ``ret``/``mkt`` are that stock's rows in date order (NaN allowed).  For each
    row: the last ``window`` rows, rows with a NaN in either column dropped,
    n >= min_periods required; weights ``decay**(n-1..0)`` = the LAST n entries
    of the 252-long table (newest valid row gets 1), un-normalised; BETA is the
    WLS slope and HSIGMA ``sqrt(scale)``."""
    ret = np.asarray(ret, dtype=np.float64)
    mkt = np.asarray(mkt, dtype=np.float64)
    L = len(ret)
    decay = 0.5 ** (1 / half_life)
    weights = decay ** np.arange(window - 1, -1, -1)
    beta = np.full(L, np.nan)
    hs = np.full(L, np.nan)
    ok = ~(np.isnan(ret) | np.isnan(mkt))
    for t in range(L):
        a = max(0, t - window + 1)
        sel = ok[a:t + 1]
        n = int(sel.sum())
        if n < min_periods:
            continue
        y = ret[a:t + 1][sel]
        x = mkt[a:t + 1][sel]
        b, sc = _wls_line(y, x, weights[-n:])
        beta[t] = b
        hs[t] = np.sqrt(sc)
    return beta, hs


def dastd(excess, window=252, half_life=42, min_periods=42):
    """Barra_factor_cal/factor_calculator.py:155-196 for ONE stock:
    ``excess = ret - market_ret``; pandas ``rolling(window, min_periods).apply``
    yields NaN unless the window holds >= min_periods non-NaN values; weights
    are the last n of the table, normalised; weighted std about the weighted
    mean."""
    x = np.asarray(excess, dtype=np.float64)
    L = len(x)
    decay = 0.5 ** (1 / half_life)
    weights = decay ** np.arange(window - 1, -1, -1)
    out = np.full(L, np.nan)
    ok = ~np.isnan(x)
    for t in range(L):
        a = max(0, t - window + 1)
        sel = ok[a:t + 1]
        n = int(sel.sum())
        if n < min_periods:
            continue
        v = x[a:t + 1][sel]
        ws = weights[-n:]
        nw = ws / np.sum(ws)
        mu = np.sum(v * nw)
        out[t] = np.sqrt(np.sum(nw * (v - mu) ** 2))
    return out


def cmra(log_ret, window=252):
    """Barra_factor_cal/factor_calculator.py:199-234 for ONE stock: pandas
    ``rolling(window)`` (min_periods = window) gives NaN unless all ``window``
    values are present; ``ln(1+max Z) - ln(1+min Z)`` with ``Z = exp(cumsum) - 1``."""
    x = np.asarray(log_ret, dtype=np.float64)
    L = len(x)
    out = np.full(L, np.nan)
    for t in range(window - 1, L):
        wv = x[t - window + 1:t + 1]
        if np.isnan(wv).any():
            continue
        z = np.exp(np.cumsum(wv)) - 1
        out[t] = np.log(1 + z.max()) - np.log(1 + z.min())
    return out


def rstr(log_ret, T=504, L=21, half_life=126, min_periods=42):
    """Barra_factor_cal/factor_calculator.py:127-153 for ONE stock (own rows in date order):
    ``x.shift(L).rolling(T - L, min_periods).apply(_rstr_calc)``.  The weight table is
    ``decay**arange(window)`` and the window takes ``weights[:len(window)]`` BY POSITION, so the OLDEST
    row of the window has weight 1 and the newest ``decay**(len-1)`` (SURVEY.md §8(f) [probed]); NaN
    rows are dropped together with their weights (no re-ranking); NaN unless >= min_periods valid."""
    x = np.asarray(log_ret, dtype=np.float64)
    n = len(x)
    window = T - L
    decay = 0.5 ** (1 / half_life)
    weights = decay ** np.arange(0, window)
    y = np.full(n, np.nan)
    if n > L:
        y[L:] = x[:n - L]
    out = np.full(n, np.nan)
    ok = ~np.isnan(y)
    for t in range(n):
        a = max(0, t - window + 1)
        sel = ok[a:t + 1]
        if int(sel.sum()) < min_periods:
            continue
        v = y[a:t + 1][sel]
        w = weights[:t + 1 - a][sel]
        out[t] = np.sum(v * (w / np.sum(w)))
    return out


def liquidity(turnover_rate, windows=((21, 15), (63, 42), (252, 126))):
    """Barra_factor_cal/factor_calculator.py:324-367 for ONE stock: ``dtv = turnover_rate / 100``;
    pandas ``rolling(window, min_periods).sum()`` (NaN skipped, NaN unless >= min_periods valid);
    ``log(sum.replace(0, nan))``.  Returns (STOM, STOQ, STOA)."""
    dtv = np.asarray(turnover_rate, dtype=np.float64) / 100.0
    n = len(dtv)
    ok = ~np.isnan(dtv)
    outs = []
    for (window, minp) in windows:
        o = np.full(n, np.nan)
        for t in range(n):
            a = max(0, t - window + 1)
            sel = ok[a:t + 1]
            if int(sel.sum()) < minp:
                continue
            sm = float(np.sum(dtv[a:t + 1][sel]))
            if sm != 0.0:
                with np.errstate(invalid="ignore"):
                    o[t] = np.log(sm)
        outs.append(o)
    return tuple(outs)


def nlsize(size):
    """Barra_factor_cal/factor_calculator.py:237-293 for ONE date: OLS (statsmodels, pinv) of
    ``SIZE**3`` on ``[1, SIZE]`` over the rows with SIZE present, NLSIZE = -residual; all-NaN when
    fewer than 2 such rows."""
    x = np.asarray(size, dtype=np.float64)
    out = np.full_like(x, np.nan)
    ok = ~np.isnan(x)
    if int(ok.sum()) < 2:
        return out
    xv = x[ok]
    yv = xv ** 3
    const = (np.ptp(xv) == 0) and np.all(xv != 0)          # add_constant(has_constant='skip')
    A = xv[:, None] if const else np.column_stack([np.ones(len(xv)), xv])
    beta = np.linalg.pinv(A) @ yv
    out[ok] = -(yv - A @ beta)
    return out


def size_factor(total_mv):
    """Barra_factor_cal/factor_calculator.py:68-77: ``SIZE = ln(total_mv)``."""
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.log(np.asarray(total_mv, dtype=np.float64))


def bp_factor(pb):
    """Barra_factor_cal/factor_calculator.py:295-321: ``BP = 1 / pb`` where ``pb > 0``, NaN otherwise."""
    pb = np.asarray(pb, dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where(pb > 0, 1.0 / pb, np.nan)


def etop_factor(pe_ttm):
    """Barra_factor_cal/factor_calculator.py:417-419: ``ETOP = 1 / pe_ttm`` where ``pe_ttm > 0``."""
    return bp_factor(pe_ttm)


def rolling_sum(x, window=4, min_periods=4):
    """``Series.rolling(window, min_periods).sum()`` over one stock's report periods
    (factor_calculator.py:394-396): NaN entries are skipped, NaN below ``min_periods`` valid ones."""
    x = np.asarray(x, dtype=np.float64)
    out = np.full_like(x, np.nan)
    for r in range(len(x)):
        w = x[max(0, r - window + 1): r + 1]
        w = w[~np.isnan(w)]
        if len(w) >= min_periods:
            out[r] = w.sum()
    return out


def cetop_factor(cash_flow_ttm, total_mv):
    """factor_calculator.py:409-415: ``ttm / total_mv`` where both are > 0, NaN otherwise."""
    a = np.asarray(cash_flow_ttm, dtype=np.float64)
    b = np.asarray(total_mv, dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where((b > 0) & (a > 0), a / b, np.nan)


def percent_factor(x):
    """factor_calculator.py:454-456: growth rates in percent -> ratios."""
    return np.asarray(x, dtype=np.float64) / 100


def mlev_factor(total_ncl, total_mv):
    """factor_calculator.py:481-484: ``(total_mv + total_ncl) / total_mv`` with +-inf replaced by NaN."""
    a = np.asarray(total_ncl, dtype=np.float64)
    b = np.asarray(total_mv, dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        r = (b + a) / b
    r[np.isinf(r)] = np.nan
    return r


def blev_factor(total_ncl, book_value):
    """factor_calculator.py:488-496: ``(book + total_ncl) / book`` where ``book > 0``, NaN otherwise."""
    a = np.asarray(total_ncl, dtype=np.float64)
    b = np.asarray(book_value, dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where(b > 0, (b + a) / b, np.nan)


def prepare_returns(close):
    """Barra_factor_cal/factor_calculator.py:50-53 for one stock's closes in
    date order: ``pct_change`` (pandas 3: no forward fill) and log-return."""
    close = np.asarray(close, dtype=np.float64)
    ret = np.full_like(close, np.nan)
    lr = np.full_like(close, np.nan)
    ret[1:] = close[1:] / close[:-1] - 1
    with np.errstate(invalid="ignore", divide="ignore"):
        lr[1:] = np.log(close[1:]) - np.log(close[:-1])
    return ret, lr


# ----------------------------------------------------------------------------
# Part 2b — per-date post-processing (1-D arrays over ONE date's rows)
# ----------------------------------------------------------------------------


def winsorize(x, n_std=2.5):
    """Barra_factor_cal/post_processing.py:7-24 for one (date, column):
    clip to mean +- n_std * std with pandas semantics (NaN skipped, ddof 1).
    With fewer than 2 valid values std is NaN and pandas' clip with NaN bounds
    leaves the values unchanged."""
    x = np.asarray(x, dtype=np.float64)
    ok = ~np.isnan(x)
    n = int(ok.sum())
    if n == 0:
        return x.copy()
    mean = x[ok].sum() / n
    if n < 2:
        return x.copy()
    std = np.sqrt(((x[ok] - mean) ** 2).sum() / (n - 1))
    lo, hi = mean - n_std * std, mean + n_std * std
    out = x.copy()
    out[ok] = np.minimum(np.maximum(x[ok], lo), hi)
    return out


def composite(components, weights):
    """Barra_factor_cal/post_processing.py:26-45: NaN-aware weighted mean,
    ``sum w_i x_i(NaN->0) / sum w_i notna(x_i)`` (0/0 -> NaN)."""
    num = np.zeros_like(np.asarray(components[0], dtype=np.float64))
    den = np.zeros_like(num)
    for c, w in zip(components, weights):
        c = np.asarray(c, dtype=np.float64)
        num = num + np.where(np.isnan(c), 0.0, c) * w
        den = den + (~np.isnan(c)) * w
    with np.errstate(invalid="ignore", divide="ignore"):
        return num / den


def orthogonalize(y, X):
    """Barra_factor_cal/post_processing.py:47-70 for ONE date: OLS residual of
    ``y ~ 1 + X`` on complete-case rows (statsmodels OLS, pinv), NaN elsewhere;
    all-NaN when fewer than k+2 complete rows (:56)."""
    y = np.asarray(y, dtype=np.float64)
    X = np.asarray(X, dtype=np.float64)
    if X.ndim == 1:
        X = X[:, None]
    k = X.shape[1]
    ok = ~(np.isnan(y) | np.isnan(X).any(axis=1))
    out = np.full_like(y, np.nan)
    n = int(ok.sum())
    if n < k + 2:
        return out
    Xv = X[ok]
    # add_constant(has_constant='skip'): unchanged if a non-zero constant column exists
    const_cols = (np.ptp(Xv, axis=0) == 0) & np.all(Xv != 0, axis=0)
    A = Xv if const_cols.any() else np.column_stack([np.ones(n), Xv])
    beta = np.linalg.pinv(A) @ y[ok]
    out[ok] = y[ok] - A @ beta
    return out


# ----------------------------------------------------------------------------
# Legacy numpy RNG restated (what the device generator implements)
# ----------------------------------------------------------------------------


class LegacyGauss:
    """``np.random.RandomState(seed).standard_normal`` restated: MT19937
    ``init_genrand`` seeding, 53-bit doubles ``((a>>5)*2^26 + (b>>6)) / 2^53``
    and the Marsaglia polar method that returns ``f*x2`` first and caches
    ``f*x1`` (numpy/random/src/legacy/legacy-distributions.c ``legacy_gauss``,
    numpy/random/src/mt19937/mt19937.c).  Pure Python — small cases only; used
    to pin the device generator's algorithm against numpy itself."""

    def __init__(self, seed):
        mt = [0] * 624
        mt[0] = seed & 0xFFFFFFFF
        for i in range(1, 624):
            mt[i] = (1812433253 * (mt[i - 1] ^ (mt[i - 1] >> 30)) + i) & 0xFFFFFFFF
        self.mt = mt
        self.pos = 624
        self.has_gauss = False
        self.gauss = 0.0

    def _gen(self):
        mt = self.mt
        for kk in range(624):
            y = (mt[kk] & 0x80000000) | (mt[(kk + 1) % 624] & 0x7FFFFFFF)
            mt[kk] = mt[(kk + 397) % 624] ^ (y >> 1) ^ (0x9908B0DF if (y & 1) else 0)
        self.pos = 0

    def u32(self):
        if self.pos == 624:
            self._gen()
        y = self.mt[self.pos]
        self.pos += 1
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        y ^= y >> 18
        return y & 0xFFFFFFFF

    def double(self):
        a = self.u32() >> 5
        b = self.u32() >> 6
        return (a * 67108864.0 + b) / 9007199254740992.0

    def normal(self):
        if self.has_gauss:
            self.has_gauss = False
            return self.gauss
        while True:
            x1 = 2.0 * self.double() - 1.0
            x2 = 2.0 * self.double() - 1.0
            r2 = x1 * x1 + x2 * x2
            if r2 < 1.0 and r2 != 0.0:
                break
        f = float(np.sqrt(-2.0 * np.log(r2) / r2))
        self.gauss = f * x1
        self.has_gauss = True
        return f * x2

    def standard_normal(self, shape):
        n = int(np.prod(shape))
        return np.array([self.normal() for _ in range(n)]).reshape(shape)
