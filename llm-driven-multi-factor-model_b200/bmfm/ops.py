"""Array-level entry points over the C-ABI (device memory owned by torch).

Everything here runs on the GPU through libbmfm.so; numpy arrays in, numpy
arrays out.  The ``mfm`` drop-in package, the tests and ``bench.py`` call these.
"""
from __future__ import annotations

import numpy as np

from . import _cabi
from .panel import pack_arrays


def _torch():
    import torch
    return torch


def _dev(a, dtype=None):
    torch = _torch()
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda(non_blocking=False)


class LinAlgError(np.linalg.LinAlgError):
    pass


def reg_by_time_device(ctx, T, P, Q, max_rows, d_off, d_cap, d_ret, d_ind, d_style, want_exposure=False):
    """Device tensors in, device tensors out (factor_ret [T,K], spec [rows], r2 [T], status [T], exposure|None)."""
    torch = _torch()
    K = 1 + P + Q
    dev = d_cap.device
    fr = torch.empty((T, K), dtype=torch.float64, device=dev)
    spec = torch.empty(d_cap.shape[0], dtype=torch.float64, device=dev)
    r2 = torch.empty(T, dtype=torch.float64, device=dev)
    status = torch.empty(T, dtype=torch.int32, device=dev)
    expo = torch.empty((T, K, K), dtype=torch.float64, device=dev) if want_exposure else None
    ctx.reg_by_time(T, P, Q, max_rows, d_off, d_cap, d_ret, d_ind, d_style, fr, spec, r2, status, expo)
    return fr, spec, r2, status, expo


def reg_by_time(date_off, cap, ret, ind, style, P, want_exposure=False, group_by_industry=True, ctx=None):
    """MFM.reg_by_time on date-major arrays (Barra-master/mfm/MFM.py:48-76).

    Inputs are in stock order inside each date; they are stored grouped by
    industry on the device and the specific returns are mapped back.
    Returns (factor_ret [T,K], specific_ret [rows], R2 [T], status [T], exposure or None).
    Raises numpy.linalg.LinAlgError where the reference's pinv would (non-finite Gram)."""
    ctx = ctx or _cabi.default_context()
    date_off = np.asarray(date_off, dtype=np.int64)
    T = len(date_off) - 1
    style = np.asarray(style, dtype=np.float64)
    Q = style.shape[1] if style.ndim == 2 else 0
    (c, r, i, s), pos = pack_arrays(date_off, cap, ret, ind, style.reshape(len(cap), Q), P, group_by_industry)
    max_rows = int(np.diff(date_off).max()) if T else 0
    fr, spec, r2, status, expo = reg_by_time_device(
        ctx, T, P, Q, max_rows, _dev(date_off), _dev(c), _dev(r), _dev(i), _dev(s), want_exposure)
    status = status.cpu().numpy()
    if (status == _cabi.ST_NONFINITE).any():
        raise LinAlgError("SVD did not converge")     # what numpy's pinv raises on a non-finite Gram
    spec = spec.cpu().numpy()[pos]
    return (fr.cpu().numpy(), spec, r2.cpu().numpy(), status,
            expo.cpu().numpy() if expo is not None else None)


# ---------------------------------------------------------------------------- part 3
_VALID = (_cabi.ST_OK, _cabi.ST_PINV)


def newey_west_by_time_device(ctx, d_fr, q, tao):
    """d_fr [T,K] device -> (cov [T,K,K] device (NaN where invalid), status [T] device)."""
    torch = _torch()
    T, K = d_fr.shape
    cov = torch.empty((T, K, K), dtype=torch.float64, device=d_fr.device)
    status = torch.empty(T, dtype=torch.int32, device=d_fr.device)
    ctx.newey_west_by_time(T, K, int(q), tao, d_fr, cov, status)
    return cov, status


def newey_west_by_time(factor_ret, q=2, tao=252, ctx=None):
    """MFM.Newey_West_by_time on a [T,K] array -> (cov [T,K,K], status [T])."""
    ctx = ctx or _cabi.default_context()
    cov, status = newey_west_by_time_device(ctx, _dev(np.asarray(factor_ret, dtype=np.float64)), q, tao)
    return cov.cpu().numpy(), status.cpu().numpy()


def vol_regime_device(ctx, d_fr, d_cov, d_status, tao):
    torch = _torch()
    T, K = d_fr.shape
    out = torch.empty_like(d_cov)
    lamb = torch.empty(T, dtype=torch.float64, device=d_fr.device)
    ctx.vol_regime_adj_by_time(T, K, tao, d_fr, d_cov, d_status, out, lamb)
    return out, lamb


def vol_regime_adj_by_time(factor_ret, cov, cov_status=None, tao=84, ctx=None):
    ctx = ctx or _cabi.default_context()
    d_st = _dev(np.asarray(cov_status, dtype=np.int32)) if cov_status is not None else None
    out, lamb = vol_regime_device(ctx, _dev(np.asarray(factor_ret, dtype=np.float64)),
                                  _dev(np.asarray(cov, dtype=np.float64)), d_st, tao)
    return out.cpu().numpy(), lamb.cpu().numpy()


_gauss_cache = {}


def gauss_cov_device(ctx, m0, nsim, T_sim, K):
    """cov(z_m) for simulations m0 .. m0+nsim-1 (seeds m+1), cached per device:
    the reference re-seeds with m+1 for EVERY date (utils.py:73), so these K x K
    sample covariances are shared by all dates."""
    torch = _torch()
    key = (ctx.device, m0, nsim, T_sim, K)
    if key not in _gauss_cache:
        if len(_gauss_cache) > 4:
            _gauss_cache.clear()
        Cm = torch.empty((nsim, K, K), dtype=torch.float64, device=f"cuda:{ctx.device}")
        ctx.gauss_cov(m0, nsim, T_sim, K, Cm)
        _gauss_cache[key] = Cm
    return _gauss_cache[key]


def eigen_risk_adj_device(ctx, d_cov, d_cov_status, T_sim, M, scale_coef, sim_range=None, allreduce=None):
    """Eigenfactor adjustment of every date's covariance.

    ``sim_range=(m_lo, m_hi)`` restricts this process to a shard of the M
    simulations; ``allreduce(tensor)`` (e.g. an NCCL all-reduce) then sums the
    per-shard accumulators before the finalisation.  Returns (out [T,K,K], status [T])."""
    torch = _torch()
    T, K = d_cov.shape[0], d_cov.shape[1]
    dev = d_cov.device
    D0 = torch.empty((T, K), dtype=torch.float64, device=dev)
    U0 = torch.empty((T, K, K), dtype=torch.float64, device=dev)
    status = torch.empty(T, dtype=torch.int32, device=dev)
    ctx.eigen_decompose(T, K, d_cov, d_cov_status, D0, U0, status)
    m_lo, m_hi = (0, M) if sim_range is None else sim_range
    vsum = torch.zeros((T, K), dtype=torch.float64, device=dev)
    if m_hi > m_lo:
        Cm = gauss_cov_device(ctx, m_lo, m_hi - m_lo, T_sim, K)
        ctx.eigen_mc_accumulate(T, K, m_hi - m_lo, Cm, D0, status, vsum, False)
    if allreduce is not None:
        allreduce(vsum)
    out = torch.empty((T, K, K), dtype=torch.float64, device=dev)
    ctx.eigen_finalize(T, K, M, scale_coef, D0, U0, vsum, status, out)
    return out, status


def eigen_risk_adj_by_time(cov, cov_status, T_sim, M=100, scale_coef=1.4, ctx=None):
    ctx = ctx or _cabi.default_context()
    d_st = _dev(np.asarray(cov_status, dtype=np.int32)) if cov_status is not None else None
    out, status = eigen_risk_adj_device(ctx, _dev(np.asarray(cov, dtype=np.float64)), d_st, T_sim, M, scale_coef)
    return out.cpu().numpy(), status.cpu().numpy()


def literal_eig_order(covmat, T_sim, M):
    """What only LAPACK can tell: the order in which ``np.linalg.eig`` (dgeev) returns the eigenpairs of F0
    and of every simulated covariance Fm of ``utils.eigen_risk_adj`` (utils.py:64,73-79), for the
    injected-literal mode.  Follows the reference's construction of Fm step by step so that dgeev sees the
    very same matrix.  Returns (D0 [K], U0 [K,K] in numpy's order, perm int32 [M,K] with perm[m][r] = the
    position of the eigenpair of descending rank r).  Host numpy: one K x K ``eig`` per simulation."""
    F0 = np.asarray(covmat, dtype=np.float64)
    K = F0.shape[0]
    D0, U0 = np.linalg.eig(F0)
    perm = np.empty((M, K), dtype=np.int32)
    for m in range(M):
        rs = np.random.RandomState(m + 1)
        bm = rs.multivariate_normal(mean=K * [0], cov=np.diag(D0), size=T_sim).T
        Dm, _ = np.linalg.eig(np.cov(U0 @ bm))
        perm[m] = np.argsort(-np.real(Dm), kind="stable")
    return np.real(D0), np.real(U0), perm


def eigen_risk_adj_literal(cov, T_sim, M, scale_coef, D0, U0, perm, ctx=None):
    """Eigenfactor adjustment with the eigen ORDER injected (SURVEY.md 8(c) mode (iii)): cov [T,K,K],
    D0 [T,K] / U0 [T,K,K] as ``np.linalg.eig`` returned them, perm [T,M,K] from ``literal_eig_order``.
    Reproduces the unpatched reference (utils.py:55-92).  Returns out [T,K,K]."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    D0 = np.asarray(D0, dtype=np.float64)
    T, K = D0.shape
    order = np.argsort(-D0, axis=1, kind="stable")
    d_desc = _dev(np.take_along_axis(D0, order, axis=1))
    status = torch.zeros(T, dtype=torch.int32, device=d_desc.device)
    vsum = torch.zeros((T, K), dtype=torch.float64, device=d_desc.device)
    Cm = gauss_cov_device(ctx, 0, M, T_sim, K)
    ctx.eigen_mc_accumulate_perm(T, K, M, Cm, d_desc, status, _dev(np.asarray(perm, dtype=np.int32)), vsum, False)
    out = torch.empty((T, K, K), dtype=torch.float64, device=d_desc.device)
    ctx.eigen_finalize(T, K, M, scale_coef, _dev(D0), _dev(np.asarray(U0, dtype=np.float64)), vsum, status, out)
    return out.cpu().numpy()


def bayes_shrink_by_time(date_off, volatility, capital, ngroup=10, q=1, ctx=None):
    """utils.bayes_shrink on every date of a date-major table -> (out [rows], status [T])."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    date_off = np.asarray(date_off, dtype=np.int64)
    T = len(date_off) - 1
    vol = _dev(np.asarray(volatility, dtype=np.float64))
    cap = _dev(np.asarray(capital, dtype=np.float64))
    out = torch.empty_like(vol)
    status = torch.empty(T, dtype=torch.int32, device=vol.device)
    max_rows = int(np.diff(date_off).max()) if T else 0
    ctx.bayes_shrink(T, max_rows, _dev(date_off), vol, cap, ngroup, q, out, status)
    return out.cpu().numpy(), status.cpu().numpy()


# ---------------------------------------------------------------------------- part 2
def prepare_returns(close, ctx=None):
    """close [L, S] (row-index-major per stock, NaN = missing) -> (ret, log_ret)."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    c = _dev(np.asarray(close, dtype=np.float64))
    L, S = c.shape
    ret = torch.empty_like(c)
    lr = torch.empty_like(c)
    ctx.prepare_returns(L, S, c, ret, lr)
    return ret.cpu().numpy(), lr.cpu().numpy()


def beta_hsigma(ret, mkt, window=252, half_life=63, min_periods=42, ctx=None):
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    r = _dev(np.asarray(ret, dtype=np.float64))
    m = _dev(np.ascontiguousarray(np.broadcast_to(np.asarray(mkt, dtype=np.float64), r.shape)))
    L, S = r.shape
    b = torch.empty_like(r)
    h = torch.empty_like(r)
    ctx.beta_hsigma(L, S, r, m, window, half_life, min_periods, b, h)
    return b.cpu().numpy(), h.cpu().numpy()


def dastd(ret, mkt, window=252, half_life=42, min_periods=42, ctx=None):
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    r = _dev(np.asarray(ret, dtype=np.float64))
    m = _dev(np.ascontiguousarray(np.broadcast_to(np.asarray(mkt, dtype=np.float64), r.shape)))
    L, S = r.shape
    o = torch.empty_like(r)
    ctx.dastd(L, S, r, m, window, half_life, min_periods, o)
    return o.cpu().numpy()


def cmra(log_ret, window=252, ctx=None):
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    r = _dev(np.asarray(log_ret, dtype=np.float64))
    L, S = r.shape
    o = torch.empty_like(r)
    ctx.cmra(L, S, r, window, o)
    return o.cpu().numpy()


def eigenfactor_bias_stat(cov, cov_status, factor_ret, predlen=1, ctx=None):
    """cov [T, K, K] (+ status [T] or None), factor_ret [T, K] -> bias_stat [K] (canonical eigen order)."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    c = _dev(np.asarray(cov, dtype=np.float64))
    f = _dev(np.asarray(factor_ret, dtype=np.float64))
    st = None if cov_status is None else _dev(np.asarray(cov_status, dtype=np.int32))
    T, K = f.shape
    out = torch.empty(K, dtype=torch.float64, device=c.device)
    ctx.eigenfactor_bias_stat(T, K, predlen, c, st, f, out)
    return out.cpu().numpy()


def rstr(log_ret, T=504, lag=21, half_life=126, min_periods=42, ctx=None):
    """compute_rstr on a [row-of-stock, stock] panel of log returns."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    r = _dev(np.asarray(log_ret, dtype=np.float64))
    L, S = r.shape
    o = torch.empty_like(r)
    ctx.rstr(L, S, r, T, lag, half_life, min_periods, o)
    return o.cpu().numpy()


def liquidity(turnover_rate, windows=((21, 15), (63, 42), (252, 126)), divisor=100.0, ctx=None):
    """compute_liquidity on a [row-of-stock, stock] panel of turnover in percent: (STOM, STOQ, STOA)."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    x = _dev(np.asarray(turnover_rate, dtype=np.float64))
    L, S = x.shape
    outs = [torch.empty_like(x) for _ in windows]
    if len(windows) == 3:
        ctx.rolling_log_sum3(L, S, x, divisor, [w for (w, _) in windows], [m for (_, m) in windows], outs)
    else:
        for (window, minp), o in zip(windows, outs):
            ctx.rolling_log_sum(L, S, x, divisor, window, minp, o)
    return tuple(o.cpu().numpy() for o in outs)


def size_factor(total_mv, ctx=None):
    """compute_size: ln(total_mv)."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    x = _dev(np.asarray(total_mv, dtype=np.float64))
    o = torch.empty_like(x)
    ctx.elementwise(_cabi.OP_LOG, x.numel(), x, o)
    return o.cpu().numpy()


def bp_factor(pb, ctx=None):
    """compute_bp: 1 / pb where pb > 0, NaN otherwise."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    x = _dev(np.asarray(pb, dtype=np.float64))
    o = torch.empty_like(x)
    ctx.elementwise(_cabi.OP_INV_POS, x.numel(), x, o)
    return o.cpu().numpy()


def _unary(op, x, ctx=None):
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    xd = _dev(np.asarray(x, dtype=np.float64))
    o = torch.empty_like(xd)
    ctx.elementwise(op, xd.numel(), xd, o)
    return o.cpu().numpy()


def _binary(op, x, y, ctx=None):
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    xd = _dev(np.asarray(x, dtype=np.float64))
    yd = _dev(np.asarray(y, dtype=np.float64))
    o = torch.empty_like(xd)
    ctx.elementwise2(op, xd.numel(), xd, yd, o)
    return o.cpu().numpy()


def etop_factor(pe_ttm, ctx=None):
    """compute_earnings_yield: ETOP = 1 / pe_ttm where pe_ttm > 0 (factor_calculator.py:417-419)."""
    return _unary(_cabi.OP_INV_POS, pe_ttm, ctx)


def cetop_factor(cash_flow_ttm, total_mv, ctx=None):
    """compute_earnings_yield: CETOP = ttm / total_mv where both > 0 (factor_calculator.py:409-415)."""
    return _binary(_cabi.OP2_RATIO_POS, cash_flow_ttm, total_mv, ctx)


def percent_factor(x, ctx=None):
    """select_growth_factors: percent -> ratio (factor_calculator.py:454-456)."""
    return _unary(_cabi.OP_PERCENT, x, ctx)


def mlev_factor(total_ncl, total_mv, ctx=None):
    """compute_leverage: MLEV = (total_mv + total_ncl) / total_mv, inf -> NaN (factor_calculator.py:481-484)."""
    return _binary(_cabi.OP2_LEV_FINITE, total_ncl, total_mv, ctx)


def blev_factor(total_ncl, book_value, ctx=None):
    """compute_leverage: BLEV = (book + total_ncl) / book where book > 0 (factor_calculator.py:488-496)."""
    return _binary(_cabi.OP2_LEV_POS, total_ncl, book_value, ctx)


def rolling_sum(x, window, min_periods, ctx=None):
    """pandas rolling(window, min_periods).sum() down the rows of a [row-of-stock, stock] panel."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    xd = _dev(np.asarray(x, dtype=np.float64))
    L, S = xd.shape
    o = torch.empty_like(xd)
    ctx.rolling_sum(L, S, xd, window, min_periods, o)
    return o.cpu().numpy()


def nlsize(date_off, size, ctx=None):
    """compute_nlsize on the long table grouped by date."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    date_off = np.asarray(date_off, dtype=np.int64)
    xd = _dev(np.asarray(size, dtype=np.float64))
    o = torch.empty_like(xd)
    ctx.nlsize(len(date_off) - 1, _dev(date_off), xd, o)
    return o.cpu().numpy()


def winsorize(date_off, x, n_std=2.5, ctx=None):
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    date_off = np.asarray(date_off, dtype=np.int64)
    xd = _dev(np.asarray(x, dtype=np.float64))
    out = torch.empty_like(xd)
    ctx.winsorize(len(date_off) - 1, _dev(date_off), xd, n_std, out)
    return out.cpu().numpy()


def composite(components, weights, ctx=None):
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    comps = [_dev(np.asarray(c, dtype=np.float64)) for c in components]
    out = torch.empty_like(comps[0])
    ctx.composite(comps[0].numel(), comps, weights, out)
    return out.cpu().numpy()


def orthogonalize(date_off, y, X, ctx=None):
    """X: list of columns (or [rows, k] array)."""
    torch = _torch()
    ctx = ctx or _cabi.default_context()
    date_off = np.asarray(date_off, dtype=np.int64)
    X = np.asarray(X, dtype=np.float64)
    cols = [X] if X.ndim == 1 else [np.ascontiguousarray(X[:, j]) for j in range(X.shape[1])]
    yd = _dev(np.asarray(y, dtype=np.float64))
    Xd = [_dev(c) for c in cols]
    out = torch.empty_like(yd)
    ctx.orthogonalize(len(date_off) - 1, _dev(date_off), yd, Xd, out)
    return out.cpu().numpy()
