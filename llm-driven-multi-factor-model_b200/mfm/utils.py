# -*- coding: utf-8 -*-
"""Drop-in for Barra-master/mfm/utils.py: same function names, arguments and
return types; Newey-West and the eigenfactor adjustment run on the GPU.

Eigen-order note: the reference calls the GENERAL eigensolver
``np.linalg.eig`` and averages simulated biases by output POSITION, so its
literal output depends on LAPACK's deflation order.  This implementation uses
the canonical convention (eigenvalues descending) and matches the reference
source run with ``np.linalg.eig`` patched to that convention to <= 1e-9.
"""
import math

import numpy as np
import pandas as pd

from bmfm import ops as _ops
from bmfm import _cabi


def Newey_West(ret, q=2, tao=252):
    '''
    Newey_West方差调整 (utils.py:16-50): ret DataFrame (rows time, cols factors) -> K x K DataFrame
    '''
    T = ret.shape[0]
    K = ret.shape[1]
    if T <= q or T <= K:
        raise Exception("T <= q or T <= K")
    names = ret.columns
    cov, status = _ops.newey_west_by_time(np.ascontiguousarray(ret.values, dtype=np.float64), q, tao)
    return (pd.DataFrame(cov[-1], columns=names, index=names))


def eigen_risk_adj(covmat, T=1000, M=100, scale_coef=1.4):
    '''
    Eigenfactor Risk Adjustment (utils.py:55-92)
    T: 序列长度   M: 模拟次数   scale_coef: scale coefficient for bias
    '''
    vals = np.ascontiguousarray(np.asarray(covmat, dtype=np.float64))
    out, status = _ops.eigen_risk_adj_by_time(vals[None], None, int(T), int(M), scale_coef)
    if status[0] == _cabi.ST_NOT_PSD:
        # the reference does raise('...') of a str, i.e. a TypeError (utils.py:67-68)
        raise TypeError('exceptions must derive from BaseException')
    if status[0] != _cabi.ST_OK:
        raise np.linalg.LinAlgError("Array must not contain infs or NaNs")
    cols = covmat.columns if hasattr(covmat, "columns") else None
    return (pd.DataFrame(out[0], columns=cols, index=cols))


def eigenfactor_bias_stat(cov, ret, predlen=1):
    '''
    计算特征因子组合的bias统计量 (utils.py:97-117) on the GPU: `cov` list of K x K DataFrames (empty frames where a
    date has none: skipped, as the reference's try/except does), `ret` DataFrame T x K.  Canonical eigen order.
    '''
    vals = ret.values if hasattr(ret, "values") else np.asarray(ret)
    vals = np.ascontiguousarray(vals, dtype=np.float64)
    T, K = vals.shape
    arr = np.full((T, K, K), np.nan)
    status = np.full(T, _cabi.ST_SHORT, dtype=np.int32)
    for i in range(min(len(cov), T)):
        c = np.asarray(cov[i], dtype=np.float64)
        if c.shape == (K, K):
            arr[i] = c
            status[i] = _cabi.ST_OK
    # the reference walks range(len(cov) - predlen): dates beyond len(cov) carry no covariance
    bias_stat = _ops.eigenfactor_bias_stat(arr[:len(cov)], status[:len(cov)], vals[:len(cov)] if len(cov) <= T else vals,
                                           predlen)
    try:
        import matplotlib.pyplot as plt
        plt.plot(bias_stat)
    except Exception:
        pass
    return (bias_stat)


def progressbar(cur, total, txt):
    '''
    显示进度条 (utils.py:123-128)
    '''
    percent = '{:.2%}'.format(cur / total)
    print("\r[%-50s] %s" % ('=' * int(math.floor(cur * 50 / total)), percent) + txt, end='')


def group_mean_std(x):
    '''
    计算一组的加权平均和波动率 (utils.py:133-139): cap-weighted mean and population std of one cap group.
    Host helper kept so that code importing it from mfm.utils keeps working; bayes_shrink below does not
    call it (the whole shrinkage runs in one kernel).  `x`: anything with .volatility and .capital.
    '''
    vol = np.asarray(x.volatility, dtype=np.float64)
    cap = np.asarray(x.capital, dtype=np.float64)
    m = float(np.dot(vol, cap) / cap.sum())
    return ([m, float(np.sqrt(np.mean(np.square(vol - m))))])


def shrink(x, group_weight_mean, q):
    '''
    计算shrink估计量 (utils.py:142-150) for one row `x` (keys 'volatility', 'group'); `group_weight_mean[g]` is
    the [mean, std] pair of group g as group_mean_std returns it.  Host helper for API compatibility only.
    '''
    m, s = group_weight_mean[x['group']][0], group_weight_mean[x['group']][1]
    dist = q * abs(x['volatility'] - m)
    v = dist / (dist + s)
    return (v * m + (1 - v) * abs(x['volatility']))


def bayes_shrink(volatility, capital, ngroup=10, q=1):
    '''
    使用市值对特异性收益率波动率进行贝叶斯收缩 (utils.py:153-168); one cross-section on the GPU
    '''
    vol = np.asarray(volatility, dtype=np.float64)
    cap = np.asarray(capital, dtype=np.float64)
    out, status = _ops.bayes_shrink_by_time(np.array([0, len(vol)], dtype=np.int64), vol, cap, ngroup, q)
    if status[0] == _cabi.ST_DUP_EDGES:
        raise ValueError("Bin edges must be unique: duplicate quantile edges of capital "
                         "(pd.qcut raises the same, utils.py:161)")
    return (out)
