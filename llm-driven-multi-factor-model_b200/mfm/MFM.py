# -*- coding: utf-8 -*-
"""Drop-in for Barra-master/mfm/MFM.py: same class, methods, attributes and
return types (MFM.py:17-167); every date loop runs as batched CUDA kernels."""
import numpy as np
import pandas as pd

from bmfm import _cabi
from bmfm import ops as _ops
from bmfm.panel import pack_frame


class MFM():
    '''
    data: DataFrame
    column1: date
    colunm2: stocknames
    colunm3: capital
    column4: ret
    style_factors: DataFrame
    industry_factors: DataFrame
    '''

    def __init__(self, data, P, Q):
        self.Q = Q
        self.P = P
        self.dates = pd.to_datetime(data.date.values)
        self.sorted_dates = pd.to_datetime(np.sort(pd.unique(self.dates)))
        self.T = len(self.sorted_dates)
        self.data = data
        self.columns = ['country']
        self.columns.extend((list(data.columns[4:])))

        self.last_capital = None
        self.factor_ret = None
        self.specific_ret = None
        self.R2 = None

        self.Newey_West_cov = None
        self.eigen_risk_adj_cov = None
        self.vol_regime_adj_cov = None

    def reg_by_time(self):
        '''
        逐时间点进行横截面多因子回归 (MFM.py:48-76) — one kernel launch for all dates
        '''
        import torch
        print('===================================逐时间点进行横截面多因子回归===================================')
        pk = pack_frame(self.data, self.P, self.Q)
        ctx = _cabi.default_context()
        dv = lambda a: torch.from_numpy(a).cuda()
        fr, spec, r2, status, _ = _ops.reg_by_time_device(
            ctx, pk.T, pk.P, pk.Q, pk.max_rows, dv(pk.date_off), dv(pk.cap), dv(pk.ret), dv(pk.ind), dv(pk.style))
        status = status.cpu().numpy()
        if (status == _cabi.ST_NONFINITE).any():
            raise np.linalg.LinAlgError("SVD did not converge")
        fr = fr.cpu().numpy()
        spec = spec.cpu().numpy()[pk.name_pos]
        r2 = r2.cpu().numpy()

        # DataFrame(factor_ret, columns=self.columns) raises ValueError on a P mismatch, like MFM.py:70
        factor_ret = pd.DataFrame(fr, columns=self.columns, index=self.sorted_dates)
        R2 = pd.DataFrame(r2, columns=['R2'], index=self.sorted_dates)
        specific_ret = []
        off = pk.date_off
        for t in range(self.T):
            a, b = off[t], off[t + 1]
            specific_ret.append(pd.DataFrame([spec[a:b]], columns=list(pk.names_sorted[a:b]),
                                             index=[self.sorted_dates[t]]))
        self.last_capital = pk.last_capital
        self._status = status
        self.factor_ret = factor_ret
        self.specific_ret = specific_ret
        self.R2 = R2
        return ((factor_ret, specific_ret, R2))
