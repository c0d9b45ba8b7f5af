# -*- coding: utf-8 -*-
"""Drop-in for Barra-master/mfm/MFM.py: same class, methods, attributes and
return types (MFM.py:17-167); every date loop runs as batched CUDA kernels."""
import numpy as np
import pandas as pd

from bmfm import _cabi
from bmfm import ops as _ops
from bmfm.panel import pack_frame_device, LazyFrameList
from mfm.utils import progressbar


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
        逐时间点进行横截面多因子回归 (MFM.py:48-76) — the table is packed on the GPU, one kernel launch for all dates
        '''
        print('===================================逐时间点进行横截面多因子回归===================================')
        ctx = _cabi.default_context()
        # a P that does not match the table's columns fails the way MFM.py:70 does (ValueError from the constructor)
        pd.DataFrame(np.empty((0, 1 + self.P + self.Q)), columns=self.columns)
        # the long table is packed on the GPU and never comes back: the regression reads it where it lies
        pk = pack_frame_device(self.data, self.P, self.Q, ctx, dates=self.dates)
        fr, spec, r2, status, _ = _ops.reg_by_time_device(
            ctx, pk.T, pk.P, pk.Q, pk.max_rows, pk.date_off, pk.cap, pk.ret, pk.ind, pk.style)
        status = status.cpu().numpy()
        if (status == _cabi.ST_NONFINITE).any():
            raise np.linalg.LinAlgError("SVD did not converge")
        fr = fr.cpu().numpy()
        r2 = r2.cpu().numpy()

        # DataFrame(factor_ret, columns=self.columns) raises ValueError on a P mismatch, like MFM.py:70
        factor_ret = pd.DataFrame(fr, columns=self.columns, index=self.sorted_dates)
        R2 = pd.DataFrame(r2, columns=['R2'], index=self.sorted_dates)
        sorted_dates = self.sorted_dates

        def fetch():
            # first access to a specific-return frame: one gather on the device, one copy, names by code
            vals = spec[pk.name_pos].cpu().numpy()
            names = pk.name_uniques[pk.name_code[pk.ref_order].cpu().numpy()]
            return vals, names, pk.date_off_host, sorted_dates

        specific_ret = LazyFrameList(self.T, fetch)          # MFM.py:66, built on demand
        self.last_capital = pk.last_capital
        self._status = status
        self.factor_ret = factor_ret
        self.specific_ret = specific_ret
        self.R2 = R2
        return ((factor_ret, specific_ret, R2))

    # ---- helpers: device arrays <-> the reference's lists of DataFrames -------------------
    def _cov_list(self, arr, status):
        """[T,K,K] array + status -> list of K x K DataFrames, empty DataFrame where the
        reference swallows an exception (MFM.py:95-96, 120-121)."""
        names = self.factor_ret.columns
        out = []
        for t in range(self.T):
            if status[t] in (_cabi.ST_OK, _cabi.ST_PINV):
                out.append(pd.DataFrame(arr[t], columns=names, index=names))
            else:
                out.append(pd.DataFrame())
        return out

    def Newey_West_by_time(self, q=2, tao=252):
        '''
        逐时间点计算协方差并进行Newey West调整 (MFM.py:80-101)
        '''
        if self.factor_ret is None:
            raise Exception('please run reg_by_time to get factor returns first')
        print('\n\n===================================逐时间点进行Newey West调整=================================')
        import torch
        ctx = _cabi.default_context()
        fr = torch.from_numpy(np.ascontiguousarray(self.factor_ret.values, dtype=np.float64)).cuda()
        cov, status = _ops.newey_west_by_time_device(ctx, fr, q, tao)
        self._d_fr, self._d_nw, self._d_nw_status = fr, cov, status
        self._nw_status = status.cpu().numpy()
        Newey_West_cov = self._cov_list(cov.cpu().numpy(), self._nw_status)
        progressbar(self.T, self.T, '   date: ' + str(self.sorted_dates[self.T - 1])[:10])
        self.Newey_West_cov = Newey_West_cov
        return (Newey_West_cov)

    def eigen_risk_adj_by_time(self, M=100, scale_coef=1.4):
        '''
        逐时间点进行Eigenfactor Risk Adjustment (MFM.py:105-126); simulation length = self.T (MFM.py:119)
        '''
        if self.Newey_West_cov is None:
            raise Exception('please run Newey_West_by_time to get factor return covariances after Newey West adjustment first')
        print('\n\n===================================逐时间点进行Eigenfactor Risk调整=================================')
        ctx = _cabi.default_context()
        out, status = _ops.eigen_risk_adj_device(ctx, self._d_nw, self._d_nw_status, self.T, M, scale_coef)
        self._d_er, self._d_er_status = out, status
        self._er_status = status.cpu().numpy()
        eigen_risk_adj_cov = self._cov_list(out.cpu().numpy(), self._er_status)
        progressbar(self.T, self.T, '   date: ' + str(self.sorted_dates[self.T - 1])[:10])
        self.eigen_risk_adj_cov = eigen_risk_adj_cov
        return (eigen_risk_adj_cov)

    def vol_regime_adj_by_time(self, tao=84):
        '''
        Volatility Regime Adjustment (MFM.py:130-167)
        '''
        if self.eigen_risk_adj_cov is None:
            raise Exception('please run eigen_risk_adj_by_time to get factor return covariances after eigenfactor risk adjustment first')
        print('\n\n==================================逐时间点进行Volatility Regime调整================================')
        ctx = _cabi.default_context()
        out, lamb = _ops.vol_regime_device(ctx, self._d_fr, self._d_er, self._d_er_status, tao)
        vol_regime_adj_cov = self._cov_list(out.cpu().numpy(), self._er_status)
        lamb = [float(x) for x in lamb.cpu().numpy()]
        progressbar(self.T, self.T, '   date: ' + str(self.sorted_dates[self.T - 1])[:10])
        self.vol_regime_adj_cov = vol_regime_adj_cov
        return ((vol_regime_adj_cov, lamb))
