# -*- coding: utf-8 -*-
"""Drop-in for Barra-master/mfm/CrossSection.py (one date's regression).

Same constructor, attributes and ``reg()`` return values as the reference
(CrossSection.py:35-108); the arithmetic runs in the CUDA regression kernel.
"""
import numpy as np
import pandas as pd

from bmfm import ops as _ops
from bmfm.panel import onehot_to_index


def style_factor_norm(factors, capital):
    """Cap-weighted mean / one scalar std standardisation (CrossSection.py:12-20).
    Kept for API parity; ``reg`` does this inside the kernel."""
    factors = np.asarray(factors, dtype=np.float64)
    capital = np.asarray(capital, dtype=np.float64)
    w_mu = np.dot(factors.T, capital) / capital.sum(0)
    return (factors - w_mu) / np.std(factors)


class CrossSection():
    '''
    base_data: DataFrame
    column1: date
    colunm2: stocknames
    colunm3: capital
    column4: ret
    style_factors: DataFrame
    industry_factors: DataFrame
    '''

    def __init__(self, base_data, style_factors=pd.DataFrame(), industry_factors=pd.DataFrame()):
        self.date = list(base_data.date)[0]
        self.stocknames = list(base_data.stocknames)
        self.capital = base_data.capital.values
        self.ret = base_data.ret.values
        self.style_factors_names = list(style_factors.columns)
        self.industry_factors_names = list(industry_factors.columns)

        self.N = base_data.shape[0]
        self.Q = style_factors.shape[1]
        self.P = industry_factors.shape[1]
        self._style_raw = np.ascontiguousarray(style_factors.values, dtype=np.float64).reshape(self.N, self.Q)
        self.style_factors = style_factor_norm(self._style_raw, self.capital) if self.Q > 0 else self._style_raw
        self.industry_factors = industry_factors.values
        self.country_factors = np.array(self.N * [[1]])

        self.W = np.sqrt(self.capital) / sum(np.sqrt(self.capital))

        print('\rCross Section Regression, ' + 'Date: ' + self.date + ', ' +
              str(self.N) + ' Stocks, ' + str(self.P) + ' Industry Facotrs, ' + str(self.Q) + ' Style Facotrs', end='')

    def reg(self):
        '''
        多因子模型求解 — returns (factor_ret, specific_ret, pure_factor_portfolio_exposure, R2)
        '''
        ind = onehot_to_index(self.industry_factors.reshape(self.N, self.P))
        date_off = np.array([0, self.N], dtype=np.int64)
        fr, spec, r2, status, expo = _ops.reg_by_time(
            date_off, self.capital, self.ret, ind, self._style_raw, self.P, want_exposure=True)
        return ((fr[0], spec, np.matrix(expo[0]), float(r2[0])))
