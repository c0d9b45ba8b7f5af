# factor_calculator.py
"""Drop-in for the hot-path methods of Barra_factor_cal/factor_calculator.py:
``FactorCalculator(prices_df, index_df)``, ``_prepare_data``,
``compute_beta_hsigma``, ``compute_dastd``, ``compute_cmra`` and ``run`` for
those factors — same DataFrame in / DataFrame out contract (results keyed by
``original_index``), rolling windows on the GPU.

The long table is pivoted once into a dense [row-of-stock, stock] panel (each
stock's own rows in date order, which is what ``groupby('ts_code').rolling``
walks); kernels write the same layout and the results are scattered back to
``original_index``.  Other ``compute_*`` methods of the reference are outside
the hot path and raise NotImplementedError.
"""
import numpy as np
import pandas as pd

from bmfm import ops as _ops


class FactorCalculator:
    def __init__(self, prices_df: pd.DataFrame, index_df: pd.DataFrame):
        print("Initializing Factor Calculator...")
        self.prices_df = prices_df
        self.index_df = index_df
        self._prepare_data()

    def _prepare_data(self):
        """factor_calculator.py:34-64: date strings, sort, per-stock returns, market merge."""
        print("Preparing base data (returns, etc.)...")
        self.index_df['trade_date'] = self.index_df['trade_date'].dt.strftime('%Y/%m/%d')
        self.prices_df['trade_date'] = self.prices_df['trade_date'].dt.strftime('%Y/%m/%d')
        self.prices_df.sort_values(by=['ts_code', 'trade_date'], inplace=True)
        self.index_df.sort_values(by='trade_date', inplace=True)

        # dense [row, stock] panel of closes -> ret / log_ret on the GPU (:50-51)
        codes, uniques = pd.factorize(self.prices_df['ts_code'].values, sort=True)
        row_in_stock = self.prices_df.groupby('ts_code', sort=False).cumcount().values
        S = len(uniques)
        L = int(row_in_stock.max()) + 1 if len(row_in_stock) else 0
        self._codes, self._rows, self._L, self._S = codes, row_in_stock, L, S
        close = np.full((L, S), np.nan)
        close[row_in_stock, codes] = self.prices_df['close'].values.astype(np.float64)
        ret, log_ret = _ops.prepare_returns(close)
        self.prices_df['ret'] = ret[row_in_stock, codes]
        self.prices_df['log_ret'] = log_ret[row_in_stock, codes]

        self.index_df['market_ret'] = self.index_df['close'].pct_change()
        self.master_df = pd.merge(self.prices_df, self.index_df[['trade_date', 'market_ret']],
                                  on='trade_date', how='left')
        self.master_df.reset_index(inplace=True)
        self.master_df.rename(columns={'index': 'original_index'}, inplace=True)
        # the merge keeps the left order, so (row, stock) coordinates carry over to master_df
        self._ret = ret
        self._log_ret = log_ret
        mkt = np.full((L, S), np.nan)
        mkt[row_in_stock, codes] = self.master_df['market_ret'].values.astype(np.float64)
        self._mkt = mkt
        print("Data preparation complete.")

    def _gather(self, panel):
        return panel[self._rows, self._codes]

    def compute_beta_hsigma(self):
        """factor_calculator.py:79-125"""
        print("Computing BETA and HSIGMA factors...")
        beta, hsigma = _ops.beta_hsigma(self._ret, self._mkt, 252, 63, 42)
        return pd.DataFrame({'original_index': self.master_df['original_index'].values,
                             'BETA': self._gather(beta), 'HSIGMA': self._gather(hsigma)})

    def compute_dastd(self):
        """factor_calculator.py:155-196"""
        print("Computing DASTD factor...")
        out = _ops.dastd(self._ret, self._mkt, 252, 42, 42)
        return pd.DataFrame({'original_index': self.master_df['original_index'].values,
                             'DASTD': self._gather(out)})

    def compute_cmra(self):
        """factor_calculator.py:199-234"""
        print("Computing CMRA factor...")
        out = _ops.cmra(self._log_ret, 252)
        return pd.DataFrame({'original_index': self.master_df['original_index'].values,
                             'CMRA': self._gather(out)})

    def run(self, factors_to_compute: list) -> pd.DataFrame:
        """factor_calculator.py:515-577 for the hot-path factors: merge results on original_index."""
        methods = {'BETA': self.compute_beta_hsigma, 'DASTD': self.compute_dastd, 'CMRA': self.compute_cmra}
        final = self.master_df[['original_index', 'ts_code', 'trade_date', 'ret']].copy()
        for name in factors_to_compute:
            key = name.upper()
            if key not in methods:
                raise NotImplementedError(f"factor {name!r} is outside the accelerated hot path (SURVEY.md §8(f))")
            final = pd.merge(final, methods[key](), on='original_index', how='left')
        return final.drop(columns=['original_index'])
