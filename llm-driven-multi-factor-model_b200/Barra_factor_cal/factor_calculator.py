# factor_calculator.py
"""Drop-in for the hot-path methods of Barra_factor_cal/factor_calculator.py:
``FactorCalculator(prices_df, index_df)``, ``_prepare_data``,
``compute_beta_hsigma``, ``compute_dastd``, ``compute_cmra``, ``compute_rstr``,
``compute_liquidity``, ``compute_nlsize``, ``compute_size``, ``compute_bp`` and ``run`` for those factors — same DataFrame in / DataFrame out contract (results keyed by
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

    def _panel(self, column):
        """[row-of-stock, stock] panel of a master_df column (NaN where the stock has no such row)."""
        p = np.full((self._L, self._S), np.nan)
        p[self._rows, self._codes] = self.master_df[column].values.astype(np.float64)
        return p

    def compute_size(self):
        """factor_calculator.py:68-77"""
        print("Computing SIZE factor...")
        return pd.DataFrame({'original_index': self.master_df['original_index'].values,
                             'SIZE': _ops.size_factor(self.master_df['total_mv'].values)})

    def compute_bp(self):
        """factor_calculator.py:295-321: None when 'pb' is missing (:305-308)"""
        print("Computing Value (BP) factor...")
        if 'pb' not in self.master_df.columns:
            print("\nERROR: 'pb' column not found in the input data.")
            print("To compute the BP factor, you must provide Price-to-Book ratio data.\n")
            return None
        return pd.DataFrame({'original_index': self.master_df['original_index'].values,
                             'BP': _ops.bp_factor(self.master_df['pb'].values)})

    def compute_rstr(self):
        """factor_calculator.py:127-153 (T 504, lag 21, half-life 126, min_periods 42; the oldest row of the
        window carries weight 1, as in the reference)"""
        print("Computing RSTR factor...")
        out = _ops.rstr(self._log_ret, 504, 21, 126, 42)
        return pd.DataFrame({'original_index': self.master_df['original_index'].values, 'RSTR': self._gather(out)})

    def compute_liquidity(self):
        """factor_calculator.py:324-367: STOM, STOQ, STOA; None when 'turnover_rate' is missing (:333-335)"""
        print("Computing base Liquidity factors (STOM, STOQ, STOA)...")
        if 'turnover_rate' not in self.master_df.columns:
            print("\nERROR: 'turnover_rate' column not found in the input data.\n")
            return None
        stom, stoq, stoa = _ops.liquidity(self._panel('turnover_rate'))
        return pd.DataFrame({'original_index': self.master_df.index,
                             'STOM': self._gather(stom), 'STOQ': self._gather(stoq), 'STOA': self._gather(stoa)})

    def compute_nlsize(self):
        """factor_calculator.py:237-293: per-date OLS of SIZE^3 on SIZE, minus the residual"""
        print("Computing NonLinear Size (NLSIZE) factor...")
        with np.errstate(invalid='ignore', divide='ignore'):
            size = np.log(self.master_df['total_mv'].values.astype(np.float64))
        # rows grouped by trade_date for the per-date kernel; results go back to master_df's row order
        codes, _ = pd.factorize(self.master_df['trade_date'].values, sort=True)
        order = np.argsort(codes, kind='stable')
        date_off = np.concatenate([[0], np.cumsum(np.bincount(codes))]).astype(np.int64)
        res = np.empty(len(size))
        res[order] = _ops.nlsize(date_off, size[order])
        return pd.DataFrame({'original_index': self.master_df['original_index'].values, 'NLSIZE': res})

    def run(self, factors_to_compute: list) -> pd.DataFrame:
        """factor_calculator.py:515-577 for the hot-path factors: merge results on original_index."""
        methods = {'BETA': self.compute_beta_hsigma, 'DASTD': self.compute_dastd, 'CMRA': self.compute_cmra,
                   'RSTR': self.compute_rstr, 'LIQUIDITY': self.compute_liquidity, 'NLSIZE': self.compute_nlsize,
                   'SIZE': self.compute_size, 'BP': self.compute_bp}
        final = self.master_df[['original_index', 'ts_code', 'trade_date', 'ret']].copy()
        for name in factors_to_compute:
            key = name.upper()
            if key not in methods:
                raise NotImplementedError(f"factor {name!r} is outside the accelerated hot path (SURVEY.md §8(f))")
            res = methods[key]()
            if res is not None:
                final = pd.merge(final, res, on='original_index', how='left')
        return final.drop(columns=['original_index'])
