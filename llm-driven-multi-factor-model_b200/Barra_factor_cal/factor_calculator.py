# factor_calculator.py
"""Drop-in for the hot-path methods of Barra_factor_cal/factor_calculator.py:
``FactorCalculator(prices_df, index_df)``, ``_prepare_data``,
``compute_beta_hsigma``, ``compute_dastd``, ``compute_cmra``, ``compute_rstr``,
``compute_liquidity``, ``compute_nlsize``, ``compute_size``, ``compute_bp``, ``compute_earnings_yield``,
``select_growth_factors``, ``compute_leverage`` and ``run`` — same DataFrame in / DataFrame out contract (results
keyed by ``original_index``), rolling windows and element-wise arithmetic on the GPU.

The long table is pivoted once into a dense [row-of-stock, stock] panel (each
stock's own rows in date order, which is what ``groupby('ts_code').rolling``
walks); kernels write the same layout and the results are scattered back to
``original_index``.  The report-period alignment of the financial-statement
factors (drop_duplicates / merge on (ts_code, end_date)) stays pandas glue as in
the reference; their arithmetic runs through bmfm_rolling_sum / bmfm_elementwise*.
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

    def compute_earnings_yield(self):
        """factor_calculator.py:371-433: CETOP = trailing-four-quarter operating cash flow / total_mv (both > 0),
        ETOP = 1 / pe_ttm (pe_ttm > 0).  None when a required column is missing (:382-387)."""
        print("Computing Earnings Yield factors (CETOP, ETOP)...")
        required_cols = ['n_cashflow_act', 'end_date', 'total_mv', 'pe_ttm']
        if not all(col in self.master_df.columns for col in required_cols):
            missing = [col for col in required_cols if col not in self.master_df.columns]
            print(f"\nERROR: Missing required columns for earnings yield calculation. "
                  f"Please ensure your merged data contains: {missing}\n")
            return None
        # one row per (stock, report period), in period order (:391-394)
        financial_data = self.master_df[['ts_code', 'end_date', 'n_cashflow_act']].drop_duplicates().copy()
        financial_data.sort_values(by=['ts_code', 'end_date'], inplace=True)
        codes, _ = pd.factorize(financial_data['ts_code'].values, sort=True)
        rows = financial_data.groupby('ts_code', sort=False).cumcount().values
        S = int(codes.max()) + 1 if len(codes) else 0
        L = int(rows.max()) + 1 if len(rows) else 0
        panel = np.full((L, S), np.nan)
        panel[rows, codes] = financial_data['n_cashflow_act'].values.astype(np.float64)
        ttm = _ops.rolling_sum(panel, 4, 4) if L and S else panel            # rolling(4, min_periods=4).sum() (:395-397)
        financial_data['n_cashflow_act_ttm'] = ttm[rows, codes]
        temp_df = pd.merge(self.master_df[['original_index', 'ts_code', 'end_date', 'total_mv', 'pe_ttm']],
                           financial_data[['ts_code', 'end_date', 'n_cashflow_act_ttm']],
                           on=['ts_code', 'end_date'], how='left')
        temp_df.sort_values(by='original_index', inplace=True)
        return pd.DataFrame({
            'original_index': temp_df['original_index'],
            'CETOP': _ops.cetop_factor(temp_df['n_cashflow_act_ttm'].values, temp_df['total_mv'].values),
            'ETOP': _ops.etop_factor(temp_df['pe_ttm'].values)})

    def select_growth_factors(self):
        """factor_calculator.py:436-462: YOYProfit / YOYSales = q_profit_yoy / 100, q_sales_yoy / 100."""
        print("Selecting Growth factors (YOYProfit, YOYSales)...")
        required_cols = ['q_profit_yoy', 'q_sales_yoy']
        if not all(col in self.master_df.columns for col in required_cols):
            print(f"\nERROR: Missing required columns for growth factors. "
                  f"Please ensure your merged data contains: {required_cols}\n")
            return None
        return pd.DataFrame({'original_index': self.master_df['original_index'].values,
                             'YOYProfit': _ops.percent_factor(self.master_df['q_profit_yoy'].values),
                             'YOYSales': _ops.percent_factor(self.master_df['q_sales_yoy'].values)})

    def compute_leverage(self):
        """factor_calculator.py:464-509: MLEV, DTOA (passed through), BLEV."""
        print("Computing Leverage factors (MLEV, DTOA, BLEV)...")
        required_cols = ['total_mv', 'total_ncl', 'total_hldr_eqy_inc_min_int', 'debt_to_assets']
        if not all(col in self.master_df.columns for col in required_cols):
            print(f"\nERROR: Missing required columns for leverage calculation. "
                  f"Please ensure your merged data contains: {required_cols}\n")
            return None
        m = self.master_df
        return pd.DataFrame({
            'original_index': m['original_index'],
            'MLEV': _ops.mlev_factor(m['total_ncl'].values, m['total_mv'].values),
            'DTOA': m['debt_to_assets'],
            'BLEV': _ops.blev_factor(m['total_ncl'].values, m['total_hldr_eqy_inc_min_int'].values)})

    def run(self, factors: list) -> pd.DataFrame:
        """factor_calculator.py:515-577: compute each requested factor and left-merge it on original_index;
        an unknown name or a method that returns None is reported and skipped, as in the reference."""
        print(f"\nStarting factor calculation for: {factors}")
        factor_methods = {
            'SIZE': self.compute_size, 'BETA': self.compute_beta_hsigma, 'RSTR': self.compute_rstr,
            'DASTD': self.compute_dastd, 'CMRA': self.compute_cmra, 'NLSIZE': self.compute_nlsize,
            'BP': self.compute_bp, 'LIQUIDITY': self.compute_liquidity, 'EARNINGS': self.compute_earnings_yield,
            'GROWTH': self.select_growth_factors, 'LEVERAGE': self.compute_leverage,
        }
        final_df = self.master_df[['ts_code', 'trade_date', 'original_index', 'ret', 'circ_mv']].copy()
        print("\nIteratively computing and merging factors...")
        for factor_name in factors:
            method = factor_methods.get(factor_name.upper())
            if not method:
                print(f"Warning: Factor '{factor_name}' not found.")
                continue
            factor_df = method()
            if factor_df is not None:
                final_df = pd.merge(final_df, factor_df, on='original_index', how='left')
            else:
                print(f"Warning: Method for '{factor_name}' returned None.")
        final_df.drop(columns='original_index', inplace=True)
        print("Factor calculation complete.")
        return final_df
