# post_processing.py
"""Drop-in for Barra_factor_cal/post_processing.py: same three functions and
signatures; per-date statistics and regressions run on the GPU."""
import numpy as np
import pandas as pd

from bmfm import ops as _ops


def _by_date(df):
    """Stable grouping of the rows by trade_date -> (order, date_off)."""
    codes, _ = pd.factorize(df['trade_date'].values, sort=True)
    order = np.argsort(codes, kind='stable')
    counts = np.bincount(codes)
    off = np.zeros(len(counts) + 1, dtype=np.int64)
    np.cumsum(counts, out=off[1:])
    return order, off


def winsorize_factors(factor_df: pd.DataFrame, factor_list: list, n_std: float = 2.5) -> pd.DataFrame:
    """post_processing.py:7-24"""
    df = factor_df.copy()
    print(f"\n--- Starting Factor Winsorization (Boundary: Mean ± {n_std} * StdDev) ---")
    order, off = _by_date(df)
    for factor in factor_list:
        if factor in df.columns:
            print(f"Processing factor: '{factor}'...")
            x = df[factor].values.astype(np.float64)[order]
            out = np.empty_like(x)
            out[order] = _ops.winsorize(off, x, n_std)
            df[factor] = out
        else:
            print(f"Warning: Factor '{factor}' not found in DataFrame. Skipping.")
    print("--- Factor Winsorization Complete ---")
    return df


def calculate_composite_factors(factor_df: pd.DataFrame, composite_factor_config: dict) -> pd.DataFrame:
    """post_processing.py:26-45"""
    df = factor_df.copy()
    for new_factor, config in composite_factor_config.items():
        print(f"Calculating composite factor: '{new_factor}'...")
        comps, ws = [], []
        for name, w in zip(config['components'], config['weights']):
            if name in df.columns:
                comps.append(df[name].values.astype(np.float64))
                ws.append(w)
            else:
                print(f"Warning: Component '{name}' not found in DataFrame. Skipping.")
        if comps:
            df[new_factor] = _ops.composite(comps, ws)
        else:
            df[new_factor] = np.nan
    print("\n--- Composite Factor Calculation Complete ---")
    return df


def orthogonalize_factors(factor_df: pd.DataFrame, ortho_config: dict) -> pd.DataFrame:
    """post_processing.py:47-70: rules are applied one after another on the same (copied) frame,
    so a later rule sees the columns an earlier rule replaced (:63-68)."""
    df = factor_df.copy()
    print("\n--- Starting Factor Orthogonalization ---")
    order, off = _by_date(df)
    for factor_to_ortho, against_factors in ortho_config.items():
        print(f"Orthogonalizing '{factor_to_ortho}' against {against_factors}...")
        y = df[factor_to_ortho].values.astype(np.float64)[order]
        X = [df[c].values.astype(np.float64)[order] for c in against_factors]
        res = np.empty_like(y)
        res[order] = _ops.orthogonalize(off, y, np.column_stack(X))
        df[factor_to_ortho] = res
    print("--- Factor Orthogonalization Complete ---")
    return df
