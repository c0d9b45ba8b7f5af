# main.py
"""The descriptor pipeline of Barra_factor_cal/main.py (steps 2-4, main.py:51-137) without its MongoDB
ends: the caller passes the three tables ``load_and_prepare_data`` would have produced and gets back the
table ``save_df_to_mongodb(db, final_barra_df, 'barra_factors')`` would have stored, plus the industry
table — so the risk model can take them straight from memory (``bmfm.handoff``) instead of through a
Mongo collection and a CSV export.  Function bodies follow main.py's call order line by line; the
arithmetic inside every call is the GPU drop-in of this package."""
import pandas as pd

try:        # run from inside the directory, like the reference's scripts (main.py:5-8) ...
    import config
    from factor_calculator import FactorCalculator
    from post_processing import winsorize_factors, calculate_composite_factors, orthogonalize_factors
except ImportError:      # ... or imported as the package Barra_factor_cal
    from . import config
    from .factor_calculator import FactorCalculator
    from .post_processing import winsorize_factors, calculate_composite_factors, orthogonalize_factors


def prepare_barra_format(processed_df, sw_industry_data):
    """main.py:95-113: industry code per stock, next-day return, rename, column order."""
    barra_df = pd.merge(processed_df, sw_industry_data[['ts_code', 'l1_code']], on='ts_code', how='left')
    barra_df['ret'] = barra_df.groupby('ts_code')['ret'].shift(-1)      # the return the exposures explain (t+1)
    barra_df.rename(columns=config.COLUMN_RENAME_MAP, inplace=True)
    final_columns = [col for col in config.BARRA_OUTPUT_COLUMNS if col in barra_df.columns]
    return barra_df[final_columns]


def build_industry_info(final_barra_df, sw_industry_data):
    """main.py:128-137: one row per industry that occurs, columns code / industry_names / start_date."""
    stk_list_df = final_barra_df[['stocknames']].drop_duplicates().rename(columns={'stocknames': 'ts_code'})
    ind_info = pd.merge(stk_list_df, sw_industry_data[['ts_code', 'l1_code', 'l1_name', 'in_date']],
                        on='ts_code', how='left')
    return ind_info.drop_duplicates(subset=['l1_code', 'l1_name']).rename(
        columns={'l1_code': 'code', 'l1_name': 'industry_names', 'in_date': 'start_date'}
    )[['code', 'industry_names', 'start_date']]


def run_factor_pipeline(stk_data, index_data, sw_industry_data, factors=None):
    """main.py:51-137 between the MongoDB load and save: (final_barra_df, ind_info)."""
    print("\nStep 2: Calculating raw factors...")
    calculator = FactorCalculator(prices_df=stk_data, index_df=index_data)
    raw_factors_df = calculator.run(config.FACTORS_TO_RUN if factors is None else factors)
    print("\nStep 3: Post-processing factors...")
    factor_columns = [col for col in raw_factors_df.columns if col not in ['ts_code', 'trade_date']]
    winsorized_df = winsorize_factors(raw_factors_df, factor_columns)
    composite_df = calculate_composite_factors(winsorized_df, config.COMPOSITE_CONFIG)
    processed_df = orthogonalize_factors(composite_df, config.ORTHO_RULES)
    print("\nStep 4: Preparing final Barra format...")
    final_barra_df = prepare_barra_format(processed_df, sw_industry_data)
    return final_barra_df, build_industry_info(final_barra_df, sw_industry_data)
