# config.py — constants of the descriptor pipeline (Barra_factor_cal/config.py:17-72).
# Paths and the MongoDB wiring of the reference are out of scope.

FACTORS_TO_RUN = [
    'SIZE', 'BETA', 'RSTR', 'DASTD', 'CMRA', 'NLSIZE', 'BP',
    'LIQUIDITY', 'EARNINGS', 'GROWTH', 'LEVERAGE'
]

COMPOSITE_CONFIG = {
    'volatility': {'components': ['DASTD', 'CMRA', 'HSIGMA'], 'weights': [0.7, 0.15, 0.15]},
    'leverage': {'components': ['MLEV', 'DTOA', 'BLEV'], 'weights': [1/3, 1/3, 1/3]},
    'liquidity': {'components': ['STOM', 'STOQ', 'STOA'], 'weights': [0.5, 0.25, 0.25]},
    'earnings': {'components': ['CETOP', 'ETOP'], 'weights': [0.5, 0.5]},
    'growth': {'components': ['YOYProfit', 'YOYSales'], 'weights': [0.5, 0.5]},
}

ORTHO_RULES = {
    'volatility': ['BETA', 'SIZE'],
    'liquidity': ['SIZE'],
}

COLUMN_RENAME_MAP = {
    'trade_date': 'date', 'ts_code': 'stocknames', 'l1_code': 'industry', 'circ_mv': 'capital',
    'SIZE': 'size', 'BETA': 'beta', 'RSTR': 'momentum', 'volatility': 'residual_volatility',
    'NLSIZE': 'non_linear_size', 'BP': 'book_to_price_ratio', 'earnings': 'earnings_yield',
}

BARRA_OUTPUT_COLUMNS = [
    'date', 'stocknames', 'capital', 'ret', 'industry',
    'size', 'beta', 'momentum', 'residual_volatility', 'non_linear_size',
    'book_to_price_ratio', 'liquidity', 'earnings_yield', 'growth', 'leverage'
]
