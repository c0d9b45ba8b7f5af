"""Drop-in for the hot-path part of the reference's Barra_factor_cal directory
(factor_calculator.py, post_processing.py, config.py).  The reference imports
these as top-level modules (main.py:5-7); put this directory on sys.path to do
the same."""
