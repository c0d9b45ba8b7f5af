# -*- coding: utf-8 -*-
"""Drop-in for the reference's ``mfm`` package (Barra-master/mfm/__init__.py:9-11):
same modules, classes, call signatures and DataFrame-in / DataFrame-out
contract, computed by hand-written sm_100a CUDA through libbmfm.so."""
__author__ = 'b200-barra'

__all__ = ["CrossSection", "utils", 'MFM']
