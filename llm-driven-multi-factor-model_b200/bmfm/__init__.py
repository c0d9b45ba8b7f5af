"""bmfm — B200-native kernels for the Barra risk-model path (Python host side).

``_cabi``   ctypes binding of libbmfm.so (include/bmfm.h); no CPU fallback
``panel``   DataFrame -> date-major packed panel (the HBM layout)
``ops``     array-level wrappers used by the drop-in ``mfm`` package, tests, bench
``synth``   seeded synthetic panels
"""
