"""Import the byte-compiled reference ``mfm`` package from oracle/_ref/ (see oracle/build_ref.py).

Returns the reference's own modules; used by bench.py's CPU legs only.  ``None`` when oracle/_ref/ has not
been built (then the callers fall back to the numpy port and say so)."""
import importlib
import os
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def load():
    if not os.path.exists(os.path.join(REF_DIR, "mfm", "MFM.pyc")):
        return None
    from oracle.ref_stubs import install_stubs
    install_stubs()
    warnings.filterwarnings("ignore", category=PendingDeprecationWarning)
    saved = {k: v for k, v in sys.modules.items() if k == "mfm" or k.startswith("mfm.")}
    for k in saved:
        del sys.modules[k]
    sys.path.insert(0, REF_DIR)
    try:
        mods = tuple(importlib.import_module(n) for n in ("mfm.MFM", "mfm.CrossSection", "mfm.utils"))
        assert all(os.path.dirname(m.__file__) == os.path.join(REF_DIR, "mfm") for m in mods)
    finally:
        sys.path.remove(REF_DIR)
        for k in [k for k in sys.modules if k == "mfm" or k.startswith("mfm.")]:
            del sys.modules[k]
        sys.modules.update(saved)          # the drop-in `mfm`, if it was imported, stays what `import mfm` gives
    return mods
