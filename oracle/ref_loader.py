"""Import the byte-compiled reference ``mfm`` package from oracle/_ref/ (see oracle/build_ref.py).

Returns the reference's own modules; used by bench.py's CPU legs only.  ``None`` when oracle/_ref/ has not
been built (then the callers fall back to the numpy port and say so)."""
import os
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def _load_module(name, path, package_dir=None):
    import importlib.machinery
    import importlib.util
    loader = importlib.machinery.SourcelessFileLoader(name, path)
    spec = importlib.util.spec_from_file_location(name, path, loader=loader,
                                                  submodule_search_locations=[package_dir] if package_dir else None)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    loader.exec_module(mod)
    return mod


def load():
    d = os.path.join(REF_DIR, "mfm")
    if not os.path.exists(os.path.join(d, "MFM.bc")):
        return None
    from oracle.ref_stubs import install_stubs
    install_stubs()
    warnings.filterwarnings("ignore", category=PendingDeprecationWarning)
    saved = {k: v for k, v in sys.modules.items() if k == "mfm" or k.startswith("mfm.")}
    for k in saved:
        del sys.modules[k]
    try:
        # the package first, then the modules in dependency order (MFM imports CrossSection and utils by name)
        _load_module("mfm", os.path.join(d, "__init__.bc"), package_dir=d)
        utils = _load_module("mfm.utils", os.path.join(d, "utils.bc"))
        cs = _load_module("mfm.CrossSection", os.path.join(d, "CrossSection.bc"))
        mfm = _load_module("mfm.MFM", os.path.join(d, "MFM.bc"))
        mods = (mfm, cs, utils)
    finally:
        for k in [k for k in sys.modules if k == "mfm" or k.startswith("mfm.")]:
            del sys.modules[k]
        sys.modules.update(saved)          # the drop-in `mfm`, if it was imported, stays what `import mfm` gives
    return mods
