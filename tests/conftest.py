"""pytest configuration: markers and import paths.

The product package directory carries the reference's own top-level module
names (``mfm``, ``Barra_factor_cal``) so that ``Barra-master/demo.py`` runs
unchanged; it is put on ``sys.path`` here.  ``oracle`` is the CPU checker and
is imported by tests only.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "llm-driven-multi-factor-model_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def pytest_sessionfinish(session, exitstatus):
    """Write the errors the parity checks observed (tests/util.py:record) where gpurun collects them."""
    try:
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        import json
        import util
        if util.OBSERVED:
            out = os.path.join(ROOT, "gpurun_out")
            os.makedirs(out, exist_ok=True)
            with open(os.path.join(out, "parity_observed.json"), "w") as f:
                json.dump(dict(sorted(util.OBSERVED.items())), f, indent=1)
    except Exception:
        pass


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture
def bmfm_option():
    """Set experiment switches on the default context for one test (bmfm_ctx_set_option); reset afterwards."""
    from bmfm import _cabi
    ctx = _cabi.default_context()
    touched = []

    def setter(name, value):
        ctx.set_option(name, int(value))
        touched.append(name)

    yield setter
    for name in touched:
        ctx.set_option(name, -1 if name == "reg_smma" else 0)
