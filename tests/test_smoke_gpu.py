"""The driver's smoke entry point, run with the GPU suite so that it cannot rot unnoticed."""
import os
import sys

import pytest


@pytest.mark.gpu
def test_graft_entry_smoke():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    import __graft_entry__ as g
    g.smoke()
