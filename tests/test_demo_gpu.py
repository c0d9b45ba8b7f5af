"""Config C1 / the north-star target: the reference's own ``Barra-master/demo.py`` executed BYTE-UNCHANGED with
the drop-in ``mfm`` package first on the import path, on stand-ins for the two CSV files it reads (they are not
in the reference tree), and its five result files compared with what the unmodified reference
wrote for the same inputs (tests/golden/demo.npz, made by tests/golden/make_golden.py with the eigen order
canonical).  ``/root/reference`` does not exist on the GPU box, so demo.py's bytes travel inside the fixture
and are checked against their SHA-256 (and, where the reference tree is present, against the file itself)."""
import hashlib
import os
import subprocess
import sys

import numpy as np
import pytest

from util import record, relerr, relerr_cols, relerr_cov

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "llm-driven-multi-factor-model_b200")
TOL = 1e-9


@pytest.fixture(scope="module")
def demo(golden_dir):
    return np.load(os.path.join(golden_dir, "demo.npz"))


def _case(demo):
    return {k[5:]: int(demo[k]) for k in demo.files if k.startswith("case_")}


def test_fixture_holds_the_reference_file_and_inputs_regenerate(demo, tmp_path):
    """CPU: the bytes in the fixture are the reference's demo.py, and the seeded CSV pair comes out identical."""
    src = demo["demo_py"].tobytes()
    assert hashlib.sha256(src).hexdigest() == str(demo["demo_py_sha256"])
    ref = "/root/reference/Barra-master/demo.py"
    if os.path.exists(ref):
        assert open(ref, "rb").read() == src
    from bmfm.synth import write_demo_csv_pair
    case = dict(_case(demo), T=40)          # the generator is prefix-stable only per call: check determinism
    a = write_demo_csv_pair(str(tmp_path / "a"), **case)
    b = write_demo_csv_pair(str(tmp_path / "b"), **case)
    assert open(a[0], "rb").read() == open(b[0], "rb").read() and open(a[1], "rb").read() == open(b[1], "rb").read()


@pytest.mark.gpu
def test_reference_demo_py_runs_unchanged_on_the_dropin(demo, tmp_path):
    from bmfm.handoff import read_results
    from bmfm.synth import write_demo_csv_pair
    bm = tmp_path / "Barra-master"
    bm.mkdir()
    src = demo["demo_py"].tobytes()
    assert hashlib.sha256(src).hexdigest() == str(demo["demo_py_sha256"])
    (bm / "demo.py").write_bytes(src)
    paths = write_demo_csv_pair(str(bm / "data_new"), **_case(demo))
    assert [hashlib.sha256(open(p, "rb").read()).hexdigest() for p in paths] == list(demo["csv_sha256"])
    env = dict(os.environ, PYTHONPATH=PKG + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, "demo.py"], cwd=str(bm), env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    got = read_results(str(bm / "results"))
    assert list(got["factor_ret_columns"]) == list(demo["factor_ret_columns"])
    assert list(got["factor_ret_index"]) == list(demo["factor_ret_index"])
    assert record("demo.factor_returns[per column]", relerr_cols(got["factor_ret"], demo["factor_ret"])) < TOL
    assert record("demo.r_squared", np.max(np.abs(got["R2"] - demo["R2"]))) < TOL
    assert list(got["specific_columns"]) == list(demo["specific_columns"])
    keep = demo["specific_keep"]
    full = got["specific_full"]
    scale = np.nanmax(np.abs(demo["specific_rows"]))
    assert record("demo.specific_returns", relerr(full[keep], demo["specific_rows"], scale=scale)) < TOL
    assert np.array_equal(np.isfinite(full).sum(axis=1), demo["specific_count"])
    assert np.max(np.abs(np.nansum(full, axis=1) - demo["specific_sum"])) < TOL * scale * full.shape[1]
    assert np.max(np.abs(np.nansum(full * full, axis=1) - demo["specific_sumsq"])) < TOL * scale * scale * full.shape[1]
    assert list(got["final_cov_labels"]) == list(demo["final_cov_labels"])
    assert record("demo.final_covariance[per element / sqrt(diag diag)]", relerr_cov(got["final_cov"], demo["final_cov"])) < TOL
    assert record("demo.lambda", relerr(got["lamb"], demo["lamb"])) < TOL
