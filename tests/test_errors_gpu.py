"""Argument and limit checks of the C-ABI: a bad call returns an error code with a message (raised as
BmfmError by the ctypes layer) and launches nothing; empty inputs are accepted."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    from bmfm import _cabi
    return torch, _cabi, _cabi.default_context()


def test_regression_limits_and_null_pointers(env):
    torch, _cabi, ctx = env
    d = lambda *s, dt=torch.float64: torch.zeros(s, dtype=dt, device="cuda")
    off = torch.tensor([0, 4], dtype=torch.int64, device="cuda")
    n0 = ctx.launches
    with pytest.raises(_cabi.BmfmError, match="compiled limits"):          # Q above the template range
        ctx.reg_by_time(1, 3, 13, 4, off, d(4), d(4), d(4, dt=torch.int32), d(4, 13), d(1, 17), d(4), d(1),
                        d(1, dt=torch.int32))
    with pytest.raises(_cabi.BmfmError, match="compiled limits"):          # more than 63 industries
        ctx.reg_by_time(1, 64, 2, 4, off, d(4), d(4), d(4, dt=torch.int32), d(4, 2), d(1, 67), d(4), d(1),
                        d(1, dt=torch.int32))
    with pytest.raises(_cabi.BmfmError, match="NULL"):
        ctx.reg_by_time(1, 3, 2, 4, off, None, d(4), d(4, dt=torch.int32), d(4, 2), d(1, 6), d(4), d(1),
                        d(1, dt=torch.int32))
    assert ctx.launches == n0
    # T = 0 is a valid empty panel
    ctx.reg_by_time(0, 3, 2, 0, torch.zeros(1, dtype=torch.int64, device="cuda"), d(0), d(0), d(0, dt=torch.int32),
                    d(0, 2), d(0, 6), d(0), d(0), d(0, dt=torch.int32))
    assert ctx.launches == n0


def test_covariance_entry_points_reject_bad_sizes(env):
    torch, _cabi, ctx = env
    d = lambda *s, dt=torch.float64: torch.zeros(s, dtype=dt, device="cuda")
    T, K = 20, 4
    fr, cov, st = d(T, K), d(T, K, K), d(T, dt=torch.int32)
    n0 = ctx.launches
    with pytest.raises(_cabi.BmfmError, match="exceeds compiled limit"):   # more than 4 lags
        ctx.newey_west_by_time(T, K, 5, 10.0, fr, cov, st)
    with pytest.raises(_cabi.BmfmError, match="bad sizes"):
        ctx.newey_west_by_time(T, K, 2, -1.0, fr, cov, st)
    with pytest.raises(_cabi.BmfmError, match="bad sizes"):
        ctx.newey_west_by_time_range(T, K, 2, 10.0, fr, 12, 8, cov, st)    # t1 < t0
    with pytest.raises(_cabi.BmfmError, match="bad sizes"):
        ctx.vol_regime_adj_by_time(T, K, 0.0, fr, cov, st, cov.clone(), d(T))
    with pytest.raises(_cabi.BmfmError):
        ctx.rstr(10, 3, d(10, 3), 21, 21, 126.0, 42, d(10, 3))             # window T - lag must be positive
    with pytest.raises(_cabi.BmfmError):
        ctx.rolling_log_sum(10, 3, d(10, 3), 0.0, 21, 15, d(10, 3))        # zero divisor
    assert ctx.launches == n0
    ctx.newey_west_by_time(0, K, 2, 10.0, fr, cov, st)                     # empty series: nothing to do
    assert ctx.launches == n0


def test_short_series_status_not_error(env):
    """A series too short for any window is not an error: every date gets BMFM_ST_SHORT and a NaN matrix."""
    torch, _cabi, ctx = env
    T, K = 3, 5
    fr = torch.randn(T, K, dtype=torch.float64, device="cuda")
    cov = torch.zeros(T, K, K, dtype=torch.float64, device="cuda")
    st = torch.zeros(T, dtype=torch.int32, device="cuda")
    ctx.newey_west_by_time(T, K, 2, 10.0, fr, cov, st)
    assert (st.cpu().numpy() == _cabi.ST_SHORT).all() and bool(torch.isnan(cov).all())


@pytest.mark.parametrize("mode", ["1", "2", "3", "4", "5", "6"])
def test_empty_date_inside_a_panel(env, mode, bmfm_option):
    """A date with no rows gets NaN factor returns and BMFM_ST_NONFINITE; its neighbours are unaffected."""
    torch, _cabi, ctx = env
    bmfm_option("reg_mode", mode)
    from bmfm.panel import pack_arrays
    from bmfm.synth import make_panel
    pan = make_panel(4, 90, 6, 9, seed=31)
    csr = pan.to_csr()
    (cap, ret, ind, style), _ = pack_arrays(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P)
    off = csr["date_off"]
    off2 = np.concatenate([off[:2], [off[1]], off[2:]]).astype(np.int64)      # an empty date after the first one
    T, K = len(off2) - 1, 1 + pan.P + pan.Q
    dv = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()

    def run(o):
        n = len(o) - 1
        fr = torch.empty((n, K), dtype=torch.float64, device="cuda")
        spec = torch.empty(len(cap), dtype=torch.float64, device="cuda")
        r2 = torch.empty(n, dtype=torch.float64, device="cuda")
        st = torch.empty(n, dtype=torch.int32, device="cuda")
        ctx.reg_by_time(n, pan.P, pan.Q, pan.N, dv(o), dv(cap), dv(ret), dv(ind), dv(style), fr, spec, r2, st)
        return fr.cpu().numpy(), spec.cpu().numpy(), r2.cpu().numpy(), st.cpu().numpy()

    fr0, spec0, r20, st0 = run(off)
    fr1, spec1, r21, st1 = run(off2)
    assert (st0 == 0).all()
    assert st1[1] == _cabi.ST_NONFINITE and np.isnan(fr1[1]).all() and np.isnan(r21[1])
    keep = [0, 2, 3, 4]
    assert (st1[keep] == 0).all()
    assert np.array_equal(fr1[keep], fr0) and np.array_equal(spec1, spec0) and np.array_equal(r21[keep], r20)
