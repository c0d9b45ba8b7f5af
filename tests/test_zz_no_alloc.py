"""The allocation-free Newey-West entry point (bmfm_newey_west_by_time_noalloc).

Who needs it: a host process whose GPU memory was divided up front among pools this library
does not own (a library-side cudaMalloc could fail there, or eat someone else's share), and
loops that cannot afford the device-wide stall that cudaMalloc / cudaFree cause.  The regular
entry point grows a workspace on first use; this one keeps its chunk states in a fixed array
inside the library and must give the same covariances.

The file name sorts last on purpose: a run that stops at the first failure has then already
been through every other test.  It passes on its own in a fresh process as well (the first call
on a device does the one-time address look-up).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(ctx, fn, fr, q, tao):
    import torch
    T, K = fr.shape
    d_fr = torch.from_numpy(fr).cuda()
    cov = torch.empty((T, K, K), dtype=torch.float64, device="cuda")
    st = torch.empty(T, dtype=torch.int32, device="cuda")
    fn(T, K, q, tao, d_fr, cov, st)
    torch.cuda.synchronize()
    return cov.cpu().numpy(), st.cpu().numpy()


def test_noalloc_matches_regular_entry_twice_in_a_row():
    from bmfm import _cabi
    from bmfm.synth import make_factor_returns
    ctx = _cabi.default_context()
    # two calls in a row with different sizes: the second one finds the first one's leftovers in
    # the shared array and must not be affected by them
    for (T, K, q, tao, seed) in [(900, 42, 2, 252.0, 1), (300, 17, 3, 40.0, 2)]:
        fr = make_factor_returns(T, K, seed=seed) + 0.002
        got, got_st = _run(ctx, ctx.newey_west_by_time_noalloc, fr, q, tao)
        want, want_st = _run(ctx, ctx.newey_west_by_time, fr, q, tao)
        assert np.array_equal(got_st, want_st)
        ok = want_st == 0
        assert ok.any() and np.isnan(got[~ok]).all()
        scale = np.abs(want[ok]).max()
        assert np.abs(got[ok] - want[ok]).max() <= 1e-11 * scale     # same tolerance as test_risk_gpu.py


def test_noalloc_leaves_free_device_memory_unchanged():
    """cudaMemGetInfo before and after: with the outputs allocated up front, a call of the allocation-free
    entry point must not change the free device memory (its first call on the device is made beforehand: that
    one loads the kernels and looks the array's address up).  The regular entry point on a fresh context is
    measured the same way to show that the probe does see a library-side cudaMalloc."""
    import torch
    from bmfm import _cabi
    from bmfm.synth import make_factor_returns
    T, K, q, tao = 2500, 42, 2, 252.0
    d_fr = torch.from_numpy(make_factor_returns(T, K, seed=4)).cuda()
    cov = torch.empty((T, K, K), dtype=torch.float64, device="cuda")
    st = torch.empty(T, dtype=torch.int32, device="cuda")
    ctx = _cabi.Context(torch.cuda.current_device())          # fresh: no workspace yet
    ctx.newey_west_by_time_noalloc(T, K, q, tao, d_fr, cov, st)   # first call: module load + address look-up
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(2):
        ctx.newey_west_by_time_noalloc(T, K, q, tao, d_fr, cov, st)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0
    ctx.newey_west_by_time(T, K, q, tao, d_fr, cov, st)       # grows the context workspace with cudaMalloc
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] < free0
    ctx.close()


def test_noalloc_long_series_uses_longer_chunks():
    """T = 30 000 at K = 42, q = 2: the ordinary chunk length would need more chunk states than the fixed array
    holds, so the entry point lengthens the chunks (more shared memory per CTA of the emit kernel) — the one path
    on which its limits matter.  Checked against the regular entry point on a sample of dates."""
    import torch
    from bmfm import _cabi
    from bmfm.synth import make_factor_returns
    ctx = _cabi.default_context()
    T, K, q, tao = 30000, 42, 2, 252.0
    d_fr = torch.from_numpy(make_factor_returns(T, K, seed=6) + 0.001).cuda()
    st_a = torch.empty(T, dtype=torch.int32, device="cuda")
    st_b = torch.empty(T, dtype=torch.int32, device="cuda")
    cov_a = torch.empty((T, K, K), dtype=torch.float64, device="cuda")
    ctx.newey_west_by_time_noalloc(T, K, q, tao, d_fr, cov_a, st_a)
    pick = torch.tensor(sorted(set(list(range(0, T, 997)) + [K, K + 1, T - 1])), device="cuda")
    got = cov_a[pick].cpu().numpy()
    del cov_a
    cov_b = torch.empty((T, K, K), dtype=torch.float64, device="cuda")
    ctx.newey_west_by_time(T, K, q, tao, d_fr, cov_b, st_b)
    want = cov_b[pick].cpu().numpy()
    assert torch.equal(st_a, st_b)
    ok = st_b[pick].cpu().numpy() == 0
    assert ok.any()
    assert np.abs(got[ok] - want[ok]).max() <= 1e-11 * np.abs(want[ok]).max()


def test_noalloc_refuses_series_past_its_limit():
    """Just past the documented limit (about T = 85 000 at K = 42, q = 2) the call is refused with a message and
    nothing is launched."""
    import torch
    from bmfm import _cabi
    ctx = _cabi.default_context()
    T, K = 90000, 42
    fr = torch.zeros((T, K), dtype=torch.float64, device="cuda")
    cov = torch.empty((1,), dtype=torch.float64, device="cuda")      # never touched: the call is refused first
    st = torch.empty(T, dtype=torch.int32, device="cuda")
    n0 = ctx.launches
    with pytest.raises(_cabi.BmfmError, match="fixed"):
        ctx.newey_west_by_time_noalloc(T, K, 2, 252.0, fr, cov, st)
    assert ctx.launches == n0


def test_noalloc_rejects_oversized_problems():
    import torch
    from bmfm import _cabi
    ctx = _cabi.default_context()
    K, T = 400, 500      # (1+q)(K^2+1) doubles per chunk: too many chunks of shared-memory size for 16 MiB
    fr = torch.zeros((T, K), dtype=torch.float64, device="cuda")
    cov = torch.empty((1,), dtype=torch.float64, device="cuda")
    st = torch.empty(T, dtype=torch.int32, device="cuda")
    with pytest.raises(_cabi.BmfmError, match="fixed"):
        ctx.newey_west_by_time_noalloc(T, K, 2, 252.0, fr, cov, st)
