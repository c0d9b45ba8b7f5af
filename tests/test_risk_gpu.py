"""GPU parity of the risk-model adjustments (Newey-West, eigenfactor Monte-Carlo,
vol-regime) against outputs of the unmodified reference (tests/golden/*.npz;
eigen step: reference source with np.linalg.eig patched to the canonical order)
and against the CPU oracle on seeded inputs."""
import os

import numpy as np
import pytest

from util import record, relerr, relerr_cols, relerr_cov

pytestmark = pytest.mark.gpu
TOL = 1e-9


@pytest.fixture(scope="module")
def ops():
    from bmfm import ops
    return ops


@pytest.fixture(scope="module")
def risk(golden_dir):
    return np.load(os.path.join(golden_dir, "risk.npz"))


@pytest.fixture(scope="module")
def pipe(golden_dir):
    return np.load(os.path.join(golden_dir, "pipeline.npz"))


@pytest.mark.parametrize("q,tao", [(2, 10), (0, 252), (3, 40)])
def test_newey_west_golden(ops, risk, q, tao):
    fr = risk["nw_factor_ret"]
    want, ok = risk[f"nw_q{q}_tao{tao}"], risk[f"nw_q{q}_tao{tao}_ok"]
    cov, status = ops.newey_west_by_time(fr, q, tao)
    assert np.array_equal(status == 0, ok)
    assert (status[~ok] == 1).all()   # BMFM_ST_SHORT
    for t in np.nonzero(ok)[0]:
        assert record("nw.golden[per element / sqrt(diag diag)]", relerr_cov(cov[t], want[t])) < 1e-11, t
    assert np.isnan(cov[~ok]).all()


@pytest.mark.parametrize("T,K,q,tao", [(700, 42, 2, 252), (300, 12, 4, 30), (50, 3, 1, 5),
                                       (2500, 42, 2, 252), (3750, 42, 2, 252)])     # configs C2 and C3 / C4
def test_newey_west_oracle(ops, T, K, q, tao):
    """Chunked recurrence (several chunks, chunk boundaries crossing the lags) vs the direct sums."""
    from bmfm.synth import make_factor_returns
    from oracle import barra_oracle as orc
    fr = make_factor_returns(T, K, seed=T + K) + 0.003     # non-zero means
    cov, status = ops.newey_west_by_time(fr, q, tao)
    for t in list(range(0, T, 37)) + [T - 1, K, K + 1, max(q, K) - 1]:
        if t + 1 <= q or t + 1 <= K:
            assert status[t] == 1
        else:
            assert status[t] == 0
            assert record(f"nw.oracle.T{T}_K{K}[per element / sqrt(diag diag)]",
                          relerr_cov(cov[t], orc.newey_west(fr[:t + 1], q, tao))) < 1e-11, t


def test_vol_regime_golden(ops, pipe):
    fr, er, ok = pipe["factor_ret"], pipe["er"], pipe["er_ok"]
    status = np.where(ok, 0, 1).astype(np.int32)
    out, lamb = ops.vol_regime_adj_by_time(fr, np.nan_to_num(er), status, tao=12)
    assert relerr(lamb, pipe["lamb"]) < 1e-12
    assert np.array_equal(~np.isnan(out[:, 0, 0]), pipe["vr_ok"])
    assert relerr(out, pipe["vr"]) < 1e-12


def test_gauss_cov_legacy_stream(ops):
    """cov(z) of the seeded legacy numpy stream (MT19937 + polar): the device generator must
    reproduce np.random.seed(m+1); standard_normal to rounding of log/sqrt (<= few ulp)."""
    import torch
    from bmfm import _cabi
    ctx = _cabi.default_context()
    for (m0, nsim, T, K) in [(0, 5, 40, 6), (7, 3, 252, 42), (0, 2, 1001, 3)]:
        Cm = torch.empty((nsim, K, K), dtype=torch.float64, device="cuda")
        ctx.gauss_cov(m0, nsim, T, K, Cm)
        Cm = Cm.cpu().numpy()
        for j in range(nsim):
            z = np.random.RandomState(m0 + j + 1).standard_normal((T, K))
            assert relerr(Cm[j], np.cov(z.T)) < 1e-13


@pytest.mark.parametrize("T,M", [(40, 5), (252, 8)])
@pytest.mark.parametrize("solver", ["0", "1", "2", "3"])     # 0: register-tile solver (default), 1: block Jacobi, 2: fused QL, 3: split QL
def test_eigen_risk_adj_golden_canonical(ops, risk, T, M, solver, bmfm_option):
    bmfm_option("eig_mode", solver)
    cov = risk["eig_cov_in"]
    out, status = ops.eigen_risk_adj_by_time(cov[None], None, T, M, 1.4)
    assert status[0] == 0
    assert record(f"eigen.golden.T{T}_M{M}[per element / sqrt(diag diag)]",
                  relerr_cov(out[0], risk[f"eig_T{T}_M{M}_canonical"])) < TOL
    # the literal reference output depends on LAPACK's eigenvalue order:
    # report the gap, do not assert on it
    gap = relerr(out[0], risk[f"eig_T{T}_M{M}_literal"])
    print(f"literal-vs-canonical gap T={T} M={M}: {gap:.3e}")


@pytest.mark.parametrize("solver", ["0", "1", "2", "3"])
def test_eigen_by_time_oracle(ops, solver, bmfm_option):
    """K = 42 factor covariances from an expanding Newey-West run, simulation length 252."""
    bmfm_option("eig_mode", solver)
    from bmfm.synth import make_factor_returns
    from oracle import barra_oracle as orc
    K, T = 42, 140
    fr = make_factor_returns(T, K, seed=5)
    nw, nw_status = ops.newey_west_by_time(fr, 2, 60)
    sel = [60, 99, 139]
    out, status = ops.eigen_risk_adj_by_time(nw[sel], nw_status[sel], 252, 12, 1.4)
    for j, t in enumerate(sel):
        assert status[j] == 0
        want = orc.eigen_risk_adj(nw[t], 252, 12, 1.4)
        assert record("eigen.oracle.K42_M12[per element / sqrt(diag diag)]", relerr_cov(out[j], want)) < TOL, t
    # short windows stay invalid, non-PSD input is gated (utils.py:67-68)
    bad = -np.eye(K)[None]
    o2, s2 = ops.eigen_risk_adj_by_time(np.concatenate([np.nan_to_num(nw[:2]), bad]),
                                        np.array([1, 1, 0], dtype=np.int32), 50, 3, 1.4)
    assert list(s2) == [1, 1, 2] and np.isnan(o2).all()


@pytest.mark.parametrize("T,M", [(40, 5), (252, 8)])
def test_eigen_risk_adj_golden_literal_injected_order(ops, risk, T, M):
    """SURVEY.md 8(c) mode (iii): with the eigen order of numpy's general solver injected (D0 / U0 of F0 as
    np.linalg.eig returns them, and for every simulation the positions dgeev left the eigenpairs at), the GPU
    reproduces the UNPATCHED reference output."""
    cov = risk["eig_cov_in"]
    D0, U0, perm = ops.literal_eig_order(cov, T, M)
    out = ops.eigen_risk_adj_literal(cov[None], T, M, 1.4, D0[None], U0[None], perm[None])
    assert record(f"eigen.golden.literal.T{T}_M{M}[per element / sqrt(diag diag)]",
                  relerr_cov(out[0], risk[f"eig_T{T}_M{M}_literal"])) < TOL


def test_eigen_literal_injected_order_pipeline(ops, pipe):
    """The same on the pipeline panel: every valid date of the unpatched reference run (M = 6)."""
    nw, ok = pipe["nw"], pipe["er_literal_ok"].astype(bool)
    sel = np.nonzero(ok)[0]
    T_sim = len(pipe["factor_ret"])                   # MFM.py:119 passes self.T
    parts = [ops.literal_eig_order(nw[t], T_sim, 6) for t in sel]
    out = ops.eigen_risk_adj_literal(nw[sel], T_sim, 6, 1.4, np.stack([p[0] for p in parts]),
                                     np.stack([p[1] for p in parts]), np.stack([p[2] for p in parts]))
    assert record("eigen.pipeline.literal[per element / sqrt(diag diag)]", relerr_cov(out, pipe["er_literal"][sel])) < TOL


def test_eigen_retry_path_for_unconverged_problems(ops, bmfm_option):
    """A problem whose QL iteration does not converge is redone by the Jacobi kernel; the test hook marks
    every third problem as unconverged, the result must not change beyond rounding."""
    from bmfm.synth import make_factor_returns
    K, T = 42, 110
    fr = make_factor_returns(T, K, seed=29)
    nw, st = ops.newey_west_by_time(fr, 2, 60)
    want, s0 = ops.eigen_risk_adj_by_time(nw[-2:], st[-2:], 252, 20, 1.4)
    bmfm_option("eig_fail_every", 3)
    got, s1 = ops.eigen_risk_adj_by_time(nw[-2:], st[-2:], 252, 20, 1.4)
    assert (s0 == 0).all() and (s1 == 0).all()
    assert record("eigen.retry_vs_direct", relerr_cov(got, want)) < 1e-11


def test_eigen_config_size_simulation_count(ops):
    """Seeds far beyond the small cases (m = 1 .. 2000: the legacy stream of every simulation is its own
    MT19937 sequence) and a problem count that spans several batches of the solver: two dates x 2000
    simulations, K = 42, window 252, against the oracle."""
    from bmfm.synth import make_factor_returns
    from oracle import barra_oracle as orc
    K, T = 42, 120
    fr = make_factor_returns(T, K, seed=17)
    nw, nw_status = ops.newey_west_by_time(fr, 2, 60)
    sel = [80, 119]
    out, status = ops.eigen_risk_adj_by_time(nw[sel], nw_status[sel], 252, 2000, 1.4)
    assert (status == 0).all()
    for j, t in enumerate(sel):
        want = orc.eigen_risk_adj(nw[t], 252, 2000, 1.4)
        assert record("eigen.oracle.K42_M2000[per element / sqrt(diag diag)]", relerr_cov(out[j], want)) < TOL, t


def test_mfm_dropin_full_pipeline(pipe):
    """demo.py's four calls (demo.py:38-42) through the drop-in classes."""
    import contextlib
    import io
    from bmfm.synth import make_panel
    from mfm.MFM import MFM
    pan = make_panel(T=90, N=48, P=4, Q=3, seed=41, ragged=0.05, scale_log10=(-1, 1))
    df = pan.to_frame(shuffle_seed=3)
    with contextlib.redirect_stdout(io.StringIO()):
        model = MFM(df, pan.P, pan.Q)
        with pytest.raises(Exception):
            model.Newey_West_by_time()
        model.reg_by_time()
        with pytest.raises(Exception):
            model.eigen_risk_adj_by_time()
        nw = model.Newey_West_by_time(q=2, tao=30)
        with pytest.raises(Exception):
            model.vol_regime_adj_by_time()
        er = model.eigen_risk_adj_by_time(M=6, scale_coef=1.4)
        vr, lamb = model.vol_regime_adj_by_time(tao=12)
    assert [len(c) > 0 for c in nw] == list(pipe["nw_ok"])
    assert [len(c) > 0 for c in er] == list(pipe["er_ok"])
    assert [len(c) > 0 for c in vr] == list(pipe["vr_ok"])
    for t in range(pan.T):
        if pipe["nw_ok"][t]:
            assert record("pipeline.nw", relerr_cov(nw[t].values, pipe["nw"][t])) < TOL
            assert list(nw[t].columns) == list(model.factor_ret.columns)
        if pipe["er_ok"][t]:
            assert record("pipeline.er", relerr_cov(er[t].values, pipe["er"][t])) < TOL
            assert record("pipeline.vr", relerr_cov(vr[t].values, pipe["vr"][t])) < TOL
    assert relerr(np.array(lamb), pipe["lamb"]) < TOL
    assert isinstance(lamb, list) and isinstance(lamb[0], float)


def test_date_sharded_entry_points_match_full(ops):
    """bmfm_newey_west_by_time_range and the split vol-regime pass, run shard by shard on one GPU
    (three 'ranks' of dates), must reproduce the un-sharded calls."""
    import torch
    from bmfm import _cabi, dist as bd
    from bmfm.synth import make_factor_returns
    ctx = _cabi.default_context()
    T, K, q, tao, tao_vr = 500, 17, 2, 60.0, 21.0
    fr = make_factor_returns(T, K, seed=11)
    cov_full, st_full = ops.newey_west_by_time(fr, q, tao)
    vr_full, lamb_full = ops.vol_regime_adj_by_time(fr, np.nan_to_num(cov_full), st_full, tao_vr)
    d_fr = torch.from_numpy(fr).cuda()
    be = bd.GpuBackend(ctx)
    world = 3
    counts = bd.shard_counts(T, world)
    blocks, wbs, wvs = [], [], []
    for r in range(world):
        lo, hi = bd.shard_range(T, r, world)
        cov, st = be.newey_west_range(d_fr, lo, hi, q, tao)
        assert np.array_equal(st.cpu().numpy(), st_full[lo:hi])
        got = cov.cpu().numpy()
        ok = st_full[lo:hi] == 0
        assert np.array_equal(got[ok], cov_full[lo:hi][ok])       # same kernels, same chunks: bit-identical
        wb, wv = be.vr_bias(T, lo, d_fr[lo:hi].contiguous(), cov, st, tao_vr)
        blocks.append((lo, hi, cov, st))
        wbs.append(wb)
        wvs.append(wv)
    wb_all, wv_all = torch.cat(wbs), torch.cat(wvs)
    for (lo, hi, cov, st) in blocks:
        out, lamb = be.vr_scale(T, lo, wb_all, wv_all, cov, st)
        assert relerr(lamb.cpu().numpy(), lamb_full) < 1e-14
        assert relerr(out.cpu().numpy(), vr_full[lo:hi]) < 1e-14
    # the orchestration class with world = 1
    from bmfm.synth import make_panel
    from bmfm.panel import pack_arrays
    pan = make_panel(T=60, N=50, P=4, Q=3, seed=5)
    csr = pan.to_csr()
    (cap, ret, ind, style), pos = pack_arrays(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P)
    dv = lambda a: torch.from_numpy(a).cuda()
    res = bd.RiskPipeline(be).run([pan.T], pan.P, pan.Q, pan.N, dv(csr["date_off"]), dv(cap), dv(ret), dv(ind),
                                  dv(style), q=2, tao_nw=20, tao_vr=10, eigen=dict(T_sim=60, M=3, scale_coef=1.4))
    from oracle import barra_oracle as orc
    wfr, _, _ = orc.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P)
    nw = orc.newey_west_by_time(wfr, 2, 20)
    er = orc.eigen_risk_adj_by_time(nw, 60, 3, 1.4)
    vr, lamb = orc.vol_regime_adj_by_time(wfr, er, 10)
    assert relerr(res["factor_ret"].cpu().numpy(), wfr) < TOL
    assert relerr(res["lamb_all"].cpu().numpy(), np.array(lamb)) < TOL
    assert relerr(res["vr"].cpu().numpy()[-1], vr[-1]) < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("T,K,q,tao,world", [(500, 17, 2, 60.0, 3), (2000, 42, 2, 252.0, 8), (90, 6, 1, 20.0, 4),
                                             (300, 12, 0, 40.0, 2)])
def test_sharded_newey_west_exchanges_totals(ops, T, K, q, tao, world):
    """bmfm_newey_west_shard_totals / _emit: every 'rank' reduces only its own dates and the carry-in comes
    from the other ranks' end-of-range states.  Emulated on one GPU with one context per rank and a single
    shared totals buffer; must agree with the un-sharded pass (different chunking, so not bit-identical)."""
    import torch
    from bmfm import _cabi, dist as bd
    from bmfm.synth import make_factor_returns
    fr = make_factor_returns(T, K, seed=3 + T)
    cov_full, st_full = ops.newey_west_by_time(fr, q, tao)
    d_fr = torch.from_numpy(fr).cuda()
    ctxs = [_cabi.Context(0) for _ in range(world)]
    ns = ctxs[0].newey_west_state_doubles(K, q)
    totals = torch.full((world, ns), float("nan"), dtype=torch.float64, device="cuda")
    ranges = [bd.shard_range(T, r, world) for r in range(world)]
    for r, (lo, hi) in enumerate(ranges):
        ctxs[r].newey_west_shard_totals(T, K, q, tao, d_fr, lo, hi, r, [totals.data_ptr()])
    torch.cuda.synchronize()
    assert bool(torch.isfinite(totals).all())
    ends = [hi for (_, hi) in ranges]
    scale = np.nanmax(np.abs(cov_full))
    for r, (lo, hi) in enumerate(ranges):
        cov = torch.empty((hi - lo, K, K), dtype=torch.float64, device="cuda")
        st = torch.empty(hi - lo, dtype=torch.int32, device="cuda")
        ctxs[r].newey_west_shard_emit(T, K, q, tao, d_fr, lo, hi, r, ends[:r], totals, cov, st)
        torch.cuda.synchronize()
        assert np.array_equal(st.cpu().numpy(), st_full[lo:hi])
        ok = st_full[lo:hi] == 0
        got = cov.cpu().numpy()
        assert np.isnan(got[~ok]).all()
        assert np.max(np.abs(got[ok] - cov_full[lo:hi][ok])) < 1e-12 * scale


@pytest.mark.gpu
def test_eigenfactor_bias_stat(ops, golden_dir):
    """bmfm_eigenfactor_bias_stat against the reference's output on the pipeline panel (dates without a
    covariance are skipped as the reference's try/except does) and against the oracle on a longer series;
    then through the mfm.utils drop-in with a list of DataFrames."""
    import pandas as pd
    from oracle import barra_oracle as orc
    from bmfm.synth import make_factor_returns
    g = np.load(os.path.join(golden_dir, "pipeline.npz"))
    ok = g["nw_ok"][40:].astype(bool)
    st = np.where(ok, 0, 1).astype(np.int32)
    got = ops.eigenfactor_bias_stat(np.nan_to_num(g["nw"][40:]), st, g["factor_ret"][40:], predlen=5)
    assert record("bias_stat.golden", relerr_cols(got[None], g["bias_stat_nw"][None])) < TOL
    T, K = 400, 17
    fr = make_factor_returns(T, K, seed=23)
    cov, status = ops.newey_west_by_time(fr, 2, 90.0)
    covs = [cov[t] if status[t] == 0 else None for t in range(T)]
    for predlen in (1, 21):
        want = orc.eigenfactor_bias_stat(covs, fr, predlen=predlen)
        got = ops.eigenfactor_bias_stat(np.nan_to_num(cov), status, fr, predlen=predlen)
        assert record(f"bias_stat.oracle.predlen{predlen}", relerr_cols(got[None], want[None])) < TOL
    from mfm import utils as mu
    frames = [pd.DataFrame(cov[t]) if status[t] == 0 else pd.DataFrame() for t in range(T)]
    got = mu.eigenfactor_bias_stat(frames[100:], pd.DataFrame(fr[100:]), predlen=21)
    want = orc.eigenfactor_bias_stat(covs[100:], fr[100:], predlen=21)
    assert relerr_cols(got[None], want[None]) < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("K", [13, 17, 5])
def test_eigen_by_time_odd_factor_counts(ops, K):
    """Odd K: the shared-memory layouts of the split solver must keep their 16-byte pieces aligned."""
    from bmfm.synth import make_factor_returns
    from oracle import barra_oracle as orc
    T = 3 * K + 8
    fr = make_factor_returns(T, K, seed=K)
    cov, st = ops.newey_west_by_time(fr, 1, 30.0)
    got, est = ops.eigen_risk_adj_by_time(cov[-3:], st[-3:], 60, 5, 1.4)
    assert (est == 0).all()
    for i in range(3):
        assert relerr(got[i], orc.eigen_risk_adj(cov[T - 3 + i], 60, 5, 1.4)) < TOL


@pytest.mark.gpu
def test_newey_west_and_vol_regime_shape_sweep(ops):
    """Odd / even K, every supported lag count, series shorter than one chunk and not a multiple of it."""
    from bmfm.synth import make_factor_returns
    from oracle import barra_oracle as orc
    for (T, K, q, tao) in [(9, 3, 0, 5.0), (30, 5, 4, 12.0), (77, 13, 3, 40.0), (150, 8, 1, 60.0), (301, 21, 2, 90.0),
                           (1000, 7, 2, 252.0)]:
        fr = make_factor_returns(T, K, seed=T + K) + 0.002
        cov, st = ops.newey_west_by_time(fr, q, tao)
        want = orc.newey_west_by_time(fr, q, tao)
        for t in list(range(0, T, max(1, T // 25))) + [T - 1]:
            if want[t] is None:
                assert st[t] != 0, (T, K, q, t)
            else:
                assert st[t] == 0 and relerr(cov[t], want[t]) < TOL, (T, K, q, t)
        vr, lamb = ops.vol_regime_adj_by_time(fr, np.nan_to_num(cov), st, 21.0)
        wv, wl = orc.vol_regime_adj_by_time(fr, want, 21.0)
        assert relerr(np.asarray(lamb), np.asarray(wl)) < TOL, (T, K, q)
        for t in (T - 1, T // 2):
            if wv[t] is not None:
                assert relerr(vr[t], wv[t]) < TOL, (T, K, q, t)


@pytest.mark.gpu
def test_backend_record_exchange_path(ops):
    """GpuBackend.nw_shard_total / nw_shard_emit (what RiskPipeline.run(shard_nw=True) drives over NCCL or gloo),
    three emulated ranks with a context each, against the un-sharded pass."""
    import torch
    from bmfm import _cabi, dist as bd
    from bmfm.synth import make_factor_returns
    T, K, q, tao, world = 420, 11, 2, 50.0, 3
    fr = make_factor_returns(T, K, seed=9)
    cov_full, st_full = ops.newey_west_by_time(fr, q, tao)
    d_fr = torch.from_numpy(fr).cuda()
    bes = [bd.GpuBackend(_cabi.Context(0)) for _ in range(world)]
    ranges = [bd.shard_range(T, r, world) for r in range(world)]
    ends = [hi for (_, hi) in ranges]
    totals = torch.stack([bes[r].nw_shard_total(d_fr, lo, hi, q, tao, r, world) for r, (lo, hi) in enumerate(ranges)])
    scale = np.nanmax(np.abs(cov_full))
    for r, (lo, hi) in enumerate(ranges):
        cov, st = bes[r].nw_shard_emit(d_fr, lo, hi, q, tao, r, ends, totals)
        assert np.array_equal(st.cpu().numpy(), st_full[lo:hi])
        ok = st_full[lo:hi] == 0
        assert np.max(np.abs(cov.cpu().numpy()[ok] - cov_full[lo:hi][ok])) < 1e-12 * scale


@pytest.mark.gpu
@pytest.mark.parametrize("T,K,world", [(500, 17, 3), (2000, 42, 8), (90, 6, 4)])
def test_sharded_vol_regime_exchanges_two_totals(ops, T, K, world):
    """bmfm_vol_regime_shard_totals / _scale: every 'rank' scans only its own dates, the carry-in is the sum of the
    earlier ranks' (sum wb, sum wv) pairs.  Emulated on one GPU with a single shared record buffer; must agree with
    the un-sharded pass (a different association of the prefix sums: rounding only)."""
    import torch
    from bmfm import _cabi, dist as bd
    from bmfm.synth import make_factor_returns
    ctx = _cabi.default_context()
    fr = make_factor_returns(T, K, seed=T)
    cov, st = ops.newey_west_by_time(fr, 2, 60.0)
    want, lamb_want = ops.vol_regime_adj_by_time(fr, np.nan_to_num(cov), st, 21.0)
    dv = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_fr, d_cov, d_st = dv(fr), dv(np.nan_to_num(cov)), dv(st.astype(np.int32))
    totals = torch.full((world, 2), float("nan"), dtype=torch.float64, device="cuda")
    ranges = [bd.shard_range(T, r, world) for r in range(world)]
    terms = []
    for r, (lo, hi) in enumerate(ranges):
        wb = torch.empty(hi - lo, dtype=torch.float64, device="cuda")
        wv = torch.empty(hi - lo, dtype=torch.float64, device="cuda")
        ctx.vol_regime_shard_totals(T, lo, hi - lo, K, 21.0, d_fr[lo:hi].contiguous(), d_cov[lo:hi].contiguous(),
                                    d_st[lo:hi].contiguous(), wb, wv, r, [totals.data_ptr()])
        terms.append((wb, wv))
    torch.cuda.synchronize()
    assert bool(torch.isfinite(totals).all())
    for r, (lo, hi) in enumerate(ranges):
        out = torch.empty((hi - lo, K, K), dtype=torch.float64, device="cuda")
        lamb = torch.empty(hi - lo, dtype=torch.float64, device="cuda")
        ctx.vol_regime_shard_scale(hi - lo, K, terms[r][0], terms[r][1], totals, r, d_cov[lo:hi].contiguous(),
                                   d_st[lo:hi].contiguous(), out, lamb)
        assert relerr(lamb.cpu().numpy(), lamb_want[lo:hi]) < 1e-13
        assert relerr(out.cpu().numpy(), want[lo:hi]) < 1e-13


@pytest.mark.gpu
def test_sharded_exchanges_through_per_producer_flags(ops):
    """bmfm_ctx_set_peer_flags: with the ranks' flag blocks registered, the sharded Newey-West and vol-regime calls
    publish their records' epochs to the later ranks and wait for the earlier ones inside the kernels (no caller-side
    barrier).  Three emulated ranks with a context and a flag block each on one GPU, two steps in a row (epochs 1
    and 2); results must equal the barrier form's."""
    import torch
    from bmfm import _cabi, dist as bd
    from bmfm.synth import make_factor_returns
    T, K, q, tao, world = 420, 11, 2, 50.0, 3
    fr = make_factor_returns(T, K, seed=19)
    cov_full, st_full = ops.newey_west_by_time(fr, q, tao)
    vr_full, lamb_full = ops.vol_regime_adj_by_time(fr, np.nan_to_num(cov_full), st_full, 21.0)
    d_fr = torch.from_numpy(fr).cuda()
    ctxs = [_cabi.Context(0) for _ in range(world)]
    flags = [torch.zeros(64, dtype=torch.int64, device="cuda") for _ in range(world)]
    for r in range(world):
        ctxs[r].set_peer_flags(r, [f.data_ptr() for f in flags])
    ns = ctxs[0].newey_west_state_doubles(K, q)
    nw_tot = torch.zeros((world, ns), dtype=torch.float64, device="cuda")
    vr_tot = torch.zeros((world, 2), dtype=torch.float64, device="cuda")
    ranges = [bd.shard_range(T, r, world) for r in range(world)]
    ends = [hi for (_, hi) in ranges]
    scale = np.nanmax(np.abs(cov_full))
    for step in (1, 2):
        covs = []
        for r, (lo, hi) in enumerate(ranges):          # producers first: on ONE GPU a consumer must not spin before them
            ctxs[r].newey_west_shard_totals(T, K, q, tao, d_fr, lo, hi, r, [nw_tot.data_ptr()] * world)
        for r, (lo, hi) in enumerate(ranges):
            cov = torch.empty((hi - lo, K, K), dtype=torch.float64, device="cuda")
            st = torch.empty(hi - lo, dtype=torch.int32, device="cuda")
            ctxs[r].newey_west_shard_emit(T, K, q, tao, d_fr, lo, hi, r, ends[:r], nw_tot, cov, st)
            covs.append((cov, st))
        terms = []
        for r, (lo, hi) in enumerate(ranges):
            wb = torch.empty(hi - lo, dtype=torch.float64, device="cuda")
            wv = torch.empty(hi - lo, dtype=torch.float64, device="cuda")
            ctxs[r].vol_regime_shard_totals(T, lo, hi - lo, K, 21.0, d_fr[lo:hi].contiguous(), torch.nan_to_num(covs[r][0]),
                                            covs[r][1], wb, wv, r, [vr_tot.data_ptr()] * world)
            terms.append((wb, wv))
        for r, (lo, hi) in enumerate(ranges):
            out = torch.empty((hi - lo, K, K), dtype=torch.float64, device="cuda")
            lamb = torch.empty(hi - lo, dtype=torch.float64, device="cuda")
            ctxs[r].vol_regime_shard_scale(hi - lo, K, terms[r][0], terms[r][1], vr_tot, r, torch.nan_to_num(covs[r][0]),
                                           covs[r][1], out, lamb)
            torch.cuda.synchronize()
            ok = st_full[lo:hi] == 0
            assert np.array_equal(covs[r][1].cpu().numpy(), st_full[lo:hi])
            assert np.max(np.abs(covs[r][0].cpu().numpy()[ok] - cov_full[lo:hi][ok])) < 1e-12 * scale
            assert relerr(lamb.cpu().numpy(), lamb_full[lo:hi]) < 1e-13
        for r in range(world):
            f = flags[r].cpu().numpy()
            assert f[48] == 0                                           # no wait gave up
            assert f[32] == step and f[33] == step                      # own epochs on both channels
            assert all(f[s] == step for s in range(r)) and all(f[8 + s] == step for s in range(r))
    for c in ctxs:
        c.close()


@pytest.mark.gpu
def test_vol_regime_negative_variance_poisons_later_dates(ops):
    """A date whose mean f^2 / var is negative (a negative variance on the diagonal): the reference's sqrt makes its
    bias NaN and every later lamb with it (MFM.py:149,160); dates before it are untouched."""
    from bmfm.synth import make_factor_returns
    from oracle import barra_oracle as orc
    T, K = 40, 5
    fr = make_factor_returns(T, K, seed=2)
    cov = np.stack([np.cov(fr.T) for _ in range(T)])
    cov[25, 0, 0] = -1e-12                       # f^2 / var hugely negative for factor 0 on date 25
    out, lamb = ops.vol_regime_adj_by_time(fr, cov, np.zeros(T, dtype=np.int32), 9.0)
    wv, wl = orc.vol_regime_adj_by_time(fr, [cov[t] for t in range(T)], 9.0)
    assert np.isnan(lamb[25:]).all() and np.isfinite(lamb[:25]).all()
    assert relerr(lamb, np.array(wl)) < TOL
    assert relerr(out[10], wv[10]) < TOL and np.isnan(out[30]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("T,K,cuts", [(97, 6, (0, 1, 40, 41, 97)), (300, 42, (0, 128, 300)), (50, 5, (0, 50))])
def test_vol_regime_range_blocks_equal_whole_series(ops, T, K, cuts):
    """bmfm_vol_regime_adj_range over consecutive blocks of dates (what the pipelined host step calls) gives the rows
    of the whole-series call; one date without a covariance in the middle."""
    import torch
    from bmfm import _cabi
    from bmfm.synth import make_factor_returns
    from oracle import barra_oracle as orc
    fr = make_factor_returns(T, K, seed=11)
    covs, st = ops.newey_west_by_time(fr, 1, 30.0)
    st = st.copy()
    st[T // 2] = 1                                     # an empty DataFrame in the reference
    want_cov, want_l = orc.vol_regime_adj_by_time(fr, [covs[t] if st[t] == 0 else None for t in range(T)], 12.0)
    ctx = _cabi.default_context()
    dv = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_fr, d_cov, d_st = dv(fr), dv(covs), dv(st.astype(np.int32))
    wb = torch.full((T,), float("nan"), dtype=torch.float64, device="cuda")     # nothing has to be cleared
    wv = torch.full((T,), float("nan"), dtype=torch.float64, device="cuda")
    out = torch.empty_like(d_cov)
    lamb = torch.empty(T, dtype=torch.float64, device="cuda")
    for t0, t1 in zip(cuts[:-1], cuts[1:]):
        ctx.vol_regime_adj_range(T, t0, t1, K, 12.0, d_fr, d_cov, d_st, wb, wv, out, lamb)
    torch.cuda.synchronize()
    whole_cov, whole_l = ops.vol_regime_adj_by_time(fr, covs, st.astype(np.int32), 12.0)
    got_l, got = lamb.cpu().numpy(), out.cpu().numpy()
    assert record(f"vol_regime.range.lamb.T{T}", relerr(got_l, np.array(want_l))) < TOL
    assert relerr(got_l, whole_l) < 1e-14
    for t in range(T):
        if st[t] == 0 and np.isfinite(covs[t]).all():
            assert relerr(got[t], want_cov[t]) < TOL
        else:
            assert np.isnan(got[t]).all()
