"""GPU parity of the regression kernel (through the C-ABI) against
(1) outputs of the unmodified reference (tests/golden/reg.npz) and
(2) the CPU oracle on seeded panels, plus size-independent properties."""
import os

import numpy as np
import pytest

from util import record, relerr, relerr_cols

pytestmark = pytest.mark.gpu

TOL = 1e-9   # north_star: <= 1e-9 relative, fp64


@pytest.fixture(scope="module")
def ops():
    from bmfm import ops
    return ops


@pytest.fixture(scope="module")
def reg_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "reg.npz"))


def _case(reg, case):
    return {k.split("__", 1)[1]: reg[k] for k in reg.files if k.startswith(case + "__")}


@pytest.mark.parametrize("case", ["small", "csi", "nostyle_scale"])
@pytest.mark.parametrize("grouped", [True, False])
@pytest.mark.parametrize("mode", ["1", "2", "3", "4", "5", "6"])     # 1: CTA-per-date kernel, 2: warp-per-date streaming kernel, 3: warp-per-date FP64-MMA kernel
def test_golden_reference_outputs(ops, reg_golden, case, grouped, mode, bmfm_option):
    bmfm_option("reg_mode", mode)
    g = _case(reg_golden, case)
    fr, spec, r2, status, _ = ops.reg_by_time(g["date_off"], g["cap"], g["ret"], g["ind"], g["style"],
                                              int(g["P"]), group_by_industry=grouped)
    assert record(f"reg.golden.{case}.factor_ret[per column]", relerr_cols(fr, g["factor_ret"])) < TOL
    assert record(f"reg.golden.{case}.specific_ret", relerr(spec, g["specific_ret"], scale=np.max(np.abs(g["ret"])))) < TOL
    assert record(f"reg.golden.{case}.R2", np.max(np.abs(r2 - g["R2"]))) < TOL


def _numpy_sensitivity(g, P):
    """How far numpy's own answer moves when the rows of each date are permuted (same problem, different
    summation order): the yardstick for rank-deficient dates, where pinv cuts at 1e-15 of the largest
    singular value and the answer is not determined to 1e-9 by the inputs."""
    from oracle import barra_oracle as orc
    off = g["date_off"]
    rng = np.random.default_rng(0)
    perm = np.concatenate([off[t] + rng.permutation(int(off[t + 1] - off[t])) for t in range(len(off) - 1)])
    fr, spec, r2 = orc.reg_by_time(off, g["cap"], g["ret"], g["ind"], g["style"], P)
    fr2, spec2, r22 = orc.reg_by_time(off, g["cap"][perm], g["ret"][perm], g["ind"][perm], g["style"][perm], P)
    return relerr_cols(fr2, fr), relerr(spec2, spec[perm], scale=np.max(np.abs(g["ret"])))


def test_golden_rank_deficient_pinv(ops, reg_golden):
    """N < K: the reference's pinv picks the minimum-norm solution (R2 = 1)."""
    g = _case(reg_golden, "n_lt_k")
    fr, spec, r2, status, _ = ops.reg_by_time(g["date_off"], g["cap"], g["ret"], g["ind"], g["style"], int(g["P"]))
    assert (status == 4).all()          # BMFM_ST_PINV: the eigen/pinv path was taken
    # N < K: the Gram has N - K zero singular values and numpy's own answer moves by `sens` under a row
    # permutation; the kernel is held to the north-star bound or ten times that movement, whichever is larger
    sens_fr, sens_spec = _numpy_sensitivity(g, int(g["P"]))
    record("reg.golden.n_lt_k.numpy_row_permutation_sensitivity[per column]", sens_fr)
    assert record("reg.golden.n_lt_k.factor_ret[per column]", relerr_cols(fr, g["factor_ret"])) < max(TOL, 10 * sens_fr)
    assert record("reg.golden.n_lt_k.specific_ret", relerr(spec, g["specific_ret"], scale=np.max(np.abs(g["ret"])))) < max(TOL, 10 * sens_spec)
    assert record("reg.golden.n_lt_k.R2", np.max(np.abs(r2 - g["R2"]))) < TOL


@pytest.mark.parametrize("case", ["small", "csi"])
def test_golden_exposure(ops, reg_golden, case):
    g = _case(reg_golden, case)
    a, b = int(g["date_off"][0]), int(g["date_off"][1])
    off = np.array([0, b - a], dtype=np.int64)
    fr, spec, r2, status, expo = ops.reg_by_time(off, g["cap"][a:b], g["ret"][a:b], g["ind"][a:b],
                                                 g["style"][a:b], int(g["P"]), want_exposure=True)
    assert record(f"reg.golden.{case}.exposure", relerr(expo[0], g["exposure0"])) < TOL
    assert relerr(fr[0], g["factor_ret"][0]) < TOL


@pytest.mark.parametrize("T,N,P,Q,ragged,scale", [
    (12, 300, 31, 10, 0.05, (-2, 2)),     # CSI300-like, un-standardised column scales
    (6, 800, 31, 10, 0.0, (-1, 1)),       # CSI800 cross-sections (config C2)
    (5, 97, 7, 3, 0.2, (-1, 1)),
    (4, 64, 0, 5, 0.0, (-1, 1)),          # P = 0 branch (CrossSection.py:95-98)
    (3, 70, 63, 12, 0.0, (0, 0)),         # compiled limits; N < K -> pinv path
    (2, 5000, 31, 10, 0.02, (-2, 2)),     # full A-share cross-section (config C3): multi-tile path
])
@pytest.mark.parametrize("mode", ["0", "1", "2", "3", "4", "5", "6"])   # auto / CTA-per-date / warp-per-date streaming / warp-per-date FP64-MMA
def test_oracle_parity(ops, T, N, P, Q, ragged, scale, mode, bmfm_option):
    bmfm_option("reg_mode", mode)
    from bmfm.synth import make_panel
    from oracle import barra_oracle as orc
    pan = make_panel(T, N, P, Q, seed=100 + N, ragged=ragged, scale_log10=scale)
    csr = pan.to_csr()
    ind = csr["ind"].copy()
    if P > 0:
        ind[5::37] = -1            # some rows with no industry column set
    want_fr, want_spec, want_r2 = orc.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], ind, csr["style"], P)
    fr, spec, r2, status, _ = ops.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], ind, csr["style"], P)
    tol_fr = tol_spec = TOL
    if N < 1 + P + Q:      # rank deficient: see test_golden_rank_deficient_pinv
        sens_fr, sens_spec = _numpy_sensitivity(dict(csr, ind=ind), P)
        tol_fr, tol_spec = max(TOL, 10 * sens_fr), max(TOL, 10 * sens_spec)
    tag = f"reg.oracle.T{T}_N{N}_P{P}_Q{Q}"
    assert record(tag + ".factor_ret[per column]", relerr_cols(fr, want_fr)) < tol_fr
    assert record(tag + ".specific_ret", relerr(spec, want_spec, scale=np.max(np.abs(csr["ret"])))) < tol_spec
    assert record(tag + ".R2", np.max(np.abs(r2 - want_r2))) < TOL


@pytest.mark.parametrize("T,N,P,Q", [
    (6, 800, 31, 10),     # ~20 industries in every 32-row chunk: the match.any group path
    (5, 400, 5, 9),       # at most 6 industries per chunk, rows interleaved: indicator-column products
    (4, 1500, 12, 7),     # two warps per date
    (3, 333, 40, 8),      # more than 32 industries
])
@pytest.mark.parametrize("mode", ["3", "4", "5", "6"])
def test_oracle_parity_rows_not_grouped(ops, T, N, P, Q, mode, bmfm_option):
    """The C-ABI accepts any row order inside a date; the packer's grouping by industry is an optimisation only."""
    bmfm_option("reg_mode", mode)
    from bmfm.synth import make_panel
    from oracle import barra_oracle as orc
    pan = make_panel(T, N, P, Q, seed=7 + N, ragged=0.03, scale_log10=(-1, 1))
    csr = pan.to_csr()
    want_fr, want_spec, want_r2 = orc.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], P)
    fr, spec, r2, status, _ = ops.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], P,
                                              group_by_industry=False)
    assert (status == 0).all()
    assert record("reg.oracle.ungrouped.factor_ret[per column]", relerr_cols(fr, want_fr)) < TOL
    assert relerr(spec, want_spec, scale=np.max(np.abs(csr["ret"]))) < TOL
    assert np.max(np.abs(r2 - want_r2)) < TOL


def test_host_buffer_entry_point(ops):
    """bmfm_reg_by_time_host: host pointers in/out (the e2e path of bench.py)."""
    from bmfm import _cabi
    from bmfm.synth import make_panel
    pan = make_panel(5, 200, 31, 10, seed=7)
    csr = pan.to_csr()
    T, K = pan.T, 1 + pan.P + pan.Q
    fr = np.empty((T, K)); spec = np.empty(len(csr["cap"])); r2 = np.empty(T)
    st = np.empty(T, dtype=np.int32)
    ctx = _cabi.default_context()
    ctx.reg_by_time_host(T, pan.P, pan.Q, csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"],
                         fr, spec, r2, st)
    fr2, spec2, r22, _, _ = ops.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P,
                                            group_by_industry=False)
    # same kernels, same order of rows; the ungrouped path sums with shared-memory atomics, so compare to rounding
    assert relerr(fr, fr2) < 1e-12 and relerr(spec, spec2) < 1e-12 and np.max(np.abs(r2 - r22)) < 1e-12


def test_host_buffer_whole_step(ops):
    """bmfm_risk_step_host (pinned host buffers, copies in blocks of dates overlapped with the kernels) against
    the three device-pointer calls it chains; run twice to cover the reuse of its staging buffers."""
    import torch
    from bmfm import _cabi
    from bmfm.synth import make_panel
    pan = make_panel(40, 200, 31, 10, seed=17, ragged=0.04)
    csr = pan.to_csr()
    T, K = pan.T, 1 + pan.P + pan.Q
    rows = len(csr["cap"])
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    h = {k: pin(v) for k, v in csr.items()}
    out = dict(fr=torch.empty((T, K), dtype=torch.float64).pin_memory(), spec=torch.empty(rows, dtype=torch.float64).pin_memory(),
               r2=torch.empty(T, dtype=torch.float64).pin_memory(), st=torch.empty(T, dtype=torch.int32).pin_memory(),
               nw=torch.empty((T, K, K), dtype=torch.float64).pin_memory(), nwst=torch.empty(T, dtype=torch.int32).pin_memory(),
               vr=torch.empty((T, K, K), dtype=torch.float64).pin_memory(), lamb=torch.empty(T, dtype=torch.float64).pin_memory())
    ctx = _cabi.default_context()
    fr, spec, r2, status, _ = ops.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P,
                                              group_by_industry=False)
    nw, nwst = ops.newey_west_by_time(fr, 2, 10.0)
    vr, lamb = ops.vol_regime_adj_by_time(fr, np.nan_to_num(nw), nwst, 7.0)
    for _ in range(2):
        for v in out.values():
            v.fill_(-7)
        ctx.risk_step_host(T, pan.P, pan.Q, h["date_off"], h["cap"], h["ret"], h["ind"], h["style"], 2, 10.0, 7.0,
                           out["fr"], out["spec"], out["r2"], out["st"], out["nw"], out["nwst"], out["vr"], out["lamb"])
        assert np.array_equal(out["st"].numpy(), status) and np.array_equal(out["nwst"].numpy(), nwst)
        assert relerr(out["fr"].numpy(), fr) < 1e-12 and relerr(out["spec"].numpy(), spec) < 1e-12
        assert relerr(out["r2"].numpy(), r2) < 1e-12
        assert relerr(out["nw"].numpy(), nw) < 1e-11 and relerr(out["vr"].numpy(), vr) < 1e-11
        assert relerr(out["lamb"].numpy(), lamb) < 1e-11


def test_last_industry_empty_raises(ops):
    """-cap/0 in the constraint row (CrossSection.py:70): numpy's pinv raises LinAlgError."""
    from bmfm.synth import make_panel
    pan = make_panel(2, 60, 5, 3, seed=3)
    csr = pan.to_csr()
    ind = csr["ind"].copy()
    ind[ind == 4] = 0
    with pytest.raises(np.linalg.LinAlgError):
        ops.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], ind, csr["style"], pan.P)


def test_properties_full_size(ops):
    """Config-C3-sized cross-sections: properties that need no oracle run.
    (a) the cap-weighted industry returns sum to zero (the constraint of CrossSection.py:66-71);
    (b) the sqrt-cap weighted residual is orthogonal to country and every style column;
    (c) regressing again on the residuals returns zero factor returns (idempotence)."""
    from bmfm.synth import make_panel
    pan = make_panel(8, 5000, 31, 10, seed=20240003, scale_log10=(-1, 1))
    csr = pan.to_csr()
    P, Q = pan.P, pan.Q
    fr, spec, r2, status, _ = ops.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], P)
    assert (status == 0).all()
    off = csr["date_off"]
    for t in range(pan.T):
        a, b = off[t], off[t + 1]
        cap = csr["cap"][a:b]
        indcap = np.bincount(csr["ind"][a:b], cap, P)
        assert abs(indcap @ fr[t, 1:1 + P]) < 1e-12 * indcap.sum() * np.abs(fr[t]).max()
        w = np.sqrt(cap)
        e = spec[a:b]
        scale = (w * np.abs(csr["ret"][a:b])).sum()
        assert abs((w * e).sum()) < 1e-10 * scale
        s = csr["style"][a:b]
        sc = s - s.mean(0)
        assert np.max(np.abs((w * e) @ sc) / (np.abs(sc).max(0) * scale)) < 1e-9
    fr2, spec2, _, _, _ = ops.reg_by_time(off, csr["cap"], spec, csr["ind"], csr["style"], P)
    assert np.max(np.abs(fr2)) < 1e-9 * np.max(np.abs(fr))
    assert relerr(spec2, spec) < 1e-9


def test_mfm_dropin_reg_by_time(golden_dir):
    """The drop-in MFM class on the DataFrame demo.py builds (MFM.py:28-76)."""
    import contextlib, io
    from bmfm.synth import make_panel
    from mfm.MFM import MFM
    g = np.load(os.path.join(golden_dir, "pipeline.npz"))
    pan = make_panel(T=90, N=48, P=4, Q=3, seed=41, ragged=0.05, scale_log10=(-1, 1))
    df = pan.to_frame(shuffle_seed=3)
    with contextlib.redirect_stdout(io.StringIO()):
        model = MFM(df, pan.P, pan.Q)
        fr, spec, r2 = model.reg_by_time()
    assert list(fr.columns) == ['country'] + pan.industry_names + pan.style_names
    assert relerr_cols(fr.values, g["factor_ret"]) < TOL
    assert relerr(np.concatenate([s.values[0] for s in spec]), g["specific_ret"]) < TOL
    assert np.max(np.abs(r2.values[:, 0] - g["R2"])) < TOL
    assert relerr(np.asarray(model.last_capital), g["last_capital"]) == 0.0
    assert list(spec[0].columns) == sorted(spec[0].columns)


def test_device_packer_matches_host_packer():
    """pack_frame_device (row work on the GPU) against pack_frame (host numpy / pandas) on a shuffled ragged panel
    with rows that have no industry: same storage order, same arrays bit for bit, same name mapping; the checks on the
    one-hot block raise the same errors; a P that does not fit the table raises ValueError as MFM.py:70 does."""
    import contextlib
    import io
    from bmfm.panel import pack_frame, pack_frame_device
    from bmfm.synth import make_panel
    from mfm.MFM import MFM
    pan = make_panel(T=25, N=90, P=6, Q=4, seed=77, ragged=0.1)
    df = pan.to_frame(shuffle_seed=9)
    df.loc[df.index[::13], pan.industry_names] = 0               # rows with no industry column set
    h = pack_frame(df, pan.P, pan.Q)
    d = pack_frame_device(df, pan.P, pan.Q)
    assert (d.T, d.rows, d.max_rows) == (h.T, h.rows, h.max_rows)
    assert np.array_equal(d.date_off_host, h.date_off) and d.sorted_dates.equals(h.sorted_dates)
    for name in ("cap", "ret", "ind", "style"):
        assert np.array_equal(getattr(d, name).cpu().numpy(), getattr(h, name)), name
    assert np.array_equal(d.name_pos.cpu().numpy(), h.name_pos)
    assert np.array_equal(d.name_uniques[d.name_code[d.ref_order].cpu().numpy()], h.names_sorted)
    assert np.array_equal(d.last_capital, h.last_capital)
    bad = df.copy()
    bad.iloc[3, 5] = 2
    with pytest.raises(NotImplementedError, match="0/1"):
        pack_frame_device(bad, pan.P, pan.Q)
    bad = df.copy()
    bad.iloc[3, 4:4 + pan.P] = 1
    with pytest.raises(NotImplementedError, match="more than one"):
        pack_frame_device(bad, pan.P, pan.Q)
    with contextlib.redirect_stdout(io.StringIO()), pytest.raises(ValueError):
        MFM(df, pan.P - 1, pan.Q).reg_by_time()


def test_lazy_specific_return_list(golden_dir):
    """reg_by_time's list of one-row DataFrames is built on demand: it must behave like the reference's list
    (len, truth value, indexing incl. negative and slices, iteration, pd.concat as demo.py:78 uses it)."""
    import contextlib
    import io
    import pandas as pd
    from bmfm.synth import make_panel
    from mfm.MFM import MFM
    g = np.load(os.path.join(golden_dir, "pipeline.npz"))
    pan = make_panel(T=90, N=48, P=4, Q=3, seed=41, ragged=0.05, scale_log10=(-1, 1))
    with contextlib.redirect_stdout(io.StringIO()):
        fr, spec, r2 = MFM(pan.to_frame(shuffle_seed=3), pan.P, pan.Q).reg_by_time()
    assert isinstance(spec, list) and len(spec) == 90 and bool(spec)
    last = spec[-1]
    assert isinstance(last, pd.DataFrame) and last.shape[0] == 1 and last.index[0] == fr.index[-1]
    assert spec[89] is last and spec[-1] is last                    # built once
    assert [f.shape[1] for f in spec[3:6]] == [int(n) for n in np.diff(g["date_off"])[3:6]]
    off = g["date_off"]
    assert relerr(spec[7].values[0], g["specific_ret"][off[7]:off[8]]) < TOL
    allf = pd.concat(spec)
    assert allf.shape[0] == 90 and list(allf.index) == list(fr.index)
    flat = np.concatenate([f.values[0] for f in spec])
    assert relerr(flat, g["specific_ret"]) < TOL
    assert sum(1 for _ in reversed(spec)) == 90


def test_fused_gather_entry_points_single_gpu(ops):
    """bmfm_reg_by_time_gather / bmfm_vol_regime_bias_gather with this GPU as its own (only) peer: the
    kernels' peer stores must land in the gathered buffers at the rank offset."""
    import torch
    from bmfm import _cabi
    from bmfm.panel import pack_arrays
    from bmfm.synth import make_panel
    ctx = _cabi.default_context()
    pan = make_panel(7, 120, 6, 4, seed=21)
    csr = pan.to_csr()
    (cap, ret, ind, style), pos = pack_arrays(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P)
    dv = lambda a: torch.from_numpy(a).cuda()
    T, K, row0, T_all = pan.T, 1 + pan.P + pan.Q, 5, 20
    off, dcap, dret, dind, dsty = dv(csr["date_off"]), dv(cap), dv(ret), dv(ind), dv(style)
    fr = torch.empty((T, K), dtype=torch.float64, device="cuda")
    fr2 = torch.empty_like(fr)
    spec = torch.empty(len(cap), dtype=torch.float64, device="cuda")
    r2 = torch.empty(T, dtype=torch.float64, device="cuda")
    st = torch.empty(T, dtype=torch.int32, device="cuda")
    gathered = torch.full((T_all, K), -7.0, dtype=torch.float64, device="cuda")
    ctx.reg_by_time(T, pan.P, pan.Q, pan.N, off, dcap, dret, dind, dsty, fr, spec, r2, st)
    ctx.reg_by_time_gather(T, pan.P, pan.Q, pan.N, off, dcap, dret, dind, dsty, fr2, spec, r2, st,
                           [gathered.data_ptr()], row0)
    torch.cuda.synchronize()
    assert torch.equal(fr, fr2) and torch.equal(gathered[row0:row0 + T], fr)
    assert (gathered[:row0] == -7.0).all() and (gathered[row0 + T:] == -7.0).all()
    cov = torch.eye(K, dtype=torch.float64, device="cuda").repeat(T, 1, 1) * 1e-4
    cst = torch.zeros(T, dtype=torch.int32, device="cuda")
    wb = torch.empty(T, dtype=torch.float64, device="cuda")
    wv = torch.empty(T, dtype=torch.float64, device="cuda")
    ctx.vol_regime_bias(T_all, row0, T, K, 10.0, fr, cov, cst, wb, wv)
    wba = torch.zeros(T_all, dtype=torch.float64, device="cuda")
    wva = torch.zeros(T_all, dtype=torch.float64, device="cuda")
    ctx.vol_regime_bias_gather(T_all, row0, T, K, 10.0, fr, cov, cst, [wba.data_ptr()], [wva.data_ptr()])
    torch.cuda.synchronize()
    assert torch.equal(wba[row0:row0 + T], wb) and torch.equal(wva[row0:row0 + T], wv)
    assert wba.sum() == wb.sum()


@pytest.mark.parametrize("mode", ["3", "4", "5", "6"])
def test_oracle_parity_shape_sweep(ops, mode, bmfm_option):
    """Odd and even style counts on both sides of the one-tile / two-tile split of the MMA kernel, industry
    counts around the 32-lane boundary, cross-sections shorter than a chunk, ragged dates (unaligned starts)."""
    bmfm_option("reg_mode", mode)
    from bmfm.synth import make_panel
    from oracle import barra_oracle as orc
    shapes = [(3, 45, 1, 1), (3, 33, 2, 2), (2, 130, 31, 6), (2, 129, 32, 7), (2, 200, 33, 9), (2, 257, 5, 11),
              (2, 90, 3, 4), (2, 64, 12, 8), (4, 31, 2, 0), (2, 300, 50, 10)]
    for (T, N, P, Q) in shapes:
        for grouped in (True, False):
            pan = make_panel(T, N, P, Q, seed=N + P + Q, ragged=0.07, scale_log10=(-1, 1))
            csr = pan.to_csr()
            want_fr, want_spec, want_r2 = orc.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"],
                                                          csr["style"], P)
            fr, spec, r2, status, _ = ops.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"],
                                                      csr["style"], P, group_by_industry=grouped)
            tag = (T, N, P, Q, grouped)
            assert record("reg.oracle.shape_sweep.factor_ret[per column]", relerr_cols(fr, want_fr)) < TOL, tag
            assert relerr(spec, want_spec, scale=np.max(np.abs(csr["ret"]))) < TOL, tag
            assert np.max(np.abs(r2 - want_r2)) < TOL, tag


@pytest.mark.parametrize("case", ["small", "csi", "nostyle_scale"])
def test_crosssection_dropin(reg_golden, case):
    """The drop-in CrossSection class itself (CrossSection.py:35-53, 108): constructor attributes and the
    reg() return tuple against what the reference's own object held on the first date of each golden case."""
    import contextlib
    import io
    import pandas as pd
    from mfm.CrossSection import CrossSection, style_factor_norm
    g = _case(reg_golden, case)
    P = int(g["P"])
    a, b = int(g["date_off"][0]), int(g["date_off"][1])
    n, Q = b - a, g["style"].shape[1]
    names = [f"{i:06d}.SZ" for i in range(n)]
    base = pd.DataFrame({"date": ["2010/01/04"] * n, "stocknames": names, "capital": g["cap"][a:b], "ret": g["ret"][a:b]})
    style = pd.DataFrame(g["style"][a:b], columns=[f"style_{q:02d}" for q in range(Q)])
    onehot = (g["ind"][a:b][:, None] == np.arange(P)[None, :]).astype(np.int64)
    industry = pd.DataFrame(onehot, columns=[f"ind_{p:02d}" for p in range(P)])
    with contextlib.redirect_stdout(io.StringIO()) as out:
        cs = CrossSection(base, style, industry)
        f, spec, expo, r2 = cs.reg()
    assert "Cross Section Regression" in out.getvalue()      # the status line of CrossSection.py:52-53
    assert (cs.N, cs.P, cs.Q) == (n, P, Q) and cs.date == "2010/01/04" and cs.stocknames == names
    assert cs.style_factors_names == list(style.columns) and cs.industry_factors_names == list(industry.columns)
    assert cs.country_factors.shape == (n, 1) and (cs.country_factors == 1).all()
    assert np.array_equal(cs.industry_factors, onehot)
    assert record(f"crosssection.{case}.W", relerr(cs.W, g["W0"])) < 1e-14
    assert record(f"crosssection.{case}.style_factors", relerr(cs.style_factors, g["style_norm0"])) < 1e-12
    assert relerr(style_factor_norm(g["style"][a:b], g["cap"][a:b]), g["style_norm0"]) < 1e-12
    assert isinstance(expo, np.matrix) and expo.shape == (1 + P + Q, 1 + P + Q) and isinstance(r2, float)
    assert record(f"crosssection.{case}.factor_ret[per column]", relerr_cols(f[None], g["factor_ret"][:1])) < TOL
    assert record(f"crosssection.{case}.exposure", relerr(np.asarray(expo), g["exposure0"])) < TOL
    assert relerr(spec, g["specific_ret"][a:b], scale=np.max(np.abs(g["ret"]))) < TOL
    assert abs(r2 - g["R2"][0]) < TOL
