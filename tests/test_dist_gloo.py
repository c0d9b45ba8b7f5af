"""World-size-2 `gloo` tests (CPU) of the multi-GPU orchestration in bmfm/dist.py: the exchange logic
(all-gather of factor returns and vol-regime terms, all-reduce of the eigen accumulators) is run with
an oracle-backed backend and must reproduce the single-process oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bmfm import dist as bd
from bmfm.synth import make_panel
from oracle import barra_oracle as orc


class OracleBackend:
    """Same interface as bmfm.dist.GpuBackend, arithmetic from the CPU oracle (test only)."""

    def reg(self, T, P, Q, max_rows, off, cap, ret, ind, style):
        fr, spec, r2 = orc.reg_by_time(off.numpy(), cap.numpy(), ret.numpy(), ind.numpy(), style.numpy(), P)
        return torch.from_numpy(fr), torch.from_numpy(spec), torch.from_numpy(r2), torch.zeros(T, dtype=torch.int32)

    def newey_west_range(self, fr_all, t0, t1, q, tao):
        fr = fr_all.numpy()
        K = fr.shape[1]
        cov = np.full((t1 - t0, K, K), np.nan)
        st = np.ones(t1 - t0, dtype=np.int32)
        for t in range(t0, t1):
            try:
                cov[t - t0] = orc.newey_west(fr[:t + 1], q, tao)
                st[t - t0] = 0
            except Exception:
                pass
        return torch.from_numpy(cov), torch.from_numpy(st)

    # -- Newey-West by exchanged state records: the recurrence form of csrc/newey_west.cu restated in numpy --
    @staticmethod
    def _nw_zero(K, q):
        return {"A": np.zeros((1 + q, K, K)), "S": np.zeros(K), "R": np.zeros((q, K)), "P": np.zeros((q, K)),
                "Z": 0.0, "Zi": np.zeros(q)}

    @staticmethod
    def _nw_advance(s, fr, r, lam, q):
        x = fr[r]
        s["A"][0] = lam * s["A"][0] + np.outer(x, x)
        s["S"] = lam * s["S"] + x
        s["Z"] = lam * s["Z"] + 1.0
        for i in range(1, q + 1):
            xl = fr[r - i] if r - i >= 0 else np.zeros_like(x)
            s["A"][i] = lam * s["A"][i] + np.outer(xl, x)
            s["P"][i - 1] = lam * s["P"][i - 1] + xl
            s["R"][i - 1] = lam * s["R"][i - 1] + (x if r >= i else 0.0)
            s["Zi"][i - 1] = lam * s["Zi"][i - 1] + (1.0 if r >= i else 0.0)

    @staticmethod
    def _nw_cov(s, q):
        Z = s["Z"]
        m = s["S"] / Z
        V = s["A"][0] / Z - np.outer(m, m)
        for i in range(1, q + 1):
            G = (s["A"][i] - np.outer(s["P"][i - 1], m) - np.outer(m, s["R"][i - 1]) + np.outer(m, m) * s["Zi"][i - 1]) / Z
            V = V + (1 - i / (1 + q)) * (G + G.T)
        return V

    @staticmethod
    def _nw_flat(s):
        return np.concatenate([s["A"].ravel(), s["S"], s["R"].ravel(), s["P"].ravel(), [s["Z"]], s["Zi"]])

    @staticmethod
    def _nw_unflat(v, K, q):
        s = OracleBackend._nw_zero(K, q)
        o = 0
        for key in ("A", "S", "R", "P"):
            n = s[key].size
            s[key] = v[o:o + n].reshape(s[key].shape).copy()
            o += n
        s["Z"] = float(v[o])
        s["Zi"] = v[o + 1:o + 1 + q].copy()
        return s

    def nw_shard_total(self, fr_all, t0, t1, q, tao, rank, world):
        fr = fr_all.numpy()
        lam = 0.5 ** (1.0 / tao)
        s = self._nw_zero(fr.shape[1], q)
        for r in range(t0, t1):
            self._nw_advance(s, fr, r, lam, q)
        return torch.from_numpy(self._nw_flat(s))

    def nw_shard_emit(self, fr_all, t0, t1, q, tao, rank, range_ends, totals):
        fr, tot = fr_all.numpy(), totals.numpy()
        K = fr.shape[1]
        lam = 0.5 ** (1.0 / tao)
        carry = np.zeros(tot.shape[1])
        for r in range(rank):
            carry += lam ** (t0 - range_ends[r]) * tot[r]
        s = self._nw_unflat(carry, K, q)
        cov = np.full((t1 - t0, K, K), np.nan)
        st = np.ones(t1 - t0, dtype=np.int32)
        for r in range(t0, t1):
            self._nw_advance(s, fr, r, lam, q)
            if r + 1 > max(q, K):
                cov[r - t0] = self._nw_cov(s, q)
                st[r - t0] = 0
        return torch.from_numpy(cov), torch.from_numpy(st)

    def eigen(self, cov, status, T_sim, M, scale, sim_range=None, allreduce=None):
        cov, status = cov.numpy(), status.numpy()
        T, K = cov.shape[0], cov.shape[1]
        m_lo, m_hi = (0, M) if sim_range is None else sim_range
        vsum = np.zeros((T, K))
        eig = {}
        for t in range(T):
            if status[t] != 0:
                continue
            D0, U0 = orc.canonical_eig(cov[t])
            eig[t] = (D0, U0)
            _, _, sims = orc.eigen_risk_adj(cov[t], T_sim, m_hi, scale, return_sims=True)
            for (Dm, dg) in sims[m_lo:m_hi]:
                vsum[t] += dg / Dm
        vs = torch.from_numpy(vsum)
        if allreduce is not None:
            allreduce(vs)
        out = np.full_like(cov, np.nan)
        for t, (D0, U0) in eig.items():
            v = scale * (np.sqrt(vs.numpy()[t] / M) - 1) + 1
            out[t] = U0 @ np.diag(v ** 2 * D0) @ U0.T
        return torch.from_numpy(out), torch.from_numpy(status.copy())

    def vr_bias(self, T_total, t0, fr_block, cov_block, status_block, tao):
        fr, cov, st = fr_block.numpy(), cov_block.numpy(), status_block.numpy()
        n = len(fr)
        wb, wv = np.zeros(n), np.zeros(n)
        for i in range(n):
            if st[i] != 0:
                continue
            var = np.diag(cov[i])
            b2 = np.mean(fr[i] ** 2 / var)
            w = 0.5 ** ((T_total - 1 - (t0 + i)) / tao)
            wb[i], wv[i] = w * b2, w
        return torch.from_numpy(wb), torch.from_numpy(wv)

    def vr_scale(self, T_total, t0, wb_all, wv_all, cov_block, status_block):
        nb, nv = np.cumsum(wb_all.numpy()), np.cumsum(wv_all.numpy())
        with np.errstate(invalid="ignore", divide="ignore"):
            lamb = np.where(nv > 0, np.sqrt(nb / np.where(nv > 0, nv, 1)), 0.0)
        cov = cov_block.numpy()
        n = cov.shape[0]
        out = cov * (lamb[t0:t0 + n] ** 2)[:, None, None]
        return torch.from_numpy(out), torch.from_numpy(lamb)


    # -- vol-regime by exchanged block totals (the form of bmfm_vol_regime_shard_totals / _scale) --
    def vr_shard_totals(self, T_total, t0, fr_block, cov_block, status_block, tao, rank, world):
        wb, wv = self.vr_bias(T_total, t0, fr_block, cov_block, status_block, tao)
        return wb, wv, torch.stack([wb.sum(), wv.sum()])

    def vr_shard_scale(self, wb, wv, totals, rank, cov_block, status_block):
        cb, cv = totals[:rank, 0].sum().item(), totals[:rank, 1].sum().item()
        nb, nv = cb + np.cumsum(wb.numpy()), cv + np.cumsum(wv.numpy())
        with np.errstate(invalid="ignore", divide="ignore"):
            lamb = np.where(nv > 0, np.sqrt(nb / np.where(nv > 0, nv, 1)), 0.0)
        return torch.from_numpy(cov_block.numpy() * (lamb ** 2)[:, None, None]), torch.from_numpy(lamb)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _panel():
    return make_panel(T=40, N=30, P=3, Q=2, seed=77, ragged=0.1)


def _worker(rank, world, port, q, shard_nw=False, shard_vr=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pan = _panel()
        csr = pan.to_csr()
        T = pan.T
        counts = bd.shard_counts(T, world)
        lo, hi = bd.shard_range(T, rank, world)
        a, b = int(csr["date_off"][lo]), int(csr["date_off"][hi])
        off = torch.from_numpy(csr["date_off"][lo:hi + 1] - csr["date_off"][lo])
        tt = lambda x: torch.from_numpy(np.ascontiguousarray(x))
        pipe = bd.RiskPipeline(OracleBackend(), rank, world)
        res = pipe.run(counts, pan.P, pan.Q, pan.N, off, tt(csr["cap"][a:b]), tt(csr["ret"][a:b]),
                       tt(csr["ind"][a:b]), tt(csr["style"][a:b]), q=2, tao_nw=15, tao_vr=8,
                       eigen=dict(T_sim=30, M=4, scale_coef=1.4), shard_nw=shard_nw, shard_vr=shard_vr)
        # config-4 style: all dates on every rank, simulations sharded
        fr_all = res["factor_ret_all"]
        nw_all, st_all = OracleBackend().newey_west_range(fr_all, 0, T, 2, 15)
        er_sims, _ = bd.eigen_sharded_by_sims(OracleBackend(), nw_all, st_all, 30, 4, 1.4, rank, world)
        q.put((rank, {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in res.items()}, er_sims.numpy()))
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions():
    for n in (0, 1, 7, 2500, 3750):
        for world in (1, 2, 3, 8):
            edges = [bd.shard_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1 and sizes == bd.shard_counts(n, world)


def test_newey_west_recurrence_and_record_exchange():
    """The recurrence form behind the date-sharded Newey-West (one state record per rank, decayed and summed
    into the next ranks' carry-in) against the reference's from-scratch weighted covariance, in numpy."""
    from bmfm.synth import make_factor_returns
    ob = OracleBackend()
    for (T, K, q, tao, world) in [(60, 4, 2, 9.0, 3), (45, 3, 0, 20.0, 2), (80, 5, 3, 14.0, 4)]:
        fr = torch.from_numpy(make_factor_returns(T, K, seed=T + K) + 0.004)
        ranges = [bd.shard_range(T, r, world) for r in range(world)]
        ends = [hi for (_, hi) in ranges]
        totals = torch.stack([ob.nw_shard_total(fr, lo, hi, q, tao, r, world) for r, (lo, hi) in enumerate(ranges)])
        for r, (lo, hi) in enumerate(ranges):
            cov, st = ob.nw_shard_emit(fr, lo, hi, q, tao, r, ends, totals)
            for t in range(lo, hi):
                if t + 1 > max(q, K):
                    want = orc.newey_west(fr.numpy()[:t + 1], q, tao)
                    assert st[t - lo] == 0
                    assert np.max(np.abs(cov[t - lo].numpy() - want)) < 1e-12 * np.max(np.abs(want))
                else:
                    assert st[t - lo] != 0


@pytest.mark.timeout(300)
@pytest.mark.parametrize("shard_nw,shard_vr", [(False, False), (True, False), (True, True)])
def test_sharded_pipeline_matches_single_process(shard_nw, shard_vr):
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, shard_nw, shard_vr)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        rank, res, er_sims = q.get(timeout=240)
        got[rank] = (res, er_sims)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0

    pan = _panel()
    csr = pan.to_csr()
    fr, spec, r2 = orc.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P)
    nw = orc.newey_west_by_time(fr, 2, 15)
    er = orc.eigen_risk_adj_by_time(nw, 30, 4, 1.4)
    vr, lamb = orc.vol_regime_adj_by_time(fr, er, 8)

    def close(a, b, tol=1e-11):
        a, b = np.asarray(a), np.asarray(b)
        assert np.array_equal(np.isnan(a), np.isnan(b))
        return np.nanmax(np.abs(a - b)) <= tol * max(np.nanmax(np.abs(b)), 1e-300)

    for rank in range(world):
        res, er_sims = got[rank]
        t0, n = res["t0"], res["n"]
        assert close(res["factor_ret"], fr[t0:t0 + n])
        assert close(res["factor_ret_all"], fr)
        if shard_vr:       # the two-totals exchange leaves every rank with the lamb of its own dates
            assert close(res["lamb_all"][t0:t0 + n], np.array(lamb)[t0:t0 + n], 1e-10)
            assert np.isnan(np.delete(res["lamb_all"], np.arange(t0, t0 + n))).all()
        else:
            assert close(res["lamb_all"], np.array(lamb), 1e-10)
        for i in range(n):
            t = t0 + i
            assert (res["nw_status"][i] == 0) == (nw[t] is not None)
            if nw[t] is not None:
                assert close(res["nw"][i], nw[t])
            if er[t] is not None:
                assert close(res["cov"][i], er[t], 1e-10)
                assert close(res["vr"][i], vr[t], 1e-10)
        for t in range(pan.T):       # simulations sharded, all-reduced
            if er[t] is not None:
                assert close(er_sims[t], er[t], 1e-10)


# ---- descriptors sharded by stock: the host logic of winsorize_sharded / orthogonalize_sharded ----------------------
class OraclePostprocCtx:
    """The six pass entry points of the C-ABI (bmfm_winsorize_sums / _sqdev / _apply, bmfm_orthogonalize_pass) restated
    in numpy on CPU tensors (test only): same arguments, same per-date sums."""

    @staticmethod
    def _dates(date_off):
        off = date_off.numpy()
        return [(int(off[t]), int(off[t + 1])) for t in range(len(off) - 1)]

    def winsorize_sums(self, T, date_off, x, sums):
        for t, (a, b) in enumerate(self._dates(date_off)):
            v = x.numpy()[a:b]
            ok = ~np.isnan(v)
            sums[t, 0], sums[t, 1] = float(v[ok].sum()), float(ok.sum())

    def winsorize_sqdev(self, T, date_off, x, mean, sqdev):
        for t, (a, b) in enumerate(self._dates(date_off)):
            v = x.numpy()[a:b]
            ok = ~np.isnan(v)
            sqdev[t] = float(((float(mean[t]) - v[ok]) ** 2).sum())

    def winsorize_apply(self, T, date_off, x, lo, hi, out):
        for t, (a, b) in enumerate(self._dates(date_off)):
            v = x.numpy()[a:b].copy()
            l, h = float(lo[t]), float(hi[t])
            if not (np.isnan(l) or np.isnan(h)):
                ok = ~np.isnan(v)
                v[ok] = np.minimum(np.maximum(v[ok], l), h)
            out.numpy()[a:b] = v

    def orthogonalize_pass(self, which, T, date_off, y, Xs, out, shift=None, gram=None, count=None):
        k = len(Xs)
        for t, (a, b) in enumerate(self._dates(date_off)):
            yy = y.numpy()[a:b]
            X = np.column_stack([c.numpy()[a:b] for c in Xs])
            ok = ~(np.isnan(yy) | np.isnan(X).any(axis=1))
            if which == 0:
                out[t] = torch.from_numpy(np.concatenate([[ok.sum()], X[ok].sum(0), [yy[ok].sum()]]))
                continue
            sh = shift[t].numpy()
            A = np.column_stack([np.ones(ok.sum()), X[ok] - sh[:k]])
            r = yy[ok] - sh[k]
            if which == 1:
                G = A.T @ A
                out[t] = torch.from_numpy(np.concatenate([G[np.triu_indices(k + 1)], A.T @ r]))
                continue
            res = np.full(b - a, np.nan)
            if float(count[t]) >= k + 2:
                g = gram[t].numpy()
                G = np.zeros((k + 1, k + 1))
                G[np.triu_indices(k + 1)] = g[:(k + 1) * (k + 2) // 2]
                G = G + np.triu(G, 1).T
                coef = np.linalg.pinv(G) @ g[(k + 1) * (k + 2) // 2:]
                Afull = np.column_stack([np.ones(b - a), X - sh[:k]])
                res = (yy - sh[k]) - Afull @ coef
            out.numpy()[a:b] = res


def _postproc_table():
    rng = np.random.default_rng(31)
    T, S = 12, 23
    cols = [rng.normal(2.0 * j, 1.0 + j, size=(T, S)) for j in range(3)]
    for c in cols:
        c[rng.random((T, S)) < 0.08] = np.nan
    cols[1][4, 2:] = np.nan            # fewer than k + 2 complete rows on date 4
    return T, S, cols


def _postproc_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        T, S, cols = _postproc_table()
        lo, hi = S * rank // world, S * (rank + 1) // world
        mine = [torch.from_numpy(np.ascontiguousarray(c[:, lo:hi]).reshape(-1)) for c in cols]
        off = torch.arange(T + 1, dtype=torch.int64) * (hi - lo)
        ctx = OraclePostprocCtx()
        wz = [torch.empty_like(x) for x in mine]
        bd.winsorize_sharded(ctx, T, off, mine, 2.5, wz, lambda t: dist.all_reduce(t))
        res = torch.empty_like(mine[0])
        bd.orthogonalize_sharded(ctx, T, off, mine[0], [mine[1], mine[2]], res, lambda t: dist.all_reduce(t))
        q.put((rank, [w.numpy().reshape(T, -1) for w in wz], res.numpy().reshape(T, -1)))
    finally:
        dist.destroy_process_group()


def test_stock_sharded_post_processing_gloo():
    """Two ranks, each with half of the stocks of every date: the drivers of bmfm/dist.py with gloo all-reduces
    reproduce the oracle's winsorize / orthogonalize of the whole table."""
    world, port = 2, _free_port()
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    procs = [ctxm.Process(target=_postproc_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict()
    for _ in range(world):
        r, wz, res = q.get(timeout=120)
        got[r] = (wz, res)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    T, S, cols = _postproc_table()
    for j in range(3):
        have = np.concatenate([got[r][0][j] for r in range(world)], axis=1)
        want = np.stack([orc.winsorize(cols[j][t]) for t in range(T)])
        assert np.array_equal(np.isnan(have), np.isnan(want))
        assert np.nanmax(np.abs(have - want)) < 1e-12
    have = np.concatenate([got[r][1] for r in range(world)], axis=1)
    want = np.stack([orc.orthogonalize(cols[0][t], np.column_stack([cols[1][t], cols[2][t]])) for t in range(T)])
    assert np.array_equal(np.isnan(have), np.isnan(want))
    assert np.isnan(have[4]).all()
    assert np.nanmax(np.abs(have - want)) < 1e-10
