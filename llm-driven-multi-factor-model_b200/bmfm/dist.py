"""Multi-GPU plumbing for the risk-model path: one process per GPU,
``torch.distributed`` (NCCL on GPUs, gloo in the CPU tests) for the few small
exchanges the path has:

* dates shard for the regression — no exchange;
* factor returns are all-gathered (T x K doubles) so that every rank's
  expanding Newey-West windows see the earlier ranks' dates;
* the vol-regime sums need every date's weighted bias term: all-gather of
  2 x T doubles;
* the eigenfactor simulations can shard instead of the dates: all-reduce of
  the T x K bias accumulators.

``RiskPipeline`` holds only this orchestration; the arithmetic comes from a
backend object (``GpuBackend`` = the C-ABI kernels).  The CPU tests run the same
orchestration over gloo with an oracle-backed backend to check the exchange
logic without a GPU.
"""
from __future__ import annotations

import numpy as np


def shard_range(n, rank, world):
    """Contiguous balanced block [lo, hi) of n items for `rank` of `world`."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_counts(n, world):
    return [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]


def all_gather_rows(x, counts, group=None):
    """Concatenate per-rank row blocks (rank r contributes counts[r] rows) in rank order."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return x
    world = dist.get_world_size(group)
    mx = max(counts)
    pad = torch.zeros((mx,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    pad[: x.shape[0]] = x
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    return torch.cat([outs[r][: counts[r]] for r in range(world)], dim=0)


def all_reduce_sum(x, group=None):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(x, group=group)
    return x


class SymmetricGather:
    """Gathered series in symmetric (peer-mapped) memory: the producing kernels store each rank's block
    straight into every rank's copy over NVLink (bmfm_reg_by_time_gather, bmfm_vol_regime_bias_gather) and
    a device-side barrier replaces the all-gather.  Needs torch symmetric memory on an NVLink/NVSwitch box;
    RiskPipeline falls back to NCCL all-gathers when it is not given one of these."""

    def __init__(self, T_total, K, device, group=None, nw_state_doubles=0):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        g = group or dist.group.WORLD
        self.world = dist.get_world_size(g)
        # sharded Newey-West: one end-of-range state record per rank (bmfm_newey_west_shard_totals / _emit)
        self.nw_totals = self.h_tot = None
        if nw_state_doubles:
            self.nw_totals = symm.empty((self.world, int(nw_state_doubles)), dtype=torch.float64, device=device)
            self.h_tot = symm.rendezvous(self.nw_totals, group=g)
            self.tot_peers = [int(x) for x in self.h_tot.buffer_ptrs]
        self.fr_all = symm.empty((T_total, K), dtype=torch.float64, device=device)
        self.h_fr = symm.rendezvous(self.fr_all, group=g)
        # sharded vol-regime: two totals per rank (bmfm_vol_regime_shard_totals / _scale)
        self.vr_totals = symm.empty((self.world, 2), dtype=torch.float64, device=device)
        self.h_vrt = symm.rendezvous(self.vr_totals, group=g)
        self.vrt_peers = [int(x) for x in self.h_vrt.buffer_ptrs]
        self.wbv = symm.empty((2, T_total), dtype=torch.float64, device=device)
        self.h_wbv = symm.rendezvous(self.wbv, group=g)
        self.fr_peers = [int(x) for x in self.h_fr.buffer_ptrs]
        self.wb_peers = [int(x) for x in self.h_wbv.buffer_ptrs]
        self.wv_peers = [int(x) + T_total * 8 for x in self.h_wbv.buffer_ptrs]


class GpuBackend:
    """The arithmetic of the pipeline on this process's GPU (C-ABI kernels)."""

    def __init__(self, ctx=None):
        from . import _cabi
        self.ctx = ctx or _cabi.default_context()

    def reg(self, T, P, Q, max_rows, off, cap, ret, ind, style):
        from .ops import reg_by_time_device
        fr, spec, r2, status, _ = reg_by_time_device(self.ctx, T, P, Q, max_rows, off, cap, ret, ind, style)
        return fr, spec, r2, status

    def newey_west_range(self, fr_all, t0, t1, q, tao):
        import torch
        T, K = fr_all.shape
        cov = torch.empty((t1 - t0, K, K), dtype=torch.float64, device=fr_all.device)
        st = torch.empty(t1 - t0, dtype=torch.int32, device=fr_all.device)
        self.ctx.newey_west_by_time_range(T, K, q, tao, fr_all, t0, t1, cov, st)
        return cov, st

    def nw_shard_total(self, fr_all, t0, t1, q, tao, rank, world):
        """State record (1-D tensor) of this rank's dates reduced from a zero state
        (bmfm_newey_west_shard_totals with this GPU's own buffer as the only destination)."""
        import torch
        T, K = fr_all.shape
        ns = self.ctx.newey_west_state_doubles(K, q)
        buf = torch.zeros((world, ns), dtype=torch.float64, device=fr_all.device)
        self.ctx.newey_west_shard_totals(T, K, q, tao, fr_all, t0, t1, rank, [buf.data_ptr()])
        return buf[rank]

    def nw_shard_emit(self, fr_all, t0, t1, q, tao, rank, range_ends, totals):
        """Covariances of [t0, t1) from the chunk states left by nw_shard_total and every rank's record."""
        import torch
        T, K = fr_all.shape
        cov = torch.empty((t1 - t0, K, K), dtype=torch.float64, device=fr_all.device)
        st = torch.empty(t1 - t0, dtype=torch.int32, device=fr_all.device)
        self.ctx.newey_west_shard_emit(T, K, q, tao, fr_all, t0, t1, rank, range_ends[:rank], totals.contiguous(), cov, st)
        return cov, st

    def eigen(self, cov, status, T_sim, M, scale, sim_range=None, allreduce=None):
        from .ops import eigen_risk_adj_device
        return eigen_risk_adj_device(self.ctx, cov, status, T_sim, M, scale, sim_range, allreduce)

    def vr_bias(self, T_total, t0, fr_block, cov_block, status_block, tao):
        import torch
        n, K = fr_block.shape
        wb = torch.empty(n, dtype=torch.float64, device=fr_block.device)
        wv = torch.empty(n, dtype=torch.float64, device=fr_block.device)
        self.ctx.vol_regime_bias(T_total, t0, n, K, tao, fr_block, cov_block, status_block, wb, wv)
        return wb, wv

    def vr_scale(self, T_total, t0, wb_all, wv_all, cov_block, status_block):
        import torch
        n, K = cov_block.shape[0], cov_block.shape[1]
        out = torch.empty_like(cov_block)
        lamb = torch.empty(T_total, dtype=torch.float64, device=cov_block.device)
        self.ctx.vol_regime_scale(T_total, t0, n, K, wb_all, wv_all, cov_block, status_block, out, lamb)
        return out, lamb

    def vr_shard_totals(self, T_total, t0, fr_block, cov_block, status_block, tao, rank, world):
        """Per-date terms of this rank's block (kept) and their two sums as a [2] tensor
        (bmfm_vol_regime_shard_totals with this GPU's own buffer as the only destination)."""
        import torch
        n, K = fr_block.shape
        wb = torch.empty(n, dtype=torch.float64, device=fr_block.device)
        wv = torch.empty(n, dtype=torch.float64, device=fr_block.device)
        buf = torch.zeros((world, 2), dtype=torch.float64, device=fr_block.device)
        self.ctx.vol_regime_shard_totals(T_total, t0, n, K, tao, fr_block, cov_block, status_block, wb, wv, rank,
                                         [buf.data_ptr()])
        return wb, wv, buf[rank]

    def vr_shard_scale(self, wb, wv, totals, rank, cov_block, status_block):
        """lamb of this rank's dates from its own terms and the earlier ranks' totals ([world, 2]); scaled block."""
        import torch
        n, K = cov_block.shape[0], cov_block.shape[1]
        out = torch.empty_like(cov_block)
        lamb = torch.empty(n, dtype=torch.float64, device=cov_block.device)
        self.ctx.vol_regime_shard_scale(n, K, wb, wv, totals.contiguous(), rank, cov_block, status_block, out, lamb)
        return out, lamb


class RiskPipeline:
    """reg_by_time -> Newey_West_by_time -> [eigen_risk_adj_by_time] -> vol_regime_adj_by_time with the
    dates of the panel sharded over the ranks of `group` (demo.py:38-42, MFM.py:48-167)."""

    def __init__(self, backend, rank=0, world=1, group=None):
        self.b = backend
        self.rank, self.world, self.group = rank, world, group

    def run(self, counts, P, Q, max_rows, off, cap, ret, ind, style, q=2, tao_nw=252, tao_vr=84,
            eigen=None, symmetric=None, shard_nw=False, shard_vr=False):
        """`counts[r]` = number of dates owned by rank r (date order = rank order); the panel arrays are
        this rank's block.  `eigen` = None (vol-regime directly on the Newey-West covariances, BASELINE
        config 2) or dict(T_sim=, M=, scale_coef=).  `shard_nw`: Newey-West by exchanged state records (see
        bmfm_newey_west_shard_totals) instead of the whole-series pass.  `shard_vr`: vol-regime by two exchanged
        block totals per rank (bmfm_vol_regime_shard_totals) instead of gathered per-date terms; lamb_all then holds
        this rank's dates only (NaN elsewhere).  Returns a dict of this rank's blocks + lamb_all."""
        T_total = int(sum(counts))
        t0 = int(sum(counts[: self.rank]))
        n = int(counts[self.rank])
        if symmetric is not None:
            self._range_ends = [int(sum(counts[: r + 1])) for r in range(len(counts))]
            return self._run_fused(symmetric, T_total, t0, n, P, Q, max_rows, off, cap, ret, ind, style, q, tao_nw,
                                   tao_vr, eigen)
        fr, spec, r2, status = self.b.reg(n, P, Q, max_rows, off, cap, ret, ind, style)
        fr_all = all_gather_rows(fr, counts, self.group)
        if shard_nw and self.world > 1:
            # every rank reduces its own dates only and the ranks exchange one state record each (instead of
            # every rank reducing all earlier dates of the gathered series)
            rec = self.b.nw_shard_total(fr_all, t0, t0 + n, q, tao_nw, self.rank, self.world)
            totals = all_gather_rows(rec.reshape(1, -1), [1] * self.world, self.group)
            ends = [int(sum(counts[: r + 1])) for r in range(self.world)]
            nw, nw_st = self.b.nw_shard_emit(fr_all, t0, t0 + n, q, tao_nw, self.rank, ends, totals)
        else:
            nw, nw_st = self.b.newey_west_range(fr_all, t0, t0 + n, q, tao_nw)
        cov, cov_st = nw, nw_st
        if eigen is not None:
            cov, cov_st = self.b.eigen(nw, nw_st, eigen["T_sim"], eigen["M"], eigen["scale_coef"])
        if shard_vr and self.world > 1:
            import torch
            wb, wv, tot = self.b.vr_shard_totals(T_total, t0, fr, cov, cov_st, tao_vr, self.rank, self.world)
            totals = all_gather_rows(tot.reshape(1, 2), [1] * self.world, self.group)
            vr, lamb = self.b.vr_shard_scale(wb, wv, totals, self.rank, cov, cov_st)
            lamb_all = torch.full((T_total,), float("nan"), dtype=lamb.dtype, device=lamb.device)
            lamb_all[t0:t0 + n] = lamb
        else:
            wb, wv = self.b.vr_bias(T_total, t0, fr, cov, cov_st, tao_vr)
            wb_all = all_gather_rows(wb, counts, self.group)
            wv_all = all_gather_rows(wv, counts, self.group)
            vr, lamb_all = self.b.vr_scale(T_total, t0, wb_all, wv_all, cov, cov_st)
        return {"factor_ret": fr, "factor_ret_all": fr_all, "specific_ret": spec, "R2": r2, "status": status,
                "nw": nw, "nw_status": nw_st, "cov": cov, "cov_status": cov_st, "vr": vr,
                "lamb_all": lamb_all, "t0": t0, "n": n}


    def _run_fused(self, sg, T_total, t0, n, P, Q, max_rows, off, cap, ret, ind, style, q, tao_nw, tao_vr, eigen):
        """Same pipeline with both gathers done by the kernels over peer memory (GpuBackend only)."""
        import torch
        ctx = self.b.ctx
        K = 1 + P + Q
        dev = cap.device
        fr = torch.empty((n, K), dtype=torch.float64, device=dev)
        spec = torch.empty(cap.shape[0], dtype=torch.float64, device=dev)
        r2 = torch.empty(n, dtype=torch.float64, device=dev)
        status = torch.empty(n, dtype=torch.int32, device=dev)
        ctx.reg_by_time_gather(n, P, Q, max_rows, off, cap, ret, ind, style, fr, spec, r2, status, sg.fr_peers, t0)
        sg.h_fr.barrier()
        if sg.nw_totals is not None:
            # every rank reduces only its own dates; one state record per rank crosses NVLink
            nw = torch.empty((n, K, K), dtype=torch.float64, device=dev)
            nw_st = torch.empty(n, dtype=torch.int32, device=dev)
            ctx.newey_west_shard_totals(T_total, K, q, tao_nw, sg.fr_all, t0, t0 + n, self.rank, sg.tot_peers)
            sg.h_tot.barrier()
            ctx.newey_west_shard_emit(T_total, K, q, tao_nw, sg.fr_all, t0, t0 + n, self.rank,
                                      self._range_ends[: self.rank], sg.nw_totals, nw, nw_st)
        else:
            nw, nw_st = self.b.newey_west_range(sg.fr_all, t0, t0 + n, q, tao_nw)
        cov, cov_st = nw, nw_st
        if eigen is not None:
            cov, cov_st = self.b.eigen(nw, nw_st, eigen["T_sim"], eigen["M"], eigen["scale_coef"])
        # lamb is a ratio of two prefix sums: every rank keeps its per-date terms and only the two block totals cross
        # NVLink; lamb_all holds this rank's dates (NaN elsewhere)
        wb = torch.empty(n, dtype=torch.float64, device=dev)
        wv = torch.empty(n, dtype=torch.float64, device=dev)
        ctx.vol_regime_shard_totals(T_total, t0, n, K, tao_vr, fr, cov, cov_st, wb, wv, self.rank, sg.vrt_peers)
        sg.h_vrt.barrier()
        vr = torch.empty_like(cov)
        lamb_all = torch.full((T_total,), float("nan"), dtype=torch.float64, device=dev)
        ctx.vol_regime_shard_scale(n, K, wb, wv, sg.vr_totals, self.rank, cov, cov_st, vr, lamb_all[t0:t0 + n])
        return {"factor_ret": fr, "factor_ret_all": sg.fr_all, "specific_ret": spec, "R2": r2, "status": status,
                "nw": nw, "nw_status": nw_st, "cov": cov, "cov_status": cov_st, "vr": vr,
                "lamb_all": lamb_all, "t0": t0, "n": n}


def eigen_sharded_by_sims(backend, cov_all, status_all, T_sim, M, scale_coef, rank, world, group=None):
    """BASELINE config 4: every rank holds all dates' covariances and runs M/world simulations;
    one all-reduce of the T x K accumulators, then every rank finalises (utils.py:71-92)."""
    m_lo, m_hi = shard_range(M, rank, world)
    return backend.eigen(cov_all, status_all, T_sim, M, scale_coef, sim_range=(m_lo, m_hi),
                         allreduce=lambda t: all_reduce_sum(t, group))


# ---- descriptors: stocks shard, per-date moments are all-reduced -------------------------------
# The rolling kernels write [days][my stocks] panels, which already are "long tables grouped by date" with this rank's
# stocks of every date, so the cross-sectional stages run on them in place: every pass leaves per-date sums (KB), one
# all-reduce per pass makes them the sums over ALL stocks, the next pass uses them.  No bulk transpose of the panels.

def winsorize_sharded(ctx, T, date_off, cols, n_std, outs, allreduce):
    """winsorize_factors (post_processing.py:7-24) for several columns of a stock-sharded table at once:
    two all-reduces in total (sums of all columns, then squared deviations of all columns)."""
    import torch
    nc, dev = len(cols), cols[0].device
    sums = torch.empty((nc, T, 2), dtype=torch.float64, device=dev)
    for i, x in enumerate(cols):
        ctx.winsorize_sums(T, date_off, x, sums[i])
    allreduce(sums)
    cnt = sums[..., 1].contiguous()
    mean = (sums[..., 0] / cnt).contiguous()
    q = torch.empty((nc, T), dtype=torch.float64, device=dev)
    for i, x in enumerate(cols):
        ctx.winsorize_sqdev(T, date_off, x, mean[i], q[i])
    allreduce(q)
    nan = torch.full_like(q, float("nan"))
    sd = torch.where(cnt > 1.0, torch.sqrt(q / (cnt - 1.0)), nan)        # pandas std, ddof = 1
    lo, hi = (mean - n_std * sd).contiguous(), (mean + n_std * sd).contiguous()
    for i, x in enumerate(cols):
        ctx.winsorize_apply(T, date_off, x, lo[i], hi[i], outs[i])
    return outs


def orthogonalize_sharded(ctx, T, date_off, y, Xs, out, allreduce):
    """orthogonalize_factors (post_processing.py:47-70) for one dependent column of a stock-sharded table: the
    complete-row count and column sums, then the Gram about the column means — two all-reduces."""
    import torch
    k, dev = len(Xs), y.device
    sums = torch.empty((T, k + 2), dtype=torch.float64, device=dev)
    ctx.orthogonalize_pass(0, T, date_off, y, Xs, sums)
    allreduce(sums)
    count = sums[:, 0].contiguous()
    shift = torch.nan_to_num(sums[:, 1:] / count[:, None]).contiguous()    # column means (0 where a date has no complete row)
    gram = torch.empty((T, (k + 1) * (k + 2) // 2 + k + 1), dtype=torch.float64, device=dev)
    ctx.orthogonalize_pass(1, T, date_off, y, Xs, gram, shift=shift)
    allreduce(gram)
    ctx.orthogonalize_pass(2, T, date_off, y, Xs, out, shift=shift, gram=gram, count=count)
    return out
