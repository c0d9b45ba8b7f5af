#!/usr/bin/env python
"""bench.py — the Barra risk-model hot path on B200, next to the CPU port.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1], "C2"): synthetic CSI800 panel, 800 stocks x
2500 days, P = 31 industries, Q = 10 styles.  One step = MFM.reg_by_time +
Newey_West_by_time(q=2, tao=252) + vol_regime_adj_by_time(tao=42) over the whole
panel.  Metric: regression dates per second (dates processed by the full step /
time).  With N ranks every rank owns its own 2500-date block of a 2500*N-date
panel (weak scaling); factor returns are all-gathered so each rank's expanding
Newey-West windows and vol-regime sums cover the earlier ranks' dates.

A second block ("eigen_mc") reports eigen-adjustment Monte-Carlo simulations per
second on a bounded slice of configs[3] (K = 42, 252-day simulations).

One JSON line on stdout (rank 0).  --impl reference times the CPU port of the
reference's algorithm (oracle/, numpy) on the host cores with the same metric.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))
sys.path.insert(0, ROOT)

N_STOCKS, N_DATES, P, Q = 800, 2500, 31, 10
K = 1 + P + Q
NW_Q, NW_TAO, VR_TAO = 2, 252.0, 42.0
EIG_DATES, EIG_SIMS, EIG_TSIM = 64, 512, 252
BYTES_PER_STOCK_DATE = 108   # cap 8 + ret 8 + 10 styles 80 + industry id 4 + specific return 8


def log(*a):
    print(*a, file=sys.stderr, flush=True)


_RESULT_FD = None


def claim_stdout():
    """Keep stdout to exactly one JSON line: libraries (NCCL's version banner, ...) write to fd 1 directly, so
    fd 1 is pointed at stderr for the whole run and the result line goes to a saved copy of the real stdout."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    fd = _RESULT_FD if _RESULT_FD is not None else 1
    while data:
        data = data[os.write(fd, data):]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("reg_fast_kernel_dram_bytes_per_launch")
        except Exception:
            return None
    return None


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region through NVML.  Each NVML query was
    seen to stall the polled GPU for milliseconds, so the region is sampled sparsely: one query right after
    the timed launches are enqueued (the GPU is then in the middle of them) and one every 250 ms after that
    (the recipe's nvidia-smi loop uses 200 ms)."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.sm, self.reasons = [], set()
        self.sm_max = None
        self._stop = threading.Event()
        self.thread = None
        self.err = None
        self.h = None
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            ent = vis.split(",")[gpu_index].strip() if vis else str(gpu_index)
            self.h = (pynvml.nvmlDeviceGetHandleByIndex(int(ent)) if ent.isdigit()
                      else pynvml.nvmlDeviceGetHandleByUUID(ent.encode()))
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:      # noqa: BLE001
            self.err = repr(e)

    def sample(self):
        if self.h is None:
            return
        try:
            nv = self.nv
            self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
            try:
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for bit, name in self.REASONS.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception as e:      # noqa: BLE001
            self.err = repr(e)

    def _loop(self):
        while not self._stop.wait(0.25):
            self.sample()

    def start(self):
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()

    def stop(self):
        self._stop.set()
        if self.thread is not None:
            self.thread.join(timeout=2)
        out = {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.sm_max,
               "samples": len(self.sm), "reasons": sorted(self.reasons)}
        if self.err:
            out["error"] = self.err
        return out


# ------------------------------------------------------------------------------------------------
# CPU port (oracle) — the reference arm and the cpu_baseline leg
# ------------------------------------------------------------------------------------------------
def cpu_sample(panel_csr, n_dates_sample=24, vr_dates=None):
    """Time the numpy restatement of the reference on a bounded sample of the C2 step and
    return (dates_per_sec, description).  One 'date' of the step costs: its cross-sectional
    regression + its expanding-window Newey-West + its share of the vol-regime loop.
    `vr_dates`: run the O(T^2) vol-regime loop on that many dates only and scale its per-date
    cost linearly to the full T (None = the full loop)."""
    from oracle import barra_oracle as orc
    from bmfm.synth import make_factor_returns
    off = panel_csr["date_off"]
    T = len(off) - 1
    pick = np.linspace(K + 1, T - 1, n_dates_sample).astype(int)
    fr = make_factor_returns(T, K, seed=20240002)
    ar = np.arange(P)
    t0 = time.perf_counter()
    for t in pick:
        a, b = int(off[t]), int(off[t + 1])
        onehot = (panel_csr["ind"][a:b, None] == ar[None, :]).astype(np.int64)
        orc.cross_section_reg(panel_csr["cap"][a:b], panel_csr["ret"][a:b], onehot, panel_csr["style"][a:b])
    t_reg = (time.perf_counter() - t0) / len(pick)
    t0 = time.perf_counter()
    covs = {}
    for t in pick:
        covs[t] = orc.newey_west(fr[:t + 1], NW_Q, NW_TAO)
    t_nw = (time.perf_counter() - t0) / len(pick)
    # vol-regime: the reference's O(T^2) loop on a covariance series that repeats the sampled ones
    Tv = T if vr_dates is None else min(T, int(vr_dates))
    cov_list = [covs[pick[min(np.searchsorted(pick, t), len(pick) - 1)]] for t in range(Tv)]
    t0 = time.perf_counter()
    orc.vol_regime_adj_by_time(fr[:Tv], cov_list, VR_TAO)
    t_vr = (time.perf_counter() - t0) / Tv * (T / Tv)      # per-date cost grows linearly with T
    per_date = t_reg + t_nw + t_vr
    vr_txt = "full vol-regime loop" if Tv == T else f"vol-regime loop on {Tv} dates scaled x{T / Tv:.1f} to T={T}"
    desc = (f"numpy port of the reference (oracle/barra_oracle.py; vectorised Newey-West, no N x N diag(W)): "
            f"{len(pick)} evenly spaced dates of the {N_STOCKS}x{T} panel for reg ({t_reg*1e3:.2f} ms/date) and "
            f"expanding Newey-West ({t_nw*1e3:.2f} ms/date), {vr_txt} ({t_vr*1e3:.3f} ms/date); "
            f"dates/s = 1 / sum")
    return 1.0 / per_date, desc


def cpu_eigen_sample(n_sims=24):
    from oracle import barra_oracle as orc
    from bmfm.synth import make_factor_returns
    fr = make_factor_returns(600, K, seed=20240004)
    cov = orc.newey_west(fr, NW_Q, NW_TAO)
    t0 = time.perf_counter()
    orc.eigen_risk_adj(cov, EIG_TSIM, n_sims, 1.4)
    dt = time.perf_counter() - t0
    return n_sims / dt


def blas_threads():
    try:
        from threadpoolctl import threadpool_info
        return max([i.get("num_threads", 1) for i in threadpool_info()] or [1])
    except Exception:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from bmfm.synth import make_panel
    pan = make_panel(N_DATES, N_STOCKS, P, Q, seed=20240002, scale_log10=(-2, 2))
    csr = pan.to_csr()
    vals = []
    # every step is a bounded sample; the sample shrinks with the step count so that the whole
    # --steps K --warmup W run stays within a few minutes
    n_dates = int(max(8, min(600, 12000 // max(1, args.steps + args.warmup))))
    vr_dates = None if args.steps + args.warmup <= 8 else 400
    for i in range(args.warmup):
        cpu_sample(csr, n_dates, vr_dates)
    t0 = time.perf_counter()
    desc = ""
    for i in range(args.steps):
        v, desc = cpu_sample(csr, n_dates, vr_dates)
        vals.append(v)
    dt = time.perf_counter() - t0
    value = float(np.mean(vals))
    eig = cpu_eigen_sample(24)
    cores = blas_threads()
    line = {
        "impl": "reference", "metric": "regression_dates_per_sec", "value": value, "unit": "dates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / max(1, args.steps) * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"C2 synthetic CSI800 panel {N_STOCKS}x{N_DATES}, P={P}, Q={Q}: reg_by_time + "
                               f"Newey_West_by_time(q={NW_Q},tao={int(NW_TAO)}) + vol_regime_adj_by_time(tao={int(VR_TAO)})",
                   "note": "each step is a bounded CPU sample of the workload, see cpu_baseline.sample"},
        "cpu_baseline": {"value": value, "unit": "dates/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "dates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "eigen_mc": {"value": eig, "unit": "sims/s", "config": f"K={K}, T_sim={EIG_TSIM}, 24 sims of one date (numpy port)"},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import torch.distributed as dist
    from bmfm import _cabi
    from bmfm.panel import pack_arrays
    from bmfm.synth import make_panel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep NCCL's version banner off stdout (one JSON line)
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    # everything runs on one side stream, so that the multi-GPU step can be captured in a CUDA graph
    gs = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(gs)
    ctx = _cabi.Context(local_rank, stream=gs.cuda_stream)
    nocoll = os.environ.get("BENCH_NO_COLL") == "1"     # timing experiment only (results are wrong)
    use_graph = os.environ.get("BENCH_GRAPH", "1") == "1"     # one captured step replayed (BENCH_GRAPH=0: eager launches)

    # this rank's block of dates (weak scaling: 2500 dates per rank)
    pan = make_panel(N_DATES, N_STOCKS, P, Q, seed=20240002 + rank, scale_log10=(-2, 2))
    csr = pan.to_csr()
    (cap, ret, ind, style), pos = pack_arrays(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], P)
    date_off = csr["date_off"]
    rows = int(date_off[-1])
    T = N_DATES
    T_all = T * world
    host = {"off": torch.from_numpy(date_off).pin_memory(), "cap": torch.from_numpy(cap).pin_memory(),
            "ret": torch.from_numpy(ret).pin_memory(), "ind": torch.from_numpy(ind).pin_memory(),
            "style": torch.from_numpy(style).pin_memory()}
    d = {k: v.to(dev) for k, v in host.items()}
    fr = torch.empty((T, K), dtype=torch.float64, device=dev)
    # multi-GPU: the gathered factor returns and vol-regime terms live in symmetric (peer-mapped) memory and
    # are filled by the producing kernels themselves over NVLink (fused all-gather); NCCL is the fallback
    fused = world > 1 and os.environ.get("BENCH_FUSED", "1") == "1"
    nw_shard = os.environ.get("BENCH_NW_SHARD", "1") == "1"
    h_fr = h_wbv = h_tot = None
    if fused:
        try:
            import torch.distributed._symmetric_memory as symm
            fr_all = symm.empty((T_all, K), dtype=torch.float64, device=dev)
            h_fr = symm.rendezvous(fr_all, group=dist.group.WORLD)
            wbv_sym = symm.empty((2, T_all), dtype=torch.float64, device=dev)
            h_wbv = symm.rendezvous(wbv_sym, group=dist.group.WORLD)
            fr_peers = [int(x) for x in h_fr.buffer_ptrs]
            wb_peers = [int(x) for x in h_wbv.buffer_ptrs]
            wv_peers = [int(x) + T_all * 8 for x in h_wbv.buffer_ptrs]
            # sharded Newey-West: one end-of-range state record per rank instead of every rank reducing all
            # earlier dates
            nw_ns = ctx.newey_west_state_doubles(K, NW_Q)
            tot_sym = symm.empty((world, nw_ns), dtype=torch.float64, device=dev)
            h_tot = symm.rendezvous(tot_sym, group=dist.group.WORLD)
            tot_peers = [int(x) for x in h_tot.buffer_ptrs]
        except Exception as e:      # noqa: BLE001
            log(f"symmetric memory unavailable ({e!r}); using NCCL all-gathers")
            fused = False
    if not fused:
        fr_all = torch.empty((T_all, K), dtype=torch.float64, device=dev) if world > 1 else fr
    spec = torch.empty(rows, dtype=torch.float64, device=dev)
    r2 = torch.empty(T, dtype=torch.float64, device=dev)
    st = torch.empty(T, dtype=torch.int32, device=dev)
    # every rank keeps the covariance matrices of ITS dates only
    nw = torch.empty((T, K, K), dtype=torch.float64, device=dev)
    nw_st = torch.empty(T, dtype=torch.int32, device=dev)
    vr = torch.empty((T, K, K), dtype=torch.float64, device=dev)
    lamb = torch.empty(T_all, dtype=torch.float64, device=dev)
    # vol-regime terms of this rank's dates, packed [wb | wv] so that one all-gather moves both
    wbv = torch.empty((2, T), dtype=torch.float64, device=dev)
    wbv_all = torch.empty((world, 2, T), dtype=torch.float64, device=dev)
    if fused:
        wb_all, wv_all = wbv_sym[0], wbv_sym[1]
    else:
        wb_all = torch.empty(T_all, dtype=torch.float64, device=dev)
        wv_all = torch.empty(T_all, dtype=torch.float64, device=dev)
    out_host = {"fr": torch.empty((T, K), dtype=torch.float64).pin_memory(),
                "spec": torch.empty(rows, dtype=torch.float64).pin_memory(),
                "r2": torch.empty(T, dtype=torch.float64).pin_memory(),
                "vr": torch.empty((T, K, K), dtype=torch.float64).pin_memory(),
                "lamb": torch.empty(T, dtype=torch.float64).pin_memory()}
    lo = rank * T

    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps + args.warmup)]

    def step(i):
        e = ev[i]
        e[0].record()
        regress()
        e[1].record()
        gather_fr()
        newey_west()
        e[2].record()
        vol_regime()
        e[3].record()

    def regress():
        if fused:
            ctx.reg_by_time_gather(T, P, Q, N_STOCKS, d["off"], d["cap"], d["ret"], d["ind"], d["style"], fr, spec, r2,
                                   st, fr_peers, lo)
        else:
            ctx.reg_by_time(T, P, Q, N_STOCKS, d["off"], d["cap"], d["ret"], d["ind"], d["style"], fr, spec, r2, st)

    def gather_fr():
        if fused:
            h_fr.barrier()            # the factor returns were stored into every rank's buffer by the kernel
        elif world > 1 and not nocoll:
            dist.all_gather_into_tensor(fr_all, fr)

    def newey_west():
        if fused and nw_shard:
            # every rank reduces its own dates only; the ranks exchange one 44 KB state record each
            ctx.newey_west_shard_totals(T_all, K, NW_Q, NW_TAO, fr_all, lo, lo + T, rank, tot_peers)
            h_tot.barrier()
            ctx.newey_west_shard_emit(T_all, K, NW_Q, NW_TAO, fr_all, lo, lo + T, rank,
                                      [(r + 1) * T for r in range(rank)], tot_sym, nw, nw_st)
        else:
            # the recurrence reads the whole gathered series (T_all*K doubles) and emits this rank's dates only
            ctx.newey_west_by_time_range(T_all, K, NW_Q, NW_TAO, fr_all, lo, lo + T, nw, nw_st)

    def step_body():      # the same step without timing events (graph capture)
        regress()
        gather_fr()
        newey_west()
        vol_regime()

    def vol_regime():
        if world == 1:
            ctx.vol_regime_adj_by_time(T, K, VR_TAO, fr, nw, nw_st, vr, lamb)
        elif fused:
            ctx.vol_regime_bias_gather(T_all, lo, T, K, VR_TAO, fr, nw, nw_st, wb_peers, wv_peers)
            h_wbv.barrier()
            ctx.vol_regime_scale(T_all, lo, T, K, wb_all, wv_all, nw, nw_st, vr, lamb)
        else:
            ctx.vol_regime_bias(T_all, lo, T, K, VR_TAO, fr, nw, nw_st, wbv[0], wbv[1])
            if not nocoll:
                dist.all_gather_into_tensor(wbv_all, wbv)
            wb_all.view(world, T).copy_(wbv_all[:, 0])
            wv_all.view(world, T).copy_(wbv_all[:, 1])
            ctx.vol_regime_scale(T_all, lo, T, K, wb_all, wv_all, nw, nw_st, vr, lamb)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    l0 = ctx.launches
    for i in range(args.warmup):
        step(i)
    launches_per_step = (ctx.launches - l0) // max(1, args.warmup)
    barrier()
    graph = None
    n_phase = min(3, args.steps)
    if use_graph:
        for i in range(args.warmup, args.warmup + n_phase):    # eager steps that only feed phase_ms / the roofline
            step(i)
        barrier()
        # one step (kernels + NCCL collectives) captured once and replayed: removes the host-side launch and
        # collective-enqueue gaps that dominate a sub-millisecond multi-GPU step
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=gs):
                step_body()
            graph.replay()
            barrier()
        except Exception as e:      # noqa: BLE001
            log(f"CUDA graph capture failed ({e!r}); timing eager steps")
            graph = None
            torch.cuda.set_stream(gs)
    sampler = ClockSampler(local_rank)
    if rank == 0 and os.environ.get("BENCH_NO_SAMPLER") != "1":
        sampler.start()
    launches0 = ctx.launches
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    if graph is not None:
        for i in range(args.steps):
            graph.replay()
    else:
        for i in range(args.warmup, args.warmup + args.steps):
            step(i)
    t_end.record()
    if rank == 0 and sampler.thread is not None:
        sampler.sample()          # the launches above are asynchronous: the GPU is inside the timed region now
    barrier()
    clocks = sampler.stop() if (rank == 0 and sampler.thread is not None) else None
    launches = ctx.launches - launches0
    if graph is not None:
        launches = launches_per_step * args.steps      # replays do not go through the C-ABI counters
    ms_total = t_start.elapsed_time(t_end)
    tm = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    ms_total = float(tm.item())
    ms_step = ms_total / args.steps
    value = T_all / (ms_step * 1e-3)
    # per-phase times: from the timed steps (eager) or, in graph mode, from the eager warm-up steps
    timed = ev[args.warmup:] if graph is None else ev[args.warmup:args.warmup + n_phase]
    ms_reg = float(np.mean([e[0].elapsed_time(e[1]) for e in timed]))
    ms_nw = float(np.mean([e[1].elapsed_time(e[2]) for e in timed]))
    ms_vr = float(np.mean([e[2].elapsed_time(e[3]) for e in timed]))

    # parity spot-check of what was timed (not a substitute for tests/)
    status_ok = bool((st == 0).all().item())
    log(f"rank {rank}: phases ms reg {ms_reg:.3f} nw {ms_nw:.3f} vr {ms_vr:.3f}; step {ms_step:.3f}; "
        f"status!=0 on {int((st != 0).sum().item())} dates")
    if world > 1:
        ok_t = torch.tensor([1 if status_ok else 0], device=dev)
        dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
        status_ok = bool(ok_t.item())

    # multi-GPU: the fused (peer-memory) gather must give what the NCCL all-gather gives
    gather_check = None
    if fused:
        lamb_fused = lamb.clone()
        fr_all_fused = fr_all.clone()
        ref_all = torch.empty((T_all, K), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(ref_all, fr)
        w2 = torch.stack([wb_all[lo:lo + T], wv_all[lo:lo + T]]).contiguous()
        w2_all = torch.empty((world, 2, T), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(w2_all, w2)
        ok = bool(torch.equal(ref_all, fr_all_fused)) and bool(torch.equal(w2_all[:, 0].reshape(-1), wb_all)) \
            and bool(torch.equal(w2_all[:, 1].reshape(-1), wv_all))
        ok_t = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
        gather_check = "peer-memory gather == NCCL all-gather (bit-identical on every rank)" if ok_t.item() else "MISMATCH"
        del lamb_fused
        if nw_shard:
            # the sharded Newey-West against the pass that reduces the whole gathered series on every rank
            nw_ref = torch.empty_like(nw)
            st_ref = torch.empty_like(nw_st)
            ctx.newey_west_by_time_range(T_all, K, NW_Q, NW_TAO, fr_all, lo, lo + T, nw_ref, st_ref)
            okm = st_ref == 0
            dmax = (nw[okm] - nw_ref[okm]).abs().max() / nw_ref[okm].abs().max()
            same = torch.equal(st_ref, nw_st) and bool(dmax.item() < 1e-12)
            ok_t = torch.tensor([1 if same else 0], device=dev)
            dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
            gather_check += "; sharded Newey-West == whole-series pass (<1e-12 rel)" if ok_t.item() else "; sharded Newey-West MISMATCH"

    # ---- end to end: pinned host buffers -> device -> results back in pinned host buffers ----
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())

    # Steps are pipelined the way a host feeding a stream of panels would: inputs are double-buffered on the
    # device and copied on their own stream, so the H2D copy of step i+1 overlaps the kernels and the D2H
    # copy of step i (PCIe is full duplex).  Every step still moves all of its inputs and results.
    s_h2d = torch.cuda.Stream(device=dev)
    s_d2h = torch.cuda.Stream(device=dev)
    d_in = [d, {k: torch.empty_like(v) for k, v in d.items()}]
    ev_in = [torch.cuda.Event() for _ in range(2)]      # inputs of buffer b have landed
    ev_free = [torch.cuda.Event() for _ in range(2)]    # the kernels have finished reading buffer b
    ev_done, ev_out, ev_go = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()

    def e2e_step(i):
        b = i & 1
        with torch.cuda.stream(s_h2d):
            s_h2d.wait_event(ev_free[b])
            for k in host:
                d_in[b][k].copy_(host[k], non_blocking=True)
            ev_in[b].record(s_h2d)
        gs.wait_event(ev_in[b])
        gs.wait_event(ev_out)                 # the previous step's results have left the output buffers
        db = d_in[b]
        ctx.reg_by_time(T, P, Q, N_STOCKS, db["off"], db["cap"], db["ret"], db["ind"], db["style"], fr, spec, r2, st)
        if world > 1 and not nocoll:
            dist.all_gather_into_tensor(fr_all, fr)
        ctx.newey_west_by_time_range(T_all, K, NW_Q, NW_TAO, fr_all, lo, lo + T, nw, nw_st)
        vol_regime()
        ev_free[b].record(gs)
        ev_done.record(gs)
        with torch.cuda.stream(s_d2h):
            s_d2h.wait_event(ev_done)
            out_host["fr"].copy_(fr, non_blocking=True)
            out_host["spec"].copy_(spec, non_blocking=True)
            out_host["r2"].copy_(r2, non_blocking=True)
            out_host["vr"].copy_(vr, non_blocking=True)
            out_host["lamb"].copy_(lamb[lo:lo + T], non_blocking=True)
            ev_out.record(s_d2h)

    def e2e_run(nsteps):
        ev_go.record(gs)
        s_h2d.wait_event(ev_go)
        s_d2h.wait_event(ev_go)
        for i in range(nsteps):
            e2e_step(i)
        gs.wait_event(ev_out)

    e2e_run(max(2, args.warmup // 2))
    barrier()
    n_e2e = max(4, min(args.steps, 12))
    t_start.record()
    e2e_run(n_e2e)
    t_end.record()
    barrier()
    tm = torch.tensor([t_start.elapsed_time(t_end)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    e2e_value = T_all / (float(tm.item()) / n_e2e * 1e-3)

    # ---- eigen Monte-Carlo block (bounded slice of C4), sims sharded over ranks ----
    from bmfm import ops
    # weak scaling like the main step: every rank runs EIG_SIMS simulations (C4 spreads 10 000 over the ranks,
    # 1250 each at 8 GPUs), one all-reduce of the T x K accumulators
    sims_per_rank = EIG_SIMS
    m_lo = rank * sims_per_rank
    covs = nw[T - EIG_DATES:].contiguous()      # the same rank-local dates on every rank: only the sims are sharded
    csts = nw_st[T - EIG_DATES:].contiguous()
    if world > 1:
        dist.broadcast(covs, 0)
        dist.broadcast(csts, 0)
    D0 = torch.empty((EIG_DATES, K), dtype=torch.float64, device=dev)
    U0 = torch.empty((EIG_DATES, K, K), dtype=torch.float64, device=dev)
    est = torch.empty(EIG_DATES, dtype=torch.int32, device=dev)
    vsum = torch.zeros((EIG_DATES, K), dtype=torch.float64, device=dev)
    eout = torch.empty((EIG_DATES, K, K), dtype=torch.float64, device=dev)
    Cm = ops.gauss_cov_device(ctx, m_lo, sims_per_rank, EIG_TSIM, K)

    def eig_step():
        ctx.eigen_decompose(EIG_DATES, K, covs, csts, D0, U0, est)
        ctx.eigen_mc_accumulate(EIG_DATES, K, sims_per_rank, Cm, D0, est, vsum, False)
        if world > 1:
            dist.all_reduce(vsum)
        ctx.eigen_finalize(EIG_DATES, K, sims_per_rank * world, 1.4, D0, U0, vsum, est, eout)

    eig_step()
    barrier()
    n_eig = 3
    t_start.record()
    for _ in range(n_eig):
        eig_step()
    t_end.record()
    barrier()
    tm = torch.tensor([t_start.elapsed_time(t_end)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    eig_value = EIG_DATES * sims_per_rank * world / (float(tm.item()) / n_eig * 1e-3)
    eig_ok = bool((est == 0).all().item())

    if rank == 0:
        peak, peak_src = load_peaks()
        achieved = BYTES_PER_STOCK_DATE * rows / (ms_reg * 1e-3) / 1e9
        dominant = max((ms_reg, "reg_by_time (reg_mma_kernel)"), (ms_nw, "newey_west (nw_emit)"), (ms_vr, "vol_regime (vr_scale_kernel)"))[1]
        t0 = time.perf_counter()
        cpu_v, cpu_desc = cpu_sample(csr, 600)
        cpu_eig = cpu_eigen_sample(16)
        log(f"cpu baseline took {time.perf_counter() - t0:.1f}s")
        line = {
            "metric": "regression_dates_per_sec", "value": value, "unit": "dates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"C2 synthetic CSI800 panel {N_STOCKS}x{N_DATES} per GPU, P={P}, Q={Q}: reg_by_time + "
                                   f"Newey_West_by_time(q={NW_Q},tao={int(NW_TAO)}) + vol_regime_adj_by_time(tao={int(VR_TAO)})",
                       "parallelism": (f"dates sharded over {world} GPU(s); factor returns and vol-regime terms gathered "
                                       + ("by the producing kernels over peer memory (symmetric memory + device barrier)" if fused
                                          else "with NCCL all-gather")) if world > 1 else "1 GPU",
                       "timed_as": "CUDA graph replays of one captured step (phase_ms from 3 extra eager steps after the warm-up)" if graph is not None else "eager launches",
                       "l2": f"inputs larger than L2: panel {h2d / 1e6:.0f} MB + {2 * T * K * K * 8 / 1e6:.0f} MB of covariance output per GPU per step vs 126 MB L2",
                       "phase_ms": {"reg_by_time": ms_reg, "newey_west": ms_nw, "vol_regime": ms_vr},
                       "all_status_ok": status_ok, "gather_check": gather_check},
            "roofline": {"kernel": "reg_mma_kernel<10,3,1> (per-date weighted Gram on FP64 MMA + constrained solve + residuals, one warp per date)",
                         "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": load_traffic(), "traffic_source": "ncu --set full capture of one launch, profiles/roofline_traffic.json",
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": BYTES_PER_STOCK_DATE * rows,
                         "dominant_phase": dominant},
            "cpu_baseline": {"value": cpu_v, "unit": "dates/s", "cores": blas_threads(), "kind": "port", "sample": cpu_desc},
            "e2e": {"value": e2e_value, "unit": "dates/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "how": "pinned host buffers; every step copies all inputs in and all results out; inputs "
                           "double-buffered so the H2D copy of step i+1 overlaps the kernels and D2H of step i"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "eigen_mc": {"value": eig_value, "unit": "sims/s",
                         "config": f"K={K}, T_sim={EIG_TSIM}, {EIG_DATES} dates x {sims_per_rank * world} sims ({sims_per_rank} per rank, sims sharded over ranks, one all-reduce), canonical eigen order",
                         "all_status_ok": eig_ok,
                         "cpu_baseline": {"value": cpu_eig, "unit": "sims/s", "kind": "port", "sample": "16 sims of one date"}},
        }
        emit(line)
    if world > 1:
        # leave without tearing NCCL down: destroying the process group while a captured graph still holds
        # NCCL work was seen to hang at exit; everything is flushed and the numbers are out
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    claim_stdout()
    if args.warmup < 3 and args.impl != "reference":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
