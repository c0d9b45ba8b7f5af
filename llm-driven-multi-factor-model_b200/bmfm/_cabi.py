"""ctypes binding of libbmfm.so (the C-ABI declared in include/bmfm.h).

There is no CPU fallback: importing this module without the built library, or
creating a context without a CUDA device, raises.  PyTorch is used only to own
device memory and streams; the signatures below are plain pointers and sizes.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbmfm.so")

# status codes (include/bmfm.h)
ST_OK, ST_SHORT, ST_NOT_PSD, ST_NONFINITE, ST_PINV, ST_DUP_EDGES = 0, 1, 2, 3, 4, 5
OP_LOG, OP_INV_POS, OP_PERCENT = 0, 1, 2
OP2_RATIO_POS, OP2_LEV_FINITE, OP2_LEV_POS = 10, 11, 12

_lib = None
_lock = threading.Lock()

_P = C.c_void_p
_I = C.c_int
_D = C.c_double

_SIGNATURES = {
    "bmfm_ctx_create": ([C.POINTER(_P), _I, _P], _I),
    "bmfm_ctx_destroy": ([_P], _I),
    "bmfm_ctx_sync": ([_P], _I),
    "bmfm_last_error": ([_P], C.c_char_p),
    "bmfm_launch_count": ([_P], C.c_longlong),
    "bmfm_ctx_set_option": ([_P, C.c_char_p, _I], _I),
    "bmfm_ctx_get_option": ([_P, C.c_char_p, C.POINTER(_I)], _I),
    "bmfm_ctx_set_peer_flags": ([_P, _I, _I, _P], _I),
    "bmfm_limits": ([C.POINTER(_I), C.POINTER(_I), C.POINTER(_I)], None),
    "bmfm_version": ([], C.c_char_p),
    "bmfm_reg_by_time": ([_P, _I, _I, _I, _I] + [_P] * 10, _I),
    "bmfm_reg_by_time_gather": ([_P, _I, _I, _I, _I] + [_P] * 9 + [_I, _P, C.c_int64], _I),
    "bmfm_vol_regime_bias_gather": ([_P, _I, _I, _I, _I, _D, _P, _P, _P, _I, _P, _P], _I),
    "bmfm_reg_by_time_host": ([_P, _I, _I, _I] + [_P] * 10, _I),
    "bmfm_onehot_to_index": ([_P, C.c_int64, _I, _P, C.c_int64, C.c_int64, _P, _P], _I),
    "bmfm_pack_gather": ([_P, C.c_int64, _I, _P, _P, _P, _P, _P, C.c_int64, C.c_int64, _P, _P, _P, _P], _I),
    "bmfm_risk_step_host": ([_P, _I, _I, _I] + [_P] * 5 + [_I, _D, _D] + [_P] * 8, _I),
    "bmfm_newey_west_by_time": ([_P, _I, _I, _I, _D, _P, _P, _P], _I),
    "bmfm_newey_west_scratch_bytes": ([_I, _I, _I], C.c_size_t),
    "bmfm_newey_west_by_time_range": ([_P, _I, _I, _I, _D, _P, _I, _I, _P, _P], _I),
    "bmfm_newey_west_state_doubles": ([_I, _I], C.c_size_t),
    "bmfm_newey_west_shard_totals": ([_P, _I, _I, _I, _D, _P, _I, _I, _I, _I, _P], _I),
    "bmfm_newey_west_shard_emit": ([_P, _I, _I, _I, _D, _P, _I, _I, _I, _P, _P, _P, _P], _I),
    "bmfm_vol_regime_bias": ([_P, _I, _I, _I, _I, _D, _P, _P, _P, _P, _P], _I),
    "bmfm_vol_regime_scale": ([_P, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P], _I),
    "bmfm_vol_regime_shard_totals": ([_P, _I, _I, _I, _I, _D, _P, _P, _P, _P, _P, _I, _I, _P], _I),
    "bmfm_vol_regime_shard_scale": ([_P, _I, _I, _P, _P, _P, _I, _P, _P, _P, _P], _I),
    "bmfm_newey_west_by_time_noalloc": ([_P, _I, _I, _I, _D, _P, _P, _P], _I),
    "bmfm_vol_regime_adj_by_time": ([_P, _I, _I, _D, _P, _P, _P, _P, _P], _I),
    "bmfm_vol_regime_adj_range": ([_P, _I, _I, _I, _I, _D, _P, _P, _P, _P, _P, _P, _P], _I),
    "bmfm_gauss_cov": ([_P, _I, _I, _I, _I, _P], _I),
    "bmfm_eigen_decompose": ([_P, _I, _I, _P, _P, _P, _P, _P], _I),
    "bmfm_eigen_mc_accumulate": ([_P, _I, _I, _I, _P, _P, _P, _P, _I], _I),
    "bmfm_eigen_mc_accumulate_perm": ([_P, _I, _I, _I, _P, _P, _P, _P, _P, _I], _I),
    "bmfm_eigen_finalize": ([_P, _I, _I, _I, _D, _P, _P, _P, _P, _P], _I),
    "bmfm_eigenfactor_bias_stat": ([_P, _I, _I, _I, _P, _P, _P, _P, _P], _I),
    "bmfm_bayes_shrink": ([_P, _I, _I, _P, _P, _P, _I, _D, _P, _P], _I),
    "bmfm_prepare_returns": ([_P, _I, _I, _P, _P, _P], _I),
    "bmfm_rstr": ([_P, _I, _I, _P, _I, _I, _D, _I, _P], _I),
    "bmfm_rolling_log_sum": ([_P, _I, _I, _P, _D, _I, _I, _P], _I),
    "bmfm_rolling_log_sum3": ([_P, _I, _I, _P, _D, _P, _P, _P], _I),
    "bmfm_nlsize": ([_P, _I, _P, _P, _P], _I),
    "bmfm_elementwise": ([_P, _I, C.c_int64, _P, _P], _I),
    "bmfm_elementwise2": ([_P, _I, C.c_int64, _P, _P, _P], _I),
    "bmfm_rolling_sum": ([_P, _I, _I, _P, _I, _I, _P], _I),
    "bmfm_beta_hsigma": ([_P, _I, _I, _P, _P, _I, _D, _I, _P, _P], _I),
    "bmfm_dastd": ([_P, _I, _I, _P, _P, _I, _D, _I, _P], _I),
    "bmfm_cmra": ([_P, _I, _I, _P, _I, _P], _I),
    "bmfm_winsorize": ([_P, _I, _P, _P, _D, _P], _I),
    "bmfm_composite": ([_P, C.c_int64, _I, _P, _P, _P], _I),
    "bmfm_orthogonalize": ([_P, _I, _P, _P, _I, _P, _P], _I),
    "bmfm_winsorize_sums": ([_P, _I, _P, _P, _P], _I),
    "bmfm_winsorize_sqdev": ([_P, _I, _P, _P, _P, _P], _I),
    "bmfm_winsorize_apply": ([_P, _I, _P, _P, _P, _P, _P], _I),
    "bmfm_orthogonalize_pass": ([_P, _I, _I, _P, _P, _I, _P, _P, _P, _P, _P], _I),
}


def exported_symbols():
    """Every symbol include/bmfm.h declares (checked by the CPU test-suite)."""
    return sorted(_SIGNATURES)


def lib():
    """Load libbmfm.so once; raise loudly if it has not been built."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: build it with "
                    "`python llm-driven-multi-factor-model_b200/csrc/build.py` "
                    "(there is no CPU fallback)")
            L = C.CDLL(LIB_PATH)
            for name, (argt, rest) in _SIGNATURES.items():
                fn = getattr(L, name)   # AttributeError if the .so is stale
                fn.argtypes = argt
                fn.restype = rest
            _lib = L
    return _lib


class BmfmError(RuntimeError):
    pass


def ptr(t):
    """Device (or host) address of a torch tensor / numpy array, or NULL."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        assert t.is_contiguous()
        return C.c_void_p(t.data_ptr())
    assert t.flags["C_CONTIGUOUS"]
    return C.c_void_p(t.ctypes.data)


def _raw(t):
    """Address of a (possibly strided) torch tensor, or NULL."""
    return None if t is None else C.c_void_p(t.data_ptr())


class Context:
    """One bmfm_ctx bound to a CUDA device and (by default) torch's current stream."""

    def __init__(self, device=None, stream="torch"):
        import torch

        if not torch.cuda.is_available():
            raise RuntimeError("bmfm needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.device = torch.cuda.current_device() if device is None else int(device)
        L = lib()
        sp = None
        if stream == "torch":
            with torch.cuda.device(self.device):
                sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        elif stream is not None:
            sp = C.c_void_p(int(stream))
        h = C.c_void_p()
        rc = L.bmfm_ctx_create(C.byref(h), self.device, sp)
        if rc != 0:
            raise BmfmError(f"bmfm_ctx_create failed: {rc}")
        self._h = h
        self._L = L

    def check(self, rc):
        if rc != 0:
            raise BmfmError(f"libbmfm error {rc}: {self._L.bmfm_last_error(self._h).decode()}")

    def sync(self):
        self.check(self._L.bmfm_ctx_sync(self._h))

    def set_peer_flags(self, rank, flag_ptrs):
        """Register the ranks' flag blocks (device addresses, 64 zeroed int64 each, peer-mapped): the sharded
        Newey-West / vol-regime calls then synchronise through them instead of caller-side barriers."""
        n = len(flag_ptrs)
        arr = (C.c_void_p * max(1, n))(*[int(x) for x in flag_ptrs])
        self.check(self._L.bmfm_ctx_set_peer_flags(self._h, int(rank), n, arr))

    def set_option(self, name, value):
        """Experiment switch of this context (include/bmfm.h: bmfm_ctx_set_option)."""
        self.check(self._L.bmfm_ctx_set_option(self._h, name.encode(), int(value)))

    def get_option(self, name):
        v = C.c_int(0)
        self.check(self._L.bmfm_ctx_get_option(self._h, name.encode(), C.byref(v)))
        return int(v.value)

    @property
    def launches(self):
        return int(self._L.bmfm_launch_count(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._L.bmfm_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ part 1
    def reg_by_time(self, T, P, Q, max_rows, date_off, cap, ret, ind, style,
                    factor_ret, spec, r2, status, exposure=None):
        self.check(self._L.bmfm_reg_by_time(self._h, T, P, Q, max_rows, ptr(date_off), ptr(cap), ptr(ret),
                                            ptr(ind), ptr(style), ptr(factor_ret), ptr(spec), ptr(r2),
                                            ptr(exposure), ptr(status)))

    def reg_by_time_gather(self, T, P, Q, max_rows, date_off, cap, ret, ind, style, factor_ret, spec, r2, status,
                           peer_ptrs, row0):
        """Regression + fused all-gather: `peer_ptrs` = device addresses (ints) of every rank's gathered
        [T_total, K] buffer (peer-mapped, e.g. symmetric memory), written at row row0 + t."""
        arr = (C.c_void_p * len(peer_ptrs))(*[int(x) for x in peer_ptrs])
        self.check(self._L.bmfm_reg_by_time_gather(self._h, T, P, Q, max_rows, ptr(date_off), ptr(cap), ptr(ret),
                                                   ptr(ind), ptr(style), ptr(factor_ret), ptr(spec), ptr(r2),
                                                   ptr(status), len(peer_ptrs), arr, int(row0)))

    def vol_regime_bias_gather(self, T_total, t0, n, K, tao, fr_block, cov_block, status_block, peer_wb, peer_wv):
        a = (C.c_void_p * len(peer_wb))(*[int(x) for x in peer_wb])
        b = (C.c_void_p * len(peer_wv))(*[int(x) for x in peer_wv])
        self.check(self._L.bmfm_vol_regime_bias_gather(self._h, T_total, t0, n, K, float(tao), ptr(fr_block),
                                                       ptr(cov_block), ptr(status_block), len(peer_wb), a, b))

    def reg_by_time_host(self, T, P, Q, date_off, cap, ret, ind, style,
                         factor_ret, spec, r2, status, exposure=None):
        self.check(self._L.bmfm_reg_by_time_host(self._h, T, P, Q, ptr(date_off), ptr(cap), ptr(ret),
                                                 ptr(ind), ptr(style), ptr(factor_ret), ptr(spec), ptr(r2),
                                                 ptr(exposure), ptr(status)))


    def onehot_to_index(self, rows, P, onehot, row_stride, col_stride, ind, flags):
        self.check(self._L.bmfm_onehot_to_index(self._h, int(rows), int(P), _raw(onehot), int(row_stride), int(col_stride),
                                                ptr(ind), ptr(flags)))

    def pack_gather(self, rows, Q, order, cap, ret, ind, style, srs, scs, cap_o, ret_o, ind_o, style_o):
        self.check(self._L.bmfm_pack_gather(self._h, int(rows), int(Q), ptr(order), ptr(cap), ptr(ret), ptr(ind), _raw(style),
                                            int(srs), int(scs), ptr(cap_o), ptr(ret_o), ptr(ind_o), ptr(style_o)))

    def risk_step_host(self, T, P, Q, date_off, cap, ret, ind, style, q, tao_nw, tao_vr,
                       factor_ret, spec, r2, status, nw_cov, nw_status, vr_cov, lamb):
        """reg_by_time + Newey_West_by_time + vol-regime on HOST buffers, copies overlapped with the kernels."""
        self.check(self._L.bmfm_risk_step_host(self._h, T, P, Q, ptr(date_off), ptr(cap), ptr(ret), ptr(ind), ptr(style),
                                               int(q), float(tao_nw), float(tao_vr), ptr(factor_ret), ptr(spec), ptr(r2),
                                               ptr(status), ptr(nw_cov), ptr(nw_status), ptr(vr_cov), ptr(lamb)))

    # ------------------------------------------------------------------ part 3
    def newey_west_by_time(self, T, K, q, tao, fr, cov, status):
        self.check(self._L.bmfm_newey_west_by_time(self._h, T, K, q, float(tao), ptr(fr), ptr(cov), ptr(status)))

    def newey_west_by_time_range(self, T, K, q, tao, fr, t0, t1, cov, status):
        self.check(self._L.bmfm_newey_west_by_time_range(self._h, T, K, q, float(tao), ptr(fr), t0, t1, ptr(cov),
                                                         ptr(status)))

    def newey_west_state_doubles(self, K, q):
        return int(self._L.bmfm_newey_west_state_doubles(K, q))

    def newey_west_shard_totals(self, T_total, K, q, tao, fr_all, t0, t1, rank, peer_totals):
        """Sharded Newey-West, step 1: reduce this rank's dates and store its end-of-range state into slot
        `rank` of every buffer in `peer_totals` (device addresses of [n_rank, state_doubles] buffers)."""
        arr = (C.c_void_p * len(peer_totals))(*[int(x) for x in peer_totals])
        self.check(self._L.bmfm_newey_west_shard_totals(self._h, T_total, K, q, float(tao), ptr(fr_all), t0, t1,
                                                        rank, len(peer_totals), arr))

    def newey_west_shard_emit(self, T_total, K, q, tao, fr_all, t0, t1, rank, range_ends, totals, cov, status):
        """Step 2 (after a cross-rank barrier): covariances of dates [t0, t1) with the earlier ranks' totals
        as carry-in.  `range_ends[r]` = one past the last date of rank r (host ints, ranks before this one)."""
        ends = (C.c_int64 * max(1, len(range_ends)))(*[int(x) for x in range_ends])
        self.check(self._L.bmfm_newey_west_shard_emit(self._h, T_total, K, q, float(tao), ptr(fr_all), t0, t1, rank,
                                                      ends, ptr(totals), ptr(cov), ptr(status)))

    def vol_regime_bias(self, T_total, t0, n, K, tao, fr_block, cov_block, status_block, wb, wv):
        self.check(self._L.bmfm_vol_regime_bias(self._h, T_total, t0, n, K, float(tao), ptr(fr_block),
                                                ptr(cov_block), ptr(status_block), ptr(wb), ptr(wv)))

    def vol_regime_scale(self, T_total, t0, n, K, wb_all, wv_all, cov_block, status_block, out_block, lamb_all):
        self.check(self._L.bmfm_vol_regime_scale(self._h, T_total, t0, n, K, ptr(wb_all), ptr(wv_all),
                                                 ptr(cov_block), ptr(status_block), ptr(out_block), ptr(lamb_all)))

    def vol_regime_shard_totals(self, T_total, t0, n, K, tao, fr_block, cov_block, status_block, wb, wv, rank, peer_totals):
        arr = (C.c_void_p * len(peer_totals))(*[int(x) for x in peer_totals])
        self.check(self._L.bmfm_vol_regime_shard_totals(self._h, T_total, t0, n, K, float(tao), ptr(fr_block),
                                                        ptr(cov_block), ptr(status_block), ptr(wb), ptr(wv), rank,
                                                        len(peer_totals), arr))

    def vol_regime_shard_scale(self, n, K, wb, wv, totals, rank, cov_block, status_block, out_block, lamb_block):
        self.check(self._L.bmfm_vol_regime_shard_scale(self._h, n, K, ptr(wb), ptr(wv), ptr(totals), rank, ptr(cov_block),
                                                       ptr(status_block), ptr(out_block), ptr(lamb_block)))

    def newey_west_by_time_noalloc(self, T, K, q, tao, fr, cov, status):
        """Newey-West with the library's built-in fixed scratch: for callers that may not allocate
        device memory at call time.  Refuses sizes beyond the fixed array (BmfmError)."""
        self.check(self._L.bmfm_newey_west_by_time_noalloc(self._h, T, K, q, float(tao), ptr(fr), ptr(cov),
                                                           ptr(status)))

    def vol_regime_adj_by_time(self, T, K, tao, fr, cov, cov_status, out_cov, lamb):
        self.check(self._L.bmfm_vol_regime_adj_by_time(self._h, T, K, float(tao), ptr(fr), ptr(cov),
                                                       ptr(cov_status), ptr(out_cov), ptr(lamb)))

    def vol_regime_adj_range(self, T, t0, t1, K, tao, fr, cov, cov_status, wb, wv, out_cov, lamb):
        """Dates [t0, t1) of vol_regime_adj_by_time; wb / wv [T] persist between the calls (full-series pointers)."""
        self.check(self._L.bmfm_vol_regime_adj_range(self._h, T, t0, t1, K, float(tao), ptr(fr), ptr(cov),
                                                     ptr(cov_status), ptr(wb), ptr(wv), ptr(out_cov), ptr(lamb)))

    def eigenfactor_bias_stat(self, T, K, predlen, cov, cov_status, fr, bias_stat, b_out=None):
        self.check(self._L.bmfm_eigenfactor_bias_stat(self._h, T, K, int(predlen), ptr(cov), ptr(cov_status), ptr(fr),
                                                      ptr(bias_stat), ptr(b_out)))

    def gauss_cov(self, m0, nsim, T_sim, K, Cmat):
        self.check(self._L.bmfm_gauss_cov(self._h, m0, nsim, T_sim, K, ptr(Cmat)))

    def eigen_decompose(self, T, K, cov, cov_status, D0, U0, status):
        self.check(self._L.bmfm_eigen_decompose(self._h, T, K, ptr(cov), ptr(cov_status), ptr(D0), ptr(U0),
                                                ptr(status)))

    def eigen_mc_accumulate(self, T, K, nsim, Cmat, D0, status, vsum, accumulate=False):
        self.check(self._L.bmfm_eigen_mc_accumulate(self._h, T, K, nsim, ptr(Cmat), ptr(D0), ptr(status),
                                                    ptr(vsum), int(bool(accumulate))))

    def eigen_mc_accumulate_perm(self, T, K, nsim, Cmat, D0, status, perm, vsum, accumulate=False):
        self.check(self._L.bmfm_eigen_mc_accumulate_perm(self._h, T, K, nsim, ptr(Cmat), ptr(D0), ptr(status),
                                                         ptr(perm), ptr(vsum), int(bool(accumulate))))

    def eigen_finalize(self, T, K, M, scale_coef, D0, U0, vsum, status, out):
        self.check(self._L.bmfm_eigen_finalize(self._h, T, K, M, float(scale_coef), ptr(D0), ptr(U0), ptr(vsum),
                                               ptr(status), ptr(out)))


    def bayes_shrink(self, T, max_rows, date_off, vol, cap, ngroup, q, out, status):
        self.check(self._L.bmfm_bayes_shrink(self._h, T, max_rows, ptr(date_off), ptr(vol), ptr(cap), int(ngroup),
                                             float(q), ptr(out), ptr(status)))

    # ------------------------------------------------------------------ part 2
    def prepare_returns(self, L, S, close, ret, log_ret):
        self.check(self._L.bmfm_prepare_returns(self._h, L, S, ptr(close), ptr(ret), ptr(log_ret)))

    def beta_hsigma(self, L, S, ret, mkt, window, half_life, min_periods, beta, hsigma):
        self.check(self._L.bmfm_beta_hsigma(self._h, L, S, ptr(ret), ptr(mkt), int(window), float(half_life),
                                            int(min_periods), ptr(beta), ptr(hsigma)))

    def dastd(self, L, S, ret, mkt, window, half_life, min_periods, out):
        self.check(self._L.bmfm_dastd(self._h, L, S, ptr(ret), ptr(mkt), int(window), float(half_life),
                                      int(min_periods), ptr(out)))

    def cmra(self, L, S, log_ret, window, out):
        self.check(self._L.bmfm_cmra(self._h, L, S, ptr(log_ret), int(window), ptr(out)))

    def rstr(self, L, S, log_ret, T, lag, half_life, min_periods, out):
        self.check(self._L.bmfm_rstr(self._h, L, S, ptr(log_ret), int(T), int(lag), float(half_life),
                                     int(min_periods), ptr(out)))

    def rolling_log_sum(self, L, S, x, divisor, window, min_periods, out):
        self.check(self._L.bmfm_rolling_log_sum(self._h, L, S, ptr(x), float(divisor), int(window),
                                                int(min_periods), ptr(out)))

    def rolling_log_sum3(self, L, S, x, divisor, windows, min_periods, outs):
        w = (C.c_int * 3)(*[int(v) for v in windows])
        m = (C.c_int * 3)(*[int(v) for v in min_periods])
        o = (C.c_void_p * 3)(*[int(t.data_ptr()) for t in outs])
        self.check(self._L.bmfm_rolling_log_sum3(self._h, L, S, ptr(x), float(divisor), w, m, o))

    def elementwise(self, op, n, x, out):
        self.check(self._L.bmfm_elementwise(self._h, int(op), int(n), ptr(x), ptr(out)))

    def elementwise2(self, op, n, x, y, out):
        self.check(self._L.bmfm_elementwise2(self._h, int(op), int(n), ptr(x), ptr(y), ptr(out)))

    def rolling_sum(self, L, S, x, window, min_periods, out):
        self.check(self._L.bmfm_rolling_sum(self._h, L, S, ptr(x), int(window), int(min_periods), ptr(out)))

    def nlsize(self, T, date_off, size, out):
        self.check(self._L.bmfm_nlsize(self._h, T, ptr(date_off), ptr(size), ptr(out)))

    def winsorize(self, T, date_off, x, n_std, out):
        self.check(self._L.bmfm_winsorize(self._h, T, ptr(date_off), ptr(x), float(n_std), ptr(out)))

    def composite(self, rows, comps, weights, out):
        n = len(comps)
        arr = (C.c_void_p * n)(*[t.data_ptr() for t in comps])
        w = (C.c_double * n)(*[float(x) for x in weights])
        self.check(self._L.bmfm_composite(self._h, rows, n, arr, w, ptr(out)))

    def winsorize_sums(self, T, date_off, x, sums):
        self.check(self._L.bmfm_winsorize_sums(self._h, T, ptr(date_off), ptr(x), ptr(sums)))

    def winsorize_sqdev(self, T, date_off, x, mean, sqdev):
        self.check(self._L.bmfm_winsorize_sqdev(self._h, T, ptr(date_off), ptr(x), ptr(mean), ptr(sqdev)))

    def winsorize_apply(self, T, date_off, x, lo, hi, out):
        self.check(self._L.bmfm_winsorize_apply(self._h, T, ptr(date_off), ptr(x), ptr(lo), ptr(hi), ptr(out)))

    def orthogonalize_pass(self, which, T, date_off, y, Xs, out, shift=None, gram=None, count=None):
        """Stock-sharded orthogonalize: pass 0 (sums), 1 (Gram about `shift`), 2 (residuals from gram / count)."""
        k = len(Xs)
        arr = (C.c_void_p * k)(*[t.data_ptr() for t in Xs])
        self.check(self._L.bmfm_orthogonalize_pass(self._h, int(which), T, ptr(date_off), ptr(y), k, arr, ptr(shift),
                                                   ptr(gram), ptr(count), ptr(out)))

    def orthogonalize(self, T, date_off, y, Xs, out):
        k = len(Xs)
        arr = (C.c_void_p * k)(*[t.data_ptr() for t in Xs])
        self.check(self._L.bmfm_orthogonalize(self._h, T, ptr(date_off), ptr(y), k, arr, ptr(out)))


_default_ctx = {}


def default_context():
    """The context of (current device, torch's CURRENT stream), created on first use.

    A context enqueues on the stream it was created with, while the wrappers in ops.py / mfm allocate, upload
    and read back on torch's current stream at call time; keying the cache on the stream keeps kernels and
    copies on one stream whatever ``torch.cuda.stream(...)`` block the caller is in."""
    import torch

    if not torch.cuda.is_available():
        return Context(None)        # raises: there is no CPU fallback
    dev = torch.cuda.current_device()
    key = (dev, int(torch.cuda.current_stream().cuda_stream))
    ctx = _default_ctx.get(key)
    if ctx is None:
        if len(_default_ctx) >= 16:      # side streams come and go: keep the table small
            for k in [k for k in _default_ctx if k[0] == dev and k != (dev, 0)][:8]:
                _default_ctx.pop(k).close()
        ctx = _default_ctx[key] = Context(dev)
    return ctx
