/*
 * bmfm.h — C-ABI of the B200-native Barra risk-model path (libbmfm.so).
 *
 * The reference (Izumighj/LLM-Driven-Multi-factor-Model) is pure Python and
 * has no FFI layer; its boundary is the Python API of Barra-master/mfm and
 * Barra_factor_cal.  Each entry point below replaces the
 * arithmetic of one reference function and is what a ctypes binding inside
 * that function would call (INTEGRATION.md shows the stub).  Citations are
 * file:line under the reference tree.
 *
 * Conventions
 *   - every function returns 0 on success, <0 for a bad argument
 *     (BMFM_E_*), >0 = a cudaError_t from the CUDA runtime;
 *     bmfm_last_error(ctx) gives the message.
 *   - `*_dev` style pointers (no suffix) are DEVICE pointers owned by the
 *     caller; the `_host` variants take HOST pointers and do the H2D/D2H copies
 *     themselves on the context's stream (pinned memory recommended).
 *   - all work is enqueued on the context's stream and is asynchronous for the
 *     device-pointer entry points; `_host` variants synchronise before return.
 *   - fp64 throughout; industry membership is an int32 index per row
 *     (-1 or >= P: the row has no industry column set).
 *   - one bmfm_ctx per host thread; a context is bound to one device.
 */
#ifndef BMFM_H
#define BMFM_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bmfm_ctx bmfm_ctx;

enum {
  BMFM_OK = 0,
  BMFM_E_ARG = -1,          /* NULL pointer / negative size */
  BMFM_E_UNSUPPORTED = -2,  /* outside compiled limits (see bmfm_limits) */
  BMFM_E_NOMEM = -3,
  BMFM_E_STATE = -4
};

/* per-date status written by the *_by_time entry points */
enum {
  BMFM_ST_OK = 0,
  BMFM_ST_SHORT = 1,     /* t <= q or t <= K: the reference raises, caller stores an empty DataFrame (utils.py:29, MFM.py:95-96) */
  BMFM_ST_NOT_PSD = 2,   /* eigen_risk_adj PSD gate failed (utils.py:67-68) -> empty DataFrame */
  BMFM_ST_NONFINITE = 3, /* non-finite Gram (e.g. last industry empty, CrossSection.py:70) -> numpy raises LinAlgError */
  BMFM_ST_PINV = 4,      /* informational: rank-deficient Gram, eigen-based pinv path was used (CrossSection.py:76) */
  BMFM_ST_DUP_EDGES = 5  /* bayes_shrink: duplicate decile edges, pd.qcut raises ValueError (utils.py:161) */
};

/* ---- context ------------------------------------------------------------ */
/* stream: a cudaStream_t (as void*) to enqueue on; NULL = the default stream. */
int bmfm_ctx_create(bmfm_ctx** out, int device, void* stream);
int bmfm_ctx_destroy(bmfm_ctx* ctx);
int bmfm_ctx_sync(bmfm_ctx* ctx);
const char* bmfm_last_error(const bmfm_ctx* ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
long long bmfm_launch_count(const bmfm_ctx* ctx);
/* Experiment switches of ONE context (no counterpart in the reference; they select alternative kernels
 * for comparisons and parity cross-checks and never change results beyond rounding).  Names:
 *   "reg_mode"  0 = automatic, 1 CTA-per-date kernel, 2 register-Gram warp kernel, 3 / 5 / 4 = FP64-MMA
 *               kernel with 1 / 2 / 4 warps per date;   "reg_smma" 0/1;   "reg_warps", "reg_ctas";
 *   "nw_chunks" target number of time chunks of the Newey-West pass (0 = default);
 *   "eig_mode"  eigen Monte-Carlo solver (0 = default).
 * bmfm_ctx_get_option also reads "last_reg_mode": the kernel form the last regression call launched. */
int bmfm_ctx_set_option(bmfm_ctx* ctx, const char* name, int value);
int bmfm_ctx_get_option(bmfm_ctx* ctx, const char* name, int* value);
/* Per-producer flags for the date-sharded exchanges (multi-GPU; no counterpart in the reference).  flag_bufs: HOST
 * array of n_rank device pointers, one ZEROED block of 64 int64 per rank in peer-mapped memory (flag_bufs[rank] = this
 * rank's own).  Once registered, bmfm_newey_west_shard_totals / _emit and bmfm_vol_regime_shard_totals / _scale
 * publish their records' epochs to the later ranks and wait only for the earlier ranks, inside the kernels; the
 * caller drops its cross-rank barriers between those calls and keeps ONE all-rank barrier per step (so that no rank
 * gets a whole step ahead).  n_rank = 0 unregisters. */
int bmfm_ctx_set_peer_flags(bmfm_ctx* ctx, int rank, int n_rank, int64_t* const* flag_bufs);
/* compiled limits: max styles Q, max industries P, max factors K */
void bmfm_limits(int* max_Q, int* max_P, int* max_K);
const char* bmfm_version(void);

/* Device side of the panel packer (replaces the host passes of demo.py:33-35 and
 * MFM.py:58-61 over the long table).
 * bmfm_onehot_to_index: the P one-hot industry columns (int64, element (r, p) at onehot[r * row_stride +
 *   p * col_stride]) -> ind int32 [rows] (-1: no column set), bit-exact.  flags int32 [2] (device) receives the
 *   number of thread blocks that saw an entry other than 0 / 1 and that saw a row with more than one 1.
 * bmfm_pack_gather: out[i] = in[order[i]] for cap, ret, ind and the Q style columns (element (r, q) at
 *   style[r * style_row_stride + q * style_col_stride]); style_out row-major [rows * Q]: the layout
 *   bmfm_reg_by_time reads.  order int64 [rows] = the (date, industry, stock name) sort permutation. */
int bmfm_onehot_to_index(bmfm_ctx* ctx, int64_t rows, int P, const int64_t* onehot, int64_t row_stride,
                         int64_t col_stride, int32_t* ind, int32_t* flags);
int bmfm_pack_gather(bmfm_ctx* ctx, int64_t rows, int Q, const int64_t* order, const double* cap, const double* ret,
                     const int32_t* ind, const double* style, int64_t style_row_stride, int64_t style_col_stride,
                     double* cap_out, double* ret_out, int32_t* ind_out, double* style_out);

/* One whole step with HOST buffers (bench config C2; the e2e leg of bench.py): MFM.reg_by_time (MFM.py:48-76),
 * MFM.Newey_West_by_time (MFM.py:80-101) and the volatility-regime adjustment of MFM.vol_regime_adj_by_time
 * (MFM.py:130-167) applied to the Newey-West covariances.  Inputs as bmfm_reg_by_time_host; outputs
 * factor_ret [T*K], specific_ret [rows], r2 [T], status [T], nw_cov [T*K*K] + nw_status [T], vr_cov [T*K*K],
 * lamb [T].  The panel is copied in blocks of dates on a copy stream while earlier blocks are regressed and
 * their results copied back on another (use pinned host memory); synchronises before returning. */
int bmfm_risk_step_host(bmfm_ctx* ctx, int T, int P, int Q, const int64_t* date_off, const double* cap,
                        const double* ret, const int32_t* ind, const double* style, int q, double tao_nw,
                        double tao_vr, double* factor_ret, double* specific_ret, double* r2, int32_t* status,
                        double* nw_cov, int32_t* nw_status, double* vr_cov, double* lamb);

/* ---- part 1: cross-sectional regression ---------------------------------- */
/*
 * Replaces MFM.reg_by_time (Barra-master/mfm/MFM.py:48-76) and, per date,
 * CrossSection.__init__ + CrossSection.reg (CrossSection.py:35-108) with
 * style_factor_norm (CrossSection.py:12-20).
 *
 * Rows are grouped by date: date t owns rows [date_off[t], date_off[t+1]).
 * Any row order inside a date is correct; the kernel is fastest when a date's
 * rows are grouped by industry index.  Outputs keep the input row order.
 *
 *   date_off     int64  [T+1]
 *   cap, ret     double [rows]
 *   ind          int32  [rows]      industry index, -1 = none
 *   style        double [rows * Q]  row-major
 *   factor_ret   double [T * K]     K = 1 + P + Q, order country, industries, styles
 *   specific_ret double [rows]
 *   r2           double [T]
 *   exposure     double [T * K * K] or NULL (pure-factor-portfolio exposure, CrossSection.py:104)
 *   status       int32  [T]         BMFM_ST_*
 *   max_rows_per_date  the largest date's row count (sizes the shared-memory
 *                tile so a date is read from HBM once); 0 = unknown.
 */
int bmfm_reg_by_time(bmfm_ctx* ctx, int T, int P, int Q, int max_rows_per_date,
                     const int64_t* date_off, const double* cap, const double* ret,
                     const int32_t* ind, const double* style,
                     double* factor_ret, double* specific_ret, double* r2,
                     double* exposure, int32_t* status);

/* Same regression with the all-gather of the factor returns fused into the kernel (multi-GPU, dates
 * sharded): every date's K factor returns are also stored into each rank's gathered series over peer
 * memory — peer_bufs is a HOST array of n_peer (<= 8) device pointers to [T_total * K] buffers mapped
 * for peer access (e.g. torch symmetric memory), written at row peer_row0 + t.  Replaces the
 * all-gather after MFM.reg_by_time in a sharded run; the caller issues one cross-rank barrier. */
int bmfm_reg_by_time_gather(bmfm_ctx* ctx, int T, int P, int Q, int max_rows_per_date,
                            const int64_t* date_off, const double* cap, const double* ret,
                            const int32_t* ind, const double* style,
                            double* factor_ret, double* specific_ret, double* r2, int32_t* status,
                            int n_peer, double* const* peer_bufs, int64_t peer_row0);

/* Same, HOST pointers; copies inputs to the device, runs, copies results back. */
int bmfm_reg_by_time_host(bmfm_ctx* ctx, int T, int P, int Q,
                          const int64_t* date_off, const double* cap, const double* ret,
                          const int32_t* ind, const double* style,
                          double* factor_ret, double* specific_ret, double* r2,
                          double* exposure, int32_t* status);

/* ---- part 3: risk-model adjustments -------------------------------------- */
/*
 * Replaces MFM.Newey_West_by_time (Barra-master/mfm/MFM.py:80-101) = utils.Newey_West
 * (Barra-master/mfm/utils.py:16-50) on every expanding window factor_ret[:t], t = 1..T.
 *   factor_ret double [T * K];  cov double [T * K * K] (NaN-filled where status != 0)
 *   status int32 [T]: BMFM_ST_SHORT where the reference raises (t <= q or t <= K, utils.py:29)
 * q <= 4.  Uses bmfm_newey_west_scratch_bytes(T, K, q) bytes of context workspace.
 */
int bmfm_newey_west_by_time(bmfm_ctx* ctx, int T, int K, int q, double tao,
                            const double* factor_ret, double* cov, int32_t* status);
size_t bmfm_newey_west_scratch_bytes(int T, int K, int q);
/* Same for the dates [t0, t1) only (date-sharded runs): factor_ret is still the WHOLE series
 * [T * K]; cov [(t1-t0) * K * K] and status [t1-t0] hold this block. */
int bmfm_newey_west_by_time_range(bmfm_ctx* ctx, int T, int K, int q, double tao,
                                  const double* factor_ret, int t0, int t1, double* cov,
                                  int32_t* status);
/* Date-sharded form whose cost does not grow with the number of ranks (no counterpart in the
 * reference, which is single-process; results = rows [t0, t1) of bmfm_newey_west_by_time).
 * Every rank reduces only its own dates [t0, t1) of the gathered series from a zero state
 * (rows t0-q .. t0-1 are read for the lagged products) and the ranks exchange one state record
 * each, bmfm_newey_west_state_doubles(K, q) doubles (44 KB at K = 42, q = 2):
 *   _shard_totals  stores this rank's end-of-range state into slot `rank` of every peer's totals
 *                  buffer: peer_totals = HOST array of n_peer (<= 8) device pointers to
 *                  [n_rank * state_doubles], mapped for peer access (e.g. symmetric memory);
 *   (the caller runs a cross-rank barrier)
 *   _shard_emit    builds the carry-in from `totals` (this rank's own copy) and range_ends
 *                  (HOST int64 [rank]: one past the last date of ranks 0 .. rank-1) and writes
 *                  cov [(t1-t0) * K * K], status [t1-t0].
 * Both calls use the same context; its workspace keeps the chunk states in between. */
size_t bmfm_newey_west_state_doubles(int K, int q);
int bmfm_newey_west_shard_totals(bmfm_ctx* ctx, int T_total, int K, int q, double tao,
                                 const double* factor_ret_all, int t0, int t1, int rank, int n_peer,
                                 double* const* peer_totals);
int bmfm_newey_west_shard_emit(bmfm_ctx* ctx, int T_total, int K, int q, double tao,
                               const double* factor_ret_all, int t0, int t1, int rank,
                               const int64_t* range_ends, const double* totals, double* cov,
                               int32_t* status);

/*
 * No counterpart in the reference: an addition for hosts that must not allocate device memory
 * at call time (GPU memory handed out in full at start-up to pools the library does not own;
 * loops that cannot absorb the device-wide stall of cudaMalloc/cudaFree).
 * Same arguments and results as bmfm_newey_west_by_time, but the chunk-state scratch is a fixed
 * 16 MiB array inside the library image (held by every GPU that loads the library, until exit)
 * and no path through the call allocates.  Sizes whose state does not fit are refused with
 * BMFM_E_UNSUPPORTED (K = 42, q = 2 fits for any T up to about 85 000).  The first call on a
 * device looks up the array's device address — make it before any stream capture.  Calls on one
 * device share the array: nothing in it needs clearing between calls, but calls that overlap on
 * the GPU race on it, so serialise them.
 */
int bmfm_newey_west_by_time_noalloc(bmfm_ctx* ctx, int T, int K, int q, double tao,
                                    const double* factor_ret, double* cov, int32_t* status);

/*
 * Replaces MFM.vol_regime_adj_by_time (Barra-master/mfm/MFM.py:130-167).
 *   cov [T*K*K] with cov_status [T] (NULL = all valid; dates with status != 0/4 are the
 *   reference's empty DataFrames); out_cov [T*K*K] = cov * lamb^2 (NaN where invalid; may be
 *   NULL to get lamb only); lamb [T] (the factor volatility multiplier, MFM.py:160).
 */
int bmfm_vol_regime_adj_by_time(bmfm_ctx* ctx, int T, int K, double tao, const double* factor_ret,
                                const double* cov, const int32_t* cov_status, double* out_cov,
                                double* lamb);
/* Dates [t0, t1) of the same adjustment (MFM.py:149-163) when the dates before t0 went through earlier calls with
 * the same wb / wv ([T] doubles each, device, kept by the caller between the calls; no need to clear them): what
 * bmfm_risk_step_host uses to adjust each block of dates as soon as its covariances exist.  All pointers address
 * the full series; calling it for consecutive blocks that cover [0, T) gives the rows of the call above. */
int bmfm_vol_regime_adj_range(bmfm_ctx* ctx, int T, int t0, int t1, int K, double tao, const double* factor_ret,
                              const double* cov, const int32_t* cov_status, double* wb, double* wv,
                              double* out_cov, double* lamb);
/* The same pass split for date-sharded runs (MFM.py:149-163): _bias gives this rank's block
 * (n dates from global date t0 of T_total) of weighted bias terms wb = w_t B_t^2, wv = w_t; the
 * caller gathers them over ranks in date order; _scale turns the gathered terms into lamb_all
 * [T_total] and scales this rank's block (out_cov_block may be NULL). */
int bmfm_vol_regime_bias(bmfm_ctx* ctx, int T_total, int t0, int n, int K, double tao,
                         const double* factor_ret_block, const double* cov_block,
                         const int32_t* cov_status_block, double* wb, double* wv);
/* _bias with the gather fused in: terms are stored into every rank's wb_all / wv_all [T_total] (HOST
 * arrays of n_peer peer-mapped device pointers) at offset t0. */
int bmfm_vol_regime_bias_gather(bmfm_ctx* ctx, int T_total, int t0, int n, int K, double tao,
                                const double* factor_ret_block, const double* cov_block,
                                const int32_t* cov_status_block, int n_peer, double* const* peer_wb,
                                double* const* peer_wv);
int bmfm_vol_regime_scale(bmfm_ctx* ctx, int T_total, int t0, int n, int K, const double* wb_all,
                          const double* wv_all, const double* cov_block,
                          const int32_t* cov_status_block, double* out_cov_block, double* lamb_all);
/* The same adjustment by exchanged block totals: lamb_t^2 (MFM.py:156-160) is a ratio of two prefix sums, so a rank
 * needs only TWO doubles from every earlier rank.  _shard_totals computes this rank's per-date terms wb / wv [n]
 * (kept local) and stores their sums into doubles (2 rank, 2 rank + 1) of every buffer in peer_totals (HOST array of
 * n_peer device pointers to [2 * n_rank] doubles, peer-mapped); after a cross-rank barrier _shard_scale scans this
 * rank's dates from the sum of the earlier ranks' pairs (`totals`: this rank's own buffer) and scales its block;
 * lamb_block [n]. */
int bmfm_vol_regime_shard_totals(bmfm_ctx* ctx, int T_total, int t0, int n, int K, double tao,
                                 const double* factor_ret_block, const double* cov_block,
                                 const int32_t* cov_status_block, double* wb, double* wv, int rank, int n_peer,
                                 double* const* peer_totals);
int bmfm_vol_regime_shard_scale(bmfm_ctx* ctx, int n, int K, const double* wb, const double* wv, const double* totals,
                                int rank, const double* cov_block, const int32_t* cov_status_block,
                                double* out_cov_block, double* lamb_block);

/*
 * Eigenfactor Monte-Carlo adjustment, utils.eigen_risk_adj (Barra-master/mfm/utils.py:55-92)
 * for every date, MFM.eigen_risk_adj_by_time (Barra-master/mfm/MFM.py:105-126), in four steps
 * so that simulations can be sharded across GPUs (the caller all-reduces vsum):
 *
 * bmfm_gauss_cov: C[j] = np.cov(z.T) of z = RandomState(m0+j+1).standard_normal((T_sim, K)),
 *   j < nsim — the legacy stream np.random.seed(m+1) + multivariate_normal draws from
 *   (utils.py:73-76).  C double [nsim * K * K].
 * bmfm_eigen_decompose: D0 [T*K] (descending), U0 [T*K*K] (columns, canonical sign) of every
 *   valid covariance (utils.py:64); status: input status passed through, BMFM_ST_NOT_PSD
 *   where an eigenvalue is negative (utils.py:67-68), BMFM_ST_NONFINITE for NaN input.
 * bmfm_eigen_mc_accumulate: vsum[T*K] (+)= sum over the nsim simulations in C of
 *   diag(Um' F0 Um) / Dm in descending-eigenvalue position (utils.py:79-84).
 * bmfm_eigen_finalize: v = c (sqrt(vsum / M) - 1) + 1; out = U0 diag(v^2 D0) U0' (utils.py:86-92),
 *   NaN-filled where status != 0.
 */
int bmfm_gauss_cov(bmfm_ctx* ctx, int m0, int nsim, int T_sim, int K, double* C);
int bmfm_eigen_decompose(bmfm_ctx* ctx, int T, int K, const double* cov, const int32_t* cov_status,
                         double* D0, double* U0, int32_t* status);
int bmfm_eigen_mc_accumulate(bmfm_ctx* ctx, int T, int K, int nsim, const double* C, const double* D0,
                             const int32_t* status, double* vsum, int accumulate);
int bmfm_eigen_finalize(bmfm_ctx* ctx, int T, int K, int M, double scale_coef, const double* D0,
                        const double* U0, const double* vsum, const int32_t* status, double* out);
/* Injected-literal eigen order (SURVEY.md 8(c) mode (iii)).  utils.py:64,79-84 calls the GENERAL solver
 * np.linalg.eig and averages the simulated bias by the POSITION LAPACK's dgeev left each eigenpair at, so the
 * literal output of the reference depends on that order.  This variant of bmfm_eigen_mc_accumulate takes the
 * order from the caller: D0 [T*K] descending as before, perm int32 [T * nsim * K] with perm[t][m][r] = the
 * position at which numpy's eig(Fm) returned the eigenpair of descending rank r for (date t, simulation m);
 * vsum is accumulated by position.  bmfm_eigen_finalize then takes D0 / U0 in numpy's order for F0 and the
 * result reproduces the unpatched reference.  K <= 48. */
int bmfm_eigen_mc_accumulate_perm(bmfm_ctx* ctx, int T, int K, int nsim, const double* C, const double* D0,
                                  const int32_t* status, const int32_t* perm, double* vsum, int accumulate);

/*
 * Replaces utils.eigenfactor_bias_stat (Barra-master/mfm/utils.py:97-117, used MFM.py:203-204), the
 * reference's validation metric for the eigen adjustment: for every date i < T - predlen with a covariance,
 * b_i = U' retlen_i / sigma over the eigenfactor portfolios of cov_i (U scaled to unit column sums, returns
 * compounded over the next predlen dates), bias_stat = np.std(b, axis 0) over those dates.
 *   cov double [T*K*K] with cov_status [T] (NULL = every date present); factor_ret double [T*K];
 *   bias_stat double [K], column k = k-th LARGEST eigenvalue (canonical order);
 *   b_out optional double [(T-predlen)*K] (NULL to skip).  Uses the context workspace.
 */
int bmfm_eigenfactor_bias_stat(bmfm_ctx* ctx, int T, int K, int predlen, const double* cov,
                               const int32_t* cov_status, const double* factor_ret, double* bias_stat,
                               double* b_out);

/*
 * Replaces utils.bayes_shrink with group_mean_std / shrink (Barra-master/mfm/utils.py:133-168),
 * batched over dates.  The reference defines no caller; inputs are whatever
 * specific-volatility and capital vectors the caller has, grouped by date like the regression.
 *   volatility, capital, out  double [rows];  status int32 [T]: BMFM_ST_DUP_EDGES where pd.qcut
 *   would raise "Bin edges must be unique".  ngroup <= 64; at most 16384 rows per date.
 * Decile codes are bit-exact with pd.qcut(capital, ngroup).codes.
 */
int bmfm_bayes_shrink(bmfm_ctx* ctx, int T, int max_rows_per_date, const int64_t* date_off,
                      const double* volatility, const double* capital, int ngroup, double q,
                      double* out, int32_t* status);

/* ---- part 2: descriptors -------------------------------------------------- */
/*
 * Dense [L][S] panels in row-index-major order: element (r, s) = the r-th row of stock s (its own
 * rows in date order, what groupby('ts_code').rolling walks), NaN = missing / padding.
 *
 * bmfm_prepare_returns: ret = close / close.shift(1) - 1, log_ret = log(close) - log(close.shift(1))
 *   per stock (Barra_factor_cal/factor_calculator.py:50-51); either output may be NULL.
 * bmfm_beta_hsigma: FactorCalculator.compute_beta_hsigma (factor_calculator.py:79-125): rolling
 *   `window` rows, rows with NaN in ret or mkt dropped, >= min_periods valid, weights
 *   0.5^(rank/half_life) by rank among valid rows; BETA = WLS slope of ret on mkt (with intercept),
 *   HSIGMA = sqrt(sum w e^2 / (n - 2)).  mkt is the per-row market return [L*S].
 * bmfm_dastd: compute_dastd (factor_calculator.py:155-196): same windows on ret - mkt, normalised
 *   weights, weighted standard deviation.
 * bmfm_cmra: compute_cmra (factor_calculator.py:199-234): all `window` (<= 256) rows present,
 *   ln(1 + max Z) - ln(1 + min Z) with Z = exp(cumsum(log_ret)) - 1.
 */
int bmfm_prepare_returns(bmfm_ctx* ctx, int L, int S, const double* close, double* ret, double* log_ret);
int bmfm_beta_hsigma(bmfm_ctx* ctx, int L, int S, const double* ret, const double* mkt, int window,
                     double half_life, int min_periods, double* beta, double* hsigma);
int bmfm_dastd(bmfm_ctx* ctx, int L, int S, const double* ret, const double* mkt, int window,
               double half_life, int min_periods, double* out);
int bmfm_cmra(bmfm_ctx* ctx, int L, int S, const double* log_ret, int window, double* out);
/*
 * The remaining rolling descriptors, same [L * S] panels:
 * bmfm_rstr: FactorCalculator.compute_rstr (factor_calculator.py:127-153): log_ret shifted by `lag` rows,
 *   window T - lag, weights 0.5^(pos / half_life) by POSITION in the window (the oldest row has weight 1, as
 *   the reference's weights_arr[:len(window)] does), NaN rows dropped with their weights, NaN unless
 *   >= min_periods valid rows.  Reference constants: T = 504, lag = 21, half_life = 126, min_periods = 42.
 * bmfm_rolling_log_sum: one of STOM / STOQ / STOA of compute_liquidity (factor_calculator.py:324-367):
 *   log(rolling(window, min_periods).sum() of x / divisor), NaN where the sum is 0.  Reference:
 *   divisor = 100 and (window, min_periods) = (21, 15), (63, 42), (252, 126).
 * bmfm_nlsize: compute_nlsize (factor_calculator.py:237-293) on the long table grouped by date
 *   (date_off int64 [T+1]): minus the OLS residual of size^3 on [1, size] over the rows with size present,
 *   all NaN on dates with fewer than 2 such rows.
 */
int bmfm_rstr(bmfm_ctx* ctx, int L, int S, const double* log_ret, int T, int lag, double half_life,
              int min_periods, double* out);
int bmfm_rolling_log_sum(bmfm_ctx* ctx, int L, int S, const double* x, double divisor, int window,
                         int min_periods, double* out);
/* the three windows of compute_liquidity in one pass: windows / min_periods HOST int [3], outs HOST array of
 * 3 device pointers [L * S] each (STOM, STOQ, STOA in the reference's order) */
int bmfm_rolling_log_sum3(bmfm_ctx* ctx, int L, int S, const double* x, double divisor, const int* windows,
                          const int* min_periods, double* const* outs);
int bmfm_nlsize(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* size, double* out);
/* The element-wise descriptors on a column of n values:
 *   BMFM_OP_LOG      SIZE = ln(total_mv)                      compute_size  factor_calculator.py:68-77
 *   BMFM_OP_INV_POS  BP = 1 / pb where pb > 0, NaN otherwise  compute_bp    factor_calculator.py:295-321 */
#define BMFM_OP_LOG 0
#define BMFM_OP_INV_POS 1
/*   BMFM_OP_PERCENT  YOYProfit / YOYSales = x / 100           select_growth_factors  factor_calculator.py:436-462
 *   (BMFM_OP_INV_POS is also ETOP = 1 / pe_ttm where pe_ttm > 0, compute_earnings_yield :417-419) */
#define BMFM_OP_PERCENT 2
int bmfm_elementwise(bmfm_ctx* ctx, int op, int64_t n, const double* x, double* out);
/* Two-column descriptors, out = f(x, y):
 *   BMFM_OP2_RATIO_POS   CETOP = x / y where x > 0 and y > 0 (x = TTM cash flow, y = total_mv), NaN otherwise
 *                        compute_earnings_yield  factor_calculator.py:409-415
 *   BMFM_OP2_LEV_FINITE  MLEV = (y + x) / y with +-inf -> NaN (x = total_ncl, y = total_mv)
 *                        compute_leverage  factor_calculator.py:481-484
 *   BMFM_OP2_LEV_POS     BLEV = (y + x) / y where y > 0, NaN otherwise (x = total_ncl, y = book value)  :488-496 */
#define BMFM_OP2_RATIO_POS 10
#define BMFM_OP2_LEV_FINITE 11
#define BMFM_OP2_LEV_POS 12
int bmfm_elementwise2(bmfm_ctx* ctx, int op, int64_t n, const double* x, const double* y, double* out);
/* pandas rolling(window, min_periods).sum() down the rows of an [L * S] panel (NaN rows skipped; NaN with fewer
 * than min_periods valid rows): the trailing-four-quarter cash flow of compute_earnings_yield
 * (factor_calculator.py:394-396: window = min_periods = 4 over each stock's report periods).  window <= 64. */
int bmfm_rolling_sum(bmfm_ctx* ctx, int L, int S, const double* x, int window, int min_periods, double* out);

/*
 * Per-date post-processing on the long table grouped by date (date_off int64 [T+1], one fp64 array
 * per column, NaN = missing):
 * bmfm_winsorize: winsorize_factors (Barra_factor_cal/post_processing.py:7-24) for one column:
 *   clip to mean +- n_std * std, pandas semantics (NaN skipped, ddof 1).
 * bmfm_composite: calculate_composite_factors (post_processing.py:26-45):
 *   sum_i w_i x_i(NaN->0) / sum_i w_i notna(x_i).  components / weights are HOST arrays (of device
 *   pointers / doubles), ncomp <= 8.
 * bmfm_orthogonalize: orthogonalize_factors (post_processing.py:47-70) for one dependent column:
 *   OLS residual of y ~ 1 + X on complete-case rows, NaN elsewhere, all-NaN when fewer than k + 2
 *   complete rows.  X is a HOST array of k <= 6 device pointers.
 */
int bmfm_winsorize(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, double n_std, double* out);
int bmfm_composite(bmfm_ctx* ctx, int64_t rows, int ncomp, const double* const* components,
                   const double* weights, double* out);
int bmfm_orthogonalize(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* y, int k,
                       const double* const* X, double* out);

/*
 * The same two functions when the STOCKS of every date are sharded over ranks (no counterpart in the
 * single-process reference).  A rank's table holds its own stocks of every date (date_off over its rows); each pass
 * leaves per-date sums in a small array that the caller all-reduces (sum) over the ranks before the next pass:
 *   winsorize:     _sums  -> sums [T][2] (sum of valid values, their number)            all-reduce
 *                  _sqdev -> sqdev [T] = sum (mean_t - x)^2, mean_t = sum / number       all-reduce
 *                  _apply with lo / hi [T] = mean -+ n_std sqrt(sqdev / (number - 1)) (NaN bounds: values unchanged)
 *   orthogonalize: _pass 0 -> out [T][k + 2] (complete rows, sum X_1..X_k, sum y over them)            all-reduce
 *                  _pass 1 -> out [T][(k+1)(k+2)/2 + k + 1]: upper triangle of the Gram of [1, X - shift_t] and the
 *                             right-hand side with y - shift_t[k]; shift [T][k + 1] (the column means)   all-reduce
 *                  _pass 2 -> out = this rank's residuals from the all-reduced gram and count [T] (doubles)
 * (bmfm/dist.py: winsorize_sharded / orthogonalize_sharded drive them.)
 */
int bmfm_winsorize_sums(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, double* sums);
int bmfm_winsorize_sqdev(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, const double* mean,
                         double* sqdev);
int bmfm_winsorize_apply(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, const double* lo,
                         const double* hi, double* out);
int bmfm_orthogonalize_pass(bmfm_ctx* ctx, int pass, int T, const int64_t* date_off, const double* y, int k,
                            const double* const* X, const double* shift, const double* gram, const double* count,
                            double* out);

#ifdef __cplusplus
}
#endif
#endif /* BMFM_H */
