// Rolling per-stock descriptors: returns, BETA / HSIGMA, DASTD, CMRA.
//
// Replaces  FactorCalculator._prepare_data      Barra_factor_cal/factor_calculator.py:34-64 (returns)
//           FactorCalculator.compute_beta_hsigma                              :79-125
//           FactorCalculator.compute_dastd                                    :155-196
//           FactorCalculator.compute_cmra                                     :199-234
//
// Layout: a dense [L][S] panel in ROW-INDEX-major order — element (r, s) is the r-th row of
// stock s (its own trading rows in date order, exactly what pandas' groupby(ts_code).rolling
// walks), NaN where a value is missing and for the padding after a stock's last row.  Threads
// are stocks, so every load/store of a step is one coalesced row segment.  Time is cut into
// chunks (grid.y); a chunk warms up over the preceding window, so all (stock, chunk) pairs run
// in parallel.
//
// The reference re-fits every window from scratch (statsmodels WLS per stock-day).  Here the
// exponentially weighted window sums are carried: weights are assigned by RANK among the valid
// rows of the window (newest valid row gets decay^0 — `weights_arr[-n:]`, :97, :173), so when a
// row leaves the window its weight is decay^(n-1) with n the valid count, and when a valid row
// arrives every sum is multiplied by decay once.
#include <math.h>
#include "bmfm_common.cuh"

namespace bmfm {

constexpr int DESC_CHUNK = 256;      // rows per (stock, chunk) work item
constexpr int DESC_MAX_WINDOW = 1024;
constexpr int CMRA_MAX_WINDOW = 256;
constexpr int EW_U = 4;              // steps of the EW window kernels whose loads are issued together

__device__ __forceinline__ double qnan() { return __longlong_as_double(0x7ff8000000000000LL); }

// ret = close / close.shift(1) - 1 ; log_ret = log(close) - log(close.shift(1))   (:50-51, pandas 3: no fill)
// A thread walks PR_ROWS consecutive rows of one stock, so each log(close) is computed once and used for two rows
// (the logs are most of this kernel's instructions); stocks are the fast index, every access stays coalesced.
constexpr int PR_ROWS = 8;
__global__ void __launch_bounds__(256) prepare_returns_kernel(int L, int S, const double* close, double* ret, double* log_ret) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const int r0 = blockIdx.y * PR_ROWS;
  double c[PR_ROWS + 1];
  c[0] = r0 > 0 ? close[(size_t)(r0 - 1) * S + s] : qnan();
#pragma unroll
  for (int j = 0; j < PR_ROWS; ++j) c[j + 1] = r0 + j < L ? close[(size_t)(r0 + j) * S + s] : qnan();
  double lprev = log_ret ? log(c[0]) : 0.0;
#pragma unroll
  for (int j = 0; j < PR_ROWS; ++j) {
    const int r = r0 + j;
    if (r < L) {
      double rr = qnan(), lr = qnan();
      const double lcur = log_ret ? log(c[j + 1]) : 0.0;
      if (r > 0) {
        rr = __dsub_rn(__ddiv_rn(c[j + 1], c[j]), 1.0);
        lr = __dsub_rn(lcur, lprev);
      }
      lprev = lcur;
      if (ret) st_stream(ret + (size_t)r * S + s, rr);
      if (log_ret) st_stream(log_ret + (size_t)r * S + s, lr);
    }
  }
}

// MODE 0: BETA + HSIGMA (weighted regression of ret on market), MODE 1: DASTD (weighted std of ret - market)
template <int MODE>
__global__ void __launch_bounds__(128) ew_window_kernel(int L, int S, int W, int minp, double lam,
                                                        const double* ret, const double* mkt,
                                                        double* out0, double* out1) {
  __shared__ double pw[DESC_MAX_WINDOW + 1];
  for (int i = threadIdx.x; i <= W; i += blockDim.x) pw[i] = pow(lam, (double)i);
  __syncthreads();
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const int r0 = blockIdx.y * DESC_CHUNK, r1 = min(L, r0 + DESC_CHUNK);
  const int s0 = max(0, r0 - W);        // warm-up start: the window of row r0-1 is rows [s0, r0)
  double Sw = 0, Sx = 0, Sy = 0, Sxx = 0, Sxy = 0, Syy = 0;
  int n = 0;
  // the four values of a step (arriving row, leaving row) do not depend on the running sums: the loads of
  // EW_U steps are issued together, ahead of the serial updates (one thread per stock and chunk cannot cover
  // the memory latency with a single step's loads in flight)
  const double qn = qnan();
  for (int rb = s0; rb < r1; rb += EW_U) {
    double yb[EW_U], xb[EW_U], ylb[EW_U], xlb[EW_U];
#pragma unroll
    for (int j = 0; j < EW_U; ++j) {
      const int r = rb + j, rl = r - W;
      const bool in = r < r1, lv = in && rl >= s0;
      yb[j] = in ? ret[(size_t)r * S + s] : qn;
      xb[j] = in ? mkt[(size_t)r * S + s] : qn;
      ylb[j] = lv ? ret[(size_t)rl * S + s] : qn;
      xlb[j] = lv ? mkt[(size_t)rl * S + s] : qn;
    }
#pragma unroll
    for (int j = 0; j < EW_U; ++j) {
      const int r = rb + j;
      if (r >= r1) break;
      const double y = yb[j], x = xb[j], yl = ylb[j], xl = xlb[j];
      const int rl = r - W;
      if (rl >= s0) {
        // the oldest row of the previous window leaves; among n valid rows it had rank n-1
        if (!isnan(yl) && !isnan(xl)) {
          const double wg = pw[n - 1];
          if (MODE == 0) {
            Sw -= wg; Sx = fma(-wg, xl, Sx); Sy = fma(-wg, yl, Sy);
            Sxx = fma(-wg * xl, xl, Sxx); Sxy = fma(-wg * xl, yl, Sxy); Syy = fma(-wg * yl, yl, Syy);
          } else {
            const double e = yl - xl;
            Sw -= wg; Sx = fma(-wg, e, Sx); Sxx = fma(-wg * e, e, Sxx);
          }
          n -= 1;
        }
      }
      if (!isnan(y) && !isnan(x)) {
        if (MODE == 0) {
          Sw = fma(Sw, lam, 1.0); Sx = fma(Sx, lam, x); Sy = fma(Sy, lam, y);
          Sxx = fma(Sxx, lam, x * x); Sxy = fma(Sxy, lam, x * y); Syy = fma(Syy, lam, y * y);
        } else {
          const double e = y - x;
          Sw = fma(Sw, lam, 1.0); Sx = fma(Sx, lam, e); Sxx = fma(Sxx, lam, e * e);
        }
        n += 1;
      }
      if (n == 0) { Sw = Sx = Sy = Sxx = Sxy = Syy = 0.0; }   // drop accumulated rounding when the window empties
      if (r >= r0) {
        double o0 = qnan(), o1 = qnan();
        if (n >= minp) {
          if (MODE == 0) {
            // WLS of y on [1, x] with un-normalised weights (:96-102); HSIGMA = sqrt(sum w e^2 / (n - 2))
            const double det = Sw * Sxx - Sx * Sx;
            const double beta = (Sw * Sxy - Sx * Sy) / det;
            const double alpha = (Sy - beta * Sx) / Sw;
            const double sse = Syy - alpha * Sy - beta * Sxy;
            o0 = beta;
            o1 = sqrt(fmax(sse, 0.0) / (double)(n - 2));
          } else {
            // normalised weights: sqrt(sum w (x - mu)^2), mu = sum w x (:173-180)
            const double mu = Sx / Sw;
            o0 = sqrt(fmax(Sxx / Sw - mu * mu, 0.0));
          }
        }
        st_stream(out0 + (size_t)r * S + s, o0);
        if (MODE == 0) st_stream(out1 + (size_t)r * S + s, o1);
      }
    }
  }
}

// CMRA: ln(1 + max Z) - ln(1 + min Z), Z = exp(cumsum(log_ret over the window)) - 1, all W rows present.
// = max - min of the running sum over the window (any common offset cancels).  Sliding max/min by the block
// decomposition of van Herk / Gil-Werman: a thread owns (stock, block of W rows); with the running sum measured
// from the END of the previous block, it needs the suffix extrema of the previous block (one backward pass over
// it: R_j = -(sum of the rows after j), kept in thread-local arrays that all threads index uniformly) and the
// running prefix extrema of its own block.  Loads are taken eight rows at a time so that they overlap.
constexpr int CMRA_U = 8;
__global__ void __launch_bounds__(128) cmra_kernel(int L, int S, int W, const double* log_ret, double* out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const int b = blockIdx.y;
  const int r0 = b * W, r1 = min(L, r0 + W);
  const double qn = qnan();
  double smax[CMRA_MAX_WINDOW], smin[CMRA_MAX_WINDOW];
  int lastnan = -1;          // largest row index with NaN seen so far (>= r0 - W), -1 if none
  if (b > 0) {
    const int p0 = r0 - W;
    double Q = 0.0, mx = -INFINITY, mn = INFINITY;     // Q: sum of the rows after j
    for (int jb = W - 1; jb >= 0; jb -= CMRA_U) {
      double v[CMRA_U];
#pragma unroll
      for (int u = 0; u < CMRA_U; ++u) v[u] = jb - u >= 0 ? log_ret[(size_t)(p0 + jb - u) * S + s] : qn;
#pragma unroll
      for (int u = 0; u < CMRA_U; ++u) {
        const int j = jb - u;
        if (j >= 0) {
          mx = fmax(mx, -Q);
          mn = fmin(mn, -Q);
          smax[j] = mx;
          smin[j] = mn;
          if (isnan(v[u])) { if (lastnan < 0) lastnan = p0 + j; } else Q += v[u];
        }
      }
    }
  }
  double P = 0.0, pmax = -INFINITY, pmin = INFINITY;
  for (int rb = r0; rb < r1; rb += CMRA_U) {
    double v[CMRA_U];
#pragma unroll
    for (int u = 0; u < CMRA_U; ++u) v[u] = rb + u < r1 ? log_ret[(size_t)(rb + u) * S + s] : qn;
#pragma unroll
    for (int u = 0; u < CMRA_U; ++u) {
      const int r = rb + u;
      if (r < r1) {
        if (isnan(v[u])) lastnan = r; else P += v[u];
        pmax = fmax(pmax, P);
        pmin = fmin(pmin, P);
        const int ws = r - W + 1;          // first row of the window
        double o = qn;
        if (ws >= 0 && lastnan < ws) {
          double mx = pmax, mn = pmin;
          if (ws < r0) {                   // the window reaches into the previous block
            mx = fmax(mx, smax[ws - (r0 - W)]);
            mn = fmin(mn, smin[ws - (r0 - W)]);
          }
          o = mx - mn;
        }
        st_stream(out + (size_t)r * S + s, o);
      }
    }
  }
}

}  // namespace bmfm

using namespace bmfm;

extern "C" int bmfm_prepare_returns(bmfm_ctx* ctx, int L, int S, const double* close, double* ret, double* log_ret) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, L >= 0 && S >= 0 && close && (ret || log_ret), "bad arguments");
  if (L == 0 || S == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const dim3 grid((S + 255) / 256, (L + PR_ROWS - 1) / PR_ROWS);
  prepare_returns_kernel<<<grid, 256, 0, ctx->stream>>>(L, S, close, ret, log_ret);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

static int ew_launch(bmfm_ctx* ctx, int mode, int L, int S, const double* ret, const double* mkt, int window,
                     double half_life, int min_periods, double* out0, double* out1) {
  BMFM_ARG(ctx, L >= 0 && S >= 0 && window >= 1 && half_life > 0 && min_periods >= 1, "bad sizes");
  BMFM_ARG(ctx, ret && mkt && out0 && (mode == 1 || out1), "NULL pointer");
  if (window > DESC_MAX_WINDOW) {
    snprintf(ctx->err, sizeof(ctx->err), "window=%d exceeds compiled limit %d", window, DESC_MAX_WINDOW);
    return BMFM_E_UNSUPPORTED;
  }
  if (mode == 0 && min_periods < 3) {
    snprintf(ctx->err, sizeof(ctx->err), "beta/hsigma needs min_periods >= 3");
    return BMFM_E_ARG;
  }
  if (L == 0 || S == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const double lam = pow(0.5, 1.0 / half_life);
  dim3 grid((S + 127) / 128, (L + DESC_CHUNK - 1) / DESC_CHUNK);
  if (mode == 0) ew_window_kernel<0><<<grid, 128, 0, ctx->stream>>>(L, S, window, min_periods, lam, ret, mkt, out0, out1);
  else ew_window_kernel<1><<<grid, 128, 0, ctx->stream>>>(L, S, window, min_periods, lam, ret, mkt, out0, out1);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_beta_hsigma(bmfm_ctx* ctx, int L, int S, const double* ret, const double* mkt, int window,
                                double half_life, int min_periods, double* beta, double* hsigma) {
  if (!ctx) return BMFM_E_ARG;
  return ew_launch(ctx, 0, L, S, ret, mkt, window, half_life, min_periods, beta, hsigma);
}

extern "C" int bmfm_dastd(bmfm_ctx* ctx, int L, int S, const double* ret, const double* mkt, int window,
                          double half_life, int min_periods, double* out) {
  if (!ctx) return BMFM_E_ARG;
  return ew_launch(ctx, 1, L, S, ret, mkt, window, half_life, min_periods, out, nullptr);
}

extern "C" int bmfm_cmra(bmfm_ctx* ctx, int L, int S, const double* log_ret, int window, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, L >= 0 && S >= 0 && window >= 1 && log_ret && out, "bad arguments");
  if (window > CMRA_MAX_WINDOW) {
    snprintf(ctx->err, sizeof(ctx->err), "window=%d exceeds compiled limit %d", window, CMRA_MAX_WINDOW);
    return BMFM_E_UNSUPPORTED;
  }
  if (L == 0 || S == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  dim3 grid((S + 127) / 128, (L + window - 1) / window);
  cmra_kernel<<<grid, 128, 0, ctx->stream>>>(L, S, window, log_ret, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
