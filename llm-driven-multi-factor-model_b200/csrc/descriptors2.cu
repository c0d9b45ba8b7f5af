// The remaining rolling descriptors and the per-date size regression.
//
// Replaces, for all stocks / dates at once:
//   FactorCalculator.compute_rstr        Barra_factor_cal/factor_calculator.py:127-153
//   FactorCalculator.compute_liquidity   Barra_factor_cal/factor_calculator.py:324-367
//   FactorCalculator.compute_nlsize      Barra_factor_cal/factor_calculator.py:237-293
//
// Panels are [L][S] (row of the stock's own history x stock, stock contiguous) as in descriptors.cu: one
// thread per stock keeps every load coalesced, blockIdx.y cuts the rows into chunks of DESC2_CHUNK so that all
// SMs work; each chunk rebuilds its window sums from the rows before it and then slides (add the entering row,
// subtract the leaving one), which also bounds the rounding drift of the running sums to one chunk.
#include <math.h>
#include "bmfm_common.cuh"

namespace bmfm {

constexpr int DESC2_CHUNK = 256;

__device__ __forceinline__ double qnan2() { return __longlong_as_double(0x7ff8000000000000LL); }

// RSTR: y = x.shift(lag); window of W rows ending at r; weights BY POSITION IN THE WINDOW, decay^(pos), so the
// oldest row has weight 1 (factor_calculator.py:137-144: weights_arr[:len(window)]); NaN rows are dropped with
// their weights; NaN unless >= minp valid rows.  The normalisation cancels any common factor of the weights,
// so a chunk uses decay^(row - a0) with a0 = first row of the chunk's first window.
// Rows are taken eight at a time: the sixteen loads of a batch (entering and leaving rows) are issued before the
// serial updates, so a thread keeps 128 B in flight instead of 16 B (one thread per stock and chunk is ~75 000
// threads at 5000 x 3750: a single load each does not cover the memory latency).
constexpr int DESC2_U = 8;
__global__ void __launch_bounds__(128) rstr_kernel(int L, int S, int W, int lag, int minp, double decay, int chunk,
                                                   const double* x, double* out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const int r0 = blockIdx.y * chunk, r1 = min(L, r0 + chunk);
  const int a0 = max(0, r0 - W + 1);
  const double qn = qnan2();
  double num = 0.0, den = 0.0;
  int cnt = 0;
  double wadd = 1.0;                       // decay^(row - a0) of the row being added
  for (int rb = a0; rb < r0; rb += DESC2_U) {     // rows of the first window that precede the chunk
    double v[DESC2_U];
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) {
      const int r = rb + j;
      v[j] = (r < r0 && r >= lag) ? x[(size_t)(r - lag) * S + s] : qn;
    }
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) {
      if (rb + j < r0) {
        if (!isnan(v[j])) { num = fma(wadd, v[j], num); den += wadd; ++cnt; }
        wadd *= decay;
      }
    }
  }
  double wrem = 1.0;                       // decay^(row - a0) of the row being removed
  for (int rb = r0; rb < r1; rb += DESC2_U) {
    double v[DESC2_U], u[DESC2_U];
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) {
      const int r = rb + j, gone = r - W;  // gone: row that left the window
      v[j] = (r < r1 && r >= lag) ? x[(size_t)(r - lag) * S + s] : qn;
      u[j] = (r < r1 && gone >= a0 && gone >= lag) ? x[(size_t)(gone - lag) * S + s] : qn;
    }
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) {
      const int r = rb + j;
      if (r < r1) {
        if (!isnan(v[j])) { num = fma(wadd, v[j], num); den += wadd; ++cnt; }
        wadd *= decay;
        if (r - W >= a0) {
          if (!isnan(u[j])) { num = fma(-wrem, u[j], num); den -= wrem; --cnt; }
          wrem *= decay;
        }
        st_stream(out + (size_t)r * S + s, cnt >= minp ? num / den : qn);
      }
    }
  }
}

// STOM / STOQ / STOA: log of the rolling sum of x / divisor over W rows, NaN rows skipped, NaN unless >= minp
// valid rows, and NaN where the sum is 0 (.replace(0, nan), factor_calculator.py:350-352).  A window whose valid
// values are all exact zeros (a suspended stock) must give exactly 0 whatever was added and removed before, so
// the number of non-zero values in the window is tracked beside the sum.
__global__ void __launch_bounds__(128) roll_logsum_kernel(int L, int S, int W, int minp, double divisor,
                                                          const double* x, double* out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const int r0 = blockIdx.y * DESC2_CHUNK, r1 = min(L, r0 + DESC2_CHUNK);
  const int a0 = max(0, r0 - W + 1);
  double sum = 0.0;
  int cnt = 0, nnz = 0;
  for (int r = a0; r < r0; ++r) {
    const double v = x[(size_t)r * S + s];
    if (!isnan(v)) { sum += v; ++cnt; nnz += v != 0.0; }
  }
  for (int r = r0; r < r1; ++r) {
    const double v = x[(size_t)r * S + s];
    if (!isnan(v)) { sum += v; ++cnt; nnz += v != 0.0; }
    const int gone = r - W;
    if (gone >= a0) {
      const double u = x[(size_t)gone * S + s];
      if (!isnan(u)) { sum -= u; --cnt; nnz -= u != 0.0; }
    }
    double o = qnan2();
    if (cnt >= minp && nnz > 0 && sum != 0.0) o = log(sum / divisor);   // sums in the input's units, divided once     // log of a negative sum is NaN as in numpy
    st_stream(out + (size_t)r * S + s, o);
  }
}

// STOM, STOQ and STOA in one pass: the entering row is loaded and divided once for the three windows
struct Roll3 { int W[3], minp[3]; double* out[3]; };
__global__ void __launch_bounds__(128) roll_logsum3_kernel(int L, int S, Roll3 q, double divisor, const double* x) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const int r0 = blockIdx.y * DESC2_CHUNK, r1 = min(L, r0 + DESC2_CHUNK);
  const double qn = qnan2();
  double sum[3] = {0.0, 0.0, 0.0};
  int cnt[3] = {0, 0, 0}, nnz[3] = {0, 0, 0}, a0[3];
#pragma unroll
  for (int w = 0; w < 3; ++w) a0[w] = max(0, r0 - q.W[w] + 1);
  const int amin = min(a0[0], min(a0[1], a0[2]));
  for (int rb = amin; rb < r0; rb += DESC2_U) {
    double v[DESC2_U];
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) v[j] = rb + j < r0 ? x[(size_t)(rb + j) * S + s] : qn;
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) {
      const double vv = v[j];
      if (!isnan(vv)) {
#pragma unroll
        for (int w = 0; w < 3; ++w)
          if (rb + j >= a0[w]) { sum[w] += vv; ++cnt[w]; nnz[w] += vv != 0.0; }
      }
    }
  }
  for (int rb = r0; rb < r1; rb += DESC2_U) {
    double v[DESC2_U], u[3][DESC2_U];
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) {
      const int r = rb + j;
      v[j] = r < r1 ? x[(size_t)r * S + s] : qn;
#pragma unroll
      for (int w = 0; w < 3; ++w) {
        const int gone = r - q.W[w];
        u[w][j] = (r < r1 && gone >= a0[w]) ? x[(size_t)gone * S + s] : qn;
      }
    }
#pragma unroll
    for (int j = 0; j < DESC2_U; ++j) {
      const int r = rb + j;
      if (r < r1) {
        const double vv = v[j];
        const bool vok = !isnan(vv);
#pragma unroll
        for (int w = 0; w < 3; ++w) {
          if (vok) { sum[w] += vv; ++cnt[w]; nnz[w] += vv != 0.0; }
          const double uu = u[w][j];                 // NaN when nothing leaves (r - W before the first window)
          if (!isnan(uu)) { sum[w] -= uu; --cnt[w]; nnz[w] -= uu != 0.0; }
          double o = qn;
          // the window sums are kept in the input's units and divided once here (an IEEE division per loaded
          // value was a third of this kernel's instructions)
          if (cnt[w] >= q.minp[w] && nnz[w] > 0 && sum[w] != 0.0) o = log(sum[w] / divisor);
          st_stream(q.out[w] + (size_t)r * S + s, o);
        }
      }
    }
  }
}

// NLSIZE: per date, OLS of size^3 on [1, size] over the rows with size present; out = -(residual); all NaN
// when fewer than 2 such rows.  Centred form (the raw normal equations of x ~ 15, x^3 ~ 4000 lose digits):
//   b = sum (x-mx)(y-my) / sum (x-mx)^2,  residual = (y-my) - b (x-mx);  a constant column of x gives b = 0
//   (statsmodels' pinv solution has the same residuals).
__device__ __forceinline__ double block_sum_128(double v, double* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
  return t;
}

__global__ void __launch_bounds__(128) nlsize_kernel(int T, const long long* date_off, const double* size, double* out) {
  __shared__ double red[4];
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const long long a = date_off[t];
    const int n = (int)(date_off[t + 1] - a);
    double sx = 0.0, sy = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const double x = size[a + i];
      if (!isnan(x)) { sx += x; sy += x * x * x; c += 1.0; }
    }
    sx = block_sum_128(sx, red);
    sy = block_sum_128(sy, red);
    c = block_sum_128(c, red);
    if (c < 2.0) {
      for (int i = threadIdx.x; i < n; i += blockDim.x) out[a + i] = qnan2();
      continue;
    }
    const double mx = sx / c, my = sy / c;
    double sxx = 0.0, sxy = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const double x = size[a + i];
      if (!isnan(x)) {
        const double dx = x - mx, dy = x * x * x - my;
        sxx = fma(dx, dx, sxx);
        sxy = fma(dx, dy, sxy);
      }
    }
    sxx = block_sum_128(sxx, red);
    sxy = block_sum_128(sxy, red);
    const double b = sxx > 0.0 ? sxy / sxx : 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const double x = size[a + i];
      out[a + i] = isnan(x) ? qnan2() : -((x * x * x - my) - b * (x - mx));
    }
  }
}

// SIZE = ln(total_mv) (factor_calculator.py:68-77) and BP = 1 / pb where pb > 0, NaN otherwise (:295-321):
// one pass, 16 B per element
__global__ void elementwise_kernel(int op, long long n, const double* x, double* out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double v = x[i];
    out[i] = op == BMFM_OP_LOG ? log(v) : (op == BMFM_OP_INV_POS ? (v > 0.0 ? 1.0 / v : qnan2()) : v / 100.0);
  }
}

// Two-column element-wise descriptors (earnings yield and leverage, factor_calculator.py:371-509); 24 B per element
__global__ void elementwise2_kernel(int op, long long n, const double* x, const double* y, double* out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const double a = x[i], b = y[i];
    double r;
    if (op == BMFM_OP2_RATIO_POS) {               // CETOP: a / b where both are > 0 (:411-415)
      r = (b > 0.0 && a > 0.0) ? a / b : qnan2();
    } else if (op == BMFM_OP2_LEV_FINITE) {       // MLEV: (b + a) / b, +-inf -> NaN (:483-484)
      r = (b + a) / b;
      if (isinf(r)) r = qnan2();
    } else {                                      // BLEV: (b + a) / b where b > 0 (:492-496)
      r = b > 0.0 ? (b + a) / b : qnan2();
    }
    out[i] = r;
  }
}

// pandas rolling(window, min_periods).sum() down the rows of an [L][S] panel, thread = stock: the trailing-four-
// quarter cash flow of compute_earnings_yield (factor_calculator.py:394-396).  NaN rows are skipped and the
// result is NaN with fewer than min_periods valid rows in the window; the sum is re-formed per row from the
// window's rows (windows are a few rows long), so nothing drifts.
__global__ void rolling_sum_kernel(int L, int S, int window, int minp, const double* x, double* out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  for (int r = 0; r < L; ++r) {
    double acc = 0.0;
    int cnt = 0;
    const int lo = r - window + 1 > 0 ? r - window + 1 : 0;
    for (int j = lo; j <= r; ++j) {
      const double v = x[(size_t)j * S + s];
      if (!isnan(v)) { acc += v; ++cnt; }
    }
    out[(size_t)r * S + s] = cnt >= minp ? acc : qnan2();
  }
}

}  // namespace bmfm

using namespace bmfm;

extern "C" int bmfm_rstr(bmfm_ctx* ctx, int L, int S, const double* log_ret, int T, int lag, double half_life,
                         int min_periods, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, L >= 0 && S >= 0 && T > lag && lag >= 0 && half_life > 0 && min_periods >= 1, "bad sizes");
  BMFM_ARG(ctx, (log_ret && out) || (size_t)L * S == 0, "NULL pointer");
  if ((size_t)L * S == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  // (512-row chunks, which halve the re-read warm-up rows of the 483-row window, measured slower: 0.268 against
  // 0.195 ms at 5000 x 3750 — the thread count matters more than the L2 re-reads)
  const int chunk = DESC2_CHUNK;
  const dim3 grid((S + 127) / 128, (L + chunk - 1) / chunk);
  rstr_kernel<<<grid, 128, 0, ctx->stream>>>(L, S, T - lag, lag, min_periods, pow(0.5, 1.0 / half_life), chunk, log_ret, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_rolling_log_sum(bmfm_ctx* ctx, int L, int S, const double* x, double divisor, int window,
                                    int min_periods, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, L >= 0 && S >= 0 && window >= 1 && min_periods >= 1 && divisor != 0.0, "bad sizes");
  BMFM_ARG(ctx, (x && out) || (size_t)L * S == 0, "NULL pointer");
  if ((size_t)L * S == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const dim3 grid((S + 127) / 128, (L + DESC2_CHUNK - 1) / DESC2_CHUNK);
  roll_logsum_kernel<<<grid, 128, 0, ctx->stream>>>(L, S, window, min_periods, divisor, x, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_rolling_log_sum3(bmfm_ctx* ctx, int L, int S, const double* x, double divisor,
                                     const int* windows, const int* min_periods, double* const* outs) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, L >= 0 && S >= 0 && divisor != 0.0 && windows && min_periods && outs, "bad arguments");
  if ((size_t)L * S == 0) return BMFM_OK;
  Roll3 q;
  for (int w = 0; w < 3; ++w) {
    BMFM_ARG(ctx, windows[w] >= 1 && min_periods[w] >= 1 && outs[w], "bad window");
    q.W[w] = windows[w]; q.minp[w] = min_periods[w]; q.out[w] = outs[w];
  }
  BMFM_ARG(ctx, x != nullptr, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const dim3 grid((S + 127) / 128, (L + DESC2_CHUNK - 1) / DESC2_CHUNK);
  roll_logsum3_kernel<<<grid, 128, 0, ctx->stream>>>(L, S, q, divisor, x);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_nlsize(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* size, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0, "bad sizes");
  BMFM_ARG(ctx, T == 0 || (date_off && size && out), "NULL pointer");
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  int grid = T < ctx->sm_count * 8 ? T : ctx->sm_count * 8;
  nlsize_kernel<<<grid, 128, 0, ctx->stream>>>(T, reinterpret_cast<const long long*>(date_off), size, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_elementwise(bmfm_ctx* ctx, int op, int64_t n, const double* x, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, (op == BMFM_OP_LOG || op == BMFM_OP_INV_POS || op == BMFM_OP_PERCENT) && n >= 0, "bad arguments");
  if (n == 0) return BMFM_OK;
  BMFM_ARG(ctx, x && out, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  long long blocks = (n + 255) / 256;
  const long long cap = (long long)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  elementwise_kernel<<<(int)blocks, 256, 0, ctx->stream>>>(op, (long long)n, x, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_elementwise2(bmfm_ctx* ctx, int op, int64_t n, const double* x, const double* y, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, op >= BMFM_OP2_RATIO_POS && op <= BMFM_OP2_LEV_POS && n >= 0, "bad arguments");
  if (n == 0) return BMFM_OK;
  BMFM_ARG(ctx, x && y && out, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  long long blocks = (n + 255) / 256;
  const long long cap = (long long)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  elementwise2_kernel<<<(int)blocks, 256, 0, ctx->stream>>>(op, (long long)n, x, y, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_rolling_sum(bmfm_ctx* ctx, int L, int S, const double* x, int window, int min_periods, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, L >= 0 && S >= 0 && window >= 1 && window <= 64 && min_periods >= 1, "bad sizes (window <= 64)");
  if ((size_t)L * S == 0) return BMFM_OK;
  BMFM_ARG(ctx, x && out, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  rolling_sum_kernel<<<(S + 127) / 128, 128, 0, ctx->stream>>>(L, S, window, min_periods, x, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
