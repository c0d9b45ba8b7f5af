// Per-date post-processing of descriptor columns: winsorise, composite, orthogonalise.
//
// Replaces  winsorize_factors             Barra_factor_cal/post_processing.py:7-24
//           calculate_composite_factors   Barra_factor_cal/post_processing.py:26-45
//           orthogonalize_factors         Barra_factor_cal/post_processing.py:47-70
//
// Layout: the long table grouped by date (date t owns rows [date_off[t], date_off[t+1])), one
// contiguous fp64 array per column, NaN = missing.  One CTA per date; a date's column is read
// from HBM once (later passes hit L1/L2), 16 B per (row, column) of algorithmic traffic.
#include <math.h>
#include "bmfm_common.cuh"

namespace bmfm {

constexpr int PP_MAX_K = 6;       // regressors in orthogonalize (reference uses 1 and 2)
constexpr int PP_MAX_COMP = 8;    // components of a composite

__device__ __forceinline__ double block_sum(double v, double* s_red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < nw; ++w) t += s_red[w];
  return t;
}

// clip to mean +- n_std * std; pandas semantics: NaN skipped, ddof = 1, NaN bounds leave values alone
__global__ void __launch_bounds__(256) winsorize_kernel(int T, const long long* date_off, const double* x,
                                                        double n_std, double* out) {
  __shared__ double s_red[32];
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const long long a = date_off[t];
    const int n = (int)(date_off[t + 1] - a);
    double s = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const double v = x[a + i];
      if (!isnan(v)) { s += v; c += 1.0; }
    }
    s = block_sum(s, s_red);
    c = block_sum(c, s_red);
    const double mean = s / c;
    double q = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const double v = x[a + i];
      if (!isnan(v)) { const double d = mean - v; q = fma(d, d, q); }
    }
    q = block_sum(q, s_red);
    const double sd = c > 1.0 ? sqrt(q / (c - 1.0)) : __longlong_as_double(0x7ff8000000000000LL);
    const double lo = mean - n_std * sd, hi = mean + n_std * sd;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      double v = x[a + i];
      if (!isnan(v) && !isnan(lo) && !isnan(hi)) v = fmin(fmax(v, lo), hi);
      st_stream(out + a + i, v);
    }
  }
}

struct CompParams {
  int ncomp;
  const double* comp[PP_MAX_COMP];
  double w[PP_MAX_COMP];
};

// sum_i w_i x_i (NaN -> 0) / sum_i w_i notna(x_i)
__global__ void composite_kernel(size_t rows, CompParams p, double* out) {
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < rows; e += (size_t)gridDim.x * blockDim.x) {
    double num = 0.0, den = 0.0;
    for (int i = 0; i < p.ncomp; ++i) {
      const double v = p.comp[i][e];
      const bool ok = !isnan(v);
      num = __dadd_rn(num, __dmul_rn(ok ? v : 0.0, p.w[i]));
      den = __dadd_rn(den, ok ? p.w[i] : 0.0);
    }
    out[e] = num / den;
  }
}

struct OrthoParams {
  int T, k;
  const long long* date_off;
  const double* y;
  const double* X[PP_MAX_K];
  double* out;
};

// tiny symmetric eigen-solver (cyclic Jacobi, one thread): pinv solve of G c = b, numpy cutoff
__device__ void pinv_solve_small(int m, double* G, double* b, double* c) {
  double V[(PP_MAX_K + 1) * (PP_MAX_K + 1)];
  for (int i = 0; i < m; ++i)
    for (int j = 0; j < m; ++j) V[i * m + j] = i == j ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    int rot = 0;
    for (int p = 0; p < m - 1; ++p)
      for (int q = p + 1; q < m; ++q) {
        const double apq = G[p * m + q], app = G[p * m + p], aqq = G[q * m + q];
        if (fabs(apq) > 1.1102230246251565e-16 * sqrt(fabs(app * aqq)) && fabs(apq) > 1e-290) {
          rot++;
          const double tau = (aqq - app) / (2.0 * apq);
          const double tt = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
          const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = tt * cs;
          for (int k2 = 0; k2 < m; ++k2) {
            const double gkp = G[k2 * m + p], gkq = G[k2 * m + q];
            G[k2 * m + p] = cs * gkp - sn * gkq;
            G[k2 * m + q] = sn * gkp + cs * gkq;
            const double vkp = V[k2 * m + p], vkq = V[k2 * m + q];
            V[k2 * m + p] = cs * vkp - sn * vkq;
            V[k2 * m + q] = sn * vkp + cs * vkq;
          }
          for (int k2 = 0; k2 < m; ++k2) {
            const double gpk = G[p * m + k2], gqk = G[q * m + k2];
            G[p * m + k2] = cs * gpk - sn * gqk;
            G[q * m + k2] = sn * gpk + cs * gqk;
          }
        }
      }
    if (!rot) break;
  }
  double mx = 0.0;
  for (int i = 0; i < m; ++i) mx = fmax(mx, fabs(G[i * m + i]));
  for (int i = 0; i < m; ++i) c[i] = 0.0;
  for (int e = 0; e < m; ++e) {
    const double lam = G[e * m + e];
    if (!(fabs(lam) > 1e-15 * mx)) continue;
    double dot = 0.0;
    for (int i = 0; i < m; ++i) dot += V[i * m + e] * b[i];
    for (int i = 0; i < m; ++i) c[i] += V[i * m + e] * dot / lam;
  }
}

// residual of y ~ 1 + X on complete-case rows, NaN elsewhere; all-NaN when n < k + 2 (:56)
__global__ void __launch_bounds__(256) orthogonalize_kernel(OrthoParams p) {
  __shared__ double s_red[32];
  __shared__ double s_coef[PP_MAX_K + 2];
  __shared__ double s_shift[PP_MAX_K + 1];
  __shared__ int s_first;
  const int k = p.k, m = k + 1;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    if (threadIdx.x == 0) s_first = n;
    __syncthreads();
    double cnt = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      bool ok = !isnan(p.y[a + i]);
      for (int j = 0; j < k; ++j) ok = ok && !isnan(p.X[j][a + i]);
      if (ok) { cnt += 1.0; atomicMin(&s_first, i); }
    }
    cnt = block_sum(cnt, s_red);
    if (cnt < (double)(k + 2)) {
      for (int i = threadIdx.x; i < n; i += blockDim.x) p.out[a + i] = nan;
      __syncthreads();
      continue;
    }
    // shift every column by its first complete row: residuals are unchanged (intercept), conditioning improves
    if (threadIdx.x < k) s_shift[threadIdx.x] = p.X[threadIdx.x][a + s_first];
    if (threadIdx.x == k) s_shift[k] = p.y[a + s_first];
    __syncthreads();
    // Gram of [1, X - shift] and right-hand side, upper triangle accumulated per thread
    double g[(PP_MAX_K + 1) * (PP_MAX_K + 2) / 2];
    double rb[PP_MAX_K + 1];
    for (int e = 0; e < m * (m + 1) / 2; ++e) g[e] = 0.0;
    for (int e = 0; e < m; ++e) rb[e] = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      double row[PP_MAX_K + 1];
      row[0] = 1.0;
      const double yy = p.y[a + i] - s_shift[k];
      bool ok = !isnan(yy);
      for (int j = 0; j < k; ++j) { row[1 + j] = p.X[j][a + i] - s_shift[j]; ok = ok && !isnan(row[1 + j]); }
      if (!ok) continue;
      int e = 0;
      for (int r = 0; r < m; ++r) {
        for (int c = r; c < m; ++c) { g[e] = fma(row[r], row[c], g[e]); ++e; }
        rb[r] = fma(row[r], yy, rb[r]);
      }
    }
    __shared__ double s_G[(PP_MAX_K + 1) * (PP_MAX_K + 1)];
    __shared__ double s_b[PP_MAX_K + 1];
    {
      int e = 0;
      for (int r = 0; r < m; ++r)
        for (int c = r; c < m; ++c) {
          const double v = block_sum(g[e++], s_red);
          if (threadIdx.x == 0) { s_G[r * m + c] = v; s_G[c * m + r] = v; }
        }
      for (int r = 0; r < m; ++r) {
        const double v = block_sum(rb[r], s_red);
        if (threadIdx.x == 0) s_b[r] = v;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double G[(PP_MAX_K + 1) * (PP_MAX_K + 1)], b[PP_MAX_K + 1], c[PP_MAX_K + 1];
      for (int e = 0; e < m * m; ++e) G[e] = s_G[e];
      for (int e = 0; e < m; ++e) b[e] = s_b[e];
      pinv_solve_small(m, G, b, c);
      for (int e = 0; e < m; ++e) s_coef[e] = c[e];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      double r = p.y[a + i] - s_shift[k] - s_coef[0];
      for (int j = 0; j < k; ++j) r = fma(-s_coef[1 + j], p.X[j][a + i] - s_shift[j], r);
      st_stream(p.out + a + i, r);    // NaN inputs propagate: incomplete rows come out NaN
    }
    __syncthreads();
  }
}

}  // namespace bmfm

using namespace bmfm;

extern "C" int bmfm_winsorize(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, double n_std,
                              double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && date_off && x && out, "bad arguments");
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  int grid = T < ctx->sm_count * 8 ? T : ctx->sm_count * 8;
  winsorize_kernel<<<grid, 256, 0, ctx->stream>>>(T, reinterpret_cast<const long long*>(date_off), x, n_std, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_composite(bmfm_ctx* ctx, int64_t rows, int ncomp, const double* const* components,
                              const double* weights, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, rows >= 0 && ncomp >= 1 && components && weights && out, "bad arguments");
  if (ncomp > PP_MAX_COMP) {
    snprintf(ctx->err, sizeof(ctx->err), "ncomp=%d exceeds compiled limit %d", ncomp, PP_MAX_COMP);
    return BMFM_E_UNSUPPORTED;
  }
  if (rows == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  CompParams p;
  p.ncomp = ncomp;
  for (int i = 0; i < ncomp; ++i) { p.comp[i] = components[i]; p.w[i] = weights[i]; }
  size_t nb = ((size_t)rows + 255) / 256;
  int grid = (int)(nb < (size_t)ctx->sm_count * 16 ? nb : (size_t)ctx->sm_count * 16);
  composite_kernel<<<grid, 256, 0, ctx->stream>>>((size_t)rows, p, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_orthogonalize(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* y, int k,
                                  const double* const* X, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && k >= 1 && date_off && y && X && out, "bad arguments");
  if (k > PP_MAX_K) {
    snprintf(ctx->err, sizeof(ctx->err), "k=%d exceeds compiled limit %d", k, PP_MAX_K);
    return BMFM_E_UNSUPPORTED;
  }
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  OrthoParams p;
  p.T = T; p.k = k;
  p.date_off = reinterpret_cast<const long long*>(date_off);
  p.y = y; p.out = out;
  for (int j = 0; j < k; ++j) p.X[j] = X[j];
  int grid = T < ctx->sm_count * 8 ? T : ctx->sm_count * 8;
  orthogonalize_kernel<<<grid, 256, 0, ctx->stream>>>(p);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
