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

// clip to mean +- n_std * std; pandas semantics: NaN skipped, ddof = 1, NaN bounds leave values alone.
// Three passes over the date's rows (sum and count / squared deviations / clip): the first reads HBM, the others
// hit L1/L2; loads four at a time, the two sums of the first pass reduced together.
__global__ void __launch_bounds__(256) winsorize_kernel(int T, const long long* date_off, const double* x,
                                                        double n_std, double* out) {
  __shared__ double s_a[8], s_b[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const long long a = date_off[t];
    const int n = (int)(date_off[t + 1] - a);
    const double* xs = x + a;
    double s = 0.0, c = 0.0;
    for (int i0 = tid; i0 < n; i0 += 4 * 256) {
      double v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = i0 + 256 * u < n ? xs[i0 + 256 * u] : nan;
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (!isnan(v[u])) { s += v[u]; c += 1.0; }
    }
    s = warp_sum(s);
    c = warp_sum(c);
    if (lane == 0) { s_a[warp] = s; s_b[warp] = c; }
    __syncthreads();
    s = 0.0; c = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { s += s_a[w]; c += s_b[w]; }
    __syncthreads();
    const double mean = s / c;
    double q = 0.0;
    for (int i0 = tid; i0 < n; i0 += 4 * 256) {
      double v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = i0 + 256 * u < n ? xs[i0 + 256 * u] : nan;
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (!isnan(v[u])) { const double d = mean - v[u]; q = fma(d, d, q); }
    }
    q = warp_sum(q);
    if (lane == 0) s_a[warp] = q;
    __syncthreads();
    q = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) q += s_a[w];
    __syncthreads();
    const double sd = c > 1.0 ? sqrt(q / (c - 1.0)) : nan;
    const double lo = mean - n_std * sd, hi = mean + n_std * sd;
    const bool clip = !isnan(lo) && !isnan(hi);
    for (int i0 = tid; i0 < n; i0 += 4 * 256) {
      double v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = i0 + 256 * u < n ? xs[i0 + 256 * u] : nan;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (i0 + 256 * u < n) {
          double w = v[u];
          if (!isnan(w) && clip) w = fmin(fmax(w, lo), hi);
          st_stream(out + a + i0 + 256 * u, w);
        }
      }
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

// tiny symmetric eigen-solver (cyclic Jacobi, one thread): pinv solve of G c = b, numpy cutoff.  M is a template
// parameter so that every index is a compile-time constant after unrolling and the matrices live in registers.
template <int M>
__device__ __forceinline__ void pinv_solve_small(double (&G)[M * M], const double (&b)[M], double (&c)[M]) {
  double V[M * M];
#pragma unroll
  for (int i = 0; i < M; ++i)
#pragma unroll
    for (int j = 0; j < M; ++j) V[i * M + j] = i == j ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    int rot = 0;
#pragma unroll
    for (int p = 0; p < M - 1; ++p)
#pragma unroll
      for (int q = p + 1; q < M; ++q) {
        const double apq = G[p * M + q], app = G[p * M + p], aqq = G[q * M + q];
        if (fabs(apq) > 1.1102230246251565e-16 * sqrt(fabs(app * aqq)) && fabs(apq) > 1e-290) {
          rot++;
          const double tau = (aqq - app) / (2.0 * apq);
          const double tt = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
          const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = tt * cs;
#pragma unroll
          for (int k2 = 0; k2 < M; ++k2) {
            const double gkp = G[k2 * M + p], gkq = G[k2 * M + q];
            G[k2 * M + p] = cs * gkp - sn * gkq;
            G[k2 * M + q] = sn * gkp + cs * gkq;
            const double vkp = V[k2 * M + p], vkq = V[k2 * M + q];
            V[k2 * M + p] = cs * vkp - sn * vkq;
            V[k2 * M + q] = sn * vkp + cs * vkq;
          }
#pragma unroll
          for (int k2 = 0; k2 < M; ++k2) {
            const double gpk = G[p * M + k2], gqk = G[q * M + k2];
            G[p * M + k2] = cs * gpk - sn * gqk;
            G[q * M + k2] = sn * gpk + cs * gqk;
          }
        }
      }
    if (!rot) break;
  }
  double mx = 0.0;
#pragma unroll
  for (int i = 0; i < M; ++i) mx = fmax(mx, fabs(G[i * M + i]));
#pragma unroll
  for (int i = 0; i < M; ++i) c[i] = 0.0;
#pragma unroll
  for (int e = 0; e < M; ++e) {
    const double lam = G[e * M + e];
    if (!(fabs(lam) > 1e-15 * mx)) continue;
    double dot = 0.0;
#pragma unroll
    for (int i = 0; i < M; ++i) dot += V[i * M + e] * b[i];
#pragma unroll
    for (int i = 0; i < M; ++i) c[i] += V[i * M + e] * dot / lam;
  }
}

// residual of y ~ 1 + X on complete-case rows, NaN elsewhere; all-NaN when n < k + 2 (:56).  One CTA per date,
// three passes over the date's rows (count / Gram / residuals): the first reads HBM, the other two hit L1/L2.
// KX regressors at compile time (register arrays), all sums of a pass reduced together (two barriers).
template <int KX>
__global__ void __launch_bounds__(256) orthogonalize_kernel(OrthoParams p) {
  constexpr int M = KX + 1, NG = M * (M + 1) / 2, NR = NG + M;
  __shared__ double s_part[8][NR];
  __shared__ double s_tot[NR];
  __shared__ double s_coef[M];
  __shared__ int s_cnt[8], s_fst[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    const double* y = p.y + a;
    const double* X[KX];
#pragma unroll
    for (int j = 0; j < KX; ++j) X[j] = p.X[j] + a;
    // pass 1: complete rows, and the first of them
    int cnt = 0, first = n;
    for (int i0 = tid; i0 < n; i0 += 4 * 256) {
      double v[4][M];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + 256 * u;
#pragma unroll
        for (int j = 0; j < KX; ++j) v[u][j] = i < n ? X[j][i] : nan;
        v[u][KX] = i < n ? y[i] : nan;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        bool ok = true;
#pragma unroll
        for (int j = 0; j < M; ++j) ok = ok && !isnan(v[u][j]);
        if (ok) { cnt++; first = min(first, i0 + 256 * u); }
      }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    first = __reduce_min_sync(0xffffffffu, first);
    if (lane == 0) { s_cnt[warp] = cnt; s_fst[warp] = first; }
    __syncthreads();
    cnt = 0; first = n;
#pragma unroll
    for (int w = 0; w < 8; ++w) { cnt += s_cnt[w]; first = min(first, s_fst[w]); }
    if (cnt < KX + 2) {
      for (int i = tid; i < n; i += 256) p.out[a + i] = nan;
      __syncthreads();
      continue;
    }
    // shift every column by its first complete row: residuals are unchanged (intercept), conditioning improves
    double sh[M];
#pragma unroll
    for (int j = 0; j < KX; ++j) sh[j] = X[j][first];
    sh[KX] = y[first];
    // pass 2: Gram of [1, X - shift] (upper triangle) and right-hand side
    double acc[NR];
#pragma unroll
    for (int e = 0; e < NR; ++e) acc[e] = 0.0;
    for (int i0 = tid; i0 < n; i0 += 4 * 256) {
      double v[4][M];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + 256 * u;
#pragma unroll
        for (int j = 0; j < KX; ++j) v[u][j] = i < n ? X[j][i] : nan;
        v[u][KX] = i < n ? y[i] : nan;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        double row[M];
        row[0] = 1.0;
        const double yy = v[u][KX] - sh[KX];
        bool ok = !isnan(yy);
#pragma unroll
        for (int j = 0; j < KX; ++j) { row[1 + j] = v[u][j] - sh[j]; ok = ok && !isnan(row[1 + j]); }
        if (ok) {
          int e = 0;
#pragma unroll
          for (int r = 0; r < M; ++r) {
#pragma unroll
            for (int c = r; c < M; ++c) { acc[e] = fma(row[r], row[c], acc[e]); ++e; }
            acc[NG + r] = fma(row[r], yy, acc[NG + r]);
          }
        }
      }
    }
#pragma unroll
    for (int e = 0; e < NR; ++e) {
      const double v = warp_sum(acc[e]);
      if (lane == 0) s_part[warp][e] = v;
    }
    __syncthreads();
    if (tid < NR) {
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) v += s_part[w][tid];
      s_tot[tid] = v;
    }
    __syncthreads();
    if (tid == 0) {
      double G[M * M], b[M], c[M];
      int e = 0;
#pragma unroll
      for (int r = 0; r < M; ++r)
#pragma unroll
        for (int cc = r; cc < M; ++cc) { const double v = s_tot[e++]; G[r * M + cc] = v; G[cc * M + r] = v; }
#pragma unroll
      for (int r = 0; r < M; ++r) b[r] = s_tot[NG + r];
      pinv_solve_small<M>(G, b, c);
#pragma unroll
      for (int r = 0; r < M; ++r) s_coef[r] = c[r];
    }
    __syncthreads();
    double cf[M];
#pragma unroll
    for (int r = 0; r < M; ++r) cf[r] = s_coef[r];
    for (int i0 = tid; i0 < n; i0 += 4 * 256) {
      double v[4][M];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + 256 * u;
#pragma unroll
        for (int j = 0; j < KX; ++j) v[u][j] = i < n ? X[j][i] : nan;
        v[u][KX] = i < n ? y[i] : nan;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + 256 * u;
        double r = v[u][KX] - sh[KX] - cf[0];
#pragma unroll
        for (int j = 0; j < KX; ++j) r = fma(-cf[1 + j], v[u][j] - sh[j], r);
        if (i < n) st_stream(p.out + a + i, r);    // NaN inputs propagate: incomplete rows come out NaN
      }
    }
    __syncthreads();
  }
}

// ---- stock-sharded forms (descriptors shard by stock, per-date moments are all-reduced) -----------
// A rank holds a block of STOCKS of every date (the [days][my stocks] panels the rolling kernels write are already
// "grouped by date").  The per-date statistics then need every rank's rows: each pass below leaves its per-date sums
// in a small array, the caller all-reduces it (KB), and the next pass uses the totals.  No bulk transpose.

// pass 1 of winsorize: out[2 t] = sum of the valid values, out[2 t + 1] = their number
__global__ void __launch_bounds__(256) winsorize_sums_kernel(int T, const long long* date_off, const double* x, double* out) {
  __shared__ double s_a[8], s_b[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const long long a = date_off[t];
    const int n = (int)(date_off[t + 1] - a);
    double s = 0.0, c = 0.0;
    for (int i = tid; i < n; i += 256) {
      const double v = x[a + i];
      if (!isnan(v)) { s += v; c += 1.0; }
    }
    s = warp_sum(s);
    c = warp_sum(c);
    if (lane == 0) { s_a[warp] = s; s_b[warp] = c; }
    __syncthreads();
    if (tid == 0) {
      s = 0.0; c = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { s += s_a[w]; c += s_b[w]; }
      out[2 * t] = s;
      out[2 * t + 1] = c;
    }
    __syncthreads();
  }
}

// pass 2: out[t] = sum over the valid values of (mean_t - x)^2
__global__ void __launch_bounds__(256) winsorize_sqdev_kernel(int T, const long long* date_off, const double* x,
                                                              const double* mean, double* out) {
  __shared__ double s_a[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const long long a = date_off[t];
    const int n = (int)(date_off[t + 1] - a);
    const double m = mean[t];
    double q = 0.0;
    for (int i = tid; i < n; i += 256) {
      const double v = x[a + i];
      if (!isnan(v)) { const double d = m - v; q = fma(d, d, q); }
    }
    q = warp_sum(q);
    if (lane == 0) s_a[warp] = q;
    __syncthreads();
    if (tid == 0) {
      q = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) q += s_a[w];
      out[t] = q;
    }
    __syncthreads();
  }
}

// pass 3: clip to [lo_t, hi_t]; NaN bounds leave the values alone (as the single-GPU kernel)
__global__ void __launch_bounds__(256) winsorize_apply_kernel(int T, const long long* date_off, const double* x,
                                                              const double* lo, const double* hi, double* out) {
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const long long a = date_off[t];
    const int n = (int)(date_off[t + 1] - a);
    const double l = lo[t], h = hi[t];
    const bool clip = !isnan(l) && !isnan(h);
    for (int i = threadIdx.x; i < n; i += 256) {
      double v = x[a + i];
      if (!isnan(v) && clip) v = fmin(fmax(v, l), h);
      st_stream(out + a + i, v);
    }
  }
}

// orthogonalize, pass 1 (MODE 0): out[t][0] = complete rows, out[t][1 + j] = sum of X_j, out[t][1 + KX] = sum of y over
// them.  Pass 2 (MODE 1): Gram of [1, X - shift_t] (upper triangle) and right-hand side with y - shift_t[KX], NR sums.
template <int KX, int MODE>
__global__ void __launch_bounds__(256) ortho_sums_kernel(OrthoParams p, const double* shift, double* out) {
  constexpr int M = KX + 1, NG = M * (M + 1) / 2;
  constexpr int NR = MODE == 0 ? KX + 2 : NG + M;
  __shared__ double s_part[8][NR];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    double sh[M];
#pragma unroll
    for (int j = 0; j < M; ++j) sh[j] = MODE == 1 ? shift[(size_t)t * M + j] : 0.0;
    double acc[NR];
#pragma unroll
    for (int e = 0; e < NR; ++e) acc[e] = 0.0;
    for (int i = tid; i < n; i += 256) {
      double row[M];
      row[0] = 1.0;
      const double yy = p.y[a + i] - sh[KX];
      bool ok = !isnan(yy);
#pragma unroll
      for (int j = 0; j < KX; ++j) { row[1 + j] = p.X[j][a + i] - sh[j]; ok = ok && !isnan(row[1 + j]); }
      if (ok) {
        if (MODE == 0) {
          acc[0] += 1.0;
#pragma unroll
          for (int j = 0; j < KX; ++j) acc[1 + j] += row[1 + j];
          acc[1 + KX] += yy;
        } else {
          int e = 0;
#pragma unroll
          for (int r = 0; r < M; ++r) {
#pragma unroll
            for (int c = r; c < M; ++c) { acc[e] = fma(row[r], row[c], acc[e]); ++e; }
            acc[NG + r] = fma(row[r], yy, acc[NG + r]);
          }
        }
      }
    }
#pragma unroll
    for (int e = 0; e < NR; ++e) {
      const double v = warp_sum(acc[e]);
      if (lane == 0) s_part[warp][e] = v;
    }
    __syncthreads();
    if (tid < NR) {
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) v += s_part[w][tid];
      out[(size_t)t * NR + tid] = v;
    }
    __syncthreads();
  }
}

// pass 3: solve from the all-reduced Gram (every rank the same numbers) and write this rank's residuals
template <int KX>
__global__ void __launch_bounds__(256) ortho_apply_kernel(OrthoParams p, const double* shift, const double* gram,
                                                          const double* count) {
  constexpr int M = KX + 1, NG = M * (M + 1) / 2, NR = NG + M;
  __shared__ double s_coef[M];
  const int tid = threadIdx.x;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    if (count[t] < (double)(KX + 2)) {
      for (int i = tid; i < n; i += 256) p.out[a + i] = nan;
      continue;
    }
    if (tid == 0) {
      double G[M * M], b[M], c[M];
      const double* g = gram + (size_t)t * NR;
      int e = 0;
#pragma unroll
      for (int r = 0; r < M; ++r)
#pragma unroll
        for (int cc = r; cc < M; ++cc) { const double v = g[e++]; G[r * M + cc] = v; G[cc * M + r] = v; }
#pragma unroll
      for (int r = 0; r < M; ++r) b[r] = g[NG + r];
      pinv_solve_small<M>(G, b, c);
#pragma unroll
      for (int r = 0; r < M; ++r) s_coef[r] = c[r];
    }
    __syncthreads();
    double cf[M], sh[M];
#pragma unroll
    for (int r = 0; r < M; ++r) { cf[r] = s_coef[r]; sh[r] = shift[(size_t)t * M + r]; }
    for (int i = tid; i < n; i += 256) {
      double r = p.y[a + i] - sh[KX] - cf[0];
#pragma unroll
      for (int j = 0; j < KX; ++j) r = fma(-cf[1 + j], p.X[j][a + i] - sh[j], r);
      st_stream(p.out + a + i, r);
    }
    __syncthreads();
  }
}

template <int KX>
static void ortho_shard_launch(bmfm_ctx* ctx, const OrthoParams& p, int grid, int pass, const double* shift,
                               const double* gram, const double* count, double* out) {
  if (pass == 0) ortho_sums_kernel<KX, 0><<<grid, 256, 0, ctx->stream>>>(p, nullptr, out);
  else if (pass == 1) ortho_sums_kernel<KX, 1><<<grid, 256, 0, ctx->stream>>>(p, shift, out);
  else ortho_apply_kernel<KX><<<grid, 256, 0, ctx->stream>>>(p, shift, gram, count);
}

template <int KX>
static void ortho_launch(bmfm_ctx* ctx, const OrthoParams& p, int grid) {
  orthogonalize_kernel<KX><<<grid, 256, 0, ctx->stream>>>(p);
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
  switch (k) {
    case 1: ortho_launch<1>(ctx, p, grid); break;
    case 2: ortho_launch<2>(ctx, p, grid); break;
    case 3: ortho_launch<3>(ctx, p, grid); break;
    case 4: ortho_launch<4>(ctx, p, grid); break;
    case 5: ortho_launch<5>(ctx, p, grid); break;
    default: ortho_launch<6>(ctx, p, grid); break;
  }
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// ---- stock-sharded post-processing: the three passes of winsorize and orthogonalize around the caller's all-reduce ----
extern "C" int bmfm_winsorize_sums(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, double* sums) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && date_off && x && sums, "bad arguments");
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int grid = T < ctx->sm_count * 8 ? T : ctx->sm_count * 8;
  winsorize_sums_kernel<<<grid, 256, 0, ctx->stream>>>(T, reinterpret_cast<const long long*>(date_off), x, sums);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_winsorize_sqdev(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, const double* mean,
                                    double* sqdev) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && date_off && x && mean && sqdev, "bad arguments");
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int grid = T < ctx->sm_count * 8 ? T : ctx->sm_count * 8;
  winsorize_sqdev_kernel<<<grid, 256, 0, ctx->stream>>>(T, reinterpret_cast<const long long*>(date_off), x, mean, sqdev);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_winsorize_apply(bmfm_ctx* ctx, int T, const int64_t* date_off, const double* x, const double* lo,
                                    const double* hi, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && date_off && x && lo && hi && out, "bad arguments");
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int grid = T < ctx->sm_count * 8 ? T : ctx->sm_count * 8;
  winsorize_apply_kernel<<<grid, 256, 0, ctx->stream>>>(T, reinterpret_cast<const long long*>(date_off), x, lo, hi, out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// pass 0: sums [T][k + 2] (complete rows, sum X_1..X_k, sum y);  pass 1: gram [T][(k+1)(k+2)/2 + k + 1] about
// shift [T][k + 1];  pass 2: residuals from the (all-reduced) gram and count [T] (doubles).
extern "C" int bmfm_orthogonalize_pass(bmfm_ctx* ctx, int pass, int T, const int64_t* date_off, const double* y, int k,
                                       const double* const* X, const double* shift, const double* gram,
                                       const double* count, double* out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && k >= 1 && pass >= 0 && pass <= 2 && date_off && y && X && out, "bad arguments");
  BMFM_ARG(ctx, pass == 0 || shift, "shift missing");
  BMFM_ARG(ctx, pass < 2 || (gram && count), "gram / count missing");
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
  const int grid = T < ctx->sm_count * 8 ? T : ctx->sm_count * 8;
  switch (k) {
    case 1: ortho_shard_launch<1>(ctx, p, grid, pass, shift, gram, count, out); break;
    case 2: ortho_shard_launch<2>(ctx, p, grid, pass, shift, gram, count, out); break;
    case 3: ortho_shard_launch<3>(ctx, p, grid, pass, shift, gram, count, out); break;
    case 4: ortho_shard_launch<4>(ctx, p, grid, pass, shift, gram, count, out); break;
    case 5: ortho_shard_launch<5>(ctx, p, grid, pass, shift, gram, count, out); break;
    default: ortho_shard_launch<6>(ctx, p, grid, pass, shift, gram, count, out); break;
  }
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
