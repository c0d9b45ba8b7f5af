// Bias statistic of the eigenfactor portfolios.
//
// Replaces  utils.eigenfactor_bias_stat   Barra-master/mfm/utils.py:97-117   (used MFM.py:203-204)
//
// For every date i < T - predlen with a covariance:  D, U = eig(cov_i);  U /= column sums;
//   sigma = sqrt(predlen diag(U' cov_i U));  r = U' ((1 + ret)[i+1 .. i+predlen].prod - 1);  b_i = r / sigma
// and the result is np.std(b, axis 0) over those dates (dates whose covariance is missing are skipped by the
// reference's try/except).  With unit eigenvectors u_k and s_k = sum_j u_jk the scaling cancels:
//   b_ik = sign(s_k) (u_k . retlen_i) / sqrt(predlen lambda_k),
// independent of the eigenvector's sign.  Column k is the k-th LARGEST eigenvalue (canonical order; numpy's
// eig returns LAPACK's order).  The decomposition is bmfm_eigen_decompose's.
#include <math.h>
#include "bmfm_common.cuh"

extern "C" int bmfm_eigen_decompose(bmfm_ctx* ctx, int T, int K, const double* cov, const int32_t* cov_status,
                                    double* D0, double* U0, int32_t* status);

namespace bmfm {

// one CTA per date, thread k = eigenfactor k
__global__ void bias_b_kernel(int n, int K, int predlen, const double* D0, const double* U0, const int* status,
                              const double* fr, double* b, int* valid) {
  extern __shared__ double s_ret[];      // [K] compounded returns of rows i+1 .. i+predlen
  const int i = blockIdx.x;
  if (i >= n) return;
  const int st = status[i];
  const bool have = st == BMFM_ST_OK || st == BMFM_ST_PINV || st == BMFM_ST_NOT_PSD;
  if (threadIdx.x == 0) valid[i] = have ? 1 : 0;
  if (!have) return;
  for (int j = threadIdx.x; j < K; j += blockDim.x) {
    double p = 1.0;
    for (int s = 1; s <= predlen; ++s) p *= fr[(size_t)(i + s) * K + j] + 1.0;
    s_ret[j] = p - 1.0;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const double* U = U0 + (size_t)i * K * K + k;
    double colsum = 0.0, dot = 0.0;
    for (int j = 0; j < K; ++j) {
      const double u = U[(size_t)j * K];
      colsum += u;
      dot = fma(u, s_ret[j], dot);
    }
    const double sg = colsum > 0.0 ? 1.0 : (colsum < 0.0 ? -1.0 : 0.0);
    b[(size_t)i * K + k] = sg * dot / sqrt((double)predlen * D0[(size_t)i * K + k]);
  }
}

// np.std(b, axis 0) over the valid dates: one CTA per eigenfactor, two passes (mean, then squared deviations)
__global__ void __launch_bounds__(256) bias_std_kernel(int n, int K, const double* b, const int* valid, double* out) {
  __shared__ double red[8];
  __shared__ double s_mean, s_cnt;
  const int k = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double s = 0.0, c = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x)
    if (valid[i]) { s += b[(size_t)i * K + k]; c += 1.0; }
  s = warp_sum(s);
  c = warp_sum(c);
  if (lane == 0) { red[warp] = s; }
  __syncthreads();
  if (threadIdx.x == 0) { double t = 0.0; for (int w = 0; w < 8; ++w) t += red[w]; s_mean = t; }
  __syncthreads();
  if (lane == 0) { red[warp] = c; }
  __syncthreads();
  if (threadIdx.x == 0) { double t = 0.0; for (int w = 0; w < 8; ++w) t += red[w]; s_cnt = t; s_mean = s_mean / t; }
  __syncthreads();
  const double mean = s_mean;
  double q = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x)
    if (valid[i]) { const double d = b[(size_t)i * K + k] - mean; q = fma(d, d, q); }
  q = warp_sum(q);
  if (lane == 0) red[warp] = q;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    out[k] = s_cnt > 0.0 ? sqrt(t / s_cnt) : __longlong_as_double(0x7ff8000000000000LL);
  }
}

}  // namespace bmfm

using namespace bmfm;

// cov double [T*K*K] with cov_status [T] (NULL = all present), factor_ret double [T*K]; bias_stat double [K];
// b_out (optional, may be NULL) double [(T-predlen)*K] receives the per-date b (rows of skipped dates untouched).
extern "C" int bmfm_eigenfactor_bias_stat(bmfm_ctx* ctx, int T, int K, int predlen, const double* cov,
                                          const int32_t* cov_status, const double* factor_ret, double* bias_stat,
                                          double* b_out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && K > 0 && predlen >= 1, "bad sizes");
  BMFM_ARG(ctx, cov && factor_ret && bias_stat, "NULL pointer");
  const int n = T - predlen;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n <= 0) {
    // no usable date: np.std of an empty array -> NaN
    const double nanv = nan("");
    for (int k = 0; k < K; ++k) BMFM_CUDA(ctx, cudaMemcpyAsync(bias_stat + k, &nanv, 8, cudaMemcpyHostToDevice, ctx->stream));
    BMFM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BMFM_OK;
  }
  // scratch in the context workspace (bmfm_eigen_decompose does not use it): D0 [n*K], U0 [n*K*K], b [n*K],
  // status [n], valid [n]
  const size_t nd = (size_t)n * K + (size_t)n * K * K + (b_out ? 0 : (size_t)n * K);
  int rc = bmfm_ws_reserve(ctx, nd * sizeof(double) + (size_t)2 * n * sizeof(int));
  if (rc) return rc;
  double* buf = static_cast<double*>(ctx->ws);
  double* D0 = buf;
  double* U0 = D0 + (size_t)n * K;
  double* b = b_out ? b_out : U0 + (size_t)n * K * K;
  int* status = reinterpret_cast<int*>(buf + nd);
  int* valid = status + n;
  rc = bmfm_eigen_decompose(ctx, n, K, cov, cov_status, D0, U0, status);
  if (rc == BMFM_OK) {
    const int threads = K <= 32 ? 32 : (K <= 64 ? 64 : 128);
    bias_b_kernel<<<n, threads, K * sizeof(double), ctx->stream>>>(n, K, predlen, D0, U0, status, factor_ret, b, valid);
    ctx->launches++;
    bias_std_kernel<<<K, 256, 0, ctx->stream>>>(n, K, b, valid, bias_stat);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { snprintf(ctx->err, sizeof(ctx->err), "bias_stat launch: %s", cudaGetErrorString(e)); rc = (int)e; }
  }
  return rc;
}
