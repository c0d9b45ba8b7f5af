// Bayesian shrinkage of specific volatility toward the cap-decile mean, one CTA per date.
//
// Replaces  utils.bayes_shrink (+ group_mean_std, shrink)   Barra-master/mfm/utils.py:133-168
//
//   group = pd.qcut(capital, ngroup).codes
//        edges_k  = np.percentile(capital, linspace(0,1,ngroup+1)[k] * 100)   (pandas -> numpy 'linear':
//                   virtual index (n-1) q, gamma = frac, numpy's two-sided _lerp)
//        code(x)  = searchsorted(edges, x, 'left') - 1, x == edges[0] -> 0     (right-closed, lowest included)
//        duplicate edges -> ValueError("Bin edges must be unique")
//   m_g = sum(vol cap) / sum(cap),  s_g = sqrt(mean((vol - m_g)^2))             (utils.py:137-138)
//   a = q |vol - m_g|, v = a / (a + s_g), out = v m_g + (1 - v) |vol|            (utils.py:146-149)
// The decile codes are integer work and are reproduced bit-exactly: the quantile positions use
// the same IEEE operations in the same order as numpy (no FMA contraction), on caps sorted by a
// shared-memory bitonic network.
#include <math.h>
#include "bmfm_common.cuh"

namespace bmfm {

constexpr int BS_MAX_GROUPS = 64;
constexpr int BS_ST_DUP_EDGES = 5;

struct BsParams {
  int T, ngroup;
  double q;
  const long long* date_off;
  const double* vol;
  const double* cap;
  double* out;
  int* status;
  int npad;                       // power of two >= max rows per date
  double quant[BS_MAX_GROUPS + 1];  // (linspace(0,1,ngroup+1) * 100) / 100, host arithmetic
};

__global__ void __launch_bounds__(1024) bayes_shrink_kernel(BsParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  double* s_sort = reinterpret_cast<double*>(smem);   // npad
  double* s_edges = s_sort + p.npad;                  // ngroup + 1
  double* s_gsum = s_edges + BS_MAX_GROUPS + 1;       // [ngroup][4]: sum vol*cap, sum cap, count, sum (vol-m)^2
  int* s_flag = reinterpret_cast<int*>(s_gsum + 4 * BS_MAX_GROUPS);
  const int tid = threadIdx.x, nt = blockDim.x, G = p.ngroup;
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    for (int i = tid; i < p.npad; i += nt) s_sort[i] = i < n ? p.cap[a + i] : inf;
    for (int i = tid; i < 4 * G; i += nt) s_gsum[i] = 0.0;
    if (tid == 0) *s_flag = 0;
    __syncthreads();
    // bitonic sort ascending
    for (int k = 2; k <= p.npad; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int i = tid; i < p.npad; i += nt) {
          const int ixj = i ^ j;
          if (ixj > i) {
            const double x = s_sort[i], y = s_sort[ixj];
            const bool up = (i & k) == 0;
            if ((x > y) == up) { s_sort[i] = y; s_sort[ixj] = x; }
          }
        }
        __syncthreads();
      }
    }
    // quantile edges (numpy 'linear' method, _lerp with the t >= 0.5 branch)
    if (tid <= G && n > 0) {
      const double virt = __dmul_rn((double)(n - 1), p.quant[tid]);
      double prev = floor(virt);
      const double gamma = __dsub_rn(virt, prev);
      int ip = (int)prev, in = ip + 1;
      if (virt >= (double)(n - 1)) ip = in = n - 1;
      if (virt < 0.0) ip = in = 0;
      const double lo = s_sort[ip], hi = s_sort[in];
      const double diff = __dsub_rn(hi, lo);
      double e = __dadd_rn(lo, __dmul_rn(diff, gamma));
      if (gamma >= 0.5) e = __dsub_rn(hi, __dmul_rn(diff, __dsub_rn(1.0, gamma)));
      s_edges[tid] = e;
    }
    __syncthreads();
    if (tid < G && n > 0 && !(s_edges[tid] < s_edges[tid + 1])) *s_flag = 1;   // duplicate (or NaN) edges
    __syncthreads();
    if (*s_flag || n == 0) {
      const double nan = __longlong_as_double(0x7ff8000000000000LL);
      for (int i = tid; i < n; i += nt) p.out[a + i] = nan;
      if (tid == 0) p.status[t] = n == 0 ? BMFM_ST_OK : BS_ST_DUP_EDGES;
      __syncthreads();
      continue;
    }
    // pass 1: group sums
    const double e0 = s_edges[0];
    for (int i = tid; i < n; i += nt) {
      const double c = p.cap[a + i], v = p.vol[a + i];
      int cnt = 0;
      for (int k = 0; k <= G; ++k) cnt += s_edges[k] < c;   // searchsorted(edges, c, 'left')
      int g = cnt - 1;
      if (c == e0) g = 0;
      if (g >= 0 && g < G) {
        atomicAdd(&s_gsum[4 * g + 0], v * c);
        atomicAdd(&s_gsum[4 * g + 1], c);
        atomicAdd(&s_gsum[4 * g + 2], 1.0);
      }
    }
    __syncthreads();
    // pass 2: dispersion about the group mean
    for (int i = tid; i < n; i += nt) {
      const double c = p.cap[a + i], v = p.vol[a + i];
      int cnt = 0;
      for (int k = 0; k <= G; ++k) cnt += s_edges[k] < c;
      int g = cnt - 1;
      if (c == e0) g = 0;
      if (g >= 0 && g < G) {
        const double m = s_gsum[4 * g + 0] / s_gsum[4 * g + 1];
        const double d = v - m;
        atomicAdd(&s_gsum[4 * g + 3], d * d);
      }
    }
    __syncthreads();
    for (int i = tid; i < n; i += nt) {
      const double c = p.cap[a + i], v = p.vol[a + i];
      int cnt = 0;
      for (int k = 0; k <= G; ++k) cnt += s_edges[k] < c;
      int g = cnt - 1;
      if (c == e0) g = 0;
      double o = __longlong_as_double(0x7ff8000000000000LL);
      if (g >= 0 && g < G) {
        const double m = s_gsum[4 * g + 0] / s_gsum[4 * g + 1];
        const double s = sqrt(s_gsum[4 * g + 3] / s_gsum[4 * g + 2]);
        const double aa = p.q * fabs(v - m);
        const double w = aa / (aa + s);
        o = w * m + (1.0 - w) * fabs(v);
      }
      p.out[a + i] = o;
    }
    if (tid == 0) p.status[t] = BMFM_ST_OK;
    __syncthreads();
  }
}

}  // namespace bmfm

using namespace bmfm;

extern "C" int bmfm_bayes_shrink(bmfm_ctx* ctx, int T, int max_rows_per_date, const int64_t* date_off,
                                 const double* volatility, const double* capital, int ngroup, double q,
                                 double* out, int32_t* status) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && max_rows_per_date >= 0 && ngroup >= 1, "bad sizes");
  BMFM_ARG(ctx, date_off && volatility && capital && out && status, "NULL pointer");
  if (ngroup > BS_MAX_GROUPS) {
    snprintf(ctx->err, sizeof(ctx->err), "ngroup=%d exceeds compiled limit %d", ngroup, BS_MAX_GROUPS);
    return BMFM_E_UNSUPPORTED;
  }
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  BsParams p;
  p.T = T; p.ngroup = ngroup; p.q = q;
  p.date_off = reinterpret_cast<const long long*>(date_off);
  p.vol = volatility; p.cap = capital; p.out = out; p.status = status;
  int npad = 32;
  while (npad < max_rows_per_date) npad <<= 1;
  p.npad = npad;
  // np.linspace(0, 1, ngroup + 1): k * step, last = 1; pandas multiplies by 100, numpy divides by 100
  volatile double step = 1.0 / (double)ngroup;
  for (int k = 0; k <= ngroup; ++k) {
    volatile double y = (k == ngroup) ? 1.0 : (double)k * step;
    volatile double pct = y * 100.0;
    p.quant[k] = pct / 100.0;
  }
  const size_t smem = ((size_t)npad + BS_MAX_GROUPS + 1 + 4 * BS_MAX_GROUPS) * 8 + 16;
  if (smem > (size_t)ctx->max_smem_optin) {
    snprintf(ctx->err, sizeof(ctx->err), "bayes_shrink: %d rows per date exceed the shared-memory sort (max 16384)",
             max_rows_per_date);
    return BMFM_E_UNSUPPORTED;
  }
  BMFM_CUDA(ctx, cudaFuncSetAttribute(bayes_shrink_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int nthreads = npad >= 2048 ? 1024 : (npad >= 512 ? 256 : 128);
  int per_sm = (int)((228 * 1024) / (smem + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm * nthreads > 2048) per_sm = 2048 / nthreads;
  int grid = ctx->sm_count * per_sm;
  if (grid > T) grid = T;
  bayes_shrink_kernel<<<grid, nthreads, smem, ctx->stream>>>(p);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
