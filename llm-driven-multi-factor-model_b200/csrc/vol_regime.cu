// Volatility-regime adjustment of the factor covariance series.
//
// Replaces  MFM.vol_regime_adj_by_time   Barra-master/mfm/MFM.py:130-167
//
//   var_t   = diag(cov_t), a NaN row where the date has no covariance (MFM.py:142-148)
//   B_t^2   = mean_k f_tk^2 / var_tk            (pandas mean: NaN entries skipped, :149)
//   w_s     = 0.5^((T-1-s)/tao)
//   lamb_t  = sqrt( sum_{s<=t, var_s valid} w_s B_s^2 / sum_{s<=t, var_s valid} w_s ), 0 when no valid row
//   out_t   = cov_t * lamb_t^2
// The reference's O(T^2) python loop is two prefix sums.  Two launches: the per-date bias
// statistic, then one kernel in which every CTA first adds the terms of all earlier dates and then
// streams its own dates' matrices (T*K*K doubles, HBM-bound: 16 B per matrix element).  The separate
// one-CTA scan and scale kernels remain for the gathered multi-GPU form and as a comparison switch.
#include <math.h>
#include "bmfm_common.cuh"

namespace bmfm {

// one warp per date
struct VrPeers {
  int n;
  double* wb[8];
  double* wv[8];
};

__global__ void vr_bias_kernel(int T, int K, double tao, const double* fr, const double* cov,
                               const int* cov_status, double* wb, double* wv, int t_off, int T_total,
                               VrPeers peers) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= T) return;
  const int t = warp;
  const bool have = cov_status == nullptr || cov_status[t] == BMFM_ST_OK || cov_status[t] == BMFM_ST_PINV;
  double sum = 0.0;
  int cnt = 0, nanvar = 0;
  if (have) {
    for (int k = lane; k < K; k += 32) {
      const double v = cov[((size_t)t * K + k) * K + k];
      const double f = fr[(size_t)t * K + k];
      if (isnan(v)) nanvar = 1;
      const double r = f * f / v;
      if (!isnan(r)) { sum += r; cnt++; }
    }
  } else {
    nanvar = 1;
  }
  sum = warp_sum(sum);
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  nanvar = __any_sync(0xffffffffu, nanvar);
  if (lane == 0) {
    // B_t^2; the reference takes sqrt(mean) and squares it again (MFM.py:149,160), so a negative mean (a negative
    // variance on the diagonal) is NaN there and poisons every later lamb
    double b2 = cnt > 0 ? sum / cnt : __longlong_as_double(0x7ff8000000000000LL);
    if (b2 < 0.0) b2 = __longlong_as_double(0x7ff8000000000000LL);
    const double w = exp2(-(double)(T_total - 1 - (t_off + t)) / tao);      // 0.5^((T-1-t)/tao)
    // rows whose variance vector holds a NaN are excluded from both sums (MFM.py:158)
    const double ob = nanvar ? 0.0 : w * b2, ov = nanvar ? 0.0 : w;
    if (wb) { wb[t] = ob; wv[t] = ov; }
    for (int r = 0; r < peers.n; ++r) {     // fused all-gather: straight into every rank's full-length arrays
      peers.wb[r][t_off + t] = ob;
      peers.wv[r][t_off + t] = ov;
    }
  }
}

// single CTA inclusive scan of (wb, wv) -> lamb: each thread owns a contiguous run of dates (its loads are
// independent, issued eight at a time), the 1024 run totals are scanned with warp shuffles
__global__ void __launch_bounds__(1024) vr_scan_kernel(int T, const double* wb, const double* wv, double* lamb) {
  __shared__ double s_b[32], s_v[32];
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const int per = (T + nt - 1) / nt;
  const int lo = min(T, tid * per), hi = min(T, lo + per);
  double pb = 0.0, pv = 0.0;
  for (int t = lo; t < hi; t += 8) {
    double xb[8], xv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      xb[j] = t + j < hi ? wb[t + j] : 0.0;
      xv[j] = t + j < hi ? wv[t + j] : 0.0;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { pb += xb[j]; pv += xv[j]; }
  }
  // inclusive scan of the per-thread totals: inside each warp, then over the 32 warp totals
  double ib = pb, iv = pv;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const double ab = __shfl_up_sync(0xffffffffu, ib, off), av = __shfl_up_sync(0xffffffffu, iv, off);
    if (lane >= off) { ib += ab; iv += av; }
  }
  if (lane == 31) { s_b[warp] = ib; s_v[warp] = iv; }
  __syncthreads();
  if (warp == 0) {
    double wbt = lane < (nt >> 5) ? s_b[lane] : 0.0, wvt = lane < (nt >> 5) ? s_v[lane] : 0.0;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const double ab = __shfl_up_sync(0xffffffffu, wbt, off), av = __shfl_up_sync(0xffffffffu, wvt, off);
      if (lane >= off) { wbt += ab; wvt += av; }
    }
    s_b[lane] = wbt;
    s_v[lane] = wvt;
  }
  __syncthreads();
  double rb = (warp > 0 ? s_b[warp - 1] : 0.0) + (ib - pb);   // exclusive prefix of this thread's run
  double rv = (warp > 0 ? s_v[warp - 1] : 0.0) + (iv - pv);
  for (int t = lo; t < hi; ++t) {
    rb += wb[t];
    rv += wv[t];
    lamb[t] = rv > 0.0 ? sqrt(rb / rv) : 0.0;
  }
}

// out = cov * lamb_t^2: one CTA per date (no per-element index division), 16-byte accesses when K*K is even
__global__ void __launch_bounds__(256) vr_scale_kernel(int T, int KK, const double* cov, const int* cov_status,
                                                       const double* lamb, double* out) {   // lamb offset to this block
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const bool have = cov_status == nullptr || cov_status[t] == BMFM_ST_OK || cov_status[t] == BMFM_ST_PINV;
    const double l = lamb[t], l2 = l * l;
    const double* src = cov + (size_t)t * KK;
    double* dst = out + (size_t)t * KK;
    if ((KK & 1) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
      const double2* s2 = reinterpret_cast<const double2*>(src);
      for (int e = threadIdx.x; e < (KK >> 1); e += blockDim.x) {
        double2 v = make_double2(nan, nan);
        if (have) { v = __ldcs(s2 + e); v.x *= l2; v.y *= l2; }
        st_stream2(dst + 2 * e, v.x, v.y);
      }
    } else {
      for (int e = threadIdx.x; e < KK; e += blockDim.x) st_stream(dst + e, have ? __ldcs(src + e) * l2 : nan);
    }
  }
}

// ---- date-sharded form without the replicated scan ------------------------------------------------
// lamb_t^2 is a ratio of two prefix sums over the dates up to t, so a rank needs from every EARLIER rank only the
// two totals of its block.  vr_block_totals_kernel reduces this rank's terms and stores the pair into slot `rank` of
// every rank's record buffer (peer memory); vr_scan_carry_kernel then scans the rank's own dates starting from the
// sum of the earlier ranks' pairs.
struct VrTotPeers {
  int n;
  double* tot[8];
};

__global__ void __launch_bounds__(1024) vr_block_totals_kernel(int n, const double* wb, const double* wv, int rank,
                                                               VrTotPeers peers, BmfmFlagPeers fp) {
  __shared__ double s_b[32], s_v[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double pb = 0.0, pv = 0.0;
  for (int t = tid; t < n; t += blockDim.x) { pb += wb[t]; pv += wv[t]; }
  pb = warp_sum(pb);
  pv = warp_sum(pv);
  if (lane == 0) { s_b[warp] = pb; s_v[warp] = pv; }
  __syncthreads();
  if (warp == 0) {
    pb = lane < (blockDim.x >> 5) ? s_b[lane] : 0.0;
    pv = lane < (blockDim.x >> 5) ? s_v[lane] : 0.0;
    pb = warp_sum(pb);
    pv = warp_sum(pv);
    if (lane < peers.n) {
      peers.tot[lane][2 * rank] = pb;
      peers.tot[lane][2 * rank + 1] = pv;
    }
    if (fp.n > 0) {
      __syncwarp();
      if (lane == 0) {
        volatile long long* mine = fp.f[fp.rank];
        const long long e = mine[32 + BMFM_CH_VR] + 1;
        mine[32 + BMFM_CH_VR] = e;
        bmfm_flag_publish(fp, BMFM_CH_VR, e);
      }
    }
  }
}

__global__ void __launch_bounds__(1024) vr_scan_carry_kernel(int n, const double* wb, const double* wv,
                                                             const double* totals, int rank, double* lamb,
                                                             BmfmFlagPeers fp) {
  __shared__ double s_b[32], s_v[32];
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  if (fp.n > 0) {
    if (tid == 0) bmfm_flag_wait(fp, BMFM_CH_VR);
    __syncthreads();
  }
  double cb = 0.0, cv = 0.0;
  for (int r = 0; r < rank; ++r) { cb += __ldcg(totals + 2 * r); cv += __ldcg(totals + 2 * r + 1); }
  const int per = (n + nt - 1) / nt;
  const int lo = min(n, tid * per), hi = min(n, lo + per);
  double pb = 0.0, pv = 0.0;
  for (int t = lo; t < hi; ++t) { pb += wb[t]; pv += wv[t]; }
  double ib = pb, iv = pv;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const double ab = __shfl_up_sync(0xffffffffu, ib, off), av = __shfl_up_sync(0xffffffffu, iv, off);
    if (lane >= off) { ib += ab; iv += av; }
  }
  if (lane == 31) { s_b[warp] = ib; s_v[warp] = iv; }
  __syncthreads();
  if (warp == 0) {
    double wbt = lane < (nt >> 5) ? s_b[lane] : 0.0, wvt = lane < (nt >> 5) ? s_v[lane] : 0.0;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const double ab = __shfl_up_sync(0xffffffffu, wbt, off), av = __shfl_up_sync(0xffffffffu, wvt, off);
      if (lane >= off) { wbt += ab; wvt += av; }
    }
    s_b[lane] = wbt;
    s_v[lane] = wvt;
  }
  __syncthreads();
  double rb = cb + (warp > 0 ? s_b[warp - 1] : 0.0) + (ib - pb);
  double rv = cv + (warp > 0 ? s_v[warp - 1] : 0.0) + (iv - pv);
  for (int t = lo; t < hi; ++t) {
    rb += wb[t];
    rv += wv[t];
    lamb[t] = rv > 0.0 ? sqrt(rb / rv) : 0.0;
  }
}

// Scan and scale in one launch: CTA b owns the dates [b D, (b+1) D).  lamb_t^2 is a ratio of two prefix sums, so the
// CTA first adds the terms of all earlier dates (and, date-sharded, the totals of the earlier ranks) — a few loads
// per thread from L2, all in flight together — then walks its own dates and streams their matrices.  This replaces
// a one-CTA scan kernel (6 us at T = 2500, all other SMs idle) and the launch after it.
constexpr int VR_FUSED_MAXD = 32;
__global__ void __launch_bounds__(256) vr_scan_scale_kernel(int t_first, int n, int KK, int D, const double* __restrict__ wb,
                                                            const double* __restrict__ wv, const double* totals, int rank,
                                                            BmfmFlagPeers fp, const double* cov, const int* cov_status,
                                                            double* lamb, double* out) {
  __shared__ double s_b[8], s_v[8], s_wb[VR_FUSED_MAXD], s_wv[VR_FUSED_MAXD], s_l2[VR_FUSED_MAXD];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int d0 = t_first + blockIdx.x * D, d1 = min(n, d0 + D);      // dates [t_first, n) of the arrays are this launch's
  if (fp.n > 0) {
    if (tid == 0) bmfm_flag_wait(fp, BMFM_CH_VR);
    __syncthreads();
  }
  double pb = 0.0, pv = 0.0;
  if (tid < rank) { pb = __ldcg(totals + 2 * tid); pv = __ldcg(totals + 2 * tid + 1); }
  for (int t = tid; t < d0; t += 1024) {
    double xb[4], xv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int tt = t + 256 * j;
      xb[j] = tt < d0 ? wb[tt] : 0.0;
      xv[j] = tt < d0 ? wv[tt] : 0.0;
    }
    pb += (xb[0] + xb[1]) + (xb[2] + xb[3]);
    pv += (xv[0] + xv[1]) + (xv[2] + xv[3]);
  }
  if (tid < d1 - d0) { s_wb[tid] = wb[d0 + tid]; s_wv[tid] = wv[d0 + tid]; }
  pb = warp_sum(pb);
  pv = warp_sum(pv);
  if (lane == 0) { s_b[warp] = pb; s_v[warp] = pv; }
  __syncthreads();
  if (tid == 0) {
    double rb = 0.0, rv = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { rb += s_b[w]; rv += s_v[w]; }
    for (int j = 0; j < d1 - d0; ++j) {
      rb += s_wb[j];
      rv += s_wv[j];
      const double l = rv > 0.0 ? sqrt(rb / rv) : 0.0;
      lamb[d0 + j] = l;
      s_l2[j] = l * l;
    }
  }
  __syncthreads();
  if (out == nullptr) return;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int t = d0; t < d1; ++t) {
    const bool have = cov_status == nullptr || cov_status[t] == BMFM_ST_OK || cov_status[t] == BMFM_ST_PINV;
    const double l2 = s_l2[t - d0];
    const double* src = cov + (size_t)t * KK;
    double* dst = out + (size_t)t * KK;
    if ((KK & 1) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
      const double2* s2 = reinterpret_cast<const double2*>(src);
      const int half = KK >> 1;
      for (int base = tid; base < half; base += 1024) {
        double2 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[j] = make_double2(nan, nan);
          if (have && base + 256 * j < half) v[j] = __ldcs(s2 + base + 256 * j);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (base + 256 * j < half) {
            if (have) { v[j].x *= l2; v[j].y *= l2; }
            st_stream2(dst + 2 * (base + 256 * j), v[j].x, v[j].y);
          }
      }
    } else {
      for (int e = tid; e < KK; e += 256) st_stream(dst + e, have ? __ldcs(src + e) * l2 : nan);
    }
  }
}

// dates per CTA of vr_scan_scale_kernel: about eight CTAs per SM
static inline int vr_fused_dates(const bmfm_ctx* ctx, int n) {
  const int ctas = 8 * (ctx->sm_count > 0 ? ctx->sm_count : 148);
  int D = (n + ctas - 1) / ctas;
  if (D < 1) D = 1;
  if (D > VR_FUSED_MAXD) D = VR_FUSED_MAXD;
  return D;
}

}  // namespace bmfm

using namespace bmfm;

extern "C" int bmfm_vol_regime_adj_by_time(bmfm_ctx* ctx, int T, int K, double tao, const double* factor_ret,
                                           const double* cov, const int32_t* cov_status, double* out_cov,
                                           double* lamb) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && K > 0 && tao > 0, "bad sizes");
  BMFM_ARG(ctx, factor_ret && cov && lamb, "NULL pointer");
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = bmfm_ws_reserve(ctx, (size_t)2 * T * sizeof(double));
  if (rc) return rc;
  double* wb = (double*)ctx->ws;
  double* wv = wb + T;
  VrPeers nopeers;
  nopeers.n = 0;
  vr_bias_kernel<<<(T * 32 + 255) / 256, 256, 0, ctx->stream>>>(T, K, tao, factor_ret, cov, cov_status, wb, wv, 0, T, nopeers);
  ctx->launches++;
  {
    const int D = vr_fused_dates(ctx, T);
    BmfmFlagPeers noflags;
    noflags.n = 0;
    vr_scan_scale_kernel<<<(T + D - 1) / D, 256, 0, ctx->stream>>>(0, T, K * K, D, wb, wv, nullptr, 0, noflags, cov, cov_status, lamb, out_cov);
    ctx->launches++;
  }
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// Dates [t0, t1) of a series of T dates whose earlier dates were already processed by calls of this function with
// the same wb / wv arrays ([T] each, device, kept by the caller between the calls): the pipelined host entry point
// adjusts each block of dates as soon as its covariances exist.  All pointers address the FULL series.
extern "C" int bmfm_vol_regime_adj_range(bmfm_ctx* ctx, int T, int t0, int t1, int K, double tao, const double* factor_ret,
                                         const double* cov, const int32_t* cov_status, double* wb, double* wv,
                                         double* out_cov, double* lamb) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && t0 >= 0 && t1 >= t0 && t1 <= T && K > 0 && tao > 0, "bad sizes");
  if (t1 == t0) return BMFM_OK;
  BMFM_ARG(ctx, factor_ret && cov && wb && wv && lamb, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int n = t1 - t0;
  VrPeers nopeers;
  nopeers.n = 0;
  vr_bias_kernel<<<(n * 32 + 255) / 256, 256, 0, ctx->stream>>>(n, K, tao, factor_ret + (size_t)t0 * K, cov + (size_t)t0 * K * K,
                                                              cov_status ? cov_status + t0 : nullptr, wb + t0, wv + t0, t0, T, nopeers);
  ctx->launches++;
  const int D = vr_fused_dates(ctx, n);
  BmfmFlagPeers noflags;
  noflags.n = 0;
  vr_scan_scale_kernel<<<(n + D - 1) / D, 256, 0, ctx->stream>>>(t0, t1, K * K, D, wb, wv, nullptr, 0, noflags, cov, cov_status, lamb, out_cov);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// ---- the same pass split for date-sharded runs -------------------------------------------------
// bmfm_vol_regime_bias: per-date weighted bias terms of this rank's block of n dates that starts at
//   global date t0 of T_total:  wb[i] = w_t B_t^2, wv[i] = w_t (0 where the variance row is invalid).
// The caller gathers wb / wv of all ranks in date order and calls
// bmfm_vol_regime_scale: lamb over ALL T_total dates from the gathered terms, then this rank's
//   block scaled by its own lamb values.  lamb_all [T_total]; out_cov [n*K*K] may be NULL.
extern "C" int bmfm_vol_regime_bias(bmfm_ctx* ctx, int T_total, int t0, int n, int K, double tao,
                                    const double* factor_ret_block, const double* cov_block,
                                    const int32_t* cov_status_block, double* wb, double* wv) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T_total >= 0 && t0 >= 0 && n >= 0 && t0 + n <= T_total && K > 0 && tao > 0, "bad sizes");
  BMFM_ARG(ctx, n == 0 || (factor_ret_block && cov_block && wb && wv), "NULL pointer");
  if (n == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  VrPeers nopeers;
  nopeers.n = 0;
  vr_bias_kernel<<<(n * 32 + 255) / 256, 256, 0, ctx->stream>>>(n, K, tao, factor_ret_block, cov_block,
                                                              cov_status_block, wb, wv, t0, T_total, nopeers);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// bmfm_vol_regime_bias with the all-gather fused in: the terms of this rank's block are stored into every
// rank's full-length wb_all / wv_all (HOST arrays of n_peer device pointers, peer-mapped) at offset t0.
extern "C" int bmfm_vol_regime_bias_gather(bmfm_ctx* ctx, int T_total, int t0, int n, int K, double tao,
                                           const double* factor_ret_block, const double* cov_block,
                                           const int32_t* cov_status_block, int n_peer, double* const* peer_wb,
                                           double* const* peer_wv) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T_total >= 0 && t0 >= 0 && n >= 0 && t0 + n <= T_total && K > 0 && tao > 0, "bad sizes");
  BMFM_ARG(ctx, n_peer >= 1 && n_peer <= 8 && peer_wb && peer_wv, "bad peer list");
  BMFM_ARG(ctx, n == 0 || (factor_ret_block && cov_block), "NULL pointer");
  if (n == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  VrPeers peers;
  peers.n = n_peer;
  for (int r = 0; r < 8; ++r) { peers.wb[r] = r < n_peer ? peer_wb[r] : nullptr; peers.wv[r] = r < n_peer ? peer_wv[r] : nullptr; }
  vr_bias_kernel<<<(n * 32 + 255) / 256, 256, 0, ctx->stream>>>(n, K, tao, factor_ret_block, cov_block,
                                                              cov_status_block, nullptr, nullptr, t0, T_total, peers);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_vol_regime_scale(bmfm_ctx* ctx, int T_total, int t0, int n, int K, const double* wb_all,
                                     const double* wv_all, const double* cov_block,
                                     const int32_t* cov_status_block, double* out_cov_block, double* lamb_all) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T_total >= 0 && t0 >= 0 && n >= 0 && t0 + n <= T_total && K > 0, "bad sizes");
  BMFM_ARG(ctx, wb_all && wv_all && lamb_all, "NULL pointer");
  if (T_total == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  vr_scan_kernel<<<1, 1024, 0, ctx->stream>>>(T_total, wb_all, wv_all, lamb_all);
  ctx->launches++;
  if (out_cov_block && n > 0) {
    vr_scale_kernel<<<n, 256, 0, ctx->stream>>>(n, K * K, cov_block, cov_status_block, lamb_all + t0, out_cov_block);
    ctx->launches++;
  }
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// Date-sharded vol-regime by exchanged block totals (two doubles per rank) instead of gathered per-date terms:
//   1. bmfm_vol_regime_shard_totals: the terms wb / wv [n] of this rank's block (kept local) and their two sums,
//      stored into doubles (2 rank, 2 rank + 1) of every buffer in peer_totals (HOST array of n_peer device pointers
//      to [2 * n_rank] doubles, peer-mapped);
//   2. the caller runs a cross-rank barrier;
//   3. bmfm_vol_regime_shard_scale: lamb of this rank's dates from its own terms and the totals of the ranks before
//      it (`totals`: this rank's own record buffer), then its block scaled.  lamb_block [n].
extern "C" int bmfm_vol_regime_shard_totals(bmfm_ctx* ctx, int T_total, int t0, int n, int K, double tao,
                                            const double* factor_ret_block, const double* cov_block,
                                            const int32_t* cov_status_block, double* wb, double* wv, int rank,
                                            int n_peer, double* const* peer_totals) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T_total >= 0 && t0 >= 0 && n >= 0 && t0 + n <= T_total && K > 0 && tao > 0 && rank >= 0, "bad sizes");
  BMFM_ARG(ctx, n_peer >= 1 && n_peer <= 8 && peer_totals, "bad peer list");
  BMFM_ARG(ctx, n == 0 || (factor_ret_block && cov_block && wb && wv), "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  VrPeers nopeers;
  nopeers.n = 0;
  if (n > 0) {
    vr_bias_kernel<<<(n * 32 + 255) / 256, 256, 0, ctx->stream>>>(n, K, tao, factor_ret_block, cov_block,
                                                                cov_status_block, wb, wv, t0, T_total, nopeers);
    ctx->launches++;
  }
  VrTotPeers tp;
  tp.n = n_peer;
  for (int r = 0; r < 8; ++r) tp.tot[r] = r < n_peer ? peer_totals[r] : nullptr;
  vr_block_totals_kernel<<<1, 1024, 0, ctx->stream>>>(n, wb, wv, rank, tp, bmfm_flag_peers(ctx, rank, n_peer));
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_vol_regime_shard_scale(bmfm_ctx* ctx, int n, int K, const double* wb, const double* wv,
                                           const double* totals, int rank, const double* cov_block,
                                           const int32_t* cov_status_block, double* out_cov_block, double* lamb_block) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, n >= 0 && K > 0 && rank >= 0, "bad sizes");
  if (n == 0) return BMFM_OK;
  BMFM_ARG(ctx, wb && wv && totals && lamb_block, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (ctx->opt_vr_split) {       // the separate scan and scale launches (comparison switch)
    vr_scan_carry_kernel<<<1, 1024, 0, ctx->stream>>>(n, wb, wv, totals, rank, lamb_block, bmfm_flag_peers(ctx, rank, 0));
    ctx->launches++;
    if (out_cov_block) {
      vr_scale_kernel<<<n, 256, 0, ctx->stream>>>(n, K * K, cov_block, cov_status_block, lamb_block, out_cov_block);
      ctx->launches++;
    }
  } else {
    const int D = vr_fused_dates(ctx, n);
    vr_scan_scale_kernel<<<(n + D - 1) / D, 256, 0, ctx->stream>>>(0, n, K * K, D, wb, wv, totals, rank, bmfm_flag_peers(ctx, rank, 0),
                                                                  cov_block, cov_status_block, lamb_block, out_cov_block);
    ctx->launches++;
  }
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
