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
constexpr int BS_BIN_BITS = 11, BS_BINS = 1 << BS_BIN_BITS;   // histogram of the multi-select
constexpr int BS_COLL_MAX = 2048;                              // members of needed bins ranked by counting
constexpr int BS_MAXR = 16;                                    // rows per thread: 16384 rows per date at 1024 threads
constexpr int BS_GFAST = 10;                                   // groups handled by register accumulators (deciles)

// order-preserving map double -> uint64
__device__ __forceinline__ unsigned long long bs_key(double x) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(x);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

// sum over the warp of 16 values per lane by recursive halving: lanes 2 e and 2 e + 1 both return the total of
// value e (15 + 1 shuffles instead of 16 x 5)
__device__ __forceinline__ double bs_reduce_scatter16(double (&a)[16], int lane) {
#pragma unroll
  for (int s2 = 0; s2 < 4; ++s2) {
    const int off = 16 >> s2, H = 8 >> s2;
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const double lo = a[i], hi = a[i + H];
      const double send = upper ? lo : hi, keep = upper ? hi : lo;
      a[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return a[0] + __shfl_xor_sync(0xffffffffu, a[0], 1);
}

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
  // per-warp private copies of the group sums: the ~10 groups are hit by every row, one shared table
  // would serialise thousands of shared-memory atomics on a handful of addresses
  double* s_wsum = reinterpret_cast<double*>(s_flag + 4);   // [nwarps][ngroup][4]
  const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  double* my = s_wsum + (size_t)warp * p.ngroup * 4;
  // multi-select scratch
  double* s_rval = s_wsum + (size_t)nwarps * p.ngroup * 4;            // [2 (G+1)] values at the wanted ranks
  unsigned long long* s_k = reinterpret_cast<unsigned long long*>(s_rval + 2 * BS_MAX_GROUPS + 2);   // key min / max
  int* s_hist = reinterpret_cast<int*>(s_k + 2);                      // [BS_BINS + 1] counts, then exclusive prefix
  int* s_need = s_hist + BS_BINS + 1;                                 // [BS_BINS] slot of a needed bin, -1 otherwise
  int* s_rank = s_need + BS_BINS;                                     // [2 (G+1)] wanted ranks
  int* s_rbin = s_rank + 2 * BS_MAX_GROUPS + 2;
  int* s_slotbin = s_rbin + 2 * BS_MAX_GROUPS + 2;
  int* s_slotoff = s_slotbin + 2 * BS_MAX_GROUPS + 2;                 // [+1]
  int* s_fill = s_slotoff + 2 * BS_MAX_GROUPS + 3;
  int* s_ctl = s_fill + 2 * BS_MAX_GROUPS + 2;                        // nslots, ncoll, spare
  const int tid = threadIdx.x, nt = blockDim.x, G = p.ngroup, lane = threadIdx.x & 31;
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    for (int i = tid; i < 4 * G; i += nt) s_gsum[i] = 0.0;
    for (int i = tid; i < nwarps * 4 * G; i += nt) s_wsum[i] = 0.0;
    if (tid == 0) { *s_flag = 0; s_ctl[0] = 0; s_ctl[1] = 0; s_ctl[2] = 0; s_k[0] = ~0ull; s_k[1] = 0ull; }
    for (int i = tid; i < BS_BINS + 1; i += nt) { s_hist[i] = 0; if (i < BS_BINS) s_need[i] = -1; }
    // the order statistics the quantile edges need: ranks ip / in of every edge (numpy 'linear')
    if (tid <= G && n > 0) {
      const double virt = __dmul_rn((double)(n - 1), p.quant[tid]);
      int ip = (int)floor(virt), in = ip + 1;
      if (virt >= (double)(n - 1)) ip = in = n - 1;
      if (virt < 0.0) ip = in = 0;
      s_rank[2 * tid] = ip;
      s_rank[2 * tid + 1] = in;
    }
    // ---- multi-select instead of a full sort: 2048-bin histogram on the leading bits in which the
    // ---- keys differ, then only the few bins that hold a wanted rank are collected and ranked
    unsigned long long kmin = ~0ull, kmax = 0ull;
    for (int i = tid; i < n; i += nt) {
      const unsigned long long u = bs_key(p.cap[a + i]);
      kmin = u < kmin ? u : kmin;
      kmax = u > kmax ? u : kmax;
    }
    for (int off = 16; off > 0; off >>= 1) {
      const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, kmin, off), o2 = __shfl_xor_sync(0xffffffffu, kmax, off);
      kmin = o1 < kmin ? o1 : kmin;
      kmax = o2 > kmax ? o2 : kmax;
    }
    __syncthreads();
    if ((tid & 31) == 0) { atomicMin(&s_k[0], kmin); atomicMax(&s_k[1], kmax); }
    __syncthreads();
    kmin = s_k[0]; kmax = s_k[1];
    const unsigned long long diff = kmin ^ kmax;
    const int hb = diff ? 63 - __clzll((long long)diff) : 0;
    const int shift = hb + 1 > BS_BIN_BITS ? hb + 1 - BS_BIN_BITS : 0;
    const unsigned long long base = kmin >> shift;
    for (int i = tid; i < n; i += nt) atomicAdd(&s_hist[(int)((bs_key(p.cap[a + i]) >> shift) - base)], 1);
    __syncthreads();
    // exclusive prefix over the bins (warp 0: 64 bins per lane, then a shuffle scan)
    if (tid < 32) {
      constexpr int PER = BS_BINS / 32;
      int loc = 0;
      for (int j = 0; j < PER; ++j) loc += s_hist[tid * PER + j];
      int inc = loc;
      for (int off = 1; off < 32; off <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, inc, off);
        if (tid >= off) inc += o;
      }
      int run = inc - loc;
      for (int j = 0; j < PER; ++j) {
        const int h = s_hist[tid * PER + j];
        s_hist[tid * PER + j] = run;
        run += h;
      }
      if (tid == 31) s_hist[BS_BINS] = run;
    }
    __syncthreads();
    // bins that contain a wanted rank get a slot; their sizes decide whether the short path applies
    if (tid < 2 * (G + 1) && n > 0) {
      const int r = s_rank[tid];
      int lo = 0, hi = BS_BINS;             // largest b with prefix[b] <= r
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_hist[mid] <= r) lo = mid; else hi = mid;
      }
      s_rbin[tid] = lo;
      if (atomicCAS(&s_need[lo], -1, -2) == -1) {
        const int slot = atomicAdd(&s_ctl[0], 1);
        s_need[lo] = slot;
        s_slotbin[slot] = lo;
      }
    }
    __syncthreads();
    const int nslots = s_ctl[0];
    if (tid == 0) {
      int tot = 0;
      for (int sl = 0; sl < nslots; ++sl) {
        const int bn = s_slotbin[sl];
        s_slotoff[sl] = tot;
        tot += s_hist[bn + 1] - s_hist[bn];
      }
      s_slotoff[nslots] = tot;
      s_ctl[1] = tot;
    }
    for (int i = tid; i < 2 * BS_MAX_GROUPS + 2; i += nt) s_fill[i] = 0;
    __syncthreads();
    const int ncoll = s_ctl[1];
    if (ncoll <= BS_COLL_MAX && ncoll <= p.npad) {
      // collect the members of the needed bins into s_sort (grouped by slot), rank them in place
      for (int i = tid; i < n; i += nt) {
        const double c = p.cap[a + i];
        const int bn = (int)((bs_key(c) >> shift) - base);
        const int sl = s_need[bn];
        if (sl >= 0) s_sort[s_slotoff[sl] + atomicAdd(&s_fill[sl], 1)] = c;
      }
      __syncthreads();
      for (int e = tid; e < ncoll; e += nt) {
        int sl = 0;
        while (s_slotoff[sl + 1] <= e) ++sl;
        const int lo = s_slotoff[sl], hi = s_slotoff[sl + 1];
        const double x = s_sort[e];
        int rk = 0;
        for (int i = lo; i < hi; ++i) {
          const double y = s_sort[i];
          rk += (y < x) || (y == x && i < e);
        }
        const int grank = s_hist[s_slotbin[sl]] + rk;      // position in the fully sorted order
        for (int q2 = 0; q2 < 2 * (G + 1); ++q2)
          if (s_rank[q2] == grank) s_rval[q2] = x;
      }
      __syncthreads();
    } else {
      // degenerate key distribution: full bitonic sort (ascending) of the date's caps
      for (int i = tid; i < p.npad; i += nt) s_sort[i] = i < n ? p.cap[a + i] : inf;
      __syncthreads();
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
      if (tid < 2 * (G + 1) && n > 0) s_rval[tid] = s_sort[s_rank[tid]];
      __syncthreads();
    }
    // quantile edges (numpy 'linear' method, _lerp with the t >= 0.5 branch)
    if (tid <= G && n > 0) {
      const double virt = __dmul_rn((double)(n - 1), p.quant[tid]);
      const double prev = floor(virt);
      const double gamma = __dsub_rn(virt, prev);
      const double lo = s_rval[2 * tid], hi = s_rval[2 * tid + 1];
      const double diffv = __dsub_rn(hi, lo);
      double e = __dadd_rn(lo, __dmul_rn(diffv, gamma));
      if (gamma >= 0.5) e = __dsub_rn(hi, __dmul_rn(diffv, __dsub_rn(1.0, gamma)));
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
    // The decile code of every row is found once and kept in registers (a thread owns rows tid, tid + nt, ...: at
    // most BS_MAXR of them).
    const double e0 = s_edges[0];
    signed char gcode[BS_MAXR];
#pragma unroll
    for (int j = 0; j < BS_MAXR; ++j) {
      const int i = tid + j * nt;
      int g = -1;
      if (i < n) {
        const double c = p.cap[a + i];
        int cnt = 0;
        for (int k = 0; k <= G; ++k) cnt += s_edges[k] < c;   // searchsorted(edges, c, 'left')
        g = cnt - 1;
        if (c == e0) g = 0;
        if (g >= G) g = -1;
      }
      gcode[j] = (signed char)g;
    }
    if (G <= BS_GFAST && G * nt <= p.npad) {
      // group sums without atomics and without a predicated add per (row, group): every thread owns column tid of a
      // [G][nt] table (the sort buffer, free by now; nt is a multiple of 32, so a warp's accesses never share a bank
      // whatever the groups) and adds each of its rows into the row of its group — three instructions per row where
      // one register accumulator per group costs two per (row, group).  The G column sums then go through the same
      // recursive-halving reduction inside the warp and one store per warp.
      double* col = s_sort + tid;
#pragma unroll 1
      for (int which = 0; which < 3; ++which) {
        for (int k = 0; k < G; ++k) col[k * nt] = 0.0;
#pragma unroll
        for (int j = 0; j < BS_MAXR; ++j) {
          const int i = tid + j * nt;
          const int g = gcode[j];
          if (i < n && g >= 0) {
            const double c = p.cap[a + i];
            const double x = which == 0 ? p.vol[a + i] * c : (which == 1 ? c : 1.0);
            col[g * nt] += x;
          }
        }
        double acc[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) acc[e] = e < G ? col[e * nt] : 0.0;
        const double tot = bs_reduce_scatter16(acc, lane);
        if ((lane & 1) == 0 && (lane >> 1) < G) my[4 * (lane >> 1) + which] = tot;
      }
      __syncthreads();
      for (int e = tid; e < 4 * G; e += nt) {
        if ((e & 3) == 3) continue;
        double s2 = 0.0;
        for (int w = 0; w < nwarps; ++w) s2 += s_wsum[(size_t)w * G * 4 + e];
        s_gsum[e] = s2;
      }
      __syncthreads();
      // dispersion about the group mean
      {
        for (int k = 0; k < G; ++k) col[k * nt] = 0.0;
#pragma unroll
        for (int j = 0; j < BS_MAXR; ++j) {
          const int i = tid + j * nt;
          const int g = gcode[j];
          if (i < n && g >= 0) {
            const double m = s_gsum[4 * g + 0] / s_gsum[4 * g + 1];
            const double d = p.vol[a + i] - m;
            col[g * nt] += d * d;
          }
        }
        double acc[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) acc[e] = e < G ? col[e * nt] : 0.0;
        const double tot = bs_reduce_scatter16(acc, lane);
        if ((lane & 1) == 0 && (lane >> 1) < G) my[4 * (lane >> 1) + 3] = tot;
      }
    } else if (G <= BS_GFAST) {
      // (short dates, where the sort buffer is smaller than the [G][nt] table) one register accumulator per group
      // and predicated adds, one sweep over the thread's rows per kind of sum
#pragma unroll 1
      for (int which = 0; which < 3; ++which) {
        double acc[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) acc[e] = 0.0;
#pragma unroll
        for (int j = 0; j < BS_MAXR; ++j) {
          const int i = tid + j * nt;
          if (i < n) {
            const double c = p.cap[a + i];
            const double x = which == 0 ? p.vol[a + i] * c : (which == 1 ? c : 1.0);
            const int g = gcode[j];
#pragma unroll
            for (int k = 0; k < BS_GFAST; ++k) acc[k] += g == k ? x : 0.0;
          }
        }
        const double tot = bs_reduce_scatter16(acc, lane);
        if ((lane & 1) == 0 && (lane >> 1) < G) my[4 * (lane >> 1) + which] = tot;
      }
      __syncthreads();
      for (int e = tid; e < 4 * G; e += nt) {
        if ((e & 3) == 3) continue;
        double s2 = 0.0;
        for (int w = 0; w < nwarps; ++w) s2 += s_wsum[(size_t)w * G * 4 + e];
        s_gsum[e] = s2;
      }
      __syncthreads();
      // dispersion about the group mean
      {
        double acc[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) acc[e] = 0.0;
#pragma unroll
        for (int j = 0; j < BS_MAXR; ++j) {
          const int i = tid + j * nt;
          const int g = gcode[j];
          if (i < n && g >= 0) {
            const double m = s_gsum[4 * g + 0] / s_gsum[4 * g + 1];
            const double d = p.vol[a + i] - m;
            const double dd = d * d;
#pragma unroll
            for (int k = 0; k < BS_GFAST; ++k) acc[k] += g == k ? dd : 0.0;
          }
        }
        const double tot = bs_reduce_scatter16(acc, lane);
        if ((lane & 1) == 0 && (lane >> 1) < G) my[4 * (lane >> 1) + 3] = tot;
      }
    } else {
      // many groups: per-warp shared tables and atomics
      for (int j = 0; j < BS_MAXR; ++j) {
        const int i = tid + j * nt;
        const int g = gcode[j];
        if (i < n && g >= 0) {
          const double c = p.cap[a + i], v = p.vol[a + i];
          atomicAdd(&my[4 * g + 0], v * c);
          atomicAdd(&my[4 * g + 1], c);
          atomicAdd(&my[4 * g + 2], 1.0);
        }
      }
      __syncthreads();
      for (int e = tid; e < 4 * G; e += nt) {
        if ((e & 3) == 3) continue;
        double s2 = 0.0;
        for (int w = 0; w < nwarps; ++w) s2 += s_wsum[(size_t)w * G * 4 + e];
        s_gsum[e] = s2;
      }
      __syncthreads();
      for (int j = 0; j < BS_MAXR; ++j) {
        const int i = tid + j * nt;
        const int g = gcode[j];
        if (i < n && g >= 0) {
          const double m = s_gsum[4 * g + 0] / s_gsum[4 * g + 1];
          const double d = p.vol[a + i] - m;
          atomicAdd(&my[4 * g + 3], d * d);
        }
      }
    }
    __syncthreads();
    for (int g = tid; g < G; g += nt) {
      double s2 = 0.0;
      for (int w = 0; w < nwarps; ++w) s2 += s_wsum[(size_t)w * G * 4 + 4 * g + 3];
      s_gsum[4 * g + 3] = s2;
    }
    __syncthreads();
    // per group: mean and std once, then the shrink of every row (utils.py:146-149)
    if (tid < G) {
      s_gsum[4 * tid + 0] = s_gsum[4 * tid + 0] / s_gsum[4 * tid + 1];
      s_gsum[4 * tid + 3] = sqrt(s_gsum[4 * tid + 3] / s_gsum[4 * tid + 2]);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < BS_MAXR; ++j) {
      const int i = tid + j * nt;
      if (i < n) {
        const int g = gcode[j];
        double o = __longlong_as_double(0x7ff8000000000000LL);
        if (g >= 0) {
          const double v = p.vol[a + i];
          const double m = s_gsum[4 * g + 0], sd = s_gsum[4 * g + 3];
          const double aa = p.q * fabs(v - m);
          const double w = aa / (aa + sd);
          o = w * m + (1.0 - w) * fabs(v);
        }
        p.out[a + i] = o;
      }
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
  // (two CTAs of 512 threads per SM overlap one date's barrier-separated phases with another's; a thread owns at most
  //  BS_MAXR rows)
  int nthreads = npad > 8192 ? 1024 : (npad >= 2048 ? 512 : (npad >= 512 ? 256 : 128));
  const size_t smem = ((size_t)npad + BS_MAX_GROUPS + 1 + 4 * BS_MAX_GROUPS) * 8 + 16 + (size_t)(nthreads / 32) * ngroup * 4 * 8 +
                      (size_t)(2 * BS_MAX_GROUPS + 2 + 2) * 8 + (size_t)(2 * BS_BINS + 1 + 6 * (2 * BS_MAX_GROUPS + 3) + 8) * 4;
  if (smem > (size_t)ctx->max_smem_optin) {
    snprintf(ctx->err, sizeof(ctx->err), "bayes_shrink: %d rows per date exceed the shared-memory sort (max 16384)",
             max_rows_per_date);
    return BMFM_E_UNSUPPORTED;
  }
  BMFM_CUDA(ctx, cudaFuncSetAttribute(bayes_shrink_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
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
