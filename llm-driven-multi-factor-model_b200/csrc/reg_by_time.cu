// Per-date constrained cross-sectional regression (part 1 of the Barra path).
//
// Replaces, for every date of the panel at once:
//   MFM.reg_by_time                      Barra-master/mfm/MFM.py:48-76
//   CrossSection.__init__ / .reg         Barra-master/mfm/CrossSection.py:35-108
//   style_factor_norm                    Barra-master/mfm/CrossSection.py:12-20
//
// Design (B200).  The default kernel is reg_mma_kernel (reg_mma.cuh): 1, 2 or 4 warps own a date end to end,
// rows stream through per-warp cp.async rings of 32-row chunks, the weighted Gram and the per-industry sums are
// formed by FP64 tensor-core MMA, HBM sees each input byte once (the second pass for the residuals is served
// mostly by L2).  Common to all kernels of this file:
//   sums     structured weighted Gram: style x style (packed), style x ret, cap-weighted style sums, the
//            moments of the scalar np.std, and per-industry sums (weights, w*ret, cap, w*style) — the one-hot
//            block never becomes a dense matrix; any row order inside a date is correct;
//   solve    the industry block of the normal equations is diagonal, so the K-1 = 41 unknowns reduce by block
//            elimination to a Q x Q SPD system (Gauss-Jordan in the registers of one warp) plus a 2 x 2 for
//            (country, multiplier).  Dates whose reduced system is not safely definite are flagged and redone
//            by the dense kernel below, which forms the reference's R^T G R explicitly and applies numpy's
//            pinv semantics (eigen-decomposition, cutoff 1e-15 * max) — CrossSection.py:69-76;
//   pass C   specific returns + R2.
// Two earlier designs are kept behind BMFM_REG_MODE for comparison and as parity cross-checks: reg_fast_kernel
// (1: one CTA per date with the date resident in shared memory, tiles if larger) and reg_stream_kernel (2: one
// warp per date with 97 register accumulators per lane).
#include <stdlib.h>
#include "bmfm_common.cuh"
#include "jacobi.cuh"

namespace bmfm {

constexpr int REG_MAX_Q = 12;       // register-accumulator template range
constexpr int REG_MAX_P = 63;       // P + 1 buckets <= 64
constexpr int REG_NEED_DENSE = 100; // internal status: redo this date in the dense kernel
constexpr int REG_MAX_PEERS = 8;    // GPUs of one NVSwitch box

struct RegParams {
  int T, P, Q, K;
  const long long* date_off;
  const double* cap;
  const double* ret;
  const int* ind;
  const double* style;
  double* factor_ret;
  double* spec;
  double* r2;
  double* exposure;
  int* status;
  int tile_rows;
  int all_dense;  // dense kernel: process every date (exposure requested)
  // fused all-gather: every date's factor returns are also stored straight into each rank's gathered
  // series (peer memory over NVLink), at row peer_row0 + t; n_peer = 0 disables it
  int n_peer;
  long long peer_row0;
  double* peer_fr[REG_MAX_PEERS];
  int smma;       // per-industry sums by MMA against indicator columns (else shared-memory column sums)
  int* next_date; // device counter (zeroed before the launch) that hands out the dates after each warp's first, or NULL
};

__device__ __forceinline__ void store_factor_ret(const RegParams& p, int t, int k, double v) {
  p.factor_ret[(size_t)t * p.K + k] = v;
  for (int r = 0; r < p.n_peer; ++r) p.peer_fr[r][(size_t)(p.peer_row0 + t) * p.K + k] = v;
}

__host__ __device__ constexpr int round_up32(int x) { return (x + 31) / 32 * 32; }
__host__ __device__ constexpr int na1_of(int Q) { return Q * (Q + 1) / 2 + Q; }
__host__ __device__ constexpr int na1pad_of(int Q) { return na1_of(Q) == 0 ? 32 : round_up32(na1_of(Q)); }
__host__ __device__ constexpr int na2_of(int Q) { return 3 * Q + 2; }
__host__ __device__ constexpr int na2pad_of(int Q) { return round_up32(na2_of(Q)); }

// byte offsets into dynamic shared memory
struct RegSmem {
  int style, cap, ret, w, ind, bucket, seg, acc, red, coef, fvec, solve, G, V, G0, cs, total;
};

__host__ __device__ inline int align16(int x) { return (x + 15) & ~15; }

__host__ __device__ inline RegSmem reg_smem_layout(int tile_rows, int P, int Q, int nwarps, bool dense) {
  RegSmem s;
  const int NB = P + 1, BV = 3 + Q, K = 1 + P + Q;
  const int ld = K | 1;
  int o = 0;
  s.style = o; o = align16(o + tile_rows * Q * 8);
  s.cap = o;   o = align16(o + tile_rows * 8);
  s.ret = o;   o = align16(o + tile_rows * 8);
  s.w = o;     o = align16(o + tile_rows * 8);
  s.ind = o;   o = align16(o + tile_rows * 4);
  s.bucket = o; o = align16(o + NB * BV * 8);
  s.seg = o;   o = align16(o + (3 * NB + 8) * 4);
  s.acc = o;   o = align16(o + (na1pad_of(Q) + na2pad_of(Q)) * 8);
  const int redw = na1pad_of(Q) > na2pad_of(Q) ? na1pad_of(Q) : na2pad_of(Q);
  s.red = o;   o = align16(o + nwarps * redw * 8);
  s.coef = o;  o = align16(o + (3 * Q + NB + 32) * 8);   // piv[Q] mp[Q] beta[Q] g[NB] misc[32]
  s.fvec = o;  o = align16(o + (K + 2) * 8);
  s.solve = o; o = align16(o + (Q * Q + 7 * Q + 2 * NB + 8) * 8);
  s.G = s.V = s.G0 = s.cs = o;
  if (dense) {
    s.G = o;  o = align16(o + K * ld * 8);
    s.V = o;  o = align16(o + K * ld * 8);
    s.G0 = o; o = align16(o + K * ld * 8);
    s.cs = o; o = align16(o + (4 * K + 16) * 8);
  }
  s.total = o;
  return s;
}

__device__ __forceinline__ int bucket_of(int ind, int P) { return (ind >= 0 && ind < P) ? ind : P; }

// ---- tile load: global -> shared, every byte once ---------------------------------------
template <int Q>
__device__ __forceinline__ void load_tile(const RegParams& p, long long a, int n, bool with_cap,
                                          double* s_style, double* s_cap, double* s_ret, int* s_ind) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int e = tid; e < n; e += nt) {
    if (with_cap) cp_async_8(&s_cap[e], p.cap + a + e);
    cp_async_8(&s_ret[e], p.ret + a + e);
    cp_async_4(&s_ind[e], p.ind + a + e);
  }
  if (Q > 0) {
    const double* src = p.style + a * Q;
    const int tot = n * Q;
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
      const int n16 = tot >> 1;
      for (int e = tid; e < n16; e += nt) cp_async_16(&s_style[2 * e], src + 2 * e);
      if ((tot & 1) && tid == 0) cp_async_8(&s_style[tot - 1], src + tot - 1);
    } else {
      for (int e = tid; e < tot; e += nt) cp_async_8(&s_style[e], src + e);
    }
  }
}

template <int Q>
__device__ __forceinline__ void load_row_styles(const double* s_style, int i, const double* piv, double (&d)[Q > 0 ? Q : 1]) {
  if (Q == 0) return;
  const double* row = s_style + (size_t)i * Q;
  if (Q % 2 == 0) {
#pragma unroll
    for (int q = 0; q < Q; q += 2) {
      const double2 v = *reinterpret_cast<const double2*>(row + q);
      d[q] = v.x - piv[q];
      d[q + 1] = v.y - piv[q + 1];
    }
  } else {
#pragma unroll
    for (int q = 0; q < Q; ++q) d[q] = row[q] - piv[q];
  }
}

// ---- accumulation over one date (all tiles) ----------------------------------------------
// On return (after the trailing __syncthreads):
//   s_acc[0 .. NA1PAD)            S2 packed (q(q+1)/2 + r, r <= q) = sum w d_q d_r,  then BY[q] = sum w d_q y
//   s_acc[NA1PAD ..)              CD[q] = sum cap d_q, D1[q] = sum d_q, D2[q] = sum d_q^2, Y1, Y2 (shifted ret moments)
//   s_bucket[b*(3+Q) + v]         per bucket b: sum w, sum w y, sum cap, sum w d_q
// with d_q = style_q - piv_q (piv = the date's first row: a cheap shift that removes the
// cancellation a raw-moment formulation would have) and the LAST tile resident in shared memory.
template <int Q>
__device__ void accumulate_date(const RegParams& p, const RegSmem& L, unsigned char* smem, long long a, int n,
                                int& ntiles_out) {
  constexpr int NA1 = na1_of(Q), NA1PAD = na1pad_of(Q), NA2 = na2_of(Q), NA2PAD = na2pad_of(Q);
  constexpr int QQ = Q > 0 ? Q : 1;
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarps = nt >> 5;
  const int P = p.P, NB = P + 1, BV = 3 + Q;
  double* s_style = reinterpret_cast<double*>(smem + L.style);
  double* s_cap = reinterpret_cast<double*>(smem + L.cap);
  double* s_ret = reinterpret_cast<double*>(smem + L.ret);
  double* s_w = reinterpret_cast<double*>(smem + L.w);
  int* s_ind = reinterpret_cast<int*>(smem + L.ind);
  double* s_bucket = reinterpret_cast<double*>(smem + L.bucket);
  int* s_seg = reinterpret_cast<int*>(smem + L.seg);   // start[NB] end[NB] runs[NB] flag
  double* s_acc = reinterpret_cast<double*>(smem + L.acc);
  double* s_red = reinterpret_cast<double*>(smem + L.red);
  double* s_piv = reinterpret_cast<double*>(smem + L.coef);   // piv[Q], then ypiv at misc
  double* s_misc = s_piv + 3 * Q + NB;

  const int tile_rows = p.tile_rows;
  const int ntiles = (n + tile_rows - 1) / tile_rows;
  ntiles_out = ntiles;

  for (int e = tid; e < NB * BV; e += nt) s_bucket[e] = 0.0;
  if (tid < Q) s_piv[tid] = p.style[a * Q + tid];
  if (tid == 0) s_misc[0] = p.ret[a];   // ypiv
  __syncthreads();
  const double ypiv = s_misc[0];

  double acc1[NA1PAD];
  double acc2[NA2PAD];
#pragma unroll
  for (int i = 0; i < NA1PAD; ++i) acc1[i] = 0.0;
#pragma unroll
  for (int i = 0; i < NA2PAD; ++i) acc2[i] = 0.0;

  for (int tile = 0; tile < ntiles; ++tile) {
    const int r0 = tile * tile_rows;
    const int m = min(tile_rows, n - r0);
    if (tile > 0) __syncthreads();   // previous tile fully consumed
    load_tile<Q>(p, a + r0, m, true, s_style, s_cap, s_ret, s_ind);
    for (int e = tid; e < 3 * NB + 1; e += nt) s_seg[e] = 0;
    cp_async_wait_all();
    __syncthreads();

    // pass A1: weighted style Gram + style x ret; also stores w = sqrt(cap)
    for (int i = tid; i < m; i += nt) {
      const double cap = s_cap[i];
      const double w = sqrt(cap);
      s_w[i] = w;
      if (Q > 0) {
        const double y = s_ret[i];
        double d[QQ];
        load_row_styles<Q>(s_style, i, s_piv, d);
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          const double wd = w * d[q];
#pragma unroll
          for (int r = 0; r <= q; ++r) acc1[q * (q + 1) / 2 + r] = fma(wd, d[r], acc1[q * (q + 1) / 2 + r]);
          acc1[Q * (Q + 1) / 2 + q] = fma(wd, y, acc1[Q * (Q + 1) / 2 + q]);
        }
      }
    }
    // pass A2: cap-weighted style sums (means of style_factor_norm), moments of the scalar std, ret moments
    for (int i = tid; i < m; i += nt) {
      const double cap = s_cap[i];
      const double u = s_ret[i] - ypiv;
      acc2[3 * Q] += u;
      acc2[3 * Q + 1] = fma(u, u, acc2[3 * Q + 1]);
      if (Q > 0) {
        double d[QQ];
        load_row_styles<Q>(s_style, i, s_piv, d);
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          acc2[q] = fma(cap, d[q], acc2[q]);
          acc2[Q + q] += d[q];
          acc2[2 * Q] = fma(d[q], d[q], acc2[2 * Q]);   // only the total over q enters the scalar std
        }
      }
    }
    // industry runs inside the tile
    int* seg_start = s_seg;
    int* seg_end = s_seg + NB;
    int* seg_runs = s_seg + 2 * NB;
    int* seg_flag = s_seg + 3 * NB;
    for (int i = tid; i < m; i += nt) {
      const int b = bucket_of(s_ind[i], P);
      const int bp = i > 0 ? bucket_of(s_ind[i - 1], P) : -1;
      const int bn = i + 1 < m ? bucket_of(s_ind[i + 1], P) : -1;
      if (b != bp) {
        if (atomicAdd(&seg_runs[b], 1) > 0) *seg_flag = 1;   // a second run: rows not grouped
        seg_start[b] = i;
      }
      if (b != bn) seg_end[b] = i + 1;
    }
    __syncthreads();   // s_w, runs visible
    if (*seg_flag == 0) {
      // grouped rows: a half-warp per bucket, lane = value column
      const int v = tid & 15, grp = tid >> 4, ngrp = nt >> 4;
      if (v < BV) {
        for (int b = grp; b < NB; b += ngrp) {
          if (seg_runs[b] == 0) continue;
          const int lo = seg_start[b], hi = seg_end[b];
          double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
          int i = lo;
          if (v == 0) {
            for (; i + 3 < hi; i += 4) { s0 += s_w[i]; s1 += s_w[i + 1]; s2 += s_w[i + 2]; s3 += s_w[i + 3]; }
            for (; i < hi; ++i) s0 += s_w[i];
          } else if (v == 1) {
            for (; i + 3 < hi; i += 4) {
              s0 = fma(s_w[i], s_ret[i], s0); s1 = fma(s_w[i + 1], s_ret[i + 1], s1);
              s2 = fma(s_w[i + 2], s_ret[i + 2], s2); s3 = fma(s_w[i + 3], s_ret[i + 3], s3);
            }
            for (; i < hi; ++i) s0 = fma(s_w[i], s_ret[i], s0);
          } else if (v == 2) {
            for (; i + 3 < hi; i += 4) { s0 += s_cap[i]; s1 += s_cap[i + 1]; s2 += s_cap[i + 2]; s3 += s_cap[i + 3]; }
            for (; i < hi; ++i) s0 += s_cap[i];
          } else {
            const int q = v - 3;
            const double pv = s_piv[q];
            const double* col = s_style + q;
            for (; i + 3 < hi; i += 4) {
              s0 = fma(s_w[i], col[(size_t)i * Q] - pv, s0);
              s1 = fma(s_w[i + 1], col[(size_t)(i + 1) * Q] - pv, s1);
              s2 = fma(s_w[i + 2], col[(size_t)(i + 2) * Q] - pv, s2);
              s3 = fma(s_w[i + 3], col[(size_t)(i + 3) * Q] - pv, s3);
            }
            for (; i < hi; ++i) s0 = fma(s_w[i], col[(size_t)i * Q] - pv, s0);
          }
          const double sum = (s0 + s1) + (s2 + s3);
          s_bucket[b * BV + v] += sum;
        }
      }
    } else {
      // arbitrary row order: shared-memory atomics (correct, slower)
      for (int i = tid; i < m; i += nt) {
        const int b = bucket_of(s_ind[i], P);
        const double w = s_w[i];
        atomicAdd(&s_bucket[b * BV + 0], w);
        atomicAdd(&s_bucket[b * BV + 1], w * s_ret[i]);
        atomicAdd(&s_bucket[b * BV + 2], s_cap[i]);
        for (int q = 0; q < Q; ++q)
          atomicAdd(&s_bucket[b * BV + 3 + q], w * (s_style[(size_t)i * Q + q] - s_piv[q]));
      }
    }
  }

  // CTA reduction of the register accumulators: warp recursive-halving, then across warps
  warp_reduce_scatter<NA1PAD>(acc1, lane);
#pragma unroll
  for (int i = 0; i < NA1PAD / 32; ++i) s_red[warp * NA1PAD + reduce_scatter_index<NA1PAD>(i, lane)] = acc1[i];
  __syncthreads();
  for (int j = tid; j < NA1PAD; j += nt) {
    double s = 0.0;
    for (int w2 = 0; w2 < nwarps; ++w2) s += s_red[w2 * NA1PAD + j];
    s_acc[j] = s;
  }
  __syncthreads();
  warp_reduce_scatter<NA2PAD>(acc2, lane);
#pragma unroll
  for (int i = 0; i < NA2PAD / 32; ++i) s_red[warp * NA2PAD + reduce_scatter_index<NA2PAD>(i, lane)] = acc2[i];
  __syncthreads();
  for (int j = tid; j < NA2PAD; j += nt) {
    double s = 0.0;
    for (int w2 = 0; w2 < nwarps; ++w2) s += s_red[w2 * NA2PAD + j];
    s_acc[NA1PAD + j] = s;
  }
  __syncthreads();
  (void)NA1; (void)NA2;
}

// ---- fast solve: block elimination of the diagonal industry block -------------------------
// Minimise sum_i w_i (y_i - f_c - g_{p(i)} - beta . x_i)^2  s.t.  sum_p c_p g_p = 0,
// x_i = d_i - m' (the sigma of style_factor_norm only rescales beta).  Eliminating g_p through
// the diagonal block leaves  [w0 gs0' C; gs0 St t; C t' -D] (f_c, beta, mu') = (b0, bt, cy).
// Equivalent to R pinv(R'GR) R'b of CrossSection.py:69-76 whenever R'GR is non-singular;
// empty non-last industries (zero row/column of R'GR, pinv -> 0) are skipped the same way.
// The reduced matrices are formed by all threads; the Q x (Q+3) Gauss-Jordan elimination runs
// in the registers of warp 0 (lane = row, pivot rows broadcast by shuffle, no pivoting: St is
// SPD).  Returns 0, REG_NEED_DENSE if a pivot is not safely positive, or BMFM_ST_NONFINITE.
template <bool WARP>
__device__ __forceinline__ void group_sync() {
  if (WARP) __syncwarp(); else __syncthreads();
}

// WARP = true: the calling warp owns `smem` (per-warp layout) and solves on its own;
// WARP = false: the whole CTA cooperates (barriers are __syncthreads).
template <int Q, bool WARP>
__device__ int solve_fast_cta(const RegParams& p, const RegSmem& L, unsigned char* smem, int n) {
  constexpr int QQ = Q > 0 ? Q : 1;
  const int lane = threadIdx.x & 31;
  const int tid = WARP ? lane : (int)threadIdx.x, nt = WARP ? 32 : (int)blockDim.x;
  const int P = p.P, NB = P + 1, BV = 3 + Q, K = p.K;
  constexpr int NA1PAD = na1pad_of(Q);
  const double* s_bucket = reinterpret_cast<const double*>(smem + L.bucket);
  const double* s_acc = reinterpret_cast<const double*>(smem + L.acc);
  double* s_piv = reinterpret_cast<double*>(smem + L.coef);
  double* s_mp = s_piv + Q;
  double* s_beta = s_mp + Q;
  double* s_g = s_beta + Q;
  double* s_misc = s_g + NB;
  double* s_f = reinterpret_cast<double*>(smem + L.fvec);
  double* St = reinterpret_cast<double*>(smem + L.solve);   // Q*Q (lower triangle used)
  double* rhs = St + Q * Q;                                 // gs0[Q] t[Q] bt[Q]
  double* winv = rhs + 3 * Q;                               // NB
  int* s_flag = reinterpret_cast<int*>(smem + L.seg) + 3 * NB + 2;
  const double* S2 = s_acc;
  const double* BY = s_acc + Q * (Q + 1) / 2;
  const double* CD = s_acc + NA1PAD;
  const double* D1 = CD + Q;
  const double* D2 = D1 + Q;
#define BW(b) s_bucket[(b)*BV + 0]
#define BWY(b) s_bucket[(b)*BV + 1]
#define BC(b) s_bucket[(b)*BV + 2]
#define BD(b, q) s_bucket[(b)*BV + 3 + (q)]
  // S1: reciprocal industry weights (all threads); scalar sums, style means, sigma (warp 0)
  for (int b = tid; b < P; b += nt) {
    const double wv = BW(b);
    winv[b] = wv > 0.0 ? 1.0 / wv : 0.0;
  }
  if (tid < 32) {
    double sC = 0.0, Dm = 0.0, cy = 0.0, capall = 0.0;
    for (int b = lane; b < NB; b += 32) {
      const double wv = BW(b), cb = BC(b);
      capall += cb;
      if (b < P && wv > 0.0) {
        const double wi = 1.0 / wv;
        sC += cb;
        Dm = fma(cb * cb, wi, Dm);
        cy = fma(cb * wi, BWY(b), cy);
      }
    }
    sC = warp_sum(sC); Dm = warp_sum(Dm); cy = warp_sum(cy); capall = warp_sum(capall);
    const double dn = (double)n;
    double sigma = 1.0;
    if (Q > 0) {
      if (lane < Q) s_mp[lane] = CD[lane] / capall;
      const double v = lane < Q ? D1[lane] + dn * s_piv[lane] : 0.0;
      const double mu = warp_sum(v) / (dn * Q);
      double t2 = 0.0;
      if (lane < Q) {
        const double dq = s_piv[lane] - mu;
        t2 = D2[lane] + 2.0 * dq * D1[lane] + dn * dq * dq;
      }
      sigma = sqrt(warp_sum(t2) / (dn * Q));
    }
    if (lane == 0) {
      s_misc[2] = sigma; s_misc[5] = sC; s_misc[6] = Dm; s_misc[7] = cy;
      *s_flag = (P > 0 && !(BC(P - 1) != 0.0)) ? BMFM_ST_NONFINITE : 0;   // -cap/0 in R (CrossSection.py:70)
    }
  }
  group_sync<WARP>();
  // S2: reduced matrices in d-units, one task per thread: St lower triangle, then t, bt, gs0
  {
    const double W0 = BW(P), b0 = BWY(P);
    constexpr int NTRI = Q * (Q + 1) / 2;
    for (int e = tid; e < NTRI + 2 * Q; e += nt) {
      if (e < NTRI) {
        int q = 0;
        while ((q + 1) * (q + 2) / 2 <= e) ++q;
        const int r = e - q * (q + 1) / 2;    // r <= q
        double a0 = S2[e], a1 = 0.0;
        int b = 0;
        for (; b + 1 < P; b += 2) {
          a0 = fma(-BD(b, q) * winv[b], BD(b, r), a0);
          a1 = fma(-BD(b + 1, q) * winv[b + 1], BD(b + 1, r), a1);
        }
        if (b < P) a0 = fma(-BD(b, q) * winv[b], BD(b, r), a0);
        const double mq = s_mp[q], mr = s_mp[r];
        St[q * Q + r] = a0 + a1 - mq * BD(P, r) - mr * BD(P, q) + mq * mr * W0;
      } else if (e < NTRI + Q) {
        const int q = e - NTRI;
        const double mq = s_mp[q];
        double t = 0.0;
        for (int b = 0; b < P; ++b)
          if (winv[b] > 0.0) t = fma(BD(b, q) * winv[b] - mq, BC(b), t);
        rhs[Q + q] = t;
        rhs[q] = BD(P, q) - mq * W0;
      } else {
        const int q = e - NTRI - Q;
        double bt = BY[q];
        for (int b = 0; b < P; ++b) bt = fma(-BD(b, q) * winv[b], BWY(b), bt);
        rhs[2 * Q + q] = bt - s_mp[q] * b0;
      }
    }
  }
  group_sync<WARP>();
  // S3: warp 0 — Gauss-Jordan in registers, 2x2 Schur system, coefficients
  if (tid < 32) {
    int fail = *s_flag;
    const double sigma = s_misc[2], sC = s_misc[5], Dm = s_misc[6], cy = s_misc[7];
    const double W0 = BW(P), b0 = BWY(P);
    double a[QQ + 3];
    double og0 = 0.0, ot = 0.0, d0 = 1.0;
    const int row = lane < Q ? lane : 0;
#pragma unroll
    for (int j = 0; j < Q; ++j) a[j] = (j <= row) ? St[row * Q + j] : St[j * Q + row];
    if (Q > 0) {
      a[Q] = rhs[row]; a[Q + 1] = rhs[Q + row]; a[Q + 2] = rhs[2 * Q + row];
      og0 = a[Q]; ot = a[Q + 1];
#pragma unroll
      for (int j = 0; j < Q; ++j) if (j == row) d0 = a[j];
    }
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      const double piv = __shfl_sync(0xffffffffu, a[k], k);
      const double dk0 = __shfl_sync(0xffffffffu, d0, k);
      if (!(piv > 1e-11 * dk0) || !(dk0 > 0.0)) { if (!fail) fail = REG_NEED_DENSE; }
      const double inv = 1.0 / piv;
      const double f = (lane == k) ? 0.0 : a[k] * inv;
#pragma unroll
      for (int j = k + 1; j < Q + 3; ++j) {
        const double pr = __shfl_sync(0xffffffffu, a[j], k);
        a[j] = (lane == k) ? pr * inv : fma(-f, pr, a[j]);
      }
    }
    // a[Q + r] on lane i is now x_r[i] = (St^-1 rhs_r)[i]
    double x0 = 0.0, x1 = 0.0, x2 = 0.0;
    if (Q > 0 && lane < Q) { x0 = a[Q]; x1 = a[Q + 1]; x2 = a[Q + 2]; } else { og0 = 0.0; ot = 0.0; }
    const double g0x0 = warp_sum(og0 * x0), g0x1 = warp_sum(og0 * x1), g0x2 = warp_sum(og0 * x2);
    const double tx1 = warp_sum(ot * x1), tx2 = warp_sum(ot * x2);
    const double M00 = W0 - g0x0, M01 = sC - g0x1, M11 = -Dm - tx1;
    const double r0 = b0 - g0x2, r1 = cy - tx2;
    double fc, mu;
    if (P > 0) {
      const double det = M00 * M11 - M01 * M01;
      if (!(fabs(det) > 1e-12 * (fabs(M00 * M11) + M01 * M01)) && !fail) fail = REG_NEED_DENSE;
      fc = (r0 * M11 - M01 * r1) / det;
      mu = (M00 * r1 - M01 * r0) / det;
    } else {
      if (!(M00 > 1e-12 * W0) && !fail) fail = REG_NEED_DENSE;
      fc = r0 / M00;
      mu = 0.0;
    }
    double bsum = 0.0;
    if (lane < Q) {
      const double be = x2 - x0 * fc - x1 * mu;
      s_beta[lane] = be;
      s_f[1 + P + lane] = be * sigma;
      bsum = be * s_mp[lane];
    }
    bsum = warp_sum(bsum);
    __syncwarp();
    int bad = 0;
    for (int b = lane; b < NB; b += 32) {
      double g = 0.0;
      if (b < P && winv[b] > 0.0) {
        double eb = 0.0;
#pragma unroll
        for (int q = 0; q < Q; ++q) eb = fma(BD(b, q) * winv[b] - s_mp[q], s_beta[q], eb);
        g = BWY(b) * winv[b] - eb + mu * BC(b) * winv[b] - fc;
      }
      s_g[b] = g;
      if (b < P) s_f[1 + b] = g;
      bad |= !isfinite(g);
    }
    if (lane < Q) bad |= !isfinite(s_beta[lane] * sigma);
    bad |= !isfinite(fc);
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) {
      s_f[0] = fc;
      s_misc[1] = fc - bsum;   // c0: intercept on the pivot-shifted styles
      if (bad && fail != REG_NEED_DENSE) fail = BMFM_ST_NONFINITE;
      *s_flag = fail;
    }
  }
#undef BW
#undef BWY
#undef BC
#undef BD
  group_sync<WARP>();
  (void)K;
  return *s_flag;
}

// ---- pass C: specific returns and R2 from coefficients in shared memory -------------------
// spec_i = y_i - c0 - g[bucket_i] - sum_q beta_q d_iq   (CrossSection.py:105),
// R2 = 1 - var(spec)/var(ret), np.var ddof 0 (CrossSection.py:106).
template <int Q>
__device__ void residual_date(const RegParams& p, const RegSmem& L, unsigned char* smem, long long a, int n,
                              int ntiles, int t) {
  constexpr int QQ = Q > 0 ? Q : 1;
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarps = nt >> 5;
  const int P = p.P, NB = P + 1;
  constexpr int NA1PAD = na1pad_of(Q);
  double* s_style = reinterpret_cast<double*>(smem + L.style);
  double* s_cap = reinterpret_cast<double*>(smem + L.cap);
  double* s_ret = reinterpret_cast<double*>(smem + L.ret);
  int* s_ind = reinterpret_cast<int*>(smem + L.ind);
  double* s_red = reinterpret_cast<double*>(smem + L.red);
  const double* s_acc = reinterpret_cast<const double*>(smem + L.acc);
  double* s_piv = reinterpret_cast<double*>(smem + L.coef);
  double* s_beta = s_piv + 2 * Q;
  double* s_g = s_beta + Q;
  double* s_misc = s_g + NB;
  const double c0 = s_misc[1];
  double se = 0.0, see = 0.0;
  for (int tile = ntiles - 1; tile >= 0; --tile) {
    const int r0 = tile * p.tile_rows;
    const int m = min(p.tile_rows, n - r0);
    if (tile != ntiles - 1) {
      __syncthreads();
      load_tile<Q>(p, a + r0, m, false, s_style, s_cap, s_ret, s_ind);
      cp_async_wait_all();
      __syncthreads();
    }
    for (int i = tid; i < m; i += nt) {
      double e = s_ret[i] - c0 - s_g[bucket_of(s_ind[i], P)];
      if (Q > 0) {
        double d[QQ];
        load_row_styles<Q>(s_style, i, s_piv, d);
#pragma unroll
        for (int q = 0; q < Q; ++q) e = fma(-s_beta[q], d[q], e);
      }
      st_stream(p.spec + a + r0 + i, e);
      se += e;
      see = fma(e, e, see);
    }
  }
  se = warp_sum(se);
  see = warp_sum(see);
  if (lane == 0) { s_red[2 * warp] = se; s_red[2 * warp + 1] = see; }
  __syncthreads();
  if (tid == 0) {
    double a1 = 0.0, a2 = 0.0;
    for (int w = 0; w < nwarps; ++w) { a1 += s_red[2 * w]; a2 += s_red[2 * w + 1]; }
    const double dn = (double)n;
    const double me = a1 / dn;
    const double var_e = a2 / dn - me * me;
    const double Y1 = s_acc[NA1PAD + 3 * Q], Y2 = s_acc[NA1PAD + 3 * Q + 1];
    const double my = Y1 / dn;
    const double var_y = Y2 / dn - my * my;
    p.r2[t] = 1.0 - var_e / var_y;
  }
}

// ---- dense solve: the reference's formulation, numpy pinv semantics (whole CTA) -----------
// Builds G = X'WX and b = X'Wy from the same accumulators (standardised styles), applies the
// constraint matrix R of CrossSection.py:69-71 (row P <- -cap_p/cap_last, column P deleted),
// eigen-decomposes R'GR and keeps eigenvalues with |lambda| > 1e-15 * max|lambda| — numpy's
// default pinv cutoff on singular values (CrossSection.py:76; :98 when P == 0).
template <int Q>
__device__ int solve_dense_cta(const RegParams& p, const RegSmem& L, unsigned char* smem, int n_rows, int t) {
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31;
  const int P = p.P, NB = P + 1, BV = 3 + Q, K = p.K, ld = K | 1;
  constexpr int NA1PAD = na1pad_of(Q);
  const double* s_bucket = reinterpret_cast<const double*>(smem + L.bucket);
  const double* s_acc = reinterpret_cast<const double*>(smem + L.acc);
  double* s_piv = reinterpret_cast<double*>(smem + L.coef);
  double* s_mp = s_piv + Q;
  double* s_beta = s_mp + Q;
  double* s_g = s_beta + Q;
  double* s_misc = s_g + NB;
  double* s_f = reinterpret_cast<double*>(smem + L.fvec);
  double* G = reinterpret_cast<double*>(smem + L.G);
  double* V = reinterpret_cast<double*>(smem + L.V);
  double* G0 = reinterpret_cast<double*>(smem + L.G0);
  double* cs = reinterpret_cast<double*>(smem + L.cs);    // K+2 rotation scratch
  double* bvec = cs + K + 2;                               // K
  double* rr = bvec + K;                                   // K
  double* xr = rr + K;                                     // K (+ few)
  double* SD = reinterpret_cast<double*>(smem + L.solve);  // Q
  int* flag = reinterpret_cast<int*>(smem + L.seg);
  const double* S2 = s_acc;
  const double* BY = s_acc + Q * (Q + 1) / 2;
  const double* CD = s_acc + NA1PAD;
  const double* D1 = CD + Q;
  const double* D2 = D1 + Q;
#define BW(b) s_bucket[(b)*BV + 0]
#define BWY(b) s_bucket[(b)*BV + 1]
#define BC(b) s_bucket[(b)*BV + 2]
#define BD(b, q) s_bucket[(b)*BV + 3 + (q)]
  if (tid < 32) {
    double sw = 0, capall = 0, bc = 0;
    for (int b = lane; b < NB; b += 32) { sw += BW(b); capall += BC(b); bc += BWY(b); }
    sw = warp_sum(sw); capall = warp_sum(capall); bc = warp_sum(bc);
    const double dn = (double)n_rows;
    double sigma = 1.0;
    if (Q > 0) {
      if (lane < Q) {
        s_mp[lane] = CD[lane] / capall;
        double sd = 0;
        for (int b = 0; b < NB; ++b) sd += BD(b, lane);
        SD[lane] = sd;
      }
      const double v = lane < Q ? D1[lane] + dn * s_piv[lane] : 0.0;
      const double mu = warp_sum(v) / (dn * Q);
      double t2 = 0.0;
      if (lane < Q) {
        const double dq = s_piv[lane] - mu;
        t2 = D2[lane] + 2.0 * dq * D1[lane] + dn * dq * dq;
      }
      sigma = sqrt(warp_sum(t2) / (dn * Q));
    }
    if (lane == 0) { s_misc[2] = sigma; s_misc[3] = sw; s_misc[4] = bc; }
  }
  __syncthreads();
  const double sigma = s_misc[2], sw = s_misc[3], bc = s_misc[4];
  const double is = 1.0 / sigma;
  for (int e = tid; e < K * K; e += nt) {
    int i = e / K, j = e - i * K;
    if (i > j) { const int tt = i; i = j; j = tt; }
    double v;
    if (i == 0) {
      if (j == 0) v = sw;
      else if (j <= P) v = BW(j - 1);
      else { const int q = j - 1 - P; v = (SD[q] - s_mp[q] * sw) * is; }
    } else if (i <= P) {
      if (j <= P) v = (i == j) ? BW(i - 1) : 0.0;
      else { const int q = j - 1 - P; v = (BD(i - 1, q) - s_mp[q] * BW(i - 1)) * is; }
    } else {
      const int q = i - 1 - P, r = j - 1 - P;   // q <= r
      v = (S2[r * (r + 1) / 2 + q] - s_mp[q] * SD[r] - s_mp[r] * SD[q] + s_mp[q] * s_mp[r] * sw) * is * is;
    }
    const int ii = e / K, jj = e - ii * K;
    G[ii * ld + jj] = v;
    G0[ii * ld + jj] = v;
  }
  for (int k = tid; k < K; k += nt) {
    double v;
    if (k == 0) v = bc;
    else if (k <= P) v = BWY(k - 1);
    else { const int q = k - 1 - P; v = (BY[q] - s_mp[q] * bc) * is; }
    bvec[k] = v;
    rr[k] = (k >= 1 && k < P) ? BC(k - 1) / BC(P - 1) : 0.0;
  }
  __syncthreads();
  int n = K;
  double* A = G;
  double* EV = V;
  if (P > 0) {
    n = K - 1;
    // G R : columns 1..P-1 -= rr * column P
    for (int e = tid; e < K * (P - 1); e += nt) {
      const int i = e / (P - 1), c = 1 + e % (P - 1);
      G[i * ld + c] = fma(-rr[c], G[i * ld + P], G[i * ld + c]);
    }
    __syncthreads();
    // R' (G R): rows 1..P-1 -= rr * row P ; same for b
    for (int e = tid; e < K * (P - 1); e += nt) {
      const int j = e / (P - 1), c = 1 + e % (P - 1);
      G[c * ld + j] = fma(-rr[c], G[P * ld + j], G[c * ld + j]);
    }
    if (tid >= 1 && tid < P) bvec[tid] = fma(-rr[tid], bvec[P], bvec[tid]);
    for (int c = P + nt; c < P; c += nt) {}   // (P-1 <= 63 < blockDim)
    __syncthreads();
    // drop row/column P -> compact into V's storage
    for (int e = tid; e < n * n; e += nt) {
      const int i = e / n, j = e - i * n;
      V[i * ld + j] = G[(i < P ? i : i + 1) * ld + (j < P ? j : j + 1)];
    }
    if (tid < n) xr[tid] = bvec[tid < P ? tid : tid + 1];
    __syncthreads();
    if (tid < n) bvec[tid] = xr[tid];
    A = V;
    EV = G;
    __syncthreads();
  }
  // non-finite Gram: numpy's SVD raises LinAlgError
  if (tid == 0) flag[4] = 0;
  __syncthreads();
  {
    int bad = 0;
    for (int e = tid; e < n * n; e += nt) bad |= !isfinite(A[(e / n) * ld + e % n]);
    for (int k = tid; k < n; k += nt) bad |= !isfinite(bvec[k]);
    if (bad) flag[4] = 1;
  }
  __syncthreads();
  if (flag[4]) return BMFM_ST_NONFINITE;

  jacobi_eig_cta(A, EV, n, ld, cs, flag);
  __syncthreads();
  // pinv: keep |lambda| > 1e-15 * max|lambda|
  double* lam = cs;            // n  (rotation scratch is free now)
  double* coef = rr;           // reuse after saving ratios
  double* ratio = xr + K;      // not available -> keep ratios in registers instead
  (void)ratio;
  double myratio = 0.0;
  if (tid >= 1 && tid < P) myratio = rr[tid];
  __syncthreads();
  if (tid < n) lam[tid] = A[tid * ld + tid];
  __syncthreads();
  double mx = 0.0;
  for (int k = 0; k < n; ++k) mx = fmax(mx, fabs(lam[k]));
  const double cutoff = 1e-15 * mx;
  int dropped = 0;
  for (int k = 0; k < n; ++k) dropped |= !(fabs(lam[k]) > cutoff);
  if (tid < n) {
    double dot = 0.0;
    for (int j = 0; j < n; ++j) dot = fma(EV[j * ld + tid], bvec[j], dot);
    coef[tid] = (fabs(lam[tid]) > cutoff) ? dot / lam[tid] : 0.0;
  }
  __syncthreads();
  if (tid < n) {
    double x = 0.0;
    for (int k = 0; k < n; ++k) x = fma(EV[tid * ld + k], coef[k], x);
    xr[tid] = x;
  }
  __syncthreads();
  // f = R x
  if (tid < K) {
    double f;
    if (P == 0) f = xr[tid];
    else if (tid < P) f = xr[tid];
    else if (tid > P) f = xr[tid - 1];
    else f = 0.0;
    s_f[tid] = f;
  }
  // the ratios were overwritten by coef: recompute the constrained entry from registers
  if (P > 0) {
    double part = (tid >= 1 && tid < P) ? -myratio * xr[tid] : 0.0;
    part = warp_sum(part);
    double* s_red = reinterpret_cast<double*>(smem + L.red);
    if (lane == 0) s_red[tid >> 5] = part;
    __syncthreads();
    if (tid == 0) {
      double s = 0.0;
      for (int w = 0; w < (nt >> 5); ++w) s += s_red[w];
      s_f[P] = s;
    }
  }
  __syncthreads();

  if (p.exposure != nullptr) {
    // exposure = R pinv(R'GR) R' G   (CrossSection.py:104; weight normalisation cancels)
    double* s_red = reinterpret_cast<double*>(smem + L.red);
    (void)s_red;
    // restore ratios
    if (tid >= 1 && tid < P) rr[tid] = myratio;
    if (tid == 0 || tid >= P) { if (tid < K) rr[tid] = 0.0; }
    __syncthreads();
    // M1 = R' G0 : rows 1..P-1 -= ratio * row P (row P then ignored)
    if (P > 0) {
      for (int e = tid; e < K * (P - 1); e += nt) {
        const int j = e / (P - 1), c = 1 + e % (P - 1);
        G0[c * ld + j] = fma(-rr[c], G0[P * ld + j], G0[c * ld + j]);
      }
    }
    __syncthreads();
    // Ginv = EV diag(inv) EV'  -> A's storage
    for (int e = tid; e < n * n; e += nt) {
      const int i = e / n, j = e - i * n;
      double s = 0.0;
      for (int k = 0; k < n; ++k) {
        const double inv = (fabs(lam[k]) > cutoff) ? 1.0 / lam[k] : 0.0;
        s = fma(EV[i * ld + k] * inv, EV[j * ld + k], s);
      }
      A[i * ld + j] = s;
    }
    __syncthreads();
    // M2 = Ginv M1 (n x K) -> EV's storage
    for (int e = tid; e < n * K; e += nt) {
      const int i = e / K, j = e - i * K;
      double s = 0.0;
      for (int k = 0; k < n; ++k) s = fma(A[i * ld + k], G0[((P > 0 && k >= P) ? k + 1 : k) * ld + j], s);
      EV[i * ld + j] = s;
    }
    __syncthreads();
    double* out = p.exposure + (size_t)t * K * K;
    for (int e = tid; e < K * K; e += nt) {
      const int i = e / K, j = e - i * K;
      double v;
      if (P == 0 || i < P) v = EV[i * ld + j];
      else if (i > P) v = EV[(i - 1) * ld + j];
      else {
        v = 0.0;
        for (int c = 1; c < P; ++c) v = fma(-rr[c], EV[c * ld + j], v);
      }
      out[e] = v;
    }
    __syncthreads();
  }
  // coefficients for pass C
  if (tid < Q) s_beta[tid] = s_f[1 + P + tid] * is;
  for (int b = tid; b < NB; b += nt) s_g[b] = b < P ? s_f[1 + b] : 0.0;
  __syncthreads();
  if (tid == 0) {
    double c0 = s_f[0];
    for (int q = 0; q < Q; ++q) c0 -= s_beta[q] * s_mp[q];
    s_misc[1] = c0;
  }
  int bad = 0;
  for (int k = tid; k < K; k += nt) bad |= !isfinite(s_f[k]);
  bad = __syncthreads_or(bad);
#undef BW
#undef BWY
#undef BC
#undef BD
  if (bad) return BMFM_ST_NONFINITE;
  return dropped ? BMFM_ST_PINV : BMFM_ST_OK;
}

// ---- kernels ---------------------------------------------------------------------------------
__device__ __forceinline__ void write_nan_outputs(const RegParams& p, long long a, int n, int t) {
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int k = threadIdx.x; k < p.K; k += blockDim.x) store_factor_ret(p, t, k, nan);
  for (int i = threadIdx.x; i < n; i += blockDim.x) p.spec[a + i] = nan;
  if (threadIdx.x == 0) p.r2[t] = nan;
}

template <int Q>
__global__ void __launch_bounds__(256) reg_fast_kernel(RegParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const RegSmem L = reg_smem_layout(p.tile_rows, p.P, Q, blockDim.x >> 5, false);
  const double* s_f = reinterpret_cast<const double*>(smem + L.fvec);
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    if (n <= 0) {
      write_nan_outputs(p, a, 0, t);
      if (threadIdx.x == 0) p.status[t] = BMFM_ST_NONFINITE;
      continue;
    }
    int ntiles;
    accumulate_date<Q>(p, L, smem, a, n, ntiles);
    const int st = solve_fast_cta<Q, false>(p, L, smem, n);
    if (st == BMFM_ST_OK) {
      for (int k = threadIdx.x; k < p.K; k += blockDim.x) store_factor_ret(p, t, k, s_f[k]);
      residual_date<Q>(p, L, smem, a, n, ntiles, t);
      if (threadIdx.x == 0) p.status[t] = BMFM_ST_OK;
    } else if (st == REG_NEED_DENSE) {
      if (threadIdx.x == 0) p.status[t] = REG_NEED_DENSE;
    } else {
      write_nan_outputs(p, a, n, t);
      if (threadIdx.x == 0) p.status[t] = st;
    }
    __syncthreads();
  }
}

template <int Q>
__global__ void __launch_bounds__(256) reg_dense_kernel(RegParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const RegSmem L = reg_smem_layout(p.tile_rows, p.P, Q, blockDim.x >> 5, true);
  const double* s_f = reinterpret_cast<const double*>(smem + L.fvec);
  // one parallel look at the statuses of this CTA's dates (a serial walk costs a global-memory latency per date)
  if (!p.all_dense) {
    int any = 0;
    for (int t = blockIdx.x + (int)threadIdx.x * (int)gridDim.x; t < p.T; t += (int)(blockDim.x * gridDim.x))
      any |= p.status[t] == REG_NEED_DENSE;
    if (!__syncthreads_or(any)) return;
  }
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    if (!p.all_dense && p.status[t] != REG_NEED_DENSE) continue;
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    if (n <= 0) {
      write_nan_outputs(p, a, 0, t);
      if (threadIdx.x == 0) p.status[t] = BMFM_ST_NONFINITE;
      continue;
    }
    int ntiles;
    accumulate_date<Q>(p, L, smem, a, n, ntiles);
    const int st = solve_dense_cta<Q>(p, L, smem, n, t);
    if (st == BMFM_ST_NONFINITE) {
      write_nan_outputs(p, a, n, t);
    } else {
      for (int k = threadIdx.x; k < p.K; k += blockDim.x) store_factor_ret(p, t, k, s_f[k]);
      residual_date<Q>(p, L, smem, a, n, ntiles, t);
    }
    if (threadIdx.x == 0) p.status[t] = st;
    __syncthreads();
  }
}

// ---- streaming kernel: one WARP per date (small cross-sections) ---------------------------------
// For N of a few hundred to ~2000 stocks a CTA per date spends most of its time in barriers around
// short serial phases.  Here every warp owns a date end to end and never meets a CTA barrier: rows
// flow through a per-warp cp.async ring (32-row chunks, 4 stages), pass A accumulates in registers
// (25 rows per lane at N = 800, one warp reduction per date), the per-industry sums are a running
// walk (lane = value column, flush on industry change), the solve is the same block elimination,
// and pass C replays the date through the ring — the second read is served by the 126 MB L2 (the
// in-flight footprint of 8 warps x 148 SMs x 80 KB is below it), so HBM still sees ~one read.
constexpr int RS_STAGES = 4;
constexpr int RS_WARPS = 4;   // warps per CTA of the streaming kernel
constexpr int RS_CTAS = 2;    // CTAs per SM

template <int Q>
__host__ __device__ constexpr int rs_stage_bytes() { return 32 * (Q * 8 + 8 + 8 + 4); }
// per-row staging of the 3+Q per-industry terms; odd stride (in doubles) keeps the row-per-lane
// stores and the column-per-lane loads free of bank conflicts
constexpr int RS_ZS = 17;
template <int Q>
__host__ __device__ inline int rs_per_warp_bytes(int fixed) {
  return align16(fixed + RS_STAGES * rs_stage_bytes<Q>() + 32 * RS_ZS * 8 + 16);
}

template <int Q>
__device__ __forceinline__ void rs_issue_chunk(const RegParams& p, long long row0, int m, bool with_cap,
                                               unsigned char* stage, int lane) {
  double* st_style = reinterpret_cast<double*>(stage);
  double* st_cap = st_style + 32 * Q;
  double* st_ret = st_cap + 32;
  int* st_ind = reinterpret_cast<int*>(st_ret + 32);
  if (lane < m) {
    if (with_cap) cp_async_8(&st_cap[lane], p.cap + row0 + lane);
    cp_async_8(&st_ret[lane], p.ret + row0 + lane);
    cp_async_4(&st_ind[lane], p.ind + row0 + lane);
  }
  if (Q > 0) {
    const double* src = p.style + row0 * Q;
    const int tot = m * Q;
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
      if (m == 32) {
        // full chunk: 16*Q pieces of 16 B, a compile-time trip count per lane
#pragma unroll
        for (int j = 0; j < (16 * Q + 31) / 32; ++j) {
          const int e = lane + 32 * j;
          if (e < 16 * Q) cp_async_16(&st_style[2 * e], src + 2 * e);
        }
      } else {
        const int n16 = tot >> 1;
        for (int e = lane; e < n16; e += 32) cp_async_16(&st_style[2 * e], src + 2 * e);
        if ((tot & 1) && lane == 0) cp_async_8(&st_style[tot - 1], src + tot - 1);
      }
    } else {
      for (int e = lane; e < tot; e += 32) cp_async_8(&st_style[e], src + e);
    }
  }
  asm volatile("cp.async.commit_group;\n" ::: "memory");
}

template <int Q>
// 128 threads x 2 CTAs = 8 warps per SM at 255 registers (97 fp64 accumulators); 9 warps at 224 registers
// was tried and is 1.9x slower (spills inside the accumulation loop)
__global__ void __launch_bounds__(RS_WARPS * 32, RS_CTAS) reg_stream_kernel(RegParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int NA1PAD = na1pad_of(Q), NA2PAD = na2pad_of(Q);
  constexpr int QQ = Q > 0 ? Q : 1;
  constexpr int SB = rs_stage_bytes<Q>();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int P = p.P, NB = P + 1, BV = 3 + Q, K = p.K;
  const RegSmem L = reg_smem_layout(0, P, Q, 1, false);
  const int per_warp = rs_per_warp_bytes<Q>(L.total);
  unsigned char* wsm = smem + (size_t)warp * per_warp;
  unsigned char* ring = wsm + L.total;
  double* s_z = reinterpret_cast<double*>(ring + RS_STAGES * SB);   // [32][RS_ZS] per-row bucket terms
  double* s_bucket = reinterpret_cast<double*>(wsm + L.bucket);
  double* s_acc = reinterpret_cast<double*>(wsm + L.acc);
  double* s_piv = reinterpret_cast<double*>(wsm + L.coef);
  double* s_beta = s_piv + 2 * Q;
  double* s_g = s_beta + Q;
  double* s_misc = s_g + NB;
  const double* s_f = reinterpret_cast<const double*>(wsm + L.fvec);
  const int gw = blockIdx.x * nwarps + warp, tw = gridDim.x * nwarps;
  // lane roles of the per-industry walk: two row streams x 16 value columns
  const int h = lane >> 4, v = lane & 15;
  const bool vact = v < BV;

  for (int t = gw; t < p.T; t += tw) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (n <= 0) {
      for (int k = lane; k < K; k += 32) store_factor_ret(p, t, k, nan);
      if (lane == 0) { p.r2[t] = nan; p.status[t] = BMFM_ST_NONFINITE; }
      continue;
    }
    const int nchunks = (n + 31) >> 5;
    for (int e = lane; e < NB * BV; e += 32) s_bucket[e] = 0.0;
    // prologue of the ring
#pragma unroll
    for (int c = 0; c < RS_STAGES - 1; ++c) {
      if (c < nchunks) rs_issue_chunk<Q>(p, a + 32 * c, min(32, n - 32 * c), true, ring + c * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    // the pivot shift is the date's first row: take it from chunk 0 once it has landed
    asm volatile("cp.async.wait_group %0;\n" ::"n"(RS_STAGES - 2) : "memory");
    __syncwarp();
    {
      const double* st0 = reinterpret_cast<const double*>(ring);
      if (lane < Q) s_piv[lane] = st0[lane];
      if (lane == 0) s_misc[0] = st0[32 * Q + 32];
    }
    __syncwarp();
    const double ypiv = s_misc[0];
    double acc1[NA1PAD];
    double acc2[NA2PAD];
#pragma unroll
    for (int i = 0; i < NA1PAD; ++i) acc1[i] = 0.0;
#pragma unroll
    for (int i = 0; i < NA2PAD; ++i) acc2[i] = 0.0;
    int curb = -1;
    double bacc = 0.0;

    for (int c = 0; c < nchunks; ++c) {
      const int cn = c + RS_STAGES - 1;
      if (cn < nchunks) rs_issue_chunk<Q>(p, a + 32 * cn, min(32, n - 32 * cn), true, ring + (cn % RS_STAGES) * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
      asm volatile("cp.async.wait_group %0;\n" ::"n"(RS_STAGES - 1) : "memory");
      __syncwarp();
      unsigned char* stage = ring + (c % RS_STAGES) * SB;
      const double* st_style = reinterpret_cast<const double*>(stage);
      const double* st_cap = st_style + 32 * Q;
      const double* st_ret = st_cap + 32;
      const int* st_ind = reinterpret_cast<const int*>(st_ret + 32);
      const int m = min(32, n - 32 * c);
      if (lane < m) {
        const double cap = st_cap[lane];
        const double w = sqrt(cap);
        const double y = st_ret[lane];
        double* zr = s_z + lane * RS_ZS;
        zr[0] = w;
        zr[1] = w * y;
        zr[2] = cap;
        const double u = y - ypiv;
        acc2[3 * Q] += u;
        acc2[3 * Q + 1] = fma(u, u, acc2[3 * Q + 1]);
        if (Q > 0) {
          double d[QQ];
          load_row_styles<Q>(st_style, lane, s_piv, d);
#pragma unroll
          for (int q = 0; q < Q; ++q) {
            const double wd = w * d[q];
            zr[3 + q] = wd;
#pragma unroll
            for (int r = 0; r <= q; ++r) acc1[q * (q + 1) / 2 + r] = fma(wd, d[r], acc1[q * (q + 1) / 2 + r]);
            acc1[Q * (Q + 1) / 2 + q] = fma(wd, y, acc1[Q * (Q + 1) / 2 + q]);
            acc2[q] = fma(cap, d[q], acc2[q]);
            acc2[Q + q] += d[q];
            acc2[2 * Q] = fma(d[q], d[q], acc2[2 * Q]);   // only the total over q enters the scalar std
          }
        }
      }
      __syncwarp();
      // per-industry running sums: lane = value column (two row streams of 16), flush when the
      // industry changes.  Run starts come from one ballot; the terms were staged by pass A.
      {
        const int myb = lane < m ? bucket_of(st_ind[lane], P) : -2;
        const int prevb = __shfl_up_sync(0xffffffffu, myb, 1);
        const unsigned runstart = __ballot_sync(0xffffffffu, lane > 0 && lane < m && myb != prevb);
        const int rlo = h * 16, rhi = min(m, rlo + 16);
        const int bfirst = __shfl_sync(0xffffffffu, myb, rlo);
        if (rlo < rhi) {
          unsigned flush = (runstart >> rlo) & 0xffffu;
          flush = (flush & ~1u) | (bfirst != curb ? 1u : 0u);
          const double* zc = s_z + rlo * RS_ZS + v;
          for (int r = 0; r < rhi - rlo; ++r) {
            if ((flush >> r) & 1u) {
              if (curb >= 0 && vact) atomicAdd(&s_bucket[curb * BV + v], bacc);
              bacc = 0.0;
              curb = bucket_of(st_ind[rlo + r], P);
            }
            if (vact) bacc += zc[r * RS_ZS];
          }
        }
      }
      __syncwarp();   // the stage may be refilled by the next iteration's loads
    }
    if (curb >= 0 && vact) atomicAdd(&s_bucket[curb * BV + v], bacc);
    // warp reduction of the register accumulators through the (now idle) ring: every lane stores
    // its values as a column, then each lane sums one row — no shuffles, ~1/7 of the instructions
    {
      constexpr int NA1 = na1_of(Q), NA2 = na2_of(Q), NTOT = NA1 + NA2;
      constexpr int RWX = (RS_STAGES * SB) / (33 * 8), RW = RWX < 48 ? RWX : 48;   // rows of 33 doubles that fit the ring
      static_assert(RW >= 1, "ring too small for the reduction buffer");
      double* red = reinterpret_cast<double*>(ring);
#pragma unroll
      for (int r0 = 0; r0 < NTOT; r0 += RW) {
#pragma unroll
        for (int j = 0; j < RW; ++j) {
          const int i = r0 + j;
          if (i < NTOT) red[j * 33 + lane] = (i < NA1) ? acc1[i < NA1 ? i : 0] : acc2[i >= NA1 ? i - NA1 : 0];
        }
        __syncwarp();
        for (int j = lane; j < RW && r0 + j < NTOT; j += 32) {
          const double* row = red + j * 33;
          double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
          for (int l = 0; l < 32; l += 4) { s0 += row[l]; s1 += row[l + 1]; s2 += row[l + 2]; s3 += row[l + 3]; }
          const int i = r0 + j;
          s_acc[i < NA1 ? i : NA1PAD + (i - NA1)] = (s0 + s1) + (s2 + s3);
        }
        __syncwarp();
      }
    }
    // start refilling the ring for pass C while the solve runs (served by L2)
#pragma unroll
    for (int c = 0; c < RS_STAGES - 1; ++c) {
      if (c < nchunks) rs_issue_chunk<Q>(p, a + 32 * c, min(32, n - 32 * c), false, ring + c * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    const int st = solve_fast_cta<Q, true>(p, L, wsm, n);
    if (st != BMFM_ST_OK) {
      // drain the prefetched ring before leaving the date
      asm volatile("cp.async.wait_group 0;\n" ::: "memory");
      __syncwarp();
      if (st == REG_NEED_DENSE) {
        if (lane == 0) p.status[t] = REG_NEED_DENSE;
      } else {
        for (int k = lane; k < K; k += 32) store_factor_ret(p, t, k, nan);
        for (int i = lane; i < n; i += 32) p.spec[a + i] = nan;
        if (lane == 0) { p.r2[t] = nan; p.status[t] = st; }
      }
      continue;
    }
    for (int k = lane; k < K; k += 32) store_factor_ret(p, t, k, s_f[k]);
    // pass C
    const double c0 = s_misc[1];
    double se = 0.0, see = 0.0;
    for (int c = 0; c < nchunks; ++c) {
      const int cn = c + RS_STAGES - 1;
      if (cn < nchunks) rs_issue_chunk<Q>(p, a + 32 * cn, min(32, n - 32 * cn), false, ring + (cn % RS_STAGES) * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
      asm volatile("cp.async.wait_group %0;\n" ::"n"(RS_STAGES - 1) : "memory");
      __syncwarp();
      unsigned char* stage = ring + (c % RS_STAGES) * SB;
      const double* st_style = reinterpret_cast<const double*>(stage);
      const double* st_ret = st_style + 32 * Q + 32;
      const int* st_ind = reinterpret_cast<const int*>(st_ret + 32);
      const int m = min(32, n - 32 * c);
      if (lane < m) {
        double e = st_ret[lane] - c0 - s_g[bucket_of(st_ind[lane], P)];
        if (Q > 0) {
          double d[QQ];
          load_row_styles<Q>(st_style, lane, s_piv, d);
#pragma unroll
          for (int q = 0; q < Q; ++q) e = fma(-s_beta[q], d[q], e);
        }
        st_stream(p.spec + a + 32 * c + lane, e);
        se += e;
        see = fma(e, e, see);
      }
      __syncwarp();
    }
    se = warp_sum(se);
    see = warp_sum(see);
    if (lane == 0) {
      const double dn = (double)n;
      const double me = se / dn;
      const double var_e = see / dn - me * me;
      const double my = s_acc[NA1PAD + 3 * Q] / dn;
      const double var_y = s_acc[NA1PAD + 3 * Q + 1] / dn - my * my;
      p.r2[t] = 1.0 - var_e / var_y;
      p.status[t] = BMFM_ST_OK;
    }
    __syncwarp();
  }
}

#include "reg_mma.cuh"
#include "reg_bulk.cuh"

// ---- host side ---------------------------------------------------------------------------------
template <int Q, int NSTAGE, int WPD>
static int launch_reg_mma(bmfm_ctx* ctx, const RegParams& p) {
  int warps = WPD > 1 ? WPD : RM_WARPS;
  if (WPD == 1)
    if (ctx->opt_reg_warps >= 1 && ctx->opt_reg_warps <= RM_WARPS) warps = ctx->opt_reg_warps;
  const int smem_s = rm_cta_bytes<Q, NSTAGE>(p.P, WPD, warps);
  BMFM_CUDA(ctx, cudaFuncSetAttribute(reg_mma_kernel<Q, NSTAGE, WPD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_s));
  int per_sm = (228 * 1024) / (smem_s + 1024);
  if (per_sm < 1) per_sm = 1;
  if (ctx->opt_reg_ctas > 0) per_sm = ctx->opt_reg_ctas;
  int grid = ctx->sm_count * per_sm;
  const int dates_per_cta = WPD > 1 ? 1 : warps;
  if (grid * dates_per_cta > p.T) grid = (p.T + dates_per_cta - 1) / dates_per_cta;
  RegParams pk = p;
  pk.next_date = nullptr;
  // measured: with three or more rounds of dates per slot the counter wins 5% (5000 x 3750: 0.858 against 0.904 ms);
  // with fewer rounds the static stride is as good or better (800 x 2500: 0.146 against 0.162 ms)
  const long long slots_total = (long long)grid * dates_per_cta;
  const bool dynamic = ctx->opt_reg_static == 2 || (ctx->opt_reg_static == 0 && p.T >= 3 * slots_total);
  if (dynamic && slots_total < p.T) {
    int rc = bmfm_stage_reserve(ctx, 15, 256);
    if (rc) return rc;
    pk.next_date = (int*)ctx->stage[15];
    BMFM_CUDA(ctx, cudaMemsetAsync(pk.next_date, 0, sizeof(int), ctx->stream));
  }
  reg_mma_kernel<Q, NSTAGE, WPD><<<grid, warps * 32, smem_s, ctx->stream>>>(pk);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// the bulk-copy form (reg_mode 6): four warps per date, even Q in 8..10 (others fall back to reg_mode 4)
template <int Q>
static int launch_reg_bulk(bmfm_ctx* ctx, const RegParams& p) {
  if constexpr (Q % 2 == 0 && rm_has_s(Q)) {
    constexpr int NS = 2;
    const int smem_s = rb_cta_bytes<Q, NS>(p.P);
    BMFM_CUDA(ctx, cudaFuncSetAttribute(reg_bulk_kernel<Q, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_s));
    int per_sm = (228 * 1024) / (smem_s + 1024);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    if (ctx->opt_reg_ctas > 0) per_sm = ctx->opt_reg_ctas;
    int grid = ctx->sm_count * per_sm;
    if (grid > p.T) grid = p.T;
    reg_bulk_kernel<Q, NS><<<grid, RB_WARPS * 32, smem_s, ctx->stream>>>(p);
    ctx->launches++;
    BMFM_CUDA(ctx, cudaGetLastError());
    return BMFM_OK;
  } else {
    return launch_reg_mma<Q, 2, RM_WARPS>(ctx, p);
  }
}

template <int Q>
static int launch_reg(bmfm_ctx* ctx, RegParams p, int max_rows) {
  const int SMEM_MAX = ctx->max_smem_optin;           // 227 KB on B200
  const int SM_TOTAL = 228 * 1024;
  // fast kernel: size the tile so a whole date stays resident; more CTAs/SM for small dates
  int rows = max_rows > 0 ? (max_rows + 31) / 32 * 32 : 1024;
  int nthreads = rows <= 1536 ? 128 : 256;
  RegSmem L = reg_smem_layout(rows, p.P, Q, nthreads / 32, false);
  if (L.total > SMEM_MAX) {
    // a date does not fit one CTA: stream it in tiles (pass C re-reads all but the last)
    nthreads = 256;
    const int fixed = reg_smem_layout(0, p.P, Q, nthreads / 32, false).total;
    rows = (SMEM_MAX - fixed - 64) / (Q * 8 + 28) / 32 * 32;
    L = reg_smem_layout(rows, p.P, Q, nthreads / 32, false);
  }
  int per_sm = SM_TOTAL / (L.total + 1024);
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  p.tile_rows = rows;
  // Default: the FP64-MMA kernel with 1, 2 or 4 warps per date (modes 3, 5, 4): one warp per date is the
  // best grain for the serial solve, more warps per date when the cross-section is long (per-date latency,
  // L2 footprint of the dates in flight) or when there are too few dates to fill the SMs.  BMFM_REG_MODE
  // forces a kernel: 1 = CTA kernel with the date resident in shared memory, 2 = warp-per-date kernel with
  // register accumulators (the two earlier designs, kept for comparison).
  int mode = ctx->opt_reg_mode;
  if (mode < 1 || mode > 6) {
    int wpd = (max_rows > 0 && max_rows <= 1024) ? 1 : 2;
    const long long slots = (long long)ctx->sm_count * 12;      // warps the SMs hold with one warp per date
    // pass C re-reads a date when its solve is done: with one warp per date the dates in flight (slots x rows x
    // ~100 B) must stay in L2 for that.  At 800 x 2500 they are 142 MB and 64% of the second pass goes back to HBM
    // (ncu: 341 MB of DRAM traffic for 216 MB of algorithmic bytes); two warps per date halve the dates in flight:
    // 241 MB (1.12x) at the same kernel time.
    {
      const long long inflight = (p.T < slots ? p.T : slots) * (long long)(max_rows > 0 ? max_rows : 1024) * (Q * 8 + 28);
      if (wpd == 1 && inflight > (64LL << 20)) wpd = 2;
    }
    while (wpd < 4 && (long long)p.T * wpd * 2 <= slots) wpd *= 2;
    mode = wpd == 1 ? 3 : (wpd == 2 ? 5 : 4);
  }
  const bool use_stream = mode == 2;
  ctx->last_reg_mode = p.all_dense ? 0 : mode;
  if (!p.all_dense && mode >= 3 && mode <= 6) {
    const int rc = mode == 3 ? launch_reg_mma<Q, 3, 1>(ctx, p)
                 : mode == 5 ? launch_reg_mma<Q, 2, 2>(ctx, p)
                 : mode == 6 ? launch_reg_bulk<Q>(ctx, p) : launch_reg_mma<Q, 2, RM_WARPS>(ctx, p);
    if (rc != BMFM_OK) return rc;
  } else if (!p.all_dense && use_stream) {
    const RegSmem Lw = reg_smem_layout(0, p.P, Q, 1, false);
    const int per_warp = rs_per_warp_bytes<Q>(Lw.total);
    const int smem_s = RS_WARPS * per_warp;
    BMFM_CUDA(ctx, cudaFuncSetAttribute(reg_stream_kernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_s));
    int grid = ctx->sm_count * RS_CTAS;
    if (grid * RS_WARPS > p.T) grid = (p.T + RS_WARPS - 1) / RS_WARPS;
    reg_stream_kernel<Q><<<grid, RS_WARPS * 32, smem_s, ctx->stream>>>(p);
    ctx->launches++;
    BMFM_CUDA(ctx, cudaGetLastError());
  } else if (!p.all_dense) {
    BMFM_CUDA(ctx, cudaFuncSetAttribute(reg_fast_kernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
    int grid = ctx->sm_count * per_sm;
    if (grid > p.T) grid = p.T;
    reg_fast_kernel<Q><<<grid, nthreads, L.total, ctx->stream>>>(p);
    ctx->launches++;
    BMFM_CUDA(ctx, cudaGetLastError());
  }
  // dense kernel: flagged dates only (or all when the exposure matrix is wanted)
  {
    const int nthreads_d = 256;
    const int fixed = reg_smem_layout(0, p.P, Q, nthreads_d / 32, true).total;
    int rows_d = (SMEM_MAX - fixed - 64) / (Q * 8 + 28) / 32 * 32;
    if (rows_d < 32) {
      snprintf(ctx->err, sizeof(ctx->err), "K=%d too large for the dense path", p.K);
      return BMFM_E_UNSUPPORTED;
    }
    if (max_rows > 0 && rows_d > (max_rows + 31) / 32 * 32) rows_d = (max_rows + 31) / 32 * 32;
    RegSmem Ld = reg_smem_layout(rows_d, p.P, Q, nthreads_d / 32, true);
    RegParams pd = p;
    pd.tile_rows = rows_d;
    BMFM_CUDA(ctx, cudaFuncSetAttribute(reg_dense_kernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, Ld.total));
    int grid = ctx->sm_count;
    if (grid > p.T) grid = p.T;
    reg_dense_kernel<Q><<<grid, nthreads_d, Ld.total, ctx->stream>>>(pd);
    ctx->launches++;
    BMFM_CUDA(ctx, cudaGetLastError());
  }
  return BMFM_OK;
}

}  // namespace bmfm

using namespace bmfm;

extern "C" void bmfm_limits(int* max_Q, int* max_P, int* max_K) {
  if (max_Q) *max_Q = REG_MAX_Q;
  if (max_P) *max_P = REG_MAX_P;
  if (max_K) *max_K = 1 + REG_MAX_P + REG_MAX_Q;
}

static int reg_by_time_impl(bmfm_ctx* ctx, int T, int P, int Q, int max_rows_per_date,
                            const int64_t* date_off, const double* cap, const double* ret,
                            const int32_t* ind, const double* style,
                            double* factor_ret, double* specific_ret, double* r2,
                            double* exposure, int32_t* status, int n_peer, double* const* peer_bufs,
                            int64_t peer_row0) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && P >= 0 && Q >= 0, "T, P, Q must be non-negative");
  if (T == 0 && Q <= REG_MAX_Q && P <= REG_MAX_P) return BMFM_OK;   // an empty panel needs no buffers
  BMFM_ARG(ctx, date_off && cap && ret && ind && (style || Q == 0), "NULL input");
  BMFM_ARG(ctx, factor_ret && specific_ret && r2 && status, "NULL output");
  BMFM_ARG(ctx, n_peer >= 0 && n_peer <= REG_MAX_PEERS && (n_peer == 0 || peer_bufs) && peer_row0 >= 0, "bad peer list");
  if (Q > REG_MAX_Q || P > REG_MAX_P) {
    snprintf(ctx->err, sizeof(ctx->err), "P=%d Q=%d outside compiled limits (P<=%d, Q<=%d)", P, Q, REG_MAX_P, REG_MAX_Q);
    return BMFM_E_UNSUPPORTED;
  }
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  RegParams p;
  p.T = T; p.P = P; p.Q = Q; p.K = 1 + P + Q;
  p.date_off = reinterpret_cast<const long long*>(date_off);
  p.cap = cap; p.ret = ret; p.ind = ind; p.style = style;
  p.factor_ret = factor_ret; p.spec = specific_ret; p.r2 = r2; p.exposure = exposure; p.status = status;
  p.tile_rows = 0;
  p.all_dense = exposure != nullptr;
  p.n_peer = n_peer;
  p.peer_row0 = peer_row0;
  p.smma = 1;
  p.next_date = nullptr;
  if (ctx->opt_reg_smma >= 0) p.smma = ctx->opt_reg_smma;
  for (int r = 0; r < REG_MAX_PEERS; ++r) p.peer_fr[r] = r < n_peer ? peer_bufs[r] : nullptr;
  switch (Q) {
#define CASE(q) case q: return launch_reg<q>(ctx, p, max_rows_per_date);
    CASE(0) CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8) CASE(9) CASE(10) CASE(11) CASE(12)
#undef CASE
  }
  return BMFM_E_UNSUPPORTED;
}

extern "C" int bmfm_reg_by_time(bmfm_ctx* ctx, int T, int P, int Q, int max_rows_per_date,
                                const int64_t* date_off, const double* cap, const double* ret,
                                const int32_t* ind, const double* style,
                                double* factor_ret, double* specific_ret, double* r2,
                                double* exposure, int32_t* status) {
  return reg_by_time_impl(ctx, T, P, Q, max_rows_per_date, date_off, cap, ret, ind, style, factor_ret, specific_ret,
                          r2, exposure, status, 0, nullptr, 0);
}

// The same regression with the all-gather of the factor returns fused in: each date's K factor returns
// are stored by the regression kernel itself into the gathered series of every rank (`peer_bufs`: HOST
// array of n_peer device pointers to [T_total * K] buffers mapped for peer access, e.g. symmetric
// memory), at row peer_row0 + t.  The caller only needs a cross-rank barrier afterwards.
extern "C" int bmfm_reg_by_time_gather(bmfm_ctx* ctx, int T, int P, int Q, int max_rows_per_date,
                                       const int64_t* date_off, const double* cap, const double* ret,
                                       const int32_t* ind, const double* style,
                                       double* factor_ret, double* specific_ret, double* r2, int32_t* status,
                                       int n_peer, double* const* peer_bufs, int64_t peer_row0) {
  return reg_by_time_impl(ctx, T, P, Q, max_rows_per_date, date_off, cap, ret, ind, style, factor_ret, specific_ret,
                          r2, nullptr, status, n_peer, peer_bufs, peer_row0);
}
