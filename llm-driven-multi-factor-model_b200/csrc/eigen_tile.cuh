// Register-tile eigen solver of the Monte-Carlo loop (default solver of bmfm_eigen_mc_accumulate, K <= 48).
//
// Per (date, simulation) the adjustment needs, for every eigenpair (lambda_j, w_j) of the K x K matrix
//   S = D0^1/2 C_m D0^1/2        (orthogonally similar to the reference's Fm, utils.py:74-79),
// the ratio  v_j = (w_j' D0 w_j) / lambda_j  (utils.py:80-84).  Three kernels per batch of problems, the
// intermediate data (about 15 KB per problem) sized to stay in L2:
//
//   A  et_tridiag_kernel<NT>   one WARP per problem.  S lives in REGISTERS as NT x NT tiles of 8 x 8 in the
//                              layout of an FP64 MMA accumulator (lane = (row g = lane / 4, column pair
//                              c = lane % 4), two doubles per tile), full symmetric storage, so every lane
//                              holds the same share and all tile indices are compile-time.  Householder
//                              reduction to tridiagonal form: per step the column is passed through shared
//                              memory to get it in both the row and the column layout, the symmetric
//                              matrix-vector product is 2 FMAs per tile and a 4-lane reduction, the rank-2
//                              update 4 FMAs per tile.  The step loop is unrolled in phases of 8 columns,
//                              each phase touching only the tiles that are still active.  Output: the
//                              tridiagonal (d, e) and the reflectors (v_k, 1 / h_k) — Q is never formed.
//   B  et_chain_kernel         one THREAD per problem: square-root-free QL (Pal-Walker-Kahan, as LAPACK's dsterf)
//                              on (d, e^2), eigenvalues only (the chain is serial per problem, 32 problems per
//                              warp).  A problem that does not converge is put on a retry list.
//   C  et_vectors_kernel<NT,N> one THREAD per eigenpair, floor(128 / K) problems per CTA: the eigenvector of
//                              the tridiagonal by twisted factorisation (forward and backward pivots, twist at
//                              the smallest |gamma|; O(K) per vector, reciprocals instead of divisions), the
//                              reflectors applied to it from shared memory (broadcast loads), then
//                              v_j = sum_i D0_i w_ij^2 / lambda_j accumulated at the eigenvalue's rank
//                              (descending; or at an injected position, SURVEY.md 8(c) mode (iii)).
//   R  et_retry_kernel         problems on the retry list are redone with the shared-memory Jacobi solver.
//
// Measured against numpy's eigh on factor covariances of regression panels the ratios agree to 1e-13
// (tests/test_risk_gpu.py pins the end result against the reference to 1e-9).
#pragma once
#include "bmfm_common.cuh"
#include "jacobi.cuh"

namespace bmfm {

constexpr int ET_MAX_NT = 6;          // K <= 48
constexpr int ET_WARPS = 4;           // warps (= problems in flight) per CTA of kernel A
constexpr int ET_CTHREADS = 128;      // threads per CTA of kernel C
constexpr int ET_RETRY_CAP = 1 << 16;

struct EtBatch {
  int K, LD;               // LD = 8 * NT
  int nsim;                // simulations in this call
  int m0, nm;              // simulations [m0, m0 + nm) of the call are in this batch
  int t0, nd;              // dates [t0, t0 + nd)
  int P;                   // nd * nm problems; problem p = dt * nm + (m - m0)
  const double* C;         // [nsim][K][K]
  const double* D0;        // [T][K]
  const int* status;       // [T]
  double* de;              // [2 * LD][P]   d then e, problem-minor
  double* lam;             // [LD][P]       eigenvalues (unordered), problem-minor
  double* V;               // [P][VSZ]      reflector k packed by phase: entries 8 * (k / 8) .. LD - 1 (see et_voff)
  double* hinv;            // [P][LD]       1 / h_k (0: identity reflector)
  int* retry;              // [0] = count, then ET_RETRY_CAP entries (t * nsim + m)
  int fail_every;          // test hook: treat every fail_every-th problem as not converged (0 = off)
};

// reciprocal to ~1 ulp without the division's special-case path: hardware seed + two Newton steps.  The callers
// keep x away from 0, inf and the denormal range.
__device__ __forceinline__ double et_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}

// Reflector k is stored from index 8 * TK (TK = k / 8, the first row of its phase's tiles) to LD - 1:
// doubles per problem and the offset of v_k[0] (an address such that v_k[i] is at et_voff + i).
__host__ __device__ constexpr int et_vsz(int NT) { return 32 * NT * (NT + 1); }
__host__ __device__ constexpr int et_vrow(int NT, int k) {
  // start of row k's storage
  return 64 * ((k / 8) * NT - (k / 8) * ((k / 8) - 1) / 2) + (k % 8) * (8 * NT - 8 * (k / 8));
}
__host__ __device__ constexpr int et_voff(int NT, int k) { return et_vrow(NT, k) - 8 * (k / 8); }

// ---------------------------------------------------------------------------------------------------------
// A: Householder tridiagonalisation, matrix in registers
// ---------------------------------------------------------------------------------------------------------
template <int NT>
struct EtTiles {
  double a[NT][NT][2];
};

// one Householder step: column k = 8 * TK + 2 * kp + EK (EK compile-time so that the tile element is a register)
template <int NT, int TK, int EK>
__device__ __forceinline__ void et_step(EtTiles<NT>& M, int kp, int n, int lane, double* sx, double* sw, double* Vp,
                                        double* hinv_p, double* e_out /* smem [LD] */) {
  const int g = lane >> 2, c = lane & 3;
  const int kk = 2 * kp + EK, k = 8 * TK + kk;
  double* vout = Vp + et_voff(NT, 8 * TK) + kk * (8 * NT - 8 * TK);
  // 1. column k below the diagonal -> shared memory (rows <= k as zeros)
  if (c == kp) {
#pragma unroll
    for (int ti = TK; ti < NT; ++ti) sx[8 * ti + g] = (8 * ti + g > k) ? M.a[ti][TK][EK] : 0.0;
  }
  __syncwarp();
  double vr[NT], vc[NT][2];
  double s = 0.0;
#pragma unroll
  for (int ti = TK; ti < NT; ++ti) {
    vr[ti] = sx[8 * ti + g];
    const double2 t2 = *reinterpret_cast<const double2*>(sx + 8 * ti + 2 * c);
    vc[ti][0] = t2.x; vc[ti][1] = t2.y;
    s = fma(vr[ti], vr[ti], s);
  }
  const double alpha = sx[k + 1];
  if (k == n - 2) {          // the last sub-diagonal element is never eliminated: it only has to be read out
    if (lane == 0) e_out[k] = alpha;
    __syncwarp();
    return;
  }
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  s += __shfl_xor_sync(0xffffffffu, s, 8);
  s += __shfl_xor_sync(0xffffffffu, s, 16);
  const double sigma = s - alpha * alpha;
  if (!(sigma > 0.0)) {     // nothing below the sub-diagonal: identity reflector
    if (lane == 0) { e_out[k] = alpha; hinv_p[k] = 0.0; }
    if (c == 0) {
#pragma unroll
      for (int ti = TK; ti < NT; ++ti) vout[8 * ti + g] = 0.0;
    }
    __syncwarp();
    return;
  }
  const double rn = rsqrt(s);
  double norm = s * rn;
  norm = fma(fma(-norm, norm, s), 0.5 * rn, norm);       // one Newton step: sqrt(s) to full precision
  const double beta = alpha >= 0.0 ? -norm : norm;
  const double v0 = alpha - beta;
  const double h = fma(fabs(alpha), norm, s);            // v'v / 2
  const double ih = et_rcp(h);
  if (lane == 0) { e_out[k] = beta; hinv_p[k] = ih; }
  // the leading entry of v: index k + 1 lies in tile row TK or TK + 1
#pragma unroll
  for (int ti = TK; ti < NT && ti <= TK + 1; ++ti) {
    if (8 * ti + g == k + 1) vr[ti] = v0;
    if (8 * ti + 2 * c == k + 1) vc[ti][0] = v0;
    if (8 * ti + 2 * c + 1 == k + 1) vc[ti][1] = v0;
  }
  // the reflector goes out in the row layout (lanes c == 0: 8 consecutive doubles per tile row)
  if (c == 0) {
#pragma unroll
    for (int ti = TK; ti < NT; ++ti) vout[8 * ti + g] = vr[ti];
  }
  // 2. p = S v / h   (rows <= k masked), K = v'p / (2h), w = p - K v
  double pr[NT];
  double vp = 0.0;
#pragma unroll
  for (int ti = TK; ti < NT; ++ti) {
    double acc = 0.0;
#pragma unroll
    for (int tj = TK; tj < NT; ++tj) {
      acc = fma(M.a[ti][tj][0], vc[tj][0], acc);
      acc = fma(M.a[ti][tj][1], vc[tj][1], acc);
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    pr[ti] = (8 * ti + g > k) ? acc * ih : 0.0;
    vp = fma(vr[ti], pr[ti], vp);
  }
  vp += __shfl_xor_sync(0xffffffffu, vp, 4);
  vp += __shfl_xor_sync(0xffffffffu, vp, 8);
  vp += __shfl_xor_sync(0xffffffffu, vp, 16);
  const double kc = 0.5 * vp * ih;
  double wr[NT], wc[NT][2];
#pragma unroll
  for (int ti = TK; ti < NT; ++ti) {
    wr[ti] = fma(-kc, vr[ti], pr[ti]);
    if (c == 0) sw[8 * ti + g] = wr[ti];
  }
  __syncwarp();
#pragma unroll
  for (int tj = TK; tj < NT; ++tj) {
    const double2 t2 = *reinterpret_cast<const double2*>(sw + 8 * tj + 2 * c);
    wc[tj][0] = t2.x; wc[tj][1] = t2.y;
  }
  // 3. S -= v w' + w v'
#pragma unroll
  for (int ti = TK; ti < NT; ++ti) {
#pragma unroll
    for (int tj = TK; tj < NT; ++tj) {
      M.a[ti][tj][0] = fma(-vr[ti], wc[tj][0], fma(-wr[ti], vc[tj][0], M.a[ti][tj][0]));
      M.a[ti][tj][1] = fma(-vr[ti], wc[tj][1], fma(-wr[ti], vc[tj][1], M.a[ti][tj][1]));
    }
  }
  __syncwarp();     // sx / sw are rewritten by the next step
}

// one phase: columns k = 8 * TK .. 8 * TK + 7 (those <= n - 2; column n - 2 is only read out)
template <int NT, int TK>
__device__ __forceinline__ void et_phase(EtTiles<NT>& M, int n, int lane, double* sx, double* sw, double* Vp,
                                         double* hinv_p, double* e_out) {
#pragma unroll 1
  for (int kp = 0; kp < 4; ++kp) {
    const int k = 8 * TK + 2 * kp;
    if (k > n - 2) break;
    et_step<NT, TK, 0>(M, kp, n, lane, sx, sw, Vp, hinv_p, e_out);
    if (k + 1 > n - 2) break;
    et_step<NT, TK, 1>(M, kp, n, lane, sx, sw, Vp, hinv_p, e_out);
  }
}

template <int NT, int TK>
struct EtPhases {
  static __device__ __forceinline__ void run(EtTiles<NT>& M, int n, int lane, double* sx, double* sw, double* Vp,
                                             double* hinv_p, double* e_out) {
    if (8 * TK <= n - 2) {
      et_phase<NT, TK>(M, n, lane, sx, sw, Vp, hinv_p, e_out);
      EtPhases<NT, TK + 1>::run(M, n, lane, sx, sw, Vp, hinv_p, e_out);
    }
  }
};
template <int NT>
struct EtPhases<NT, NT> {
  static __device__ __forceinline__ void run(EtTiles<NT>&, int, int, double*, double*, double*, double*, double*) {}
};

template <int NT>
__global__ void __launch_bounds__(ET_WARPS * 32) et_tridiag_kernel(EtBatch b) {
  constexpr int LD = 8 * NT;
  __shared__ __align__(16) double s_buf[ET_WARPS][4 * LD];     // per warp: x, w, e, d
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, c = lane & 3;
  const int n = b.K;
  double* sx = s_buf[warp];
  double* sw = sx + LD;
  double* se = sw + LD;
  double* sd = se + LD;
  const bool vec2 = (n & 1) == 0;
  for (int p = blockIdx.x * ET_WARPS + warp; p < b.P; p += gridDim.x * ET_WARPS) {
    const int dt = p / b.nm, mi = p - dt * b.nm;
    const int t = b.t0 + dt, m = b.m0 + mi;
    if (b.status[t] != BMFM_ST_OK) continue;
    const double* D0 = b.D0 + (size_t)t * n;
    const double* Cm = b.C + (size_t)m * n * n;
    double sr[NT], sc[NT][2];
#pragma unroll
    for (int ti = 0; ti < NT; ++ti) {
      const int r = 8 * ti + g, q = 8 * ti + 2 * c;
      sr[ti] = r < n ? sqrt(D0[r]) : 0.0;
      sc[ti][0] = q < n ? sqrt(D0[q]) : 0.0;
      sc[ti][1] = q + 1 < n ? sqrt(D0[q + 1]) : 0.0;
    }
    EtTiles<NT> M;
#pragma unroll
    for (int ti = 0; ti < NT; ++ti) {
      const int r = 8 * ti + g;
#pragma unroll
      for (int tj = 0; tj < NT; ++tj) {
        const int q = 8 * tj + 2 * c;
        double x0 = 0.0, x1 = 0.0;
        if (r < n) {
          if (vec2) {
            if (q < n) { const double2 t2 = *reinterpret_cast<const double2*>(Cm + (size_t)r * n + q); x0 = t2.x; x1 = t2.y; }
          } else {
            if (q < n) x0 = Cm[(size_t)r * n + q];
            if (q + 1 < n) x1 = Cm[(size_t)r * n + q + 1];
          }
        }
        M.a[ti][tj][0] = sr[ti] * x0 * sc[tj][0];
        M.a[ti][tj][1] = sr[ti] * x1 * sc[tj][1];
      }
    }
    double* Vp = b.V + (size_t)p * et_vsz(NT);
    double* hp = b.hinv + (size_t)p * LD;
    EtPhases<NT, 0>::run(M, n, lane, sx, sw, Vp, hp, se);
    // the diagonal
#pragma unroll
    for (int ti = 0; ti < NT; ++ti) {
      if (2 * c == g) sd[8 * ti + g] = M.a[ti][ti][0];
      if (2 * c + 1 == g) sd[8 * ti + g] = M.a[ti][ti][1];
    }
    __syncwarp();
    for (int k = lane; k < LD; k += 32) {
      b.de[(size_t)k * b.P + p] = k < n ? sd[k] : 0.0;
      b.de[(size_t)(LD + k) * b.P + p] = k < n - 1 ? se[k] : 0.0;
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------------
// B: implicit QL, eigenvalues only, one thread per problem; (d, e) in shared memory [k][thread]
// ---------------------------------------------------------------------------------------------------------
// Square-root-free QL (Pal-Walker-Kahan, the iteration of LAPACK's dsterf) on d and the SQUARED off-diagonal e2:
// eigenvalues only, two reciprocals per rotation and no square root inside the chain.
__device__ inline bool et_ql_values(double* d, double* e2, int st, int n) {
  const double eps2 = 2.220446049250313e-16 * 2.220446049250313e-16;
  int budget = 60 * n;
  int m = -1;          // end of the unreduced block that starts at l (first negligible off-diagonal), -1: not known
  for (int l = 0; l < n;) {
    if (m < l) {       // a scan per converged eigenvalue; between the sweeps of one block the split is tracked below
      for (m = l; m < n - 1; ++m) {
        const double dd = fabs(d[m * st]) + fabs(d[(m + 1) * st]);
        if (e2[m * st] <= eps2 * dd * dd) break;
      }
    }
    if (m < n - 1) e2[m * st] = 0.0;
    if (m == l) { ++l; continue; }
    double p = d[l * st];
    if (m == l + 1) {                 // 2 x 2 block: closed form (dlae2)
      const double a = p, c2 = d[(l + 1) * st], b2 = e2[l * st];
      const double sm = a + c2, df = a - c2;
      const double rt = sqrt(fma(df, df, 4.0 * b2));
      double rt1, rt2;
      const double amx = fabs(a) > fabs(c2) ? a : c2, amn = fabs(a) > fabs(c2) ? c2 : a;
      if (sm < 0.0) { rt1 = 0.5 * (sm - rt); rt2 = (amx / rt1) * amn - (b2 / rt1); }
      else if (sm > 0.0) { rt1 = 0.5 * (sm + rt); rt2 = (amx / rt1) * amn - (b2 / rt1); }
      else { rt1 = 0.5 * rt; rt2 = -0.5 * rt; }
      d[l * st] = rt1;
      d[(l + 1) * st] = rt2;
      e2[l * st] = 0.0;
      l += 2;
      continue;
    }
    if (--budget < 0) return false;
    // shift
    const double rte = sqrt(e2[l * st]);
    double sigma = (d[(l + 1) * st] - p) / (2.0 * rte);
    const double r0 = sqrt(fma(sigma, sigma, 1.0));
    sigma = p - rte / (sigma + (sigma >= 0.0 ? r0 : -r0));
    double c = 1.0, sn = 0.0;
    double gamma = d[m * st] - sigma;
    p = gamma * gamma;
    double* pd = d + (size_t)(m - 1) * st;
    double* pe = e2 + (size_t)(m - 1) * st;
    double bb = *pe, alpha = *pd;
    double d_up = 0.0;                // final d[i + 2] of the previous step (for the split test)
    int msplit = -1;
    for (int i = m - 1; i >= l; --i, pd -= st, pe -= st) {
      const double bb_nx = i > l ? pe[-st] : 0.0, alpha_nx = i > l ? pd[-st] : 0.0;
      const double r = p + bb;
      const double e_new = sn * r;                // e2[i + 1]
      const double oldc = c;
      const double ir = r > 0.0 ? et_rcp(r) : 0.0;
      c = p * ir;
      sn = bb * ir;
      const double oldgam = gamma;
      gamma = c * (alpha - sigma) - sn * oldgam;
      const double d_new = oldgam + (alpha - gamma);     // d[i + 1]
      pd[st] = d_new;
      if (i != m - 1) {
        pe[st] = e_new;
        const double dd = fabs(d_new) + fabs(d_up);
        if (e_new <= eps2 * dd * dd) msplit = i + 1;      // i decreases: the last hit is the smallest index
      }
      d_up = d_new;
      p = c != 0.0 ? gamma * gamma * et_rcp(c) : oldc * bb;
      bb = bb_nx;
      alpha = alpha_nx;
    }
    {
      const double e_new = sn * p, d_new = sigma + gamma;
      e2[l * st] = e_new;
      d[l * st] = d_new;
      const double dd = fabs(d_new) + fabs(d_up);
      if (e_new <= eps2 * dd * dd) msplit = l;
    }
    if (msplit >= 0) m = msplit;
  }
  return true;
}

constexpr int ET_BTHREADS = 64;

__global__ void __launch_bounds__(ET_BTHREADS) et_chain_kernel(EtBatch b) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= b.P) return;
  const int n = b.K, st = blockDim.x, LD = b.LD;
  const int dt = p / b.nm, mi = p - dt * b.nm;
  const int t = b.t0 + dt;
  if (b.status[t] != BMFM_ST_OK) return;
  double* d = reinterpret_cast<double*>(smem) + threadIdx.x;
  double* e = d + (size_t)n * st;
  for (int k = 0; k < n; ++k) {
    d[k * st] = b.de[(size_t)k * b.P + p];
    const double ek = b.de[(size_t)(LD + k) * b.P + p];
    e[k * st] = ek * ek;
  }
  bool ok = et_ql_values(d, e, st, n);
  if (b.fail_every > 0 && (p % b.fail_every) == 0) ok = false;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int k = 0; k < n; ++k) b.lam[(size_t)k * b.P + p] = ok ? d[k * st] : nan;
  if (!ok) {
    const int slot = atomicAdd(b.retry, 1);
    if (slot < ET_RETRY_CAP) b.retry[1 + slot] = t * b.nsim + (b.m0 + mi);
  }
}

// ---------------------------------------------------------------------------------------------------------
// C: eigenvectors by twisted factorisation + reflectors + ratio; one thread per eigenpair
// ---------------------------------------------------------------------------------------------------------
// Shared memory per problem slot: d[LD] e[LD] d0[LD] lam[LD] hinv[LD] ratio[LD] rank(int)[LD] V[VSZ]
template <int NT>
__host__ __device__ constexpr int et_slot_doubles() { return 8 * NT * 7 + et_vsz(NT); }

// N = the factor count as a compile-time constant (the common ones are instantiated), or 0: run-time count,
// loops padded to LD
template <int NT, int N, int MINB>
__global__ void __launch_bounds__(ET_CTHREADS, MINB) et_vectors_kernel(EtBatch b, int nblk, const int* perm, double* partial) {
  constexpr int LD = 8 * NT;
  constexpr int NL = N > 0 ? N : LD;                  // trip count of the unrolled loops
  constexpr int SLOT = et_slot_doubles<NT>() + 2;     // +2 doubles: shifts the next slot by two banks
  extern __shared__ __align__(16) unsigned char smem[];
  const int n = N > 0 ? N : b.K;
  const int G = ET_CTHREADS / n;                       // problems per round (>= 2 for K <= 48)
  const int tid = threadIdx.x;
  const int slot = tid / n, j = tid - slot * n;        // threads >= G * n idle in the per-eigenpair part
  const bool active = slot < G;
  // two sets of G slots: while a round is computed from one set the next round's problems arrive in the other
  // (cp.async), so the ~12 KB per problem that come from L2 / HBM are not waited for
  double* S0 = reinterpret_cast<double*>(smem);
  double* vacc = S0 + (size_t)2 * G * SLOT;            // [G][LD] accumulators
  const int per = (b.nm + nblk - 1) / nblk;
  for (int w = blockIdx.x; w < b.nd * nblk; w += gridDim.x) {
    const int dt = w / nblk, blk = w - dt * nblk, t = b.t0 + dt;
    double* out = partial + ((size_t)dt * nblk + blk) * n;      // partial sums are per batch
    if (b.status[t] != BMFM_ST_OK) continue;                   // (never read for an invalid date)
    __syncthreads();
    for (int e = tid; e < G * LD; e += ET_CTHREADS) vacc[e] = 0.0;
    const int mi_lo = blk * per, mi_hi = min(b.nm, mi_lo + per);
    auto stage = [&](int mi0, int buf) {
      // d, e, lam, hinv, V of the round's problems and D0 of the date -> slot set `buf` (asynchronous copies)
      for (int sl = 0; sl < G; ++sl) {
        const int mi = mi0 + sl;
        if (mi >= mi_hi) break;
        const int p = dt * b.nm + mi;
        double* dst = S0 + (size_t)(buf * G + sl) * SLOT;
        for (int k = tid; k < LD; k += ET_CTHREADS) {
          cp_async_8(dst + k, b.de + (size_t)k * b.P + p);
          cp_async_8(dst + LD + k, b.de + (size_t)(LD + k) * b.P + p);
          if (k < n) cp_async_8(dst + 2 * LD + k, b.D0 + (size_t)t * n + k);
          else dst[2 * LD + k] = 0.0;
          cp_async_8(dst + 3 * LD + k, b.lam + (size_t)k * b.P + p);
          cp_async_8(dst + 4 * LD + k, b.hinv + (size_t)p * LD + k);
        }
        const double* src = b.V + (size_t)p * et_vsz(NT);
        double* dv = dst + 7 * LD;
        const int nv = n >= 3 ? et_vrow(NT, n - 2) / 2 : 0;         // reflector rows 0 .. n-3, in 16-byte pieces
        for (int e = tid; e < nv; e += ET_CTHREADS) cp_async_16(dv + 2 * e, src + 2 * e);
      }
      asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    int buf = 0;
    stage(mi_lo, 0);
    for (int mi0 = mi_lo; mi0 < mi_hi; mi0 += G, buf ^= 1) {
      if (mi0 + G < mi_hi) {
        stage(mi0 + G, buf ^ 1);                         // the other set was released by the barrier that ended the last round
        asm volatile("cp.async.wait_group 1;\n" ::: "memory");
      } else {
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
      }
      __syncthreads();
      double* my = S0 + (size_t)(buf * G + (active ? slot : 0)) * SLOT;
      const double* s_d = my;
      const double* s_e = s_d + LD;
      const double* s_d0 = s_e + LD;
      const double* s_lam = s_d0 + LD;
      const double* s_hinv = s_lam + LD;
      double* s_ratio = my + 5 * LD;
      int* s_rank = reinterpret_cast<int*>(s_ratio + LD);   // LD ints
      const double* s_V = s_ratio + 2 * LD;
      const int mi = mi0 + slot;
      const bool work = active && mi < mi_hi;
      double ratio = 0.0, lam = 0.0;
      if (work) {
        lam = s_lam[j];
        // Twisted factorisation of T - lam I.  z[i] first holds 1 / D+_i (forward pivots); the backward pivots
        // D-_i are streamed once to find the twist r = argmin |gamma_i| and a second time to overwrite z[i],
        // i > r, with 1 / D-_i — one array instead of two, for one more pass of cheap reciprocals.
        double z[NL];
        double nrm = 0.0;
#pragma unroll
        for (int i = 0; i < NL; ++i) nrm = fmax(nrm, fmax(fabs(s_d[i]), fabs(s_e[i])));
        const double tiny = fmax(nrm * 4.9e-32, 1e-290);          // pivot floor: eps^2 |T|
        double dp = s_d[0] - lam;
#pragma unroll
        for (int i = 0; i < NL - 1; ++i) {
          if (N > 0 || i < n - 1) {
            if (fabs(dp) < tiny) dp = dp < 0.0 ? -tiny : tiny;
            const double r = et_rcp(dp);
            z[i] = r;
            const double ei = s_e[i];
            dp = fma(-ei * ei, r, s_d[i + 1] - lam);
          }
        }
        // gamma_i = (d_i - lam) - e_{i-1}^2 / D+_{i-1} - e_i^2 / D-_{i+1};  at the last row it is D+_{n-1} = dp
        double gbest = fabs(dp);
        int rbest = n - 1;
        {
          double dm = s_d[n - 1] - lam;
#pragma unroll
          for (int i = NL - 2; i >= 0; --i) {
            if (N > 0 || i < n - 1) {
              if (fabs(dm) < tiny) dm = dm < 0.0 ? -tiny : tiny;
              const double r = et_rcp(dm);
              const double ei = s_e[i];
              dm = fma(-ei * ei, r, s_d[i] - lam);               // D-_i
              double gam = dm;
              if (i > 0) { const double em = s_e[i - 1]; gam = fma(-em * em, z[i - 1], gam); }
              if (fabs(gam) < gbest) { gbest = fabs(gam); rbest = i; }
            }
          }
        }
        // second backward pass: z_{i+1} = -e_i / D-_{i+1} z_i above the twist needs 1 / D-_{i+1} for i >= r
        {
          double dm = s_d[n - 1] - lam;
#pragma unroll
          for (int i = NL - 2; i >= 0; --i) {
            if (N > 0 || i < n - 1) {
              if (fabs(dm) < tiny) dm = dm < 0.0 ? -tiny : tiny;
              const double r = et_rcp(dm);
              if (i + 1 > rbest) z[i + 1] = r;                   // 1 / D-_{i+1}
              const double ei = s_e[i];
              dm = fma(-ei * ei, r, s_d[i] - lam);
            }
          }
        }
        // z: 1 at the twist, z_i = -e_i / D+_i z_{i+1} below it, z_{i+1} = -e_i / D-_{i+1} z_i above it
        {
          double cur = 1.0;
#pragma unroll
          for (int i = NL - 2; i >= 0; --i) {
            if (N > 0 || i < n - 1) {
              const double nxt = (i + 1 == rbest) ? 1.0 : cur;
              cur = -s_e[i] * z[i] * nxt;
              if (i < rbest) z[i] = cur;
            }
          }
          cur = 1.0;
#pragma unroll
          for (int i = 0; i < NL - 1; ++i) {
            if (N > 0 || i < n - 1) {
              const double prv = (i == rbest) ? 1.0 : cur;
              cur = -s_e[i] * z[i + 1] * prv;
              if (i + 1 > rbest) z[i + 1] = cur;
            }
          }
#pragma unroll
          for (int i = 0; i < NL; ++i) {
            if (i == rbest) z[i] = 1.0;
            if (N == 0 && i >= n) z[i] = 0.0;
          }
        }
        // keep the numbers in range (the ratio below divides by sum w^2, so the scale itself drops out)
        {
          double nz = 0.0;
#pragma unroll
          for (int i = 0; i < NL; ++i) nz = fma(z[i], z[i], nz);
          const double sc = rsqrt(nz);
#pragma unroll
          for (int i = 0; i < NL; ++i) z[i] *= sc;
        }
        // reflectors: w = H_0 H_1 ... H_{n-3} z, H_k = I - v_k v_k' / h_k, v_k in row k of V (entries i > k)
#pragma unroll
        for (int k = NL - 3; k >= 0; --k) {
          if (N > 0 || k <= n - 3) {
            const double* vk = s_V + et_voff(NT, k);
            double dot = 0.0;
#pragma unroll
            for (int i = (k + 1) & ~1; i < NL; i += 2) {
              const double2 v2 = *reinterpret_cast<const double2*>(vk + i);
              if (i > k) dot = fma(v2.x, z[i], dot);
              if (i + 1 < NL) dot = fma(v2.y, z[i + 1], dot);
            }
            const double f = dot * s_hinv[k];
#pragma unroll
            for (int i = (k + 1) & ~1; i < NL; i += 2) {
              const double2 v2 = *reinterpret_cast<const double2*>(vk + i);
              if (i > k) z[i] = fma(-f, v2.x, z[i]);
              if (i + 1 < NL) z[i + 1] = fma(-f, v2.y, z[i + 1]);
            }
          }
        }
        double num = 0.0, den = 0.0;
#pragma unroll
        for (int i = 0; i < NL; ++i) {
          const double zz = z[i] * z[i];
          num = fma(s_d0[i], zz, num);
          den += zz;
        }
        ratio = num / (den * lam);
      }
      // rank of every eigenvalue of the problem (descending, ties by index), then accumulate by position
      if (work) {
        int r = 0;
        for (int i = 0; i < n; ++i) {
          const double o = s_lam[i];
          r += (o > lam) || (o == lam && i < j);
        }
        if (isnan(lam)) { r = j; ratio = 0.0; }       // not converged: the retry kernel adds this problem
        int pos = r;
        if (perm) pos = perm[((size_t)t * b.nsim + (b.m0 + mi)) * n + r];
        s_rank[j] = pos;
        s_ratio[j] = ratio;
      }
      __syncthreads();
      if (work) {
        // thread j collects position j's contribution (a permutation: exactly one source)
        for (int i = 0; i < n; ++i)
          if (s_rank[i] == j) vacc[slot * LD + j] += s_ratio[i];
      }
      __syncthreads();          // the round's slot set may be refilled
    }
    __syncthreads();
    for (int k = tid; k < n; k += ET_CTHREADS) {
      double sacc = 0.0;
      for (int q = 0; q < G; ++q) sacc += vacc[q * LD + k];
      out[k] = sacc;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// R: problems whose QL iteration did not converge, redone by the shared-memory Jacobi solver (one CTA each)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) et_retry_kernel(int K, int nsim, const double* C, const double* D0,
                                                       const int* retry, const int* perm, double* vsum) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int ld = K | 1, tid = threadIdx.x, nt = blockDim.x;
  double* A = reinterpret_cast<double*>(smem);
  double* V = A + K * ld;
  double* cs = V + K * ld;          // K + 2
  double* d0 = cs + K + 2;          // K
  double* lam = d0 + K;             // K
  int* flag = reinterpret_cast<int*>(lam + K);   // 8
  int cnt = retry[0];
  if (cnt > ET_RETRY_CAP) cnt = ET_RETRY_CAP;
  for (int q = blockIdx.x; q < cnt; q += gridDim.x) {
    const int id = retry[1 + q], t = id / nsim, m = id - t * nsim;
    __syncthreads();
    if (tid < K) d0[tid] = D0[(size_t)t * K + tid];
    __syncthreads();
    const double* Cm = C + (size_t)m * K * K;
    for (int e = tid; e < K * K; e += nt) {
      const int a = e / K, bcol = e - a * K;
      A[a * ld + bcol] = sqrt(d0[a]) * Cm[e] * sqrt(d0[bcol]);
    }
    __syncthreads();
    jacobi_eig_cta(A, V, K, ld, cs, flag);
    __syncthreads();
    if (tid < K) lam[tid] = A[tid * ld + tid];
    __syncthreads();
    if (tid < K) {
      const double me = lam[tid];
      int r = 0;
      for (int i = 0; i < K; ++i) r += (lam[i] > me) || (lam[i] == me && i < tid);
      double num = 0.0;
      for (int i = 0; i < K; ++i) {
        const double v = V[i * ld + tid];
        num = fma(d0[i] * v, v, num);
      }
      const int pos = perm ? perm[((size_t)t * nsim + m) * K + r] : r;
      atomicAdd(&vsum[(size_t)t * K + pos], num / me);
    }
  }
}

}  // namespace bmfm
