// CTA-cooperative symmetric eigensolver in shared memory: two-sided Jacobi with
// round-robin (tournament) ordering, so the n/2 rotations of one round touch
// disjoint row/column pairs and run in parallel across the CTA.  fp64.
//
// Used where the reference calls LAPACK on a small dense symmetric matrix:
//   np.linalg.pinv of the transformed Gram   (Barra-master/mfm/CrossSection.py:76,98)
//   np.linalg.eig of factor covariances      (Barra-master/mfm/utils.py:64,79,105)
// Jacobi is chosen over tridiagonal-QR because every step is a batch of
// independent 2x2 rotations (no serial bulge chase) and it is accurate in the
// relative sense on the graded matrices this path produces.
#pragma once
#include "bmfm_common.cuh"

namespace bmfm {

// circle-method pairing: m even, round r in [0, m-1), slot i in [0, m/2)
__device__ __forceinline__ void rr_pair(int m, int r, int i, int& p, int& q) {
  const int mm = m - 1;
  if (i == 0) {
    p = mm;
    q = r;
  } else {
    p = r + i;
    if (p >= mm) p -= mm;
    q = r - i;
    if (q < 0) q += mm;
  }
  if (p > q) {
    const int t = p;
    p = q;
    q = t;
  }
}

// A: n x n symmetric, row stride ld (choose ld odd to keep column walks free of
// bank conflicts).  V: n x n, same stride, receives the eigenvectors as COLUMNS.
// cs: >= 2*(n/2+1) doubles scratch, flag: 2 ints scratch.  All threads of the
// CTA must call.  On return diag(A) holds the eigenvalues (unsorted).
// Returns the number of sweeps used.
__device__ inline int jacobi_eig_cta(double* A, double* V, int n, int ld, double* cs, int* flag,
                                     int max_sweeps = 30) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int m = (n + 1) & ~1;   // even number of players; player n (if any) is a bye
  const int half = m >> 1;
  for (int e = tid; e < n * n; e += nt) {
    const int i = e / n, j = e - i * n;
    V[i * ld + j] = (i == j) ? 1.0 : 0.0;
  }
  __syncthreads();
  int sweep = 0;
  for (; sweep < max_sweeps; ++sweep) {
    if (tid == 0) flag[0] = 0;
    __syncthreads();
    for (int r = 0; r < m - 1; ++r) {
      // phase 1: rotation parameters for the round's disjoint pairs
      if (tid < half) {
        int p, q;
        rr_pair(m, r, tid, p, q);
        double c = 1.0, s = 0.0;
        if (q < n) {
          const double apq = A[p * ld + q];
          const double app = A[p * ld + p], aqq = A[q * ld + q];
          const double mag = fabs(apq);
          if (mag > 1.1102230246251565e-16 * sqrt(fabs(app * aqq)) && mag > 1e-290) {
            const double tau = (aqq - app) / (2.0 * apq);
            const double t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
            c = 1.0 / sqrt(1.0 + t * t);
            s = t * c;
            atomicAdd(&flag[0], 1);
          }
        }
        cs[2 * tid] = c;
        cs[2 * tid + 1] = s;
      }
      __syncthreads();
      // phase 2: column rotations of A and V (k runs fastest across lanes)
      for (int e = tid; e < half * n; e += nt) {
        const int i = e / n, k = e - i * n;
        const double s = cs[2 * i + 1];
        if (s != 0.0) {
          const double c = cs[2 * i];
          int p, q;
          rr_pair(m, r, i, p, q);
          const double akp = A[k * ld + p], akq = A[k * ld + q];
          A[k * ld + p] = c * akp - s * akq;
          A[k * ld + q] = s * akp + c * akq;
          const double vkp = V[k * ld + p], vkq = V[k * ld + q];
          V[k * ld + p] = c * vkp - s * vkq;
          V[k * ld + q] = s * vkp + c * vkq;
        }
      }
      __syncthreads();
      // phase 3: row rotations of A
      for (int e = tid; e < half * n; e += nt) {
        const int i = e / n, k = e - i * n;
        const double s = cs[2 * i + 1];
        if (s != 0.0) {
          const double c = cs[2 * i];
          int p, q;
          rr_pair(m, r, i, p, q);
          const double apk = A[p * ld + k], aqk = A[q * ld + k];
          A[p * ld + k] = c * apk - s * aqk;
          A[q * ld + k] = s * apk + c * aqk;
        }
      }
      __syncthreads();
    }
    const int rotated = flag[0];
    __syncthreads();
    if (rotated == 0) break;
  }
  return sweep;
}

}  // namespace bmfm

// ---------------------------------------------------------------------------------------------
// Block form of the same parallel Jacobi, used by the Monte-Carlo hot loop (eigen_mc.cu).
// With the round's disjoint pairs the update A <- J'AJ factors into independent 2 x 2 blocks
// B_ab <- J_a' B_ab J_b (rows from pair a, columns from pair b).  Only blocks with a < b are
// computed and mirrored (symmetry halves the work and removes the separate row phase);
// diagonal blocks use the closed form app -= t apq, aqq += t apq, apq = 0.  Pair and block
// index tables are built once per CTA, so the inner loops have no integer division.
// N > 0: compile-time matrix order (even); N == 0: run-time even order `n_rt`.
// ---------------------------------------------------------------------------------------------
namespace bmfm {

struct JacobiTables {
  unsigned short* pair;   // [(m-1)][half]  p | q << 8  (p < q)
  unsigned short* blk;    // [half*(half-1)/2]  a | b << 8  (a < b)
};

__device__ inline void jacobi_build_tables(int n, JacobiTables tb) {
  const int half = n >> 1;
  for (int e = threadIdx.x; e < (n - 1) * half; e += blockDim.x) {
    const int r = e / half, i = e - r * half;
    int p, q;
    rr_pair(n, r, i, p, q);
    tb.pair[e] = (unsigned short)(p | (q << 8));
  }
  int idx = 0;
  for (int a = 0; a < half; ++a)
    for (int b = a + 1; b < half; ++b, ++idx)
      if (idx % blockDim.x == threadIdx.x) tb.blk[idx] = (unsigned short)(a | (b << 8));
  __syncthreads();
}

template <int N>
__device__ inline int jacobi_eig_blocks(double* A, double* V, int n_rt, int ld, double* cs, JacobiTables tb,
                                        int max_sweeps = 30) {
  const int n = N > 0 ? N : n_rt;
  const int half = n >> 1, nblk = half * (half - 1) / 2;
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int e = tid; e < n * n; e += nt) {
    const int i = e / n, j = e - i * n;
    V[i * ld + j] = (i == j) ? 1.0 : 0.0;
  }
  __syncthreads();
  int sweep = 0;
  for (; sweep < max_sweeps; ++sweep) {
    int rotated = 0;
    for (int r = 0; r < n - 1; ++r) {
      const unsigned short* pr = tb.pair + r * half;
      // phase 1: rotation of every pair + closed-form update of its diagonal block
      int rot = 0;
      if (tid < half) {
        const int p = pr[tid] & 0xff, q = pr[tid] >> 8;
        const double apq = A[p * ld + q];
        const double app = A[p * ld + p], aqq = A[q * ld + q];
        const double mag = fabs(apq);
        double c = 1.0, s = 0.0;
        if (mag > 1.1102230246251565e-16 * sqrt(fabs(app * aqq)) && mag > 1e-290) {
          const double tau = (aqq - app) / (2.0 * apq);
          const double t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
          c = 1.0 / sqrt(1.0 + t * t);
          s = t * c;
          A[p * ld + p] = app - t * apq;
          A[q * ld + q] = aqq + t * apq;
          A[p * ld + q] = 0.0;
          A[q * ld + p] = 0.0;
          rot = 1;
        }
        cs[2 * tid] = c;
        cs[2 * tid + 1] = s;
      }
      rotated += __syncthreads_count(rot);
      // phase 2a: off-diagonal blocks (a < b), mirrored
      for (int idx = tid; idx < nblk; idx += nt) {
        const int a = tb.blk[idx] & 0xff, b = tb.blk[idx] >> 8;
        const double sa = cs[2 * a + 1], sb = cs[2 * b + 1];
        if (sa == 0.0 && sb == 0.0) continue;
        const double ca = cs[2 * a], cb = cs[2 * b];
        const int pa = pr[a] & 0xff, qa = pr[a] >> 8, pb = pr[b] & 0xff, qb = pr[b] >> 8;
        const double m00 = A[pa * ld + pb], m01 = A[pa * ld + qb], m10 = A[qa * ld + pb], m11 = A[qa * ld + qb];
        const double t00 = ca * m00 - sa * m10, t01 = ca * m01 - sa * m11;
        const double t10 = sa * m00 + ca * m10, t11 = sa * m01 + ca * m11;
        const double n00 = cb * t00 - sb * t01, n01 = sb * t00 + cb * t01;
        const double n10 = cb * t10 - sb * t11, n11 = sb * t10 + cb * t11;
        A[pa * ld + pb] = n00; A[pb * ld + pa] = n00;
        A[pa * ld + qb] = n01; A[qb * ld + pa] = n01;
        A[qa * ld + pb] = n10; A[pb * ld + qa] = n10;
        A[qa * ld + qb] = n11; A[qb * ld + qa] = n11;
      }
      // phase 2b: eigenvector columns V <- V J (row k, pair i); k runs fastest across lanes
      for (int e = tid; e < half * n; e += nt) {
        const int i = e / n, k = e - i * n;      // compile-time n when N > 0
        const double s = cs[2 * i + 1];
        if (s == 0.0) continue;
        const double c = cs[2 * i];
        const int p = pr[i] & 0xff, q = pr[i] >> 8;
        const double vkp = V[k * ld + p], vkq = V[k * ld + q];
        V[k * ld + p] = c * vkp - s * vkq;
        V[k * ld + q] = s * vkp + c * vkq;
      }
      __syncthreads();
    }
    if (rotated == 0) break;
  }
  return sweep;
}

}  // namespace bmfm
