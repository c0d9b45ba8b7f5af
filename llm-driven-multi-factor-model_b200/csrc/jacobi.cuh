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
