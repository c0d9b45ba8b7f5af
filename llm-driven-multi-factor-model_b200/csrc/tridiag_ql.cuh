// CTA-cooperative symmetric eigensolver: Householder tridiagonalisation + implicit QL.
//
// The Monte-Carlo loop of the eigenfactor adjustment (utils.py:71-84) needs T*M full
// eigendecompositions of K x K matrices.  Jacobi in shared memory makes ~330 read-modify-write
// passes over the matrix and its eigenvectors and is bound by shared-memory bandwidth
// (profiles/: 217 k wavefronts per 42 x 42 problem); the classic two-stage method touches the
// matrix ~40 times during the reduction and then only two columns of Z per QL rotation, about
// an eighth of the shared-memory traffic and a fifth of the instructions.
//   stage 1  tred2: A -> Q T Q', Q accumulated in place (A becomes Q)
//   stage 2  tql2 : per QL sweep one thread runs the (serial, Z-independent) chain of rotation
//            parameters, then every thread applies the whole sweep to its own row of Z with no
//            barrier in between (row k only ever touches Z[k][*]).
// Many problems per SM hide the serial chain: one CTA = 64 threads = one problem, ~16 KB.
#pragma once
#include "bmfm_common.cuh"

namespace bmfm {

// sum over the CTA (one or two warps; with one warp no shared memory or barrier is needed)
__device__ __forceinline__ double block_sum64(double v, double* red) {
  v = warp_sum(v);
  if (blockDim.x <= 32) return v;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  return red[0] + red[1];
}

// stage 1 only: on exit Z = Q (accumulated Householder transform), d = diagonal, e[0..n-2] =
// sub-diagonal (already shifted for the QL stage), e[n-1] = 0.
// Works for blockDim.x = 32 or 64 (n <= 64): thread t owns rows t, t + blockDim.x.  One warp per
// problem is the efficient shape: with two warps the second one carries n - 32 live lanes through
// every loop, and every step pays CTA barriers instead of warp barriers.
__device__ inline void tred2_cta(double* Z, double* d, double* e, double* red, int n, int ld) {
  const int tid = threadIdx.x, nt = blockDim.x;
  // ---- stage 1: Householder reduction (EISPACK tred2, lower triangle), parallel over rows j ----
  for (int i = n - 1; i >= 1; --i) {
    const int l = i - 1;
    double h = 0.0;
    const double* zi = Z + i * ld;
    if (l > 0) {
      double sc = 0.0;
      for (int k = tid; k <= l; k += nt) sc += fabs(zi[k]);
      const double scale = block_sum64(sc, red);
      if (scale == 0.0) {
        if (tid == 0) e[i] = zi[l];
      } else {
        double hs = 0.0;
        for (int k = tid; k <= l; k += nt) {
          const double v = zi[k] / scale;
          Z[i * ld + k] = v;
          hs = fma(v, v, hs);
        }
        h = block_sum64(hs, red);      // (its barriers also publish the scaled row)
        if (nt <= 32) __syncwarp();
        const double f0 = zi[l];
        const double g0 = f0 >= 0.0 ? -sqrt(h) : sqrt(h);
        h -= f0 * g0;
        __syncthreads();
        if (tid == 0) {
          e[i] = scale * g0;
          Z[i * ld + l] = f0 - g0;
        }
        __syncthreads();
        // p = A u / h  (row j up to the diagonal, column j below it)
        double fsum = 0.0;
        double ejv[2] = {0.0, 0.0};
        int jj = 0;
        for (int j = tid; j <= l; j += nt, ++jj) {
          const double uj = zi[j];
          Z[j * ld + i] = uj / h;
          double g0a = 0.0, g1a = 0.0;
          const double* zj = Z + j * ld;
          int k = 0;
          for (; k + 1 <= j; k += 2) { g0a = fma(zj[k], zi[k], g0a); g1a = fma(zj[k + 1], zi[k + 1], g1a); }
          for (; k <= j; ++k) g0a = fma(zj[k], zi[k], g0a);
          k = j + 1;
          for (; k + 1 <= l; k += 2) { g0a = fma(Z[k * ld + j], zi[k], g0a); g1a = fma(Z[(k + 1) * ld + j], zi[k + 1], g1a); }
          for (; k <= l; ++k) g0a = fma(Z[k * ld + j], zi[k], g0a);
          const double ej = (g0a + g1a) / h;
          ejv[jj] = ej;
          fsum = fma(ej, uj, fsum);
        }
        const double f = block_sum64(fsum, red);
        const double hh = f / (h + h);
        jj = 0;
        for (int j = tid; j <= l; j += nt, ++jj) e[j] = ejv[jj] - hh * zi[j];
        __syncthreads();
        // rank-2 update of the leading lower triangle
        for (int j = tid; j <= l; j += nt) {
          const double fu = zi[j], gq = e[j];
          double* zj = Z + j * ld;
          for (int k = 0; k <= j; ++k) zj[k] -= fma(fu, e[k], gq * zi[k]);
        }
        __syncthreads();
      }
    } else {
      if (tid == 0) e[i] = zi[l];
    }
    if (tid == 0) d[i] = h;
    __syncthreads();
  }
  if (tid == 0) { d[0] = 0.0; e[0] = 0.0; }
  __syncthreads();
  // accumulate the transformation: Z <- Q (thread = column j)
  for (int i = 0; i < n; ++i) {
    const int l = i - 1;
    if (d[i] != 0.0) {
      const double* zi = Z + i * ld;
      for (int j = tid; j <= l; j += nt) {
        double g0a = 0.0, g1a = 0.0;
        int k = 0;
        for (; k + 1 <= l; k += 2) { g0a = fma(zi[k], Z[k * ld + j], g0a); g1a = fma(zi[k + 1], Z[(k + 1) * ld + j], g1a); }
        for (; k <= l; ++k) g0a = fma(zi[k], Z[k * ld + j], g0a);
        const double g = g0a + g1a;
        for (k = 0; k <= l; ++k) Z[k * ld + j] = fma(-g, Z[k * ld + i], Z[k * ld + j]);
      }
    }
    __syncthreads();
    if (tid == 0) { d[i] = Z[i * ld + i]; Z[i * ld + i] = 1.0; }
    for (int j = tid; j <= l; j += nt) { Z[j * ld + i] = 0.0; Z[i * ld + j] = 0.0; }
    __syncthreads();
  }
  if (tid == 0) {
    for (int i = 1; i < n; ++i) e[i - 1] = e[i];
    e[n - 1] = 0.0;
  }
  __syncthreads();
}

__device__ inline int tridiag_ql_cta(double* Z, double* d, double* e, double* cs, double* red, int n, int ld, int dbg = 0) {
  const int tid = threadIdx.x;
  tred2_cta(Z, d, e, red, n, ld);
  if (dbg == 1) return 0;
  // ---- stage 2: implicit QL on (d, e) with the rotations of each sweep applied row-wise ----
  int* ictl = reinterpret_cast<int*>(red + 2);   // [0] = m, [1] = lowest rotation index applied, [2] = fail
  if (tid == 0) ictl[2] = 0;
  for (int l = 0; l < n; ++l) {
    for (int iter = 0;; ++iter) {
      if (tid == 0) {
        int m = l;
        for (; m < n - 1; ++m) {
          const double dd = fabs(d[m]) + fabs(d[m + 1]);
          if (fabs(e[m]) <= 2.220446049250313e-16 * dd) break;
        }
        int ilo = m;          // no rotation
        if (m != l) {
          if (iter >= 60) ictl[2] = 1;
          double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
          double r = sqrt(g * g + 1.0);
          g = d[m] - d[l] + e[l] / (g + (g >= 0.0 ? fabs(r) : -fabs(r)));
          double s = 1.0, c = 1.0, p = 0.0;
          int i = m - 1;
          bool brk = false;
          // the chain is one long dependency; the operands of the NEXT step are fetched a step
          // ahead and d[i+1] is carried in a register so no shared-memory latency sits on it
          double e_i = e[i], d_i = d[i], d_ip1 = d[i + 1];
          for (; i >= l; --i) {
            const double e_nx = i > l ? e[i - 1] : 0.0, d_nx = i > l ? d[i - 1] : 0.0;
            const double f = s * e_i;
            const double b = c * e_i;
            const double h2 = f * f + g * g;
            const double ir = rsqrt(h2);     // one reciprocal square root instead of a sqrt and two divisions
            r = h2 * ir;
            e[i + 1] = h2 > 0.0 ? r : 0.0;
            if (!(h2 > 0.0)) {
              d[i + 1] = d_ip1 - p;
              e[m] = 0.0;
              brk = true;
              break;
            }
            s = f * ir;
            c = g * ir;
            g = d_ip1 - p;
            r = (d_i - g) * s + 2.0 * c * b;
            p = s * r;
            d[i + 1] = g + p;
            g = c * r - b;
            cs[2 * i] = c;
            cs[2 * i + 1] = s;
            d_ip1 = d_i;
            d_i = d_nx;
            e_i = e_nx;
          }
          ilo = brk ? i + 1 : l;
          if (!brk) {
            d[l] = d_ip1 - p;     // d_ip1 now holds the untouched d[l]
            e[l] = g;
            e[m] = 0.0;
          }
        }
        ictl[0] = m;
        ictl[1] = ilo;
      }
      __syncthreads();
      const int m = ictl[0], ilo = ictl[1], fail = ictl[2];
      if (dbg != 2 && m != l && ilo < m) {
        for (int row = tid; row < n; row += blockDim.x) {
          double* zr = Z + row * ld;
          double z1 = zr[m];
          for (int i = m - 1; i >= ilo; --i) {
            const double c = cs[2 * i], s = cs[2 * i + 1];
            const double zi = zr[i];
            zr[i + 1] = fma(s, zi, c * z1);
            z1 = fma(c, zi, -s * z1);
          }
          zr[ilo] = z1;
        }
      }
      __syncthreads();
      if (m == l || fail) break;
    }
    if (ictl[2]) return 1;
  }
  return 0;
}


// ---- stage 2, split form: one THREAD runs the whole QL iteration of one problem on private
// copies of (d, e) and records every sweep — header (m | ilo << 8) and the (c, s) of its rotations
// in application order — so that the eigenvector update can be replayed later by a CTA.  The
// rotation chain is serial per problem but independent across problems: 32 problems per warp.
// Returns the number of sweeps, or -1 on overflow / non-convergence.
// d, e: private copies with element stride `st` (shared memory laid out [k][thread], so the lanes of
// a warp touch consecutive banks).
__device__ inline int ql_chain_record(double* d, double* e, int st, int n, double2* cs_out, int cs_cap,
                                      int* hdr_out, int hdr_cap, int* n_rot_out) {
  int nrot = 0, nsw = 0;
  for (int l = 0; l < n; ++l) {
    for (int iter = 0;; ++iter) {
      int m = l;
      for (; m < n - 1; ++m) {
        const double dd = fabs(d[m * st]) + fabs(d[(m + 1) * st]);
        if (fabs(e[m * st]) <= 2.220446049250313e-16 * dd) break;
      }
      if (m == l) break;
      if (iter >= 60) return -1;
      double g = (d[(l + 1) * st] - d[l * st]) / (2.0 * e[l * st]);
      double r = sqrt(g * g + 1.0);
      g = d[m * st] - d[l * st] + e[l * st] / (g + (g >= 0.0 ? fabs(r) : -fabs(r)));
      double s = 1.0, c = 1.0, p = 0.0;
      int i = m - 1;
      bool brk = false;
      if (nrot + (m - l) > cs_cap || nsw >= hdr_cap) return -1;
      // operands of the next step are fetched one step ahead; d[i+1] is carried in a register
      double e_i = e[i * st], d_i = d[i * st], d_ip1 = d[(i + 1) * st];
      for (; i >= l; --i) {
        const double e_nx = i > l ? e[(i - 1) * st] : 0.0, d_nx = i > l ? d[(i - 1) * st] : 0.0;
        const double f = s * e_i;
        const double b = c * e_i;
        const double h2 = f * f + g * g;
        const double ir = rsqrt(h2);
        r = h2 * ir;
        e[(i + 1) * st] = h2 > 0.0 ? r : 0.0;
        if (!(h2 > 0.0)) {
          d[(i + 1) * st] = d_ip1 - p;
          e[m * st] = 0.0;
          brk = true;
          break;
        }
        s = f * ir;
        c = g * ir;
        g = d_ip1 - p;
        r = (d_i - g) * s + 2.0 * c * b;
        p = s * r;
        d[(i + 1) * st] = g + p;
        g = c * r - b;
        cs_out[nrot++] = make_double2(c, s);
        d_ip1 = d_i;
        d_i = d_nx;
        e_i = e_nx;
      }
      const int ilo = brk ? i + 1 : l;
      if (!brk) {
        d[l * st] = d_ip1 - p;
        e[l * st] = g;
        e[m * st] = 0.0;
      }
      hdr_out[nsw++] = m | (ilo << 8);
    }
  }
  *n_rot_out = nrot;
  return nsw;
}

}  // namespace bmfm
