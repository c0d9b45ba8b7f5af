// Expanding-window exponentially weighted Newey-West factor covariance for every date.
//
// Replaces  MFM.Newey_West_by_time   Barra-master/mfm/MFM.py:80-101
//           utils.Newey_West         Barra-master/mfm/utils.py:16-50
//
// The reference recomputes the whole weighted covariance from scratch for each of the T
// expanding windows (O(T^2 K^2 (q+1)) python-level outer products).  With decay
// lam = 0.5^(1/tao) every moment it needs obeys a first-order recurrence in the window
// length L:
//     A0_L   = lam A0_{L-1}   + x_{L-1} x_{L-1}'          (lag-0 raw second moment)
//     B_i,L  = lam B_i,{L-1}  + x_{L-1-i} x_{L-1}'        (lag-i raw cross moment, weight of the later row)
//     S_L    = lam S_{L-1}    + x_{L-1},  Z_L = lam Z_{L-1} + 1
//     R_i,L  = lam R_i,{L-1}  + x_{L-1} [L-1 >= i],  Zi_L likewise      (sums over rows r >= i)
//     P_i,L  = lam P_i,{L-1}  + x_{L-1-i}                 (= S_{L-i}, the partner-row sum)
// and with m = S_L / Z_L (the EW mean, utils.py:36-37)
//     Gamma_0 = A0/Z - m m'
//     Gamma_i = ( B_i - P_i m' - m R_i' + m m' Zi ) / Z              (utils.py:46)
//     V = Gamma_0 + sum_i (1 - i/(1+q)) (Gamma_i + Gamma_i')             (utils.py:48)
// O(T K^2 (q+1)) work.  Time is cut into chunks so all SMs work: kernel 1 reduces each
// chunk from a zero start, kernel 2 scans the chunk totals (tiny), kernel 3 replays each
// chunk from its true start state and streams out K x K matrices, one coalesced row-major
// write per element.  Scratch: nchunks * ((1+q)(K^2+1) + (1+2q)K) doubles.
#include <math.h>
#include <stdlib.h>
#include <mutex>
#include "bmfm_common.cuh"

namespace bmfm {

constexpr int NW_MAX_Q = 4;
constexpr int NW_LOCAL_THREADS = 256;   // chunk-total kernel: one 2 x 2 block of every lag matrix (or one vector element) per thread
constexpr int NW_EMIT_THREADS = 128;    // emit kernel: one 2 x 2 block of the covariance per thread
constexpr int NW_TARGET_CHUNKS = 144;   // chunks per series aimed at (state traffic ~ nchunks * 44 KB at K = 42)
constexpr int NW_EMIT_SMEM = 56 * 1024; // shared memory the emit kernel may use (several CTAs per SM)

struct NwParams {
  int T, K, q, C, nchunks;
  int t0, t1, c_first;   // emit only dates [t0, t1) (chunks from c_first on), written at row t - t0
  int c_used;            // chunks 0 .. c_used-1 are needed (those up to the last emitted one)
  int row0;              // global row of chunk 0's first row (0 unless the series is sharded by dates)
  // date-sharded form: the chunks start at this rank's first date from a ZERO state; the decayed state of all
  // earlier dates (`carry`, nstate doubles, as of row row0 - 1) is added when a chunk's start state is read,
  // and the zero-start state after the rank's last row goes to slot `rank` of every peer's totals buffer
  const double* carry;
  int n_peer, rank;
  double* peer_totals[8];
  BmfmFlagPeers flags;   // per-producer flags (n = 0: the caller runs a barrier instead)
  double lam, log_lam;
  double fC, fLast, fPer;   // lam^C, lam^(length of the last chunk), lam^(C * chunks per scan part): set by nw_run_totals
  const double* fr;   // [T][K]
  double* cov;        // [T][K][K]
  int* status;        // [T]
  double* state;      // [nchunks][nstate]   chunk totals, then chunk start states
};

__host__ __device__ inline int nw_nstate(int K, int q) { return (1 + q) * (K * K + 1) + (1 + 2 * q) * K; }
// layout of one state record:
//   mats[(1+q)][K][K] | vecs[(1+2q)][K] (S, R_1..R_q, P_1..P_q) | scal[(1+q)] (Z, Z_1..Z_q)
__host__ __device__ inline size_t nw_emit_smem(int C, int K, int q) {
  return ((size_t)(C + q) * K + 2 * (size_t)C * K + 2 * (size_t)C) * sizeof(double);
}
// chunk length: about NW_TARGET_CHUNKS chunks, at least 8 rows, short enough for the emit kernel's shared memory
__host__ inline int nw_pick_chunk(int T, int K, int q, int target_opt = 0) {
  const int target = target_opt > 0 ? target_opt : NW_TARGET_CHUNKS;   // bmfm_ctx_set_option "nw_chunks": tuning knob
  int C = (T + target - 1) / target;
  if (C < 8) C = 8;
  while (C > 1 && nw_emit_smem(C, K, q) > (size_t)NW_EMIT_SMEM) --C;
  return C;
}

// stage rows r0-q .. r1-1 of the series (zeros before row 0) in shared memory: issue only, the caller waits
// (cp_async_wait_all + __syncthreads) after whatever else it can do meanwhile
__device__ __forceinline__ void nw_stage_rows_issue(const NwParams& p, double* sx, int r0, int len) {
  // rows r0-q .. r0+len-1 are one contiguous piece of the series; cp.async keeps every copy in flight at once
  // (a load-then-store loop would pay one global-memory latency per row)
  const int K = p.K, q = p.q;
  const int rlo = max(0, r0 - q);                    // first real row
  const int zero = (rlo - (r0 - q)) * K;             // leading elements that stand for rows before row 0
  const int n = (r0 + len - rlo) * K;                // real elements
  for (int e = threadIdx.x; e < zero; e += blockDim.x) sx[e] = 0.0;
  const double* src = p.fr + (size_t)rlo * K;
  double* dst = sx + zero;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
    for (int e = threadIdx.x; e < (n >> 1); e += blockDim.x) cp_async_16(dst + 2 * e, src + 2 * e);
    if ((n & 1) && threadIdx.x == 0) cp_async_8(dst + n - 1, src + n - 1);
  } else {
    for (int e = threadIdx.x; e < n; e += blockDim.x) cp_async_8(dst + e, src + e);
  }
}
__device__ __forceinline__ void nw_stage_rows(const NwParams& p, double* sx, int r0, int len) {
  nw_stage_rows_issue(p, sx, r0, len);
  cp_async_wait_all();
}

// kernel 1: chunk totals from a zero start (grid: chunks x item blocks).  Work items of a chunk: the KB x KB
// 2 x 2 blocks of the moment matrices (KB = ceil(K/2); one thread carries its block of ALL 1+q lag matrices, so the
// rows it reads are shared between 4 (1+q) accumulators), then the (1+2q) K vector elements, then the 1+q scalars.
template <int QL>
__global__ void __launch_bounds__(NW_LOCAL_THREADS) nw_chunk_local(NwParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int q = QL;
  double* sx = reinterpret_cast<double*>(smem_raw);   // [(C+q)][K]
  const int c = blockIdx.x, K = p.K;
  const int r0 = p.row0 + c * p.C, r1 = min(p.T, r0 + p.C), len = r1 - r0;
  double* st = p.state + (size_t)c * nw_nstate(K, q);
  nw_stage_rows(p, sx, r0, len);
  __syncthreads();
  const double lam = p.lam;
  const int KB = (K + 1) >> 1, nblk = KB * KB;
  const int nm = (1 + q) * K * K, nv = (1 + 2 * q) * K;
  int e = blockIdx.y * blockDim.x + threadIdx.x;
  if (e < nblk) {
    const int ba = e / KB, bb = e - ba * KB;
    const int a = 2 * ba, b = 2 * bb;
    const bool aok = a + 1 < K, bok = b + 1 < K;
    const int a1 = aok ? a + 1 : a, b1 = bok ? b + 1 : b;
    double s[QL + 1][2][2];
#pragma unroll
    for (int i = 0; i <= QL; ++i) s[i][0][0] = s[i][0][1] = s[i][1][0] = s[i][1][1] = 0.0;
    const double* xr = sx + q * K;            // row j; row j - i is i rows above (zero rows before row 0)
    for (int j = 0; j < len; ++j, xr += K) {
      const double vb0 = xr[b], vb1 = xr[b1];
#pragma unroll
      for (int i = 0; i <= QL; ++i) {
        const double la0 = xr[a - i * K], la1 = xr[a1 - i * K];
        s[i][0][0] = fma(s[i][0][0], lam, la0 * vb0);
        s[i][0][1] = fma(s[i][0][1], lam, la0 * vb1);
        s[i][1][0] = fma(s[i][1][0], lam, la1 * vb0);
        s[i][1][1] = fma(s[i][1][1], lam, la1 * vb1);
      }
    }
#pragma unroll
    for (int i = 0; i <= QL; ++i) {
      double* m = st + (size_t)i * K * K;
      m[a * K + b] = s[i][0][0];
      if (bok) m[a * K + b1] = s[i][0][1];
      if (aok) {
        m[a1 * K + b] = s[i][1][0];
        if (bok) m[a1 * K + b1] = s[i][1][1];
      }
    }
    return;
  }
  e -= nblk;
  if (e < nv) {
    const int v = e / K, k = e - v * K;
    double s = 0.0;
    for (int j = 0; j < len; ++j) {
      const int r = r0 + j;
      double add;
      if (v == 0) add = sx[(q + j) * K + k];
      else if (v <= q) add = r >= v ? sx[(q + j) * K + k] : 0.0;
      else add = sx[(q + j - (v - q)) * K + k];
      s = fma(s, lam, add);
    }
    st[nm + e] = s;
  } else if (e < nv + 1 + q) {
    const int i = e - nv;
    double s = 0.0;
    for (int r = r0; r < r1; ++r) {
      s *= lam;
      if (r >= i) s += 1.0;
    }
    st[nm + e] = s;
  }
}

// kernel 2: exclusive scan over chunks.  Every chunk but the last has length C, so the recurrence over chunks
// is start_{c+1} = lam^C start_c + total_c.  Sixteen threads share a state element: each walks one sixteenth of the
// chunks from zero, the sixteen partial results are chained through shared memory, then each walks its part
// again from its true start and overwrites the totals with the start states.  (One thread per element and a
// serial walk over all chunks is a chain of global-memory latencies.)
constexpr int NW_SCAN_GROUPS = 16;
__global__ void __launch_bounds__(32 * NW_SCAN_GROUPS) nw_chunk_scan(NwParams p) {
  __shared__ double s_part[NW_SCAN_GROUPS][32];
  const int ns = nw_nstate(p.K, p.q);
  const int ex = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int e = blockIdx.x * 32 + ex;
  const bool act = e < ns;
  const int per = max(1, (p.nchunks + NW_SCAN_GROUPS - 1) / NW_SCAN_GROUPS);   // from nchunks, not c_used: a date range
                                                                                 // then adds in the same order as the whole series
  const int c_lo = min(p.c_used, g * per), c_hi = min(p.c_used, c_lo + per);
  // lam^C, and lam^len for the last chunk, which may be shorter than C (computed once on the host: a pow per
  // thread was a third of this kernel's instructions)
  const double fC = p.fC, fLast = p.fLast;
  double* col = p.state + (act ? e : 0);
  double run = 0.0;
  for (int c0 = c_lo; c0 < c_hi; c0 += 8) {
    double loc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) loc[j] = (act && c0 + j < c_hi) ? col[(size_t)(c0 + j) * ns] : 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (c0 + j < c_hi) run = fma(run, c0 + j == p.nchunks - 1 ? fLast : fC, loc[j]);
  }
  s_part[g][ex] = run;
  __syncthreads();
  // state at the start of this thread's part: every earlier part holds exactly `per` full chunks
  double start = 0.0;
  {
    const double fPer = p.fPer;
    for (int h = 0; h < g; ++h) start = fma(start, fPer, s_part[h][ex]);
  }
  run = start;
  for (int c0 = c_lo; c0 < c_hi; c0 += 8) {
    double loc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) loc[j] = (act && c0 + j < c_hi) ? col[(size_t)(c0 + j) * ns] : 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (c0 + j < c_hi) {
        if (act) col[(size_t)(c0 + j) * ns] = run;
        run = fma(run, c0 + j == p.nchunks - 1 ? fLast : fC, loc[j]);
      }
    }
  }
  // the state after the last row (date-sharded form: all chunks are scanned): held by the last non-empty part
  const int g_last = p.c_used > 0 ? (p.c_used - 1) / per : 0;
  if (p.n_peer > 0 && act && g == g_last)
    for (int r = 0; r < p.n_peer; ++r) p.peer_totals[r][(size_t)p.rank * ns + e] = run;
  if (p.n_peer > 0 && p.flags.n > 0) {
    // the CTA that finishes last publishes the record's epoch to the later ranks
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      long long* mine = p.flags.f[p.flags.rank];
      const unsigned long long done = atomicAdd(reinterpret_cast<unsigned long long*>(mine + 40 + BMFM_CH_NW), 1ULL) + 1ULL;
      if (done % gridDim.x == 0) {
        __threadfence_system();
        const long long ep = (long long)(done / gridDim.x);
        *reinterpret_cast<volatile long long*>(mine + 32 + BMFM_CH_NW) = ep;
        bmfm_flag_publish(p.flags, BMFM_CH_NW, ep);
      }
    }
  }
}

// carry-in of a date-sharded rank: decayed sum of the earlier ranks' zero-start totals,
//   carry = sum_{r' < rank} lam^(t0 - end_{r'}) total_{r'}      (end_{r'} = one past rank r' 's last row)
struct NwEnds { long long v[8]; };
__global__ void nw_carry_kernel(int ns, int rank, double lam, const double* totals, NwEnds ends, long long t0,
                                double* carry, BmfmFlagPeers fp) {
  if (fp.n > 0) {
    if (threadIdx.x == 0) bmfm_flag_wait(fp, BMFM_CH_NW);
    __syncthreads();
  }
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= ns) return;
  double s = 0.0;
  for (int r = 0; r < rank; ++r) s = fma(pow(lam, (double)(t0 - ends.v[r])), __ldcg(totals + (size_t)r * ns + e), s);
  carry[e] = s;
}

// kernel 3: replay the chunk from its start state and emit V for every date in it.  Grid: chunks x element
// blocks, one covariance element (a, b) per thread.  With m = S/Z, w_i = 1 - i/(1+q) and
//   U_k = sum_i w_i (P_i,k + R_i,k) / Z,   cc = 1 - 2 sum_i w_i Z_i / Z
// the combination of utils.py:46-48 is
//   V_ab = [A0_ab + sum_i w_i (B_i,ab + B_i,ba)] / Z - m_a m_b cc - m_b U_a - m_a U_b,
// so the per-date vectors m, U and scalars 1/Z, cc are prepared once per CTA and the element loop is a few
// FMAs per date.
template <int QL>   // QL = q, the number of lags (compile-time so that the lag loops carry no predicates)
__global__ void __launch_bounds__(NW_EMIT_THREADS) nw_emit(NwParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int q = QL;
  constexpr int QN = QL > 0 ? QL : 1;
  const int c = p.c_first + blockIdx.x, K = p.K, C = p.C;
  const int r0 = p.row0 + c * C, r1 = min(p.T, r0 + C), len = r1 - r0;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int nm = (1 + q) * K * K;
  const double* st = p.state + (size_t)c * nw_nstate(K, q);
  double* sx = reinterpret_cast<double*>(smem_raw);   // [(C+q)][K]  rows r0-q .. r1-1 (zeros before row 0)
  double* sm = sx + (C + q) * K;                       // [C][K]  EW mean after row r0+j
  double* su = sm + C * K;                             // [C][K]  U
  double* siz = su + C * K;                            // [C]     1/Z
  double* scc = siz + C;                               // [C]     cc
  const double lam = p.lam;
  const double* cr = p.carry;                                           // sharded form: decayed state of earlier ranks
  const double fcar = cr ? pow(lam, (double)(r0 - p.row0)) : 0.0;
#define NW_LD(idx) (cr ? fma(fcar, cr[idx], st[idx]) : st[idx])
  nw_stage_rows_issue(p, sx, r0, len);
  // ---- while the rows are in flight: everything that does not need them ----
  // this thread's 2 x 2 block of V (rows a, a+1 and columns b, b+1 share their loads).  V is symmetric: only blocks
  // on or above the block diagonal are computed (row-major enumeration of the upper triangle of the KB x KB block
  // grid), each thread also stores the transpose of its block
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  const int KB = (K + 1) >> 1;
  const int e = blockIdx.y * nt + tid;
  const bool eact = e < KB * (KB + 1) / 2;
  int ba = 0, rem = eact ? e : 0;
  while (rem >= KB - ba) { rem -= KB - ba; ++ba; }
  const int a = 2 * ba, b = 2 * (ba + rem);
  const int a1 = a + 1 < K ? a + 1 : a, b1 = b + 1 < K ? b + 1 : b;     // clamped partner (stores are guarded)
  const bool aok = a + 1 < K, bok = b + 1 < K;
  const int ia[2] = {a, a1}, ib[2] = {b, b1};
  double a0[2][2], bab[QN][2][2], bba[QN][2][2], wl[QN];
  if (eact) {          // start state of the block (global loads, L2)
#pragma unroll
    for (int x = 0; x < 2; ++x)
#pragma unroll
      for (int y = 0; y < 2; ++y) {
        a0[x][y] = NW_LD(ia[x] * K + ib[y]);
#pragma unroll
        for (int i = 0; i < QL; ++i) {
          bab[i][x][y] = NW_LD((1 + i) * K * K + ia[x] * K + ib[y]);
          bba[i][x][y] = NW_LD((1 + i) * K * K + ib[y] * K + ia[x]);
        }
      }
  }
  // start state of the vector sums of factor k = tid (further k of a thread, K > blockDim, are loaded in the loop)
  double S0 = 0.0, R0[QN], P0[QN];
  if (tid < K) {
    S0 = NW_LD(nm + tid);
#pragma unroll
    for (int i = 0; i < QL; ++i) {
      R0[i] = NW_LD(nm + (1 + i) * K + tid);
      P0[i] = NW_LD(nm + (1 + q + i) * K + tid);
    }
  }
  // the scalar sums Z, Z_i do not depend on the data: closed form per date (one thread per date).  With L = r + 1
  // rows so far, Z = (1 - lam^L) / (1 - lam) and Z_i the same with L - i rows (those with index >= i);
  // 1 - lam^n = -expm1(n log lam) keeps full precision when n log lam is small
  for (int j = tid; j < len; j += nt) {
    const int L = r0 + j + 1;
    const double inv1 = -1.0 / expm1(p.log_lam);
    const double iz = 1.0 / (-expm1((double)L * p.log_lam) * inv1);
    double zs = 0.0;
#pragma unroll
    for (int i = 0; i < QL; ++i) {
      const int lag = i + 1;
      const double Zl = L > lag ? -expm1((double)(L - lag) * p.log_lam) * inv1 : 0.0;
      zs = fma(1.0 - (double)lag / (1.0 + q), Zl, zs);
    }
    siz[j] = iz;
    scc[j] = 1.0 - 2.0 * zs * iz;
  }
  cp_async_wait_all();
  __syncthreads();
  for (int k = tid; k < K; k += nt) {
    double S = S0;
    double R[QN], Pv[QN];
#pragma unroll
    for (int i = 0; i < QL; ++i) { R[i] = R0[i]; Pv[i] = P0[i]; }
    if (k != tid) {
      S = NW_LD(nm + k);
#pragma unroll
      for (int i = 0; i < QL; ++i) {
        R[i] = NW_LD(nm + (1 + i) * K + k);
        Pv[i] = NW_LD(nm + (1 + q + i) * K + k);
      }
    }
    for (int j = 0; j < len; ++j) {
      const int r = r0 + j;
      const double x = sx[(q + j) * K + k];
      S = fma(S, lam, x);
      double u = 0.0;
#pragma unroll
      for (int i = 0; i < QL; ++i) {
        const int lag = i + 1;
        R[i] = fma(R[i], lam, r >= lag ? x : 0.0);
        Pv[i] = fma(Pv[i], lam, sx[(q + j - lag) * K + k]);      // zero-filled before row 0
        u = fma(1.0 - (double)lag / (1.0 + q), Pv[i] + R[i], u);
      }
      const double iz = siz[j];
      sm[j * K + k] = S * iz;
      su[j * K + k] = u * iz;
    }
  }
  __syncthreads();
  if (eact) {
    const bool keven = (K & 1) == 0 && (reinterpret_cast<uintptr_t>(p.cov) & 15) == 0;
#pragma unroll
    for (int i = 0; i < QL; ++i) wl[i] = 1.0 - (double)(i + 1) / (1.0 + q);
    const int jlo = max(0, p.t0 - r0), jhi = min(len, p.t1 - r0);   // dates of this chunk that are emitted
    const int jnan = min(jhi, max(q, K) - r0);                      // dates with L = r + 1 <= max(q, K): NaN
    const long long kk = (long long)K * K;
    double* dst = p.cov + (((long long)(r0 - p.t0) * K + a) * K + b);   // date r0 (dereferenced for emitted dates only)
    double* dstT = p.cov + (((long long)(r0 - p.t0) * K + b) * K + a);
    for (int j = 0; j < jhi; ++j) {
      const double* xr = sx + (q + j) * K;
      const double va[2] = {xr[a], xr[a1]}, vb[2] = {xr[b], xr[b1]};
      double acc[2][2];
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int y = 0; y < 2; ++y) {
          a0[x][y] = fma(a0[x][y], lam, va[x] * vb[y]);
          acc[x][y] = a0[x][y];
        }
#pragma unroll
      for (int i = 0; i < QL; ++i) {
        const double* xl = xr - (i + 1) * K;                         // zero rows before row 0
        const double la[2] = {xl[a], xl[a1]}, lb[2] = {xl[b], xl[b1]};
#pragma unroll
        for (int x = 0; x < 2; ++x)
#pragma unroll
          for (int y = 0; y < 2; ++y) {
            bab[i][x][y] = fma(bab[i][x][y], lam, la[x] * vb[y]);
            bba[i][x][y] = fma(bba[i][x][y], lam, lb[y] * va[x]);
            acc[x][y] = fma(wl[i], bab[i][x][y] + bba[i][x][y], acc[x][y]);
          }
      }
      if (j >= jlo) {
        double o00 = nan, o01 = nan, o10 = nan, o11 = nan;
        if (j >= jnan) {
          const double* mj = sm + j * K;
          const double* uj = su + j * K;
          const double iz = siz[j], cc = scc[j];
          const double ma[2] = {mj[a], mj[a1]}, mb[2] = {mj[b], mj[b1]};
          const double ua[2] = {uj[a], uj[a1]}, ub[2] = {uj[b], uj[b1]};
          o00 = acc[0][0] * iz - ma[0] * mb[0] * cc - mb[0] * ua[0] - ma[0] * ub[0];
          o01 = acc[0][1] * iz - ma[0] * mb[1] * cc - mb[1] * ua[0] - ma[0] * ub[1];
          o10 = acc[1][0] * iz - ma[1] * mb[0] * cc - mb[0] * ua[1] - ma[1] * ub[0];
          o11 = acc[1][1] * iz - ma[1] * mb[1] * cc - mb[1] * ua[1] - ma[1] * ub[1];
        }
        if (keven) {                       // a, b even and K even: 16-byte aligned pairs
          st_stream2(dst, o00, o01);
          st_stream2(dst + K, o10, o11);
          if (a != b) {                    // the transposed block, rows b, b+1 and columns a, a+1
            st_stream2(dstT, o00, o10);
            st_stream2(dstT + K, o01, o11);
          }
        } else {
          st_stream(dst, o00);
          if (bok) st_stream(dst + 1, o01);
          if (aok) {
            st_stream(dst + K, o10);
            if (bok) st_stream(dst + K + 1, o11);
          }
          if (a != b) {
            st_stream(dstT, o00);
            if (aok) st_stream(dstT + 1, o10);
            if (bok) {
              st_stream(dstT + K, o01);
              if (aok) st_stream(dstT + K + 1, o11);
            }
          }
        }
      }
      dst += kk;
      dstT += kk;
    }
  }
  if (blockIdx.y == 0)
    for (int j = tid; j < len; j += nt) {
      const int r = r0 + j, L = r + 1;
      if (r >= p.t0 && r < p.t1) p.status[r - p.t0] = (L <= q || L <= K) ? BMFM_ST_SHORT : BMFM_ST_OK;
    }
}

#undef NW_LD

template <int QL>
static int nw_launch_emit(bmfm_ctx* ctx, const NwParams& p, int nchunks_emit, size_t smem_emit) {
  if (smem_emit > 48 * 1024)
    BMFM_CUDA(ctx, cudaFuncSetAttribute(nw_emit<QL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_emit));
  nw_emit<QL><<<dim3(nchunks_emit, (((p.K + 1) / 2) * ((p.K + 1) / 2 + 1) / 2 + NW_EMIT_THREADS - 1) / NW_EMIT_THREADS), NW_EMIT_THREADS, smem_emit, ctx->stream>>>(p);
  return BMFM_OK;
}

template <int QL>
static int nw_launch_local(bmfm_ctx* ctx, const NwParams& p, size_t smem_local) {
  if (smem_local > 48 * 1024)
    BMFM_CUDA(ctx, cudaFuncSetAttribute(nw_chunk_local<QL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_local));
  const int KB = (p.K + 1) / 2;
  const int items = KB * KB + (1 + 2 * QL) * p.K + 1 + QL;
  nw_chunk_local<QL><<<dim3(p.c_used, (items + NW_LOCAL_THREADS - 1) / NW_LOCAL_THREADS), NW_LOCAL_THREADS, smem_local, ctx->stream>>>(p);
  return BMFM_OK;
}

// chunk totals + scan for a filled-in NwParams (C, nchunks, state, row0 and the date range set by the caller)
static int nw_run_totals(bmfm_ctx* ctx, NwParams& p) {
  const int K = p.K, q = p.q;
  p.c_first = (p.t0 - p.row0) / p.C;
  const int c_last = p.t1 > p.t0 ? (p.t1 - 1 - p.row0) / p.C : p.c_first - 1;
  p.c_used = c_last + 1;   // later chunks feed nothing this call emits
  const size_t smem_local = (size_t)(p.C + q) * K * sizeof(double);
  if (smem_local > (size_t)ctx->max_smem_optin || nw_emit_smem(p.C, K, q) > (size_t)ctx->max_smem_optin) {
    snprintf(ctx->err, sizeof(ctx->err), "Newey-West chunk does not fit shared memory (K=%d q=%d)", K, q);
    return BMFM_E_UNSUPPORTED;
  }
  if (c_last < p.c_first && p.n_peer == 0) return BMFM_OK;
  const int ns = nw_nstate(K, q);
  p.fC = pow(p.lam, (double)p.C);
  p.fLast = pow(p.lam, (double)(p.T - p.row0 - (p.nchunks - 1) * p.C));
  {
    const int per = (p.nchunks + NW_SCAN_GROUPS - 1) / NW_SCAN_GROUPS;      // chunks per scan part, as in nw_chunk_scan
    p.fPer = pow(p.fC, (double)(per > 1 ? per : 1));
  }
  if (p.c_used > 0) {
    int rc = BMFM_E_UNSUPPORTED;
    switch (q) {
      case 0: rc = nw_launch_local<0>(ctx, p, smem_local); break;
      case 1: rc = nw_launch_local<1>(ctx, p, smem_local); break;
      case 2: rc = nw_launch_local<2>(ctx, p, smem_local); break;
      case 3: rc = nw_launch_local<3>(ctx, p, smem_local); break;
      case 4: rc = nw_launch_local<4>(ctx, p, smem_local); break;
    }
    if (rc) return rc;
    ctx->launches++;
  }
  nw_chunk_scan<<<(ns + 31) / 32, 32 * NW_SCAN_GROUPS, 0, ctx->stream>>>(p);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

// emit for the same NwParams (after nw_run_totals)
static int nw_run_emit(bmfm_ctx* ctx, NwParams& p) {
  const int q = p.q;
  p.c_first = (p.t0 - p.row0) / p.C;
  const int c_last = p.t1 > p.t0 ? (p.t1 - 1 - p.row0) / p.C : p.c_first - 1;
  p.c_used = c_last + 1;
  if (c_last < p.c_first) return BMFM_OK;
  const size_t smem_emit = nw_emit_smem(p.C, p.K, q);
  const int ne = c_last - p.c_first + 1;
  int rc = BMFM_E_UNSUPPORTED;
  switch (q) {
    case 0: rc = nw_launch_emit<0>(ctx, p, ne, smem_emit); break;
    case 1: rc = nw_launch_emit<1>(ctx, p, ne, smem_emit); break;
    case 2: rc = nw_launch_emit<2>(ctx, p, ne, smem_emit); break;
    case 3: rc = nw_launch_emit<3>(ctx, p, ne, smem_emit); break;
    case 4: rc = nw_launch_emit<4>(ctx, p, ne, smem_emit); break;
  }
  if (rc) return rc;
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

static int nw_run(bmfm_ctx* ctx, NwParams& p) {
  const int rc = nw_run_totals(ctx, p);
  return rc ? rc : nw_run_emit(ctx, p);
}

}  // namespace bmfm

using namespace bmfm;

// Scratch (device) bytes bmfm_newey_west_by_time needs for given sizes.
extern "C" size_t bmfm_newey_west_scratch_bytes(int T, int K, int q) {
  if (T <= 0) return 0;
  const int C = nw_pick_chunk(T, K, q);
  const int nchunks = (T + C - 1) / C;
  return (size_t)nchunks * nw_nstate(K, q) * sizeof(double);
}

static size_t nw_scratch_bytes_ctx(const bmfm_ctx* ctx, int T, int K, int q) {
  const int C = nw_pick_chunk(T, K, q, ctx->opt_nw_chunks);
  return (size_t)((T + C - 1) / C) * nw_nstate(K, q) * sizeof(double);
}

static int nw_launch(bmfm_ctx* ctx, int T, int K, int q, double tao, const double* fr, double* cov,
                     int32_t* status, double* scratch, int t0, int t1) {
  NwParams p;
  p.T = T; p.K = K; p.q = q;
  p.C = nw_pick_chunk(T, K, q, ctx->opt_nw_chunks);
  p.nchunks = (T + p.C - 1) / p.C;
  p.lam = pow(0.5, 1.0 / tao);
  p.log_lam = log(p.lam);
  p.fr = fr; p.cov = cov; p.status = status; p.state = scratch;
  p.row0 = 0; p.carry = nullptr; p.n_peer = 0; p.rank = 0; p.flags.n = 0;
  if (t0 < 0) t0 = 0;
  if (t1 > T) t1 = T;
  p.t0 = t0; p.t1 = t1;
  return nw_run(ctx, p);
}

// dates [t0, t1) only: cov [(t1-t0)*K*K], status [t1-t0].  The decayed moments of ALL earlier dates are
// still accumulated (chunk totals + scan read the whole series, T*K doubles), so a rank that owns a block
// of dates gets exactly the expanding-window result without any exchange beyond the factor returns.
extern "C" int bmfm_newey_west_by_time_range(bmfm_ctx* ctx, int T, int K, int q, double tao,
                                             const double* factor_ret, int t0, int t1, double* cov,
                                             int32_t* status) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && K > 0 && q >= 0 && tao > 0 && t0 >= 0 && t1 >= t0 && t1 <= T, "bad sizes");
  BMFM_ARG(ctx, factor_ret && (t1 == t0 || (cov && status)), "NULL pointer");
  if (q > NW_MAX_Q) {
    snprintf(ctx->err, sizeof(ctx->err), "q=%d exceeds compiled limit %d", q, NW_MAX_Q);
    return BMFM_E_UNSUPPORTED;
  }
  if (T == 0 || t1 == t0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = bmfm_ws_reserve(ctx, nw_scratch_bytes_ctx(ctx, T, K, q));
  if (rc) return rc;
  return nw_launch(ctx, T, K, q, tao, factor_ret, cov, status, (double*)ctx->ws, t0, t1);
}

extern "C" int bmfm_newey_west_by_time(bmfm_ctx* ctx, int T, int K, int q, double tao,
                                       const double* factor_ret, double* cov, int32_t* status) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && K > 0 && q >= 0 && tao > 0, "bad sizes");
  if (q > NW_MAX_Q) {
    snprintf(ctx->err, sizeof(ctx->err), "q=%d exceeds compiled limit %d", q, NW_MAX_Q);
    return BMFM_E_UNSUPPORTED;
  }
  if (T == 0) return BMFM_OK;      // an empty series needs no buffers
  BMFM_ARG(ctx, factor_ret && cov && status, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = bmfm_ws_reserve(ctx, nw_scratch_bytes_ctx(ctx, T, K, q));
  if (rc) return rc;
  return nw_launch(ctx, T, K, q, tao, factor_ret, cov, status, (double*)ctx->ws, 0, T);
}

// ---------------------------------------------------------------------------------------------
// Newey-West without any allocation
// ---------------------------------------------------------------------------------------------
// Some hosts of this library are not allowed to allocate device memory when they call it: a risk
// service whose GPU memory was carved up at start-up (every byte belongs to a pool it does not
// own, so a cudaMalloc from inside a library can fail or push another tenant over its budget), or
// a latency-bound loop that cannot tolerate the device-wide synchronisation a cudaMalloc/cudaFree
// implies.  bmfm_newey_west_by_time() grows the context workspace with cudaMalloc on first use;
// the entry point below never does.  Its chunk-state scratch is this fixed array, which lives in
// the library image: every GPU that loads libbmfm.so holds these 16 MiB until the process exits.
constexpr size_t NW_STATIC_DOUBLES = (size_t)1 << 21;   // 2 Mi doubles = 16 MiB
static __device__ __align__(256) double nw_static_scratch[NW_STATIC_DOUBLES];

// In host code `nw_static_scratch` is only a handle for the runtime, not a device address, so the
// address is looked up once per device (with that device current) and cached here.
constexpr int NW_MAX_DEVICES = 64;
static double* nw_static_ptr[NW_MAX_DEVICES];
static std::mutex nw_static_mutex;

// chunk length that keeps nchunks * nstate inside the static array (0 if impossible)
static int nw_static_chunk(int T, int K, int q) {
  const size_t ns = (size_t)nw_nstate(K, q);
  const size_t max_chunks = NW_STATIC_DOUBLES / ns;
  if (max_chunks == 0) return 0;
  int C = nw_pick_chunk(T, K, q);
  const size_t need = ((size_t)T + max_chunks - 1) / max_chunks;   // smallest C with ceil(T/C) <= max_chunks
  if ((size_t)C < need) C = (int)need;
  return C;
}

// Same computation and results as bmfm_newey_west_by_time, with no cudaMalloc, cudaFree or other
// allocating call on any path.
//   * Limits: q <= 4, and (1+q)(K^2+1) + (1+2q)K doubles of state per chunk must fit the 16 MiB
//     array with chunks short enough for shared memory (K = 42, q = 2: any T up to ~85 000);
//     larger problems are rejected with BMFM_E_UNSUPPORTED and a message, nothing is launched.
//   * The first call on a device resolves the array's device address (cudaGetSymbolAddress);
//     make that first call before starting any stream capture.
//   * All calls on a device share the one array and it keeps whatever the previous call left.
//     Nothing in it has to start at zero — the first kernel of a call writes every cell that the
//     later kernels read — but two calls whose kernels overlap on the GPU (different streams or
//     contexts, same device) race on it; serialise them.
extern "C" int bmfm_newey_west_by_time_noalloc(bmfm_ctx* ctx, int T, int K, int q, double tao,
                                               const double* factor_ret, double* cov, int32_t* status) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && K > 0 && q >= 0 && tao > 0, "bad sizes");
  BMFM_ARG(ctx, factor_ret && cov && status, "NULL pointer");
  if (q > NW_MAX_Q) {
    snprintf(ctx->err, sizeof(ctx->err), "q=%d exceeds compiled limit %d", q, NW_MAX_Q);
    return BMFM_E_UNSUPPORTED;
  }
  if (T == 0) return BMFM_OK;
  if (ctx->device < 0 || ctx->device >= NW_MAX_DEVICES) {
    snprintf(ctx->err, sizeof(ctx->err), "device %d outside the no-alloc table (%d devices)", ctx->device, NW_MAX_DEVICES);
    return BMFM_E_UNSUPPORTED;
  }
  NwParams p;
  p.T = T; p.K = K; p.q = q;
  p.C = nw_static_chunk(T, K, q);
  const size_t smem = p.C > 0 ? nw_emit_smem(p.C, K, q) : 0;
  if (p.C == 0 || smem > (size_t)ctx->max_smem_optin) {
    snprintf(ctx->err, sizeof(ctx->err),
             "newey_west noalloc: T=%d K=%d q=%d needs more than the fixed %zu-byte scratch; "
             "use bmfm_newey_west_by_time", T, K, q, NW_STATIC_DOUBLES * sizeof(double));
    return BMFM_E_UNSUPPORTED;
  }
  p.nchunks = (T + p.C - 1) / p.C;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  {
    // contexts are one per host thread, so two threads may make their first call on a device together
    std::lock_guard<std::mutex> guard(nw_static_mutex);
    if (nw_static_ptr[ctx->device] == nullptr) {
      double* dptr = nullptr;
      BMFM_CUDA(ctx, cudaGetSymbolAddress((void**)&dptr, nw_static_scratch));
      nw_static_ptr[ctx->device] = dptr;
    }
  }
  p.lam = pow(0.5, 1.0 / tao);
  p.log_lam = log(p.lam);
  p.fr = factor_ret; p.cov = cov; p.status = status;
  p.t0 = 0; p.t1 = T;
  p.row0 = 0; p.carry = nullptr; p.n_peer = 0; p.rank = 0; p.flags.n = 0;
  p.state = nw_static_ptr[ctx->device];
  return nw_run(ctx, p);
}

// ---------------------------------------------------------------------------------------------
// Date-sharded Newey-West (one rank per block of dates) without the growing prefix
// ---------------------------------------------------------------------------------------------
// bmfm_newey_west_by_time_range lets a rank emit its own dates of a gathered series, but every rank still
// reduces ALL earlier dates, so that part grows with the number of ranks.  Here every rank reduces only its
// own dates from a zero state and the ranks exchange what the recurrences need — one state record
// (nstate = (1+q)(K^2+1) + (1+2q)K doubles, 44 KB at K = 42, q = 2) per rank:
//   1. bmfm_newey_west_shard_totals: chunk totals + scan of this rank's dates [t0, t1) of the gathered
//      series (rows t0-q .. t0-1 are read for the lagged products); the state after its last row is stored
//      into slot `rank` of every peer's totals buffer (peer_totals: HOST array of n_peer device pointers to
//      [n_rank * nstate] doubles, peer-mapped, e.g. symmetric memory).
//   2. the caller runs a cross-rank barrier;
//   3. bmfm_newey_west_shard_emit: the decayed sum of the earlier ranks' totals (this rank's own copy
//      `totals`, range_ends[r] = one past rank r's last date, HOST array) is this rank's carry-in; the
//      covariances of [t0, t1) follow exactly as in the single-GPU pass.
// Both calls must use the same context (the chunk states stay in its workspace in between).
static int nw_shard_params(bmfm_ctx* ctx, NwParams& p, int T_total, int K, int q, double tao, const double* fr_all,
                           int t0, int t1, double** carry_out) {
  if (q > NW_MAX_Q) {
    snprintf(ctx->err, sizeof(ctx->err), "q=%d exceeds compiled limit %d", q, NW_MAX_Q);
    return BMFM_E_UNSUPPORTED;
  }
  const int n = t1 - t0;
  p.T = t1; p.K = K; p.q = q;
  p.C = nw_pick_chunk(n > 0 ? n : 1, K, q);
  p.nchunks = (n + p.C - 1) / p.C;
  p.lam = pow(0.5, 1.0 / tao);
  p.log_lam = log(p.lam);
  p.fr = fr_all; p.cov = nullptr; p.status = nullptr;
  p.row0 = t0; p.t0 = t0; p.t1 = t1;
  p.carry = nullptr; p.n_peer = 0; p.rank = 0; p.flags.n = 0;
  const size_t ns = (size_t)nw_nstate(K, q);
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int rc = bmfm_ws_reserve(ctx, ((size_t)(p.nchunks > 0 ? p.nchunks : 1) * ns + ns) * sizeof(double));
  if (rc) return rc;
  p.state = static_cast<double*>(ctx->ws);
  *carry_out = p.state + (size_t)(p.nchunks > 0 ? p.nchunks : 1) * ns;
  (void)T_total;
  return BMFM_OK;
}

extern "C" size_t bmfm_newey_west_state_doubles(int K, int q) { return (size_t)nw_nstate(K, q); }

extern "C" int bmfm_newey_west_shard_totals(bmfm_ctx* ctx, int T_total, int K, int q, double tao,
                                            const double* factor_ret_all, int t0, int t1, int rank, int n_peer,
                                            double* const* peer_totals) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T_total >= 0 && K > 0 && q >= 0 && tao > 0 && t0 >= 0 && t1 >= t0 && t1 <= T_total, "bad sizes");
  BMFM_ARG(ctx, factor_ret_all && n_peer >= 1 && n_peer <= 8 && peer_totals && rank >= 0, "bad peer list");
  NwParams p;
  double* carry = nullptr;
  int rc = nw_shard_params(ctx, p, T_total, K, q, tao, factor_ret_all, t0, t1, &carry);
  if (rc) return rc;
  p.n_peer = n_peer; p.rank = rank;
  for (int r = 0; r < 8; ++r) p.peer_totals[r] = r < n_peer ? peer_totals[r] : nullptr;
  p.flags = bmfm_flag_peers(ctx, rank, n_peer);
  return nw_run_totals(ctx, p);
}

extern "C" int bmfm_newey_west_shard_emit(bmfm_ctx* ctx, int T_total, int K, int q, double tao,
                                          const double* factor_ret_all, int t0, int t1, int rank,
                                          const int64_t* range_ends, const double* totals, double* cov,
                                          int32_t* status) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T_total >= 0 && K > 0 && q >= 0 && tao > 0 && t0 >= 0 && t1 >= t0 && t1 <= T_total, "bad sizes");
  BMFM_ARG(ctx, factor_ret_all && rank >= 0 && rank < 8 && (rank == 0 || (range_ends && totals)), "bad rank data");
  BMFM_ARG(ctx, t1 == t0 || (cov && status), "NULL pointer");
  if (t1 == t0) return BMFM_OK;
  NwParams p;
  double* carry = nullptr;
  int rc = nw_shard_params(ctx, p, T_total, K, q, tao, factor_ret_all, t0, t1, &carry);
  if (rc) return rc;
  p.cov = cov; p.status = status;
  if (rank > 0) {
    NwEnds ends;
    for (int r = 0; r < 8; ++r) ends.v[r] = r < rank ? (long long)range_ends[r] : 0;
    const int ns = nw_nstate(K, q);
    nw_carry_kernel<<<(ns + 127) / 128, 128, 0, ctx->stream>>>(ns, rank, p.lam, totals, ends, (long long)t0, carry,
                                                               bmfm_flag_peers(ctx, rank, 0));
    ctx->launches++;
    p.carry = carry;
  }
  return nw_run_emit(ctx, p);
}
