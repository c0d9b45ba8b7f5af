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
#include "bmfm_common.cuh"

namespace bmfm {

constexpr int NW_MAX_Q = 4;
constexpr int NW_EMIT_SPLIT = 4;   // CTAs per chunk in the emit kernel (the K*K elements are split)

struct NwParams {
  int T, K, q, C, nchunks;
  int t0, t1, c_first;   // emit only dates [t0, t1) (chunks from c_first on), written at row t - t0
  int c_used;            // chunks 0 .. c_used-1 are needed (those up to the last emitted one)
  double lam;
  const double* fr;   // [T][K]
  double* cov;        // [T][K][K]
  int* status;        // [T]
  double* state;      // [nchunks][nstate]   chunk totals, then chunk start states
};

__host__ __device__ inline int nw_nstate(int K, int q) { return (1 + q) * (K * K + 1) + (1 + 2 * q) * K; }
// layout of one state record:
//   mats[(1+q)][K][K] | vecs[(1+2q)][K] (S, R_1..R_q, P_1..P_q) | scal[(1+q)] (Z, Z_1..Z_q)
__device__ __forceinline__ double nw_vec_term(const double* fr, int K, int q, int v, int r, int k) {
  // the value row r adds to vector v
  if (v == 0) return fr[(size_t)r * K + k];
  if (v <= q) return r >= v ? fr[(size_t)r * K + k] : 0.0;
  const int i = v - q;
  return r - i >= 0 ? fr[(size_t)(r - i) * K + k] : 0.0;
}

// kernel 1: chunk totals from a zero start.  The chunk's rows (plus q rows of history for the lagged
// products) are staged in shared memory once.
__global__ void __launch_bounds__(256) nw_chunk_local(NwParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sx = reinterpret_cast<double*>(smem_raw);   // [(C+q)][K], rows r0-q .. r1-1, zeros before row 0
  const int c = blockIdx.x, K = p.K, q = p.q;
  const int r0 = c * p.C, r1 = min(p.T, r0 + p.C), len = r1 - r0;
  double* st = p.state + (size_t)c * nw_nstate(K, q);
  for (int e = threadIdx.x; e < (len + q) * K; e += blockDim.x) {
    const int j = e / K, k = e - j * K;
    const int r = r0 - q + j;
    sx[e] = r >= 0 ? p.fr[(size_t)r * K + k] : 0.0;
  }
  __syncthreads();
  const double lam = p.lam;
  const int nm = (1 + q) * K * K;
  for (int e = threadIdx.x; e < nm; e += blockDim.x) {
    const int i = e / (K * K), ab = e - i * K * K, a = ab / K, b = ab - a * K;
    double s = 0.0;
    for (int j = 0; j < len; ++j) s = fma(s, lam, sx[(q + j - i) * K + a] * sx[(q + j) * K + b]);   // zero rows before row 0
    st[e] = s;
  }
  const int nv = (1 + 2 * q) * K;
  for (int e = threadIdx.x; e < nv; e += blockDim.x) {
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
  }
  if (threadIdx.x <= q) {
    const int i = threadIdx.x;
    double s = 0.0;
    for (int r = r0; r < r1; ++r) {
      s *= lam;
      if (r >= i) s += 1.0;
    }
    st[nm + nv + i] = s;
  }
}

// kernel 2: exclusive scan over chunks, one thread per state element.  Every chunk but the
// last has length C (the total after the last chunk is never used), so the carry factor is
// the constant lam^C; loads are issued 32 chunks at a time to overlap their latency.
__global__ void nw_chunk_scan(NwParams p) {
  const int ns = nw_nstate(p.K, p.q);
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= ns) return;
  const double fC = pow(p.lam, (double)p.C);
  double run = 0.0;
  for (int c0 = 0; c0 < p.c_used; c0 += 32) {
    double loc[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) loc[j] = (c0 + j < p.c_used) ? p.state[(size_t)(c0 + j) * ns + e] : 0.0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (c0 + j < p.c_used) {
        p.state[(size_t)(c0 + j) * ns + e] = run;
        run = fma(run, fC, loc[j]);
      }
    }
  }
}

// kernel 3: replay the chunk from its start state and emit V for every date in it
__global__ void __launch_bounds__(256) nw_emit(NwParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int c = p.c_first + blockIdx.x, K = p.K, q = p.q, C = p.C;
  const int r0 = c * C, r1 = min(p.T, r0 + C), len = r1 - r0;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int nm = (1 + q) * K * K;
  const double* st = p.state + (size_t)c * nw_nstate(K, q);
  const int NV = 1 + 2 * q, nv = NV * K;
  double* sx = reinterpret_cast<double*>(smem_raw);   // [(C+q)][K]  rows r0-q .. r1-1 (zeros before row 0)
  double* sv = sx + (C + q) * K;                       // [C][NV][K]  vectors after row r0+j
  double* sz = sv + C * nv;                            // [C][(1+q)]
  const double lam = p.lam;
  for (int e = tid; e < (len + q) * K; e += nt) {
    const int j = e / K, k = e - j * K;
    const int r = r0 - q + j;
    sx[e] = r >= 0 ? p.fr[(size_t)r * K + k] : 0.0;
  }
  __syncthreads();
  for (int e = tid; e < nv; e += nt) {
    const int v = e / K, k = e - v * K;
    double s = st[nm + e];
    for (int j = 0; j < len; ++j) {
      const int r = r0 + j;
      double add;
      if (v == 0) add = sx[(q + j) * K + k];
      else if (v <= q) add = r >= v ? sx[(q + j) * K + k] : 0.0;
      else add = sx[(q + j - (v - q)) * K + k];      // zero-filled before row 0
      s = fma(s, lam, add);
      sv[(size_t)j * nv + e] = s;
    }
  }
  if (tid <= q) {
    const int i = tid;
    double s = st[nm + nv + i];
    for (int j = 0; j < len; ++j) {
      const int r = r0 + j;
      s *= lam;
      if (r >= i) s += 1.0;
      sz[j * (1 + q) + i] = s;
    }
  }
  __syncthreads();
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int e = blockIdx.y * nt + tid; e < K * K; e += gridDim.y * nt) {
    const int a = e / K, b = e - a * K;
    double a0 = st[e];
    double bab[NW_MAX_Q], bba[NW_MAX_Q];
#pragma unroll
    for (int i = 0; i < NW_MAX_Q; ++i) {
      bab[i] = i < q ? st[(1 + i) * K * K + a * K + b] : 0.0;
      bba[i] = i < q ? st[(1 + i) * K * K + b * K + a] : 0.0;
    }
    for (int j = 0; j < len; ++j) {
      const int r = r0 + j;       // date index; window length L = r + 1
      const double* xr = sx + (q + j) * K;
      const double xa = xr[a], xb = xr[b];
      a0 = fma(a0, lam, xa * xb);
#pragma unroll
      for (int i = 0; i < NW_MAX_Q; ++i) {
        if (i < q) {
          bab[i] *= lam;
          bba[i] *= lam;
          if (r - (i + 1) >= 0) {
            const double* xl = sx + (q + j - (i + 1)) * K;
            bab[i] = fma(xl[a], xb, bab[i]);
            bba[i] = fma(xl[b], xa, bba[i]);
          }
        }
      }
      const int L = r + 1;
      double out;
      if (L <= q || L <= K) {
        out = nan;
      } else {
        const double* vec = sv + (size_t)j * nv;
        const double Z = sz[j * (1 + q)];
        const double iz = 1.0 / Z;
        const double ma = vec[a] * iz, mb = vec[b] * iz;
        out = a0 * iz - ma * mb;
#pragma unroll
        for (int i = 0; i < NW_MAX_Q; ++i) {
          if (i < q) {
            const int lag = i + 1;
            const double Pa = vec[(q + lag) * K + a], Pb = vec[(q + lag) * K + b];
            const double Ra = vec[lag * K + a], Rb = vec[lag * K + b];
            const double Zi = sz[j * (1 + q) + lag];
            const double gab = (bab[i] - Pa * mb - ma * Rb + ma * mb * Zi) * iz;
            const double gba = (bba[i] - Pb * ma - mb * Ra + ma * mb * Zi) * iz;
            out = fma(1.0 - (double)lag / (1.0 + q), gab + gba, out);
          }
        }
      }
      if (r >= p.t0 && r < p.t1) st_stream(p.cov + ((size_t)(r - p.t0) * K + a) * K + b, out);
    }
  }
  if (blockIdx.y == 0)
  for (int j = tid; j < len; j += nt) {
    const int r = r0 + j, L = r + 1;
    if (r >= p.t0 && r < p.t1) p.status[r - p.t0] = (L <= q || L <= K) ? BMFM_ST_SHORT : BMFM_ST_OK;
  }
}

}  // namespace bmfm

using namespace bmfm;

// Scratch (device) bytes bmfm_newey_west_by_time needs for given sizes.
extern "C" size_t bmfm_newey_west_scratch_bytes(int T, int K, int q) {
  if (T <= 0) return 0;
  int C = (T + 295) / 296;
  if (C < 8) C = 8;
  const int nchunks = (T + C - 1) / C;
  return (size_t)nchunks * nw_nstate(K, q) * sizeof(double);
}

static int nw_launch(bmfm_ctx* ctx, int T, int K, int q, double tao, const double* fr, double* cov,
                     int32_t* status, double* scratch, int t0, int t1) {
  NwParams p;
  p.T = T; p.K = K; p.q = q;
  p.C = (T + 295) / 296;
  if (p.C < 8) p.C = 8;
  p.nchunks = (T + p.C - 1) / p.C;
  p.lam = pow(0.5, 1.0 / tao);
  p.fr = fr; p.cov = cov; p.status = status; p.state = scratch;
  if (t0 < 0) t0 = 0;
  if (t1 > T) t1 = T;
  p.t0 = t0; p.t1 = t1;
  p.c_first = t0 / p.C;
  const int c_last = t1 > t0 ? (t1 - 1) / p.C : p.c_first - 1;
  p.c_used = c_last + 1;   // later chunks feed nothing this call emits
  {
    const size_t smem_local = (size_t)(p.C + q) * K * 8;
    if (smem_local > (size_t)ctx->max_smem_optin) {
      snprintf(ctx->err, sizeof(ctx->err), "Newey-West chunk does not fit shared memory (K=%d q=%d)", K, q);
      return BMFM_E_UNSUPPORTED;
    }
    if (smem_local > 48 * 1024)
      BMFM_CUDA(ctx, cudaFuncSetAttribute(nw_chunk_local, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_local));
    nw_chunk_local<<<p.c_used, 256, smem_local, ctx->stream>>>(p);
  }
  ctx->launches++;
  const int ns = nw_nstate(K, q);
  nw_chunk_scan<<<(ns + 127) / 128, 128, 0, ctx->stream>>>(p);
  ctx->launches++;
  const size_t smem = ((size_t)(p.C + q) * K + (size_t)p.C * (1 + 2 * q) * K + (size_t)p.C * (1 + q)) * 8;
  if (smem > (size_t)ctx->max_smem_optin) {
    snprintf(ctx->err, sizeof(ctx->err), "Newey-West chunk does not fit shared memory (K=%d q=%d)", K, q);
    return BMFM_E_UNSUPPORTED;
  }
  BMFM_CUDA(ctx, cudaFuncSetAttribute(nw_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (c_last >= p.c_first) {
    nw_emit<<<dim3(c_last - p.c_first + 1, NW_EMIT_SPLIT), 256, smem, ctx->stream>>>(p);
    ctx->launches++;
  }
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
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
  int rc = bmfm_ws_reserve(ctx, bmfm_newey_west_scratch_bytes(T, K, q));
  if (rc) return rc;
  return nw_launch(ctx, T, K, q, tao, factor_ret, cov, status, (double*)ctx->ws, t0, t1);
}

extern "C" int bmfm_newey_west_by_time(bmfm_ctx* ctx, int T, int K, int q, double tao,
                                       const double* factor_ret, double* cov, int32_t* status) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && K > 0 && q >= 0 && tao > 0, "bad sizes");
  BMFM_ARG(ctx, factor_ret && cov && status, "NULL pointer");
  if (q > NW_MAX_Q) {
    snprintf(ctx->err, sizeof(ctx->err), "q=%d exceeds compiled limit %d", q, NW_MAX_Q);
    return BMFM_E_UNSUPPORTED;
  }
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = bmfm_ws_reserve(ctx, bmfm_newey_west_scratch_bytes(T, K, q));
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

// chunk length that keeps nchunks * nstate inside the static array (0 if impossible)
static int nw_static_chunk(int T, int K, int q) {
  const size_t ns = (size_t)nw_nstate(K, q);
  const size_t max_chunks = NW_STATIC_DOUBLES / ns;
  if (max_chunks == 0) return 0;
  int C = (T + 295) / 296;
  if (C < 8) C = 8;
  const size_t need = ((size_t)T + max_chunks - 1) / max_chunks;   // smallest C with ceil(T/C) <= max_chunks
  if ((size_t)C < need) C = (int)need;
  return C;
}

// Same computation and results as bmfm_newey_west_by_time, with no cudaMalloc, cudaFree or other
// allocating call on any path.
//   * Limits: q <= 4, and (1+q)(K^2+1) + (1+2q)K doubles of state per chunk must fit the 16 MiB
//     array with chunks short enough for shared memory (K = 42, q = 2: any T up to ~35 000);
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
  const size_t smem = p.C > 0 ? ((size_t)(p.C + q) * K + (size_t)p.C * (1 + 2 * q) * K + (size_t)p.C * (1 + q)) * 8 : 0;
  if (p.C == 0 || smem > (size_t)ctx->max_smem_optin) {
    snprintf(ctx->err, sizeof(ctx->err),
             "newey_west noalloc: T=%d K=%d q=%d needs more than the fixed %zu-byte scratch; "
             "use bmfm_newey_west_by_time", T, K, q, NW_STATIC_DOUBLES * sizeof(double));
    return BMFM_E_UNSUPPORTED;
  }
  p.nchunks = (T + p.C - 1) / p.C;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (nw_static_ptr[ctx->device] == nullptr) {
    double* dptr = nullptr;
    BMFM_CUDA(ctx, cudaGetSymbolAddress((void**)&dptr, nw_static_scratch));
    nw_static_ptr[ctx->device] = dptr;
  }
  p.lam = pow(0.5, 1.0 / tao);
  p.fr = factor_ret; p.cov = cov; p.status = status;
  p.t0 = 0; p.t1 = T; p.c_first = 0; p.c_used = p.nchunks;
  p.state = nw_static_ptr[ctx->device];
  {
    const size_t smem_local = (size_t)(p.C + q) * K * 8;
    if (smem_local > (size_t)ctx->max_smem_optin) {
      snprintf(ctx->err, sizeof(ctx->err), "Newey-West chunk does not fit shared memory (K=%d q=%d)", K, q);
      return BMFM_E_UNSUPPORTED;
    }
    if (smem_local > 48 * 1024)
      BMFM_CUDA(ctx, cudaFuncSetAttribute(nw_chunk_local, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_local));
    nw_chunk_local<<<p.c_used, 256, smem_local, ctx->stream>>>(p);
  }
  ctx->launches++;
  const int ns = nw_nstate(K, q);
  nw_chunk_scan<<<(ns + 127) / 128, 128, 0, ctx->stream>>>(p);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaFuncSetAttribute(nw_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  nw_emit<<<dim3(p.nchunks, NW_EMIT_SPLIT), 256, smem, ctx->stream>>>(p);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
