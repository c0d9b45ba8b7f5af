// Tensor-core streaming kernel of the per-date regression: one WARP per date, like reg_stream_kernel,
// but the weighted Gram of pass A is formed by FP64 MMA (mma.sync m8n8k4 -> DMMA.8x8x4) instead of 97
// register accumulators per lane.  Included by reg_by_time.cu (inside namespace bmfm).
//
// Per 32-row chunk of the cp.async ring:
//   stage   lane = row: w = sqrt(cap), z = [w, w y, cap, w d_1..w d_Q] (d = style - first row) -> shared
//   gram    8 k-steps of 4 rows.  Left vector  L = [w d_1..w d_Q, cap, 1]   (A fragments, read from z)
//                                 right vector R = [d_1..d_Q, u, 1]         (B fragments, raw stage - shift)
//           laid out in 8-wide tiles so that three 8x8 products (one when Q <= 6) hold every sum the solve
//           needs: w d d' (lower triangle), w d u, w d, cap d, sum d, sum u.  Accumulators: 6 doubles a lane.
//   buckets lane = row: rows of the chunk that share an industry are found with one match.any; the first
//           lane of each group adds the group's z rows and updates the per-warp table — no atomics, and
//           the summation order is fixed (deterministic results).
// The solve and pass C are the ones of reg_stream_kernel.  ~90 registers instead of 255, so the SM holds
// 12-16 warps instead of 8 (shared memory per warp is the limit) and each warp has far more independent
// work in flight.
#pragma once

constexpr int RM_WARPS = 4;   // warps per CTA (= warps per date in the CTA-per-date form)

__host__ __device__ constexpr int rm_zs(int Q) { return (3 + Q) | 1; }   // odd row stride of the z staging

// Shared memory: a "solve" block (accumulator vector, coefficients, factor returns, flag) and a "warp" block
// (bucket table, z staging, constants, ring).  WPD = warps per date: with WPD == 1 every warp has both blocks and
// the solve scratch aliases the z staging; with WPD > 1 the CTA has one solve block and WPD warp blocks.
struct RmSolveBlk { int acc, coef, fvec, flag, solve, red, total; };
struct RmWarpBlk { int bucket, z, cst, ring, total; };

template <int Q>
__host__ __device__ inline RmSolveBlk rm_solve_blk(int P, int wpd) {
  const int NB = P + 1, K = 1 + P + Q;
  RmSolveBlk s;
  int o = 0;
  s.acc = o;    o = align16(o + (na1pad_of(Q) + na2pad_of(Q)) * 8);
  s.coef = o;   o = align16(o + (3 * Q + NB + 32) * 8);
  s.fvec = o;   o = align16(o + (K + 2) * 8);
  s.flag = o;   o += 16;
  s.solve = o;  if (wpd > 1) o = align16(o + (Q * Q + 7 * Q + 2 * NB + 8) * 8);
  s.red = o;    if (wpd > 1) o = align16(o + 2 * wpd * 8);
  if (o < (3 * NB + 2) * 4) o = align16((3 * NB + 2) * 4);   // keeps the flag's RegSmem-style offset non-negative
  s.total = o;
  return s;
}

template <int Q, int NSTAGE>
__host__ __device__ inline RmWarpBlk rm_warp_blk(int P) {
  const int NB = P + 1, BV = 3 + Q;
  RmWarpBlk s;
  int o = 0;
  s.bucket = o; o = align16(o + NB * BV * 8);
  const int solve = (Q * Q + 7 * Q + 2 * NB + 8) * 8, zst = 32 * rm_zs(Q) * 8, cf = 8 * 32 * 8;
  int u = solve > zst ? solve : zst;
  if (cf > u) u = cf;
  s.z = o;      o = align16(o + u);                     // z staging; later solve scratch (WPD 1) / fragment hand-over (WPD > 1)
  s.cst = o;    o = align16(o + 16 + rm_zs(Q) * 8 + 32);     // {1.0, 0.0}, a z row of zeros, bucket of local id [8]
  s.ring = o;   o = align16(o + NSTAGE * rs_stage_bytes<Q>());
  s.total = o;
  return s;
}

template <int Q, int NSTAGE>
__host__ __device__ inline int rm_cta_bytes(int P, int wpd, int warps) {
  const int sb = rm_solve_blk<Q>(P, wpd).total, wb = rm_warp_blk<Q, NSTAGE>(P).total;
  return wpd > 1 ? sb + wpd * wb : warps * (sb + wb);
}

__device__ __forceinline__ double lds_f64(unsigned addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(v) : "r"(addr) : "memory");
  return v;
}

__device__ __forceinline__ void dmma_884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// what row m of left tile i / column n of right tile j stands for
enum { RM_NONE = -1, RM_CAP = -2, RM_ONE = -3, RM_U = -4, RM_W = -5, RM_WY = -6 };   // >= 0: style index
// 7 <= Q <= 10: the second left tile has room for w and w y, and the per-industry sums come from two more
// 8x8 products with the chunk's industry indicator columns (see the kernel)
__host__ __device__ constexpr bool rm_has_s(int Q) { return Q >= 7 && Q <= 10; }
template <int Q>
__device__ __forceinline__ int rm_left(int i, int m) {
  constexpr int Q0 = Q < 6 ? Q : 6, Q1 = Q - Q0;
  constexpr bool TWO = Q > 6;
  if (i == 0) {
    if (m < Q0) return m;
    if (!TWO && m == 6) return RM_CAP;
    if (!TWO && m == 7) return RM_ONE;
    return RM_NONE;
  }
  if (m < Q1) return 6 + m;
  if (m == Q1) return RM_CAP;
  if (m == Q1 + 1) return RM_ONE;
  if (rm_has_s(Q) && m == Q1 + 2) return RM_W;
  if (rm_has_s(Q) && m == Q1 + 3) return RM_WY;
  return RM_NONE;
}
template <int Q>
__device__ __forceinline__ int rm_right(int j, int n) {
  constexpr int Q0 = Q < 6 ? Q : 6, Q1 = Q - Q0;
  if (j == 0) {
    if (n < Q0) return n;
    if (n == 6) return RM_U;
    if (n == 7) return RM_ONE;
    return RM_NONE;
  }
  return n < Q1 ? 6 + n : RM_NONE;
}

// one lane's two accumulators of tile (i, j) -> the s_acc slots the solve reads
template <int Q>
__device__ __forceinline__ void rm_scatter(int i, int j, int lane, double c0, double c1, double ypiv, double* s_acc) {
  constexpr int NA1PAD = na1pad_of(Q);
  double* S2 = s_acc;
  double* BY = s_acc + Q * (Q + 1) / 2;
  double* CD = s_acc + NA1PAD;
  double* D1 = CD + Q;
  const int m = lane >> 2, n0 = 2 * (lane & 3);
  const int lt = rm_left<Q>(i, m);
  if (lt == RM_NONE || lt == RM_W || lt == RM_WY) return;
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int rt = rm_right<Q>(j, n0 + e);
    const double v = e ? c1 : c0;
    if (lt >= 0) {
      if (rt >= 0) { if (rt <= lt) S2[lt * (lt + 1) / 2 + rt] = v; }
      else if (rt == RM_U) BY[lt] = c0 + ypiv * c1;      // sum w d y = sum w d u + ypiv sum w d (columns 6, 7)
    } else if (lt == RM_CAP) {
      if (rt >= 0) CD[rt] = v;
    } else {   // RM_ONE
      if (rt >= 0) D1[rt] = v;
      else if (rt == RM_U) D1[2 * Q] = v;                // Y1 = sum u, at s_acc[NA1PAD + 3Q]
    }
  }
}


template <int Q, int NSTAGE, int WPD>
__global__ void __launch_bounds__(RM_WARPS * 32, (NSTAGE >= 3 && WPD == 1) ? 3 : 4) reg_mma_kernel(RegParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr bool CTA = WPD > 1;            // WPD warps (one CTA) share a date; otherwise one warp per date
  constexpr int NA1PAD = na1pad_of(Q);
  constexpr int QQ = Q > 0 ? Q : 1;
  constexpr int SB = rs_stage_bytes<Q>();
  constexpr int ZS = rm_zs(Q), BV = 3 + Q;
  constexpr int Q0 = Q < 6 ? Q : 6, Q1 = Q - Q0;
  constexpr bool TWO = Q > 6;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int tid = CTA ? (int)threadIdx.x : lane, nt = CTA ? (int)blockDim.x : 32;
  const int P = p.P, NB = P + 1, K = p.K;
  const RmSolveBlk SK = rm_solve_blk<Q>(P, WPD);
  const RmWarpBlk WK = rm_warp_blk<Q, NSTAGE>(P);
  unsigned char* sbase = CTA ? smem : smem + (size_t)warp * (SK.total + WK.total);
  unsigned char* wbase = CTA ? smem + SK.total + (size_t)warp * WK.total : sbase + SK.total;
  RegSmem L;   // the view solve_fast_cta expects, offsets from sbase (bucket table: warp 0's in the CTA form)
  L.acc = SK.acc; L.coef = SK.coef; L.fvec = SK.fvec;
  L.seg = SK.flag - (3 * NB + 2) * 4;
  L.bucket = SK.total + WK.bucket;
  L.solve = CTA ? SK.solve : SK.total + WK.z;
  unsigned char* ring = wbase + WK.ring;
  double* s_z = reinterpret_cast<double*>(wbase + WK.z);
  double* s_cst = reinterpret_cast<double*>(wbase + WK.cst);
  double* s_bucket = reinterpret_cast<double*>(wbase + WK.bucket);
  double* s_acc = reinterpret_cast<double*>(sbase + SK.acc);
  double* s_piv = reinterpret_cast<double*>(sbase + SK.coef);
  double* s_beta = s_piv + 2 * Q;
  double* s_g = s_beta + Q;
  double* s_misc = s_g + NB;
  double* s_red = reinterpret_cast<double*>(sbase + SK.red);
  const double* s_f = reinterpret_cast<const double*>(sbase + SK.fvec);
  if (lane == 0) { s_cst[0] = 1.0; s_cst[1] = 0.0; }
  if (lane < ZS) s_cst[2 + lane] = 0.0;
  const double* s_zero = s_cst + 2;
  int* s_runb = reinterpret_cast<int*>(s_cst + 2 + ZS);
  constexpr bool HAS_S = rm_has_s(Q);

  // fragment sources of this lane (row k = lane & 3 of a k-step, element c = lane >> 2 of a tile):
  // byte offset relative to the stage (raw inputs) or an absolute shared address (z staging, constants)
  const int c = lane >> 2, kq = lane & 3;
  const unsigned z_sa = (unsigned)__cvta_generic_to_shared(s_z);
  const unsigned one_sa = (unsigned)__cvta_generic_to_shared(s_cst), zero_sa = one_sa + 8;
  constexpr int OFF_RET = 32 * Q + 32;   // doubles inside a stage
  bool b0_st; unsigned b0_a, b0_s; int b0_sub;             // sub: -1 none, -2 one (-> 0), -3 ypiv, >= 0 piv index
  if (c < Q0)       { b0_st = true;  b0_a = (c + kq * Q) * 8;      b0_s = 4 * Q * 8; b0_sub = c; }
  else if (c == 6)  { b0_st = true;  b0_a = (OFF_RET + kq) * 8;    b0_s = 4 * 8;     b0_sub = -3; }
  else if (c == 7)  { b0_st = false; b0_a = one_sa;                b0_s = 0;         b0_sub = -1; }
  else              { b0_st = false; b0_a = one_sa;                b0_s = 0;         b0_sub = -2; }
  bool b1_st; unsigned b1_a, b1_s; int b1_sub;
  if (c < Q1)       { b1_st = true;  b1_a = (6 + c + kq * Q) * 8;  b1_s = 4 * Q * 8; b1_sub = 6 + c; }
  else              { b1_st = false; b1_a = one_sa;                b1_s = 0;         b1_sub = -2; }
  unsigned a0_a, a0_s, a1_a, a1_s;
  {
    const int l0 = rm_left<Q>(0, c), l1 = rm_left<Q>(1, c);
    if (l0 >= 0)            { a0_a = z_sa + (3 + l0 + kq * ZS) * 8; a0_s = 4 * ZS * 8; }
    else if (l0 == RM_CAP)  { a0_a = z_sa + (2 + kq * ZS) * 8;      a0_s = 4 * ZS * 8; }
    else if (l0 == RM_ONE)  { a0_a = one_sa;  a0_s = 0; }
    else                    { a0_a = zero_sa; a0_s = 0; }
    if (l1 >= 0)            { a1_a = z_sa + (3 + l1 + kq * ZS) * 8; a1_s = 4 * ZS * 8; }
    else if (l1 == RM_CAP)  { a1_a = z_sa + (2 + kq * ZS) * 8;      a1_s = 4 * ZS * 8; }
    else if (l1 == RM_W)    { a1_a = z_sa + (0 + kq * ZS) * 8;      a1_s = 4 * ZS * 8; }
    else if (l1 == RM_WY)   { a1_a = z_sa + (1 + kq * ZS) * 8;      a1_s = 4 * ZS * 8; }
    else if (l1 == RM_ONE)  { a1_a = one_sa;  a1_s = 0; }
    else                    { a1_a = zero_sa; a1_s = 0; }
  }
  int val0 = -1, val1 = -1;
  {
    const int l0 = rm_left<Q>(0, c), l1 = rm_left<Q>(1, c);
    val0 = l0 >= 0 ? 3 + l0 : -1;
    val1 = l1 >= 0 ? 3 + l1 : (l1 == RM_CAP ? 2 : (l1 == RM_W ? 0 : (l1 == RM_WY ? 1 : -1)));
  }
  __syncwarp();

  const int wq = CTA ? warp : 0;                       // this warp takes chunks wq, wq + WPD, ...
  const int t_first = CTA ? (int)blockIdx.x : (int)blockIdx.x * nwarps + warp;
  const int t_step = CTA ? (int)gridDim.x : (int)gridDim.x * nwarps;
  // Dates beyond the first one of each warp (CTA) are handed out through a global counter in the order the warps
  // finish: with T not a multiple of the number of warp slots a static stride would leave some SMs with twice the
  // second-round work of others.
  int* s_next = reinterpret_cast<int*>(sbase + SK.flag) + 2;
  auto next_date = [&](int t) -> int {
    if (p.next_date == nullptr) return t + t_step;
    if (CTA) {
      __syncthreads();
      if (tid == 0) *s_next = t_step + atomicAdd(p.next_date, 1);
      __syncthreads();
      return *s_next;
    }
    int v = 0;
    if (lane == 0) v = t_step + atomicAdd(p.next_date, 1);
    return __shfl_sync(0xffffffffu, v, 0);
  };
  for (int t = t_first; t < p.T; t = next_date(t)) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    if (n <= 0) {
      for (int k = tid; k < K; k += nt) store_factor_ret(p, t, k, nan);
      if (tid == 0) { p.r2[t] = nan; p.status[t] = BMFM_ST_NONFINITE; }
      continue;
    }
    const int nchunks = (n + 31) >> 5;
    const int myn = nchunks > wq ? (nchunks - wq + WPD - 1) / WPD : 0;   // chunks of this warp
    for (int e = lane; e < NB * BV; e += 32) s_bucket[e] = 0.0;
#pragma unroll
    for (int k = 0; k < NSTAGE - 1; ++k) {
      const int cc = wq + k * WPD;
      if (k < myn) rs_issue_chunk<Q>(p, a + 32 * cc, min(32, n - 32 * cc), true, ring + k * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    // the pivot shift is the date's first row
    if (CTA) {
      if (warp == 0) {
        if (lane < Q) s_piv[lane] = p.style[a * Q + lane];
        if (lane == 0) s_misc[0] = p.ret[a];
      }
      __syncthreads();
    } else {
      asm volatile("cp.async.wait_group %0;\n" ::"n"(NSTAGE - 2) : "memory");
      __syncwarp();
      const double* st0 = reinterpret_cast<const double*>(ring);
      if (lane < Q) s_piv[lane] = st0[lane];
      if (lane == 0) s_misc[0] = st0[32 * Q + 32];
      __syncwarp();
    }
    const double ypiv = s_misc[0];
    const double sub0 = b0_sub >= 0 ? s_piv[b0_sub] : (b0_sub == -3 ? ypiv : (b0_sub == -2 ? 1.0 : 0.0));
    const double sub1 = b1_sub >= 0 ? s_piv[b1_sub] : 1.0;
    double c00a = 0.0, c00b = 0.0, c10a = 0.0, c10b = 0.0, c11a = 0.0, c11b = 0.0, sq0 = 0.0, sq1 = 0.0;

    for (int k = 0; k < myn; ++k) {
      const int cc = wq + k * WPD;
      const int kn = k + NSTAGE - 1, cn = wq + kn * WPD;
      if (kn < myn) rs_issue_chunk<Q>(p, a + 32 * cn, min(32, n - 32 * cn), true, ring + (kn % NSTAGE) * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
      asm volatile("cp.async.wait_group %0;\n" ::"n"(NSTAGE - 1) : "memory");
      __syncwarp();
      unsigned char* stage = ring + (k % NSTAGE) * SB;
      double* st_style = reinterpret_cast<double*>(stage);
      double* st_cap = st_style + 32 * Q;
      double* st_ret = st_cap + 32;
      const int* st_ind = reinterpret_cast<const int*>(st_ret + 32);
      const int m = min(32, n - 32 * cc);
      // stage: per-row terms z (lane = row); rows past the end become exact zeros in every sum
      {
        double* zr = s_z + lane * ZS;
        if (lane < m) {
          const double cap = st_cap[lane];
          const double w = sqrt(cap);
          const double y = st_ret[lane];
          zr[0] = w;
          zr[1] = w * y;
          zr[2] = cap;
          const double u = y - ypiv;
          sq1 = fma(u, u, sq1);
          if (Q > 0) {
            double d[QQ];
            load_row_styles<Q>(st_style, lane, s_piv, d);
#pragma unroll
            for (int q = 0; q < Q; ++q) { zr[3 + q] = w * d[q]; sq0 = fma(d[q], d[q], sq0); }
          }
        } else {
#pragma unroll
          for (int v = 0; v < BV; ++v) zr[v] = 0.0;
#pragma unroll
          for (int q = 0; q < Q; ++q) st_style[lane * Q + q] = s_piv[q];
          st_ret[lane] = ypiv;
        }
      }
      __syncwarp();
      // industries of the chunk: groups of equal industry (match.any), their number G and a local id per row
      const int myb = lane < m ? bucket_of(st_ind[lane], P) : 64 + lane;
      const unsigned peers = __match_any_sync(0xffffffffu, myb);
      const int myleader = __ffs(peers) - 1;
      const unsigned leaders = __ballot_sync(0xffffffffu, lane < m && myleader == lane);
      const int G = __popc(leaders);
      const bool use_s = HAS_S && p.smma && G <= 8;                // warp-uniform
      const int lid = lane < m ? __popc(leaders & ((1u << myleader) - 1u)) : 99;
      if (use_s && lane < m && myleader == lane) s_runb[lid] = myb;
      // gram: 8 k-steps of 4 rows; with use_s two more products against the indicator columns of the chunk's
      // (at most 8) industries give the per-industry sums of w d, cap, w, w y
      double cs0a = 0.0, cs0b = 0.0, cs1a = 0.0, cs1b = 0.0;
      {
        const unsigned st_sa = (unsigned)__cvta_generic_to_shared(stage);
        const unsigned pb0 = b0_st ? st_sa + b0_a : b0_a;
        const unsigned pb1 = b1_st ? st_sa + b1_a : b1_a;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const double b0 = lds_f64(pb0 + ks * b0_s) - sub0;
          const double a0 = lds_f64(a0_a + ks * a0_s);
          dmma_884(c00a, c00b, a0, b0);
          if (TWO) {
            const double b1 = lds_f64(pb1 + ks * b1_s) - sub1;
            const double a1 = lds_f64(a1_a + ks * a1_s);
            dmma_884(c10a, c10b, a1, b0);
            dmma_884(c11a, c11b, a1, b1);
            if (HAS_S) {
              const int rl = __shfl_sync(0xffffffffu, lid, 4 * ks + kq);
              if (use_s) {
                const double sb = rl == c ? 1.0 : 0.0;
                dmma_884(cs0a, cs0b, a0, sb);
                dmma_884(cs1a, cs1b, a1, sb);
              }
            }
          }
        }
      }
      if (use_s) {
        __syncwarp();
        const int n0 = 2 * kq;
        const int bn0 = s_runb[n0 & 7], bn1 = s_runb[(n0 + 1) & 7];
        if (val0 >= 0) {
          if (n0 < G) s_bucket[bn0 * BV + val0] += cs0a;
          if (n0 + 1 < G) s_bucket[bn1 * BV + val0] += cs0b;
        }
        if (val1 >= 0) {
          if (n0 < G) s_bucket[bn0 * BV + val1] += cs1a;
          if (n0 + 1 < G) s_bucket[bn1 * BV + val1] += cs1b;
        }
      } else
      // per-industry sums of the chunk's z rows, no atomics (the table is this warp's own)
      {
        const int prevb = __shfl_up_sync(0xffffffffu, myb, 1);
        const unsigned runstart = __ballot_sync(0xffffffffu, lane < m && (lane == 0 || myb != prevb));
        const int nruns = __popc(runstart);
        bool by_runs = nruns <= 3;
        if (!by_runs) {
          const int maxg = __reduce_max_sync(0xffffffffu, __popc(peers));
          by_runs = 15 * nruns + 70 <= 26 * maxg + 39;       // estimated instruction counts of the two ways
        }
        if (by_runs) {
          // rows grouped by industry (what the packer produces): column sums over each run of equal
          // industries, lane = (row parity, value column)
          const int h = lane >> 4, v = lane & 15;
          const bool vact = v < BV;
          unsigned rs = runstart;
          while (rs) {
            const int s = __ffs(rs) - 1;
            rs &= rs - 1;
            const int e = rs ? __ffs(rs) - 1 : m;
            double acc = 0.0;
            const double* zc = s_z + (s + h) * ZS + (vact ? v : 0);
            const int half = (e - s + 1) >> 1;
            for (int i = 0; i < half; ++i)
              if (s + h + 2 * i < e) acc += zc[2 * i * ZS];
            acc += __shfl_xor_sync(0xffffffffu, acc, 16);
            const int b = __shfl_sync(0xffffffffu, myb, s);
            if (h == 0 && vact) s_bucket[b * BV + v] += acc;
          }
        } else {
          // arbitrary row order: the first lane of every group of equal industries adds the group's rows;
          // warp-uniform trip count (the largest group), idle lanes add a row of zeros
          const bool lead = lane < m && (__ffs(peers) - 1) == lane;
          double acc[BV];
#pragma unroll
          for (int v = 0; v < BV; ++v) acc[v] = 0.0;
          unsigned rest = lead ? peers : 0u;
          while (__any_sync(0xffffffffu, rest != 0u)) {
            const double* zj = rest ? s_z + (__ffs(rest) - 1) * ZS : s_zero;
            rest &= rest - 1;
#pragma unroll
            for (int v = 0; v < BV; ++v) acc[v] += zj[v];
          }
          if (lead) {
            double* tb = s_bucket + myb * BV;
#pragma unroll
            for (int v = 0; v < BV; ++v) tb[v] += acc[v];
          }
        }
      }
      __syncwarp();   // z and the stage may be rewritten by the next iteration
    }
    // accumulators -> the layout the solve reads
    if (CTA) {
      // hand the fragments over through the (now idle) z staging; warp 0 adds them in warp order
      double* cf = s_z;
      cf[0 * 32 + lane] = c00a; cf[1 * 32 + lane] = c00b; cf[2 * 32 + lane] = c10a; cf[3 * 32 + lane] = c10b;
      cf[4 * 32 + lane] = c11a; cf[5 * 32 + lane] = c11b; cf[6 * 32 + lane] = sq0;  cf[7 * 32 + lane] = sq1;
      __syncthreads();
      if (warp == 0) {
#pragma unroll
        for (int w = 1; w < WPD; ++w) {
          const double* o = reinterpret_cast<const double*>(wbase + (size_t)w * WK.total + WK.z);
          c00a += o[0 * 32 + lane]; c00b += o[1 * 32 + lane]; c10a += o[2 * 32 + lane]; c10b += o[3 * 32 + lane];
          c11a += o[4 * 32 + lane]; c11b += o[5 * 32 + lane]; sq0 += o[6 * 32 + lane];  sq1 += o[7 * 32 + lane];
        }
      }
      {
        double* tb0 = reinterpret_cast<double*>(smem + SK.total + WK.bucket);
        for (int e = tid; e < NB * BV; e += nt) {
          double s = tb0[e];
#pragma unroll
          for (int w = 1; w < WPD; ++w) s += reinterpret_cast<const double*>(smem + SK.total + (size_t)w * WK.total + WK.bucket)[e];
          tb0[e] = s;
        }
      }
    }
    if (!CTA || warp == 0) {
      rm_scatter<Q>(0, 0, lane, c00a, c00b, ypiv, s_acc);
      if (TWO) {
        rm_scatter<Q>(1, 0, lane, c10a, c10b, ypiv, s_acc);
        rm_scatter<Q>(1, 1, lane, c11a, c11b, ypiv, s_acc);
      }
      const double d2 = warp_sum(sq0);    // sum over rows and styles of d^2 (lane = row partial sums)
      const double y2 = warp_sum(sq1);    // sum u^2
      if (Q > 0 && lane < Q) s_acc[NA1PAD + 2 * Q + lane] = lane == 0 ? d2 : 0.0;
      if (lane == 0) s_acc[NA1PAD + 3 * Q + 1] = y2;
    }
    if (CTA) __syncthreads(); else __syncwarp();
    // start refilling the ring for pass C while the solve runs (served by L2)
#pragma unroll
    for (int k = 0; k < NSTAGE - 1; ++k) {
      const int cc = wq + k * WPD;
      if (k < myn) rs_issue_chunk<Q>(p, a + 32 * cc, min(32, n - 32 * cc), false, ring + k * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    const int st = solve_fast_cta<Q, !CTA>(p, L, sbase, n);
    if (st != BMFM_ST_OK) {
      asm volatile("cp.async.wait_group 0;\n" ::: "memory");
      __syncwarp();
      if (st == REG_NEED_DENSE) {
        if (tid == 0) p.status[t] = REG_NEED_DENSE;
      } else {
        for (int k = tid; k < K; k += nt) store_factor_ret(p, t, k, nan);
        for (int i = tid; i < n; i += nt) p.spec[a + i] = nan;
        if (tid == 0) { p.r2[t] = nan; p.status[t] = st; }
      }
      continue;
    }
    for (int k = tid; k < K; k += nt) store_factor_ret(p, t, k, s_f[k]);
    // pass C
    const double c0 = s_misc[1];
    double se = 0.0, see = 0.0;
    for (int k = 0; k < myn; ++k) {
      const int cc = wq + k * WPD;
      const int kn = k + NSTAGE - 1, cn = wq + kn * WPD;
      if (kn < myn) rs_issue_chunk<Q>(p, a + 32 * cn, min(32, n - 32 * cn), false, ring + (kn % NSTAGE) * SB, lane);
      else asm volatile("cp.async.commit_group;\n" ::: "memory");
      asm volatile("cp.async.wait_group %0;\n" ::"n"(NSTAGE - 1) : "memory");
      __syncwarp();
      unsigned char* stage = ring + (k % NSTAGE) * SB;
      const double* st_style = reinterpret_cast<const double*>(stage);
      const double* st_ret = st_style + 32 * Q + 32;
      const int* st_ind = reinterpret_cast<const int*>(st_ret + 32);
      const int m = min(32, n - 32 * cc);
      if (lane < m) {
        double e = st_ret[lane] - c0 - s_g[bucket_of(st_ind[lane], P)];
        if (Q > 0) {
          double d[QQ];
          load_row_styles<Q>(st_style, lane, s_piv, d);
#pragma unroll
          for (int q = 0; q < Q; ++q) e = fma(-s_beta[q], d[q], e);
        }
        st_stream(p.spec + a + 32 * cc + lane, e);
        se += e;
        see = fma(e, e, see);
      }
      __syncwarp();
    }
    se = warp_sum(se);
    see = warp_sum(see);
    if (CTA) {
      if (lane == 0) { s_red[2 * warp] = se; s_red[2 * warp + 1] = see; }
      __syncthreads();
      se = 0.0; see = 0.0;
      if (tid == 0) {
#pragma unroll
        for (int w = 0; w < WPD; ++w) { se += s_red[2 * w]; see += s_red[2 * w + 1]; }
      }
    }
    if (tid == 0) {
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
