// The four-warps-per-date form of reg_mma_kernel with the rows brought in by the bulk-copy engine
// (cp.async.bulk + mbarrier, SASS UBLKCP) instead of per-lane cp.async pieces.  Included by reg_by_time.cu
// after reg_mma.cuh (inside namespace bmfm).
//
// One CTA (4 warps) per date.  The date's rows move in SUPER-CHUNKS of 128 rows: one elected thread arms the
// stage's mbarrier with the byte count and issues four bulk copies (styles 10 KB, cap, ret 1 KB each, industry
// ids 512 B at Q = 10) — ~10 instructions per 128 rows where the per-lane ring needs ~100 per 32 rows and pass.
// Warp w then works on rows 32 w .. 32 w + 31 of the super-chunk exactly as reg_mma_kernel does on a chunk of
// its own ring; a CTA barrier per super-chunk releases the stage for the copy of super-chunk k + NS - 1.
//
// Bulk copies need 16-byte aligned global addresses and sizes.  A date may start at any row, so the 8-byte
// (cap, ret) and 4-byte (industry) columns are copied from the enclosing 16-byte aligned range and read with a
// constant per-date shift; style rows are Q * 8 bytes, a multiple of 16 for even Q (the kernel is built for even
// Q only).  The aligned range can reach up to 12 bytes past the super-chunk's last row: a super-chunk whose range
// would leave the panel (the last rows of the last dates), and panels whose base pointers are not 16-byte aligned,
// go through plain loads instead.
#pragma once

constexpr int RB_WARPS = 4;
constexpr int RB_SROWS = 32 * RB_WARPS;      // rows of a super-chunk

template <int Q> __host__ __device__ constexpr int rb_off_cap() { return (RB_SROWS * Q * 8 + 15) & ~15; }
template <int Q> __host__ __device__ constexpr int rb_off_ret() { return rb_off_cap<Q>() + RB_SROWS * 8 + 16; }
template <int Q> __host__ __device__ constexpr int rb_off_ind() { return rb_off_ret<Q>() + RB_SROWS * 8 + 16; }
template <int Q> __host__ __device__ constexpr int rb_stage_bytes() { return rb_off_ind<Q>() + RB_SROWS * 4 + 16; }

struct RbWarpBlk { int bucket, z, cst, total; };
template <int Q>
__host__ __device__ inline RbWarpBlk rb_warp_blk(int P) {
  const int NB = P + 1, BV = 3 + Q;
  RbWarpBlk s;
  int o = 0;
  s.bucket = o; o = align16(o + NB * BV * 8);
  const int zst = 32 * rm_zs(Q) * 8, cf = 8 * 32 * 8;
  s.z = o;      o = align16(o + (zst > cf ? zst : cf));       // z staging; later the fragment hand-over
  s.cst = o;    o = align16(o + 16 + rm_zs(Q) * 8 + 32);      // {1.0, 0.0}, a z row of zeros, bucket of local id [8]
  s.total = o;
  return s;
}
template <int Q, int NS>
__host__ __device__ inline int rb_cta_bytes(int P) {
  return rm_solve_blk<Q>(P, RB_WARPS).total + RB_WARPS * rb_warp_blk<Q>(P).total + NS * rb_stage_bytes<Q>() + 16 * NS;
}

__device__ __forceinline__ void rb_mbar_init(unsigned a, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(a), "r"(count) : "memory");
}
__device__ __forceinline__ void rb_mbar_expect_tx(unsigned a, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool rb_mbar_try_wait(unsigned a, unsigned parity) {
  unsigned ok;
  asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
               : "=r"(ok) : "r"(a), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void rb_mbar_wait(unsigned a, unsigned parity) {
  // the hardware wait sleeps up to a time limit and returns; a copy that never lands must not hang the GPU
  for (int spin = 0; !rb_mbar_try_wait(a, parity); ++spin)
    if (spin > (1 << 20)) asm volatile("trap;\n");
}
__device__ __forceinline__ void rb_bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}

template <int Q, int NS>
__global__ void __launch_bounds__(RB_WARPS * 32, 4) reg_bulk_kernel(RegParams p) {
  static_assert(Q % 2 == 0 && rm_has_s(Q), "even Q in 8..10");
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int NA1PAD = na1pad_of(Q);
  constexpr int ZS = rm_zs(Q), BV = 3 + Q;
  constexpr int Q0 = 6, Q1 = Q - Q0;
  constexpr int SSB = rb_stage_bytes<Q>();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int P = p.P, NB = P + 1, K = p.K;
  const RmSolveBlk SK = rm_solve_blk<Q>(P, RB_WARPS);
  const RbWarpBlk WK = rb_warp_blk<Q>(P);
  unsigned char* sbase = smem;
  unsigned char* wbase = smem + SK.total + (size_t)warp * WK.total;
  unsigned char* ring = smem + SK.total + (size_t)RB_WARPS * WK.total;
  const unsigned ring_sa = (unsigned)__cvta_generic_to_shared(ring);
  const unsigned mbar_sa = ring_sa + NS * SSB;
  RegSmem L;   // the view solve_fast_cta expects (bucket table: warp 0's)
  L.acc = SK.acc; L.coef = SK.coef; L.fvec = SK.fvec;
  L.seg = SK.flag - (3 * NB + 2) * 4;
  L.bucket = SK.total + WK.bucket;
  L.solve = SK.solve;
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
  int* s_runb = reinterpret_cast<int*>(s_cst + 2 + ZS);
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NS; ++s) rb_mbar_init(mbar_sa + 16 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();

  // fragment sources of this lane (row kq = lane & 3 of a k-step, element c = lane >> 2 of a tile)
  const int c = lane >> 2, kq = lane & 3;
  const unsigned z_sa = (unsigned)__cvta_generic_to_shared(s_z);
  const unsigned one_sa = (unsigned)__cvta_generic_to_shared(s_cst), zero_sa = one_sa + 8;
  // right operands: tile 0 = [d_0..d_5, u, 1], tile 1 = [d_6..d_{Q-1}, 0...]; sources relative to the warp's rows
  // of the stage (styles / ret) or constants
  int b0_kind, b1_kind;          // 0 style, 1 ret, 2 const one (sub 0), 3 const one (sub 1 -> 0)
  unsigned b0_off = 0, b1_off = 0;
  if (c < Q0)      { b0_kind = 0; b0_off = (c + kq * Q) * 8; }
  else if (c == 6) { b0_kind = 1; b0_off = kq * 8; }
  else             { b0_kind = 2; }
  if (c < Q1)      { b1_kind = 0; b1_off = (6 + c + kq * Q) * 8; }
  else             { b1_kind = 3; }
  unsigned a0_a, a0_s, a1_a, a1_s;
  {
    const int l0 = rm_left<Q>(0, c), l1 = rm_left<Q>(1, c);
    if (l0 >= 0)            { a0_a = z_sa + (3 + l0 + kq * ZS) * 8; a0_s = 4 * ZS * 8; }
    else                    { a0_a = zero_sa; a0_s = 0; }
    if (l1 >= 0)            { a1_a = z_sa + (3 + l1 + kq * ZS) * 8; a1_s = 4 * ZS * 8; }
    else if (l1 == RM_CAP)  { a1_a = z_sa + (2 + kq * ZS) * 8;      a1_s = 4 * ZS * 8; }
    else if (l1 == RM_W)    { a1_a = z_sa + (0 + kq * ZS) * 8;      a1_s = 4 * ZS * 8; }
    else if (l1 == RM_WY)   { a1_a = z_sa + (1 + kq * ZS) * 8;      a1_s = 4 * ZS * 8; }
    else if (l1 == RM_ONE)  { a1_a = one_sa;  a1_s = 0; }
    else                    { a1_a = zero_sa; a1_s = 0; }
  }
  int val0, val1;
  {
    const int l0 = rm_left<Q>(0, c), l1 = rm_left<Q>(1, c);
    val0 = l0 >= 0 ? 3 + l0 : -1;
    val1 = l1 >= 0 ? 3 + l1 : (l1 == RM_CAP ? 2 : (l1 == RM_W ? 0 : (l1 == RM_WY ? 1 : -1)));
  }
  const bool aligned16 = ((reinterpret_cast<uintptr_t>(p.cap) | reinterpret_cast<uintptr_t>(p.ret) |
                           reinterpret_cast<uintptr_t>(p.ind) | reinterpret_cast<uintptr_t>(p.style)) & 15) == 0;
  const long long rows_total = p.date_off[p.T];
  unsigned n_issued = 0, n_used = 0;      // super-chunks issued (thread 0) / consumed (all threads), in the same order

  // rows [128 k, 128 k + m) of the date that starts at row a -> stage (n_issued % NS)
  auto issue = [&](long long a, int n, int k, bool with_cap) {
    const int s = n_issued % NS;
    ++n_issued;
    const int r0 = RB_SROWS * k, m = min(RB_SROWS, n - r0);
    const long long g0 = a + r0;
    unsigned char* st = ring + (size_t)s * SSB;
    const unsigned st_sa = ring_sa + s * SSB, mb = mbar_sa + 16 * s;
    const bool bulk = aligned16 && ((g0 + m + 3) & ~3LL) <= rows_total;     // the aligned ranges stay inside the panel
    if (bulk) {
      if (tid == 0) {
        const long long c0 = g0 & ~1LL, c1 = (g0 + m + 1) & ~1LL;       // doubles, 16-byte aligned range
        const long long i0 = g0 & ~3LL, i1 = (g0 + m + 3) & ~3LL;       // ints
        const unsigned by_s = (unsigned)m * Q * 8, by_c = (unsigned)(c1 - c0) * 8, by_i = (unsigned)(i1 - i0) * 4;
        rb_mbar_expect_tx(mb, by_s + (with_cap ? 2 : 1) * by_c + by_i);
        rb_bulk_g2s(st_sa, p.style + g0 * Q, by_s, mb);
        if (with_cap) rb_bulk_g2s(st_sa + rb_off_cap<Q>(), p.cap + c0, by_c, mb);
        rb_bulk_g2s(st_sa + rb_off_ret<Q>(), p.ret + c0, by_c, mb);
        rb_bulk_g2s(st_sa + rb_off_ind<Q>(), p.ind + i0, by_i, mb);
      }
    } else {
      // plain loads into the same places (same shifts), then the barrier is completed by hand
      double* d_style = reinterpret_cast<double*>(st);
      double* d_cap = reinterpret_cast<double*>(st + rb_off_cap<Q>()) + (g0 & 1);
      double* d_ret = reinterpret_cast<double*>(st + rb_off_ret<Q>()) + (g0 & 1);
      int* d_ind = reinterpret_cast<int*>(st + rb_off_ind<Q>()) + (g0 & 3);
      for (int e = tid; e < m * Q; e += nt) d_style[e] = p.style[g0 * Q + e];
      for (int e = tid; e < m; e += nt) {
        if (with_cap) d_cap[e] = p.cap[g0 + e];
        d_ret[e] = p.ret[g0 + e];
        d_ind[e] = p.ind[g0 + e];
      }
      __syncthreads();
      if (tid == 0) rb_mbar_expect_tx(mb, 0);
    }
  };
  auto wait_stage = [&]() -> unsigned char* {
    const int s = n_used % NS;
    rb_mbar_wait(mbar_sa + 16 * s, (n_used / NS) & 1);
    ++n_used;
    return ring + (size_t)s * SSB;
  };

  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (int t = blockIdx.x; t < p.T; t += gridDim.x) {
    const long long a = p.date_off[t];
    const int n = (int)(p.date_off[t + 1] - a);
    if (n <= 0) {
      for (int k = tid; k < K; k += nt) store_factor_ret(p, t, k, nan);
      if (tid == 0) { p.r2[t] = nan; p.status[t] = BMFM_ST_NONFINITE; }
      continue;
    }
    const int nsup = (n + RB_SROWS - 1) / RB_SROWS;
    const int sh_c = (int)(a & 1), sh_i = (int)(a & 3);      // 128 k is a multiple of 4: the shifts hold for the whole date
    for (int e = lane; e < NB * BV; e += 32) s_bucket[e] = 0.0;
    if (warp == 0) {
      if (lane < Q) s_piv[lane] = p.style[a * Q + lane];
      if (lane == 0) s_misc[0] = p.ret[a];
    }
    // (the issue() fallback path contains a CTA barrier: every thread calls issue)
#pragma unroll
    for (int k = 0; k < NS - 1; ++k)
      if (k < nsup) issue(a, n, k, true);
    __syncthreads();
    const double ypiv = s_misc[0];
    const double sub0 = b0_kind == 0 ? s_piv[c] : (b0_kind == 1 ? ypiv : 0.0);
    const double sub1 = b1_kind == 0 ? s_piv[6 + c] : 1.0;
    double c00a = 0.0, c00b = 0.0, c10a = 0.0, c10b = 0.0, c11a = 0.0, c11b = 0.0, sq0 = 0.0, sq1 = 0.0;

    for (int k = 0; k < nsup; ++k) {
      if (k + NS - 1 < nsup) issue(a, n, k + NS - 1, true);
      unsigned char* stage = wait_stage();
      const int m = min(32, n - RB_SROWS * k - 32 * warp);      // rows of this warp's chunk (<= 0: none)
      if (m > 0) {
        double* st_style = reinterpret_cast<double*>(stage) + 32 * warp * Q;
        double* st_cap = reinterpret_cast<double*>(stage + rb_off_cap<Q>()) + sh_c + 32 * warp;
        double* st_ret = reinterpret_cast<double*>(stage + rb_off_ret<Q>()) + sh_c + 32 * warp;
        const int* st_ind = reinterpret_cast<const int*>(stage + rb_off_ind<Q>()) + sh_i + 32 * warp;
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
            double d[Q];
            load_row_styles<Q>(st_style, lane, s_piv, d);
#pragma unroll
            for (int q = 0; q < Q; ++q) { zr[3 + q] = w * d[q]; sq0 = fma(d[q], d[q], sq0); }
          } else {
#pragma unroll
            for (int v = 0; v < BV; ++v) zr[v] = 0.0;
#pragma unroll
            for (int q = 0; q < Q; ++q) st_style[lane * Q + q] = s_piv[q];
            st_ret[lane] = ypiv;
            asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");   // before the stage's next bulk copy
          }
        }
        __syncwarp();
        const int myb = lane < m ? bucket_of(st_ind[lane], P) : 64 + lane;
        const unsigned peers = __match_any_sync(0xffffffffu, myb);
        const int myleader = __ffs(peers) - 1;
        const unsigned leaders = __ballot_sync(0xffffffffu, lane < m && myleader == lane);
        const int G = __popc(leaders);
        const bool use_s = (p.smma & 1) && G <= 8;                // warp-uniform
        const int lid = lane < m ? __popc(leaders & ((1u << myleader) - 1u)) : 99;
        if (use_s && lane < m && myleader == lane) s_runb[lid] = myb;
        double cs0a = 0.0, cs0b = 0.0, cs1a = 0.0, cs1b = 0.0;
        {
          const unsigned sty_sa = (unsigned)__cvta_generic_to_shared(st_style);
          const unsigned ret_sa = (unsigned)__cvta_generic_to_shared(st_ret);
          const unsigned pb0 = b0_kind == 0 ? sty_sa + b0_off : (b0_kind == 1 ? ret_sa + b0_off : one_sa);
          const unsigned sb0 = b0_kind == 0 ? 4 * Q * 8 : (b0_kind == 1 ? 4 * 8 : 0);
          const unsigned pb1 = b1_kind == 0 ? sty_sa + b1_off : one_sa;
          const unsigned sb1 = b1_kind == 0 ? 4 * Q * 8 : 0;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const double b0 = lds_f64(pb0 + ks * sb0) - sub0;
            const double a0 = lds_f64(a0_a + ks * a0_s);
            dmma_884(c00a, c00b, a0, b0);
            const double b1 = lds_f64(pb1 + ks * sb1) - sub1;
            const double a1 = lds_f64(a1_a + ks * a1_s);
            dmma_884(c10a, c10b, a1, b0);
            dmma_884(c11a, c11b, a1, b1);
            const int rl = __shfl_sync(0xffffffffu, lid, 4 * ks + kq);
            if (use_s) {
              const double sb = rl == c ? 1.0 : 0.0;
              dmma_884(cs0a, cs0b, a0, sb);
              dmma_884(cs1a, cs1b, a1, sb);
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
        } else {
          // arbitrary row order: the first lane of every group of equal industries adds the group's rows
          const bool lead = lane < m && (__ffs(peers) - 1) == lane;
          double acc[BV];
#pragma unroll
          for (int v = 0; v < BV; ++v) acc[v] = 0.0;
          unsigned rest = lead ? peers : 0u;
          const double* s_zero = s_cst + 2;
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
      __syncthreads();   // every warp is done with the stage: it may be refilled
    }
    // accumulators -> the layout the solve reads: fragments handed over through the z staging, warp 0 adds them
    {
      double* cf = s_z;
      cf[0 * 32 + lane] = c00a; cf[1 * 32 + lane] = c00b; cf[2 * 32 + lane] = c10a; cf[3 * 32 + lane] = c10b;
      cf[4 * 32 + lane] = c11a; cf[5 * 32 + lane] = c11b; cf[6 * 32 + lane] = sq0;  cf[7 * 32 + lane] = sq1;
      __syncthreads();
      if (warp == 0) {
#pragma unroll
        for (int w = 1; w < RB_WARPS; ++w) {
          const double* o = reinterpret_cast<const double*>(smem + SK.total + (size_t)w * WK.total + WK.z);
          c00a += o[0 * 32 + lane]; c00b += o[1 * 32 + lane]; c10a += o[2 * 32 + lane]; c10b += o[3 * 32 + lane];
          c11a += o[4 * 32 + lane]; c11b += o[5 * 32 + lane]; sq0 += o[6 * 32 + lane];  sq1 += o[7 * 32 + lane];
        }
      }
      double* tb0 = reinterpret_cast<double*>(smem + SK.total + WK.bucket);
      for (int e = tid; e < NB * BV; e += nt) {
        double s = tb0[e];
#pragma unroll
        for (int w = 1; w < RB_WARPS; ++w) s += reinterpret_cast<const double*>(smem + SK.total + (size_t)w * WK.total + WK.bucket)[e];
        tb0[e] = s;
      }
    }
    if (warp == 0) {
      rm_scatter<Q>(0, 0, lane, c00a, c00b, ypiv, s_acc);
      rm_scatter<Q>(1, 0, lane, c10a, c10b, ypiv, s_acc);
      rm_scatter<Q>(1, 1, lane, c11a, c11b, ypiv, s_acc);
      const double d2 = warp_sum(sq0);
      const double y2 = warp_sum(sq1);
      if (lane < Q) s_acc[NA1PAD + 2 * Q + lane] = lane == 0 ? d2 : 0.0;
      if (lane == 0) s_acc[NA1PAD + 3 * Q + 1] = y2;
    }
    __syncthreads();
    // pass C's first super-chunks come in (from L2) while the solve runs
#pragma unroll
    for (int k = 0; k < NS - 1; ++k)
      if (k < nsup) issue(a, n, k, false);
    const int st = solve_fast_cta<Q, false>(p, L, sbase, n);
    if (st != BMFM_ST_OK) {
      // drain what was issued so that the stage / barrier bookkeeping stays in step
      for (int k = 0; k < NS - 1; ++k)
        if (k < nsup) wait_stage();
      __syncthreads();
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
    for (int k = 0; k < nsup; ++k) {
      if (k + NS - 1 < nsup) issue(a, n, k + NS - 1, false);
      unsigned char* stage = wait_stage();
      const int m = min(32, n - RB_SROWS * k - 32 * warp);
      if (lane < m) {
        const double* st_style = reinterpret_cast<const double*>(stage) + 32 * warp * Q;
        const double* st_ret = reinterpret_cast<const double*>(stage + rb_off_ret<Q>()) + sh_c + 32 * warp;
        const int* st_ind = reinterpret_cast<const int*>(stage + rb_off_ind<Q>()) + sh_i + 32 * warp;
        double e = st_ret[lane] - c0 - s_g[bucket_of(st_ind[lane], P)];
        double d[Q];
        load_row_styles<Q>(st_style, lane, s_piv, d);
#pragma unroll
        for (int q = 0; q < Q; ++q) e = fma(-s_beta[q], d[q], e);
        st_stream(p.spec + a + RB_SROWS * k + 32 * warp + lane, e);
        se += e;
        see = fma(e, e, see);
      }
      __syncthreads();
    }
    se = warp_sum(se);
    see = warp_sum(see);
    if (lane == 0) { s_red[2 * warp] = se; s_red[2 * warp + 1] = see; }
    __syncthreads();
    if (tid == 0) {
      se = 0.0; see = 0.0;
#pragma unroll
      for (int w = 0; w < RB_WARPS; ++w) { se += s_red[2 * w]; see += s_red[2 * w + 1]; }
      const double dn = (double)n;
      const double me = se / dn;
      const double var_e = see / dn - me * me;
      const double my = s_acc[NA1PAD + 3 * Q] / dn;
      const double var_y = s_acc[NA1PAD + 3 * Q + 1] / dn - my * my;
      p.r2[t] = 1.0 - var_e / var_y;
      p.status[t] = BMFM_ST_OK;
    }
    __syncthreads();
  }
}
