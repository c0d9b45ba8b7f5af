// Shared host/device helpers for libbmfm (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/bmfm.h"

struct bmfm_ctx {
  int device;
  cudaStream_t stream;
  bool own_stream;
  int sm_count;
  int max_smem_optin;
  // grow-only device workspace (scratch beyond caller inputs/outputs)
  void* ws;
  size_t ws_bytes;
  // grow-only staging buffers for the *_host entry points
  void* stage[16];
  size_t stage_bytes[16];
  long long launches;
  // copy streams and events of the pipelined host entry point (bmfm_risk_step_host), created on first use
  cudaStream_t copy_in, copy_out;
  cudaEvent_t pipe_ev[40];
  bool pipe_ready;
  // per-producer flags of the date-sharded exchanges (bmfm_ctx_set_peer_flags); n_flag_peers = 0: not registered
  int flag_rank, n_flag_peers;
  long long* peer_flags[8];
  // per-context experiment switches (bmfm_ctx_set_option); 0 / negative = library default
  int opt_reg_mode, opt_reg_smma, opt_reg_warps, opt_reg_ctas, opt_nw_chunks, opt_eig_mode, opt_eig_fail_every, opt_eig_c_blocks, opt_reg_static, opt_vr_split;
  int last_reg_mode;   // kernel form the last bmfm_reg_by_time call launched (bmfm_ctx_get_option "last_reg_mode")
  char err[512];
};

#define BMFM_CUDA(ctx, call)                                                        \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d %s: %s", __FILE__, __LINE__,  \
               #call, cudaGetErrorString(e__));                                     \
      return (int)e__;                                                              \
    }                                                                               \
  } while (0)

#define BMFM_ARG(ctx, cond, msg)                                                    \
  do {                                                                              \
    if (!(cond)) {                                                                  \
      snprintf((ctx)->err, sizeof((ctx)->err), "bad argument: %s", msg);            \
      return BMFM_E_ARG;                                                            \
    }                                                                               \
  } while (0)

// ---- per-producer flags of the date-sharded exchanges -------------------------------------------------------
// Every rank owns a block of BMFM_FLAG_WORDS 64-bit words in peer-mapped memory:
//   [ch * 8 + s]   epoch of the newest record that rank s has published on channel ch (written by rank s)
//   [32 + ch]      this rank's own epoch on channel ch            (local)
//   [40 + ch]      CTAs of this rank's producer kernel that have finished, all steps   (local)
//   [48]           set when a wait gave up                        (local)
// A producer bumps its epoch and writes it into the flag word of every LATER rank after its record stores; a
// consumer waits until the words of all EARLIER ranks have reached its own epoch.  Nobody waits for a later rank.
constexpr int BMFM_FLAG_WORDS = 64;
constexpr int BMFM_CH_NW = 0, BMFM_CH_VR = 1;
struct BmfmFlagPeers {
  int rank, n;
  long long* f[8];     // f[rank] is this rank's own block
};

__device__ __forceinline__ void bmfm_flag_publish(const BmfmFlagPeers& fp, int ch, long long epoch) {
  __threadfence_system();
  for (int r = fp.rank + 1; r < fp.n; ++r) {
    volatile long long* w = fp.f[r] + ch * 8 + fp.rank;
    *w = epoch;
  }
  __threadfence_system();
}

// one thread: wait for the records of all earlier ranks (bounded: a lost producer must not hang the GPU)
__device__ __forceinline__ void bmfm_flag_wait(const BmfmFlagPeers& fp, int ch) {
  volatile long long* mine = fp.f[fp.rank];
  const long long want = mine[32 + ch];
  for (int s = 0; s < fp.rank; ++s) {
    long long spins = 0;
    while (mine[ch * 8 + s] < want) {
      if (++spins > (1LL << 26)) { mine[48] = 1; break; }
      __nanosleep(64);
    }
  }
  __threadfence_system();
}

// the registered flag blocks as a kernel argument (n = 0 when none are registered or the rank does not match)
inline BmfmFlagPeers bmfm_flag_peers(const bmfm_ctx* ctx, int rank, int n_peer_expected) {
  BmfmFlagPeers fp;
  fp.rank = rank;
  fp.n = 0;
  for (int r = 0; r < 8; ++r) fp.f[r] = nullptr;
  if (ctx->n_flag_peers > 1 && ctx->flag_rank == rank && (n_peer_expected == 0 || n_peer_expected == ctx->n_flag_peers)) {
    fp.n = ctx->n_flag_peers;
    for (int r = 0; r < fp.n; ++r) fp.f[r] = ctx->peer_flags[r];
  }
  return fp;
}

// host helpers (ctx.cu)
int bmfm_ws_reserve(bmfm_ctx* ctx, size_t bytes);             // ctx->ws >= bytes
int bmfm_stage_reserve(bmfm_ctx* ctx, int slot, size_t bytes);  // ctx->stage[slot] >= bytes
static inline size_t bmfm_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

#ifdef __CUDACC__
// ---- device helpers -------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}

// Reduce NPAD (multiple of 32) per-lane values across the warp by recursive
// halving: 5 exchange steps move NPAD/2 + NPAD/4 + ... values instead of 5*NPAD.
// Afterwards slot i (< NPAD/32) of this lane holds the warp total of original
// index reduce_scatter_index<NPAD>(i, lane).
template <int NPAD>
__device__ __forceinline__ void warp_reduce_scatter(double (&a)[NPAD], int lane) {
  static_assert(NPAD % 32 == 0, "pad to a multiple of 32");
#pragma unroll
  for (int s = 0; s < 5; ++s) {
    const int off = 16 >> s;
    const int H = NPAD >> (s + 1);
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const double lo = a[i], hi = a[i + H];
      const double send = upper ? lo : hi;
      const double keep = upper ? hi : lo;
      a[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
}
template <int NPAD>
__device__ __forceinline__ int reduce_scatter_index(int i, int lane) {
  int idx = i;
#pragma unroll
  for (int s = 0; s < 5; ++s)
    if (lane & (16 >> s)) idx += NPAD >> (s + 1);
  return idx;
}

// cp.async (LDGSTS): element-granular global -> shared copies, any 4/8/16-byte alignment
__device__ __forceinline__ void cp_async_4(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_8(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.wait_all;\n" ::: "memory");
}
// streaming store: results written once and not re-read by this kernel
__device__ __forceinline__ void st_stream(double* p, double v) {
  asm volatile("st.global.cs.f64 [%0], %1;\n" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ void st_stream2(double* p, double v0, double v1) {   // p 16-byte aligned
  asm volatile("st.global.cs.v2.f64 [%0], {%1, %2};\n" ::"l"(p), "d"(v0), "d"(v1) : "memory");
}
#endif
