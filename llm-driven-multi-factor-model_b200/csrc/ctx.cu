// Context, workspace and the host-buffer (H2D / D2H) variants of the entry points.
#include <stdlib.h>
#include "bmfm_common.cuh"

extern "C" const char* bmfm_version(void) { return "bmfm 0.1 (sm_100a)"; }

extern "C" int bmfm_ctx_create(bmfm_ctx** out, int device, void* stream) {
  if (!out) return BMFM_E_ARG;
  *out = nullptr;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  bmfm_ctx* c = (bmfm_ctx*)calloc(1, sizeof(bmfm_ctx));
  if (!c) return BMFM_E_NOMEM;
  c->device = device;
  // NULL = the default (legacy) stream, which is also torch's default stream
  c->stream = (cudaStream_t)stream;
  c->own_stream = false;
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  cudaDeviceGetAttribute(&c->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
  c->opt_reg_smma = -1;
  *out = c;
  return BMFM_OK;
}

// Experiment switches, per context (so one caller's choice never changes another's kernels).
static int* bmfm_opt_slot(bmfm_ctx* c, const char* name) {
  if (!strcmp(name, "reg_mode")) return &c->opt_reg_mode;
  if (!strcmp(name, "reg_smma")) return &c->opt_reg_smma;
  if (!strcmp(name, "reg_warps")) return &c->opt_reg_warps;
  if (!strcmp(name, "reg_ctas")) return &c->opt_reg_ctas;
  if (!strcmp(name, "reg_static")) return &c->opt_reg_static;
  if (!strcmp(name, "vr_split")) return &c->opt_vr_split;
  if (!strcmp(name, "nw_chunks")) return &c->opt_nw_chunks;
  if (!strcmp(name, "eig_mode")) return &c->opt_eig_mode;
  if (!strcmp(name, "eig_fail_every")) return &c->opt_eig_fail_every;
  if (!strcmp(name, "eig_c_blocks")) return &c->opt_eig_c_blocks;
  if (!strcmp(name, "last_reg_mode")) return &c->last_reg_mode;
  return nullptr;
}

extern "C" int bmfm_ctx_set_option(bmfm_ctx* c, const char* name, int value) {
  if (!c || !name) return BMFM_E_ARG;
  int* slot = bmfm_opt_slot(c, name);
  BMFM_ARG(c, slot && slot != &c->last_reg_mode, "unknown option name");
  *slot = value;
  return BMFM_OK;
}

// Per-producer flags for the date-sharded Newey-West / vol-regime exchanges: flag_bufs = HOST array of n_rank device
// pointers, one zero-initialised block of 64 int64 per rank in peer-mapped memory (flag_bufs[rank] = this rank's own).
// Once registered, bmfm_newey_west_shard_totals / _emit and bmfm_vol_regime_shard_totals / _scale publish and wait
// through these flags and the caller drops its cross-rank barriers between them (it keeps ONE all-rank barrier per
// step, e.g. the one after the regression, so that no rank runs a whole step ahead).  n_rank = 0 unregisters.
extern "C" int bmfm_ctx_set_peer_flags(bmfm_ctx* c, int rank, int n_rank, int64_t* const* flag_bufs) {
  if (!c) return BMFM_E_ARG;
  BMFM_ARG(c, n_rank >= 0 && n_rank <= 8 && (n_rank == 0 || (flag_bufs && rank >= 0 && rank < n_rank)), "bad peer flags");
  c->flag_rank = rank;
  c->n_flag_peers = n_rank;
  for (int r = 0; r < 8; ++r) c->peer_flags[r] = r < n_rank ? (long long*)flag_bufs[r] : nullptr;
  return BMFM_OK;
}

extern "C" int bmfm_ctx_get_option(bmfm_ctx* c, const char* name, int* value) {
  if (!c || !name || !value) return BMFM_E_ARG;
  int* slot = bmfm_opt_slot(c, name);
  BMFM_ARG(c, slot, "unknown option name");
  *value = *slot;
  return BMFM_OK;
}

extern "C" int bmfm_ctx_destroy(bmfm_ctx* c) {
  if (!c) return BMFM_E_ARG;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->ws) cudaFree(c->ws);
  for (int i = 0; i < 16; ++i)
    if (c->stage[i]) cudaFree(c->stage[i]);
  if (c->pipe_ready) {
    cudaStreamDestroy(c->copy_in);
    cudaStreamDestroy(c->copy_out);
    for (int i = 0; i < 40; ++i) cudaEventDestroy(c->pipe_ev[i]);
  }
  if (c->own_stream) cudaStreamDestroy(c->stream);
  free(c);
  return BMFM_OK;
}

extern "C" int bmfm_ctx_sync(bmfm_ctx* c) {
  if (!c) return BMFM_E_ARG;
  BMFM_CUDA(c, cudaStreamSynchronize(c->stream));
  return BMFM_OK;
}

extern "C" const char* bmfm_last_error(const bmfm_ctx* c) { return c ? c->err : "null context"; }
extern "C" long long bmfm_launch_count(const bmfm_ctx* c) { return c ? c->launches : 0; }

int bmfm_ws_reserve(bmfm_ctx* c, size_t bytes) {
  if (c->ws_bytes >= bytes) return BMFM_OK;
  BMFM_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->ws) BMFM_CUDA(c, cudaFree(c->ws));
  c->ws = nullptr;
  c->ws_bytes = 0;
  BMFM_CUDA(c, cudaMalloc(&c->ws, bytes));
  c->ws_bytes = bytes;
  return BMFM_OK;
}

int bmfm_stage_reserve(bmfm_ctx* c, int slot, size_t bytes) {
  if (bytes == 0) bytes = 16;
  if (c->stage_bytes[slot] >= bytes) return BMFM_OK;
  BMFM_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->stage[slot]) BMFM_CUDA(c, cudaFree(c->stage[slot]));
  c->stage[slot] = nullptr;
  c->stage_bytes[slot] = 0;
  BMFM_CUDA(c, cudaMalloc(&c->stage[slot], bytes));
  c->stage_bytes[slot] = bytes;
  return BMFM_OK;
}

#define H2D(slot, src, bytes)                                                              \
  do {                                                                                     \
    int rc__ = bmfm_stage_reserve(ctx, slot, bytes);                                       \
    if (rc__) return rc__;                                                                 \
    if ((bytes) > 0)                                                                       \
      BMFM_CUDA(ctx, cudaMemcpyAsync(ctx->stage[slot], src, bytes, cudaMemcpyHostToDevice, ctx->stream)); \
  } while (0)
#define DEVBUF(slot, bytes)                                  \
  do {                                                       \
    int rc__ = bmfm_stage_reserve(ctx, slot, bytes);         \
    if (rc__) return rc__;                                   \
  } while (0)
#define D2H(dst, slot, bytes)                                                              \
  do {                                                                                     \
    if ((bytes) > 0)                                                                       \
      BMFM_CUDA(ctx, cudaMemcpyAsync(dst, ctx->stage[slot], bytes, cudaMemcpyDeviceToHost, ctx->stream)); \
  } while (0)

extern "C" int bmfm_reg_by_time_host(bmfm_ctx* ctx, int T, int P, int Q,
                                     const int64_t* date_off, const double* cap, const double* ret,
                                     const int32_t* ind, const double* style,
                                     double* factor_ret, double* specific_ret, double* r2,
                                     double* exposure, int32_t* status) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && P >= 0 && Q >= 0 && date_off, "bad sizes");
  if (T == 0) return BMFM_OK;
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t rows = (size_t)date_off[T];
  const size_t K = 1 + (size_t)P + Q;
  int max_rows = 0;
  for (int t = 0; t < T; ++t) {
    const int64_t n = date_off[t + 1] - date_off[t];
    if (n > max_rows) max_rows = (int)n;
  }
  H2D(0, date_off, (T + 1) * sizeof(int64_t));
  H2D(1, cap, rows * 8);
  H2D(2, ret, rows * 8);
  H2D(3, ind, rows * 4);
  H2D(4, style, rows * Q * 8);
  DEVBUF(5, T * K * 8);
  DEVBUF(6, rows * 8);
  DEVBUF(7, T * 8);
  DEVBUF(8, T * 4);
  if (exposure) DEVBUF(9, T * K * K * 8);
  int rc = bmfm_reg_by_time(ctx, T, P, Q, max_rows, (const int64_t*)ctx->stage[0], (const double*)ctx->stage[1],
                            (const double*)ctx->stage[2], (const int32_t*)ctx->stage[3],
                            (const double*)ctx->stage[4], (double*)ctx->stage[5], (double*)ctx->stage[6],
                            (double*)ctx->stage[7], exposure ? (double*)ctx->stage[9] : nullptr,
                            (int32_t*)ctx->stage[8]);
  if (rc) return rc;
  D2H(factor_ret, 5, T * K * 8);
  D2H(specific_ret, 6, rows * 8);
  D2H(r2, 7, T * 8);
  D2H(status, 8, T * 4);
  if (exposure) D2H(exposure, 9, T * K * K * 8);
  BMFM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return BMFM_OK;
}

// ---------------------------------------------------------------------------------------------------------
// One whole step with HOST buffers, copies overlapped with the kernels
// ---------------------------------------------------------------------------------------------------------
// MFM.reg_by_time (MFM.py:48-76) + MFM.Newey_West_by_time (MFM.py:80-101) + the volatility-regime adjustment of
// MFM.vol_regime_adj_by_time (MFM.py:130-167) applied to the Newey-West covariances — bench config C2 — for a
// caller whose panel and results live in host memory.  The panel goes to the device in blocks of dates on a
// copy stream; each block's regression starts as soon as its rows have landed and its specific returns leave
// on a second copy stream while the next block is still arriving (PCIe is full duplex); the 2 x T x K x K
// covariance outputs leave while the following kernel runs.  Synchronises before returning.  Pinned host
// memory is needed for the copies to overlap; pageable memory works but serialises.
extern "C" int bmfm_risk_step_host(bmfm_ctx* ctx, int T, int P, int Q, const int64_t* date_off, const double* cap,
                                   const double* ret, const int32_t* ind, const double* style, int q, double tao_nw,
                                   double tao_vr, double* factor_ret, double* specific_ret, double* r2, int32_t* status,
                                   double* nw_cov, int32_t* nw_status, double* vr_cov, double* lamb) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, T >= 0 && P >= 0 && Q >= 0 && date_off && q >= 0 && tao_nw > 0 && tao_vr > 0, "bad sizes");
  if (T == 0) return BMFM_OK;
  BMFM_ARG(ctx, cap && ret && ind && (style || Q == 0) && factor_ret && specific_ret && r2 && status && nw_cov &&
                nw_status && vr_cov && lamb, "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!ctx->pipe_ready) {
    BMFM_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
    BMFM_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
    for (int i = 0; i < 40; ++i) BMFM_CUDA(ctx, cudaEventCreateWithFlags(&ctx->pipe_ev[i], cudaEventDisableTiming));
    ctx->pipe_ready = true;
  }
  const size_t rows = (size_t)date_off[T];
  const size_t K = 1 + (size_t)P + Q;
  int max_rows = 0;
  for (int t = 0; t < T; ++t) {
    const int64_t n = date_off[t + 1] - date_off[t];
    if (n > max_rows) max_rows = (int)n;
  }
  DEVBUF(0, (T + 1) * sizeof(int64_t));
  DEVBUF(1, rows * 8);
  DEVBUF(2, rows * 8);
  DEVBUF(3, rows * 4);
  DEVBUF(4, rows * Q * 8);
  DEVBUF(5, T * K * 8);
  DEVBUF(6, rows * 8);
  DEVBUF(7, T * 8);
  DEVBUF(8, T * 4);
  DEVBUF(10, T * K * K * 8);
  DEVBUF(11, T * 4);
  DEVBUF(12, T * K * K * 8);
  DEVBUF(13, T * 8);
  DEVBUF(9, 2 * T * 8);
  const int64_t* d_off = (const int64_t*)ctx->stage[0];
  double* d_cap = (double*)ctx->stage[1];
  double* d_ret = (double*)ctx->stage[2];
  int32_t* d_ind = (int32_t*)ctx->stage[3];
  double* d_style = (double*)ctx->stage[4];
  double* d_fr = (double*)ctx->stage[5];
  double* d_spec = (double*)ctx->stage[6];
  double* d_r2 = (double*)ctx->stage[7];
  int32_t* d_st = (int32_t*)ctx->stage[8];
  double* d_nw = (double*)ctx->stage[10];
  int32_t* d_nwst = (int32_t*)ctx->stage[11];
  double* d_vr = (double*)ctx->stage[12];
  double* d_lamb = (double*)ctx->stage[13];
  double* d_wb = (double*)ctx->stage[9];        // vol-regime bias terms of the dates done so far (wb | wv)
  cudaEvent_t* ev = ctx->pipe_ev;
  // whatever the compute stream still does with the staging buffers comes first
  BMFM_CUDA(ctx, cudaEventRecord(ev[39], ctx->stream));
  BMFM_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_in, ev[39], 0));
  BMFM_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, ev[39], 0));
  BMFM_CUDA(ctx, cudaMemcpyAsync((void*)d_off, date_off, (T + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->copy_in));
  const int nchunk = T < 16 ? 1 : 16;
  for (int c = 0; c < nchunk; ++c) {
    const int t0 = (int)((long long)T * c / nchunk), t1 = (int)((long long)T * (c + 1) / nchunk);
    const size_t r0 = (size_t)date_off[t0], r1 = (size_t)date_off[t1];
    if (r1 > r0) {
      BMFM_CUDA(ctx, cudaMemcpyAsync(d_cap + r0, cap + r0, (r1 - r0) * 8, cudaMemcpyHostToDevice, ctx->copy_in));
      BMFM_CUDA(ctx, cudaMemcpyAsync(d_ret + r0, ret + r0, (r1 - r0) * 8, cudaMemcpyHostToDevice, ctx->copy_in));
      BMFM_CUDA(ctx, cudaMemcpyAsync(d_ind + r0, ind + r0, (r1 - r0) * 4, cudaMemcpyHostToDevice, ctx->copy_in));
      if (Q > 0)
        BMFM_CUDA(ctx, cudaMemcpyAsync(d_style + r0 * Q, style + r0 * Q, (r1 - r0) * Q * 8, cudaMemcpyHostToDevice, ctx->copy_in));
    }
    BMFM_CUDA(ctx, cudaEventRecord(ev[c], ctx->copy_in));
    BMFM_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ev[c], 0));
    int rc = bmfm_reg_by_time(ctx, t1 - t0, P, Q, max_rows, d_off + t0, d_cap, d_ret, d_ind, d_style, d_fr + (size_t)t0 * K,
                              d_spec, d_r2 + t0, nullptr, d_st + t0);
    if (rc) return rc;
    // the covariances of a date need only the factor returns up to it: Newey-West and the vol-regime adjustment of
    // this block run now (the range forms re-reduce the earlier dates' moments, microseconds against the 200+ us a
    // block takes to arrive), and the block's results leave while later blocks are still coming in
    rc = bmfm_newey_west_by_time_range(ctx, T, (int)K, q, tao_nw, d_fr, t0, t1, d_nw + (size_t)t0 * K * K, d_nwst + t0);
    if (rc) return rc;
    rc = bmfm_vol_regime_adj_range(ctx, T, t0, t1, (int)K, tao_vr, d_fr, d_nw, d_nwst, d_wb, d_wb + T, d_vr, d_lamb);
    if (rc) return rc;
    BMFM_CUDA(ctx, cudaEventRecord(ev[16 + c], ctx->stream));
    BMFM_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, ev[16 + c], 0));
    if (r1 > r0)
      BMFM_CUDA(ctx, cudaMemcpyAsync(specific_ret + r0, d_spec + r0, (r1 - r0) * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
    if (t1 > t0) {
      BMFM_CUDA(ctx, cudaMemcpyAsync(nw_cov + (size_t)t0 * K * K, d_nw + (size_t)t0 * K * K, (size_t)(t1 - t0) * K * K * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
      BMFM_CUDA(ctx, cudaMemcpyAsync(vr_cov + (size_t)t0 * K * K, d_vr + (size_t)t0 * K * K, (size_t)(t1 - t0) * K * K * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
    }
  }
  BMFM_CUDA(ctx, cudaMemcpyAsync(factor_ret, d_fr, T * K * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
  BMFM_CUDA(ctx, cudaMemcpyAsync(r2, d_r2, T * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
  BMFM_CUDA(ctx, cudaMemcpyAsync(status, d_st, T * 4, cudaMemcpyDeviceToHost, ctx->copy_out));
  BMFM_CUDA(ctx, cudaMemcpyAsync(nw_status, d_nwst, T * 4, cudaMemcpyDeviceToHost, ctx->copy_out));
  BMFM_CUDA(ctx, cudaMemcpyAsync(lamb, d_lamb, T * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
  BMFM_CUDA(ctx, cudaStreamSynchronize(ctx->copy_out));
  BMFM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return BMFM_OK;
}
