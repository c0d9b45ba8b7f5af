// Device side of the panel packer: what demo.py:33-35 and MFM.reg_by_time (MFM.py:58-61) do
// to the long table before the first regression, without a host pass over the rows.
//
//   bmfm_onehot_to_index   the P one-hot industry columns (the `1*(industry == code)` block of demo.py:33, int64,
//                          any strides: pandas keeps it column-major) -> one int32 industry index per row, bit-exact,
//                          with the checks the host packer makes (entries other than 0/1, more than one 1 in a row);
//   bmfm_pack_gather       rows permuted into the storage order (date, industry, stock name): cap, ret, industry
//                          index and the Q style columns (any strides) gathered through the sort permutation into the
//                          date-major layout the regression kernel reads (style row-major [rows][Q]).
// The sort itself (one 64-bit key per row) is a library radix sort on the caller's side; these two kernels are the
// HBM-bound passes around it: (8 P + 4) B and (8 + 24 + 16 Q + 8) B per row.
#include "bmfm_common.cuh"

namespace bmfm {

__global__ void onehot_to_index_kernel(long long rows, int P, const long long* onehot, long long rs, long long cs,
                                       int* ind, int* flags) {
  int bad_val = 0, bad_cnt = 0;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
    int first = -1, cnt = 0;
    for (int p = 0; p < P; ++p) {
      const long long v = onehot[r * rs + p * cs];
      if (v != 0 && v != 1) bad_val = 1;
      if (v != 0) { if (first < 0) first = p; ++cnt; }
    }
    if (cnt > 1) bad_cnt = 1;
    ind[r] = cnt == 0 ? -1 : first;
  }
  if (bad_val) atomicAdd(&flags[0], 1);
  if (bad_cnt) atomicAdd(&flags[1], 1);
}

__global__ void pack_gather_kernel(long long rows, int Q, const long long* order, const double* cap, const double* ret,
                                   const int* ind, const double* style, long long srs, long long scs, double* cap_o,
                                   double* ret_o, int* ind_o, double* style_o) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += (long long)gridDim.x * blockDim.x) {
    const long long s = order[i];
    cap_o[i] = cap[s];
    ret_o[i] = ret[s];
    ind_o[i] = ind[s];
    for (int q = 0; q < Q; ++q) style_o[i * Q + q] = style[s * srs + q * scs];
  }
}

}  // namespace bmfm

using namespace bmfm;

extern "C" int bmfm_onehot_to_index(bmfm_ctx* ctx, int64_t rows, int P, const int64_t* onehot, int64_t row_stride,
                                    int64_t col_stride, int32_t* ind, int32_t* flags) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, rows >= 0 && P >= 0 && flags, "bad arguments");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  BMFM_CUDA(ctx, cudaMemsetAsync(flags, 0, 2 * sizeof(int32_t), ctx->stream));
  if (rows == 0) return BMFM_OK;
  BMFM_ARG(ctx, ind && (onehot || P == 0), "NULL pointer");
  long long blocks = (rows + 255) / 256;
  const long long cap = (long long)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  onehot_to_index_kernel<<<(int)blocks, 256, 0, ctx->stream>>>((long long)rows, P, (const long long*)onehot,
                                                               (long long)row_stride, (long long)col_stride, ind, flags);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}

extern "C" int bmfm_pack_gather(bmfm_ctx* ctx, int64_t rows, int Q, const int64_t* order, const double* cap,
                                const double* ret, const int32_t* ind, const double* style, int64_t style_row_stride,
                                int64_t style_col_stride, double* cap_out, double* ret_out, int32_t* ind_out,
                                double* style_out) {
  if (!ctx) return BMFM_E_ARG;
  BMFM_ARG(ctx, rows >= 0 && Q >= 0, "bad sizes");
  if (rows == 0) return BMFM_OK;
  BMFM_ARG(ctx, order && cap && ret && ind && cap_out && ret_out && ind_out && ((style && style_out) || Q == 0), "NULL pointer");
  BMFM_CUDA(ctx, cudaSetDevice(ctx->device));
  long long blocks = (rows + 255) / 256;
  const long long capb = (long long)ctx->sm_count * 16;
  if (blocks > capb) blocks = capb;
  pack_gather_kernel<<<(int)blocks, 256, 0, ctx->stream>>>((long long)rows, Q, (const long long*)order, cap, ret, ind, style,
                                                           (long long)style_row_stride, (long long)style_col_stride,
                                                           cap_out, ret_out, ind_out, style_out);
  ctx->launches++;
  BMFM_CUDA(ctx, cudaGetLastError());
  return BMFM_OK;
}
