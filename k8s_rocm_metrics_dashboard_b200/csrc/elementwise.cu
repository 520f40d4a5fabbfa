// elementwise.cu -- the small per-card steps of the reference's instant path.
//   pivot_gather : long -> wide (app.py:207 df.pivot): cell (card, metric) <- sample or NaN
//   ratio_pct    : used/total*100 per card (app.py:210-212); any zero denominator is reported
//                  so the adapter can raise ZeroDivisionError like Python's float division
//   where_gt     : keep x > threshold else NaN (app.py:343 power_values[power_values > 0])
#include "pmx_common.cuh"

namespace {

__global__ void pivot_gather_kernel(const double* __restrict__ val, const int32_t* __restrict__ src,
                                    int64_t n_cells, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_cells) return;
  const int32_t k = src[i];
  out[i] = k >= 0 ? val[k] : pmx_nan();
}

__global__ void ratio_pct_kernel(const double* __restrict__ num, const double* __restrict__ den,
                                 int64_t n, int64_t num_stride, int64_t den_stride, int64_t out_stride,
                                 double* __restrict__ out, int32_t* zero_flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double d = den[i * den_stride];
  if (d == 0.0) atomicExch(zero_flag, 1);
  // two roundings, in this order, like (used / total) * 100 on Python floats
  out[i * out_stride] = __dmul_rn(__ddiv_rn(num[i * num_stride], d), 100.0);
}

__global__ void where_gt_kernel(const double* __restrict__ x, int64_t n, double thr, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double v = x[i];
  out[i] = v > thr ? v : pmx_nan();
}

}  // namespace

extern "C" {

int pmx_pivot_gather(pmx_ctx* ctx, const double* d_val, const int32_t* d_cell_src, int64_t n_cells,
                     double* d_out, pmx_stream stream) {
  PMX_REQUIRE(ctx && n_cells >= 0, "bad argument");
  if (n_cells == 0) return PMX_OK;
  PMX_REQUIRE(d_cell_src && d_out, "NULL buffer");
  pivot_gather_kernel<<<(unsigned)((n_cells + 255) / 256), 256, 0, pmx_st(stream)>>>(d_val, d_cell_src, n_cells, d_out);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_ratio_pct(pmx_ctx* ctx, const double* d_num, int64_t num_stride, const double* d_den,
                  int64_t den_stride, int64_t n, double* d_out, int64_t out_stride,
                  int32_t* d_zero_flag, pmx_stream stream) {
  PMX_REQUIRE(ctx && n >= 0 && d_zero_flag, "bad argument");
  PMX_CUDA(cudaMemsetAsync(d_zero_flag, 0, sizeof(int32_t), pmx_st(stream)));
  if (n == 0) return PMX_OK;
  PMX_REQUIRE(d_num && d_den && d_out, "NULL buffer");
  ratio_pct_kernel<<<(unsigned)((n + 255) / 256), 256, 0, pmx_st(stream)>>>(d_num, d_den, n, num_stride, den_stride,
                                                                           out_stride, d_out, d_zero_flag);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_where_gt(pmx_ctx* ctx, const double* d_x, int64_t n, double threshold, double* d_out,
                 pmx_stream stream) {
  PMX_REQUIRE(ctx && n >= 0, "bad argument");
  if (n == 0) return PMX_OK;
  PMX_REQUIRE(d_x && d_out, "NULL buffer");
  where_gt_kernel<<<(unsigned)((n + 255) / 256), 256, 0, pmx_st(stream)>>>(d_x, n, threshold, d_out);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
