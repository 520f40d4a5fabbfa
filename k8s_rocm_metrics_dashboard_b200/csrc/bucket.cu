// bucket.cu -- X3, second form: bucket down-sampling of ragged series onto a coarser grid.
//
// Series.resample('<step>') bins [start + j*step, start + (j+1)*step) labelled
// left; per bin the LAST non-NaN sample, the MEAN of the non-NaN samples and their COUNT (a stale
// marker is a NaN to pandas and is skipped like one).  It is the align kernel with a bin test
// instead of a floor: one thread per (series, bin) finds the bin's first sample by binary search in
// the series' timestamps (its last is the next bin's first: one shuffle, the warp's last lane
// searches twice) and folds the bin's samples in time order, so the mean's sum is the sequential
// sum the oracle forms.  Bins of one series are consecutive threads: the searches of a warp land in
// neighbouring samples and the three output rows are written coalesced.
#include "pmx_common.cuh"

namespace {

// first index in [0, n) with ts[i] >= t (n if none)
__device__ __forceinline__ int64_t lower_bound(const int64_t* __restrict__ ts, int64_t n, int64_t t) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (ts[mid] < t) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(256)
bucket_kernel(const int64_t* __restrict__ ts, const double* __restrict__ val, const int64_t* __restrict__ off,
              int64_t start_ms, int64_t step_ms, int64_t n_steps, double* __restrict__ o_last,
              double* __restrict__ o_mean, int32_t* __restrict__ o_cnt) {
  const int64_t s = blockIdx.y;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t base = off[s], n = off[s + 1] - base;
  const int64_t* t = ts + base;
  const double* v = val + base;
  const bool live = j < n_steps;
  const int64_t t0 = start_ms + (live ? j : n_steps) * step_ms;
  const int64_t lo = lower_bound(t, n, t0);
  int64_t hi = __shfl_down_sync(0xffffffffu, lo, 1);
  if ((threadIdx.x & 31) == 31) hi = lower_bound(t, n, t0 + step_ms);
  if (!live) return;
  double sum = 0.0, last = pmx_nan();
  int cnt = 0;
  for (int64_t i = lo; i < hi; ++i) {
    const double x = v[i];
    if (x == x) { sum += x; last = x; ++cnt; }
  }
  const int64_t o = s * n_steps + j;
  if (o_last) o_last[o] = last;
  if (o_mean) o_mean[o] = cnt > 0 ? sum / (double)cnt : pmx_nan();
  if (o_cnt) o_cnt[o] = cnt;
}

}  // namespace

extern "C" {

int pmx_bucket(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off, int64_t n_series,
               int64_t start_ms, int64_t step_ms, int64_t n_steps, double* d_last, double* d_mean,
               int32_t* d_count, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0 && step_ms > 0, "bad sizes");
  if (n_series == 0 || n_steps == 0) return PMX_OK;
  PMX_REQUIRE(d_off && (d_last || d_mean || d_count), "NULL buffer");
  PMX_REQUIRE(n_series <= 65535, "at most 65535 series per call (grid.y); split the call");
  dim3 grid((unsigned)((n_steps + 255) / 256), (unsigned)n_series);
  bucket_kernel<<<grid, 256, 0, pmx_st(stream)>>>(d_ts, d_val, d_off, start_ms, step_ms, n_steps, d_last, d_mean, d_count);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
