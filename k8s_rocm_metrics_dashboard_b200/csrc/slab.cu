// slab.cu -- the series cache's tail gather (cache.py).  Every series keeps
// its recent samples in a row of a [rows][cap] slab, ascending, count[s] of them held.  A refresh
// needs, per series, the samples newer than a threshold as one CSR for the fused kernel:
//   pmx_slab_tail_lengths : first[s] = number of held samples with ts <= threshold (binary search,
//                           one thread per series), len[s] = count[s] - first[s];
//   (the caller turns len into offsets with one cumulative sum)
//   pmx_slab_tail_gather  : one warp per series copies ts/val[s][first[s] .. count[s]) to
//                           out[off[s] ..).
#include "pmx_common.cuh"

namespace {

__global__ void tail_lengths_kernel(const int64_t* __restrict__ ts, const int64_t* __restrict__ count, int64_t n_series,
                                    int64_t cap, int64_t threshold, int64_t* __restrict__ first,
                                    int64_t* __restrict__ len) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_series) return;
  const int64_t* row = ts + s * cap;
  int64_t lo = 0, hi = count[s];
  const int64_t n = hi;
  while (lo < hi) {                                          // first index with ts > threshold
    const int64_t mid = (lo + hi) >> 1;
    if (row[mid] <= threshold) lo = mid + 1; else hi = mid;
  }
  first[s] = lo;
  len[s] = n - lo;
}

__global__ void __launch_bounds__(256)
tail_gather_kernel(const int64_t* __restrict__ ts, const double* __restrict__ val, const int64_t* __restrict__ first,
                   const int64_t* __restrict__ off, int64_t n_series, int64_t cap,
                   int64_t* __restrict__ out_ts, double* __restrict__ out_val) {
  const int64_t s = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (s >= n_series) return;
  const int64_t a = off[s], n = off[s + 1] - a, f = first[s];
  const int64_t* rt = ts + s * cap + f;
  const double* rv = val + s * cap + f;
  for (int64_t i = lane; i < n; i += 32) { out_ts[a + i] = rt[i]; out_val[a + i] = rv[i]; }
}

}  // namespace

extern "C" {

int pmx_slab_tail_lengths(pmx_ctx* ctx, const int64_t* d_ts, const int64_t* d_count, int64_t n_series, int64_t cap,
                          int64_t threshold_ms, int64_t* d_first, int64_t* d_len, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && cap >= 1, "bad sizes");
  if (n_series == 0) return PMX_OK;
  PMX_REQUIRE(d_ts && d_count && d_first && d_len, "NULL buffer");
  tail_lengths_kernel<<<(unsigned)((n_series + 255) / 256), 256, 0, pmx_st(stream)>>>(d_ts, d_count, n_series, cap,
                                                                                    threshold_ms, d_first, d_len);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_slab_tail_gather(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_first,
                         const int64_t* d_off, int64_t n_series, int64_t cap, int64_t* d_out_ts, double* d_out_val,
                         pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && cap >= 1, "bad sizes");
  if (n_series == 0) return PMX_OK;
  PMX_REQUIRE(d_ts && d_val && d_first && d_off && d_out_ts && d_out_val, "NULL buffer");
  const int64_t threads = n_series * 32;
  tail_gather_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, pmx_st(stream)>>>(d_ts, d_val, d_first, d_off, n_series,
                                                                                  cap, d_out_ts, d_out_val);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
