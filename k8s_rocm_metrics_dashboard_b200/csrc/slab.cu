// slab.cu -- the series cache's tail gather (cache.py).  Every series keeps
// its recent samples in a row of a [rows][cap] slab, ascending, count[s] of them held.  A refresh
// needs, per series, the samples newer than a threshold as one CSR for the fused kernel:
//   pmx_slab_tail_lengths : first[s] = number of held samples with ts <= threshold (binary search,
//                           one thread per series), len[s] = count[s] - first[s];
//   (the caller turns len into offsets with one cumulative sum)
//   pmx_slab_tail_offsets : the same search and off[0 .. S] (exclusive sums of len) in ONE launch;
//   pmx_slab_tail_gather  : one warp per series copies ts/val[s][first[s] .. count[s]) to
//                           out[off[s] ..);
//   pmx_slab_append       : the samples of a new answer go to the ends of their rows (skipping what
//                           a row already holds, dropping its oldest samples when it is full);
//   pmx_csr_copy          : series of one CSR copied to given offsets of another (range client).
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

// The same search and the offsets in one launch: every CTA searches 1024 series and parks their tail
// lengths in off[s + 1]; the CTA that finishes last (a counter that it leaves at zero again) turns them
// into running sums, 1024 at a time.  A refresh is latency-bound: the launches saved are what counts.
__global__ void __launch_bounds__(1024)
tail_offsets_kernel(const int64_t* __restrict__ ts, const int64_t* __restrict__ count, int64_t n_series, int64_t cap,
                    int64_t threshold, int64_t* __restrict__ first, int64_t* __restrict__ off, unsigned int* __restrict__ done) {
  __shared__ long long s_warp[32];
  __shared__ long long s_total;
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n_series) {
      const int64_t* row = ts + s * cap;
      int64_t lo = 0, hi = count[s];
      const int64_t n = hi;
      while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (row[mid] <= threshold) lo = mid + 1; else hi = mid; }
      first[s] = lo;
      off[s + 1] = n - lo;
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x == 0) { s_total = 0; off[0] = 0; *done = 0; }
  __syncthreads();
  for (int64_t s0 = 0; s0 < n_series; s0 += blockDim.x) {
    const int64_t s = s0 + threadIdx.x;
    const long long len = s < n_series ? (long long)__ldcg(off + s + 1) : 0;
    long long inc = len;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const long long v = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += v; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    long long base = s_total;
    for (int w = 0; w < warp; ++w) base += s_warp[w];
    if (s < n_series) off[s + 1] = base + inc;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) s_total = base + inc;
    __syncthreads();
  }
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


// ---- append (cache.py SeriesCache.append): one warp per incoming series.
// plan: how many of the new samples the row already holds (they are ascending, so it is a prefix:
// binary search against the row's newest sample), how many old samples fall out, and whether one of
// those is still needed (newer than needed_from) -- then the whole append must not happen.
// status[0] = rows that would lose a needed sample, status[1] = samples stored.
__global__ void slab_plan_kernel(const int64_t* __restrict__ ts, const int64_t* __restrict__ count, int64_t cap,
                                 const int64_t* __restrict__ new_ts, const int64_t* __restrict__ in_off,
                                 const int64_t* __restrict__ in_row, int64_t n_in, int64_t needed_from,
                                 int64_t* __restrict__ skip, int64_t* __restrict__ drop,
                                 unsigned long long* __restrict__ status) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_in) return;
  const int64_t r = in_row[i];
  if (r < 0) { skip[i] = 0; drop[i] = -1; return; }          // not this cache's series (another rank's shard)
  const int64_t a = in_off[i], n = in_off[i + 1] - a, c = count[r];
  const int64_t* row = ts + r * cap;
  int64_t lo = 0;
  if (c > 0) {
    const int64_t last = row[c - 1];
    int64_t hi = n;
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (new_ts[a + mid] <= last) lo = mid + 1; else hi = mid; }
  }
  const int64_t add = n - lo, total = c + add, d = total > cap ? total - cap : 0;
  skip[i] = lo; drop[i] = d;
  if (d > 0 && (d > c || row[d - 1] > needed_from)) atomicAdd(&status[0], 1ULL);
  if (add > 0) atomicAdd(&status[1], (unsigned long long)add);
}

// apply: row <- (old samples from `drop` on) ++ (new samples from `skip` on), in place.  A warp walks
// the row upwards 32 positions at a time; what it reads lies at or above what it writes.
__global__ void __launch_bounds__(256)
slab_apply_kernel(int64_t* __restrict__ ts, double* __restrict__ val, int64_t* __restrict__ count, int64_t cap,
                  const int64_t* __restrict__ new_ts, const double* __restrict__ new_val,
                  const int64_t* __restrict__ in_off, const int64_t* __restrict__ in_row, int64_t n_in,
                  const int64_t* __restrict__ skip, const int64_t* __restrict__ drop,
                  const unsigned long long* __restrict__ status) {
  if (status[0] != 0) return;                                // a needed sample would be lost: nothing is touched
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= n_in) return;
  const int64_t r = in_row[i], d = drop[i];
  if (r < 0 || d < 0) return;
  const int64_t a = in_off[i] + skip[i], add = in_off[i + 1] - a, c = count[r];
  if (add <= 0) return;
  const int64_t total = c + add - d;                         // = min(c + add, cap)
  int64_t* rt = ts + r * cap; double* rv = val + r * cap;
  const int64_t from = d > 0 ? 0 : c;                        // nothing moves when nothing falls out
  for (int64_t p0 = from; p0 < total; p0 += 32) {
    const int64_t p = p0 + lane, q = p + d;
    int64_t t = 0; double v = 0.0;
    if (p < total) {
      if (q < c) { t = rt[q]; v = rv[q]; } else { t = new_ts[a + (q - c)]; v = new_val[a + (q - c)]; }
    }
    __syncwarp();
    if (p < total) { rt[p] = t; rv[p] = v; }
    __syncwarp();
  }
  if (lane == 0) count[r] = total;
}

// copy the series of one CSR to given positions of another (range client: the answers to consecutive
// sub-ranges concatenated per label set).  One warp per series.
__global__ void __launch_bounds__(256)
csr_copy_kernel(const int64_t* __restrict__ ts, const double* __restrict__ val, const int64_t* __restrict__ off,
                const int64_t* __restrict__ dst, int64_t n_series, int64_t* __restrict__ out_ts,
                double* __restrict__ out_val) {
  const int64_t s = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (s >= n_series) return;
  const int64_t a = off[s], n = off[s + 1] - a, b = dst[s];
  for (int64_t i = lane; i < n; i += 32) { out_ts[b + i] = ts[a + i]; out_val[b + i] = val[a + i]; }
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

int pmx_slab_tail_offsets(pmx_ctx* ctx, const int64_t* d_ts, const int64_t* d_count, int64_t n_series, int64_t cap,
                          int64_t threshold_ms, int64_t* d_first, int64_t* d_off, uint32_t* d_work, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && cap >= 1, "bad sizes");
  PMX_REQUIRE(d_off && d_work, "NULL buffer");
  if (n_series == 0) { PMX_CUDA(cudaMemsetAsync(d_off, 0, sizeof(int64_t), pmx_st(stream))); return PMX_OK; }
  PMX_REQUIRE(d_ts && d_count && d_first, "NULL buffer");
  tail_offsets_kernel<<<(unsigned)((n_series + 1023) / 1024), 1024, 0, pmx_st(stream)>>>(d_ts, d_count, n_series, cap, threshold_ms,
                                                                                       d_first, d_off, d_work);
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

int pmx_slab_append(pmx_ctx* ctx, int64_t* d_ts, double* d_val, int64_t* d_count, int64_t cap,
                    const int64_t* d_new_ts, const double* d_new_val, const int64_t* d_in_off, const int64_t* d_in_row,
                    int64_t n_in, int64_t needed_from_ms, int64_t* d_work, int64_t* d_status, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_in >= 0 && cap >= 1, "bad sizes");
  PMX_REQUIRE(d_status, "d_status is required");
  cudaStream_t st = pmx_st(stream);
  PMX_CUDA(cudaMemsetAsync(d_status, 0, 2 * sizeof(int64_t), st));
  if (n_in == 0) return PMX_OK;
  PMX_REQUIRE(d_ts && d_val && d_count && d_new_ts && d_new_val && d_in_off && d_in_row && d_work, "NULL buffer");
  int64_t* skip = d_work; int64_t* drop = d_work + n_in;
  slab_plan_kernel<<<(unsigned)((n_in + 255) / 256), 256, 0, st>>>(d_ts, d_count, cap, d_new_ts, d_in_off, d_in_row, n_in,
                                                                  needed_from_ms, skip, drop, (unsigned long long*)d_status);
  PMX_LAUNCH_CHECK();
  slab_apply_kernel<<<(unsigned)((n_in * 32 + 255) / 256), 256, 0, st>>>(d_ts, d_val, d_count, cap, d_new_ts, d_new_val, d_in_off,
                                                                        d_in_row, n_in, skip, drop,
                                                                        (const unsigned long long*)d_status);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_csr_copy(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off, const int64_t* d_dst,
                 int64_t n_series, int64_t* d_out_ts, double* d_out_val, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0, "bad sizes");
  if (n_series == 0) return PMX_OK;
  PMX_REQUIRE(d_off && d_dst && d_out_ts && d_out_val, "NULL buffer");
  csr_copy_kernel<<<(unsigned)((n_series * 32 + 255) / 256), 256, 0, pmx_st(stream)>>>(d_ts, d_val, d_off, d_dst, n_series,
                                                                                      d_out_ts, d_out_val);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
