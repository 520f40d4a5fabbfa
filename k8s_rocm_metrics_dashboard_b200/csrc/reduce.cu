// reduce.cu -- K5: NaN-skipping group-by reductions (app.py:216-221, :335-345).
//
// Groups arrive as a CSR over series (perm/goff) so that every sum visits its inputs in a
// fixed order: results are deterministic and, for the per-step form, strictly sequential in
// series order -- the order pandas' object-dtype column reductions use at app.py:218-220.
#include <math.h>
#include "pmx_common.cuh"
#include "reduce_part.cuh"

namespace {

__device__ __forceinline__ void part_add(Part& p, double v) {
  if (v == v) { p.sum += v; p.mn = fmin(p.mn, v); p.mx = fmax(p.mx, v); p.cnt++; }
}

// Scratch of pmx_group_reduce_noalloc: lives in the library image, so no allocation is ever made
// for it.  Every GPU that loads libpmx.so holds these 2 MiB until the process exits.
constexpr int64_t NOALLOC_MAX_SERIES = 65536;
static __device__ __align__(256) unsigned char g_noalloc_partials[NOALLOC_MAX_SERIES * sizeof(Part)];
constexpr int NOALLOC_MAX_DEVICES = 64;
static Part* g_noalloc_addr[NOALLOC_MAX_DEVICES];      // device address of the array, per CUDA device

// one CTA per series: partial (sum, min, max, count) over all steps -> part[s]  (reduce_part.cuh)
__global__ void __launch_bounds__(PMX_ROW_THREADS)
series_partial_kernel(const double* __restrict__ x, int64_t n_steps, Part* __restrict__ part) {
  const int64_t s = blockIdx.x;
  __shared__ Part sh[8];
  const Part r = cta_part(row_partial(x + s * n_steps, n_steps, [](double) {}, []() {}), sh);
  if (threadIdx.x == 0) part[s] = r;
}

// one warp per group (group_fold_warp, reduce_part.cuh)
__global__ void __launch_bounds__(128)
group_fold_kernel(const Part* __restrict__ part, const int32_t* __restrict__ perm,
                  const int64_t* __restrict__ goff, int64_t n_groups,
                  double* o_sum, int64_t* o_cnt, double* o_min, double* o_max, double* o_mean) {
  const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= n_groups) return;
  group_fold_warp(part, perm, goff, g, threadIdx.x & 31, o_sum, o_cnt, o_min, o_max, o_mean);
}

// per-step form: thread per (group, step), series visited sequentially in group order
__global__ void __launch_bounds__(256)
group_step_kernel(const double* __restrict__ x, int64_t n_steps, const int32_t* __restrict__ perm,
                  const int64_t* __restrict__ goff, double* o_sum, int64_t* o_cnt, double* o_min,
                  double* o_max, double* o_mean) {
  const int64_t g = blockIdx.y;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_steps) return;
  Part r = part_identity();
  // The next eight values are already on their way while the current eight are added in series
  // order (the order of the additions is unchanged).
  int64_t k = goff[g];
  const int64_t k1 = goff[g + 1];
  if (k + 8 <= k1) {
    double v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = x[(int64_t)perm[k + u] * n_steps + j];
    k += 8;
    for (; k + 8 <= k1; k += 8) {
      double w[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) w[u] = x[(int64_t)perm[k + u] * n_steps + j];
#pragma unroll
      for (int u = 0; u < 8; ++u) part_add(r, v[u]);
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = w[u];
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) part_add(r, v[u]);
  }
  for (; k < k1; ++k) part_add(r, x[(int64_t)perm[k] * n_steps + j]);
  const bool any = r.cnt > 0;
  const int64_t o = g * n_steps + j;
  if (o_sum) o_sum[o] = r.sum;
  if (o_cnt) o_cnt[o] = r.cnt;
  if (o_min) o_min[o] = any ? r.mn : pmx_nan();
  if (o_max) o_max[o] = any ? r.mx : pmx_nan();
  if (o_mean) o_mean[o] = any ? r.sum / (double)r.cnt : pmx_nan();
}


// Multi-GPU exchange of group partials: pack puts one rank's (sum, count, min, max)
// into one buffer of 4 x n doubles (empty groups: min +inf, max -inf; counts are exact in a
// double below 2^53), one all-gather moves the buffers, combine folds the ranks in rank order
// from identical data on every rank, so every rank ends with the same bits.
__global__ void group_pack_kernel(const double* __restrict__ sum, const int64_t* __restrict__ cnt,
                                  const double* __restrict__ mn, const double* __restrict__ mx, int64_t n,
                                  double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int64_t c = cnt[i];
  out[i] = sum[i]; out[n + i] = (double)c;
  out[2 * n + i] = c > 0 ? mn[i] : INF; out[3 * n + i] = c > 0 ? mx[i] : -INF;
}
__global__ void group_combine_kernel(const double* __restrict__ all, int world, int64_t n, double* o_sum,
                                     int64_t* o_cnt, double* o_min, double* o_max, double* o_mean) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Part r = part_identity();
  double c = 0.0;
  for (int w = 0; w < world; ++w) {
    const double* b = all + (size_t)w * 4 * n;
    r.sum += b[i]; c += b[n + i]; r.mn = fmin(r.mn, b[2 * n + i]); r.mx = fmax(r.mx, b[3 * n + i]);
  }
  r.cnt = (long long)(c + 0.5);
  const bool any = r.cnt > 0;
  if (o_sum) o_sum[i] = r.sum;
  if (o_cnt) o_cnt[i] = r.cnt;
  if (o_min) o_min[i] = any ? r.mn : pmx_nan();
  if (o_max) o_max[i] = any ? r.mx : pmx_nan();
  if (o_mean) o_mean[i] = any ? r.sum / (double)r.cnt : pmx_nan();
}

}  // namespace

extern "C" {

int pmx_group_index(const int32_t* h_group_id, int64_t n_series, int64_t n_groups, int32_t* h_perm,
                    int64_t* h_goff) {
  if (!h_goff || n_series < 0 || n_groups < 0 || (n_series > 0 && (!h_group_id || !h_perm)))
    return pmx_fail(PMX_E_INVALID, "pmx_group_index: bad argument");
  if (n_series > 0x7fffffffLL) return pmx_fail(PMX_E_LIMIT, "pmx_group_index: more than 2^31-1 series");
  for (int64_t g = 0; g <= n_groups; ++g) h_goff[g] = 0;
  for (int64_t s = 0; s < n_series; ++s) {
    const int32_t g = h_group_id[s];
    if (g >= n_groups) return pmx_fail(PMX_E_INVALID, "pmx_group_index: group id %d >= n_groups", g);
    if (g >= 0) h_goff[g + 1]++;
  }
  for (int64_t g = 0; g < n_groups; ++g) h_goff[g + 1] += h_goff[g];
  // counting sort, stable: series stay in ascending order inside each group
  int64_t* cursor = new int64_t[n_groups > 0 ? n_groups : 1];
  for (int64_t g = 0; g < n_groups; ++g) cursor[g] = h_goff[g];
  for (int64_t s = 0; s < n_series; ++s) {
    const int32_t g = h_group_id[s];
    if (g >= 0) h_perm[cursor[g]++] = (int32_t)s;
  }
  delete[] cursor;
  return PMX_OK;
}

int pmx_group_reduce(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                     const int32_t* d_perm, const int64_t* d_goff, int64_t n_groups, int per_step,
                     double* d_sum, int64_t* d_count, double* d_min, double* d_max, double* d_mean,
                     pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0 && n_groups >= 0, "bad sizes");
  if (n_groups == 0) return PMX_OK;
  PMX_REQUIRE(d_goff, "d_goff is NULL");
  PMX_REQUIRE(n_series == 0 || (d_x && d_perm), "NULL buffer");
  cudaStream_t st = pmx_st(stream);
  if (per_step) {
    if (n_steps == 0) return PMX_OK;
    PMX_REQUIRE(n_groups <= 65535, "per_step form supports at most 65535 groups");
    dim3 grid((unsigned)((n_steps + 255) / 256), (unsigned)n_groups);
    group_step_kernel<<<grid, 256, 0, st>>>(d_x, n_steps, d_perm, d_goff, d_sum, d_count, d_min, d_max, d_mean);
    PMX_LAUNCH_CHECK();
    return PMX_OK;
  }
  Part* part = nullptr;
  if (n_series > 0) {
    part = (Part*)pmx_scratch(ctx, PMX_SCR_PARTIALS, (size_t)n_series * sizeof(Part));
    if (!part) return PMX_E_CUDA;
    series_partial_kernel<<<(unsigned)n_series, PMX_ROW_THREADS, 0, st>>>(d_x, n_steps, part);
    PMX_LAUNCH_CHECK();
  }
  group_fold_kernel<<<(unsigned)((n_groups + 3) / 4), 128, 0, st>>>(part, d_perm, d_goff, n_groups, d_sum, d_count,
                                                                      d_min, d_max, d_mean);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

// Same result as pmx_group_reduce(per_step = 0), for callers that cannot allocate device memory
// when the call is made: programs whose memory budget is already handed out in full (another
// allocator owns the rest of the GPU, so a late cudaMalloc can fail), and latency-bound refresh
// loops where the implicit device synchronisation of cudaMalloc/cudaFree is not acceptable.
// The per-series partials go to a fixed array inside the library instead of ctx scratch:
//   * at most 65536 series; more is refused with PMX_E_LIMIT;
//   * the first call on a device resolves the array's device address (cudaGetSymbolAddress, with
//     that device current) and keeps it; make that first call outside of any stream capture;
//   * no cudaMalloc / cudaFree / allocating library call happens here, and no ctx is needed;
//   * every call uses the same array.  Each call overwrites the slots it reads, so nothing has to
//     be cleared, but two calls that overlap on the GPU (different streams) race on it.
int pmx_group_reduce_noalloc(const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_perm,
                             const int64_t* d_goff, int64_t n_groups, double* d_sum, int64_t* d_count,
                             double* d_min, double* d_max, double* d_mean, pmx_stream stream) {
  if (n_series < 0 || n_steps < 0 || n_groups < 0)
    return pmx_fail(PMX_E_INVALID, "pmx_group_reduce_noalloc: negative size");
  if (n_series > NOALLOC_MAX_SERIES)
    return pmx_fail(PMX_E_LIMIT, "pmx_group_reduce_noalloc: %lld series exceed the fixed scratch of %lld; "
                    "use pmx_group_reduce", (long long)n_series, (long long)NOALLOC_MAX_SERIES);
  if (n_groups == 0) return PMX_OK;
  if (!d_goff || (n_series > 0 && (!d_x || !d_perm)))
    return pmx_fail(PMX_E_INVALID, "pmx_group_reduce_noalloc: NULL buffer");
  int dev = -1;
  PMX_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
    return pmx_fail(PMX_E_LIMIT, "pmx_group_reduce_noalloc: device ordinal %d not supported", dev);
  if (!g_noalloc_addr[dev]) {
    // in host code the array's name is only a handle; kernels need its address on THIS device
    void* p = nullptr;
    PMX_CUDA(cudaGetSymbolAddress(&p, g_noalloc_partials));
    g_noalloc_addr[dev] = static_cast<Part*>(p);
  }
  Part* part = g_noalloc_addr[dev];
  cudaStream_t st = pmx_st(stream);
  if (n_series > 0) {
    series_partial_kernel<<<(unsigned)n_series, PMX_ROW_THREADS, 0, st>>>(d_x, n_steps, part);
    PMX_LAUNCH_CHECK();
  }
  group_fold_kernel<<<(unsigned)((n_groups + 3) / 4), 128, 0, st>>>(part, d_perm, d_goff, n_groups, d_sum, d_count,
                                                                  d_min, d_max, d_mean);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_group_pack(const double* d_sum, const int64_t* d_count, const double* d_min, const double* d_max,
                   int64_t n, double* d_packed, pmx_stream stream) {
  if (n < 0) return pmx_fail(PMX_E_INVALID, "pmx_group_pack: negative size");
  if (n == 0) return PMX_OK;
  if (!d_sum || !d_count || !d_min || !d_max || !d_packed) return pmx_fail(PMX_E_INVALID, "pmx_group_pack: NULL buffer");
  group_pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, pmx_st(stream)>>>(d_sum, d_count, d_min, d_max, n, d_packed);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_group_combine(const double* d_gathered, int n_ranks, int64_t n, double* d_sum, int64_t* d_count,
                      double* d_min, double* d_max, double* d_mean, pmx_stream stream) {
  if (n < 0 || n_ranks < 1) return pmx_fail(PMX_E_INVALID, "pmx_group_combine: bad sizes");
  if (n == 0) return PMX_OK;
  if (!d_gathered) return pmx_fail(PMX_E_INVALID, "pmx_group_combine: NULL buffer");
  group_combine_kernel<<<(unsigned)((n + 255) / 256), 256, 0, pmx_st(stream)>>>(d_gathered, n_ranks, n, d_sum, d_count,
                                                                                d_min, d_max, d_mean);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
