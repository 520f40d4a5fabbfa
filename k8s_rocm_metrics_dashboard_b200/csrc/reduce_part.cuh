// reduce_part.cuh -- the per-series partial (sum, min, max, count) of K5 and the fixed order in which
// one CTA walks a row.  pmx_group_reduce (reduce.cu) and pmx_summary_quantiles (bracket_select.cu)
// both go through row_partial(), so their sums agree bit for bit.
#pragma once
#include "pmx_common.cuh"

struct Part { double sum; double mn; double mx; long long cnt; };

__device__ __forceinline__ void part_merge(Part& a, const Part& b) {
  a.sum += b.sum; a.mn = fmin(a.mn, b.mn); a.mx = fmax(a.mx, b.mx); a.cnt += b.cnt;
}
__device__ __forceinline__ Part part_identity() {
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  return Part{0.0, INF, -INF, 0};
}
__device__ __forceinline__ Part part_shfl_down(const Part& p, int d) {
  Part q;
  q.sum = __shfl_down_sync(0xffffffffu, p.sum, d);
  q.mn = __shfl_down_sync(0xffffffffu, p.mn, d);
  q.mx = __shfl_down_sync(0xffffffffu, p.mx, d);
  q.cnt = __shfl_down_sync(0xffffffffu, p.cnt, d);
  return q;
}
__device__ __forceinline__ void ld4(const double* p, double& a, double& b, double& c, double& d) {
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}

// Branch-free accumulation: a NaN adds +0.0 (the sum starts at +0.0 and can never be -0.0, so that
// changes no bit), fails both comparisons and is not counted -- 11 instructions per value.
struct Acc { double sum, mn, mx; int cnt; };
__device__ __forceinline__ void acc_add(Acc& p, double v) {
  const bool ok = v == v;
  p.sum += ok ? v : 0.0;
  p.mn = v < p.mn ? v : p.mn;
  p.mx = v > p.mx ? v : p.mx;
  p.cnt += ok ? 1 : 0;
}

// One CTA (blockDim.x threads) walks the row xr[0, n_steps): 32-byte groups, two per thread in
// flight, between a scalar head and tail; the order in which a thread meets its values is fixed by
// the row's address and length alone.  `each(v)` sees every value the thread adds, in that order;
// all lanes of a warp call it together (positions beyond the row pass NaN), so it may use warp
// collectives; `after()` runs once per trip of the main loop (at most 8 values per lane since the
// last one, plus one before the loop), also warp-convergent.  Returns this thread's partial;
// combine with cta_part().
template <class F, class G>
__device__ __forceinline__ Part row_partial(const double* __restrict__ xr, int64_t n_steps, F&& each, G&& after) {
  const double qnan = __longlong_as_double(0x7FF8000000000000LL);
  Acc a{0.0, __longlong_as_double(0x7FF0000000000000LL), __longlong_as_double((long long)0xFFF0000000000000ULL), 0};
  long long cnt = 0;
  int64_t head = (int64_t)((0 - (reinterpret_cast<uintptr_t>(xr) >> 3)) & 3);   // values before a 32-byte boundary
  if (head > n_steps) head = n_steps;
  if (head > 0) {                                                               // uniform
    const double v = (int64_t)threadIdx.x < head ? xr[threadIdx.x] : qnan;
    acc_add(a, v); each(v);
  }
  const double* xb = xr + head;
  const int64_t nb = (n_steps - head) >> 2;                                     // whole groups of four
  const int64_t per_trip = 2 * (int64_t)blockDim.x;
  // counts stay in 32 bits for 2^28 groups at a time
  for (int64_t g0 = 0; g0 < nb; g0 += (int64_t)1 << 28) {
    const int64_t g1 = g0 + ((int64_t)1 << 28) < nb ? g0 + ((int64_t)1 << 28) : nb;
    for (int64_t gt = g0; gt < g1; gt += per_trip) {                            // uniform trip count
      const int64_t g = gt + threadIdx.x, g2 = g + blockDim.x;
      double a0 = qnan, a1 = qnan, a2 = qnan, a3 = qnan, b0 = qnan, b1 = qnan, b2 = qnan, b3 = qnan;
      if (g < g1) ld4(xb + 4 * g, a0, a1, a2, a3);
      if (g2 < g1) ld4(xb + 4 * g2, b0, b1, b2, b3);
      acc_add(a, a0); each(a0); acc_add(a, a1); each(a1); acc_add(a, a2); each(a2); acc_add(a, a3); each(a3);
      acc_add(a, b0); each(b0); acc_add(a, b1); each(b1); acc_add(a, b2); each(b2); acc_add(a, b3); each(b3);
      after();
    }
    cnt += a.cnt; a.cnt = 0;
  }
  if (head + 4 * nb < n_steps) {                                                // uniform
    const int64_t tail = head + 4 * nb + threadIdx.x;
    const double v = tail < n_steps ? xr[tail] : qnan;
    acc_add(a, v); each(v);
  }
  cnt += a.cnt;
  return Part{a.sum, a.mn, a.mx, cnt};
}

// fixed shuffle tree + warp order: thread 0 returns the CTA's partial (sh: one Part per warp)
__device__ __forceinline__ Part cta_part(Part p, Part* sh) {
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) { Part q = part_shfl_down(p, d); part_merge(p, q); }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = p;
  __syncthreads();
  Part r = sh[0];
  if (threadIdx.x == 0)
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) part_merge(r, sh[w]);
  return r;
}

// One warp folds the partials of group g: lane l takes the group's series l, l + 32, ... in order
// (eight on their way at a time), then a fixed shuffle tree combines the 32 lanes -- the order never
// depends on timing.  Lane 0 writes the outputs that are not NULL.
__device__ __forceinline__ void group_fold_warp(const Part* __restrict__ part, const int32_t* __restrict__ perm,
                                                const int64_t* __restrict__ goff, int64_t g, int lane, double* o_sum,
                                                int64_t* o_cnt, double* o_min, double* o_max, double* o_mean) {
  Part r = part_identity();
  const int64_t k1 = goff[g + 1];
  int64_t k = goff[g] + lane;
  for (; k + 7 * 32 < k1; k += 8 * 32) {
    int32_t idx[8]; Part p[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) idx[u] = perm[k + 32 * u];
#pragma unroll
    for (int u = 0; u < 8; ++u) p[u] = part[idx[u]];
#pragma unroll
    for (int u = 0; u < 8; ++u) part_merge(r, p[u]);
  }
  for (; k < k1; k += 32) part_merge(r, part[perm[k]]);
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) { Part q = part_shfl_down(r, d); part_merge(r, q); }
  if (lane != 0) return;
  const bool any = r.cnt > 0;
  if (o_sum) o_sum[g] = r.sum;
  if (o_cnt) o_cnt[g] = r.cnt;
  if (o_min) o_min[g] = any ? r.mn : pmx_nan();
  if (o_max) o_max[g] = any ? r.mx : pmx_nan();
  if (o_mean) o_mean[g] = any ? r.sum / (double)r.cnt : pmx_nan();
}

// threads of a CTA that walks one row (series_partial_kernel and bs_sweep_kernel: the same walk)
constexpr int PMX_ROW_THREADS = 128;
