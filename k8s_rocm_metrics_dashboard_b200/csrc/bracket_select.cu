// bracket_select.cu -- per-segment summary (sum / count / min / max / mean) AND exact quantiles of an
// [S, T] matrix in ONE sweep of the data (SURVEY X6 + X7 on one GPU; the pandas forms are
// groupby(...).agg(['sum','count','min','max','mean']) and Series.quantile, which the reference's
// app.py:216-221 / :335-345 do for the instant vector).
//
// A radix selection reads the data once per digit, after a first read for the summary.  Here a
// small sample (one value per 64 steps of every row, at a hashed position) fixes, per quantile, a
// bracket [lo, hi] of two sample VALUES that holds the wanted order statistics unless the sample is
// off by more than five standard deviations.  One sweep of the data then
//   (a) forms the per-series partial sums exactly as pmx_group_reduce does (reduce_part.cuh),
//   (b) counts, per bracket, the values below lo, equal to lo and equal to hi, and
//   (c) copies the keys of the values strictly inside (lo, hi) -- a few per cent -- to a list.
// The order statistic of rank k is lo, hi, or the (k - below - equal_lo)-th smallest key of the list:
// exact whenever the counts place k inside the bracket, which the last kernel checks.  A segment
// where they do not (or whose list overflowed) is counted in *d_failed and gets NaN; the caller then
// runs the per-pass selection (pmx_group_reduce + pmx_quantile_*), which needs no luck.
// Repeated values cost nothing: bounds are data values, and what equals a bound is counted, not
// copied.  -0.0 and +0.0 compare equal in (b); when a bound is zero the sweep also counts the
// negative zeros so that the run of zeros can be split in key order (-0.0 first).
#include <math.h>
#include "pmx_common.cuh"
#include "quantile_keys.cuh"
#include "reduce_part.cuh"
#include <cooperative_groups.h>

namespace {

constexpr int BS_MAX_Q = 4;
constexpr int BS_BINS = 2048;
constexpr int BS_STRIDE = 64;                    // one sampled value per this many steps of a row
constexpr double BS_SIGMAS = 5.0;
constexpr int BS_SMALL = 256;                    // a bucket this small is ranked directly
constexpr int BS_BOUNDS_W = 2 * BS_MAX_Q + 4;    // several ranks, per segment: lo[4], hi[4], sample size, smallest and largest sample value, pad
constexpr int BS_WBUF = 576;                     // values a warp can stage: 256 kept + a trip of 8 per lane + head and tail

struct BsSeg {                                   // per segment
  double lo[BS_MAX_Q], hi[BS_MAX_Q];             // bracket bounds: sample values, or -inf / +inf
  unsigned long long below[BS_MAX_Q], eq_lo[BS_MAX_Q], eq_hi[BS_MAX_Q];
  unsigned long long neg_zero;                   // -0.0 values (counted only when a bound is zero)
  unsigned int list_n;                           // keys appended to the list (may exceed its capacity)
  unsigned int zero_bound;
  unsigned int sample_m;                         // valid sample values the bounds were taken from
  unsigned int overflow;                         // several ranks: some rank's list overflowed
  double smin, smax;                             // smallest / largest sample value of this rank
  double gmin, gmax;                             // several ranks: smallest / largest sample value over all ranks
  unsigned long long n_valid;                    // values that are not NaN (this rank's rows)
};

// Several ranks take two digits only, so they have to spread the keys inside the bracket: the digits
// are taken relative to the bracket clipped to the range of the sample values of all ranks (an open
// side would otherwise stretch the key range to +-inf and put every key into one bucket).  Keys
// beyond the clipped range (values the samples did not see) fall into the first / last bucket: any
// monotone map of the keys partitions them correctly.
struct BsGeom { uint64_t kbase, range; int Z; };
__device__ __forceinline__ BsGeom bs_digit_geom(const BsSeg& B, int j) {
  const double lo = B.lo[j] > B.gmin ? B.lo[j] : B.gmin, hi = B.hi[j] < B.gmax ? B.hi[j] : B.gmax;
  const uint64_t a = f64_key(lo), b = f64_key(hi);
  BsGeom G;
  G.kbase = a; G.range = b > a ? b - a : 0;
  G.Z = G.range ? __clzll((long long)G.range) : 0;
  return G;
}
__device__ __forceinline__ uint64_t bs_rel(const BsGeom& G, uint64_t key) {      // 64 bits, the digits from the top
  uint64_t r = key > G.kbase ? key - G.kbase : 0;
  r = r > G.range ? G.range : r;
  return r << G.Z;
}

__device__ __forceinline__ uint64_t bs_hash(uint64_t h) {
  h ^= h >> 33; h *= 0xff51afd7ed558ccdULL; h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL; h ^= h >> 33;
  return h;
}

// sample slot (row, i): one hashed position in the i-th stretch of BS_STRIDE steps of the row
__global__ void __launch_bounds__(256)
bs_sample_kernel(const double* __restrict__ x, int64_t T, int64_t n_series, int64_t P, double* __restrict__ samp) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_series * P) return;
  const int64_t row = idx / P, i = idx - row * P;
  const int64_t first = i * BS_STRIDE;
  const int64_t span = T - first < BS_STRIDE ? T - first : BS_STRIDE;
  const int64_t pos = first + (int64_t)(bs_hash((uint64_t)row * 0x9E3779B97F4A7C15ULL + (uint64_t)i) % (uint64_t)span);
  samp[idx] = x[row * T + pos];
}

// The same slots filled from the series' raw samples (CSR) instead of the aligned matrix: slot i of
// a row takes the sample at the same relative position in the series.  The brackets are only hints
// that the sweep verifies, and a series' samples have the distribution of its grid points, so the
// brackets can be ready before the matrix exists (the aggregate step computes them beside the fused kernel).
__global__ void __launch_bounds__(256)
bs_sample_csr_kernel(const double* __restrict__ val, const int64_t* __restrict__ off, int64_t T, int64_t n_series, int64_t P,
                     double* __restrict__ samp) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_series * P) return;
  const int64_t row = idx / P, i = idx - row * P;
  const int64_t a = off[row], len = off[row + 1] - a;
  if (len <= 0) { samp[idx] = pmx_nan(); return; }
  const int64_t first = i * BS_STRIDE;
  const int64_t span = T - first < BS_STRIDE ? T - first : BS_STRIDE;
  const int64_t step = first + (int64_t)(bs_hash((uint64_t)row * 0x9E3779B97F4A7C15ULL + (uint64_t)i) % (uint64_t)span);
  int64_t k = (int64_t)((double)step * (double)len / (double)T);
  k = k < 0 ? 0 : (k > len - 1 ? len - 1 : k);
  samp[idx] = val[a + k];
}

// block-wide exclusive scan of BS_BINS counters by 1024 threads, two bins (c0, c1) each; returns
// this thread's exclusive base.  s_warp: 32 words.
__device__ __forceinline__ unsigned bs_scan2(unsigned c0, unsigned c1, unsigned* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned sum = c0 + c1;
  unsigned inc = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += v; }
  __syncthreads();                               // s_warp may still be read from the previous scan
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {                               // exclusive scan of the 32 warp totals
    const unsigned t = s_warp[lane];
    unsigned w = t;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += v; }
    s_warp[lane] = w - t;
  }
  __syncthreads();
  return s_warp[warp] + inc - sum;
}
// A cluster of BS_CL CTAs per segment: the sample values of the brackets.  Every CTA keeps an
// eighth of the segment's sample as keys in its shared memory; histograms and extremes are
// exchanged through distributed shared memory, and every CTA repeats the (cheap) scans and picks,
// so nothing has to be broadcast.  For quantile q of m valid sample values the bracket is
// [value of sample rank floor(q(m-1) - d), value of sample rank ceil(q(m-1) + d) + 1],
// d = 5 sqrt(m q (1-q)) + 3, rounded outwards to the smallest / largest sample value that shares the
// first 22 significant key bits with it (two histogram sweeps instead of six); a rank outside the
// sample leaves that side open (-inf / +inf).  Also clears the segment's counters.
constexpr int BS_CL = 8;
constexpr unsigned BS_KCAP = 12288;              // keys per CTA (96 KiB); a larger sample is thinned
__global__ void __cluster_dims__(BS_CL, 1, 1) __launch_bounds__(1024)
bs_bounds_kernel(const double* __restrict__ samp, int64_t P, const int32_t* __restrict__ perm,
                 const int64_t* __restrict__ goff, const double* __restrict__ q, int n_q, BsSeg* __restrict__ segs,
                 double* __restrict__ exported) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) unsigned char bs_smem[];
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(bs_smem);            // [BS_KCAP]
  unsigned int* sh_hist = reinterpret_cast<unsigned int*>(sk + BS_KCAP);              // [2 n_q][BS_BINS]
  __shared__ unsigned int s_warp[32];
  __shared__ unsigned long long x_min, x_max, x_key[2 * BS_MAX_Q];                    // read by the other CTAs
  __shared__ unsigned int x_n;
  __shared__ unsigned long long s_key[2 * BS_MAX_Q];
  __shared__ long long s_rank[2 * BS_MAX_Q];     // sample rank wanted (-1: open side)
  __shared__ unsigned int s_bin1[2 * BS_MAX_Q], s_bin2[2 * BS_MAX_Q];
  __shared__ int s_u[2 * BS_MAX_Q], s_nu;
  __shared__ unsigned int s_ubin[2 * BS_MAX_Q];
  const int g = blockIdx.x / BS_CL;
  const unsigned rank = cluster.block_rank();
  const int n_t = 2 * n_q;
  const int64_t k0 = goff[g], rows = goff[g + 1] - k0;
  // slots of the segment: (row r of the group, slot i of the row); thinned to BS_CL * BS_KCAP at most
  const uint64_t slots = (uint64_t)rows * (uint64_t)P;
  const uint64_t skip = (slots + (uint64_t)BS_CL * BS_KCAP - 1) / ((uint64_t)BS_CL * BS_KCAP) > 0
                            ? (slots + (uint64_t)BS_CL * BS_KCAP - 1) / ((uint64_t)BS_CL * BS_KCAP) : 1;
  const uint64_t used = (slots + skip - 1) / skip;                                   // <= BS_CL * BS_KCAP
  const unsigned per = (unsigned)((used + BS_CL - 1) / BS_CL);                        // <= BS_KCAP
  if (threadIdx.x == 0) { x_min = ~0ULL; x_max = 0; x_n = 0; }
  __syncthreads();
  {
    unsigned long long mn = ~0ULL, mx = 0;
    const uint64_t a = (uint64_t)rank * per, b = a + per < used ? a + per : used;
    auto slot_value = [&](uint64_t u) -> double {
      // a hashed slot of every `skip` consecutive ones: a fixed phase would take the same stretches of
      // time from every row, and rows that move together (counters) would leave gaps in the sample
      uint64_t idx = u * skip + (skip > 1 ? bs_hash(u + ((uint64_t)g << 40)) % skip : 0);
      if (idx >= slots) idx = slots - 1;
      uint64_t r, i;
      if (slots <= 0xffffffffULL) { r = (unsigned)idx / (unsigned)P; i = (unsigned)idx - (unsigned)r * (unsigned)P; }
      else { r = idx / (uint64_t)P; i = idx - r * (uint64_t)P; }
      return samp[(int64_t)perm[k0 + (int64_t)r] * P + (int64_t)i];
    };
    for (uint64_t u0 = a + threadIdx.x; u0 < b; u0 += 4 * (uint64_t)blockDim.x) {          // four loads in flight
      double v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) { const uint64_t u = u0 + (uint64_t)k * blockDim.x; v[k] = u < b ? slot_value(u) : pmx_nan(); }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (v[k] != v[k]) continue;
        const uint64_t key = f64_key(v[k]);
        sk[atomicAdd(&x_n, 1u)] = key;
        mn = key < mn ? key : mn; mx = key > mx ? key : mx;
      }
    }
    // one pair of 64-bit atomics per warp
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
      const unsigned long long a2 = __shfl_xor_sync(0xffffffffu, mn, d), b2 = __shfl_xor_sync(0xffffffffu, mx, d);
      mn = a2 < mn ? a2 : mn; mx = b2 > mx ? b2 : mx;
    }
    if ((threadIdx.x & 31) == 0 && mn != ~0ULL) { atomicMin(&x_min, mn); atomicMax(&x_max, mx); }
  }
  cluster.sync();
  const unsigned n_local = x_n;
  unsigned m = 0; uint64_t kmin = ~0ULL, kmax = 0;
  for (int r = 0; r < BS_CL; ++r) {
    m += *cluster.map_shared_rank(&x_n, r);
    const uint64_t a = *cluster.map_shared_rank(&x_min, r), b = *cluster.map_shared_rank(&x_max, r);
    kmin = a < kmin ? a : kmin; kmax = b > kmax ? b : kmax;
  }
  if (threadIdx.x < n_t) {
    const int j = threadIdx.x >> 1;
    long long r = -1;
    if (m > 0) {
      const double qq = q[j], a = qq * (double)(m - 1), d = BS_SIGMAS * sqrt((double)m * qq * (1.0 - qq)) + 3.0;
      if (threadIdx.x & 1) { const double h = ceil(a + d) + 1.0; r = h > (double)(m - 1) ? -1 : (long long)h; }
      else { const double l = floor(a - d); r = l < 0.0 ? -1 : (long long)l; }
    }
    s_rank[threadIdx.x] = r;
    x_key[threadIdx.x] = (threadIdx.x & 1) ? 0ULL : ~0ULL;       // max / min accumulators of the last sweep
    s_bin1[threadIdx.x] = 0; s_bin2[threadIdx.x] = 0;
  }
  __syncthreads();
  const bool spread = m > 0 && kmax > kmin;
  const int Z = spread ? __clzll((long long)(kmax - kmin)) : 0;
  if (spread) {
    // first 11 significant bits: one histogram for all ranks
    for (int i = threadIdx.x; i < BS_BINS; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    for (unsigned i = threadIdx.x; i < n_local; i += blockDim.x) atomicAdd(&sh_hist[(unsigned)(((sk[i] - kmin) << Z) >> 53)], 1u);
    cluster.sync();
    {
      long long rr[2 * BS_MAX_Q];                // read before the scan's barriers: the finder of a rank rewrites it
#pragma unroll
      for (int t = 0; t < 2 * BS_MAX_Q; ++t) rr[t] = t < n_t ? s_rank[t] : -1;
      unsigned c0 = 0, c1 = 0;
      for (int r = 0; r < BS_CL; ++r) {
        const uint2 c = *reinterpret_cast<const uint2*>(cluster.map_shared_rank(sh_hist, r) + 2 * threadIdx.x);
        c0 += c.x; c1 += c.y;
      }
      const unsigned base = bs_scan2(c0, c1, s_warp);
#pragma unroll
      for (int t = 0; t < 2 * BS_MAX_Q; ++t) {
        const long long r = rr[t];
        if (r < 0) continue;
        if ((unsigned long long)r >= base && (unsigned long long)r < base + c0) { s_bin1[t] = 2 * threadIdx.x; s_rank[t] = r - base; }
        else if ((unsigned long long)r >= base + c0 && (unsigned long long)r < base + c0 + c1) { s_bin1[t] = 2 * threadIdx.x + 1; s_rank[t] = r - base - c0; }
      }
    }
    cluster.sync();                              // every CTA has read every histogram
    // next 11 bits: one histogram per distinct first digit among the ranks
    if (threadIdx.x == 0) {
      int nu = 0;
      for (int t = 0; t < n_t; ++t) {
        s_u[t] = -1;
        if (s_rank[t] < 0) continue;
        int u = 0;
        while (u < nu && s_ubin[u] != s_bin1[t]) ++u;
        if (u == nu) s_ubin[nu++] = s_bin1[t];
        s_u[t] = u;
      }
      s_nu = nu;
    }
    __syncthreads();
    const int nu = s_nu;
    for (int i = threadIdx.x; i < nu * BS_BINS; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    unsigned ub[2 * BS_MAX_Q];
#pragma unroll
    for (int u = 0; u < 2 * BS_MAX_Q; ++u) ub[u] = u < nu ? s_ubin[u] : 0xffffffffu;
    for (unsigned i = threadIdx.x; i < n_local; i += blockDim.x) {
      const uint64_t rel = (sk[i] - kmin) << Z;
      const unsigned d1 = (unsigned)(rel >> 53), d2 = (unsigned)(rel >> 42) & (BS_BINS - 1);
#pragma unroll
      for (int u = 0; u < 2 * BS_MAX_Q; ++u) if (d1 == ub[u]) atomicAdd(&sh_hist[u * BS_BINS + d2], 1u);
    }
    cluster.sync();
    for (int u = 0; u < nu; ++u) {
      unsigned c0 = 0, c1 = 0;
      for (int r = 0; r < BS_CL; ++r) {
        const uint2 c = *reinterpret_cast<const uint2*>(cluster.map_shared_rank(sh_hist, r) + u * BS_BINS + 2 * threadIdx.x);
        c0 += c.x; c1 += c.y;
      }
      const unsigned base = bs_scan2(c0, c1, s_warp);
      for (int t = 0; t < n_t; ++t) {
        if (s_u[t] != u) continue;
        const unsigned long long r = (unsigned long long)s_rank[t];
        if (r >= base && r < base + c0) s_bin2[t] = 2 * threadIdx.x;
        else if (r >= base + c0 && r < base + c0 + c1) s_bin2[t] = 2 * threadIdx.x + 1;
      }
    }
    __syncthreads();
    // outwards to a sample value: the smallest key of the bucket for a lower bound, the largest for an upper
    unsigned tp[2 * BS_MAX_Q];                    // 22-bit prefix of each closed rank (an open one matches nothing)
#pragma unroll
    for (int t = 0; t < 2 * BS_MAX_Q; ++t) tp[t] = (t < n_t && s_rank[t] >= 0) ? ((s_bin1[t] << 11) | s_bin2[t]) : 0xffffffffu;
    for (unsigned i = threadIdx.x; i < n_local; i += blockDim.x) {
      const uint64_t key = sk[i];
      const unsigned pf = (unsigned)(((key - kmin) << Z) >> 42);
#pragma unroll
      for (int t = 0; t < 2 * BS_MAX_Q; ++t) {
        if (pf != tp[t]) continue;
        if (t & 1) { if (key > *(volatile unsigned long long*)&x_key[t]) atomicMax(&x_key[t], key); }
        else { if (key < *(volatile unsigned long long*)&x_key[t]) atomicMin(&x_key[t], key); }
      }
    }
    cluster.sync();
    if (threadIdx.x < n_t) {
      unsigned long long k = (threadIdx.x & 1) ? 0ULL : ~0ULL;
      for (int r = 0; r < BS_CL; ++r) {
        const unsigned long long o = *cluster.map_shared_rank(&x_key[threadIdx.x], r);
        k = (threadIdx.x & 1) ? (o > k ? o : k) : (o < k ? o : k);
      }
      s_key[threadIdx.x] = k;
    }
  } else if (threadIdx.x < n_t) {
    s_key[threadIdx.x] = kmin;                   // every sample value is the same (or there is none)
  }
  cluster.sync();                                // no CTA leaves while its shared memory may still be read
  if (rank == 0 && threadIdx.x == 0) {
    BsSeg& b = segs[g];
    const double INF = __longlong_as_double(0x7FF0000000000000LL);
    unsigned zb = 0;
    for (int j = 0; j < BS_MAX_Q; ++j) {
      double lo = pmx_nan(), hi = pmx_nan();
      if (j < n_q) {
        lo = s_rank[2 * j] < 0 ? -INF : key_f64(s_key[2 * j]);
        hi = s_rank[2 * j + 1] < 0 ? INF : key_f64(s_key[2 * j + 1]);
        if (lo == 0.0 || hi == 0.0) zb = 1;
      }
      b.lo[j] = lo; b.hi[j] = hi; b.below[j] = 0; b.eq_lo[j] = 0; b.eq_hi[j] = 0;
    }
    b.neg_zero = 0; b.list_n = 0; b.zero_bound = zb; b.sample_m = m; b.overflow = 0; b.n_valid = 0;
    b.smin = m ? key_f64(kmin) : INF; b.smax = m ? key_f64(kmax) : -INF; b.gmin = -INF; b.gmax = INF;
    if (exported) {                              // several ranks: what this rank proposes (lo[4], hi[4], sample size, sample range)
      double* o = exported + (size_t)g * BS_BOUNDS_W;
      for (int j = 0; j < BS_MAX_Q; ++j) { o[j] = b.lo[j]; o[BS_MAX_Q + j] = b.hi[j]; }
      o[2 * BS_MAX_Q] = (double)m; o[2 * BS_MAX_Q + 1] = b.smin; o[2 * BS_MAX_Q + 2] = b.smax; o[2 * BS_MAX_Q + 3] = 0.0;
    }
  }
}

// The sweep: one CTA per series.  Partial sums in pmx_group_reduce's order, bracket counters, and the
// keys strictly inside a bracket appended to the segment's list (staged per warp).
template <int NB>
__global__ void __launch_bounds__(PMX_ROW_THREADS)
bs_sweep_kernel(const double* __restrict__ x, int64_t T, const int32_t* __restrict__ seg, Part* __restrict__ part,
                BsSeg* __restrict__ segs, unsigned long long* __restrict__ lists, unsigned cap) {
  __shared__ Part sh_part[8];
  constexpr int NW = PMX_ROW_THREADS / 32;
  __shared__ unsigned long long s_buf[NW][BS_WBUF];
  __shared__ unsigned int s_cnt[NW][3 * NB + 1];
  const int64_t s = blockIdx.x;
  const int g = seg[s];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double lo[NB], hi[NB];
  bool zb = false;
#pragma unroll
  for (int j = 0; j < NB; ++j) { lo[j] = pmx_nan(); hi[j] = pmx_nan(); }
  if (g >= 0) {
#pragma unroll
    for (int j = 0; j < NB; ++j) { lo[j] = segs[g].lo[j]; hi[j] = segs[g].hi[j]; }
    zb = segs[g].zero_bound != 0;
  }
  unsigned c_lt[NB], c_el[NB], c_eh[NB], c_nz = 0;
#pragma unroll
  for (int j = 0; j < NB; ++j) { c_lt[j] = 0; c_el[j] = 0; c_eh[j] = 0; }
  unsigned long long* wbuf = s_buf[warp];
  unsigned wcnt = 0;                             // warp-uniform
  // The warp's staged values (all inside a closed bracket [lo, hi]) are sorted out 32 at a time, every
  // lane busy: equal to a bound -> counted; strictly inside -> its key stays, compacted in place;
  // then one atomic reserves the room in the segment's list and the keys leave coalesced.
  auto flush = [&]() {
    if (wcnt == 0) return;
    unsigned nmid = 0;
    for (unsigned i0 = 0; i0 < wcnt; i0 += 32) {
      const unsigned i = i0 + lane;
      const double v = i < wcnt ? __longlong_as_double((long long)wbuf[i]) : pmx_nan();
      bool mid = false;
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        if (v == lo[j]) ++c_el[j];
        if (v == hi[j]) ++c_eh[j];
        mid |= v > lo[j] && v < hi[j];
      }
      if (zb && (unsigned long long)__double_as_longlong(v) == 0x8000000000000000ULL) ++c_nz;
      const unsigned any = __ballot_sync(0xffffffffu, mid);
      __syncwarp();                              // every lane has read its value: slots <= i0 + 31 may be rewritten
      if (mid) wbuf[nmid + (unsigned)__popc(any & ((1u << lane) - 1u))] = f64_key(v);
      nmid += (unsigned)__popc(any);
    }
    __syncwarp();
    if (nmid) {
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(&segs[g].list_n, nmid);
      base = __shfl_sync(0xffffffffu, base, 0);
      for (unsigned i = lane; i < nmid; i += 32)
        if (base + i < cap) lists[(size_t)g * cap + base + i] = wbuf[i];
    }
    __syncwarp();
    wcnt = 0;
  };
  // Per value: one comparison and one counter per bracket (below lo) and one comparison more to see
  // whether it lies in a closed bracket [lo, hi] -- a few per cent of the values, or the repeats of
  // a bound; those are staged as they are.  NaN fails every comparison.
  const Part mine = row_partial(x + s * T, T, [&](double v) {
    bool in = false;
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const bool lt = v < lo[j];
      if (lt) ++c_lt[j];
      in |= !lt && v <= hi[j];
    }
    const unsigned any = __ballot_sync(0xffffffffu, in);
    if (any) {
      if (in) wbuf[wcnt + (unsigned)__popc(any & ((1u << lane) - 1u))] = (unsigned long long)__double_as_longlong(v);
      wcnt += (unsigned)__popc(any);
    }
  }, [&]() { if (wcnt > 256) flush(); });
  const Part r = cta_part(mine, sh_part);
  if (threadIdx.x == 0) part[s] = r;
  if (g < 0) return;
  if (threadIdx.x == 0 && r.cnt) atomicAdd(&segs[g].n_valid, (unsigned long long)r.cnt);
  flush();
  // counters: warp sums, then one 64-bit atomic per counter and CTA
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    const unsigned a = __reduce_add_sync(0xffffffffu, c_lt[j]), b = __reduce_add_sync(0xffffffffu, c_el[j]),
                   c = __reduce_add_sync(0xffffffffu, c_eh[j]);
    if (lane == 0) { s_cnt[warp][3 * j] = a; s_cnt[warp][3 * j + 1] = b; s_cnt[warp][3 * j + 2] = c; }
  }
  {
    const unsigned z = __reduce_add_sync(0xffffffffu, c_nz);
    if (lane == 0) s_cnt[warp][3 * NB] = z;
  }
  __syncthreads();
  if (threadIdx.x < 3 * NB + 1) {
    unsigned long long tot = 0;
    for (int w = 0; w < NW; ++w) tot += s_cnt[w][threadIdx.x];
    if (tot) {
      BsSeg& b = segs[g];
      unsigned long long* dst = threadIdx.x == 3 * NB ? &b.neg_zero
                               : (threadIdx.x % 3 == 0 ? &b.below[threadIdx.x / 3]
                                  : threadIdx.x % 3 == 1 ? &b.eq_lo[threadIdx.x / 3] : &b.eq_hi[threadIdx.x / 3]);
      atomicAdd(dst, tot);
    }
  }
}

// A cluster of BS_RC CTAs per (quantile j, segment g): place the two ranks of quantile j in their
// bracket and, where they fall strictly inside, select them from the list; write the outputs of
// pmx_quantile_finish.  Every CTA sweeps a quarter of the list; counts and histograms are summed over
// distributed shared memory and every CTA repeats the scans and picks; CTA 0 ranks the last few keys.
constexpr int BS_RC = 4;
__global__ void __cluster_dims__(BS_RC, 1, 1) __launch_bounds__(1024)
bs_resolve_kernel(const BsSeg* __restrict__ segs, const unsigned long long* __restrict__ lists, unsigned cap,
                  const double* __restrict__ q, int n_q,
                  int64_t* o_n, int64_t* o_lo, int64_t* o_hi, double* o_xlo, double* o_xhi, double* o_val,
                  unsigned int* __restrict__ failed, const Part* __restrict__ part, const int32_t* __restrict__ perm,
                  const int64_t* __restrict__ goff, double* o_sum, int64_t* o_cnt, double* o_min, double* o_max,
                  double* o_mean) {
  __shared__ unsigned int sh_hist[2][BS_BINS];
  __shared__ unsigned int s_warp[32];
  __shared__ unsigned long long s_keys[2 * BS_SMALL];
  __shared__ unsigned int s_nk, s_mid;          // read by the other CTAs of the cluster, as are sh_hist and s_keys
  __shared__ int s_zone[2];                      // 0 fail, 1 lo, 2 inside, 3 hi
  __shared__ unsigned long long s_r[2];          // rank inside the list (zone 2) or inside the run of zeros
  __shared__ unsigned long long s_pf[2], s_ans[2];
  __shared__ unsigned int s_cb[2];               // keys in the rank's current bucket
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned crank = cluster.block_rank();
  const int j = blockIdx.x / BS_RC, g = blockIdx.y;
  if (j == n_q) {
    // the last cluster of every segment folds the per-series partials into the summary (one warp,
    // the order of pmx_group_reduce) while the others select: it costs no time of its own
    if (crank == 0 && threadIdx.x < 32) group_fold_warp(part, perm, goff, g, threadIdx.x, o_sum, o_cnt, o_min, o_max, o_mean);
    return;
  }
  const int64_t out = (int64_t)g * n_q + j;
  const long long n = (long long)segs[g].n_valid;
  if (j == 0 && crank == 0 && threadIdx.x == 0 && o_n) o_n[g] = n;
  auto write = [&](long long klo, long long khi, double a, double b, double v) {
    if (o_lo) o_lo[out] = klo;
    if (o_hi) o_hi[out] = khi;
    if (o_xlo) o_xlo[out] = a;
    if (o_xhi) o_xhi[out] = b;
    if (o_val) o_val[out] = v;
  };
  if (n <= 0) { if (crank == 0 && threadIdx.x == 0) write(-1, -1, pmx_nan(), pmx_nan(), pmx_nan()); return; }
  const BsSeg& B = segs[g];
  const double lo = B.lo[j], hi = B.hi[j];
  const uint64_t klo = f64_key(lo), khi = f64_key(hi);
  const unsigned long long c_lt = B.below[j], c_el = B.eq_lo[j], c_eh = hi == lo ? 0 : B.eq_hi[j], c_nz = B.neg_zero;
  const unsigned list_n = B.list_n;
  const bool overflow = list_n > cap;
  const unsigned n_list = overflow ? cap : list_n;
  const unsigned long long* keys = lists + (size_t)g * cap;
  const double pos = (double)(n - 1) * q[j];
  unsigned long long k0 = (unsigned long long)floor(pos);
  if (k0 > (unsigned long long)n - 1) k0 = (unsigned long long)n - 1;
  const unsigned long long k1 = k0 + 1 < (unsigned long long)n ? k0 + 1 : (unsigned long long)n - 1;
  const double tt = pos - (double)k0;
  auto inside = [&](uint64_t key) -> bool { const double v = key_f64(key); return v > lo && v < hi; };   // as the sweep saw it
  // every key of this CTA's part of the list, eight loads in flight per thread
  const unsigned part_n = (n_list + BS_RC - 1) / BS_RC;
  const unsigned k_a = crank * part_n < n_list ? crank * part_n : n_list;
  const unsigned k_b = k_a + part_n < n_list ? k_a + part_n : n_list;
  auto for_keys = [&](auto&& f) {
    for (unsigned i0 = k_a; i0 < k_b; i0 += blockDim.x * 8) {
      unsigned long long kk[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) { const unsigned i = i0 + u * blockDim.x + threadIdx.x; kk[u] = i < k_b ? keys[i] : 0ULL; }
#pragma unroll
      for (int u = 0; u < 8; ++u) { const unsigned i = i0 + u * blockDim.x + threadIdx.x; if (i < k_b) f(kk[u]); }
    }
  };
  // the histogram h summed over the cluster: this thread's two bins
  auto hist2 = [&](int h, unsigned& c0, unsigned& c1) {
    c0 = 0; c1 = 0;
#pragma unroll
    for (int r = 0; r < BS_RC; ++r) {
      const uint2 c = *reinterpret_cast<const uint2*>(cluster.map_shared_rank(&sh_hist[h][0], r) + 2 * threadIdx.x);
      c0 += c.x; c1 += c.y;
    }
  };
  const int Z = khi > klo ? __clzll((long long)(khi - klo)) : 0;
  // first sweep of the list: the keys strictly inside this bracket, and their first digit
  for (int i = threadIdx.x; i < BS_BINS; i += blockDim.x) sh_hist[0][i] = 0;
  if (threadIdx.x == 0) { s_mid = 0; s_nk = 0; }
  __syncthreads();
  {
    unsigned c = 0;
    for_keys([&](uint64_t key) {
      if (inside(key)) { ++c; atomicAdd(&sh_hist[0][(unsigned)(((key - klo) << Z) >> 53)], 1u); }
    });
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_mid, c);
  }
  cluster.sync();
  unsigned long long c_mid = 0;
  for (int r = 0; r < BS_RC; ++r) c_mid += *cluster.map_shared_rank(&s_mid, r);
  if (threadIdx.x < 2) {
    const unsigned long long k = threadIdx.x ? k1 : k0;
    int zone = 0; unsigned long long r = 0;
    if (k < c_lt) zone = 0;
    else if (k < c_lt + c_el) { zone = 1; r = k - c_lt; }
    else if (overflow) zone = 0;                 // the list is cut short: nothing beyond lo can be placed
    else if (k < c_lt + c_el + c_mid) { zone = 2; r = k - c_lt - c_el; }
    else if (k < c_lt + c_el + c_mid + c_eh) { zone = 3; r = k - c_lt - c_el - c_mid; }
    s_zone[threadIdx.x] = zone; s_r[threadIdx.x] = r; s_pf[threadIdx.x] = 0; s_cb[threadIdx.x] = 0;
    // a bound that is zero stands for the run of -0.0 (first in key order) and +0.0 values
    if (zone == 1) s_ans[threadIdx.x] = lo == 0.0 ? f64_key(r < c_nz ? -0.0 : 0.0) : klo;
    if (zone == 3) s_ans[threadIdx.x] = hi == 0.0 ? f64_key(r < c_nz ? -0.0 : 0.0) : khi;
  }
  __syncthreads();
  if (s_zone[0] == 0 || s_zone[1] == 0) {        // the same in every CTA of the cluster
    if (crank == 0 && threadIdx.x == 0) { atomicAdd(failed, 1u); write((long long)k0, (long long)k1, pmx_nan(), pmx_nan(), pmx_nan()); }
    cluster.sync();                              // nobody leaves while its counters may still be read
    return;
  }
  if (s_zone[0] == 2 || s_zone[1] == 2) {
    // radix passes over the inside keys until every wanted rank sits in a bucket of <= BS_SMALL keys
    int used = 0;                                // digit bits consumed so far
    for (int pass = 0;; ++pass) {
      const int bits = 64 - used >= 11 ? 11 : 64 - used;             // of this pass (the low Z bits of rel are zero)
      const int n_h = (pass == 0 || s_pf[0] == s_pf[1] || s_zone[0] != 2 || s_zone[1] != 2) ? 1 : 2;
      if (pass > 0) {
        for (int i = threadIdx.x; i < n_h * BS_BINS; i += blockDim.x) (&sh_hist[0][0])[i] = 0;
        __syncthreads();
        const uint64_t p0 = s_zone[0] == 2 ? s_pf[0] : s_pf[1], p1 = s_zone[1] == 2 ? s_pf[1] : s_pf[0];
        for_keys([&](uint64_t key) {
          if (!inside(key)) return;
          const uint64_t rel = (key - klo) << Z;
          const uint64_t top = rel >> (64 - used);
          const unsigned d = (unsigned)((rel << used) >> (64 - bits));
          if (top == p0) atomicAdd(&sh_hist[0][d], 1u);
          else if (n_h == 2 && top == p1) atomicAdd(&sh_hist[1][d], 1u);
        });
        cluster.sync();
      }
      for (int h = 0; h < n_h; ++h) {
        // read before the scan's barriers: the thread that finds a rank's bin rewrites rank and prefix
        const unsigned long long r0 = s_r[0], r1 = s_r[1], f0 = s_pf[0], f1 = s_pf[1];
        unsigned c0, c1;
        hist2(h, c0, c1);
        const unsigned base = bs_scan2(c0, c1, s_warp);
        for (int t = 0; t < 2; ++t) {
          if (s_zone[t] != 2) continue;
          const int ht = (n_h == 2 && t == 1) ? 1 : 0;               // the histogram that serves rank t
          if (ht != h) continue;
          const unsigned long long r = t ? r1 : r0, f = t ? f1 : f0;
          if (r >= base && r < (unsigned long long)base + c0) { s_pf[t] = (f << bits) | (2 * threadIdx.x); s_r[t] = r - base; s_cb[t] = c0; }
          else if (r >= (unsigned long long)base + c0 && r < (unsigned long long)base + c0 + c1) {
            s_pf[t] = (f << bits) | (2 * threadIdx.x + 1); s_r[t] = r - base - c0; s_cb[t] = c1;
          }
        }
        __syncthreads();
      }
      used += bits;
      cluster.sync();                            // every CTA has read every histogram before any is cleared
      const bool small = (s_zone[0] != 2 || s_cb[0] <= BS_SMALL) && (s_zone[1] != 2 || s_cb[1] <= BS_SMALL);
      if (small || used >= 64 - Z) break;
    }
    if (used >= 64 - Z) {
      // every bit is settled: the prefix is the key (all keys of the bucket are equal)
      if (threadIdx.x < 2 && s_zone[threadIdx.x] == 2) s_ans[threadIdx.x] = ((s_pf[threadIdx.x] << (64 - used)) >> Z) + klo;
    } else {
      // gather the buckets and rank them directly
      const uint64_t p0 = s_zone[0] == 2 ? s_pf[0] : s_pf[1], p1 = s_zone[1] == 2 ? s_pf[1] : s_pf[0];
      for_keys([&](uint64_t key) {
        if (!inside(key)) return;
        const uint64_t top = ((key - klo) << Z) >> (64 - used);
        if (top == p0 || top == p1) { const unsigned at = atomicAdd(&s_nk, 1u); if (at < 2 * BS_SMALL) s_keys[at] = key; }
      });
      cluster.sync();
      // CTA 0 takes the other CTAs' keys behind its own (the buckets hold <= 2 BS_SMALL keys together)
      unsigned nk = s_nk < 2 * BS_SMALL ? s_nk : 2 * BS_SMALL;
      if (crank == 0) {
        for (int r = 1; r < BS_RC; ++r) {
          const unsigned nr = *cluster.map_shared_rank(&s_nk, r);
          const unsigned long long* rk = cluster.map_shared_rank(&s_keys[0], r);
          for (unsigned i = threadIdx.x; i < nr; i += blockDim.x) if (nk + i < 2 * BS_SMALL) s_keys[nk + i] = rk[i];
          nk = nk + nr < 2 * BS_SMALL ? nk + nr : 2 * BS_SMALL;
        }
        __syncthreads();
      } else {
        nk = 0;
      }
      for (unsigned i = threadIdx.x; i < nk; i += blockDim.x) {
        const uint64_t ki = s_keys[i], top = ((ki - klo) << Z) >> (64 - used);
        unsigned less = 0, eq = 0;
        for (unsigned u = 0; u < nk; ++u) {
          const uint64_t ku = s_keys[u];
          if ((((ku - klo) << Z) >> (64 - used)) == top) { less += ku < ki; eq += ku == ki; }
        }
        for (int t = 0; t < 2; ++t)
          if (s_zone[t] == 2 && s_pf[t] == top && s_r[t] >= less && s_r[t] < (unsigned long long)less + eq) s_ans[t] = ki;
      }
    }
    __syncthreads();
  }
  cluster.sync();                                // nobody leaves while its shared memory may still be read
  if (crank == 0 && threadIdx.x == 0) {
    const double a = key_f64(s_ans[0]), b = key_f64(s_ans[1]);
    // numpy _lerp: a + (b-a)*t, and b - (b-a)*(1-t) when t >= 0.5 (no FMA contraction)
    const double d = __dsub_rn(b, a);
    double v = __dadd_rn(a, __dmul_rn(d, tt));
    if (tt >= 0.5) v = __dsub_rn(b, __dmul_rn(d, __dsub_rn(1.0, tt)));
    write((long long)k0, (long long)k1, a, b, v);
  }
}

// ------------------------------------------------------------------------------------------------
// Several ranks (one process per GPU, rows sharded): every rank samples its own rows and proposes
// brackets; the hull of the proposals (the global quantile lies between the smallest and the largest
// local one) is what all ranks sweep with; counters, summary partials and the first digit's histogram
// of the keys inside every bracket are gathered and combined in rank order; the second digit's
// histograms are summed; every rank then contributes the few keys of the 22-bit buckets that hold
// the wanted ranks, and all ranks settle every quantile alike from the gathered rows.  Four small
// exchanges (NCCL calls, or stores into peer memory: peer.cu), one sweep of the data.
constexpr int BS_PACK_W = 4 + 3 * BS_MAX_Q + 2;  // sum, count, min, max | below, eq_lo, eq_hi | -0.0 count, overflow
constexpr unsigned BS_ROWK = 1024;               // keys a rank can contribute per (segment, quantile)
constexpr unsigned BS_ROW_W = 5 + BS_ROWK + 1;   // count, (min, max) of either rank's bucket, keys, pad (rows stay multiples of 16 bytes)

struct BsPick {                                  // per (segment, quantile): where its two ranks fall
  int zone[2];                                   // -1 empty segment, 0 fail, 1 lo, 2 inside, 3 hi
  unsigned int bin[2];                           // zone 2: first two digits (22 bits) of the bucket
  unsigned long long r[2];                       // zone 2: rank inside the bucket
  unsigned int bin2[2];                          // ... after the second digit (22 bits)
  unsigned long long r2[2];
  unsigned long long ans[2];                     // zones 1 and 3: the key
  unsigned long long k0, k1;
  double tt;
};

// the counters of a sweep (bs_bounds_kernel clears them too; a sweep that reuses brackets starts here)
__global__ void bs_clear_kernel(BsSeg* __restrict__ segs, int64_t n_seg) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_seg) return;
  for (int j = 0; j < BS_MAX_Q; ++j) { segs[g].below[j] = 0; segs[g].eq_lo[j] = 0; segs[g].eq_hi[j] = 0; }
  segs[g].neg_zero = 0; segs[g].list_n = 0; segs[g].n_valid = 0; segs[g].overflow = 0;
}

// hull over the ranks that hold sample values of the segment; counters were cleared by bs_bounds_kernel
__global__ void bs_merge_bounds_kernel(const double* __restrict__ all, int world, int64_t n_seg, int n_q,
                                       BsSeg* __restrict__ segs) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_seg) return;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  unsigned zb = 0;
  for (int j = 0; j < n_q; ++j) {
    double lo = INF, hi = -INF; bool any = false;
    for (int w = 0; w < world; ++w) {
      const double* b = all + ((size_t)w * n_seg + g) * BS_BOUNDS_W;
      if (b[2 * BS_MAX_Q] > 0.0) { any = true; lo = b[j] < lo ? b[j] : lo; hi = b[BS_MAX_Q + j] > hi ? b[BS_MAX_Q + j] : hi; }
    }
    if (!any) { lo = -INF; hi = INF; }
    segs[g].lo[j] = lo; segs[g].hi[j] = hi;
    if (lo == 0.0 || hi == 0.0) zb = 1;
  }
  segs[g].zero_bound = zb;
  double gmin = INF, gmax = -INF;
  for (int w = 0; w < world; ++w) {
    const double* b = all + ((size_t)w * n_seg + g) * BS_BOUNDS_W;
    if (b[2 * BS_MAX_Q] > 0.0) { gmin = b[2 * BS_MAX_Q + 1] < gmin ? b[2 * BS_MAX_Q + 1] : gmin; gmax = b[2 * BS_MAX_Q + 2] > gmax ? b[2 * BS_MAX_Q + 2] : gmax; }
  }
  if (!(gmin <= gmax)) { gmin = -INF; gmax = INF; }
  segs[g].gmin = gmin; segs[g].gmax = gmax;
}

// one warp folds this rank's per-series partials of segment g (the order of pmx_group_reduce), then packs
struct BsPackArgs {
  const Part* part; const int32_t* perm; const int64_t* goff; double* sum; int64_t* cnt; double* mn; double* mx;
  unsigned cap; unsigned long long* out;
};
__device__ __forceinline__ void bs_fold_pack_warp(const BsSeg* __restrict__ segs, const BsPackArgs& A, int64_t g) {
  const Part* part = A.part; const int32_t* perm = A.perm; const int64_t* goff = A.goff;
  double* sum = A.sum; int64_t* cnt = A.cnt; double* mn = A.mn; double* mx = A.mx;
  const unsigned cap = A.cap; unsigned long long* out = A.out;
  group_fold_warp(part, perm, goff, g, threadIdx.x & 31, sum, cnt, mn, mx, nullptr);
  if ((threadIdx.x & 31) != 0) return;           // lane 0 wrote the fold's results and reads them back
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  unsigned long long* o = out + g * BS_PACK_W;
  const int64_t c = cnt[g];
  o[0] = (unsigned long long)__double_as_longlong(sum[g]);
  o[1] = (unsigned long long)c;
  o[2] = (unsigned long long)__double_as_longlong(c > 0 ? mn[g] : INF);
  o[3] = (unsigned long long)__double_as_longlong(c > 0 ? mx[g] : -INF);
  for (int j = 0; j < BS_MAX_Q; ++j) {
    o[4 + j] = segs[g].below[j]; o[4 + BS_MAX_Q + j] = segs[g].eq_lo[j]; o[4 + 2 * BS_MAX_Q + j] = segs[g].eq_hi[j];
  }
  o[4 + 3 * BS_MAX_Q] = segs[g].neg_zero;
  o[5 + 3 * BS_MAX_Q] = segs[g].list_n > cap ? 1 : 0;
}

// A digit of this rank's keys strictly inside bracket j: grid (4 n_q, n_seg).  SECOND = false: the first
// 11 bits, one histogram per (segment, quantile).  SECOND = true: the next 11 bits of the keys whose
// first digit is that of rank t's bucket, one histogram per (segment, quantile, t).
template <bool SECOND>
__global__ void __launch_bounds__(1024)
bs_midhist_kernel(const BsSeg* __restrict__ segs, const unsigned long long* __restrict__ lists, unsigned cap, int n_q,
                  const BsPick* __restrict__ picks, unsigned int* __restrict__ hist, const BsPackArgs pack) {
  __shared__ unsigned int sh[(SECOND ? 2 : 1) * BS_BINS];
  const int j = blockIdx.x >> 2, quarter = blockIdx.x & 3, g = blockIdx.y;
  if (!SECOND && j == n_q) {
    // one more block per segment (first histogram only): its first warp folds and packs this rank's
    // partial summary and counters while the others histogram -- no launch and no time of its own
    if (quarter == 0 && threadIdx.x < 32) bs_fold_pack_warp(segs, pack, g);
    return;
  }
  const BsSeg& B = segs[g];
  const double lo = B.lo[j], hi = B.hi[j];
  const BsGeom G = bs_digit_geom(B, j);
  const unsigned n_list = B.list_n > cap ? cap : B.list_n;
  const unsigned long long* keys = lists + (size_t)g * cap;
  const unsigned part_n = (n_list + 3) / 4;
  const unsigned k_a = quarter * part_n < n_list ? quarter * part_n : n_list;
  const unsigned k_b = k_a + part_n < n_list ? k_a + part_n : n_list;
  if (k_a >= k_b) return;
  unsigned b0 = 0xffffffffu, b1 = 0xffffffffu;
  if (SECOND) {
    const BsPick& P = picks[(size_t)g * n_q + j];
    if (P.zone[0] == 2) b0 = P.bin[0];
    if (P.zone[1] == 2) b1 = P.bin[1];
    if (b0 == 0xffffffffu && b1 == 0xffffffffu) return;
  }
  for (int i = threadIdx.x; i < (SECOND ? 2 : 1) * BS_BINS; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (unsigned i0 = k_a; i0 < k_b; i0 += blockDim.x * 8) {
    unsigned long long kk[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { const unsigned i = i0 + u * blockDim.x + threadIdx.x; kk[u] = i < k_b ? keys[i] : 0ULL; }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const unsigned i = i0 + u * blockDim.x + threadIdx.x;
      if (i >= k_b) continue;
      const double v = key_f64(kk[u]);
      if (!(v > lo && v < hi)) continue;
      const uint64_t rel = bs_rel(G, kk[u]);
      const unsigned d1 = (unsigned)(rel >> 53);
      if (!SECOND) { atomicAdd(&sh[d1], 1u); continue; }
      const unsigned d2 = (unsigned)(rel >> 42) & (BS_BINS - 1);
      if (d1 == b0) atomicAdd(&sh[d2], 1u);
      if (d1 == b1) atomicAdd(&sh[BS_BINS + d2], 1u);
    }
  }
  __syncthreads();
  constexpr int NH = SECOND ? 2 : 1;
  unsigned int* dst = hist + ((size_t)g * n_q + j) * NH * BS_BINS;
  for (int i = threadIdx.x; i < NH * BS_BINS; i += blockDim.x) { const unsigned c = sh[i]; if (c) atomicAdd(&dst[i], c); }
}

// CTA (j, g), after the all-reduce of the first histograms: place the two ranks of quantile j (as
// bs_resolve_kernel does, from the combined counters) and pick the first digit of their buckets.
__global__ void __launch_bounds__(1024)
bs_pick1_kernel(BsSeg* __restrict__ segs, const unsigned long long* __restrict__ all, int world, int64_t n_seg,
                const double* __restrict__ q, int n_q, BsPick* __restrict__ picks,
                double* o_sum, int64_t* o_cnt, double* o_min, double* o_max, double* o_mean) {
  __shared__ unsigned int s_warp[32];
  __shared__ BsPick s_p;
  __shared__ unsigned int s_tot;
  __shared__ unsigned long long s_cn[3 * BS_MAX_Q + 2];
  __shared__ long long s_n;
  const int j = blockIdx.x, g = blockIdx.y;
  BsPick* pick = picks + (size_t)g * n_q + j;
  // what every rank packed, in rank order: the counters (one thread each), the summary (one thread),
  // and this thread's two bins of the first histogram -- every CTA of the segment gets the same numbers
  const size_t per_rank = (size_t)n_seg * BS_PACK_W + (size_t)n_seg * n_q * BS_BINS / 2;   // 64-bit words
  if (threadIdx.x < 3 * BS_MAX_Q + 2) {
    unsigned long long c = 0;
    for (int w = 0; w < world; ++w) c += all[(size_t)w * per_rank + (size_t)g * BS_PACK_W + 4 + threadIdx.x];
    s_cn[threadIdx.x] = c;
  } else if (threadIdx.x == 32) {
    Part r = part_identity();
    for (int w = 0; w < world; ++w) {
      const unsigned long long* b = all + (size_t)w * per_rank + (size_t)g * BS_PACK_W;
      r.sum += __longlong_as_double((long long)b[0]); r.cnt += (long long)b[1];
      r.mn = fmin(r.mn, __longlong_as_double((long long)b[2])); r.mx = fmax(r.mx, __longlong_as_double((long long)b[3]));
    }
    s_n = r.cnt;
    if (j == 0) {
      const bool any = r.cnt > 0;
      if (o_sum) o_sum[g] = r.sum;
      if (o_cnt) o_cnt[g] = r.cnt;
      if (o_min) o_min[g] = any ? r.mn : pmx_nan();
      if (o_max) o_max[g] = any ? r.mx : pmx_nan();
      if (o_mean) o_mean[g] = any ? r.sum / (double)r.cnt : pmx_nan();
    }
  }
  unsigned c0 = 0, c1 = 0;
  for (int w = 0; w < world; ++w) {
    const uint2 c = *reinterpret_cast<const uint2*>(reinterpret_cast<const unsigned int*>(all + (size_t)w * per_rank + (size_t)n_seg * BS_PACK_W) +
                                                    ((size_t)g * n_q + j) * BS_BINS + 2 * threadIdx.x);
    c0 += c.x; c1 += c.y;
  }
  __syncthreads();
  const long long n = s_n;
  if (n <= 0) { if (threadIdx.x == 0) { pick->zone[0] = pick->zone[1] = -1; } return; }
  const BsSeg& B = segs[g];
  const double lo = B.lo[j], hi = B.hi[j];
  const uint64_t klo = f64_key(lo), khi = f64_key(hi);
  const unsigned long long c_lt = s_cn[j], c_el = s_cn[BS_MAX_Q + j], c_eh = hi == lo ? 0 : s_cn[2 * BS_MAX_Q + j],
                           c_nz = s_cn[3 * BS_MAX_Q];
  const bool overflow = s_cn[3 * BS_MAX_Q + 1] != 0;
  const unsigned base = bs_scan2(c0, c1, s_warp);
  if (threadIdx.x == blockDim.x - 1) s_tot = base + c0 + c1;      // all keys inside the bracket, over all ranks
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long c_mid = s_tot;
    const double pos = (double)(n - 1) * q[j];
    unsigned long long k0 = (unsigned long long)floor(pos);
    if (k0 > (unsigned long long)n - 1) k0 = (unsigned long long)n - 1;
    const unsigned long long k1 = k0 + 1 < (unsigned long long)n ? k0 + 1 : (unsigned long long)n - 1;
    s_p.k0 = k0; s_p.k1 = k1; s_p.tt = pos - (double)k0;
    for (int t = 0; t < 2; ++t) {
      const unsigned long long k = t ? k1 : k0;
      int zone = 0; unsigned long long r = 0, ans = 0;
      if (k < c_lt) zone = 0;
      else if (k < c_lt + c_el) { zone = 1; r = k - c_lt; ans = lo == 0.0 ? f64_key(r < c_nz ? -0.0 : 0.0) : klo; }
      else if (overflow) zone = 0;
      else if (k < c_lt + c_el + c_mid) { zone = 2; r = k - c_lt - c_el; }
      else if (k < c_lt + c_el + c_mid + c_eh) { zone = 3; r = k - c_lt - c_el - c_mid; ans = hi == 0.0 ? f64_key(r < c_nz ? -0.0 : 0.0) : khi; }
      s_p.zone[t] = zone; s_p.r[t] = r; s_p.ans[t] = ans; s_p.bin[t] = 0xffffffffu;
    }
  }
  __syncthreads();
  const unsigned long long r0 = s_p.r[0], r1 = s_p.r[1];
  const int z0 = s_p.zone[0], z1 = s_p.zone[1];
  __syncthreads();
  for (int t = 0; t < 2; ++t) {
    if ((t ? z1 : z0) != 2) continue;
    const unsigned long long r = t ? r1 : r0;
    if (r >= base && r < (unsigned long long)base + c0) { s_p.bin[t] = 2 * threadIdx.x; s_p.r[t] = r - base; }
    else if (r >= (unsigned long long)base + c0 && r < (unsigned long long)base + c0 + c1) { s_p.bin[t] = 2 * threadIdx.x + 1; s_p.r[t] = r - base - c0; }
  }
  __syncthreads();
  if (threadIdx.x == 0) *pick = s_p;
}

// Four CTAs per (j, g), after the all-reduce of the second histograms: the second digit of either
// bucket, and this rank's keys of the buckets (22 bits now) into its row, which the caller cleared:
// count | ~min, max of rank 0's bucket | ~min, max of rank 1's bucket | keys  (the minimum is kept
// complemented so that zero is neutral for all four).
__global__ void __launch_bounds__(1024)
bs_pick2_rows_kernel(const BsSeg* __restrict__ segs, const unsigned long long* __restrict__ lists, unsigned cap,
                     const unsigned int* __restrict__ hist2, int n_hist, int64_t n_seg, int n_q, BsPick* __restrict__ picks,
                     unsigned long long* __restrict__ rows) {
  __shared__ unsigned int s_warp[32];
  __shared__ unsigned int s_bin[2];
  __shared__ unsigned long long s_r[2], s_mm[4];
  const int j = blockIdx.x >> 2, quarter = blockIdx.x & 3, g = blockIdx.y;
  unsigned long long* row = rows + ((size_t)g * n_q + j) * BS_ROW_W;
  BsPick* pick = picks + (size_t)g * n_q + j;
  const int z0 = pick->zone[0], z1 = pick->zone[1];
  if (z0 != 2 && z1 != 2) return;
  const unsigned bin_a = pick->bin[0], bin_b = pick->bin[1];
  const unsigned long long ra = pick->r[0], rb = pick->r[1];
  if (threadIdx.x == 0) { s_mm[0] = s_mm[1] = s_mm[2] = s_mm[3] = 0; s_bin[0] = s_bin[1] = 0xffffffffu; }
  __syncthreads();
  for (int t = 0; t < 2; ++t) {
    if ((t ? z1 : z0) != 2) continue;                               // uniform
    // the second histograms: already summed (n_hist = 1: an all-reduce did it) or one copy per rank
    unsigned c0 = 0, c1 = 0;
    for (int w = 0; w < n_hist; ++w) {
      const uint2 c = *reinterpret_cast<const uint2*>(hist2 + (size_t)w * n_seg * n_q * 2 * BS_BINS +
                                                      (((size_t)g * n_q + j) * 2 + t) * BS_BINS + 2 * threadIdx.x);
      c0 += c.x; c1 += c.y;
    }
    const unsigned base = bs_scan2(c0, c1, s_warp);
    const unsigned long long r = t ? rb : ra;
    const unsigned b1st = t ? bin_b : bin_a;
    if (r >= base && r < (unsigned long long)base + c0) { s_bin[t] = (b1st << 11) | (2 * threadIdx.x); s_r[t] = r - base; }
    else if (r >= (unsigned long long)base + c0 && r < (unsigned long long)base + c0 + c1) { s_bin[t] = (b1st << 11) | (2 * threadIdx.x + 1); s_r[t] = r - base - c0; }
    __syncthreads();
  }
  const unsigned b0 = s_bin[0], b1 = s_bin[1];
  if (quarter == 0 && threadIdx.x == 0) {
    if (z0 == 2) { pick->bin2[0] = b0; pick->r2[0] = s_r[0]; }
    if (z1 == 2) { pick->bin2[1] = b1; pick->r2[1] = s_r[1]; }
  }
  const BsSeg& B = segs[g];
  const double lo = B.lo[j], hi = B.hi[j];
  const BsGeom G = bs_digit_geom(B, j);
  const unsigned n_list = B.list_n > cap ? cap : B.list_n;
  const unsigned long long* keys = lists + (size_t)g * cap;
  const unsigned part_n = (n_list + 3) / 4;
  const unsigned k_a = quarter * part_n < n_list ? quarter * part_n : n_list;
  const unsigned k_b = k_a + part_n < n_list ? k_a + part_n : n_list;
  for (unsigned i0 = k_a; i0 < k_b; i0 += blockDim.x * 8) {
    unsigned long long kk[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { const unsigned i = i0 + u * blockDim.x + threadIdx.x; kk[u] = i < k_b ? keys[i] : 0ULL; }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const unsigned i = i0 + u * blockDim.x + threadIdx.x;
      if (i >= k_b) continue;
      const double v = key_f64(kk[u]);
      if (!(v > lo && v < hi)) continue;
      const unsigned d = (unsigned)(bs_rel(G, kk[u]) >> 42);
      if (d != b0 && d != b1) continue;
      const unsigned long long at = atomicAdd(&row[0], 1ULL);
      if (at < BS_ROWK) row[5 + at] = kk[u];
      const int t = d == b0 ? 0 : 1;
      if (~kk[u] > *(volatile unsigned long long*)&s_mm[2 * t]) atomicMax(&s_mm[2 * t], ~kk[u]);
      if (kk[u] > *(volatile unsigned long long*)&s_mm[2 * t + 1]) atomicMax(&s_mm[2 * t + 1], kk[u]);
    }
  }
  __syncthreads();
  if (threadIdx.x < 4 && s_mm[threadIdx.x]) atomicMax(&row[1 + threadIdx.x], s_mm[threadIdx.x]);
}

// CTA (j, g), after the all-gather of the rows ([world][n_seg][n_q][ROW]): every rank settles every
// quantile alike.  A bucket whose keys are all equal needs no keys; otherwise the gathered keys are
// ranked directly.  What does not fit (a rank kept more than BS_ROWK keys of distinct values) fails.
__global__ void __launch_bounds__(256)
bs_final_kernel(const BsSeg* __restrict__ segs, const BsPick* __restrict__ picks, const unsigned long long* __restrict__ rows_all,
                int world, int64_t n_seg, int n_q, const int64_t* __restrict__ count,
                int64_t* o_n, int64_t* o_lo, int64_t* o_hi, double* o_xlo, double* o_xhi, double* o_val,
                unsigned int* __restrict__ failed) {
  constexpr unsigned KMAX = 4096;
  __shared__ unsigned long long s_keys[KMAX];
  __shared__ unsigned long long s_ans[2];
  __shared__ int s_ok[2];
  const int j = blockIdx.x, g = blockIdx.y;
  const int64_t out = (int64_t)g * n_q + j;
  const BsPick P = picks[out];
  if (j == 0 && threadIdx.x == 0 && o_n) o_n[g] = count[g];
  auto write = [&](long long klo, long long khi, double a, double b, double v) {
    if (o_lo) o_lo[out] = klo;
    if (o_hi) o_hi[out] = khi;
    if (o_xlo) o_xlo[out] = a;
    if (o_xhi) o_xhi[out] = b;
    if (o_val) o_val[out] = v;
  };
  if (P.zone[0] < 0) { if (threadIdx.x == 0) write(-1, -1, pmx_nan(), pmx_nan(), pmx_nan()); return; }
  bool fail = P.zone[0] == 0 || P.zone[1] == 0;
  const BsSeg& B = segs[g];
  const BsGeom G = bs_digit_geom(B, j);
  if (threadIdx.x < 2) {
    const int t = threadIdx.x;
    s_ok[t] = 1; s_ans[t] = P.ans[t];
    if (P.zone[t] == 2) {
      unsigned long long mn = ~0ULL, mx = 0;
      const int ts = (t == 1 && P.zone[0] == 2 && P.bin2[0] == P.bin2[1]) ? 0 : t;   // one bucket for both ranks: filed under rank 0
      for (int w = 0; w < world; ++w) {
        const unsigned long long* row = rows_all + (((size_t)w * n_seg + g) * n_q + j) * BS_ROW_W;
        mn = ~row[1 + 2 * ts] < mn ? ~row[1 + 2 * ts] : mn; mx = row[2 + 2 * ts] > mx ? row[2 + 2 * ts] : mx;
      }
      if (mn == mx) s_ans[t] = mn; else s_ok[t] = 0;             // all keys of the bucket are one value
    }
  }
  __syncthreads();
  if (!fail && (!s_ok[0] || !s_ok[1])) {
    // the keys themselves: every rank's row, in rank order
    unsigned total = 0; bool fits = true;
    for (int w = 0; w < world; ++w) {
      const unsigned long long c = rows_all[(((size_t)w * n_seg + g) * n_q + j) * BS_ROW_W];
      if (c > BS_ROWK) fits = false; else total += (unsigned)c;
    }
    if (!fits || total > KMAX) fail = true;
    else {
      unsigned off = 0;
      for (int w = 0; w < world; ++w) {
        const unsigned long long* row = rows_all + (((size_t)w * n_seg + g) * n_q + j) * BS_ROW_W;
        const unsigned c = (unsigned)row[0];
        for (unsigned i = threadIdx.x; i < c; i += blockDim.x) s_keys[off + i] = row[5 + i];
        off += c;
      }
      __syncthreads();
      for (unsigned i = threadIdx.x; i < total; i += blockDim.x) {
        const uint64_t ki = s_keys[i];
        const unsigned d = (unsigned)(bs_rel(G, ki) >> 42);
        unsigned less = 0, eq = 0;
        for (unsigned u = 0; u < total; ++u) {
          const uint64_t ku = s_keys[u];
          if ((unsigned)(bs_rel(G, ku) >> 42) == d) { less += ku < ki; eq += ku == ki; }
        }
        for (int t = 0; t < 2; ++t)
          if (P.zone[t] == 2 && !s_ok[t] && P.bin2[t] == d && P.r2[t] >= less && P.r2[t] < (unsigned long long)less + eq) s_ans[t] = ki;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (fail) { atomicAdd(failed, 1u); write((long long)P.k0, (long long)P.k1, pmx_nan(), pmx_nan(), pmx_nan()); return; }
    const double a = key_f64(s_ans[0]), b = key_f64(s_ans[1]);
    const double d = __dsub_rn(b, a);
    double v = __dadd_rn(a, __dmul_rn(d, P.tt));
    if (P.tt >= 0.5) v = __dsub_rn(b, __dmul_rn(d, __dsub_rn(1.0, P.tt)));
    write((long long)P.k0, (long long)P.k1, a, b, v);
  }
}

}  // namespace

// scratch of one selection: per-series partials (slot PMX_SCR_PARTIALS); per-segment brackets and
// counters | local fold outputs | picks | sample | lists (slot PMX_SCR_BRACKET)
struct BsScratch {
  Part* part; BsSeg* segs; double* f_sum; int64_t* f_cnt; double* f_min; double* f_max; BsPick* picks;
  double* samp; unsigned long long* lists; int64_t cap; int64_t P;
};
static int bs_scratch(pmx_ctx* ctx, int64_t n_series, int64_t n_steps, int64_t n_seg, int64_t n_q, BsScratch* S) {
  S->P = (n_steps + BS_STRIDE - 1) / BS_STRIDE;
  S->part = nullptr;
  if (n_series > 0) {
    S->part = (Part*)pmx_scratch(ctx, PMX_SCR_PARTIALS, (size_t)n_series * sizeof(Part));
    if (!S->part) return PMX_E_CUDA;
  }
  int64_t cap = ((int64_t)256 << 20) / 8 / n_seg;
  if (cap > (1 << 22)) cap = 1 << 22;
  if (cap < 4096) cap = 4096;
  S->cap = cap;
  auto up = [](size_t b) { return (b + 255) & ~size_t(255); };
  const size_t seg_bytes = up((size_t)n_seg * sizeof(BsSeg)), fold_bytes = up((size_t)n_seg * 8);
  const size_t pick_bytes = up((size_t)n_seg * BS_MAX_Q * sizeof(BsPick));
  const size_t samp_bytes = up((size_t)(n_series > 0 ? n_series : 1) * (S->P > 0 ? S->P : 1) * 8);
  unsigned char* scr = (unsigned char*)pmx_scratch(ctx, PMX_SCR_BRACKET, seg_bytes + 4 * fold_bytes + pick_bytes + samp_bytes +
                                                                         (size_t)n_seg * cap * 8);
  if (!scr) return PMX_E_CUDA;
  S->segs = (BsSeg*)scr; scr += seg_bytes;
  S->f_sum = (double*)scr; scr += fold_bytes;
  S->f_cnt = (int64_t*)scr; scr += fold_bytes;
  S->f_min = (double*)scr; scr += fold_bytes;
  S->f_max = (double*)scr; scr += fold_bytes;
  S->picks = (BsPick*)scr; scr += pick_bytes;
  S->samp = (double*)scr; scr += samp_bytes;
  S->lists = (unsigned long long*)scr;
  (void)n_q;
  return PMX_OK;
}
// sample + brackets of this rank's rows
static int bs_launch_bounds(const BsScratch& S, const double* d_x, int64_t n_series, int64_t n_steps, int64_t n_seg,
                            const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, cudaStream_t st,
                            const int64_t* d_csr_off = nullptr, double* d_exported = nullptr) {
  const bool data = n_series > 0 && n_steps > 0;
  if (data) {
    const int64_t slots = n_series * S.P;
    if (d_csr_off)          // d_x holds the series' samples back to back
      bs_sample_csr_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, st>>>(d_x, d_csr_off, n_steps, n_series, S.P, S.samp);
    else
      bs_sample_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, st>>>(d_x, n_steps, n_series, S.P, S.samp);
    PMX_LAUNCH_CHECK();
  }
  const size_t smem = (size_t)BS_KCAP * 8 + (size_t)2 * n_q * BS_BINS * sizeof(unsigned int);
  PMX_CUDA(cudaFuncSetAttribute(bs_bounds_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(BS_KCAP * 8 + 2 * BS_MAX_Q * BS_BINS * sizeof(unsigned int))));
  bs_bounds_kernel<<<(unsigned)(n_seg * BS_CL), 1024, smem, st>>>(S.samp, data ? S.P : 0, d_perm, d_goff, d_q, (int)n_q, S.segs,
                                                                  d_exported);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}
static int bs_launch_sweep(const BsScratch& S, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg,
                           int64_t n_q, cudaStream_t st) {
  if (n_series <= 0 || n_steps <= 0) return PMX_OK;
  const unsigned cap = (unsigned)S.cap;
  switch (n_q) {
    case 1: bs_sweep_kernel<1><<<(unsigned)n_series, PMX_ROW_THREADS, 0, st>>>(d_x, n_steps, d_seg, S.part, S.segs, S.lists, cap); break;
    case 2: bs_sweep_kernel<2><<<(unsigned)n_series, PMX_ROW_THREADS, 0, st>>>(d_x, n_steps, d_seg, S.part, S.segs, S.lists, cap); break;
    case 3: bs_sweep_kernel<3><<<(unsigned)n_series, PMX_ROW_THREADS, 0, st>>>(d_x, n_steps, d_seg, S.part, S.segs, S.lists, cap); break;
    default: bs_sweep_kernel<4><<<(unsigned)n_series, PMX_ROW_THREADS, 0, st>>>(d_x, n_steps, d_seg, S.part, S.segs, S.lists, cap); break;
  }
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

static int bs_summary(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps,
                      const int32_t* d_seg, int64_t n_seg, const int32_t* d_perm, const int64_t* d_goff,
                      const double* d_q, int64_t n_q, double* d_sum, int64_t* d_count, double* d_min,
                      double* d_max, double* d_mean, int64_t* d_n, int64_t* d_lo, int64_t* d_hi,
                      double* d_x_lo, double* d_x_hi, double* d_value, int32_t* d_failed,
                      pmx_stream stream, bool brackets_ready) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0 && n_seg >= 0 && n_seg <= 65535, "bad sizes (n_seg <= 65535)");
  if (n_q < 1 || n_q > BS_MAX_Q) return pmx_fail(PMX_E_LIMIT, "pmx_summary_quantiles: n_q must be in [1, %d]; use pmx_quantile_select", BS_MAX_Q);
  PMX_REQUIRE(d_failed, "d_failed is NULL");
  cudaStream_t st = pmx_st(stream);
  PMX_CUDA(cudaMemsetAsync(d_failed, 0, sizeof(int32_t), st));
  if (n_seg == 0) return PMX_OK;
  PMX_REQUIRE(d_goff && d_q && d_count, "NULL buffer");
  PMX_REQUIRE(n_series == 0 || (d_x && d_seg && d_perm), "NULL buffer");
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  if (!brackets_ready) {
    if (int rc = bs_launch_bounds(S, d_x, n_series, n_steps, n_seg, d_perm, d_goff, d_q, n_q, st)) return rc;
  } else {
    bs_clear_kernel<<<(unsigned)((n_seg + 127) / 128), 128, 0, st>>>(S.segs, n_seg);
    PMX_LAUNCH_CHECK();
  }
  if (int rc = bs_launch_sweep(S, d_x, n_series, n_steps, d_seg, n_q, st)) return rc;
  Part* part = S.part; BsSeg* segs = S.segs; unsigned long long* lists = S.lists; const int64_t cap = S.cap;
  dim3 grid((unsigned)(n_q + 1) * BS_RC, (unsigned)n_seg);       // n_q selecting clusters + one that folds the summary
  bs_resolve_kernel<<<grid, 1024, 0, st>>>(segs, lists, (unsigned)cap, d_q, (int)n_q, d_n, d_lo, d_hi, d_x_lo, d_x_hi,
                                           d_value, (unsigned int*)d_failed, part, d_perm, d_goff, d_sum, d_count, d_min,
                                           d_max, d_mean);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

// ---- several ranks: the same selection in five calls with four collectives between them ----------
extern "C" {

int pmx_bsel_widths(int* bounds_w, int* pack_w, int* row_w) {
  if (bounds_w) *bounds_w = BS_BOUNDS_W;
  if (pack_w) *pack_w = BS_PACK_W;
  if (row_w) *row_w = (int)BS_ROW_W;
  return PMX_OK;
}

#define BS_COMMON_CHECKS()                                                                                        \
  PMX_REQUIRE(ctx, "ctx is NULL");                                                                                \
  PMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "n_seg must be in [0, 65535]");                                        \
  if (n_q < 1 || n_q > BS_MAX_Q) return pmx_fail(PMX_E_LIMIT, "%s: n_q must be in [1, %d]", __func__, BS_MAX_Q);   \
  if (n_seg == 0) return PMX_OK

int pmx_bsel_begin(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg, int64_t n_seg,
                   const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, double* d_bounds,
                   pmx_stream stream) {
  BS_COMMON_CHECKS();
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0, "bad sizes");
  PMX_REQUIRE(d_goff && d_q && d_bounds, "NULL buffer");
  PMX_REQUIRE(n_series == 0 || (d_x && d_seg && d_perm), "NULL buffer");
  cudaStream_t st = pmx_st(stream);
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  return bs_launch_bounds(S, d_x, n_series, n_steps, n_seg, d_perm, d_goff, d_q, n_q, st, nullptr, d_bounds);
}

int pmx_bsel_begin_csr(pmx_ctx* ctx, const double* d_val, const int64_t* d_off, int64_t n_series, int64_t n_steps, int64_t n_seg,
                       const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, double* d_bounds,
                       pmx_stream stream) {
  BS_COMMON_CHECKS();
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0, "bad sizes");
  PMX_REQUIRE(d_goff && d_q && d_bounds && (n_series == 0 || (d_val && d_off && d_perm)), "NULL buffer");
  cudaStream_t st = pmx_st(stream);
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  return bs_launch_bounds(S, d_val, n_series, n_steps, n_seg, d_perm, d_goff, d_q, n_q, st, d_off, d_bounds);
}

int pmx_bsel_sweep(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg, int64_t n_seg,
                   const int32_t* d_perm, const int64_t* d_goff, int64_t n_q, const double* d_bounds_all, int n_ranks,
                   uint64_t* d_packed, pmx_stream stream) {
  BS_COMMON_CHECKS();
  PMX_REQUIRE(n_ranks >= 1 && d_bounds_all && d_packed && d_goff, "NULL buffer or bad rank count");
  cudaStream_t st = pmx_st(stream);
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  bs_merge_bounds_kernel<<<(unsigned)((n_seg + 127) / 128), 128, 0, st>>>(d_bounds_all, n_ranks, n_seg, (int)n_q, S.segs);
  PMX_LAUNCH_CHECK();
  if (int rc = bs_launch_sweep(S, d_x, n_series, n_steps, d_seg, n_q, st)) return rc;
  // first digit of this rank's keys inside every bracket, behind the packed words (it needs only the
  // brackets and the sample range, so it travels with the counters)
  unsigned int* h1 = (unsigned int*)(d_packed + (size_t)n_seg * BS_PACK_W);
  PMX_CUDA(cudaMemsetAsync(h1, 0, (size_t)n_seg * n_q * BS_BINS * sizeof(unsigned int), st));
  dim3 grid((unsigned)n_q * 4 + 1, (unsigned)n_seg);            // + one block per segment that folds and packs
  const BsPackArgs pack{S.part, d_perm, d_goff, S.f_sum, S.f_cnt, S.f_min, S.f_max, (unsigned)S.cap, (unsigned long long*)d_packed};
  bs_midhist_kernel<false><<<grid, 1024, 0, st>>>(S.segs, S.lists, (unsigned)S.cap, (int)n_q, nullptr, h1, pack);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_bsel_hist(pmx_ctx* ctx, int64_t n_series, int64_t n_steps, int64_t n_seg, const double* d_q, int64_t n_q,
                  const uint64_t* d_packed_all, int n_ranks, double* d_sum, int64_t* d_count, double* d_min, double* d_max,
                  double* d_mean, uint32_t* d_hist2, pmx_stream stream) {
  BS_COMMON_CHECKS();
  PMX_REQUIRE(n_ranks >= 1 && d_q && d_packed_all && d_count && d_hist2, "NULL buffer or bad rank count");
  cudaStream_t st = pmx_st(stream);
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  dim3 grid((unsigned)n_q, (unsigned)n_seg);
  bs_pick1_kernel<<<grid, 1024, 0, st>>>(S.segs, (const unsigned long long*)d_packed_all, n_ranks, n_seg, d_q, (int)n_q, S.picks,
                                         d_sum, d_count, d_min, d_max, d_mean);
  PMX_LAUNCH_CHECK();
  PMX_CUDA(cudaMemsetAsync(d_hist2, 0, (size_t)n_seg * n_q * 2 * BS_BINS * sizeof(uint32_t), st));
  dim3 grid4((unsigned)n_q * 4, (unsigned)n_seg);
  bs_midhist_kernel<true><<<grid4, 1024, 0, st>>>(S.segs, S.lists, (unsigned)S.cap, (int)n_q, S.picks, d_hist2, BsPackArgs{});
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_bsel_rows(pmx_ctx* ctx, int64_t n_series, int64_t n_steps, int64_t n_seg, int64_t n_q, const uint32_t* d_hist2,
                  int n_hist, uint64_t* d_rows, pmx_stream stream) {
  BS_COMMON_CHECKS();
  PMX_REQUIRE(d_hist2 && d_rows && n_hist >= 1, "NULL buffer or bad histogram count");
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  PMX_CUDA(cudaMemsetAsync(d_rows, 0, (size_t)n_seg * n_q * BS_ROW_W * 8, pmx_st(stream)));
  dim3 grid((unsigned)n_q * 4, (unsigned)n_seg);
  bs_pick2_rows_kernel<<<grid, 1024, 0, pmx_st(stream)>>>(S.segs, S.lists, (unsigned)S.cap, d_hist2, n_hist, n_seg, (int)n_q, S.picks,
                                                         (unsigned long long*)d_rows);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_bsel_finish(pmx_ctx* ctx, int64_t n_series, int64_t n_steps, int64_t n_seg, int64_t n_q, const uint64_t* d_rows_all,
                    int n_ranks, const int64_t* d_count, int64_t* d_n, int64_t* d_lo, int64_t* d_hi, double* d_x_lo,
                    double* d_x_hi, double* d_value, int32_t* d_failed, pmx_stream stream) {
  PMX_REQUIRE(d_failed, "d_failed is NULL");
  cudaStream_t st = pmx_st(stream);
  PMX_CUDA(cudaMemsetAsync(d_failed, 0, sizeof(int32_t), st));
  BS_COMMON_CHECKS();
  PMX_REQUIRE(n_ranks >= 1 && d_rows_all && d_count, "NULL buffer or bad rank count");
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  dim3 grid((unsigned)n_q, (unsigned)n_seg);
  bs_final_kernel<<<grid, 256, 0, st>>>(S.segs, S.picks, (const unsigned long long*)d_rows_all, n_ranks, n_seg, (int)n_q, d_count,
                                        d_n, d_lo, d_hi, d_x_lo, d_x_hi, d_value, (unsigned int*)d_failed);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"

extern "C" {

int pmx_summary_quantiles(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg, int64_t n_seg,
                          const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, double* d_sum,
                          int64_t* d_count, double* d_min, double* d_max, double* d_mean, int64_t* d_n, int64_t* d_lo,
                          int64_t* d_hi, double* d_x_lo, double* d_x_hi, double* d_value, int32_t* d_failed, pmx_stream stream) {
  return bs_summary(ctx, d_x, n_series, n_steps, d_seg, n_seg, d_perm, d_goff, d_q, n_q, d_sum, d_count, d_min, d_max, d_mean, d_n,
                    d_lo, d_hi, d_x_lo, d_x_hi, d_value, d_failed, stream, false);
}

int pmx_summary_brackets_csr(pmx_ctx* ctx, const double* d_val, const int64_t* d_off, int64_t n_series, int64_t n_steps,
                             int64_t n_seg, const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q,
                             pmx_stream stream) {
  BS_COMMON_CHECKS();
  PMX_REQUIRE(n_series >= 0 && n_steps >= 0, "bad sizes");
  PMX_REQUIRE(d_goff && d_q && (n_series == 0 || (d_val && d_off && d_perm)), "NULL buffer");
  BsScratch S;
  if (int rc = bs_scratch(ctx, n_series, n_steps, n_seg, n_q, &S)) return rc;
  return bs_launch_bounds(S, d_val, n_series, n_steps, n_seg, d_perm, d_goff, d_q, n_q, pmx_st(stream), d_off);
}

int pmx_summary_sweep(pmx_ctx* ctx, const double* d_x, int64_t n_series, int64_t n_steps, const int32_t* d_seg, int64_t n_seg,
                      const int32_t* d_perm, const int64_t* d_goff, const double* d_q, int64_t n_q, double* d_sum,
                      int64_t* d_count, double* d_min, double* d_max, double* d_mean, int64_t* d_n, int64_t* d_lo,
                      int64_t* d_hi, double* d_x_lo, double* d_x_hi, double* d_value, int32_t* d_failed, pmx_stream stream) {
  return bs_summary(ctx, d_x, n_series, n_steps, d_seg, n_seg, d_perm, d_goff, d_q, n_q, d_sum, d_count, d_min, d_max, d_mean, d_n,
                    d_lo, d_hi, d_x_lo, d_x_hi, d_value, d_failed, stream, true);
}

}  // extern "C"
