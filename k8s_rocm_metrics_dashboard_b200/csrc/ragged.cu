// ragged.cu -- rolling stats and rate evaluated at every sample of ragged series (config 5).
//
// One warp per series; a lane takes every 32nd sample, finds its window start by binary search
// on the series' own timestamps and folds the window directly (fresh sums, exact min/max).
// Window (t_i - W, t_i]; NaN (and therefore stale-marker) samples are skipped.
#include <math.h>
#include <stdlib.h>
#include "ragged_common.cuh"
using pmxr::RParams;

namespace {

__global__ void __launch_bounds__(256)
ragged_window_kernel(const RParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int need = p.min_periods > 1 ? p.min_periods : 1;
  const bool do_roll = p.o_min || p.o_max || p.o_mean || p.o_cnt;
  const bool do_rate = p.o_inc || p.o_rate || p.o_nres;
  for (int64_t s = warp0; s < p.n_series; s += nwarps) {
    const int64_t a = p.off[s], b = p.off[s + 1];
    const int64_t* t = p.ts + a;
    const double* v = p.val + a;
    for (int64_t i = lane; i < b - a; i += 32) {
      const int64_t ti = t[i];
      const int64_t edge = ti - p.window_ms;           // first sample with ts > edge
      int64_t lo = 0, hi = i;
      while (lo < hi) { const int64_t m = (lo + hi) >> 1; if (t[m] <= edge) lo = m + 1; else hi = m; }
      double ws = 0.0, wmn = INF, wmx = -INF; int wcf = 0, wcn = 0;
      int cnt = 0, nres = 0; double fv = 0.0, lv = 0.0, rs = 0.0; int64_t ft = 0, lt = 0;
      for (int64_t k = lo; k <= i; ++k) {
        const double u = v[k];
        if (u != u) continue;
        ++wcn;
        if (pmx_isfinite(u)) { ws += u; wmn = fmin(wmn, u); wmx = fmax(wmx, u); ++wcf; }
        if (cnt > 0) { const bool reset = u < lv; rs += reset ? u : u - lv; nres += reset ? 1 : 0; }   // sum of deltas
        else { fv = u; ft = t[k]; }
        lv = u; lt = t[k]; ++cnt;
      }
      const int64_t o = a + i;
      if (do_roll) {
        const bool ok = wcf >= need;
        if (p.o_min) p.o_min[o] = ok ? wmn : pmx_nan();
        if (p.o_max) p.o_max[o] = ok ? wmx : pmx_nan();
        if (p.o_mean) p.o_mean[o] = ok ? ws / (double)wcf : pmx_nan();
        if (p.o_cnt) p.o_cnt[o] = wcn;
      }
      if (do_rate) {
        double inc = pmx_nan(), rate = pmx_nan();
        if (cnt >= 2) {
          inc = rs;
          const double dt = (double)(lt - ft) / 1000.0;
          rate = dt != 0.0 ? inc / dt : pmx_nan();
          if (p.rate_mode == 1 && dt != 0.0)                 // window (ti - W, ti]
            pmx_extrapolate(inc, fv, cnt, dt, (double)(ft - edge) / 1000.0, (double)(ti - lt) / 1000.0,
                            (double)p.window_ms / 1000.0, &inc, &rate);
        }
        if (p.o_inc) p.o_inc[o] = inc;
        if (p.o_rate) p.o_rate[o] = rate;
        if (p.o_nres) p.o_nres[o] = cnt >= 2 ? nres : 0;
      }
    }
  }
}

int launch(pmx_ctx* ctx, const RParams& p, cudaStream_t st) {
  if (p.n_series == 0) return PMX_OK;
  {
    const int rc = pmxr::launch_ragged_fast(ctx, p, st);   // the tiled kernel when it applies
    if (rc != 1) return rc;
  }
  int64_t blocks = (p.n_series + 7) / 8;                 // 8 warps per CTA
  const int64_t cap = (int64_t)ctx->sm_count * 32;
  if (blocks > cap) blocks = cap;
  ragged_window_kernel<<<(unsigned)blocks, 256, 0, st>>>(p);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // namespace

extern "C" {

int pmx_rolling_ragged(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off,
                       int64_t n_series, int64_t n_samples, int64_t window_ms, int64_t min_periods,
                       double* d_min, double* d_max, double* d_mean, int32_t* d_count, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(n_series >= 0 && n_samples >= 0 && window_ms > 0, "bad sizes");
  PMX_REQUIRE(min_periods >= 0 && min_periods <= 0x7fffffff, "bad min_periods");
  if (n_series == 0 || n_samples == 0) return PMX_OK;
  PMX_REQUIRE(d_ts && d_val && d_off && (d_min || d_max || d_mean || d_count), "NULL buffer");
  RParams p{d_ts, d_val, d_off, n_series, n_samples, window_ms, (int32_t)min_periods, 0, d_min, d_max, d_mean, d_count,
            nullptr, nullptr, nullptr};
  return launch(ctx, p, pmx_st(stream));
}

int pmx_rate_ragged(pmx_ctx* ctx, const int64_t* d_ts, const double* d_val, const int64_t* d_off,
                    int64_t n_series, int64_t n_samples, int64_t window_ms, int mode, double* d_increase,
                    double* d_rate, int32_t* d_n_resets, pmx_stream stream) {
  PMX_REQUIRE(ctx, "ctx is NULL");
  PMX_REQUIRE(mode == 0 || mode == 1, "mode must be 0 (first..last) or 1 (extrapolated)");
  PMX_REQUIRE(n_series >= 0 && n_samples >= 0 && window_ms > 0, "bad sizes");
  if (n_series == 0 || n_samples == 0) return PMX_OK;
  PMX_REQUIRE(d_ts && d_val && d_off && (d_increase || d_rate || d_n_resets), "NULL buffer");
  RParams p{d_ts, d_val, d_off, n_series, n_samples, window_ms, 1, mode, nullptr, nullptr, nullptr, nullptr,
            d_increase, d_rate, d_n_resets};
  return launch(ctx, p, pmx_st(stream));
}

}  // extern "C"
