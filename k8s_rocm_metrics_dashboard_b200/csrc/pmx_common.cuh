// pmx_common.cuh -- shared by every translation unit of libpmx.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/pmx.h"

struct pmx_ctx {
  int device;
  int sm_count;
  // grow-only device scratch, one block per purpose so that they can live side by side:
  // 0 = per-series partials of pmx_group_reduce, 1 = state + histograms of pmx_quantile_select,
  // 2 = candidate keys kept between the passes of a quantile selection, 3 = per-chunk summaries of
  // pmx_tokenize_device, 4 = brackets, sample and lists of pmx_summary_quantiles
  void* scratch[5];
  size_t scratch_bytes[5];
};
enum { PMX_SCR_PARTIALS = 0, PMX_SCR_QSELECT = 1, PMX_SCR_QCAND = 2, PMX_SCR_TOKENIZE = 3, PMX_SCR_BRACKET = 4, PMX_SCR_SLOTS = 5 };

// error plumbing (ctx.cu)
int pmx_fail(int code, const char* fmt, ...);
int pmx_cuda_fail(cudaError_t e, const char* what);
// returns NULL (and sets the error) on failure
void* pmx_scratch(pmx_ctx* ctx, int slot, size_t bytes);

#define PMX_CUDA(call)                                                   \
  do {                                                                   \
    cudaError_t e_ = (call);                                             \
    if (e_ != cudaSuccess) return pmx_cuda_fail(e_, #call);              \
  } while (0)

#define PMX_REQUIRE(cond, msg)                                           \
  do {                                                                   \
    if (!(cond)) return pmx_fail(PMX_E_INVALID, "%s: %s", __func__, msg); \
  } while (0)

#define PMX_LAUNCH_CHECK() PMX_CUDA(cudaGetLastError())

static inline cudaStream_t pmx_st(pmx_stream s) { return (cudaStream_t)s; }

#define PMX_CANON_NAN_BITS 0x7FF8000000000000ULL

#if defined(__CUDACC__)
__device__ __forceinline__ double pmx_nan() { return __longlong_as_double(0x7FF8000000000000LL); }
__device__ __forceinline__ bool pmx_isnan(double v) { return v != v; }
__device__ __forceinline__ bool pmx_isfinite(double v) {
  return (__double_as_longlong(v) & 0x7FF0000000000000LL) != 0x7FF0000000000000LL;
}
// Rate mode 1: Prometheus' extrapolatedRate for counters (promql/functions.go).  `inc` is the
// reset-corrected increase first..last, `fv` the first sample, `cnt` >= 2 samples, `sampled` the
// seconds first..last, `dstart`/`dend` the seconds from the window start to the first sample and
// from the last sample to the window end, `range_s` the window length.  The increase is stretched
// to the window edges when they are closer than 1.1 mean sample spacings (else by half a
// spacing), never further back than the counter's zero crossing; rate = that / range.
__device__ __forceinline__ void pmx_extrapolate(double inc, double fv, int cnt, double sampled, double dstart,
                                                double dend, double range_s, double* inc_out, double* rate_out) {
  const double avg = sampled / (double)(cnt - 1);
  if (inc > 0.0 && fv >= 0.0) {
    const double dzero = sampled * (fv / inc);
    if (dzero < dstart) dstart = dzero;
  }
  const double thr = avg * 1.1;
  double ext = sampled;
  if (dstart >= thr) dstart = avg / 2.0;
  ext += dstart;
  if (dend >= thr) dend = avg / 2.0;
  ext += dend;
  const double factor = ext / sampled;
  *inc_out = inc * factor;
  *rate_out = (inc * factor) / range_s;
}
#endif
