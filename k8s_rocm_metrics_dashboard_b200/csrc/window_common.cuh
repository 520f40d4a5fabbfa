// window_common.cuh -- parameter block shared by the window kernels (window.cu, window_fast.cu).
#pragma once
#include "pmx_common.cuh"

namespace pmxw {

constexpr int E = 8;                     // grid points per thread

struct WinParams {
  const int64_t* ts; const double* val; const int64_t* off;   // ragged input (ALIGN)
  const double* x;                                            // dense [S,T] input (!ALIGN)
  int64_t n_series, n_steps;
  int64_t start_ms, step_ms, lookback_ms;
  int32_t w, min_periods;
  int32_t rate_mode;                                          // 0 = first..last, 1 = extrapolated (general kernel)
  int32_t chunks; int64_t chunk_len;
  double* o_aligned; int32_t* o_idx;
  double* o_min; double* o_max; double* o_mean; int32_t* o_cnt;
  double* o_inc; double* o_rate; int32_t* o_nres;
};


// window_fast.cu: returns PMX_OK after launching, or 1 when the request is outside what the fast
// kernel covers (the caller then uses the general kernel in window.cu).
template <bool ALIGN> int launch_window_fast(const pmx_ctx* ctx, WinParams p, cudaStream_t st);
int window_block_threads();

}  // namespace pmxw
