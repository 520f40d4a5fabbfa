// ragged_common.cuh -- parameter block shared by ragged.cu (direct kernel) and ragged_fast.cu.
#pragma once
#include "pmx_common.cuh"

namespace pmxr {

struct RParams {
  const int64_t* ts; const double* val; const int64_t* off;
  int64_t n_series; int64_t n_samples; int64_t window_ms; int32_t min_periods; int32_t rate_mode;
  double* o_min; double* o_max; double* o_mean; int32_t* o_cnt;
  double* o_inc; double* o_rate; int32_t* o_nres;
};

// PMX_OK after launching, 1 when the request is outside what the fast kernel covers
int launch_ragged_fast(const pmx_ctx* ctx, const RParams& p, cudaStream_t st);

}  // namespace pmxr
