// parse.cu -- K1: value / timestamp text decode on the GPU (SURVEY X2, app.py:188).
//
// One lane per token; a warp covers 32 consecutive samples of a series, whose tokens sit
// ~24 bytes apart in the response body, so the warp's byte reads fall into a handful of
// 128-byte lines that stay in L1 while the lanes scan them.  The arithmetic is
// strtod_core.h (Eisel-Lemire + exact big-integer fallback, all on device: no CPU path).
#include "pmx_common.cuh"
#define PMX_POW5_QUALIFIER __device__ const
#include "pow5_table.h"
#include "strtod_core.h"

namespace {

__global__ void init_first_bad(long long* first_bad) { *first_bad = 0x7fffffffffffffffLL; }

// token [off, end) where end is the first `term` byte at or after off
__device__ __forceinline__ int64_t token_len(const uint8_t* __restrict__ text, uint64_t text_len,
                                             uint64_t off, uint8_t term) {
  uint64_t i = off;
  while (i < text_len && __ldg(text + i) != term) ++i;
  return (int64_t)(i - off);
}

__global__ void __launch_bounds__(256)
parse_f64_kernel(const uint8_t* __restrict__ text, uint64_t text_len,
                 const uint32_t* __restrict__ off, int64_t n, double* __restrict__ out,
                 long long* first_bad) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t o = off[i];
    uint64_t bits = PMX_CANON_NAN_BITS;
    int st = pmx::PARSE_BAD;
    if (o < text_len) {
      const int64_t len = token_len(text, text_len, o, (uint8_t)'"');
      st = pmx::parse_f64(text + o, len, pmx_pow5_128, &bits);
    }
    out[i] = __longlong_as_double((long long)bits);
    if (st != pmx::PARSE_OK) atomicMin(first_bad, (long long)i);
  }
}

__global__ void __launch_bounds__(256)
parse_ts_kernel(const uint8_t* __restrict__ text, uint64_t text_len,
                const uint32_t* __restrict__ off, int64_t n, int64_t* __restrict__ out,
                long long* first_bad) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t o = off[i];
    int64_t v = 0;
    int st = pmx::PARSE_BAD;
    if (o < text_len) {
      const int64_t len = token_len(text, text_len, o, (uint8_t)',');
      st = pmx::parse_ts_ms(text + o, len, &v);
    }
    out[i] = v;
    if (st != pmx::PARSE_OK) atomicMin(first_bad, (long long)i);
  }
}

int grid_for(const pmx_ctx* ctx, int64_t n, int block) {
  int64_t want = (n + block - 1) / block;
  int64_t cap = (int64_t)ctx->sm_count * 16;
  return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

}  // namespace

extern "C" {

int pmx_parse_f64(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len, const uint32_t* d_val_off,
                  int64_t n, double* d_out, int64_t* d_first_bad, pmx_stream stream) {
  PMX_REQUIRE(ctx && d_first_bad, "ctx and d_first_bad are required");
  PMX_REQUIRE(n >= 0, "n must be >= 0");
  init_first_bad<<<1, 1, 0, pmx_st(stream)>>>((long long*)d_first_bad);
  PMX_LAUNCH_CHECK();
  if (n == 0) return PMX_OK;
  PMX_REQUIRE(d_text && d_val_off && d_out, "NULL buffer");
  parse_f64_kernel<<<grid_for(ctx, n, 256), 256, 0, pmx_st(stream)>>>(
      d_text, text_len, d_val_off, n, d_out, (long long*)d_first_bad);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_parse_ts_ms(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len, const uint32_t* d_ts_off,
                    int64_t n, int64_t* d_out, int64_t* d_first_bad, pmx_stream stream) {
  PMX_REQUIRE(ctx && d_first_bad, "ctx and d_first_bad are required");
  PMX_REQUIRE(n >= 0, "n must be >= 0");
  init_first_bad<<<1, 1, 0, pmx_st(stream)>>>((long long*)d_first_bad);
  PMX_LAUNCH_CHECK();
  if (n == 0) return PMX_OK;
  PMX_REQUIRE(d_text && d_ts_off && d_out, "NULL buffer");
  parse_ts_kernel<<<grid_for(ctx, n, 256), 256, 0, pmx_st(stream)>>>(
      d_text, text_len, d_ts_off, n, d_out, (long long*)d_first_bad);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
