// parse.cu -- K1: value / timestamp text decode on the GPU (SURVEY X2, app.py:188).
//
// One thread per sample decodes both of its tokens; a warp covers 32 consecutive samples of a
// series, whose tokens sit ~27 bytes apart in the response body, so a warp's 16-byte loads fall
// into a handful of 128-byte lines.  The common shapes are decoded from registers
// (strtod_fast.h: byte-parallel scan, eight digits per three multiplications, Clinger / Eisel-
// Lemire); the rest runs the byte-wise state machine of strtod_core.h (Eisel-Lemire + exact
// big-integer fallback).  All on the device: no CPU path.
#include "pmx_common.cuh"
#define PMX_POW5_QUALIFIER __device__ const
#include "pow5_table.h"
#include "strtod_core.h"
#include "strtod_fast.h"

namespace {

// token [off, end) where end is the first `term` byte at or after off (byte-wise path only)
__device__ __noinline__ int64_t token_len(const uint8_t* __restrict__ text, uint64_t text_len,
                                          uint64_t off, uint8_t term) {
  uint64_t i = off;
  while (i < text_len && __ldg(text + i) != term) ++i;
  return (int64_t)(i - off);
}

// the 32 bytes from text + off as four 64-bit words: three aligned 16-byte loads, shifted down.
// Needs off + 48 <= text_len rounded down to 16 (true for every token but the last few of a body).
__device__ __forceinline__ void load_window(const uint8_t* __restrict__ text, uint64_t off, uint64_t (&W)[4]) {
  const uint4* p = reinterpret_cast<const uint4*>(text + (off & ~(uint64_t)15));
  const uint4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  const uint64_t X[6] = {(uint64_t)a.x | ((uint64_t)a.y << 32), (uint64_t)a.z | ((uint64_t)a.w << 32),
                         (uint64_t)b.x | ((uint64_t)b.y << 32), (uint64_t)b.z | ((uint64_t)b.w << 32),
                         (uint64_t)c.x | ((uint64_t)c.y << 32), (uint64_t)c.z | ((uint64_t)c.w << 32)};
  const int s = (int)(off & 15), hi8 = s >> 3, bits = (s & 7) * 8;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint64_t lo = hi8 ? X[k + 1] : X[k], hi = hi8 ? X[k + 2] : X[k + 1];
    W[k] = bits ? (lo >> bits) | (hi << (64 - bits)) : lo;
  }
}

__device__ __noinline__ int slow_f64(const uint8_t* __restrict__ text, uint64_t text_len, uint64_t o, uint64_t* bits) {
  const int64_t len = token_len(text, text_len, o, (uint8_t)'"');
  return pmx::parse_f64(text + o, len, pmx_pow5_128, bits);
}
__device__ __noinline__ int slow_ts(const uint8_t* __restrict__ text, uint64_t text_len, uint64_t o, int64_t* v) {
  const int64_t len = token_len(text, text_len, o, (uint8_t)',');
  return pmx::parse_ts_ms(text + o, len, v);
}

// One thread per sample, both tokens.  The common shapes ([-]digits[.digits], at most 24 bytes with
// the terminator) are decoded from a register window (strtod_fast.h); everything else takes the
// byte-wise state machine of strtod_core.h.  Both offsets are fetched before either token is decoded
// (the second token's text then sits in the lines the first one brought in), and 40 registers keep
// 48 warps per SM in flight over the two dependent loads: measured 26.3 -> 28.1 -> 30.0 G samples/s.
// bad_val / bad_ts: smallest index of a malformed value / timestamp token.
__global__ void __launch_bounds__(256, 6)                     // the byte-wise path spills, it is cold
parse_samples_kernel(const uint8_t* __restrict__ text, uint64_t text_len, const uint32_t* __restrict__ ts_off,
                     const uint32_t* __restrict__ val_off, int64_t n, int64_t* __restrict__ out_ts,
                     double* __restrict__ out_val, unsigned long long* bad_val, unsigned long long* bad_ts) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bool vec = (reinterpret_cast<uintptr_t>(text) & 15) == 0;
  const uint64_t safe = text_len >= 48 ? ((text_len - 48) & ~(uint64_t)15) + 16 : 0;   // windows may start below this
  const uint32_t off_val = out_val ? __ldg(val_off + i) : 0, off_ts = out_ts ? __ldg(ts_off + i) : 0;
  if (out_val) {
    const uint64_t o = off_val;
    uint64_t bits = PMX_CANON_NAN_BITS;
    int st = pmx::PARSE_BAD;
    if (o < text_len) {
      int fast = pmx::FAST_FALLBACK;
      if (vec && o < safe) {
        uint64_t W[4];
        load_window(text, o, W);
        fast = pmx::fast_f64_scan(W, (uint8_t)'"', pmx_pow5_128, &bits);
      }
      st = fast == pmx::FAST_OK ? pmx::PARSE_OK : slow_f64(text, text_len, o, &bits);
    }
    out_val[i] = __longlong_as_double((long long)bits);
    if (st != pmx::PARSE_OK) atomicMin(bad_val, (unsigned long long)i);
  }
  if (out_ts) {
    const uint64_t o = off_ts;
    int64_t v = 0;
    int st = pmx::PARSE_BAD;
    if (o < text_len) {
      int fast = pmx::FAST_FALLBACK;
      if (vec && o < safe) {
        uint64_t W[4];
        load_window(text, o, W);
        fast = pmx::fast_ts_ms_scan(W, (uint8_t)',', &v);
      }
      st = fast == pmx::FAST_OK ? pmx::PARSE_OK : slow_ts(text, text_len, o, &v);
    }
    out_ts[i] = v;
    if (st != pmx::PARSE_OK) atomicMin(bad_ts, (unsigned long long)i);
  }
}

}  // namespace

extern "C" {

// 0x7f in every byte: a number no index reaches; the kernel lowers it to the first malformed token
static int reset_first_bad(int64_t* d, cudaStream_t st) {
  PMX_CUDA(cudaMemsetAsync(d, 0x7f, sizeof(int64_t), st));
  return PMX_OK;
}

int pmx_parse_samples(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len, const uint32_t* d_ts_off,
                      const uint32_t* d_val_off, int64_t n, int64_t* d_out_ts, double* d_out_val,
                      int64_t* d_first_bad, pmx_stream stream) {
  PMX_REQUIRE(ctx && d_first_bad, "ctx and d_first_bad are required");
  PMX_REQUIRE(n >= 0, "n must be >= 0");
  PMX_CUDA(cudaMemsetAsync(d_first_bad, 0x7f, 2 * sizeof(int64_t), pmx_st(stream)));
  if (n == 0) return PMX_OK;
  PMX_REQUIRE(d_text && (d_out_ts || d_out_val), "NULL buffer");
  PMX_REQUIRE((!d_out_ts || d_ts_off) && (!d_out_val || d_val_off), "offsets missing for a requested output");
  parse_samples_kernel<<<(unsigned)((n + 255) / 256), 256, 0, pmx_st(stream)>>>(
      d_text, text_len, d_ts_off, d_val_off, n, d_out_ts, d_out_val, (unsigned long long*)d_first_bad,
      (unsigned long long*)d_first_bad + 1);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_parse_f64(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len, const uint32_t* d_val_off,
                  int64_t n, double* d_out, int64_t* d_first_bad, pmx_stream stream) {
  PMX_REQUIRE(ctx && d_first_bad, "ctx and d_first_bad are required");
  PMX_REQUIRE(n >= 0, "n must be >= 0");
  if (int rc = reset_first_bad(d_first_bad, pmx_st(stream))) return rc;
  if (n == 0) return PMX_OK;
  PMX_REQUIRE(d_text && d_val_off && d_out, "NULL buffer");
  parse_samples_kernel<<<(unsigned)((n + 255) / 256), 256, 0, pmx_st(stream)>>>(
      d_text, text_len, nullptr, d_val_off, n, nullptr, d_out, (unsigned long long*)d_first_bad, nullptr);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

int pmx_parse_ts_ms(pmx_ctx* ctx, const uint8_t* d_text, uint64_t text_len, const uint32_t* d_ts_off,
                    int64_t n, int64_t* d_out, int64_t* d_first_bad, pmx_stream stream) {
  PMX_REQUIRE(ctx && d_first_bad, "ctx and d_first_bad are required");
  PMX_REQUIRE(n >= 0, "n must be >= 0");
  if (int rc = reset_first_bad(d_first_bad, pmx_st(stream))) return rc;
  if (n == 0) return PMX_OK;
  PMX_REQUIRE(d_text && d_ts_off && d_out, "NULL buffer");
  parse_samples_kernel<<<(unsigned)((n + 255) / 256), 256, 0, pmx_st(stream)>>>(
      d_text, text_len, d_ts_off, nullptr, n, d_out, nullptr, nullptr, (unsigned long long*)d_first_bad);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}

}  // extern "C"
