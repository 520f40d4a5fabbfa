// quantile_keys.cuh -- float64 <-> uint64 key that sorts like the number (-0.0 below +0.0; NaN is
// never keyed).  Shared by quantile.cu and bracket_select.cu.
#pragma once
#include <stdint.h>

__device__ __forceinline__ uint64_t f64_key(double v) {
  const uint64_t b = (uint64_t)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double key_f64(uint64_t k) {
  const uint64_t b = (k >> 63) ? (k & 0x7fffffffffffffffULL) : ~k;
  return __longlong_as_double((long long)b);
}
