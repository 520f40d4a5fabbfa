// byte_masks.h -- the SWAR pieces of the device tokenizer (tokenize_device.cu): byte classes of a
// 16-byte segment as 16-bit masks, and the in-string state inside the segment.  Plain integer code,
// usable on host and device; the host build exists only for the CPU test of these functions
// (tests/host/strtod_host.cpp, tests/test_host.py).
#pragma once
#include <stdint.h>
#if defined(__CUDACC__)
#define PMX_BM __host__ __device__ __forceinline__
#else
#define PMX_BM inline
#endif

namespace pmx {

// bit k (k = 0..3) set iff byte k of w equals the byte repeated in pat4.  Exact: the zero-byte test
// carries nothing from one byte into the next, and the multiply gathers bits 0, 8, 16, 24 into
// bits 24..27 with no two partial products on the same bit.
PMX_BM unsigned eq4(uint32_t w, uint32_t pat4) {
  const uint32_t x = w ^ pat4;
  const uint32_t t = (x & 0x7f7f7f7fu) + 0x7f7f7f7fu;
  const uint32_t z = ~(t | x) & 0x80808080u;
  return ((z >> 7) * 0x01020408u) >> 24;
}
PMX_BM unsigned eq16(const uint32_t* w, uint32_t pat4) {
  return eq4(w[0], pat4) | (eq4(w[1], pat4) << 4) | (eq4(w[2], pat4) << 8) | (eq4(w[3], pat4) << 12);
}
// bit k set iff bit 5 of byte k is set ('{' against '[')
PMX_BM unsigned bit5_16(const uint32_t* w) {
  unsigned m = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) m |= ((((w[j] >> 5) & 0x01010101u) * 0x01020408u) >> 24) << (4 * j);
  return m;
}
// '[' or '{': the two differ in bit 5 alone
PMX_BM unsigned open_mask(const uint32_t* w) {
  const uint32_t v[4] = {w[0] | 0x20202020u, w[1] | 0x20202020u, w[2] | 0x20202020u, w[3] | 0x20202020u};
  return eq16(v, 0x7b7b7b7bu);
}
// may this word hold the byte?  (false positives allowed, no false negatives)
PMX_BM uint32_t maybe4(uint32_t w, uint32_t pat4) {
  const uint32_t x = w ^ pat4;
  return (x - 0x01010101u) & ~x & 0x80808080u;
}

// byte k (0..15) of the segment
PMX_BM unsigned seg_byte(const uint32_t* w, int k) {
  const uint32_t lo = (k & 8) ? w[2] : w[0], hi = (k & 8) ? w[3] : w[1];
  return ((k & 4 ? hi : lo) >> (8 * (k & 3))) & 0xffu;
}

// positions of the segment that lie inside a string when the segment begins outside one; bit 16 of
// the result: the segment's own parity
PMX_BM unsigned inside_mask(unsigned q) {
  unsigned x = q;
  x ^= x << 1; x ^= x << 2; x ^= x << 4; x ^= x << 8;           // inclusive prefix parity
  return (x << 1) & 0x1ffffu;
}

}  // namespace pmx
