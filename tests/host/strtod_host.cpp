// Host build of csrc/strtod_core.h for the CPU-side fuzz test of the parser state machine
// (tests/test_strtod_host.py).  Test infrastructure: the product only runs this code on the GPU.
#include <stdint.h>
#include <stddef.h>
#define PMX_POW5_QUALIFIER static const
#include "../../k8s_rocm_metrics_dashboard_b200/csrc/pow5_table.h"
#include "../../k8s_rocm_metrics_dashboard_b200/csrc/strtod_core.h"

extern "C" int host_parse_f64(const uint8_t* p, int64_t len, uint64_t* bits) {
  return pmx::parse_f64(p, len, pmx_pow5_128, bits);
}
extern "C" int host_parse_ts_ms(const uint8_t* p, int64_t len, int64_t* out) {
  return pmx::parse_ts_ms(p, len, out);
}
// many tokens at once: offsets[n+1] into buf
extern "C" void host_parse_many(const uint8_t* buf, const int64_t* off, int64_t n, uint64_t* bits, int32_t* status) {
  for (int64_t i = 0; i < n; ++i) status[i] = pmx::parse_f64(buf + off[i], off[i + 1] - off[i], pmx_pow5_128, bits + i);
}

// the window form of the common case (csrc/strtod_fast.h): status 0 = decoded, 1 = left to the byte-wise path
#include "../../k8s_rocm_metrics_dashboard_b200/csrc/strtod_fast.h"
#include <string.h>
extern "C" int host_fast_f64(const uint8_t* p, int64_t len, uint64_t* bits) {
  uint64_t W[4] = {0, 0, 0, 0};
  if (len >= 24) return pmx::FAST_FALLBACK;
  memcpy(W, p, (size_t)len);
  return pmx::fast_f64(W, (int)len, pmx_pow5_128, bits);
}
extern "C" int host_fast_ts_ms(const uint8_t* p, int64_t len, int64_t* out) {
  uint64_t W[4] = {0, 0, 0, 0};
  if (len >= 24) return pmx::FAST_FALLBACK;
  memcpy(W, p, (size_t)len);
  return pmx::fast_ts_ms(W, (int)len, out);
}

// the scanning form: the window holds the token AND its terminator (what the kernel sees)
extern "C" int host_scan_f64(const uint8_t* p, int64_t len, uint8_t term, uint64_t* bits) {
  uint8_t buf[32];
  memset(buf, 'x', sizeof buf);                         // whatever follows in the body
  if (len >= 31) return pmx::FAST_FALLBACK;
  memcpy(buf, p, (size_t)len); buf[len] = term;
  uint64_t W[4]; memcpy(W, buf, 32);
  return pmx::fast_f64_scan(W, term, pmx_pow5_128, bits);
}
extern "C" int host_scan_ts_ms(const uint8_t* p, int64_t len, uint8_t term, int64_t* out) {
  uint8_t buf[32];
  memset(buf, '7', sizeof buf);                         // digits behind the terminator must not matter
  if (len >= 31) return pmx::FAST_FALLBACK;
  memcpy(buf, p, (size_t)len); buf[len] = term;
  uint64_t W[4]; memcpy(W, buf, 32);
  return pmx::fast_ts_ms_scan(W, term, out);
}

// the byte-class masks of the device tokenizer (csrc/byte_masks.h) over one 16-byte segment:
// out[0] quotes, out[1] '[' or '{', out[2] bit 5 of every byte, out[3] commas, out[4] inside_mask(out[0]),
// out[5] = bit k set iff word k>>2 "may hold" a backslash (maybe4), out[6..21] = seg_byte(k)
#include "../../k8s_rocm_metrics_dashboard_b200/csrc/byte_masks.h"
extern "C" void host_segment_masks(const uint8_t* seg, uint32_t* out) {
  uint32_t w[4];
  memcpy(w, seg, 16);
  out[0] = pmx::eq16(w, 0x22222222u);
  out[1] = pmx::open_mask(w);
  out[2] = pmx::bit5_16(w);
  out[3] = pmx::eq16(w, 0x2c2c2c2cu);
  out[4] = pmx::inside_mask(out[0]);
  out[5] = 0;
  for (int j = 0; j < 4; ++j) if (pmx::maybe4(w[j], 0x5c5c5c5cu)) out[5] |= 0xfu << (4 * j);
  for (int k = 0; k < 16; ++k) out[6 + k] = pmx::seg_byte(w, k);
}
extern "C" void host_segment_masks_many(const uint8_t* segs, int64_t n, uint32_t* out) {
  for (int64_t i = 0; i < n; ++i) host_segment_masks(segs + 16 * i, out + 22 * i);
}
