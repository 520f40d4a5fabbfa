// strtod_fast.h -- the common case of K1 without a byte loop (SURVEY X2, app.py:188).
// (The kernel runs the SECOND form at the end of this file -- scan_token / decimal_scanned; the first
// form below it grew out of stays for the host fuzz test, which checks both.)
//
// Prometheus writes sample values as shortest round-trip decimals and timestamps as seconds with
// up to three decimals: plain [-]digits[.digits] in all but a few tokens (NaN, +Inf, exponents).
// A token of that shape and at most 24 bytes is decoded from a 32-byte window held in four 64-bit
// words: the terminator and the decimal point are found with byte-parallel compares, the digits are
// moved to the end of a 24-byte field (closing the point up), eight digits at a time are converted
// with three multiplications, and the result is
//   * an exact integer significand m (at most 19 digits) and a decimal exponent -f;
//   * binary64: m / 10^f when m <= 2^53 and f <= 22 (both exact, one correctly rounded division:
//     Clinger's fast path), else Eisel-Lemire on (m, -f) -- both give float()'s bits;
//   * int64 milliseconds: exact.
// Everything else (exponent notation, inf/nan spellings, blanks, underscores, more than 19
// digits, more than 24 bytes) reports FALLBACK and goes to the byte-wise state machine in
// strtod_core.h, which is exact for every token.  Host and device build (the CPU suite fuzzes this
// against float()).
#pragma once
#include "strtod_core.h"

namespace pmx {

enum { FAST_OK = 0, FAST_FALLBACK = 1 };

// byte k of x is 0x80 where byte k of v equals c (exact: no false positives from borrows below it)
PMX_HD inline uint64_t eq_bytes(uint64_t v, uint8_t c) {
  const uint64_t x = v ^ (0x0101010101010101ULL * c);
  return ~(((x & 0x7f7f7f7f7f7f7f7fULL) + 0x7f7f7f7f7f7f7f7fULL) | x | 0x7f7f7f7f7f7f7f7fULL);
}
PMX_HD inline int first_set_byte(uint64_t m) {       // m has 0x80 marks; index of the lowest one
#if defined(__CUDA_ARCH__)
  return (__ffsll((long long)m) - 1) >> 3;
#else
  return __builtin_ctzll(m) >> 3;
#endif
}
// position of the first byte equal to c in the first 24 bytes of the window, or 24
PMX_HD inline int find_byte24(const uint64_t* W, uint8_t c) {
  const uint64_t m0 = eq_bytes(W[0], c), m1 = eq_bytes(W[1], c), m2 = eq_bytes(W[2], c);
  if (m0) return first_set_byte(m0);
  if (m1) return 8 + first_set_byte(m1);
  if (m2) return 16 + first_set_byte(m2);
  return 24;
}
// eight ASCII digits, first digit in the lowest byte -> their value
PMX_HD inline uint32_t eight_digits(uint64_t v) {
  v = (v & 0x0F0F0F0F0F0F0F0FULL) * 2561 >> 8;
  v = (v & 0x00FF00FF00FF00FFULL) * 6553601 >> 16;
  return (uint32_t)((v & 0x0000FFFF0000FFFFULL) * 42949672960001ULL >> 32);
}
PMX_HD inline bool all_digits8(uint64_t v) {          // every byte in '0'..'9'
  return (((v + 0x4646464646464646ULL) | (v - 0x3030303030303030ULL)) & 0x8080808080808080ULL) == 0;
}
// the window shifted down by s bytes (0 <= s < 32), zero-filled from the top.  Two select stages
// for the whole words, one funnel shift for the rest: no indexed array, everything stays in registers.
PMX_HD inline void shr_bytes(const uint64_t* W, int s, uint64_t* out) {
  const bool w1 = (s & 8) != 0, w2 = (s & 16) != 0;
  const int b = (s & 7) * 8;
  const uint64_t y0 = w1 ? W[1] : W[0], y1 = w1 ? W[2] : W[1], y2 = w1 ? W[3] : W[2], y3 = w1 ? 0 : W[3];
  const uint64_t z0 = w2 ? y2 : y0, z1 = w2 ? y3 : y1, z2 = w2 ? 0 : y2, z3 = w2 ? 0 : y3;
  out[0] = b ? (z0 >> b) | (z1 << (64 - b)) : z0;
  out[1] = b ? (z1 >> b) | (z2 << (64 - b)) : z1;
  out[2] = b ? (z2 >> b) | (z3 << (64 - b)) : z2;
  out[3] = b ? (z3 >> b) : z3;
}
// ... shifted up by s bytes, zero-filled from the bottom
PMX_HD inline void shl_bytes(const uint64_t* W, int s, uint64_t* out) {
  const bool w1 = (s & 8) != 0, w2 = (s & 16) != 0;
  const int b = (s & 7) * 8;
  const uint64_t y3 = w1 ? W[2] : W[3], y2 = w1 ? W[1] : W[2], y1 = w1 ? W[0] : W[1], y0 = w1 ? 0 : W[0];
  const uint64_t z3 = w2 ? y1 : y3, z2 = w2 ? y0 : y2, z1 = w2 ? 0 : y1, z0 = w2 ? 0 : y0;
  out[3] = b ? (z3 << b) | (z2 >> (64 - b)) : z3;
  out[2] = b ? (z2 << b) | (z1 >> (64 - b)) : z2;
  out[1] = b ? (z1 << b) | (z0 >> (64 - b)) : z1;
  out[0] = b ? (z0 << b) : z0;
}
// all-ones in the bytes below position n (0 <= n <= 32) of word k
PMX_HD inline uint64_t low_bytes_mask(int n, int k) {
  int r = n - 8 * k;
  r = r < 0 ? 0 : (r > 8 ? 8 : r);
  return r == 8 ? ~0ULL : ((1ULL << (8 * r)) - 1ULL);
}

// powers of ten: tables outside the functions (PMX_POW5_QUALIFIER is `__device__ const` in the CUDA
// build and `static const` on the host, as for pow5_table.h), so that no call builds them on its stack
PMX_POW5_QUALIFIER uint64_t pmx_pow10_u64[20] = {
    1ULL, 10ULL, 100ULL, 1000ULL, 10000ULL, 100000ULL, 1000000ULL, 10000000ULL, 100000000ULL, 1000000000ULL,
    10000000000ULL, 100000000000ULL, 1000000000000ULL, 10000000000000ULL, 100000000000000ULL, 1000000000000000ULL,
    10000000000000000ULL, 100000000000000000ULL, 1000000000000000000ULL, 10000000000000000000ULL};
PMX_POW5_QUALIFIER double pmx_pow10_f64[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13,
                                               1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
PMX_HD inline uint64_t pow10_u64(int k) { return pmx_pow10_u64[k]; }
PMX_HD inline double pow10_f64(int k) { return pmx_pow10_f64[k]; }

// [-]digits[.digits] of length L (1..24) in window W -> sign, significand (<= 19 digits), number of
// digits behind the point, number of digits before it.  FALLBACK for any other shape.
// The digits are moved to the END of a 24-byte field (the fraction by 24 - L bytes, the integer part
// one byte further to close the point up), '0' in front: then three 8-digit groups give the value
// with no division and no power-of-ten lookup.
PMX_HD inline int fast_decimal(const uint64_t* W, int L, bool* neg, uint64_t* m, int* n_frac, int* n_int) {
  if (L < 1 || L > 24) return FAST_FALLBACK;
  const bool minus = (W[0] & 0xff) == '-';
  *neg = minus;
  const int s0 = minus ? 1 : 0;
  // the decimal point, if any, inside the first L bytes
  const uint64_t d0 = eq_bytes(W[0], (uint8_t)'.'), d1 = eq_bytes(W[1], (uint8_t)'.'), d2 = eq_bytes(W[2], (uint8_t)'.');
  int dot = 24;
  if (d0) dot = first_set_byte(d0); else if (d1) dot = 8 + first_set_byte(d1); else if (d2) dot = 16 + first_set_byte(d2);
  const bool has_dot = dot < L;
  const int nf = has_dot ? L - dot - 1 : 0;
  const int ni = (has_dot ? dot : L) - s0;
  const int nd = ni + nf;
  if (ni < 1 || (has_dot && nf < 1) || nd > 19) return FAST_FALLBACK;   // "5.", ".5", long mantissas: byte-wise path
  uint64_t A[4], B[4];
  shl_bytes(W, 24 - L, A);                             // the token's last byte at position 23: fraction (or all digits) in place
  shl_bytes(W, 25 - L, B);                             // one further: the integer part next to the fraction
  uint64_t D[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const uint64_t below_frac = low_bytes_mask(24 - nf, k), below_int = low_bytes_mask(24 - nd, k);
    const uint64_t ipart = has_dot ? B[k] : A[k];
    D[k] = (A[k] & ~below_frac) | (ipart & below_frac & ~below_int) | (0x3030303030303030ULL & below_int);
  }
  if (!all_digits8(D[0]) || !all_digits8(D[1]) || !all_digits8(D[2])) return FAST_FALLBACK;
  const uint64_t c0 = eight_digits(D[0]), c1 = eight_digits(D[1]), c2 = eight_digits(D[2]);
  *m = (c0 * 100000000ULL + c1) * 100000000ULL + c2;   // c0 < 1000 (nd <= 19): no overflow
  *n_frac = nf; *n_int = ni;
  return FAST_OK;
}

// value token -> bits of float(token)
PMX_HD inline int fast_f64(const uint64_t* W, int L, const uint64_t* __restrict__ tab, uint64_t* out_bits) {
  bool neg; uint64_t m; int nf, ni;
  if (fast_decimal(W, L, &neg, &m, &nf, &ni) != FAST_OK) return FAST_FALLBACK;
  const uint64_t sign = neg ? 0x8000000000000000ULL : 0;
  if (m == 0) { *out_bits = sign; return FAST_OK; }
  if (m <= (1ULL << 53) && nf <= 22) {
    const double v = (double)m / pow10_f64(nf);        // both operands exact: one IEEE rounding
    uint64_t b;
#if defined(__CUDA_ARCH__)
    b = (uint64_t)__double_as_longlong(v);
#else
    __builtin_memcpy(&b, &v, 8);
#endif
    *out_bits = b | sign;
    return FAST_OK;
  }
  const AdjMant a = eisel_lemire(-(int64_t)nf, m, tab);
  if (!a.certain) return FAST_FALLBACK;
  *out_bits = (a.mant | ((uint64_t)a.pow2 << 52)) | sign;
  return FAST_OK;
}

// timestamp token [-]digits[.d{1,3}] seconds -> int64 ms
PMX_HD inline int fast_ts_ms(const uint64_t* W, int L, int64_t* out) {
  bool neg; uint64_t m; int nf, ni;
  if (fast_decimal(W, L, &neg, &m, &nf, &ni) != FAST_OK) return FAST_FALLBACK;
  if (nf > 3 || ni > 15) return FAST_FALLBACK;         // sub-millisecond digits / overflow: the byte-wise path decides
  const int64_t ms = (int64_t)(m * pow10_u64(3 - nf));
  *out = neg ? -ms : ms;
  return FAST_OK;
}

// ---------------------------------------------------------------------------------------------
// Second form: the window is scanned ONCE for the bytes that are not digits.  The first of them is
// the decimal point or the terminator, the second (after a point) the terminator; everything between
// is digits by construction, so neither a search per character nor a digit check is needed, and a
// token of at most eight bytes -- most gauge readings -- is decoded from the first word alone.

// a window of THREE words shifted up by s bytes (0 <= s < 24): what lies above byte 23 is dropped
PMX_HD inline void shl_bytes3(const uint64_t* W, int s, uint64_t* out) {
  const bool w1 = (s & 8) != 0, w2 = (s & 16) != 0;
  const int b = (s & 7) * 8;
  const uint64_t y2 = w1 ? W[1] : W[2], y1 = w1 ? W[0] : W[1], y0 = w1 ? 0 : W[0];
  const uint64_t z2 = w2 ? y0 : y2, z1 = w2 ? 0 : y1, z0 = w2 ? 0 : y0;
  out[2] = b ? (z2 << b) | (z1 >> (64 - b)) : z2;
  out[1] = b ? (z1 << b) | (z0 >> (64 - b)) : z1;
  out[0] = b ? (z0 << b) : z0;
}

// 0x80 in every byte of v that is not '0'..'9' (exact per byte: no carry crosses a byte)
PMX_HD inline uint64_t nondigit_bytes(uint64_t v) {
  const uint64_t t = v & 0x7f7f7f7f7f7f7f7fULL;
  const uint64_t ge3a = t + 0x4646464646464646ULL;             // bit 7 set iff t >= 0x3A
  const uint64_t ge30 = t + 0x5050505050505050ULL;             // bit 7 set iff t >= 0x30
  return (v | ge3a | ~ge30) & 0x8080808080808080ULL;
}
PMX_HD inline uint32_t byte_at24(const uint64_t* W, int p) {   // byte p (0..23) of the window
  const uint64_t w = p < 8 ? W[0] : (p < 16 ? W[1] : W[2]);
  return (uint32_t)(w >> ((p & 7) * 8)) & 0xffu;
}
struct TokenScan { int L; int dot; bool neg; };                 // dot == L: no decimal point

// [-]digits[.digits] at byte 0 of the window, ended by `term` within the first 24 bytes
PMX_HD inline bool scan_token(const uint64_t* W, uint8_t term, TokenScan* s) {
  uint64_t f0 = nondigit_bytes(W[0]), f1 = nondigit_bytes(W[1]), f2 = 0;
  const bool neg = (W[0] & 0xff) == '-';
  if (neg) f0 &= ~0x80ULL;
  // the third word only when the first two do not hold both marks (tokens beyond ~15 bytes: rare)
  {
    const uint64_t g0 = f0 & (f0 - 1);                 // f0 without its lowest mark
    if (!(g0 | (f0 && f1) | (f1 & (f1 - 1)))) f2 = nondigit_bytes(W[2]);
  }
  int p1;
  if (f0) { p1 = first_set_byte(f0); f0 &= f0 - 1; }
  else if (f1) { p1 = 8 + first_set_byte(f1); f1 &= f1 - 1; }
  else if (f2) { p1 = 16 + first_set_byte(f2); f2 &= f2 - 1; }
  else return false;
  const uint32_t c1 = byte_at24(W, p1);
  s->neg = neg;
  if (c1 == term) { s->L = p1; s->dot = p1; return true; }
  if (c1 != '.') return false;
  int p2;
  if (f0) p2 = first_set_byte(f0);
  else if (f1) p2 = 8 + first_set_byte(f1);
  else if (f2) p2 = 16 + first_set_byte(f2);
  else return false;
  if (byte_at24(W, p2) != term) return false;
  s->L = p2; s->dot = p1;
  return true;
}

// significand (<= 19 digits) and digit counts of a scanned token; FALLBACK for "5.", ".5", long mantissas
PMX_HD inline int decimal_scanned(const uint64_t* W, const TokenScan& sc, uint64_t* m, int* n_frac, int* n_int) {
  const int L = sc.L, s0 = sc.neg ? 1 : 0;
  const bool has_dot = sc.dot < L;
  const int nf = has_dot ? L - sc.dot - 1 : 0;
  const int ni = sc.dot - s0;
  const int nd = ni + nf;
  if (ni < 1 || (has_dot && nf < 1) || nd > 19) return FAST_FALLBACK;
  *n_frac = nf; *n_int = ni;
  if (L <= 8) {
    // the whole token sits in the first word: its last byte to byte 7, the integer part one byte
    // further to close the point up, '0' in front, one 8-digit conversion
    const uint64_t A = W[0] << (8 * (8 - L));
    const uint64_t B = A << 8;
    const int rf = 8 - nf, ri = 8 - nd;
    const uint64_t below_frac = rf >= 8 ? ~0ULL : ((1ULL << (8 * rf)) - 1ULL);
    const uint64_t below_int = ri >= 8 ? ~0ULL : ((1ULL << (8 * ri)) - 1ULL);
    const uint64_t D = (A & ~below_frac) | ((has_dot ? B : A) & below_frac & ~below_int) | (0x3030303030303030ULL & below_int);
    *m = eight_digits(D);
    return FAST_OK;
  }
  if (L <= 16) {
    // two words (timestamps, counters): the same with the token's last byte at position 15
    const int sh = 8 * (16 - L);                       // 0 .. 56 bits
    const uint64_t A1 = sh ? (W[1] << sh) | (W[0] >> (64 - sh)) : W[1], A0 = W[0] << sh;
    const uint64_t B1 = (A1 << 8) | (A0 >> 56), B0 = A0 << 8;
    const uint64_t bf0 = low_bytes_mask(16 - nf, 0), bf1 = low_bytes_mask(16 - nf, 1);
    const uint64_t bi0 = low_bytes_mask(16 - nd, 0), bi1 = low_bytes_mask(16 - nd, 1);
    const uint64_t D0 = (A0 & ~bf0) | ((has_dot ? B0 : A0) & bf0 & ~bi0) | (0x3030303030303030ULL & bi0);
    const uint64_t D1 = (A1 & ~bf1) | ((has_dot ? B1 : A1) & bf1 & ~bi1) | (0x3030303030303030ULL & bi1);
    *m = (uint64_t)eight_digits(D0) * 100000000ULL + eight_digits(D1);
    return FAST_OK;
  }
  uint64_t A[3], B[3];
  shl_bytes3(W, 24 - L, A);                            // the token's last byte at position 23: fraction (or all digits) in place
  shl_bytes3(W, 25 - L, B);                            // one further: the integer part next to the fraction
  uint64_t D[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const uint64_t below_frac = low_bytes_mask(24 - nf, k), below_int = low_bytes_mask(24 - nd, k);
    const uint64_t ipart = has_dot ? B[k] : A[k];
    D[k] = (A[k] & ~below_frac) | (ipart & below_frac & ~below_int) | (0x3030303030303030ULL & below_int);
  }
  const uint64_t c1 = eight_digits(D[1]), c2 = eight_digits(D[2]);
  uint64_t hi = 0;
  if (nd > 16) hi = eight_digits(D[0]);                // < 1000 (nd <= 19)
  *m = (hi * 100000000ULL + c1) * 100000000ULL + c2;
  return FAST_OK;
}

// value token ended by `term` -> bits of float(token)
PMX_HD inline int fast_f64_scan(const uint64_t* W, uint8_t term, const uint64_t* __restrict__ tab, uint64_t* out_bits) {
  TokenScan sc;
  if (!scan_token(W, term, &sc)) return FAST_FALLBACK;
  uint64_t m; int nf, ni;
  if (decimal_scanned(W, sc, &m, &nf, &ni) != FAST_OK) return FAST_FALLBACK;
  const uint64_t sign = sc.neg ? 0x8000000000000000ULL : 0;
  if (m == 0) { *out_bits = sign; return FAST_OK; }
  if (m <= (1ULL << 53) && nf <= 22) {
    const double v = (double)m / pow10_f64(nf);        // both operands exact: one IEEE rounding
    uint64_t b;
#if defined(__CUDA_ARCH__)
    b = (uint64_t)__double_as_longlong(v);
#else
    __builtin_memcpy(&b, &v, 8);
#endif
    *out_bits = b | sign;
    return FAST_OK;
  }
  const AdjMant a = eisel_lemire(-(int64_t)nf, m, tab);
  if (!a.certain) return FAST_FALLBACK;
  *out_bits = (a.mant | ((uint64_t)a.pow2 << 52)) | sign;
  return FAST_OK;
}

// timestamp token [-]digits[.d{1,3}] seconds ended by `term` -> int64 ms
PMX_HD inline int fast_ts_ms_scan(const uint64_t* W, uint8_t term, int64_t* out) {
  TokenScan sc;
  if (!scan_token(W, term, &sc)) return FAST_FALLBACK;
  uint64_t m; int nf, ni;
  if (decimal_scanned(W, sc, &m, &nf, &ni) != FAST_OK) return FAST_FALLBACK;
  if (nf > 3 || ni > 15) return FAST_FALLBACK;         // sub-millisecond digits / overflow: the byte-wise path decides
  const int64_t ms = (int64_t)(m * pow10_u64(3 - nf));
  *out = sc.neg ? -ms : ms;
  return FAST_OK;
}

}  // namespace pmx
