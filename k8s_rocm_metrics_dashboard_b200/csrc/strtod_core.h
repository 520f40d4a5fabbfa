// strtod_core.h -- decimal text -> binary64, correctly rounded, with CPython float() grammar.
//
// Takes over float(metric["value"][1]) (/root/reference/app.py:188).  Compiles for host and
// device (PMX_HD); the host build is only used by the CPU fuzz test of the state machine, the
// product runs it inside the K1 kernel (parse.cu).
//
// Pipeline per token:
//   1. scan: sign, digits (single underscores between digits), '.', exponent, or inf/nan.
//      The first 19 significant digits go into a uint64 `w`, the rest only set `truncated`.
//   2. Eisel-Lemire: w * 10^q via one or two 64x64->128 multiplies with a 128-bit power-of-five
//      table (pow5_table.h) -> correctly rounded double, or "uncertain".
//   3. truncated inputs: run step 2 for w and w+1; equal answers are final.
//   4. otherwise decide exactly with big integers (<= 770 digits): compare the decimal value
//      with the midpoints around the candidate and step the candidate until it is bracketed.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define PMX_HD __host__ __device__
#define PMX_COLD __noinline__
#else
#define PMX_HD
#define PMX_COLD __attribute__((noinline))
#endif

namespace pmx {

enum ParseStatus : int { PARSE_OK = 0, PARSE_BAD = 1 };

// ----------------------------------------------------------------------------- helpers
PMX_HD inline void mul64(uint64_t a, uint64_t b, uint64_t& hi, uint64_t& lo) {
#if defined(__CUDA_ARCH__)
  lo = a * b;
  hi = __umul64hi(a, b);
#else
  unsigned __int128 p = (unsigned __int128)a * b;
  lo = (uint64_t)p;
  hi = (uint64_t)(p >> 64);
#endif
}
PMX_HD inline int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return __builtin_clzll(x);
#endif
}
PMX_HD inline bool is_digit(int c) { return (unsigned)(c - '0') <= 9u; }
PMX_HD inline bool is_space(int c) {           // str.isspace() restricted to ASCII
  return c == ' ' || (c >= 0x09 && c <= 0x0d) || (c >= 0x1c && c <= 0x1f);
}
PMX_HD inline int lower(int c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }

struct Byte {                                   // bounded byte reader over [p, p+len)
  const uint8_t* p;
  int64_t len;
  PMX_HD int at(int64_t i) const {
#if defined(__CUDA_ARCH__)
    return i < len ? (int)__ldg(p + i) : -1;
#else
    return i < len ? (int)p[i] : -1;
#endif
  }
};

// ----------------------------------------------------------------------------- Eisel-Lemire
struct AdjMant { uint64_t mant; int32_t pow2; bool certain; };   // bits = mant | pow2 << 52

// w != 0.  Returns the correctly rounded (mant, pow2) of w * 10^q, certain=false when the
// 128-bit product cannot decide the rounding (the caller then verifies exactly).
PMX_HD inline AdjMant eisel_lemire(int64_t q, uint64_t w, const uint64_t* __restrict__ tab) {
  AdjMant am; am.certain = true;
  if (q < -342) { am.mant = 0; am.pow2 = 0; return am; }           // w < 2^64: below 2^-1075
  if (q > 308)  { am.mant = 0; am.pow2 = 0x7ff; return am; }
  int lz = clz64(w);
  w <<= lz;
  const int idx = 2 * (int)(q + 342);
  uint64_t hi, lo;
  mul64(w, tab[idx], hi, lo);
  if ((hi & 0x1ff) == 0x1ff) {                                     // need the second word
    uint64_t hi2, lo2;
    mul64(w, tab[idx + 1], hi2, lo2);
    lo += hi2;
    if (hi2 > lo) hi++;
    (void)lo2;
  }
  if (lo == 0xffffffffffffffffULL && (q < -27 || q > 55)) am.certain = false;
  const int upper = (int)(hi >> 63);
  const int shift = upper + 64 - 52 - 3;
  uint64_t mant = hi >> shift;
  int32_t pow2 = (int32_t)((((int64_t)(152170 + 65536) * q) >> 16) + 63) + upper - lz + 1023;
  if (pow2 <= 0) {                                                 // subnormal
    if (-pow2 + 1 >= 64) { am.mant = 0; am.pow2 = 0; return am; }
    mant >>= (-pow2 + 1);
    mant += (mant & 1);
    mant >>= 1;
    am.pow2 = (mant < (1ULL << 52)) ? 0 : 1;
    am.mant = mant & ~(1ULL << 52);
    return am;
  }
  if (lo <= 1 && q >= -4 && q <= 23 && (mant & 3) == 1) {          // exactly halfway: to even
    if ((mant << shift) == hi) mant &= ~1ULL;
  }
  mant += (mant & 1);
  mant >>= 1;
  if (mant >= (2ULL << 52)) { mant = 1ULL << 52; pow2++; }
  mant &= ~(1ULL << 52);
  if (pow2 >= 0x7ff) { am.mant = 0; am.pow2 = 0x7ff; return am; }
  am.mant = mant; am.pow2 = pow2;
  return am;
}

// ----------------------------------------------------------------------------- big integers
// Little-endian base-2^32, fixed capacity.  3456 bits cover 770 decimal digits (2558 bits)
// or (2^54) * 5^1100 (2609 bits) plus the alignment shift of at most a few hundred bits.
struct Big {
  static const int CAP = 108;
  uint32_t d[CAP];
  int n;                                          // used limbs; value 0 <=> n == 0
  bool overflow;
  PMX_HD void set(uint64_t v) {
    n = 0; overflow = false;
    while (v) { d[n++] = (uint32_t)v; v >>= 32; }
  }
  PMX_HD void mul_small_add(uint32_t m, uint32_t add) {
    uint64_t carry = add;
    for (int i = 0; i < n; ++i) {
      uint64_t t = (uint64_t)d[i] * m + carry;
      d[i] = (uint32_t)t; carry = t >> 32;
    }
    if (carry) { if (n < CAP) d[n++] = (uint32_t)carry; else overflow = true; }
  }
  PMX_HD void mul_pow5(int64_t k) {
    while (k >= 13) { mul_small_add(1220703125u, 0); k -= 13; }
    uint32_t m = 1;
    while (k-- > 0) m *= 5;
    if (m != 1) mul_small_add(m, 0);
  }
  PMX_HD void shl(int64_t bits) {
    if (n == 0 || bits <= 0) return;
    int64_t limbs = bits >> 5; int b = (int)(bits & 31);
    if (n + limbs + 1 > CAP) { overflow = true; return; }
    if (b) {
      uint32_t carry = 0;
      for (int i = 0; i < n; ++i) { uint32_t v = d[i]; d[i] = (v << b) | carry; carry = v >> (32 - b); }
      if (carry) d[n++] = carry;
    }
    if (limbs) {
      for (int i = n - 1; i >= 0; --i) d[i + limbs] = d[i];
      for (int i = 0; i < (int)limbs; ++i) d[i] = 0;
      n += (int)limbs;
    }
  }
  PMX_HD int cmp(const Big& o) const {
    if (n != o.n) return n < o.n ? -1 : 1;
    for (int i = n - 1; i >= 0; --i) if (d[i] != o.d[i]) return d[i] < o.d[i] ? -1 : 1;
    return 0;
  }
};

// value = D * 10^q10 compared with M * 2^e2 (M odd midpoint numerator).  Returns -1/0/+1.
PMX_HD inline int cmp_decimal_binary(const Big& D, int64_t q10, uint64_t M, int64_t e2) {
  Big L = D, R; R.set(M);
  int64_t l2 = 0, r2 = e2;
  if (q10 >= 0) { L.mul_pow5(q10); l2 = q10; }
  else          { R.mul_pow5(-q10); r2 = e2 - q10; }
  if (l2 > r2) L.shl(l2 - r2); else R.shl(r2 - l2);
  if (L.overflow) return 1;                       // cannot happen for candidates within 2 ulp
  if (R.overflow) return -1;
  return L.cmp(R);
}

// ----------------------------------------------------------------------------- scanner
struct Scan {
  uint64_t w;          // first <= 19 significant digits
  int64_t q;           // value ~= w * 10^q (exactly if !truncated)
  int64_t dp;          // significant digits before the decimal point (may be negative)
  int64_t exp10;       // explicit exponent (saturated)
  int64_t first_sig;   // byte index of the first significant digit (-1 if none)
  int64_t mant_end;    // byte index one past the mantissa
  int64_t nsig;        // number of significant digits (trailing zeros included)
  bool neg, truncated, special; uint64_t special_bits;
};

PMX_HD inline bool match_ci(const Byte& s, int64_t i, const char* word, int n) {
  for (int k = 0; k < n; ++k) if (lower(s.at(i + k)) != word[k]) return false;
  return true;
}

// Parses [begin,end) with surrounding whitespace already stripped.
PMX_HD inline int scan_number(const Byte& s, int64_t i, int64_t end, Scan& r) {
  r.neg = false; r.truncated = false; r.special = false;
  r.w = 0; r.q = 0; r.dp = 0; r.exp10 = 0; r.first_sig = -1; r.nsig = 0;
  if (i >= end) return PARSE_BAD;
  int c = s.at(i);
  if (c == '+' || c == '-') { r.neg = (c == '-'); ++i; }
  if (i >= end) return PARSE_BAD;
  c = lower(s.at(i));
  if (c == 'i' || c == 'n') {                                   // inf / infinity / nan
    int64_t rem = end - i;
    if (c == 'n') {
      if (rem == 3 && match_ci(s, i, "nan", 3)) { r.special = true; r.special_bits = 0x7ff8000000000000ULL; return PARSE_OK; }
      return PARSE_BAD;
    }
    if ((rem == 3 && match_ci(s, i, "inf", 3)) || (rem == 8 && match_ci(s, i, "infinity", 8))) {
      r.special = true; r.special_bits = 0x7ff0000000000000ULL; return PARSE_OK;
    }
    return PARSE_BAD;
  }
  // mantissa: digits [ '.' digits ], underscores only between two digits
  int64_t ndig_total = 0;            // digits seen (any)
  int64_t nsig = 0;                  // significant digits seen
  int64_t int_sig = 0;               // significant digits before the point
  int64_t frac_lead_zero = 0;        // zeros after the point before the first significant digit
  bool seen_point = false, prev_digit = false, prev_us = false;
  for (; i < end; ++i) {
    c = s.at(i);
    if (is_digit(c)) {
      ++ndig_total;
      if (c != '0' || nsig > 0) {
        if (nsig == 0) r.first_sig = i;
        if (nsig < 19) r.w = r.w * 10 + (uint64_t)(c - '0');
        else if (c != '0') r.truncated = true;
        ++nsig;
        if (!seen_point) ++int_sig;
      } else if (seen_point) {
        ++frac_lead_zero;
      }
      prev_digit = true; prev_us = false;
    } else if (c == '_') {
      if (!prev_digit) return PARSE_BAD;                         // needs a digit before ...
      prev_digit = false; prev_us = true;
    } else if (c == '.') {
      if (seen_point || prev_us) return PARSE_BAD;
      seen_point = true; prev_digit = false;
    } else break;
  }
  if (prev_us) return PARSE_BAD;                                 // ... and a digit after
  if (ndig_total == 0) return PARSE_BAD;
  r.mant_end = i;
  r.nsig = nsig;
  r.dp = seen_point ? (int_sig > 0 ? int_sig : -frac_lead_zero) : int_sig;
  if (i < end) {
    c = s.at(i);
    if (c != 'e' && c != 'E') return PARSE_BAD;
    ++i;
    bool eneg = false;
    if (i < end && (s.at(i) == '+' || s.at(i) == '-')) { eneg = (s.at(i) == '-'); ++i; }
    int64_t e = 0; int64_t nd = 0; prev_digit = false; prev_us = false;
    for (; i < end; ++i) {
      c = s.at(i);
      if (is_digit(c)) { if (e < 100000000) e = e * 10 + (c - '0'); ++nd; prev_digit = true; prev_us = false; }
      else if (c == '_') { if (!prev_digit) return PARSE_BAD; prev_digit = false; prev_us = true; }
      else return PARSE_BAD;
    }
    if (nd == 0 || prev_us) return PARSE_BAD;
    r.exp10 = eneg ? -e : e;
  }
  // value = 0.d1d2...dn * 10^(dp + exp10);  w holds min(n,19) digits
  int64_t used = nsig < 19 ? nsig : 19;
  r.q = r.dp - used + r.exp10;
  return PARSE_OK;
}

// Exact decision for the hard cases.  `bits` is a candidate (finite, non-negative).
PMX_HD PMX_COLD inline uint64_t exact_round(const Byte& s, const Scan& sc, uint64_t bits) {
  // rebuild up to 769 significant digits, then a sticky 1 if anything non-zero was dropped
  Big D; D.set(0);
  int64_t taken = 0; bool sticky = false;
  uint32_t chunk = 0; int in_chunk = 0;
  for (int64_t i = sc.first_sig; i < sc.mant_end; ++i) {
    int c = s.at(i);
    if (!is_digit(c)) continue;
    if (taken < 769) {
      chunk = chunk * 10 + (uint32_t)(c - '0'); ++in_chunk; ++taken;
      if (in_chunk == 9) { D.mul_small_add(1000000000u, chunk); chunk = 0; in_chunk = 0; }
    } else if (c != '0') sticky = true;
  }
  if (in_chunk) { uint32_t m = 1; for (int k = 0; k < in_chunk; ++k) m *= 10; D.mul_small_add(m, chunk); }
  int64_t q10 = sc.dp - taken + sc.exp10;
  if (sticky) { D.mul_small_add(10u, 1u); q10 -= 1; }
  const uint64_t INF = 0x7ff0000000000000ULL;
  for (int iter = 0; iter < 8; ++iter) {
    if (bits >= INF) bits = INF - 1;                              // start from DBL_MAX
    // midpoint above: (2m+1) * 2^(e-1)
    uint64_t ex = bits >> 52, fr = bits & ((1ULL << 52) - 1);
    uint64_t m = ex ? (fr | (1ULL << 52)) : fr;
    int64_t e = ex ? (int64_t)ex - 1075 : -1074;
    int up = cmp_decimal_binary(D, q10, 2 * m + 1, e - 1);
    if (up > 0) { bits += 1; if (bits >= INF) return INF; continue; }
    if (up == 0) return (bits & 1) ? bits + 1 : bits;             // tie -> even (may become inf)
    if (bits == 0) return 0;
    uint64_t pb = bits - 1;                                       // midpoint below
    uint64_t pex = pb >> 52, pfr = pb & ((1ULL << 52) - 1);
    uint64_t pm = pex ? (pfr | (1ULL << 52)) : pfr;
    int64_t pe = pex ? (int64_t)pex - 1075 : -1074;
    int dn = cmp_decimal_binary(D, q10, 2 * pm + 1, pe - 1);
    if (dn < 0) { bits = pb; continue; }
    if (dn == 0) return (bits & 1) ? pb : bits;
    return bits;
  }
  return bits;
}

// Full token -> bits.  [p, p+len) is the token without its terminator.
PMX_HD inline int parse_f64(const uint8_t* p, int64_t len, const uint64_t* __restrict__ tab,
                            uint64_t* out_bits) {
  Byte s; s.p = p; s.len = len;
  int64_t b = 0, e = len;
  while (b < e && is_space(s.at(b))) ++b;
  while (e > b && is_space(s.at(e - 1))) --e;
  Scan sc;
  if (scan_number(s, b, e, sc) != PARSE_OK) { *out_bits = 0x7ff8000000000000ULL; return PARSE_BAD; }
  const uint64_t sign = sc.neg ? 0x8000000000000000ULL : 0;
  if (sc.special) { *out_bits = sc.special_bits | sign; return PARSE_OK; }
  if (sc.w == 0) { *out_bits = sign; return PARSE_OK; }
  // clamp the exponent: beyond these the result is 0 / inf whatever the digits say
  int64_t q = sc.q;
  if (q > 400) q = 400;
  if (q < -400) q = -400;
  AdjMant a = eisel_lemire(q, sc.w, tab);
  uint64_t bits = a.mant | ((uint64_t)a.pow2 << 52);
  bool hard = !a.certain;
  if (sc.truncated && !hard) {
    AdjMant a2 = eisel_lemire(q, sc.w + 1, tab);
    if (!a2.certain || a2.mant != a.mant || a2.pow2 != a.pow2) hard = true;
  }
  if (hard) bits = exact_round(s, sc, bits);
  *out_bits = bits | sign;
  return PARSE_OK;
}

// Timestamp token: [+-]digits[.digits<=3] seconds -> int64 ms, exact.
PMX_HD inline int parse_ts_ms(const uint8_t* p, int64_t len, int64_t* out) {
  Byte s; s.p = p; s.len = len;
  int64_t i = 0, e = len;
  while (i < e && is_space(s.at(i))) ++i;
  while (e > i && is_space(s.at(e - 1))) --e;
  bool neg = false;
  if (i < e && s.at(i) == '-') { neg = true; ++i; }
  int64_t v = 0, nd = 0;
  for (; i < e && is_digit(s.at(i)); ++i, ++nd) {
    if (v > 922337203685476LL) return PARSE_BAD;                  // would overflow int64 ms
    v = v * 10 + (s.at(i) - '0');
  }
  if (nd == 0) return PARSE_BAD;
  int64_t frac = 0, nf = 0;
  if (i < e && s.at(i) == '.') {
    ++i;
    for (; i < e && is_digit(s.at(i)); ++i, ++nf) {
      if (nf < 3) frac = frac * 10 + (s.at(i) - '0');
      else if (s.at(i) != '0') return PARSE_BAD;                  // sub-millisecond digits
    }
    if (nf == 0) return PARSE_BAD;
  }
  if (i != e) return PARSE_BAD;
  for (int64_t k = nf; k < 3; ++k) frac *= 10;
  int64_t ms = v * 1000 + frac;
  *out = neg ? -ms : ms;
  return PARSE_OK;
}

}  // namespace pmx
