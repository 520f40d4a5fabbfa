// tokenize.cu -- X1: Prometheus response body -> per-sample token offsets (HOST code).
//
// Takes over json.loads at /root/reference/app.py:163,178 for the hot part of the body: it
// walks the JSON once, records where each sample's timestamp and value token start, the CSR
// sample offsets per series and the byte span of each series' "metric" object, and never
// materialises objects.  Numbers are decoded later on the GPU (parse.cu).
#include <string.h>
#include "pmx_common.cuh"

namespace {

struct Cur {
  const char* base;
  const char* p;
  const char* end;
  bool fail(const char* what) const {
    pmx_fail(PMX_E_JSON, "pmx_tokenize: %s at byte %lld", what, (long long)(p - base));
    return false;
  }
};

inline void ws(Cur& c) {
  while (c.p < c.end && (*c.p == ' ' || *c.p == '\n' || *c.p == '\t' || *c.p == '\r')) ++c.p;
}
inline bool expect(Cur& c, char ch) {
  ws(c);
  if (c.p >= c.end || *c.p != ch) return false;
  ++c.p;
  return true;
}
// at an opening quote: returns content span, advances past the closing quote
bool string_span(Cur& c, const char** s, size_t* n, bool* escaped) {
  ws(c);
  if (c.p >= c.end || *c.p != '"') return c.fail("expected string");
  const char* q = ++c.p;
  bool esc = false;
  for (;;) {
    const char* r = (const char*)memchr(q, '"', (size_t)(c.end - q));
    if (!r) return c.fail("unterminated string");
    // count the backslashes directly before r
    const char* b = r;
    while (b > c.p && b[-1] == '\\') --b;
    if (((r - b) & 1) == 0) {
      if (!esc) esc = memchr(c.p, '\\', (size_t)(r - c.p)) != nullptr;
      *s = c.p; *n = (size_t)(r - c.p); *escaped = esc;
      c.p = r + 1;
      return true;
    }
    esc = true;
    q = r + 1;
  }
}
bool skip_value(Cur& c, int depth);
bool skip_container(Cur& c, char open, char close, int depth) {
  if (depth > 64) return c.fail("nesting too deep");
  ++c.p;  // open
  ws(c);
  if (c.p < c.end && *c.p == close) { ++c.p; return true; }
  for (;;) {
    if (open == '{') {
      const char* s; size_t n; bool e;
      if (!string_span(c, &s, &n, &e)) return false;
      if (!expect(c, ':')) return c.fail("expected ':'");
    }
    if (!skip_value(c, depth + 1)) return false;
    ws(c);
    if (c.p >= c.end) return c.fail("unterminated container");
    if (*c.p == ',') { ++c.p; continue; }
    if (*c.p == close) { ++c.p; return true; }
    return c.fail("expected ',' or close");
  }
}
bool skip_value(Cur& c, int depth) {
  ws(c);
  if (c.p >= c.end) return c.fail("unexpected end");
  char ch = *c.p;
  if (ch == '"') { const char* s; size_t n; bool e; return string_span(c, &s, &n, &e); }
  if (ch == '{') return skip_container(c, '{', '}', depth);
  if (ch == '[') return skip_container(c, '[', ']', depth);
  const char* q = c.p;
  while (c.p < c.end && *c.p != ',' && *c.p != '}' && *c.p != ']' && *c.p != ' ' && *c.p != '\n' &&
         *c.p != '\t' && *c.p != '\r') ++c.p;
  if (c.p == q) return c.fail("expected value");
  return true;
}
inline bool key_is(const char* s, size_t n, const char* lit) { return strlen(lit) == n && memcmp(s, lit, n) == 0; }

// decode JSON escapes of [s, s+n) into out; returns bytes written or -1
long decode_escapes(const char* s, size_t n, char* out, size_t cap) {
  size_t o = 0;
  for (size_t i = 0; i < n; ++i) {
    unsigned char ch = (unsigned char)s[i];
    unsigned cp = ch;
    if (ch == '\\' && i + 1 < n) {
      char e = s[++i];
      switch (e) {
        case 'b': cp = 8; break; case 'f': cp = 12; break; case 'n': cp = 10; break;
        case 'r': cp = 13; break; case 't': cp = 9; break;
        case 'u': {
          if (i + 4 >= n) return -1;
          cp = 0;
          for (int k = 1; k <= 4; ++k) {
            char h = s[i + k]; unsigned v;
            if (h >= '0' && h <= '9') v = h - '0'; else if (h >= 'a' && h <= 'f') v = h - 'a' + 10;
            else if (h >= 'A' && h <= 'F') v = h - 'A' + 10; else return -1;
            cp = cp * 16 + v;
          }
          i += 4;
          break;
        }
        default: cp = (unsigned char)e; break;      // \" \\ \/
      }
    }
    if (cp == '"') cp = 'x';                         // keeps the token malformed, keeps '"' a terminator
    char buf[3]; int len;
    if (cp < 0x80) { buf[0] = (char)cp; len = 1; }
    else if (cp < 0x800) { buf[0] = (char)(0xC0 | (cp >> 6)); buf[1] = (char)(0x80 | (cp & 0x3F)); len = 2; }
    else { buf[0] = (char)(0xE0 | (cp >> 12)); buf[1] = (char)(0x80 | ((cp >> 6) & 0x3F)); buf[2] = (char)(0x80 | (cp & 0x3F)); len = 3; }
    if (o + len > cap) return -1;
    memcpy(out + o, buf, len); o += len;
  }
  return (long)o;
}

struct Out {
  uint32_t* ts_off; uint32_t* val_off; int64_t cap_samples;
  int64_t* series_off; uint32_t* metric_off; uint32_t* metric_len; int64_t cap_series;
  char* extra; size_t cap_extra; size_t body_len;
  int64_t n_samples, n_series; size_t extra_len;
};

// at '[' of one sample pair: [ ts , "value" ]
bool sample_pair(Cur& c, Out& o) {
  if (!expect(c, '[')) return c.fail("expected '[' of sample");
  ws(c);
  if (o.n_samples >= o.cap_samples) { pmx_fail(PMX_E_CAPACITY, "pmx_tokenize: cap_samples too small"); return false; }
  const char* t0 = c.p;
  while (c.p < c.end && *c.p != ',' && *c.p != ']') ++c.p;
  if (c.p >= c.end || *c.p != ',' || c.p == t0) return c.fail("expected timestamp");
  ++c.p;
  const char* s; size_t n; bool esc;
  if (!string_span(c, &s, &n, &esc)) return false;
  uint64_t voff = (uint64_t)(s - c.base);
  if (esc) {
    if (o.extra_len + n + 1 > o.cap_extra) { pmx_fail(PMX_E_CAPACITY, "pmx_tokenize: cap_extra too small"); return false; }
    long w = decode_escapes(s, n, o.extra + o.extra_len, o.cap_extra - o.extra_len - 1);
    if (w < 0) return c.fail("bad escape in value string");
    voff = o.body_len + o.extra_len;
    o.extra_len += (size_t)w;
    o.extra[o.extra_len++] = '"';
  }
  if (voff > 0xffffffffULL) { pmx_fail(PMX_E_LIMIT, "pmx_tokenize: body larger than 4 GiB; split the query range"); return false; }
  o.ts_off[o.n_samples] = (uint32_t)(t0 - c.base);
  o.val_off[o.n_samples] = (uint32_t)voff;
  ++o.n_samples;
  if (!expect(c, ']')) return c.fail("expected ']' of sample");
  return true;
}

bool series_object(Cur& c, Out& o) {
  if (!expect(c, '{')) return c.fail("expected series object");
  if (o.n_series >= o.cap_series) { pmx_fail(PMX_E_CAPACITY, "pmx_tokenize: cap_series too small"); return false; }
  const int64_t sidx = o.n_series;
  o.metric_off[sidx] = 0; o.metric_len[sidx] = 0;
  ws(c);
  if (c.p < c.end && *c.p == '}') { ++c.p; o.series_off[++o.n_series] = o.n_samples; return true; }
  for (;;) {
    const char* k; size_t kn; bool e;
    if (!string_span(c, &k, &kn, &e)) return false;
    if (!expect(c, ':')) return c.fail("expected ':'");
    ws(c);
    if (key_is(k, kn, "metric")) {
      const char* m0 = c.p;
      if (!skip_value(c, 0)) return false;
      o.metric_off[sidx] = (uint32_t)(m0 - c.base);
      o.metric_len[sidx] = (uint32_t)(c.p - m0);
    } else if (key_is(k, kn, "values")) {
      if (!expect(c, '[')) return c.fail("expected '[' of values");
      ws(c);
      if (c.p < c.end && *c.p == ']') { ++c.p; }
      else for (;;) {
        if (!sample_pair(c, o)) return false;
        ws(c);
        if (c.p < c.end && *c.p == ',') { ++c.p; continue; }
        if (c.p < c.end && *c.p == ']') { ++c.p; break; }
        return c.fail("expected ',' or ']' in values");
      }
    } else if (key_is(k, kn, "value")) {
      if (!sample_pair(c, o)) return false;
    } else {
      if (!skip_value(c, 0)) return false;
    }
    ws(c);
    if (c.p < c.end && *c.p == ',') { ++c.p; continue; }
    if (c.p < c.end && *c.p == '}') { ++c.p; break; }
    return c.fail("expected ',' or '}' in series");
  }
  o.series_off[++o.n_series] = o.n_samples;
  return true;
}

}  // namespace

extern "C" int pmx_tokenize(const char* h_body, size_t len, uint32_t* h_ts_off, uint32_t* h_val_off,
                            int64_t cap_samples, int64_t* h_series_off, uint32_t* h_metric_off,
                            uint32_t* h_metric_len, int64_t cap_series, char* h_extra, size_t cap_extra,
                            pmx_tokenize_result* res) {
  if (!h_body || !res || !h_series_off || cap_series < 0 || cap_samples < 0)
    return pmx_fail(PMX_E_INVALID, "pmx_tokenize: NULL argument");
  if (len > 0xffffffffULL) return pmx_fail(PMX_E_LIMIT, "pmx_tokenize: body larger than 4 GiB; split the query range");
  Cur c{h_body, h_body, h_body + len};
  Out o{h_ts_off, h_val_off, cap_samples, h_series_off, h_metric_off, h_metric_len, cap_series,
        h_extra, cap_extra, len, 0, 0, 0};
  h_series_off[0] = 0;
  bool status_ok = false, have_result = false;
  int rtype = -1;
  pmx_fail(PMX_E_JSON, "pmx_tokenize: malformed body");
  if (!expect(c, '{')) { c.fail("expected top-level object"); return PMX_E_JSON; }
  ws(c);
  if (c.p < c.end && *c.p == '}') ++c.p;
  else for (;;) {
    const char* k; size_t kn; bool e;
    if (!string_span(c, &k, &kn, &e)) return PMX_E_JSON;
    if (!expect(c, ':')) { c.fail("expected ':'"); return PMX_E_JSON; }
    if (key_is(k, kn, "status")) {
      const char* s; size_t n;
      if (!string_span(c, &s, &n, &e)) return PMX_E_JSON;
      status_ok = key_is(s, n, "success");
    } else if (key_is(k, kn, "data")) {
      if (!expect(c, '{')) { c.fail("expected data object"); return PMX_E_JSON; }
      ws(c);
      if (c.p < c.end && *c.p == '}') ++c.p;
      else for (;;) {
        const char* dk; size_t dkn;
        if (!string_span(c, &dk, &dkn, &e)) return PMX_E_JSON;
        if (!expect(c, ':')) { c.fail("expected ':'"); return PMX_E_JSON; }
        if (key_is(dk, dkn, "resultType")) {
          const char* s; size_t n;
          if (!string_span(c, &s, &n, &e)) return PMX_E_JSON;
          rtype = key_is(s, n, "matrix") ? 0 : key_is(s, n, "vector") ? 1 : 2;
        } else if (key_is(dk, dkn, "result")) {
          if (!expect(c, '[')) { c.fail("expected result array"); return PMX_E_JSON; }
          have_result = true;
          ws(c);
          if (c.p < c.end && *c.p == ']') ++c.p;
          else for (;;) {
            if (!series_object(c, o)) {
              // series_object already set the error (JSON, CAPACITY or LIMIT)
              const char* m = pmx_last_error();
              if (strstr(m, "cap_")) return PMX_E_CAPACITY;
              if (strstr(m, "4 GiB")) return PMX_E_LIMIT;
              return PMX_E_JSON;
            }
            ws(c);
            if (c.p < c.end && *c.p == ',') { ++c.p; continue; }
            if (c.p < c.end && *c.p == ']') { ++c.p; break; }
            c.fail("expected ',' or ']' in result");
            return PMX_E_JSON;
          }
        } else if (!skip_value(c, 0)) return PMX_E_JSON;
        ws(c);
        if (c.p < c.end && *c.p == ',') { ++c.p; continue; }
        if (c.p < c.end && *c.p == '}') { ++c.p; break; }
        c.fail("expected ',' or '}' in data");
        return PMX_E_JSON;
      }
    } else if (!skip_value(c, 0)) return PMX_E_JSON;
    ws(c);
    if (c.p < c.end && *c.p == ',') { ++c.p; continue; }
    if (c.p < c.end && *c.p == '}') { ++c.p; break; }
    c.fail("expected ',' or '}' at top level");
    return PMX_E_JSON;
  }
  ws(c);
  if (c.p != c.end) { c.fail("trailing bytes after the response object"); return PMX_E_JSON; }
  if (!status_ok) return pmx_fail(PMX_E_JSON, "pmx_tokenize: response status is not \"success\"");
  if (!have_result) return pmx_fail(PMX_E_KEY, "pmx_tokenize: response has no data.result");
  if (rtype != 0 && rtype != 1) return pmx_fail(PMX_E_JSON, "pmx_tokenize: resultType must be matrix or vector");
  res->n_samples = o.n_samples;
  res->n_series = o.n_series;
  res->extra_len = (int64_t)o.extra_len;
  res->result_type = rtype;
  return PMX_OK;
}
