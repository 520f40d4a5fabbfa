// tokenize.cu -- X1: Prometheus response body -> per-sample token offsets (HOST code).
//
// Takes over json.loads at /root/reference/app.py:163,178 for the hot part of the body: it
// walks the JSON once, records where each sample's timestamp and value token start, the CSR
// sample offsets per series and the byte span of each series' "metric" object, and never
// materialises objects.  Numbers are decoded later on the GPU (parse.cu).
#include <string.h>
#include <thread>
#include <vector>
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
  // multi-threaded driver: a chunk first only counts, then fills its slice of the arrays
  bool count_only = false;      // no array writes; escaped tokens reserve raw length + 1 bytes of extra
  bool reserve = false;         // fill pass of the same: keep the reserved stride so offsets match
  int64_t sample_base = 0;      // added to the CSR sample offsets written by this chunk
};

// at '[' of one sample pair: [ ts , "value" ]
bool sample_pair(Cur& c, Out& o) {
  if (!expect(c, '[')) return c.fail("expected '[' of sample");
  ws(c);
  if (!o.count_only && o.n_samples >= o.cap_samples) { pmx_fail(PMX_E_CAPACITY, "pmx_tokenize: cap_samples too small"); return false; }
  const char* t0 = c.p;
  while (c.p < c.end && *c.p != ',' && *c.p != ']') ++c.p;
  if (c.p >= c.end || *c.p != ',' || c.p == t0) return c.fail("expected timestamp");
  ++c.p;
  const char* s; size_t n; bool esc;
  if (!string_span(c, &s, &n, &esc)) return false;
  uint64_t voff = (uint64_t)(s - c.base);
  if (esc && o.count_only) {
    o.extra_len += n + 1;
  } else if (esc) {
    if (o.extra_len + n + 1 > o.cap_extra) { pmx_fail(PMX_E_CAPACITY, "pmx_tokenize: cap_extra too small"); return false; }
    long w = decode_escapes(s, n, o.extra + o.extra_len, o.cap_extra - o.extra_len - 1);
    if (w < 0) return c.fail("bad escape in value string");
    voff = o.body_len + o.extra_len;
    o.extra[o.extra_len + (size_t)w] = '"';
    o.extra_len += o.reserve ? n + 1 : (size_t)w + 1;
  }
  if (voff > 0xffffffffULL) { pmx_fail(PMX_E_LIMIT, "pmx_tokenize: body larger than 4 GiB; split the query range"); return false; }
  if (!o.count_only) {
    o.ts_off[o.n_samples] = (uint32_t)(t0 - c.base);
    o.val_off[o.n_samples] = (uint32_t)voff;
  }
  ++o.n_samples;
  if (!expect(c, ']')) return c.fail("expected ']' of sample");
  return true;
}

bool series_object(Cur& c, Out& o) {
  if (!expect(c, '{')) return c.fail("expected series object");
  if (!o.count_only && o.n_series >= o.cap_series) { pmx_fail(PMX_E_CAPACITY, "pmx_tokenize: cap_series too small"); return false; }
  const int64_t sidx = o.n_series;
  if (!o.count_only) { o.metric_off[sidx] = 0; o.metric_len[sidx] = 0; }
  ws(c);
  if (c.p < c.end && *c.p == '}') {
    ++c.p; ++o.n_series;
    if (!o.count_only) o.series_off[o.n_series] = o.n_samples + o.sample_base;
    return true;
  }
  for (;;) {
    const char* k; size_t kn; bool e;
    if (!string_span(c, &k, &kn, &e)) return false;
    if (!expect(c, ':')) return c.fail("expected ':'");
    ws(c);
    if (key_is(k, kn, "metric")) {
      const char* m0 = c.p;
      if (!skip_value(c, 0)) return false;
      if (!o.count_only) {
        o.metric_off[sidx] = (uint32_t)(m0 - c.base);
        o.metric_len[sidx] = (uint32_t)(c.p - m0);
      }
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
  ++o.n_series;
  if (!o.count_only) o.series_off[o.n_series] = o.n_samples + o.sample_base;
  return true;
}


// ---------------------------------------------------------------------------- multi-threaded driver
// A chunk is a run of whole series objects "{...},{...}," of the result array.  parse_chunk walks
// it with the ordinary series parser; `last` chunks stop at the ']' that closes the array.
struct Chunk {
  const char* begin; const char* end; bool last;
  int64_t n_samples = 0, n_series = 0; size_t extra_len = 0;
  const char* stop = nullptr;          // where the walk ended
  int rc = PMX_OK; char msg[200] = "";
};

void parse_chunk(const char* base, const char* body_end, Chunk& ch, Out o) {
  Cur c{base, ch.begin, ch.last ? body_end : ch.end};
  o.n_samples = 0; o.n_series = 0; o.extra_len = 0;
  bool ok = true;
  for (;;) {
    ws(c);
    if (c.p >= c.end) { ok = !ch.last; if (!ok) c.fail("unterminated result array"); break; }
    if (*c.p == ']' && ch.last) break;
    if (!series_object(c, o)) { ok = false; break; }
    ws(c);
    if (c.p < c.end && *c.p == ',') { ++c.p; continue; }
    if (c.p < c.end && *c.p == ']' && ch.last) break;
    if (c.p >= c.end && !ch.last) break;
    ok = false; c.fail("expected ',' or ']' in result");
    break;
  }
  ch.n_samples = o.n_samples; ch.n_series = o.n_series; ch.extra_len = o.extra_len; ch.stop = c.p;
  if (!ok) {
    const char* m = pmx_last_error();                       // thread-local message of this worker
    ch.rc = strstr(m, "cap_") ? PMX_E_CAPACITY : strstr(m, "4 GiB") ? PMX_E_LIMIT : PMX_E_JSON;
    strncpy(ch.msg, m, sizeof(ch.msg) - 1);
  }
}

// the rest of the response after the result array: "}" [, "key": value ...] "}"
bool parse_tail(Cur& c) {
  for (int level = 0; level < 2; ++level) {
    for (;;) {
      ws(c);
      if (c.p < c.end && *c.p == '}') { ++c.p; break; }
      if (c.p < c.end && *c.p == ',') {
        ++c.p;
        const char* k; size_t kn; bool e;
        if (!string_span(c, &k, &kn, &e)) return false;
        if (!expect(c, ':')) return c.fail("expected ':'");
        if (!skip_value(c, 0)) return false;
        continue;
      }
      return c.fail("expected ',' or '}' after the result array");
    }
  }
  ws(c);
  if (c.p != c.end) return c.fail("trailing bytes after the response object");
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

// Same result as pmx_tokenize, with the result array cut at series boundaries and walked by
// n_threads host threads (count pass, prefix sums, fill pass).  Takes the canonical Prometheus
// envelope {"status":"success","data":{"resultType":"matrix|vector","result":[ ... ; anything else,
// or a body that does not split cleanly, goes through pmx_tokenize.  Decoded escaped value tokens
// keep their raw length in h_extra (gaps are never referenced).
extern "C" int pmx_tokenize_mt(const char* h_body, size_t len, uint32_t* h_ts_off, uint32_t* h_val_off,
                               int64_t cap_samples, int64_t* h_series_off, uint32_t* h_metric_off,
                               uint32_t* h_metric_len, int64_t cap_series, char* h_extra, size_t cap_extra,
                               int n_threads, pmx_tokenize_result* res) {
  auto single = [&]() {
    return pmx_tokenize(h_body, len, h_ts_off, h_val_off, cap_samples, h_series_off, h_metric_off, h_metric_len,
                        cap_series, h_extra, cap_extra, res);
  };
  if (!h_body || !res || n_threads < 2 || len < (size_t(1) << 20) || len > 0xffffffffULL) return single();
  static const char P0[] = "{\"status\":\"success\",\"data\":{\"resultType\":\"";
  static const char PM[] = "matrix\",\"result\":[";
  static const char PV[] = "vector\",\"result\":[";
  const size_t n0 = sizeof(P0) - 1, nm = sizeof(PM) - 1;
  if (len < n0 + nm + 3 || memcmp(h_body, P0, n0) != 0) return single();
  int rtype;
  if (memcmp(h_body + n0, PM, nm) == 0) rtype = 0;
  else if (memcmp(h_body + n0, PV, nm) == 0) rtype = 1;
  else return single();
  const char* r0 = h_body + n0 + nm;
  const char* end = h_body + len;
  // split points: the '{' of a series object that starts with its "metric" key.  The byte string
  // below cannot occur inside a JSON string (its quotes would be escaped there).
  static const char SEP[] = "},{\"metric\":";
  const size_t nsep = sizeof(SEP) - 1;
  std::vector<const char*> starts{r0};
  for (int k = 1; k < n_threads; ++k) {
    const char* from = r0 + (size_t)(end - r0) / n_threads * k;
    if (from <= starts.back()) continue;
    const char* hit = (const char*)memmem(from, (size_t)(end - from), SEP, nsep);
    if (!hit) break;
    if (hit + 2 > starts.back()) starts.push_back(hit + 2);
  }
  const int nc = (int)starts.size();
  if (nc < 2) return single();
  std::vector<Chunk> ch(nc);
  for (int k = 0; k < nc; ++k) { ch[k].begin = starts[k]; ch[k].end = k + 1 < nc ? starts[k + 1] : end; ch[k].last = k + 1 == nc; }
  Out proto{h_ts_off, h_val_off, cap_samples, h_series_off, h_metric_off, h_metric_len, cap_series,
            h_extra, cap_extra, len, 0, 0, 0};
  auto run = [&](bool count_only, const std::vector<Out>& outs) {
    std::vector<std::thread> th;
    for (int k = 0; k < nc; ++k) th.emplace_back([&, k]() { Out o = outs[k]; o.count_only = count_only; o.reserve = true;
                                                            parse_chunk(h_body, end, ch[k], o); });
    for (auto& t : th) t.join();
  };
  run(true, std::vector<Out>(nc, proto));
  for (int k = 0; k < nc; ++k)
    if (ch[k].rc != PMX_OK || (!ch[k].last && ch[k].stop != ch[k].end)) return single();   // let the plain walk report it
  // tail and totals
  Cur tail{h_body, ch[nc - 1].stop, end};
  if (!expect(tail, ']') || !parse_tail(tail)) return single();
  int64_t ns = 0, nr = 0; size_t ne = 0;
  std::vector<Out> outs(nc, proto);
  for (int k = 0; k < nc; ++k) {
    Out& o = outs[k];
    o.ts_off = h_ts_off + ns; o.val_off = h_val_off + ns; o.cap_samples = ch[k].n_samples;
    o.series_off = h_series_off + nr; o.metric_off = h_metric_off + nr; o.metric_len = h_metric_len + nr;
    o.cap_series = ch[k].n_series; o.sample_base = ns;
    o.extra = h_extra + ne; o.cap_extra = ch[k].extra_len + 1; o.body_len = len + ne;
    ns += ch[k].n_samples; nr += ch[k].n_series; ne += ch[k].extra_len;
  }
  if (ns > cap_samples || nr > cap_series || ne + 1 > cap_extra)
    return pmx_fail(PMX_E_CAPACITY, "pmx_tokenize_mt: output capacity too small (%lld samples, %lld series, %zu extra bytes)",
                    (long long)ns, (long long)nr, ne);
  if (len + ne > 0xffffffffULL) return pmx_fail(PMX_E_LIMIT, "pmx_tokenize: body larger than 4 GiB; split the query range");
  h_series_off[0] = 0;
  run(false, outs);
  for (int k = 0; k < nc; ++k) if (ch[k].rc != PMX_OK) return pmx_fail(ch[k].rc, "%s", ch[k].msg);
  res->n_samples = ns; res->n_series = nr; res->extra_len = (int64_t)ne; res->result_type = rtype;
  return PMX_OK;
}
