// labels.cu -- X1: the interned label table (HOST code).
//
// The reference keeps every series' labels as a Python dict that json.loads built
// (/root/reference/app.py:178, read at :186-194).  With a million series that is a million
// json.loads calls and ~5 million str objects, almost all of them equal.  This walks the "metric"
// objects whose byte spans the tokenizers recorded and interns every distinct label name and
// value once: a series becomes a short run of (name id, value id) pairs, and a column such as
// __name__ or gpu_id is one int32 array.
#include <string.h>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>
#include "pmx_common.cuh"

namespace {

inline int hexv(char c) {
  if (c >= '0' && c <= '9') return c - '0';
  if (c >= 'a' && c <= 'f') return c - 'a' + 10;
  if (c >= 'A' && c <= 'F') return c - 'A' + 10;
  return -1;
}
inline void put_utf8(std::string& o, uint32_t cp) {
  if (cp < 0x80) o.push_back((char)cp);
  else if (cp < 0x800) { o.push_back((char)(0xC0 | (cp >> 6))); o.push_back((char)(0x80 | (cp & 0x3F))); }
  else if (cp < 0x10000) { o.push_back((char)(0xE0 | (cp >> 12))); o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F))); }
  else { o.push_back((char)(0xF0 | (cp >> 18))); o.push_back((char)(0x80 | ((cp >> 12) & 0x3F))); o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F))); }
}
// [s, e) is the content of a JSON string with escapes; false on a malformed escape
bool unescape(const char* s, const char* e, std::string& o) {
  o.clear();
  while (s < e) {
    if (*s != '\\') { o.push_back(*s++); continue; }
    if (++s >= e) return false;
    switch (*s++) {
      case '"': o.push_back('"'); break; case '\\': o.push_back('\\'); break; case '/': o.push_back('/'); break;
      case 'b': o.push_back('\b'); break; case 'f': o.push_back('\f'); break; case 'n': o.push_back('\n'); break;
      case 'r': o.push_back('\r'); break; case 't': o.push_back('\t'); break;
      case 'u': {
        if (e - s < 4) return false;
        uint32_t cp = 0;
        for (int i = 0; i < 4; ++i) { const int h = hexv(s[i]); if (h < 0) return false; cp = cp * 16 + (uint32_t)h; }
        s += 4;
        if (cp >= 0xD800 && cp < 0xDC00 && e - s >= 6 && s[0] == '\\' && s[1] == 'u') {   // surrogate pair
          uint32_t lo = 0; bool ok = true;
          for (int i = 0; i < 4; ++i) { const int h = hexv(s[2 + i]); if (h < 0) { ok = false; break; } lo = lo * 16 + (uint32_t)h; }
          if (ok && lo >= 0xDC00 && lo < 0xE000) { cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); s += 6; }
        }
        put_utf8(o, cp);
        break;
      }
      default: return false;
    }
  }
  return true;
}

struct Interner {
  std::unordered_map<std::string_view, int32_t> ids;
  std::vector<std::string> own;            // decoded strings (escapes) need storage of their own
  char* out; size_t cap; size_t used = 0;
  int64_t* off; int64_t cap_n; int64_t n = 0;
  // returns the id of [s, s+len) (bytes that stay valid: the body, or a string in `own`), or -1 when full
  int32_t get(const char* s, size_t len) {
    std::string_view key(s, len);
    auto it = ids.find(key);
    if (it != ids.end()) return it->second;
    if (n >= cap_n || used + len > cap) return -1;
    memcpy(out + used, s, len);
    // the table's key must outlive the call: point it at the copy in the output buffer
    std::string_view stored(out + used, len);
    off[n] = (int64_t)used;
    used += len;
    off[n + 1] = (int64_t)used;
    ids.emplace(stored, (int32_t)n);
    return (int32_t)n++;
  }
};

}  // namespace

extern "C" int pmx_labels_intern(const char* h_body, size_t len, const uint32_t* h_metric_off, const uint32_t* h_metric_len,
                                 int64_t n_series, int64_t* h_pair_off, int32_t* h_key_id, int32_t* h_val_id,
                                 int64_t cap_pairs, char* h_strings, size_t cap_strings, int64_t* h_str_off,
                                 int64_t cap_distinct, pmx_labels_result* res) {
  if (!h_body || !h_pair_off || !h_strings || !h_str_off || !res || n_series < 0 || (n_series > 0 && (!h_metric_off || !h_metric_len)))
    return pmx_fail(PMX_E_INVALID, "pmx_labels_intern: NULL argument");
  Interner in; in.out = h_strings; in.cap = cap_strings; in.off = h_str_off; in.cap_n = cap_distinct;
  if (cap_distinct >= 1) h_str_off[0] = 0;
  std::string tmp;
  int64_t np = 0;
  for (int64_t s = 0; s < n_series; ++s) {
    h_pair_off[s] = np;
    const size_t o = h_metric_off[s], n = h_metric_len[s];
    if (n == 0) continue;
    if (o + n > len) return pmx_fail(PMX_E_INVALID, "pmx_labels_intern: metric span of series %lld outside the body", (long long)s);
    const char* p = h_body + o; const char* e = p + n;
    auto skip_ws = [&]() { while (p < e && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; };
    auto str = [&](int32_t* id) -> bool {                    // at an opening quote
      skip_ws();
      if (p >= e || *p != '"') return false;
      const char* b = ++p; bool esc = false;
      while (p < e && *p != '"') { if (*p == '\\') { esc = true; ++p; } ++p; }
      if (p >= e) return false;
      const char* q = p++;
      if (!esc) { *id = in.get(b, (size_t)(q - b)); return true; }
      if (!unescape(b, q, tmp)) return false;
      *id = in.get(tmp.data(), tmp.size());
      return true;
    };
    skip_ws();
    if (p >= e || *p != '{') return pmx_fail(PMX_E_JSON, "pmx_labels_intern: series %lld: labels are not an object", (long long)s);
    ++p; skip_ws();
    if (p < e && *p == '}') continue;
    for (;;) {
      int32_t k = -2, v = -2;
      if (!str(&k)) return pmx_fail(PMX_E_JSON, "pmx_labels_intern: series %lld: malformed label name", (long long)s);
      skip_ws();
      if (p >= e || *p != ':') return pmx_fail(PMX_E_JSON, "pmx_labels_intern: series %lld: ':' expected", (long long)s);
      ++p;
      if (!str(&v)) return pmx_fail(PMX_E_JSON, "pmx_labels_intern: series %lld: label values must be strings", (long long)s);
      if (k == -1 || v == -1) return pmx_fail(PMX_E_CAPACITY, "pmx_labels_intern: string table full (%lld strings, %zu bytes)", (long long)in.n, in.used);
      if (np >= cap_pairs) return pmx_fail(PMX_E_CAPACITY, "pmx_labels_intern: cap_pairs too small");
      h_key_id[np] = k; h_val_id[np] = v; ++np;
      skip_ws();
      if (p < e && *p == ',') { ++p; continue; }
      if (p < e && *p == '}') break;
      return pmx_fail(PMX_E_JSON, "pmx_labels_intern: series %lld: ',' or '}' expected", (long long)s);
    }
  }
  h_pair_off[n_series] = np;
  res->n_pairs = np; res->n_strings = in.n; res->strings_len = (int64_t)in.used;
  return PMX_OK;
}
