// fm_decode.cu — host-only: Prometheus query_range response body -> fp32 rows (the
// step right before the hot path: foremast-service hands the brain URLs, the brain fetches
// {"status":"success","data":{"resultType":"matrix","result":[{"metric":{..},"values":[[ts,"v"],..]},..]}}
// (shape: foremast-browser/src/App.js:278-292) and must turn the quoted decimal strings into rows).
// A single forward scan, no DOM, no allocation: keys are matched where they can occur, everything
// else is skipped structurally (strings with escapes, nested objects / arrays).  Numbers go through
// std::from_chars (correctly rounded, locale-free), so a sample equals Python's
// np.float32(float(v)) bit for bit; NaN / +-Inf samples (Prometheus gaps) are dropped.
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "../../include/foremast_cuda.h"
#include "fm_internal.h"

namespace {

struct Scan {
  const char* p;
  const char* end;
  bool fail;
  void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
  bool eat(char c) { ws(); if (p < end && *p == c) { ++p; return true; } return false; }
  bool peek(char c) { ws(); return p < end && *p == c; }
  // string token: returns [b, e) of the raw contents (escapes not decoded)
  bool str(const char*& b, const char*& e) {
    ws();
    if (p >= end || *p != '"') return fail = true, false;
    b = ++p;
    while (p < end && *p != '"') { if (*p == '\\') ++p; ++p; }
    if (p >= end) return fail = true, false;
    e = p++;
    return true;
  }
  void skip_value() {
    ws();
    if (p >= end) { fail = true; return; }
    if (*p == '"') { const char *b, *e; str(b, e); return; }
    if (*p == '{' || *p == '[') {
      int depth = 0;
      while (p < end) {
        const char c = *p;
        if (c == '"') { const char *b, *e; if (!str(b, e)) return; continue; }
        if (c == '{' || c == '[') ++depth;
        else if (c == '}' || c == ']') { if (--depth == 0) { ++p; return; } }
        ++p;
      }
      fail = true;
      return;
    }
    while (p < end && *p != ',' && *p != '}' && *p != ']' && *p != ' ' && *p != '\n' && *p != '\t' && *p != '\r') ++p;
  }
  bool number(double& v) {                    // bare JSON number
    ws();
    const char* b = p;
    while (p < end && (*p == '-' || *p == '+' || *p == '.' || *p == 'e' || *p == 'E' || (*p >= '0' && *p <= '9'))) ++p;
    if (b == p) return fail = true, false;
    if (*b == '+') ++b;
    const auto r = std::from_chars(b, p, v);
    if (r.ec != std::errc() || r.ptr != p) return fail = true, false;
    return true;
  }
};

bool key_is(const char* b, const char* e, const char* k) {
  const size_t n = strlen(k);
  return (size_t)(e - b) == n && memcmp(b, k, n) == 0;
}

// "1.25", "-3e-5", "NaN", "+Inf", "-Inf": returns false for the non-finite spellings
bool sample_value(const char* b, const char* e, double& v, bool& bad) {
  if (b < e && *b == '+') ++b;
  if (b == e) { bad = true; return false; }
  const char c = (*b == '-' && b + 1 < e) ? b[1] : *b;
  if (c == 'N' || c == 'n' || c == 'I' || c == 'i') return false;
  const auto r = std::from_chars(b, e, v);
  if (r.ec == std::errc::result_out_of_range && r.ptr == e) {
    // over- or underflow: let strtod decide between +-inf (a gap) and a denormal / zero (a sample)
    char tmp[64];
    const size_t n = (size_t)(e - b) < sizeof(tmp) - 1 ? (size_t)(e - b) : sizeof(tmp) - 1;
    memcpy(tmp, b, n); tmp[n] = 0;
    v = strtod(tmp, nullptr);
    return std::isfinite(v);
  }
  if (r.ec != std::errc() || r.ptr != e) { bad = true; return false; }
  return std::isfinite(v);
}

// one result object: {"metric":{..},"values":[[ts,"v"],..]}  (or "value":[ts,"v"] of an instant query)
int64_t result_series(Scan& sc, int64_t* ts_out, float* val_out, uint64_t cap, uint64_t at) {
  int64_t n = 0;
  if (!sc.eat('{')) { sc.fail = true; return -1; }
  if (sc.eat('}')) return 0;
  do {
    const char *kb, *ke;
    if (!sc.str(kb, ke) || !sc.eat(':')) { sc.fail = true; return -1; }
    const bool many = key_is(kb, ke, "values"), one = key_is(kb, ke, "value");
    if (!many && !one) { sc.skip_value(); if (sc.fail) return -1; continue; }
    if (many && !sc.eat('[')) { sc.fail = true; return -1; }
    if (many && sc.eat(']')) continue;
    do {
      double t = 0.0, v = 0.0;
      if (!sc.eat('[') || !sc.number(t) || !sc.eat(',')) { sc.fail = true; return -1; }
      bool keep, bad = false;
      if (sc.peek('"')) {
        const char *vb, *ve;
        if (!sc.str(vb, ve)) return -1;
        keep = sample_value(vb, ve, v, bad);
      } else {
        keep = sc.number(v) && std::isfinite(v);
      }
      if (bad || sc.fail || !sc.eat(']')) { sc.fail = true; return -1; }
      if (keep) {
        if (at + (uint64_t)n >= cap) return -2;
        ts_out[at + n] = (int64_t)t;
        val_out[at + n] = (float)v;
        ++n;
      }
    } while (many && sc.eat(','));
    if (many && !sc.eat(']')) { sc.fail = true; return -1; }
  } while (sc.eat(','));
  if (!sc.eat('}')) { sc.fail = true; return -1; }
  return n;
}

}  // namespace

extern "C" int64_t fm_decode_query_range(const char* body, uint64_t len, int64_t* ts_out, float* val_out,
                                         uint64_t cap_samples, uint32_t* series_start, uint32_t cap_series,
                                         uint32_t* n_series) {
  if (!body || !n_series || (cap_samples && (!ts_out || !val_out)) || !series_start || cap_series == 0) {
    fm_set_thread_error("fm_decode_query_range: null argument");
    return FM_E_INVALID;
  }
  Scan sc{body, body + len, false};
  uint64_t total = 0;
  uint32_t ns = 0;
  bool seen_result = false;
  series_start[0] = 0;
  if (!sc.eat('{')) { fm_set_thread_error("query_range body is not a JSON object"); return FM_E_INVALID; }
  if (!sc.eat('}')) {
    do {
      const char *kb, *ke;
      if (!sc.str(kb, ke) || !sc.eat(':')) { sc.fail = true; break; }
      if (key_is(kb, ke, "status")) {
        const char *vb, *ve;
        if (!sc.str(vb, ve)) break;
        if (!key_is(vb, ve, "success")) {
          fm_set_thread_error("query_range failed: status %.*s", (int)(ve - vb), vb);
          return FM_E_INVALID;
        }
      } else if (key_is(kb, ke, "data")) {
        if (!sc.eat('{')) { sc.fail = true; break; }
        if (sc.eat('}')) continue;
        do {
          const char *db, *de;
          if (!sc.str(db, de) || !sc.eat(':')) { sc.fail = true; break; }
          if (!key_is(db, de, "result")) { sc.skip_value(); continue; }
          seen_result = true;
          if (!sc.eat('[')) { sc.fail = true; break; }
          if (sc.eat(']')) continue;
          do {
            if (ns + 1u >= cap_series) { fm_set_thread_error("fm_decode_query_range: more than %u series", cap_series - 1u); return FM_E_NOMEM; }
            const int64_t n = result_series(sc, ts_out, val_out, cap_samples, total);
            if (n == -2) { fm_set_thread_error("fm_decode_query_range: more than %llu samples", (unsigned long long)cap_samples); return FM_E_NOMEM; }
            if (n < 0) { sc.fail = true; break; }
            total += (uint64_t)n;
            series_start[++ns] = (uint32_t)total;
          } while (sc.eat(','));
          if (sc.fail || !sc.eat(']')) { sc.fail = true; break; }
        } while (sc.eat(','));
        if (sc.fail || !sc.eat('}')) { sc.fail = true; break; }
      } else {
        sc.skip_value();
      }
    } while (!sc.fail && sc.eat(','));
    if (sc.fail || !sc.eat('}')) {
      fm_set_thread_error("malformed query_range body near byte %llu", (unsigned long long)(sc.p - body));
      return FM_E_INVALID;
    }
  }
  (void)seen_result;
  *n_series = ns;
  return (int64_t)total;
}
