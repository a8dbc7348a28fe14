// fm_pool.cu — one process, several GPUs: the single-process fan-out of SURVEY.md 8(e).
//
// A pool owns one fm_ctx per device.  fm_pool_score_batch cuts a batch into one contiguous range of
// series per ctx (at service boundaries when a roll-up is asked for, so that every API and every
// service is scored on exactly one GPU), drives every ctx from its own host thread through the
// ordinary entry points (fm_score_batch / fm_score_batch_cached: chunked, three streams each) and
// lets them write into disjoint slices of the caller's output arrays — that is the whole "gather".
// No collective, no peer access: series are independent.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "fm_internal.h"

struct fm_pool {
  std::vector<fm_ctx*> ctx;
  std::vector<int32_t> device;
  std::string err;
};

namespace {
int32_t pool_fail(fm_pool* p, int32_t code, const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
  if (p) p->err = buf; else fm_set_thread_error("%s", buf);
  return code;
}
}  // namespace

extern "C" {

int32_t fm_pool_create(const int32_t* devices, int32_t n_devices, fm_pool** out) {
  if (!out) return FM_E_INVALID;
  *out = nullptr;
  if (n_devices < 0 || n_devices > 64 || (n_devices > 0 && !devices)) return pool_fail(nullptr, FM_E_INVALID, "fm_pool_create: bad device list");
  std::vector<int32_t> dev(devices, devices + n_devices);
  if (dev.empty()) {                                   // every visible GPU
    int32_t n = 0;
    fm_device_count(&n);
    for (int32_t i = 0; i < n; ++i) dev.push_back(i);
    if (dev.empty()) return pool_fail(nullptr, FM_E_NODEVICE, "fm_pool_create: no CUDA device available (no CPU fallback)");
  }
  fm_pool* p = new (std::nothrow) fm_pool();
  if (!p) return FM_E_NOMEM;
  for (int32_t d : dev) {
    fm_ctx* c = nullptr;
    const int32_t rc = fm_ctx_create(d, &c);
    if (rc != FM_OK) {
      const std::string why = fm_last_error(nullptr);
      for (fm_ctx* x : p->ctx) fm_ctx_destroy(x);
      delete p;
      return pool_fail(nullptr, rc, "fm_pool_create: device %d: %s", d, why.c_str());
    }
    p->ctx.push_back(c);
    p->device.push_back(d);
  }
  *out = p;
  return FM_OK;
}

int32_t fm_pool_destroy(fm_pool* p) {
  if (!p) return FM_OK;
  for (fm_ctx* c : p->ctx) fm_ctx_destroy(c);
  delete p;
  return FM_OK;
}

int32_t fm_pool_size(fm_pool* p) { return p ? (int32_t)p->ctx.size() : 0; }

fm_ctx* fm_pool_ctx(fm_pool* p, int32_t i) { return (p && i >= 0 && (size_t)i < p->ctx.size()) ? p->ctx[i] : nullptr; }

const char* fm_pool_last_error(fm_pool* p) { return p ? p->err.c_str() : fm_last_error(nullptr); }

int32_t fm_pool_set_params(fm_pool* p, const fm_global_params* g, const fm_type_params* table, uint32_t n_types) {
  if (!p) return FM_E_INVALID;
  for (size_t i = 0; i < p->ctx.size(); ++i) {
    const int32_t rc = fm_set_params(p->ctx[i], g, table, n_types);
    if (rc != FM_OK) return pool_fail(p, rc, "device %d: %s", p->device[i], fm_last_error(p->ctx[i]));
  }
  return FM_OK;
}

int32_t fm_pool_cache_configure(fm_pool* p, uint64_t max_bytes_per_device) {
  if (!p) return FM_E_INVALID;
  for (size_t i = 0; i < p->ctx.size(); ++i) {
    const int32_t rc = fm_cache_configure(p->ctx[i], max_bytes_per_device);
    if (rc != FM_OK) return pool_fail(p, rc, "device %d: %s", p->device[i], fm_last_error(p->ctx[i]));
  }
  return FM_OK;
}

int32_t fm_pool_score_batch(fm_pool* p, const fm_batch* b, const uint64_t* row_ids, void* out, fm_rank_aux* out_aux,
                            fm_rollup_rec* out_api, fm_rollup_rec* out_svc, fm_timing* timing_per_device) {
  if (!p) return FM_E_INVALID;
  if (!b || b->struct_size != sizeof(fm_batch)) return pool_fail(p, FM_E_INVALID, "fm_pool_score_batch: bad fm_batch");
  if (!out && b->n_series) return pool_fail(p, FM_E_INVALID, "fm_pool_score_batch: null output");
  const size_t G = p->ctx.size();
  const uint64_t S = b->n_series;
  if (timing_per_device) memset(timing_per_device, 0, G * sizeof(fm_timing));
  if (S == 0) return FM_OK;
  const bool roll = b->api_id && out_api && b->n_apis;
  const bool roll_svc = roll && out_svc && b->api_service && b->n_services;
  const size_t rsz = b->record_kind == FM_REC_COMPACT ? sizeof(fm_compact_rec) : sizeof(fm_series_rec);

  // cut points: equal shares of the series, moved forward to the next service (else API) boundary
  std::vector<uint64_t> cut(G + 1, S);
  cut[0] = 0;
  for (size_t g = 1; g < G; ++g) {
    uint64_t s = std::max(cut[g - 1], S * g / G);
    if (roll) {
      auto key = [&](uint64_t i) { return roll_svc ? b->api_service[b->api_id[i]] : b->api_id[i]; };
      while (s > 0 && s < S && key(s) == key(s - 1)) ++s;
    }
    cut[g] = std::min(s, S);
  }
  if (roll) {
    for (uint64_t i = 0; i < S; ++i)
      if (b->api_id[i] >= b->n_apis || (i && b->api_id[i] < b->api_id[i - 1]))
        return pool_fail(p, FM_E_INVALID, "fm_pool_score_batch: api_id must be non-decreasing and < n_apis");
  }

  std::vector<int32_t> rcs(G, FM_OK);
  std::vector<std::vector<uint32_t>> api_local(G), svc_local(G);
  auto work = [&](size_t g) {
    const uint64_t s0 = cut[g], n = cut[g + 1] - s0;
    if (n == 0) return;
    fm_batch sb = *b;
    sb.n_series = n;
    sb.cur = b->cur + s0 * b->cur_pitch;
    if (b->base) sb.base = b->base + s0 * b->base_pitch;
    if (b->hist) sb.hist = b->hist + s0 * b->hist_pitch;
    if (b->cur_len) sb.cur_len = b->cur_len + s0;
    if (b->base_len) sb.base_len = b->base_len + s0;
    if (b->hist_len) sb.hist_len = b->hist_len + s0;
    if (b->metric_type) sb.metric_type = b->metric_type + s0;
    fm_rollup_rec *oa = nullptr, *os = nullptr;
    if (roll) {
      // this shard's APIs (and services) are a contiguous range: rebase the ids to it
      const uint32_t a_lo = b->api_id[s0], a_hi = b->api_id[s0 + n - 1] + 1u;
      api_local[g].resize(n);
      for (uint64_t i = 0; i < n; ++i) api_local[g][i] = b->api_id[s0 + i] - a_lo;
      sb.api_id = api_local[g].data();
      sb.n_apis = a_hi - a_lo;
      oa = out_api + a_lo;
      if (roll_svc) {
        const uint32_t v_lo = b->api_service[a_lo], v_hi = b->api_service[a_hi - 1] + 1u;
        svc_local[g].resize(a_hi - a_lo);
        for (uint32_t i = 0; i < a_hi - a_lo; ++i) svc_local[g][i] = b->api_service[a_lo + i] - v_lo;
        sb.api_service = svc_local[g].data();
        sb.n_services = v_hi - v_lo;
        os = out_svc + v_lo;
      } else {
        sb.api_service = nullptr; sb.n_services = 0;
      }
    } else {
      sb.api_id = nullptr; sb.api_service = nullptr; sb.n_apis = sb.n_services = 0;
    }
    void* o = static_cast<char*>(out) + s0 * rsz;
    fm_rank_aux* oaux = out_aux ? out_aux + s0 : nullptr;
    fm_timing* t = timing_per_device ? timing_per_device + g : nullptr;
    rcs[g] = row_ids ? fm_score_batch_cached(p->ctx[g], &sb, row_ids + s0, o, oaux, oa, os, t)
                     : fm_score_batch(p->ctx[g], &sb, o, oaux, oa, os, t);
  };
  std::vector<std::thread> th;
  for (size_t g = 1; g < G; ++g) th.emplace_back(work, g);
  work(0);                                               // the calling thread drives the first ctx
  for (auto& t : th) t.join();
  for (size_t g = 0; g < G; ++g)
    if (rcs[g] != FM_OK) return pool_fail(p, rcs[g], "device %d: %s", p->device[g], fm_last_error(p->ctx[g]));
  return FM_OK;
}

}  // extern "C"
