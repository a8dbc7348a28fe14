// fm_api.cu — host side of the C ABI (include/foremast_cuda.h): contexts, device
// arenas, launch planning, and the chunked host->device pipeline.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/fm_synth.h"
#include "fm_internal.h"

namespace {

struct Arena {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    const size_t want = (bytes + 255) & ~size_t(255);
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

thread_local std::string g_create_error;

}  // namespace

// Device-resident baseline / history rows between calls (MAX_CACHE_SIZE, "max cached model size",
// foremast-brain/README.md:30).  A job is re-judged on every poll until its endTime
// (docs/guides/design.md:43) and its baseline and 7-day history do not change in between, so
// fm_score_batch_cached keeps those rows in HBM under a caller-supplied 64-bit row id and uploads
// only rows it does not hold (plus every current row).  One layout (n_base, pitch, n_hist, pitch) at
// a time; a batch of another layout starts the cache afresh.  Slots are recycled least-recently-
// scored first (a clock hand over the per-slot epoch of the last call that used it).
// chunk buffers in flight per ctx: upload of chunk i+2, kernels of chunk i+1, read-back of chunk i
#ifndef FM_PIPE_SLOTS
#define FM_PIPE_SLOTS 2
#endif
constexpr int kPipe = FM_PIPE_SLOTS;

struct RowCache {
  static constexpr uint32_t kNone = 0xFFFFFFFFu;
  uint64_t max_bytes = 0;
  uint32_t n_base = 0, base_pitch = 0, n_hist = 0, hist_pitch = 0;
  uint64_t cap = 0;                         // slots
  float* d_base = nullptr;                  // [cap x base_pitch]
  float* d_hist = nullptr;                  // [cap x hist_pitch]
  std::vector<uint64_t> slot_id;
  std::vector<uint32_t> slot_epoch, slot_blen, slot_hlen;
  std::vector<uint8_t> slot_live;
  std::unordered_map<uint64_t, uint32_t> index;
  uint64_t n_live = 0, hand = 0;
  uint32_t epoch = 0;
  std::vector<uint64_t> prev_ids;           // ids and slots of the previous call: a re-poll of the
  std::vector<uint32_t> prev_slot;          // same job set resolves without a single hash lookup
  std::vector<uint8_t> miss;                // per row of the current call: has to be uploaded
  bool prev_clean = false;                  // every row of the previous call was resident when it ended
  bool any_miss = false;
  uint32_t stamp_due = 0;                   // epoch of the last re-poll whose slots have not been stamped yet
  uint32_t* pin_slot = nullptr;             // prev_slot in pinned memory: its per-chunk uploads must not
  size_t pin_cap = 0;                       // stall the host behind the copy stream (pageable copies do)
  bool pin_fresh = false;                   // pin_slot == prev_slot
  uint64_t hits = 0, misses = 0, evictions = 0, uncached = 0;

  void release() {
    if (d_base) cudaFree(d_base);
    if (d_hist) cudaFree(d_hist);
    if (pin_slot) cudaFreeHost(pin_slot);
    d_base = d_hist = nullptr;
    pin_slot = nullptr; pin_cap = 0; pin_fresh = false;
    cap = 0; n_live = 0; hand = 0;
    slot_id.clear(); slot_epoch.clear(); slot_blen.clear(); slot_hlen.clear(); slot_live.clear();
    index.clear(); prev_ids.clear(); prev_slot.clear(); miss.clear(); prev_clean = false; stamp_due = 0;
  }
};

struct fm_ctx {
  int device = 0;
  int num_sms = 148;
  size_t smem_optin = 0;
  cudaStream_t stream = nullptr;   // compute
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_m = nullptr;
  FmParamsDev params;
  FmParamsDev* d_params = nullptr;
  // staging for fm_score_batch: kPipe chunk slots
  Arena in[kPipe], out[kPipe], roll, misc, hw;
  RowCache cache[2];               // one per row layout (canary rows, history-only rows)
  uint64_t cache_tick = 0, cache_used[2] = {0, 0};
  cudaEvent_t in_ready[kPipe] = {}, in_free[kPipe] = {}, out_ready[kPipe] = {}, out_free[kPipe] = {};
  std::vector<cudaEvent_t> kev;    // per-chunk kernel timing events
  int score_smem_set = 0, groups_smem_set = 0, sort_smem_set = 0, elect_smem_set = 0, count_smem_set = 0;
  // Diagnostic switches, read ONCE when the ctx is created (none is needed in production):
  int force_path = 0;              // FOREMAST_CUDA_PATH: generic = 1 (shared-memory sort kernel only),
                                   // sort = 2 (register-sort canary kernel), elect = 3 (CTA-per-series
                                   // election kernel), else 0 = auto (counting kernel where it applies)
  int force_stages = 0;            // FOREMAST_CUDA_STAGES=1|2: staging buffers per warp of the elect path
  int count_slots = 0;             // FOREMAST_CUDA_SLOTS: election-table slots of the counting kernel
  int count_warps = 0;             // FOREMAST_CUDA_WARPS: warps per SM of the counting kernel
  int ma_group = 0;                // FOREMAST_CUDA_MA_GROUP=4|8|16|32: lanes per series of k_ma_only
  int small_table = 0;             // FOREMAST_CUDA_SMALL_TABLE: elect path, load factor <= 0.8
  int count_sync = 4;              // FOREMAST_CUDA_COUNT_SYNC: trips between two CTA barriers of k_canary_count (power of two)
  int count_wpc = 0;               // FOREMAST_CUDA_COUNT_WPC: warps per CTA of k_canary_count (several CTAs per SM share the warp budget)
  int tie_runs = 24;               // FOREMAST_CUDA_TIE_RUNS: see FmKArgs::tie_runs (0 switches the class path off)
  int force_gather = 0;            // FOREMAST_CUDA_GATHER: cached rows are copied into the chunk even for kernels that read the cache in place
  int chunk_mb = 0;                // FOREMAST_CUDA_CHUNK_MB: rows per pipelined chunk of fm_score_groups (default 64 MB)
  std::string err;
};

namespace {

int32_t fail(fm_ctx* c, int32_t code, const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
  if (c) c->err = buf; else g_create_error = buf;
  return code;
}

}  // namespace

void fm_set_thread_error(const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
  g_create_error = buf;
}

namespace {

#define FM_CK(ctx, call)                                                              \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess)                                                           \
      return fail(ctx, e__ == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_CUDA,     \
                  "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__,  \
                  __LINE__);                                                          \
  } while (0)

uint32_t pow2_ceil(uint32_t n, uint32_t lo) { uint32_t p = lo; while (p < n) p <<= 1; return p; }
uint32_t round4(uint32_t n) { return (n + 3u) & ~3u; }

bool mode_has(uint32_t mode, uint32_t t) { return mode == t || mode == FM_PW_ALL || mode == FM_PW_ANY; }

// kind: 0 = k_score_series (generic), 2 = k_canary_sort<regs>, 3 = k_canary_elect<regs>, 4 = k_ma_only<regs>,
// 5 = k_canary_count, 6 = k_canary_team (baselines above 2048 points)
struct Plan { FmKArgs a; int wpc; int grid; bool fast; int kind; int regs; int valid; };

void fill_common(fm_ctx* c, const fm_batch* b, bool canary, FmKArgs& a) {
  a.cur = b->cur; a.cur_len = b->cur_len; a.n_cur = b->n_cur; a.cur_pitch = b->cur_pitch;
  a.base = canary ? b->base : nullptr; a.base_len = b->base_len; a.n_base = canary ? b->n_base : 0; a.base_pitch = b->base_pitch;
  a.hist = b->hist; a.hist_len = b->hist_len; a.n_hist = b->n_hist; a.hist_pitch = b->hist_pitch;
  a.metric_type = b->metric_type; a.api_id = b->api_id; a.n_apis = b->n_apis;
  a.n_series = b->n_series;
  a.P = c->d_params;
  a.record_kind = b->record_kind;
  a.strategy = b->strategy;
}

// k_canary_elect<NV> (fm_canary_elect.cu): current row <= 64 points, baseline <= 2048 points, one CTA
// of 128 threads per series.  Per CTA: two mbarriers, one or two staging buffers, the search tree,
// the current-point slots, two hand-off records, the per-thread overflow lists and the election
// table (load factor <= 0.4).  a.warp_bytes carries the bytes per CTA.
int32_t plan_canary_elect(fm_ctx* c, const fm_batch* b, Plan* pl) {
  FmKArgs& a = pl->a;
  a.cur_cap = 64;
  a.base_cap = round4(b->n_base);
  a.log_m = 8;
  while ((1u << a.log_m) < (b->n_cur + b->n_base) * 5u / 2u) ++a.log_m;
  if (c->small_table && a.log_m > 8) --a.log_m;   // experiment: load factor <= 0.8
  const int nv = std::max(1, (int)((((b->n_base + 3u) >> 2) + 127u) / 128u));
  auto bytes = [&](uint32_t stages) {
    size_t x = fm_canary_elect_fixed_bytes() + (size_t)stages * (64 + a.base_cap) * 4 +
               (size_t)(4 * nv + 2) * 128 * 2 + ((size_t)4 << a.log_m);
    return (x + 127) & ~size_t(127);
  };
  const size_t limit = c->smem_optin ? c->smem_optin : 232448;
  const size_t sm_total = 233472;
  a.n_stages = c->force_stages ? (uint32_t)c->force_stages : 2u;
  a.warp_bytes = (uint32_t)bytes(a.n_stages);
  // the kernel also has 512 B of static shared memory: the dynamic maximum is the opt-in limit minus that
  if (a.warp_bytes > limit - 1024) return fail(c, FM_E_UNSUPPORTED, "canary election path does not fit shared memory");
  int ctas = (int)std::min<size_t>(16, sm_total / ((size_t)a.warp_bytes + 1024 + 512));
  if (ctas < 1) ctas = 1;
  pl->wpc = 4;
  pl->grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * ctas, b->n_series));
  if (!c->elect_smem_set) { FM_CK(c, fm_canary_elect_smem_attr((int)limit - 1024)); c->elect_smem_set = 1; }
  fill_common(c, b, true, a);
  pl->fast = true; pl->kind = 3; pl->regs = nv;
  return FM_OK;
}

// k_canary_sort<R> (fm_canary_sort.cu): current row <= 64 points, baseline <= 32*R points sorted in
// registers.  Per warp: mbarrier, one staging buffer (it is re-armed as soon as the rows are in
// registers) and the parking area of the sorted baseline.
int32_t plan_canary_sort(fm_ctx* c, const fm_batch* b, Plan* pl) {
  FmKArgs& a = pl->a;
  const int R = b->n_base <= 512 ? 16 : (b->n_base <= 1024 ? 32 : 64);
  a.cur_cap = 64;
  a.base_cap = round4(b->n_base);
  a.n_stages = 1;
  // 1025..1536 points: the padding-free 1536-key network (fm_sort48.cuh) needs a 6 KB buffer only.
  // (18 warps per SM at 112 registers would fit then; measured 7.5 ms against 5.0: the spills meet
  // an L1 that 221 KB of shared memory have shrunk to a few KB.)
  const bool k48 = R == 64 && b->n_base <= 1536;
  const size_t buf_bytes = k48 ? 6144 : (size_t)128 * (R + 1);
  const int max_warps = 16;
  a.warp_bytes = (uint32_t)((16 + (size_t)(64 + a.base_cap) * 4 + buf_bytes + 127) & ~size_t(127));
  const size_t limit = c->smem_optin ? c->smem_optin : 232448;
  const size_t sm_total = 233472;
  // register budget: __launch_bounds__(256, 2) caps the kernel at 128 registers -> 16 warps per SM
  // (measured: 16 warps with a few spilled registers beat 12 warps at 168 registers, 7.0 vs 8.1 ms)
  int best_w = 1, best_total = 0;
  for (int w = 1; w <= max_warps / 2; ++w) {
    const size_t cta = (size_t)w * a.warp_bytes;
    if (cta > limit) break;
    int ctas = (int)(sm_total / (cta + 1024));
    ctas = std::min(ctas, max_warps / w);
    if (ctas < 1) continue;
    if (ctas * w >= best_total) { best_total = ctas * w; best_w = w; }
  }
  if (best_total == 0) return fail(c, FM_E_UNSUPPORTED, "canary sort path does not fit shared memory");
  pl->wpc = best_w;
  const uint64_t need = (b->n_series + best_w - 1) / best_w;
  pl->grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * (best_total / best_w), need));
  if (!c->sort_smem_set) { FM_CK(c, fm_canary_sort_smem_attr((int)limit)); c->sort_smem_set = 1; }
  fill_common(c, b, true, a);
  pl->fast = true; pl->kind = 2; pl->regs = R;
  pl->valid = (b->n_base <= 24u * (uint32_t)R) ? 3 * R / 4 : R;     // registers per lane that can hold keys
  return FM_OK;
}

// k_canary_count<NV, NS> (fm_canary_count.cu): current row <= 63 points, baseline <= 2048 points; the
// baseline is not sorted (U by searching every baseline point in the sorted current row, T from an
// election hash multiset).  Per warp: mbarrier, one staging buffer (rows, then the overflow lists),
// the search tables of the current row and the election table {pattern, counter} x slots.
int32_t plan_canary_count(fm_ctx* c, const fm_batch* b, Plan* pl) {
  FmKArgs& a = pl->a;
  int nv = 0, ns = 0;
  const uint32_t cap = fm_canary_count_tier(b->n_base, &nv, &ns);
  if (!cap) return fail(c, FM_E_UNSUPPORTED, "baseline too long for the counting path");
  a.cur_cap = 64;
  a.base_cap = cap;
  a.n_stages = 1;
  // slots: ~0.7 per key keeps the first round's losers near 45 % at 15 warps per SM (measured
  // trade-off between overflow rounds and resident warps: DESIGN.md section 5)
  uint32_t slots = c->count_slots ? (uint32_t)c->count_slots : std::max(64u, ((b->n_cur + cap) * 11u / 16u));
  slots = (slots + 3u) & ~3u;
  const size_t limit = c->smem_optin ? c->smem_optin : 232448;
  if (!c->count_slots && b->n_base <= 2048u) {
    // a 16th resident warp is worth more than the last slots (measured, profiles/var_slots.sh: 1440-point
    // baselines 9.85 ms per 1M series with 896-944 slots and 16 warps, 10.1 ms with 1032 and 15): take
    // what 16 warps leave if that is still 9/16 of a slot per key
    const size_t per_warp = (limit / (size_t)fm_canary_count_max_warps()) & ~size_t(127);
    const size_t fixed = fm_canary_count_fixed_bytes() + (size_t)(64 + cap) * 4;
    if (per_warp > fixed) {
      const uint32_t s16 = (uint32_t)((per_warp - fixed) / 8) & ~3u;
      if (s16 < slots && s16 >= (b->n_cur + cap) * 9u / 16u) slots = s16;
    }
  }
  a.log_m = slots;
  a.cnt_nv = (uint32_t)nv; a.cnt_ns = (uint32_t)ns;
  const size_t sm_total = 233472;
  if (!c->count_smem_set) { FM_CK(c, fm_canary_count_smem_attr((int)limit)); c->count_smem_set = 1; }
  fill_common(c, b, true, a);
  if (b->n_base > 2048u) {
    // long baselines: a team of 8 warps per series shares the stage and the table (k_canary_team);
    // its last-resort direct count parks 2 x 256 words in the table area
    if (slots < 512u) { slots = 512u; a.log_m = slots; }
    a.warp_bytes = (uint32_t)((fm_canary_team_fixed_bytes() + (size_t)cap * 4 + (size_t)slots * 8 + 127) & ~size_t(127));
    if (a.warp_bytes > limit) return fail(c, FM_E_UNSUPPORTED, "canary counting path does not fit shared memory");
    const int ctas = std::max(1, std::min(2, (int)(sm_total / ((size_t)a.warp_bytes + 1024))));
    pl->wpc = 8;
    pl->grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * ctas, b->n_series));
    pl->fast = true; pl->kind = 6; pl->regs = nv; pl->valid = ns;
    return FM_OK;
  }
  a.warp_bytes = (uint32_t)((fm_canary_count_fixed_bytes() + (size_t)(64 + cap) * 4 + (size_t)slots * 8 + 127) & ~size_t(127));
  const int built_w = fm_canary_count_max_warps();                 // 16: 128 registers per thread
  const int max_warps = c->count_warps ? std::min(c->count_warps, built_w) : built_w;
  int best_w = 1, best_total = 0;
  // ONE CTA per SM whose warps meet at a barrier per series, so that they share the instruction
  // fetches of the long loop body (+4 %); FOREMAST_CUDA_STAGES=1 (experiment): several free-running CTAs
  a.n_stages = c->force_stages == 1 ? 1u : 2u;
  a.sync_mask = (uint32_t)c->count_sync - 1u;
  a.tie_runs = (uint32_t)c->tie_runs;
  const int w_hi = a.n_stages == 2u ? built_w : 8, w_lo = a.n_stages == 2u ? 2 : 1;
  for (int w = w_lo; w <= w_hi; ++w) {
    if (c->count_wpc && w != c->count_wpc) continue;       // experiment: CTA size forced
    const size_t cta = (size_t)w * a.warp_bytes;
    if (cta > limit) break;
    int ctas = (int)(sm_total / (cta + 1024));
    ctas = std::min(ctas, max_warps / w);
    if (ctas < 1) continue;
    if (ctas * w >= best_total) { best_total = ctas * w; best_w = w; }
  }
  if (best_total == 0) return fail(c, FM_E_UNSUPPORTED, "canary counting path does not fit shared memory");
  pl->wpc = best_w;
  const uint64_t need = (b->n_series + best_w - 1) / best_w;
  pl->grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * (best_total / best_w), need));
  pl->fast = true; pl->kind = 5; pl->regs = nv; pl->valid = ns;
  return FM_OK;
}

// k_ma_only<G> (fm_ma.cu): no baseline, plain moving_average_all model, current row <= 64 points.
int32_t plan_ma_only(fm_ctx* c, const fm_batch* b, Plan* pl) {
  FmKArgs& a = pl->a;
  fill_common(c, b, false, a);
  pl->regs = b->n_hist >= 1200 ? 16 : (b->n_hist >= 400 ? 8 : 4);      // lanes per series (profiles/r02/sweep_groupsize.txt)
  if (c->ma_group) pl->regs = c->ma_group;   // experiment
  const uint64_t per_warp = 32 / pl->regs;
  const uint64_t need = (b->n_series + per_warp * 8 - 1) / (per_warp * 8);   // 8 warps per CTA
  pl->wpc = 8;
  // persistent grid = the CTAs that are resident at once (64 registers x 256 threads: 4 per SM; the
  // 2-lane form 3 per SM), so that there is no second, partial wave
  pl->grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * fm_ma_only_ctas(pl->regs), need));
  pl->fast = true; pl->kind = 4;
  return FM_OK;
}

// Shared-memory layout and launch shape for k_score_series.
int32_t plan_score(fm_ctx* c, const fm_batch* b, Plan* pl) {
  FmKArgs& a = pl->a;
  memset(&a, 0, sizeof(a));
  pl->fast = false; pl->kind = 0; pl->regs = 0;
  const bool canary = b->strategy == FM_STRAT_CANARY && b->base != nullptr;
  if (b->n_cur > FM_MAX_CUR) return fail(c, FM_E_UNSUPPORTED, "n_cur %u > %u", b->n_cur, FM_MAX_CUR);
  if (canary && (uint64_t)b->n_cur + b->n_base > FM_MAX_RANK_N)
    return fail(c, FM_E_UNSUPPORTED, "n_cur + n_base %u > %u", b->n_cur + b->n_base, FM_MAX_RANK_N);
  const bool holt_winters = c->params.g.ma_mode == FM_MODEL_HOLT_WINTERS;   // general kernel only
  if (!canary && c->force_path != 1 && b->hist && b->n_cur <= 64 && c->params.g.ma_mode == FM_MA_ALL)
    return plan_ma_only(c, b, pl);
  if (canary && c->force_path != 1 && b->n_cur <= 64 && b->n_base >= 1 && !holt_winters) {
    if (c->force_path == 3 && b->n_base <= 2048) return plan_canary_elect(c, b, pl);
    if (c->force_path != 2 && b->n_cur <= 63 && b->n_base <= 16320) return plan_canary_count(c, b, pl);
    if (b->n_base <= 2048) return plan_canary_sort(c, b, pl);
  }
  a.cur_cap = std::max(4u, round4(b->n_cur));
  a.cs_cap = pow2_ceil(b->n_cur, 64u);
  a.base_cap = canary ? pow2_ceil(b->n_base, 64u) : 0u;
  const bool wx = canary && mode_has(c->params.g.pairwise_mode, FM_PW_WILCOXON) &&
                  c->params.g.pairwise_mode != FM_PW_NONE;
  a.wx_n = wx ? ((std::min(b->n_cur, b->n_base) + 1u) & ~1u) : 0u;
  a.wx_cap = wx ? pow2_ceil(a.wx_n, 64u) : 0u;
  auto bytes = [&](uint32_t stages) {
    size_t x = 16 + (size_t)stages * (a.cur_cap + a.base_cap) * 4 + (size_t)a.cs_cap * 4 +
               (size_t)a.wx_n * 8 + (size_t)a.wx_cap * 8;
    return (x + 127) & ~size_t(127);
  };
  const size_t limit = c->smem_optin ? c->smem_optin : 232448;
  const size_t sm_total = 233472 - 1024;  // 228 KB per SM, 1 KB reserved per CTA
  a.n_stages = 2;
  if (bytes(2) > limit) a.n_stages = 1;
  // large rows: rather keep a few warps per SM than double buffer each of them
  if (a.n_stages == 2 && sm_total / bytes(2) < 4 && sm_total / bytes(1) >= 2 * (sm_total / bytes(2)))
    a.n_stages = 1;
  if (bytes(1) > limit) return fail(c, FM_E_UNSUPPORTED, "series too long for shared memory (%zu B per warp)", bytes(1));
  a.warp_bytes = (uint32_t)bytes(a.n_stages);
  // warps per CTA and CTAs per SM: fill the SM's shared memory, at most 16 warps per CTA
  int wpc = (int)std::min<size_t>(16, limit / a.warp_bytes);
  if (wpc < 1) wpc = 1;
  int ctas = (int)std::min<size_t>(4, sm_total / ((size_t)wpc * a.warp_bytes + 1024));
  if (ctas < 1) ctas = 1;
  while (ctas * wpc > 64) --ctas;
  pl->wpc = wpc;
  const uint64_t need = (b->n_series + wpc - 1) / wpc;
  pl->grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * ctas, need));
  const int smem = wpc * (int)a.warp_bytes;
  if (smem > c->score_smem_set) {
    FM_CK(c, fm_score_smem_attr((int)limit));
    c->score_smem_set = (int)limit;
  }
  a.cur = b->cur; a.cur_len = b->cur_len; a.n_cur = b->n_cur; a.cur_pitch = b->cur_pitch;
  a.base = canary ? b->base : nullptr; a.base_len = b->base_len; a.n_base = canary ? b->n_base : 0; a.base_pitch = b->base_pitch;
  a.hist = b->hist; a.hist_len = b->hist_len; a.n_hist = b->n_hist; a.hist_pitch = b->hist_pitch;
  a.metric_type = b->metric_type; a.api_id = b->api_id; a.n_apis = b->n_apis;
  a.n_series = b->n_series;
  a.P = c->d_params;
  a.record_kind = b->record_kind;
  a.strategy = b->strategy;
  if (holt_winters) {   // seasonal state: season_len doubles per resident warp
    const uint32_t m = c->params.g.season_len;
    if (m < 2u || m > (1u << 20)) return fail(c, FM_E_INVALID, "FM_MODEL_HOLT_WINTERS needs 2 <= season_len <= 2^20 (got %u)", m);
    FM_CK(c, c->hw.reserve((size_t)pl->grid * wpc * m * sizeof(double)));
    a.hw_scratch = static_cast<double*>(c->hw.p);
  }
  return FM_OK;
}

int launch_plan(const Plan& pl, cudaStream_t st) {
  if (pl.kind == 6) return fm_launch_canary_team(pl.a, pl.wpc, pl.grid, st);
  if (pl.kind == 5) return fm_launch_canary_count(pl.a, pl.wpc, pl.grid, st);
  if (pl.kind == 4) return fm_launch_ma_only(pl.a, pl.regs, pl.grid, st);
  if (pl.kind == 3) return fm_launch_canary_elect(pl.a, pl.regs, pl.grid, st);
  if (pl.kind == 2) return fm_launch_canary_sort(pl.a, pl.regs, pl.valid, pl.wpc, pl.grid, st);
  return fm_launch_score(pl.a, pl.wpc, pl.grid, st);
}

int32_t validate_batch(fm_ctx* c, const fm_batch* b, const void* out) {
  if (!c) return FM_E_INVALID;
  if (!b || !out) return fail(c, FM_E_INVALID, "null batch or output");
  if (b->struct_size != sizeof(fm_batch)) return fail(c, FM_E_INVALID, "fm_batch.struct_size %u != %zu", b->struct_size, sizeof(fm_batch));
  if (b->strategy > FM_STRAT_HISTORICAL) return fail(c, FM_E_INVALID, "unknown strategy %u", b->strategy);
  if (b->record_kind > FM_REC_COMPACT) return fail(c, FM_E_INVALID, "unknown record kind %u", b->record_kind);
  if (b->n_series && !b->cur && b->n_cur) return fail(c, FM_E_INVALID, "cur is NULL");
  auto bad = [](const float* p, uint32_t n, uint32_t pitch) {
    return p && ((reinterpret_cast<uintptr_t>(p) & 15u) || (pitch & 3u) || pitch < ((n + 3u) & ~3u));
  };
  if (bad(b->cur, b->n_cur, b->cur_pitch) || bad(b->base, b->n_base, b->base_pitch) ||
      bad(b->hist, b->n_hist, b->hist_pitch))
    return fail(c, FM_E_INVALID, "rows must be 16-byte aligned with pitch a multiple of 4 and >= roundup4(n)");
  if (b->api_id && b->n_apis == 0) return fail(c, FM_E_INVALID, "api_id given but n_apis == 0");
  return FM_OK;
}

size_t rec_size(uint32_t kind) { return kind == FM_REC_COMPACT ? sizeof(fm_compact_rec) : sizeof(fm_series_rec); }

}  // namespace

// ---------------------------------------------------------------------------
extern "C" {

int32_t fm_version(void) { return (FM_VERSION_MAJOR << 16) | FM_VERSION_MINOR; }

uint32_t fm_sizeof(uint32_t which) {
  switch (which) {
    case 0: return sizeof(fm_global_params);
    case 1: return sizeof(fm_type_params);
    case 2: return sizeof(fm_batch);
    case 3: return sizeof(fm_group_batch);
    case 4: return sizeof(fm_series_rec);
    case 5: return sizeof(fm_compact_rec);
    case 6: return sizeof(fm_rank_aux);
    case 7: return sizeof(fm_rollup_rec);
    case 8: return sizeof(fm_timing);
    case 9: return sizeof(fm_synth_desc);
    case 10: return sizeof(fm_cache_stats);
    case 11: return sizeof(fm_pair_batch);
    case 12: return sizeof(fm_pair_rec);
    default: return 0;
  }
}

int32_t fm_device_count(int32_t* out) {
  if (!out) return FM_E_INVALID;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
  *out = n;
  return FM_OK;
}

int32_t fm_default_params(fm_global_params* g, fm_type_params* table, uint32_t* n_types) {
  if (!g || !table || !n_types || *n_types < 6) return FM_E_INVALID;
  memset(g, 0, sizeof(*g));
  g->struct_size = sizeof(*g);
  g->ddof = 0; g->ma_mode = FM_MA_ALL; g->ma_window = 0; g->min_hist_points = 10;
  g->pairwise_mode = FM_PW_ANY; g->min_mw = 20; g->min_wx = 20; g->min_kw = 5;
  g->cc_mw = 1; g->cc_wx = 0; g->pair_rule = FM_PAIR_HEAD; g->min_anomaly_points = 1;
  g->alpha = 0.05; g->smooth_alpha = 0.3; g->smooth_beta = 0.1; g->smooth_gamma = 0.1; g->season_len = 0;
  // foremast-brain.yaml:26-73: default, error5xx, error4xx, latency, cpu, memory
  const double k[6] = {2.0, 2.0, 3.0, 10.0, 5.0, 5.0};
  const uint32_t bound[6] = {1, 1, 1, 3, 1, 1};
  for (int i = 0; i < 6; ++i) {
    table[i].k = k[i]; table[i].k_low = 0.5 * k[i]; table[i].min_lower_bound = 0.0;
    table[i].bound_mask = bound[i]; table[i].reserved = 0;
  }
  *n_types = 6;
  return FM_OK;
}

int32_t fm_ctx_create(int32_t device, fm_ctx** out) {
  if (!out) return FM_E_INVALID;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(nullptr, FM_E_NODEVICE, "no CUDA device available (libforemast_cuda has no CPU fallback)");
  }
  if (device < 0 || device >= n) return fail(nullptr, FM_E_INVALID, "device %d out of range [0,%d)", device, n);
  fm_ctx* c = new (std::nothrow) fm_ctx();
  if (!c) return FM_E_NOMEM;
  c->device = device;
#define FM_CKC(call)                                                                           \
  do {                                                                                         \
    cudaError_t e__ = (call);                                                                  \
    if (e__ != cudaSuccess) {                                                                  \
      fail(nullptr, FM_E_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__));               \
      fm_ctx_destroy(c);   /* streams and events created so far go with it */                  \
      return FM_E_CUDA;                                                                        \
    }                                                                                          \
  } while (0)
  FM_CKC(cudaSetDevice(device));
  cudaDeviceProp prop;
  FM_CKC(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    { fail(nullptr, FM_E_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor); delete c; return FM_E_UNSUPPORTED; }
  c->num_sms = prop.multiProcessorCount;
  c->smem_optin = prop.sharedMemPerBlockOptin;
  if (const char* e = getenv("FOREMAST_CUDA_PATH"))
    c->force_path = strcmp(e, "generic") == 0 ? 1 : (strcmp(e, "sort") == 0 ? 2 : (strcmp(e, "elect") == 0 ? 3 : 0));
  if (const char* e = getenv("FOREMAST_CUDA_STAGES")) c->force_stages = (e[0] == '1') ? 1 : (e[0] == '2' ? 2 : 0);
  if (const char* e = getenv("FOREMAST_CUDA_SLOTS")) { const int v = atoi(e); c->count_slots = (v >= 64 && v <= 8192) ? v : 0; }
  if (const char* e = getenv("FOREMAST_CUDA_WARPS")) { const int v = atoi(e); c->count_warps = (v >= 1 && v <= 32) ? v : 0; }
  if (const char* e = getenv("FOREMAST_CUDA_MA_GROUP")) { const int v = atoi(e); c->ma_group = (v == 2 || v == 4 || v == 8 || v == 16 || v == 32) ? v : 0; }
  c->small_table = getenv("FOREMAST_CUDA_SMALL_TABLE") != nullptr;
  c->force_gather = getenv("FOREMAST_CUDA_GATHER") != nullptr;
  if (const char* e = getenv("FOREMAST_CUDA_COUNT_WPC")) { const int v = atoi(e); if (v >= 2 && v <= 32) c->count_wpc = v; }
  if (const char* e = getenv("FOREMAST_CUDA_TIE_RUNS")) { const int v = atoi(e); if (v >= 0 && v <= 65) c->tie_runs = v; }
  if (const char* e = getenv("FOREMAST_CUDA_COUNT_SYNC")) { const int v = atoi(e); if (v >= 1 && v <= 1024 && (v & (v - 1)) == 0) c->count_sync = v; }
  if (const char* e = getenv("FOREMAST_CUDA_CHUNK_MB")) { const int v = atoi(e); c->chunk_mb = (v >= 1 && v <= 4096) ? v : 0; }
  FM_CKC(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  FM_CKC(cudaStreamCreateWithFlags(&c->s_h2d, cudaStreamNonBlocking));
  FM_CKC(cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
  FM_CKC(cudaEventCreate(&c->ev_a));
  FM_CKC(cudaEventCreate(&c->ev_b));
  FM_CKC(cudaEventCreate(&c->ev_m));
  for (int i = 0; i < kPipe; ++i) {
    FM_CKC(cudaEventCreateWithFlags(&c->in_ready[i], cudaEventDisableTiming));
    FM_CKC(cudaEventCreateWithFlags(&c->in_free[i], cudaEventDisableTiming));
    FM_CKC(cudaEventCreateWithFlags(&c->out_ready[i], cudaEventDisableTiming));
    FM_CKC(cudaEventCreateWithFlags(&c->out_free[i], cudaEventDisableTiming));
  }
  FM_CKC(cudaMalloc(&c->d_params, sizeof(FmParamsDev)));
#undef FM_CKC
  memset(&c->params, 0, sizeof(c->params));
  uint32_t nt = FM_MAX_TYPES;
  fm_default_params(&c->params.g, c->params.types, &nt);
  c->params.n_types = nt;
  cudaError_t e = cudaMemcpy(c->d_params, &c->params, sizeof(FmParamsDev), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { fail(nullptr, FM_E_CUDA, "params upload: %s", cudaGetErrorString(e)); fm_ctx_destroy(c); return FM_E_CUDA; }
  *out = c;
  return FM_OK;
}

int32_t fm_ctx_destroy(fm_ctx* c) {
  if (!c) return FM_OK;
  cudaSetDevice(c->device);
  if (c->s_h2d) cudaStreamSynchronize(c->s_h2d);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->s_d2h) cudaStreamSynchronize(c->s_d2h);
  for (int i = 0; i < kPipe; ++i) {
    c->in[i].release(); c->out[i].release();
    if (c->in_ready[i]) cudaEventDestroy(c->in_ready[i]);
    if (c->in_free[i]) cudaEventDestroy(c->in_free[i]);
    if (c->out_ready[i]) cudaEventDestroy(c->out_ready[i]);
    if (c->out_free[i]) cudaEventDestroy(c->out_free[i]);
  }
  for (cudaEvent_t e : c->kev) cudaEventDestroy(e);
  c->roll.release(); c->misc.release(); c->hw.release(); c->cache[0].release(); c->cache[1].release();
  if (c->d_params) cudaFree(c->d_params);
  if (c->ev_a) cudaEventDestroy(c->ev_a);
  if (c->ev_b) cudaEventDestroy(c->ev_b);
  if (c->ev_m) cudaEventDestroy(c->ev_m);
  if (c->s_h2d) cudaStreamDestroy(c->s_h2d);
  if (c->s_d2h) cudaStreamDestroy(c->s_d2h);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return FM_OK;
}

int32_t fm_set_params(fm_ctx* c, const fm_global_params* g, const fm_type_params* table, uint32_t n_types) {
  if (!c) return FM_E_INVALID;
  if (!g || !table || n_types == 0 || n_types > FM_MAX_TYPES) return fail(c, FM_E_INVALID, "bad params (n_types %u)", n_types);
  if (g->struct_size != sizeof(fm_global_params)) return fail(c, FM_E_INVALID, "fm_global_params.struct_size %u != %zu", g->struct_size, sizeof(fm_global_params));
  if (g->ddof > 1 || g->ma_mode > FM_MODEL_HOLT_WINTERS || g->pairwise_mode > FM_PW_FRIEDMAN || g->pair_rule > FM_PAIR_TAIL)
    return fail(c, FM_E_INVALID, "parameter out of range");
  FM_CK(c, cudaSetDevice(c->device));
  c->params.g = *g;
  c->params.n_types = n_types;
  memcpy(c->params.types, table, n_types * sizeof(fm_type_params));
  // a barrier on the ctx: queued launches read the old table
  FM_CK(c, cudaStreamSynchronize(c->stream));
  FM_CK(c, cudaMemcpy(c->d_params, &c->params, sizeof(FmParamsDev), cudaMemcpyHostToDevice));
  return FM_OK;
}

int32_t fm_alloc_pinned(size_t bytes, void** out) {
  if (!out) return FM_E_INVALID;
  *out = nullptr;
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);
  if (e != cudaSuccess) { cudaGetLastError(); return e == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_NODEVICE; }
  return FM_OK;
}
int32_t fm_free_pinned(void* p) { if (p) cudaFreeHost(p); return FM_OK; }

int32_t fm_device_alloc(fm_ctx* c, size_t bytes, void** out) {
  if (!c || !out) return FM_E_INVALID;
  FM_CK(c, cudaSetDevice(c->device));
  FM_CK(c, cudaMalloc(out, bytes ? bytes : 1));
  return FM_OK;
}
int32_t fm_device_free(fm_ctx* c, void* p) {
  if (!c) return FM_E_INVALID;
  FM_CK(c, cudaSetDevice(c->device));
  if (p) FM_CK(c, cudaFree(p));
  return FM_OK;
}
int32_t fm_memcpy_h2d(fm_ctx* c, void* dst, const void* src, size_t bytes) {
  if (!c) return FM_E_INVALID;
  FM_CK(c, cudaSetDevice(c->device));
  FM_CK(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream));
  FM_CK(c, cudaStreamSynchronize(c->stream));
  return FM_OK;
}
int32_t fm_memcpy_d2h(fm_ctx* c, void* dst, const void* src, size_t bytes) {
  if (!c) return FM_E_INVALID;
  FM_CK(c, cudaSetDevice(c->device));
  FM_CK(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
  FM_CK(c, cudaStreamSynchronize(c->stream));
  return FM_OK;
}
int32_t fm_ctx_sync(fm_ctx* c) {
  if (!c) return FM_E_INVALID;
  FM_CK(c, cudaSetDevice(c->device));
  FM_CK(c, cudaStreamSynchronize(c->stream));
  return FM_OK;
}

const char* fm_last_error(fm_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

// ---------------------------------------------------------------------------
int32_t fm_score_batch_device(fm_ctx* c, const fm_batch* b, void* d_out, fm_rank_aux* d_aux,
                              fm_rollup_rec* d_api, fm_rollup_rec* d_svc, fm_timing* timing) {
  int32_t rc = validate_batch(c, b, d_out);
  if (rc != FM_OK) return rc;
  FM_CK(c, cudaSetDevice(c->device));
  if (timing) memset(timing, 0, sizeof(*timing));
  if (b->n_series == 0) return FM_OK;
  Plan pl;
  rc = plan_score(c, b, &pl);
  if (rc != FM_OK) return rc;
  pl.a.out = d_out; pl.a.aux = d_aux;
  const bool roll = b->api_id && d_api;
  pl.a.api = roll ? d_api : nullptr;
  uint32_t launches = 0;
  if (roll) {
    FM_CK(c, cudaMemsetAsync(d_api, 0, (size_t)b->n_apis * sizeof(fm_rollup_rec), c->stream));
    if (d_svc && b->api_service) FM_CK(c, cudaMemsetAsync(d_svc, 0, (size_t)b->n_services * sizeof(fm_rollup_rec), c->stream));
  }
  if (timing) FM_CK(c, cudaEventRecord(c->ev_a, c->stream));
  launches += launch_plan(pl, c->stream);
  if (timing) FM_CK(c, cudaEventRecord(c->ev_m, c->stream));
  if (roll)
    launches += fm_launch_rollup_finalize(d_api, b->n_apis, b->api_service,
                                          (d_svc && b->api_service) ? d_svc : nullptr, b->n_services, c->stream);
  FM_CK(c, cudaGetLastError());
  if (timing) {
    FM_CK(c, cudaEventRecord(c->ev_b, c->stream));
    FM_CK(c, cudaEventSynchronize(c->ev_b));
    FM_CK(c, cudaEventElapsedTime(&timing->kernel_ms, c->ev_a, c->ev_b));
    FM_CK(c, cudaEventElapsedTime(&timing->main_kernel_ms, c->ev_a, c->ev_m));
    timing->launches = launches;
  }
  return FM_OK;
}

}  // extern "C"

namespace {

// (Re)build the cache storage for this batch's row layout.  Returns false when the cache is off or
// cannot hold a single row.
int32_t cache_prepare(fm_ctx* c, const fm_batch* b, bool canary, RowCache** on) {
  *on = nullptr;
  if (c->cache[0].max_bytes == 0) return FM_OK;
  const uint32_t nb = canary ? b->n_base : 0u, pb = canary ? b->base_pitch : 0u;
  const uint32_t nh = b->hist ? b->n_hist : 0u, ph = b->hist ? b->hist_pitch : 0u;
  const uint64_t row_bytes = 4ull * ((uint64_t)pb + ph);
  if (row_bytes == 0) return FM_OK;                       // nothing cacheable in this batch
  // the cache of this layout; else an unused one; else the one that was used longest ago
  int which = -1;
  for (int i = 0; i < 2; ++i) {
    const RowCache& q = c->cache[i];
    if (q.cap && q.n_base == nb && q.base_pitch == pb && q.n_hist == nh && q.hist_pitch == ph) which = i;
  }
  if (which < 0) {
    which = !c->cache[0].cap ? 0 : (!c->cache[1].cap ? 1 : (c->cache_used[0] <= c->cache_used[1] ? 0 : 1));
    c->cache[which].release();
  }
  RowCache& k = c->cache[which];
  c->cache_used[which] = ++c->cache_tick;
  if (!k.cap) {
    const uint64_t cap = std::min<uint64_t>(k.max_bytes / row_bytes, 0xFFFFFFF0ull);
    if (cap == 0) return FM_OK;
    if (pb) FM_CK(c, cudaMalloc(&k.d_base, cap * pb * 4ull));
    if (ph) {
      cudaError_t e = cudaMalloc(&k.d_hist, cap * ph * 4ull);
      if (e != cudaSuccess) { k.release(); FM_CK(c, e); }
    }
    k.cap = cap; k.n_base = nb; k.base_pitch = pb; k.n_hist = nh; k.hist_pitch = ph;
    k.slot_id.assign(cap, 0); k.slot_epoch.assign(cap, 0); k.slot_blen.assign(cap, 0); k.slot_hlen.assign(cap, 0);
    k.slot_live.assign(cap, 0);
    k.index.reserve((size_t)std::min<uint64_t>(cap, 1u << 22));
  }
  *on = &k;
  return FM_OK;
}

// Slot of every row of the call, left in k.prev_slot (kNone: the cache is full of rows of this very
// call -> that row is uploaded into the chunk buffer like an uncached one), and k.miss: whether the
// row has to be uploaded.  A re-poll of exactly the previous call's rows (the steady state of the
// job loop) costs one memcmp and one pass that stamps the slots as used.
void cache_assign(RowCache& k, const fm_batch* b, const uint64_t* ids, uint64_t S, bool canary) {
  std::vector<uint32_t>& slot = k.prev_slot;
  std::vector<uint8_t>& miss = k.miss;
  if (++k.epoch == 0) {                                   // epoch wrapped: nobody is "in use"
    std::fill(k.slot_epoch.begin(), k.slot_epoch.end(), 0u);
    k.epoch = 1;
    if (k.stamp_due) k.stamp_due = 1;
  }
  const uint64_t prev_n = k.prev_ids.size();
  if (k.prev_clean && prev_n == S && memcmp(ids, k.prev_ids.data(), S * sizeof(uint64_t)) == 0) {
    k.stamp_due = k.epoch;                                // "used in this call", written when somebody needs it
    k.hits += S;
    if (k.any_miss || miss.size() != S) miss.assign(S, 0);
    k.any_miss = false;
    return;
  }
  k.pin_fresh = false;
  if (k.stamp_due) {                                      // the re-polls since the last full pass used these slots
    for (uint64_t s = 0; s < prev_n && s < slot.size(); ++s)
      if (slot[s] != RowCache::kNone) k.slot_epoch[slot[s]] = k.stamp_due;
    k.stamp_due = 0;
  }
  if (slot.size() < S) slot.resize(S);
  miss.assign(S, 0);
  k.any_miss = false;
  // pass 1: rows that are resident (and protect their slots from this call's evictions)
  for (uint64_t s = 0; s < S; ++s) {
    const uint64_t id = ids[s];
    uint32_t sl = RowCache::kNone;
    if (s < prev_n && k.prev_ids[s] == id) {
      const uint32_t ps = slot[s];
      if (ps != RowCache::kNone && k.slot_live[ps] && k.slot_id[ps] == id) sl = ps;
    }
    if (sl == RowCache::kNone) {
      auto it = k.index.find(id);
      if (it != k.index.end()) sl = it->second;
    }
    if (sl != RowCache::kNone) { ++k.hits; k.slot_epoch[sl] = k.epoch; }
    slot[s] = sl;
  }
  // pass 2: a slot for every other row: a free one, else the least recently scored one that this
  // call does not use (clock hand, 64 candidates); none left -> the row stays uncached
  bool clean = true;
  for (uint64_t s = 0; s < S; ++s) {
    if (slot[s] != RowCache::kNone) continue;
    const uint64_t id = ids[s];
    auto it = k.index.find(id);                            // the same id twice in one call
    if (it != k.index.end()) { slot[s] = it->second; ++k.hits; continue; }
    uint32_t sl = RowCache::kNone;
    if (k.n_live < k.cap) {
      while (k.slot_live[k.hand]) k.hand = (k.hand + 1) % k.cap;
      sl = (uint32_t)k.hand;
      k.hand = (k.hand + 1) % k.cap;
    } else {
      uint64_t best = k.cap; uint32_t best_age = 0;
      for (uint64_t probe = 0; probe < std::min<uint64_t>(k.cap, 64); ++probe) {
        const uint64_t q = (k.hand + probe) % k.cap;
        if (k.slot_epoch[q] == k.epoch) continue;
        const uint32_t age = k.epoch - k.slot_epoch[q];
        if (best == k.cap || age > best_age) { best = q; best_age = age; }
      }
      if (best == k.cap) {
        for (uint64_t q = 0; q < k.cap; ++q) if (k.slot_epoch[q] != k.epoch) { best = q; break; }
      }
      if (best != k.cap) {
        k.index.erase(k.slot_id[best]);
        k.slot_live[best] = 0; --k.n_live; ++k.evictions;
        sl = (uint32_t)best;
        k.hand = (best + 1) % k.cap;
      }
    }
    miss[s] = 1;
    k.any_miss = true;
    if (sl != RowCache::kNone) {
      k.slot_live[sl] = 1; ++k.n_live;
      k.slot_id[sl] = id;
      k.slot_epoch[sl] = k.epoch;
      k.slot_blen[sl] = canary ? (b->base_len ? std::min(b->base_len[s], b->n_base) : b->n_base) : 0u;
      k.slot_hlen[sl] = b->hist ? (b->hist_len ? std::min(b->hist_len[s], b->n_hist) : b->n_hist) : 0u;
      k.index[id] = sl;
      ++k.misses;
    } else {
      ++k.uncached;
      clean = false;
    }
    slot[s] = sl;
  }
  slot.resize(S);
  k.prev_ids.assign(ids, ids + S);
  k.prev_clean = clean;
}

int32_t score_batch_impl(fm_ctx* c, const fm_batch* b, const uint64_t* row_ids, void* out, fm_rank_aux* out_aux,
                         fm_rollup_rec* out_api, fm_rollup_rec* out_svc, fm_timing* timing) {
  int32_t rc = validate_batch(c, b, out);
  if (rc != FM_OK) return rc;
  FM_CK(c, cudaSetDevice(c->device));
  if (timing) memset(timing, 0, sizeof(*timing));
  const uint64_t S = b->n_series;
  if (S == 0) return FM_OK;
  const bool canary = b->strategy == FM_STRAT_CANARY && b->base != nullptr;
  const bool roll = b->api_id && out_api && b->n_apis;
  const bool roll_svc = roll && out_svc && b->api_service && b->n_services;
  const size_t rsz = rec_size(b->record_kind);

  // row cache: slot per row, which rows still have to come over PCIe
  RowCache* kcp = nullptr;
  std::vector<uint32_t> eff_blen, eff_hlen;
  if (row_ids) {
    rc = cache_prepare(c, b, canary, &kcp);
    if (rc != FM_OK) return rc;
  }
  const bool cached = kcp != nullptr;
  RowCache& kc = cached ? *kcp : c->cache[0];
  const std::vector<uint32_t>& slot = kc.prev_slot;
  const std::vector<uint8_t>& miss = kc.miss;
  if (cached) {
    cache_assign(kc, b, row_ids, S, canary);
    if (!kc.pin_fresh) {
      if (kc.pin_cap < S) {
        if (kc.pin_slot) cudaFreeHost(kc.pin_slot);
        kc.pin_slot = nullptr; kc.pin_cap = 0;
        FM_CK(c, cudaHostAlloc(reinterpret_cast<void**>(&kc.pin_slot), S * sizeof(uint32_t), cudaHostAllocDefault));
        kc.pin_cap = S;
      }
      memcpy(kc.pin_slot, kc.prev_slot.data(), S * sizeof(uint32_t));
      kc.pin_fresh = true;
    }
    // lengths: a resident row keeps the length it was stored with
    if (canary && b->base_len) {
      eff_blen.resize(S);
      for (uint64_t s = 0; s < S; ++s) eff_blen[s] = slot[s] != RowCache::kNone ? kc.slot_blen[slot[s]] : std::min(b->base_len[s], b->n_base);
    }
    if (b->hist && b->hist_len) {
      eff_hlen.resize(S);
      for (uint64_t s = 0; s < S; ++s) eff_hlen[s] = slot[s] != RowCache::kNone ? kc.slot_hlen[slot[s]] : std::min(b->hist_len[s], b->n_hist);
    }
  }
  const uint32_t* h_blen = cached && !eff_blen.empty() ? eff_blen.data() : b->base_len;
  const uint32_t* h_hlen = cached && !eff_hlen.empty() ? eff_hlen.data() : b->hist_len;

  // chunking: ~64 MB of rows per chunk, kPipe slots in flight
  const size_t row_bytes = 4 * ((size_t)b->cur_pitch + (canary ? b->base_pitch : 0) + (b->hist ? b->hist_pitch : 0));
  // (with resident rows the copies are small and the per-chunk launches dominate: larger chunks)
  uint64_t chunk = row_bytes ? (size_t(c->chunk_mb ? c->chunk_mb : (cached ? 256 : 64)) << 20) / row_bytes : S;
  chunk = std::max<uint64_t>(chunk, 2048);
  chunk = std::min<uint64_t>(chunk, S);
  const uint64_t n_chunks = (S + chunk - 1) / chunk;

  // per-slot input layout (all offsets 256-byte aligned)
  auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
  size_t o = 0;
  const size_t o_cur = o;  o += al(chunk * b->cur_pitch * 4);
  const size_t o_base = o; o += canary ? al(chunk * b->base_pitch * 4) : 0;
  const size_t o_hist = o; o += b->hist ? al(chunk * b->hist_pitch * 4) : 0;
  const size_t o_clen = o; o += b->cur_len ? al(chunk * 4) : 0;
  const size_t o_blen = o; o += (canary && b->base_len) ? al(chunk * 4) : 0;
  const size_t o_hlen = o; o += (b->hist && b->hist_len) ? al(chunk * 4) : 0;
  const size_t o_mt = o;   o += b->metric_type ? al(chunk) : 0;
  const size_t o_api = o;  o += roll ? al(chunk * 4) : 0;
  const size_t o_slot = o; o += cached ? al(chunk * 4) : 0;
  const size_t in_bytes = o;
  const size_t o_rec = 0;
  const size_t o_aux = al(chunk * rsz);
  const size_t out_bytes = o_aux + (out_aux ? al(chunk * sizeof(fm_rank_aux)) : 0);
  const int slots = (int)std::min<uint64_t>(n_chunks, kPipe);
  for (int i = 0; i < slots; ++i) { FM_CK(c, c->in[i].reserve(in_bytes)); FM_CK(c, c->out[i].reserve(out_bytes)); }
  fm_rollup_rec *d_api = nullptr, *d_svc = nullptr;
  uint32_t* d_api_service = nullptr;
  if (roll) {
    const size_t a0 = al((size_t)b->n_apis * sizeof(fm_rollup_rec));
    const size_t a1 = roll_svc ? al((size_t)b->n_services * sizeof(fm_rollup_rec)) : 0;
    const size_t a2 = roll_svc ? al((size_t)b->n_apis * 4) : 0;
    FM_CK(c, c->roll.reserve(a0 + a1 + a2));
    d_api = static_cast<fm_rollup_rec*>(c->roll.p);
    FM_CK(c, cudaMemsetAsync(c->roll.p, 0, a0 + a1, c->stream));
    if (roll_svc) {
      d_svc = reinterpret_cast<fm_rollup_rec*>(static_cast<char*>(c->roll.p) + a0);
      d_api_service = reinterpret_cast<uint32_t*>(static_cast<char*>(c->roll.p) + a0 + a1);
      FM_CK(c, cudaMemcpyAsync(d_api_service, b->api_service, (size_t)b->n_apis * 4, cudaMemcpyHostToDevice, c->stream));
    }
  }
  while (c->kev.size() < 2 * n_chunks) {
    cudaEvent_t e; FM_CK(c, cudaEventCreate(&e)); c->kev.push_back(e);
  }

  uint32_t launches = 0;
  uint64_t h2d = 0, d2h = 0;
  // an error inside the loop must not leave copies in flight on buffers the caller owns
  auto drain = [&]() { cudaStreamSynchronize(c->s_h2d); cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->s_d2h); };
#define FM_CKL(call)                                                                  \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      drain();                                                                        \
      return fail(c, e__ == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_CUDA,       \
                  "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__,  \
                  __LINE__);                                                          \
    }                                                                                 \
  } while (0)
  for (uint64_t ci = 0; ci < n_chunks; ++ci) {
    const int s = (int)(ci % kPipe);
    const uint64_t s0 = ci * chunk, n = std::min(chunk, S - s0);
    char* din = static_cast<char*>(c->in[s].p);
    char* dout = static_cast<char*>(c->out[s].p);
    // host -> device on the copy stream, once the slot's previous kernel is done
    if (ci >= kPipe) FM_CKL(cudaStreamWaitEvent(c->s_h2d, c->in_free[s], 0));
    auto up = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
      h2d += bytes;
      return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->s_h2d);
    };
    FM_CKL(up(din + o_cur, b->cur + s0 * b->cur_pitch, n * b->cur_pitch * 4));
    if (!cached) {
      if (canary) FM_CKL(up(din + o_base, b->base + s0 * b->base_pitch, n * b->base_pitch * 4));
      if (b->hist) FM_CKL(up(din + o_hist, b->hist + s0 * b->hist_pitch, n * b->hist_pitch * 4));
    } else {
      // rows the cache does not hold yet: one copy per run of consecutive rows whose slots are
      // consecutive too (a cold cache hands out consecutive slots, so runs are long); rows that
      // found no slot go straight into the chunk buffer
      uint64_t i = kc.any_miss ? 0 : n;
      while (i < n) {
        if (!miss[s0 + i]) { ++i; continue; }
        const uint32_t sl0 = slot[s0 + i];
        uint64_t j = i + 1;
        if (sl0 == RowCache::kNone) { while (j < n && miss[s0 + j] && slot[s0 + j] == RowCache::kNone) ++j; }
        else { while (j < n && miss[s0 + j] && slot[s0 + j] == sl0 + (uint32_t)(j - i)) ++j; }
        const uint64_t r = j - i;
        if (canary) {
          float* dst = sl0 == RowCache::kNone ? reinterpret_cast<float*>(din + o_base) + i * b->base_pitch
                                              : kc.d_base + (uint64_t)sl0 * b->base_pitch;
          FM_CKL(up(dst, b->base + (s0 + i) * b->base_pitch, r * b->base_pitch * 4));
        }
        if (b->hist) {
          float* dst = sl0 == RowCache::kNone ? reinterpret_cast<float*>(din + o_hist) + i * b->hist_pitch
                                              : kc.d_hist + (uint64_t)sl0 * b->hist_pitch;
          FM_CKL(up(dst, b->hist + (s0 + i) * b->hist_pitch, r * b->hist_pitch * 4));
        }
        i = j;
      }
      FM_CKL(up(din + o_slot, kc.pin_slot + s0, n * 4));
    }
    if (b->cur_len) FM_CKL(up(din + o_clen, b->cur_len + s0, n * 4));
    if (canary && b->base_len) FM_CKL(up(din + o_blen, h_blen + s0, n * 4));
    if (b->hist && b->hist_len) FM_CKL(up(din + o_hlen, h_hlen + s0, n * 4));
    if (b->metric_type) FM_CKL(up(din + o_mt, b->metric_type + s0, n));
    if (roll) FM_CKL(up(din + o_api, b->api_id + s0, n * 4));
    FM_CKL(cudaEventRecord(c->in_ready[s], c->s_h2d));

    // score on the compute stream
    fm_batch db = *b;
    db.n_series = n;
    db.cur = reinterpret_cast<const float*>(din + o_cur);
    db.base = canary ? reinterpret_cast<const float*>(din + o_base) : nullptr;
    db.hist = b->hist ? reinterpret_cast<const float*>(din + o_hist) : nullptr;
    db.cur_len = b->cur_len ? reinterpret_cast<const uint32_t*>(din + o_clen) : nullptr;
    db.base_len = (canary && b->base_len) ? reinterpret_cast<const uint32_t*>(din + o_blen) : nullptr;
    db.hist_len = (b->hist && b->hist_len) ? reinterpret_cast<const uint32_t*>(din + o_hlen) : nullptr;
    db.metric_type = b->metric_type ? reinterpret_cast<const uint8_t*>(din + o_mt) : nullptr;
    db.api_id = roll ? reinterpret_cast<const uint32_t*>(din + o_api) : nullptr;
    Plan pl;
    rc = plan_score(c, &db, &pl);
    if (rc != FM_OK) { drain(); return rc; }
    pl.a.out = dout + o_rec;
    pl.a.aux = out_aux ? reinterpret_cast<fm_rank_aux*>(dout + o_aux) : nullptr;
    pl.a.api = roll ? d_api : nullptr;
    FM_CKL(cudaStreamWaitEvent(c->stream, c->in_ready[s], 0));
    if (ci >= kPipe) FM_CKL(cudaStreamWaitEvent(c->stream, c->out_free[s], 0));
    FM_CKL(cudaEventRecord(c->kev[2 * ci], c->stream));
    if (cached) {
      const uint32_t* d_slot = reinterpret_cast<const uint32_t*>(din + o_slot);
      if ((pl.kind == 4 || pl.kind == 5 || pl.kind == 6) && !c->force_gather) {
        // these kernels read resident rows in place (fm_base_row / fm_hist_row)
        pl.a.row_slot = d_slot; pl.a.cache_base = kc.d_base; pl.a.cache_hist = kc.d_hist;
      } else {   // resident rows -> the chunk's row matrices (HBM to HBM)
        if (canary) launches += fm_launch_gather_rows(kc.d_base, b->base_pitch, d_slot, reinterpret_cast<float*>(din + o_base), n, c->stream);
        if (b->hist) launches += fm_launch_gather_rows(kc.d_hist, b->hist_pitch, d_slot, reinterpret_cast<float*>(din + o_hist), n, c->stream);
      }
    }
    launches += launch_plan(pl, c->stream);
    FM_CKL(cudaGetLastError());
    FM_CKL(cudaEventRecord(c->kev[2 * ci + 1], c->stream));
    FM_CKL(cudaEventRecord(c->in_free[s], c->stream));
    FM_CKL(cudaEventRecord(c->out_ready[s], c->stream));

    // device -> host on the read-back stream
    FM_CKL(cudaStreamWaitEvent(c->s_d2h, c->out_ready[s], 0));
    FM_CKL(cudaMemcpyAsync(static_cast<char*>(out) + s0 * rsz, dout + o_rec, n * rsz, cudaMemcpyDeviceToHost, c->s_d2h));
    d2h += n * rsz;
    if (out_aux) {
      FM_CKL(cudaMemcpyAsync(out_aux + s0, dout + o_aux, n * sizeof(fm_rank_aux), cudaMemcpyDeviceToHost, c->s_d2h));
      d2h += n * sizeof(fm_rank_aux);
    }
    FM_CKL(cudaEventRecord(c->out_free[s], c->s_d2h));
  }
  if (roll) {
    launches += fm_launch_rollup_finalize(d_api, b->n_apis, d_api_service, d_svc, b->n_services, c->stream);
    FM_CKL(cudaGetLastError());
    FM_CKL(cudaMemcpyAsync(out_api, d_api, (size_t)b->n_apis * sizeof(fm_rollup_rec), cudaMemcpyDeviceToHost, c->stream));
    d2h += (size_t)b->n_apis * sizeof(fm_rollup_rec);
    if (roll_svc) {
      FM_CKL(cudaMemcpyAsync(out_svc, d_svc, (size_t)b->n_services * sizeof(fm_rollup_rec), cudaMemcpyDeviceToHost, c->stream));
      d2h += (size_t)b->n_services * sizeof(fm_rollup_rec);
    }
  }
#undef FM_CKL
  cudaError_t e1 = cudaStreamSynchronize(c->s_h2d), e2 = cudaStreamSynchronize(c->stream), e3 = cudaStreamSynchronize(c->s_d2h);
  FM_CK(c, e1); FM_CK(c, e2); FM_CK(c, e3);
  if (timing) {
    float tot = 0.f;
    for (uint64_t ci = 0; ci < n_chunks; ++ci) {
      float ms = 0.f;
      FM_CK(c, cudaEventElapsedTime(&ms, c->kev[2 * ci], c->kev[2 * ci + 1]));
      tot += ms;
    }
    timing->kernel_ms = tot; timing->main_kernel_ms = tot; timing->launches = launches;
    timing->h2d_bytes = h2d; timing->d2h_bytes = d2h;
  }
  return FM_OK;
}

}  // namespace

extern "C" {

int32_t fm_score_batch(fm_ctx* c, const fm_batch* b, void* out, fm_rank_aux* out_aux,
                       fm_rollup_rec* out_api, fm_rollup_rec* out_svc, fm_timing* timing) {
  return score_batch_impl(c, b, nullptr, out, out_aux, out_api, out_svc, timing);
}

int32_t fm_score_batch_cached(fm_ctx* c, const fm_batch* b, const uint64_t* row_ids, void* out, fm_rank_aux* out_aux,
                              fm_rollup_rec* out_api, fm_rollup_rec* out_svc, fm_timing* timing) {
  if (!c) return FM_E_INVALID;
  if (!row_ids && b && b->n_series) return fail(c, FM_E_INVALID, "fm_score_batch_cached: row_ids is NULL");
  return score_batch_impl(c, b, row_ids, out, out_aux, out_api, out_svc, timing);
}

int32_t fm_cache_configure(fm_ctx* c, uint64_t max_bytes) {
  if (!c) return FM_E_INVALID;
  FM_CK(c, cudaSetDevice(c->device));
  FM_CK(c, cudaStreamSynchronize(c->stream));
  for (int i = 0; i < 2; ++i) {
    RowCache& k = c->cache[i];
    k.release();
    k.max_bytes = max_bytes;
    k.hits = k.misses = k.evictions = k.uncached = 0;
  }
  return FM_OK;
}

int32_t fm_cache_get_stats(fm_ctx* c, fm_cache_stats* out) {
  if (!c || !out) return FM_E_INVALID;
  memset(out, 0, sizeof(*out));
  out->max_bytes = c->cache[0].max_bytes;
  for (int i = 0; i < 2; ++i) {
    const RowCache& k = c->cache[i];
    out->capacity_rows += k.cap; out->resident_rows += k.n_live;
    out->hits += k.hits; out->misses += k.misses; out->evictions += k.evictions; out->uncached += k.uncached;
    out->row_bytes = std::max<uint64_t>(out->row_bytes, 4ull * ((uint64_t)k.base_pitch + k.hist_pitch));
  }
  return FM_OK;
}

int32_t fm_cache_query(fm_ctx* c, const uint64_t* row_ids, uint64_t n, uint8_t* resident_out) {
  if (!c) return FM_E_INVALID;
  if (n && (!row_ids || !resident_out)) return fail(c, FM_E_INVALID, "fm_cache_query: null pointer");
  for (uint64_t i = 0; i < n; ++i) {
    resident_out[i] = 0;
    for (int q = 0; q < 2; ++q) {
      const RowCache& k = c->cache[q];
      if (k.cap && k.index.find(row_ids[i]) != k.index.end()) resident_out[i] = 1;
    }
  }
  return FM_OK;
}

// ---------------------------------------------------------------------------
static int32_t plan_groups(fm_ctx* c, const fm_group_batch* b, FmGArgs* a, int* wpc, int* grid) {
  memset(a, 0, sizeof(*a));
  if (!b || b->struct_size != sizeof(fm_group_batch)) return fail(c, FM_E_INVALID, "bad fm_group_batch");
  if (!b->group_offsets || (!b->cur && b->n_groups)) return fail(c, FM_E_INVALID, "null group_offsets / cur");
  if ((uint64_t)b->n_cur * FM_MAX_GROUPS > FM_MAX_RANK_N) return fail(c, FM_E_UNSUPPORTED, "n_cur %u too long for %u groups", b->n_cur, FM_MAX_GROUPS);
  if (c->params.g.ma_mode == FM_MODEL_HOLT_WINTERS) return fail(c, FM_E_UNSUPPORTED, "FM_MODEL_HOLT_WINTERS is a series-batch model (no pod-group form)");
  if ((b->cur_pitch & 3u) || (b->hist && (b->hist_pitch & 3u)) || (reinterpret_cast<uintptr_t>(b->hist) & 15u) ||
      (reinterpret_cast<uintptr_t>(b->cur) & 15u) || b->cur_pitch < ((b->n_cur + 3u) & ~3u) ||
      (b->hist && b->hist_pitch < ((b->n_hist + 3u) & ~3u)))
    return fail(c, FM_E_INVALID, "rows must be 16-byte aligned with pitch a multiple of 4 and >= roundup4(n)");
  a->raw_cap = std::max(64u, b->n_cur * FM_MAX_GROUPS);
  a->pool_cap = pow2_ceil(a->raw_cap, 64u);
  a->warp_bytes = (uint32_t)((((size_t)a->raw_cap + a->pool_cap) * 4 + 127) & ~size_t(127));
  const size_t limit = c->smem_optin ? c->smem_optin : 232448;
  if (a->warp_bytes > limit) return fail(c, FM_E_UNSUPPORTED, "groups too long for shared memory");
  *wpc = (int)std::max<size_t>(1, std::min<size_t>(16, limit / a->warp_bytes));
  if (c->groups_smem_set == 0) { FM_CK(c, fm_groups_smem_attr((int)limit)); c->groups_smem_set = 1; }
  const uint64_t need = (b->n_services + *wpc - 1) / *wpc;
  int ctas = (int)std::max<size_t>(1, std::min<size_t>(4, (233472 - 1024) / ((size_t)*wpc * a->warp_bytes + 1024)));
  *grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * ctas, need));
  a->group_offsets = b->group_offsets;
  a->cur = b->cur; a->cur_len = b->cur_len; a->n_cur = b->n_cur; a->cur_pitch = b->cur_pitch;
  a->hist = b->hist; a->hist_len = b->hist_len; a->n_hist = b->n_hist; a->hist_pitch = b->hist_pitch;
  a->metric_type = b->metric_type;
  a->n_services = b->n_services;
  a->P = c->d_params;
  return FM_OK;
}

// Kernels of one device-resident group batch on the compute stream (no events, no wait).
static int32_t groups_launch(fm_ctx* c, const fm_group_batch* b, fm_series_rec* d_out, fm_rollup_rec* d_svc,
                             cudaEvent_t ev_main_done, uint32_t* launches) {
  FmGArgs a; int wpc = 1, grid = 1;
  int32_t rc = plan_groups(c, b, &a, &wpc, &grid);
  if (rc != FM_OK) return rc;
  if (b->n_services == 0) return FM_OK;
  a.out = d_out; a.svc = d_svc;
  if (d_svc) FM_CK(c, cudaMemsetAsync(d_svc, 0, b->n_services * sizeof(fm_rollup_rec), c->stream));
  if (b->n_cur <= 64 && c->force_path != 1) {
    // register kernel for services of <= 8 groups; the rest (listed by it) go to the general one
    const size_t al_n = (b->n_services * 4 + 255) & ~size_t(255);
    FM_CK(c, c->misc.reserve(al_n + 256));
    uint32_t* d_list = static_cast<uint32_t*>(c->misc.p);
    uint32_t* d_cnt = reinterpret_cast<uint32_t*>(static_cast<char*>(c->misc.p) + al_n);
    FM_CK(c, cudaMemsetAsync(d_cnt, 0, 4, c->stream));
    int gw = 32, gc = 1;                                           // persistent grid: warps per CTA, CTAs per SM
    fm_groups_sort_shape(&gw, &gc);
    const uint64_t need = (b->n_services + gw - 1) / gw;
    const int g_fast = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->num_sms * gc, need));
    *launches += fm_launch_groups_sort(a, d_list, d_cnt, g_fast, c->stream);
    a.list = d_list; a.list_n = d_cnt;
    *launches += fm_launch_groups(a, wpc, std::min(grid, c->num_sms), c->stream);
  } else {
    *launches += fm_launch_groups(a, wpc, grid, c->stream);
  }
  if (ev_main_done) FM_CK(c, cudaEventRecord(ev_main_done, c->stream));
  *launches += fm_launch_rollup_finalize_one(d_svc, b->n_services, c->stream);
  FM_CK(c, cudaGetLastError());
  return FM_OK;
}

int32_t fm_score_groups_device(fm_ctx* c, const fm_group_batch* b, fm_series_rec* d_out,
                               fm_rollup_rec* d_svc, fm_timing* timing) {
  if (!c) return FM_E_INVALID;
  if (!d_out) return fail(c, FM_E_INVALID, "null output");
  FM_CK(c, cudaSetDevice(c->device));
  if (timing) memset(timing, 0, sizeof(*timing));
  uint32_t launches = 0;
  if (timing) FM_CK(c, cudaEventRecord(c->ev_a, c->stream));
  int32_t rc = groups_launch(c, b, d_out, d_svc, timing ? c->ev_m : nullptr, &launches);
  if (rc != FM_OK) return rc;
  if (timing && b->n_services) {
    FM_CK(c, cudaEventRecord(c->ev_b, c->stream));
    FM_CK(c, cudaEventSynchronize(c->ev_b));
    FM_CK(c, cudaEventElapsedTime(&timing->kernel_ms, c->ev_a, c->ev_b));
    FM_CK(c, cudaEventElapsedTime(&timing->main_kernel_ms, c->ev_a, c->ev_m));
    timing->launches = launches;
  }
  return FM_OK;
}

// Host rows: the batch is cut at service boundaries into chunks of ~64 MB of rows; the upload of
// chunk i+1 (copy stream) runs beside the kernels of chunk i (compute stream) and the read-back of
// chunk i-1 (second copy stream), kPipe buffer slots as in fm_score_batch.
int32_t fm_score_groups(fm_ctx* c, const fm_group_batch* b, fm_series_rec* out, fm_rollup_rec* out_svc,
                        fm_timing* timing) {
  if (!c) return FM_E_INVALID;
  if (!b || !out) return fail(c, FM_E_INVALID, "null batch or output");
  if (b->struct_size != sizeof(fm_group_batch)) return fail(c, FM_E_INVALID, "bad fm_group_batch.struct_size");
  FM_CK(c, cudaSetDevice(c->device));
  if (timing) memset(timing, 0, sizeof(*timing));
  auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
  const uint64_t V = b->n_services, G = b->n_groups;
  // the kernels index the output by these offsets: they must walk [0, n_groups] without a step back
  if (!b->group_offsets) return fail(c, FM_E_INVALID, "null group_offsets");
  if (b->group_offsets[0] != 0u || b->group_offsets[V] != G) return fail(c, FM_E_INVALID, "group_offsets must start at 0 and end at n_groups");
  for (uint64_t v = 0; v < V; ++v) {
    if (b->group_offsets[v + 1] < b->group_offsets[v]) return fail(c, FM_E_INVALID, "group_offsets must be non-decreasing (service %llu)", (unsigned long long)v);
    if (b->group_offsets[v + 1] - b->group_offsets[v] > FM_MAX_GROUPS) return fail(c, FM_E_UNSUPPORTED, "service %llu has more than %u groups", (unsigned long long)v, FM_MAX_GROUPS);
  }
  if (G && !b->cur) return fail(c, FM_E_INVALID, "null cur");
  if (V == 0) return FM_OK;

  // chunks: [first service, first group) of each, offsets rebased to the chunk
  const size_t svc_bytes = b->hist ? (size_t)b->hist_pitch * 4 : 0, grp_bytes = (size_t)b->cur_pitch * 4;
  const size_t target = size_t(c->chunk_mb ? c->chunk_mb : 64) << 20;
  std::vector<uint64_t> cut_v(1, 0);
  {
    size_t acc = 0;
    for (uint64_t v = 0; v < V; ++v) {
      acc += svc_bytes + grp_bytes * (b->group_offsets[v + 1] - b->group_offsets[v]);
      if (acc >= target && v + 1 < V) { cut_v.push_back(v + 1); acc = 0; }
    }
    cut_v.push_back(V);
  }
  const uint64_t n_chunks = cut_v.size() - 1;
  uint64_t max_v = 0, max_g = 0;
  std::vector<uint32_t> reb(V + n_chunks);                 // chunk ci: reb[cut_v[ci] + ci ..], V_c + 1 entries
  for (uint64_t ci = 0; ci < n_chunks; ++ci) {
    const uint64_t v0 = cut_v[ci], v1 = cut_v[ci + 1];
    const uint32_t g0 = b->group_offsets[v0];
    for (uint64_t v = v0; v <= v1; ++v) reb[v + ci] = b->group_offsets[v] - g0;
    max_v = std::max(max_v, v1 - v0);
    max_g = std::max<uint64_t>(max_g, b->group_offsets[v1] - g0);
  }
  size_t o = 0;
  const size_t o_off = o;  o += al((max_v + 1) * 4);
  const size_t o_cur = o;  o += al(max_g * b->cur_pitch * 4);
  const size_t o_clen = o; o += b->cur_len ? al(max_g * 4) : 0;
  const size_t o_hist = o; o += b->hist ? al(max_v * b->hist_pitch * 4) : 0;
  const size_t o_hlen = o; o += (b->hist && b->hist_len) ? al(max_v * 4) : 0;
  const size_t o_mt = o;   o += b->metric_type ? al(max_v) : 0;
  const size_t in_bytes = o;
  const size_t o_svc = al(max_g * sizeof(fm_series_rec));
  const size_t out_bytes = o_svc + al(max_v * sizeof(fm_rollup_rec));
  const int slots = (int)std::min<uint64_t>(n_chunks, kPipe);
  for (int i = 0; i < slots; ++i) { FM_CK(c, c->in[i].reserve(in_bytes)); FM_CK(c, c->out[i].reserve(out_bytes)); }
  while (c->kev.size() < 2 * n_chunks) {
    cudaEvent_t e; FM_CK(c, cudaEventCreate(&e)); c->kev.push_back(e);
  }
  uint32_t launches = 0;
  uint64_t h2d = 0, d2h = 0;
  auto drain = [&]() { cudaStreamSynchronize(c->s_h2d); cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->s_d2h); };
#define FM_CKL(call)                                                                  \
  do {                                                                                \
    cudaError_t e__ = (call);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      drain();                                                                        \
      return fail(c, e__ == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_CUDA,       \
                  "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__,  \
                  __LINE__);                                                          \
    }                                                                                 \
  } while (0)
  for (uint64_t ci = 0; ci < n_chunks; ++ci) {
    const int s = (int)(ci % kPipe);
    const uint64_t v0 = cut_v[ci], nv = cut_v[ci + 1] - v0;
    const uint64_t g0 = b->group_offsets[v0], ng = b->group_offsets[v0 + nv] - g0;
    char* din = static_cast<char*>(c->in[s].p);
    char* dout = static_cast<char*>(c->out[s].p);
    if (ci >= kPipe) FM_CKL(cudaStreamWaitEvent(c->s_h2d, c->in_free[s], 0));
    auto up = [&](size_t off, const void* src, size_t bytes) -> cudaError_t {
      h2d += bytes;
      return cudaMemcpyAsync(din + off, src, bytes, cudaMemcpyHostToDevice, c->s_h2d);
    };
    FM_CKL(up(o_off, reb.data() + v0 + ci, (nv + 1) * 4));
    if (ng) FM_CKL(up(o_cur, b->cur + g0 * b->cur_pitch, ng * b->cur_pitch * 4));
    if (b->cur_len && ng) FM_CKL(up(o_clen, b->cur_len + g0, ng * 4));
    if (b->hist) FM_CKL(up(o_hist, b->hist + v0 * b->hist_pitch, nv * b->hist_pitch * 4));
    if (b->hist && b->hist_len) FM_CKL(up(o_hlen, b->hist_len + v0, nv * 4));
    if (b->metric_type) FM_CKL(up(o_mt, b->metric_type + v0, nv));
    FM_CKL(cudaEventRecord(c->in_ready[s], c->s_h2d));

    fm_group_batch db = *b;
    db.n_services = nv; db.n_groups = ng;
    db.group_offsets = reinterpret_cast<const uint32_t*>(din + o_off);
    db.cur = reinterpret_cast<const float*>(din + o_cur);
    db.cur_len = b->cur_len ? reinterpret_cast<const uint32_t*>(din + o_clen) : nullptr;
    db.hist = b->hist ? reinterpret_cast<const float*>(din + o_hist) : nullptr;
    db.hist_len = (b->hist && b->hist_len) ? reinterpret_cast<const uint32_t*>(din + o_hlen) : nullptr;
    db.metric_type = b->metric_type ? reinterpret_cast<const uint8_t*>(din + o_mt) : nullptr;
    FM_CKL(cudaStreamWaitEvent(c->stream, c->in_ready[s], 0));
    if (ci >= kPipe) FM_CKL(cudaStreamWaitEvent(c->stream, c->out_free[s], 0));
    FM_CKL(cudaEventRecord(c->kev[2 * ci], c->stream));
    const int32_t rc = groups_launch(c, &db, reinterpret_cast<fm_series_rec*>(dout),
                                     out_svc ? reinterpret_cast<fm_rollup_rec*>(dout + o_svc) : nullptr, nullptr, &launches);
    if (rc != FM_OK) { drain(); return rc; }
    FM_CKL(cudaEventRecord(c->kev[2 * ci + 1], c->stream));
    FM_CKL(cudaEventRecord(c->in_free[s], c->stream));
    FM_CKL(cudaEventRecord(c->out_ready[s], c->stream));

    FM_CKL(cudaStreamWaitEvent(c->s_d2h, c->out_ready[s], 0));
    if (ng) FM_CKL(cudaMemcpyAsync(out + g0, dout, ng * sizeof(fm_series_rec), cudaMemcpyDeviceToHost, c->s_d2h));
    d2h += ng * sizeof(fm_series_rec);
    if (out_svc) {
      FM_CKL(cudaMemcpyAsync(out_svc + v0, dout + o_svc, nv * sizeof(fm_rollup_rec), cudaMemcpyDeviceToHost, c->s_d2h));
      d2h += nv * sizeof(fm_rollup_rec);
    }
    FM_CKL(cudaEventRecord(c->out_free[s], c->s_d2h));
  }
#undef FM_CKL
  cudaError_t e1 = cudaStreamSynchronize(c->s_h2d), e2 = cudaStreamSynchronize(c->stream), e3 = cudaStreamSynchronize(c->s_d2h);
  FM_CK(c, e1); FM_CK(c, e2); FM_CK(c, e3);
  if (timing) {
    float tot = 0.f;
    for (uint64_t ci = 0; ci < n_chunks; ++ci) {
      float ms = 0.f;
      FM_CK(c, cudaEventElapsedTime(&ms, c->kev[2 * ci], c->kev[2 * ci + 1]));
      tot += ms;
    }
    timing->kernel_ms = tot; timing->main_kernel_ms = tot; timing->launches = launches;
    timing->h2d_bytes = h2d; timing->d2h_bytes = d2h;
  }
  return FM_OK;
}

// ---------------------------------------------------------------------------
static int32_t validate_pairs(fm_ctx* c, const fm_pair_batch* b, const void* out) {
  if (!b || !out) return fail(c, FM_E_INVALID, "null pair batch or output");
  if (b->struct_size != sizeof(fm_pair_batch)) return fail(c, FM_E_INVALID, "fm_pair_batch.struct_size %u != %zu", b->struct_size, sizeof(fm_pair_batch));
  if (b->n_pairs && ((b->n_cur && (!b->cur_x || !b->cur_y)) || (b->n_hist && (!b->hist_x || !b->hist_y))))
    return fail(c, FM_E_INVALID, "null row pointer");
  if (b->n_cur > FM_MAX_CUR) return fail(c, FM_E_UNSUPPORTED, "n_cur %u > %u", b->n_cur, FM_MAX_CUR);
  auto bad = [](const float* p, uint32_t n, uint32_t pitch) {
    return p && ((reinterpret_cast<uintptr_t>(p) & 15u) || (pitch & 3u) || pitch < ((n + 3u) & ~3u));
  };
  if (bad(b->cur_x, b->n_cur, b->cur_pitch) || bad(b->cur_y, b->n_cur, b->cur_pitch) ||
      bad(b->hist_x, b->n_hist, b->hist_pitch) || bad(b->hist_y, b->n_hist, b->hist_pitch))
    return fail(c, FM_E_INVALID, "rows must be 16-byte aligned with pitch a multiple of 4 and >= roundup4(n)");
  return FM_OK;
}

int32_t fm_score_pairs_device(fm_ctx* c, const fm_pair_batch* b, fm_pair_rec* d_out, fm_timing* timing) {
  if (!c) return FM_E_INVALID;
  int32_t rc = validate_pairs(c, b, d_out);
  if (rc != FM_OK) return rc;
  FM_CK(c, cudaSetDevice(c->device));
  if (timing) memset(timing, 0, sizeof(*timing));
  if (b->n_pairs == 0) return FM_OK;
  if (timing) FM_CK(c, cudaEventRecord(c->ev_a, c->stream));
  const uint32_t launches = (uint32_t)fm_launch_pairs(*b, c->d_params, d_out, c->num_sms, c->stream);
  FM_CK(c, cudaGetLastError());
  if (timing) {
    FM_CK(c, cudaEventRecord(c->ev_b, c->stream));
    FM_CK(c, cudaEventSynchronize(c->ev_b));
    FM_CK(c, cudaEventElapsedTime(&timing->kernel_ms, c->ev_a, c->ev_b));
    timing->main_kernel_ms = timing->kernel_ms;
    timing->launches = launches;
  }
  return FM_OK;
}

int32_t fm_score_pairs(fm_ctx* c, const fm_pair_batch* b, fm_pair_rec* out, fm_timing* timing) {
  if (!c) return FM_E_INVALID;
  int32_t rc = validate_pairs(c, b, out);
  if (rc != FM_OK) return rc;
  FM_CK(c, cudaSetDevice(c->device));
  if (timing) memset(timing, 0, sizeof(*timing));
  const uint64_t P = b->n_pairs;
  if (P == 0) return FM_OK;
  auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t cb = al(P * b->cur_pitch * 4), hb = al(P * b->hist_pitch * 4);
  size_t o = 0;
  const size_t o_cx = o; o += cb;
  const size_t o_cy = o; o += cb;
  const size_t o_hx = o; o += hb;
  const size_t o_hy = o; o += hb;
  const size_t o_cl = o; o += b->cur_len ? al(P * 4) : 0;
  const size_t o_hl = o; o += b->hist_len ? al(P * 4) : 0;
  const size_t o_mt = o; o += b->metric_type ? al(P) : 0;
  FM_CK(c, c->in[0].reserve(o));
  FM_CK(c, c->out[0].reserve(al(P * sizeof(fm_pair_rec))));
  char* din = static_cast<char*>(c->in[0].p);
  uint64_t h2d = 0;
  auto up = [&](size_t off, const void* src, size_t bytes) -> cudaError_t {
    h2d += bytes;
    return bytes ? cudaMemcpyAsync(din + off, src, bytes, cudaMemcpyHostToDevice, c->stream) : cudaSuccess;
  };
  FM_CK(c, up(o_cx, b->cur_x, P * b->cur_pitch * 4));
  FM_CK(c, up(o_cy, b->cur_y, P * b->cur_pitch * 4));
  FM_CK(c, up(o_hx, b->hist_x, P * b->hist_pitch * 4));
  FM_CK(c, up(o_hy, b->hist_y, P * b->hist_pitch * 4));
  if (b->cur_len) FM_CK(c, up(o_cl, b->cur_len, P * 4));
  if (b->hist_len) FM_CK(c, up(o_hl, b->hist_len, P * 4));
  if (b->metric_type) FM_CK(c, up(o_mt, b->metric_type, P));
  fm_pair_batch db = *b;
  db.cur_x = reinterpret_cast<const float*>(din + o_cx); db.cur_y = reinterpret_cast<const float*>(din + o_cy);
  db.hist_x = reinterpret_cast<const float*>(din + o_hx); db.hist_y = reinterpret_cast<const float*>(din + o_hy);
  db.cur_len = b->cur_len ? reinterpret_cast<const uint32_t*>(din + o_cl) : nullptr;
  db.hist_len = b->hist_len ? reinterpret_cast<const uint32_t*>(din + o_hl) : nullptr;
  db.metric_type = b->metric_type ? reinterpret_cast<const uint8_t*>(din + o_mt) : nullptr;
  fm_timing t; memset(&t, 0, sizeof(t));
  rc = fm_score_pairs_device(c, &db, static_cast<fm_pair_rec*>(c->out[0].p), &t);
  if (rc != FM_OK) { cudaStreamSynchronize(c->stream); return rc; }
  FM_CK(c, cudaMemcpyAsync(out, c->out[0].p, P * sizeof(fm_pair_rec), cudaMemcpyDeviceToHost, c->stream));
  FM_CK(c, cudaStreamSynchronize(c->stream));
  if (timing) { *timing = t; timing->h2d_bytes = h2d; timing->d2h_bytes = P * sizeof(fm_pair_rec); }
  return FM_OK;
}

// ---------------------------------------------------------------------------
int32_t fm_rank_dump(fm_ctx* c, const float* cur, uint32_t n_cur, const float* base, uint32_t n_base,
                     uint32_t* r2_out) {
  if (!c) return FM_E_INVALID;
  const uint32_t N = n_cur + n_base;
  if (N == 0) return FM_OK;
  if (!r2_out || (n_cur && !cur) || (n_base && !base)) return fail(c, FM_E_INVALID, "null pointer");
  if (N > (1u << 16)) return fail(c, FM_E_UNSUPPORTED, "rank dump is a parity aid, limited to 65536 points");
  FM_CK(c, cudaSetDevice(c->device));
  const uint32_t cap = pow2_ceil(N, 64u);
  auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t o_cur = 0, o_base = al((size_t)n_cur * 4 + 4), o_sorted = o_base + al((size_t)n_base * 4 + 4),
               o_r2 = o_sorted + al((size_t)cap * 4);
  FM_CK(c, c->misc.reserve(o_r2 + al((size_t)N * 4)));
  char* d = static_cast<char*>(c->misc.p);
  if (n_cur) FM_CK(c, cudaMemcpyAsync(d + o_cur, cur, (size_t)n_cur * 4, cudaMemcpyHostToDevice, c->stream));
  if (n_base) FM_CK(c, cudaMemcpyAsync(d + o_base, base, (size_t)n_base * 4, cudaMemcpyHostToDevice, c->stream));
  fm_launch_rank_dump(reinterpret_cast<float*>(d + o_cur), n_cur, reinterpret_cast<float*>(d + o_base), n_base,
                      reinterpret_cast<float*>(d + o_sorted), cap, reinterpret_cast<uint32_t*>(d + o_r2), c->stream);
  FM_CK(c, cudaGetLastError());
  FM_CK(c, cudaMemcpyAsync(r2_out, d + o_r2, (size_t)N * 4, cudaMemcpyDeviceToHost, c->stream));
  FM_CK(c, cudaStreamSynchronize(c->stream));
  return FM_OK;
}

int32_t fm_ma_window_series(fm_ctx* c, const float* x, uint32_t n, uint32_t w, uint32_t ddof,
                            double* mean_out, double* sigma_out) {
  if (!c) return FM_E_INVALID;
  if (n == 0) return FM_OK;
  if (!x || !mean_out || !sigma_out) return fail(c, FM_E_INVALID, "null pointer");
  FM_CK(c, cudaSetDevice(c->device));
  auto al = [](size_t v) { return (v + 255) & ~size_t(255); };
  const size_t o_x = 0, o_p1 = al((size_t)n * 4), o_p2 = o_p1 + al((size_t)n * 8), o_m = o_p2 + al((size_t)n * 8),
               o_s = o_m + al((size_t)n * 8);
  FM_CK(c, c->misc.reserve(o_s + al((size_t)n * 8)));
  char* d = static_cast<char*>(c->misc.p);
  FM_CK(c, cudaMemcpyAsync(d + o_x, x, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
  fm_launch_ma_window(reinterpret_cast<float*>(d + o_x), n, w, ddof, reinterpret_cast<double*>(d + o_p1),
                      reinterpret_cast<double*>(d + o_p2), reinterpret_cast<double*>(d + o_m),
                      reinterpret_cast<double*>(d + o_s), c->stream);
  FM_CK(c, cudaGetLastError());
  FM_CK(c, cudaMemcpyAsync(mean_out, d + o_m, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
  FM_CK(c, cudaMemcpyAsync(sigma_out, d + o_s, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
  FM_CK(c, cudaStreamSynchronize(c->stream));
  return FM_OK;
}

// ---------------------------------------------------------------------------
int32_t fm_synth_fill_device(fm_ctx* c, const fm_synth_desc* d, float* d_cur, float* d_base,
                             float* d_hist, uint8_t* d_mt) {
  if (!c) return FM_E_INVALID;
  if (!d || d->struct_size != sizeof(fm_synth_desc)) return fail(c, FM_E_INVALID, "bad fm_synth_desc");
  FM_CK(c, cudaSetDevice(c->device));
  fm_launch_synth(*d, d_cur, d_base, d_hist, d_mt, c->stream);
  FM_CK(c, cudaGetLastError());
  return FM_OK;
}

int32_t fm_synth_fill_host(const fm_synth_desc* d, float* cur, float* base, float* hist,
                           uint8_t* mt, uint32_t n_threads) {
  if (!d || d->struct_size != sizeof(fm_synth_desc)) return FM_E_INVALID;
  if (n_threads == 0) n_threads = std::max(1u, std::thread::hardware_concurrency());
  n_threads = (uint32_t)std::min<uint64_t>(n_threads, std::max<uint64_t>(1, d->n_series));
  auto work = [&](uint64_t lo, uint64_t hi) {
    for (uint64_t r = lo; r < hi; ++r) {
      const uint64_t id = d->row_ids ? d->row_ids[r] : d->first_series + r;
      const fm_synth_series sp = fm_synth_series_params(d->seed, d->config, id, d->n_cur);
      if (cur) for (uint32_t t = 0; t < d->cur_pitch; ++t) cur[r * d->cur_pitch + t] = t < d->n_cur ? fm_synth_sample(d->seed, &sp, id, FM_ROW_CUR, t) : 0.0f;
      if (base) for (uint32_t t = 0; t < d->base_pitch; ++t) base[r * d->base_pitch + t] = t < d->n_base ? fm_synth_sample(d->seed, &sp, id, FM_ROW_BASE, t) : 0.0f;
      if (hist) for (uint32_t t = 0; t < d->hist_pitch; ++t) hist[r * d->hist_pitch + t] = t < d->n_hist ? fm_synth_sample(d->seed, &sp, id, FM_ROW_HIST, t) : 0.0f;
      if (mt) mt[r] = (uint8_t)sp.type;
    }
  };
  std::vector<std::thread> th;
  for (uint32_t t = 0; t < n_threads; ++t)
    th.emplace_back(work, d->n_series * t / n_threads, d->n_series * (t + 1) / n_threads);
  for (auto& x : th) x.join();
  return FM_OK;
}

}  // extern "C"
