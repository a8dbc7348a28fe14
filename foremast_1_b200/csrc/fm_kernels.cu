// fm_kernels.cu — sm_100a kernels of the scoring path.
//
// k_score_series: one warp per series, persistent over the batch.
//   a1  rows are staged into shared memory by the TMA unit (cp.async.bulk +
//       mbarrier), double buffered per warp: the next series lands while the
//       current one is ranked.  Historical rows that are only needed for the
//       model are streamed with 128-bit loads and never staged.
//   a2  pivoted fp64 moments, warp-shuffle reduction.
//   a4  ranks by sorting the baseline row in place (shared-memory bitonic) and
//       the 64-padded current row, then locating every current value in both:
//       2*midrank = 2*(#less) + (#equal) + 1.  Tie term from the run lengths.
//       Wilcoxon sorts the fp64 |differences|.
//   a5..a8  fp64 p-values, pairwise decision, thresholds, anomaly ballot.
//   a9  roll-up counters by atomics into the per-API output records.
#include <cstdint>
#include <mutex>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "../../include/fm_synth.h"
#include "fm_device.cuh"

namespace {

#define FM_INF CUDART_INF_F

struct WarpSmem {
  uint64_t* bar;      // [n_stages]
  float* cur[2];      // raw current rows
  float* base[2];     // baseline rows (sorted in place)
  float* cs;          // sorted current row
  double* wd;         // raw differences
  double* wk;         // sorted |differences|
};

__device__ __forceinline__ WarpSmem carve(unsigned char* p, const FmKArgs& a) {
  WarpSmem w;
  w.bar = reinterpret_cast<uint64_t*>(p);
  p += 16;
  for (uint32_t s = 0; s < a.n_stages; ++s) {
    w.cur[s] = reinterpret_cast<float*>(p);  p += (size_t)a.cur_cap * 4;
    w.base[s] = reinterpret_cast<float*>(p); p += (size_t)a.base_cap * 4;
  }
  if (a.n_stages == 1) { w.cur[1] = w.cur[0]; w.base[1] = w.base[0]; }
  w.cs = reinterpret_cast<float*>(p); p += (size_t)a.cs_cap * 4;
  w.wd = reinterpret_cast<double*>(p); p += (size_t)a.wx_n * 8;
  w.wk = reinterpret_cast<double*>(p);
  return w;
}

__device__ __forceinline__ uint32_t row_len(const uint32_t* len, uint64_t s, uint32_t n) {
  return len ? min(__ldg(len + s), n) : n;
}

// lane 0: arm the barrier and start the bulk copies of series s into stage st
__device__ __forceinline__ void issue_stage(const FmKArgs& a, const WarpSmem& w, uint32_t st,
                                            uint64_t s, uint32_t nc, uint32_t nb) {
  const uint32_t cb = ((nc + 3u) & ~3u) * 4u, bb = ((nb + 3u) & ~3u) * 4u;
  fm_mbar_expect_tx(&w.bar[st], cb + bb);
  if (cb) fm_bulk_g2s(w.cur[st], a.cur + s * a.cur_pitch, cb, &w.bar[st]);
  if (bb) fm_bulk_g2s(w.base[st], a.base + s * a.base_pitch, bb, &w.bar[st]);
}

__global__ void __launch_bounds__(512) k_score_series(const FmKArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int wpc = blockDim.x >> 5;
  const WarpSmem w = carve(smem + (size_t)warp * a.warp_bytes, a);
  const fm_global_params& g = a.P->g;
  const bool canary = (a.strategy == FM_STRAT_CANARY) && a.base != nullptr;

  if (lane == 0) {
    for (uint32_t s = 0; s < a.n_stages; ++s) fm_mbar_init(&w.bar[s], 1);
    fm_fence_mbar_init();
  }
  __syncwarp();

  const uint64_t stride = (uint64_t)gridDim.x * wpc;
  uint64_t s = (uint64_t)blockIdx.x * wpc + warp;
  uint32_t phase[2] = {0u, 0u};
  uint32_t st = 0;

  // prologue: first series of this warp
  if (s < a.n_series && lane == 0) {
    issue_stage(a, w, 0, s, row_len(a.cur_len, s, a.n_cur),
                canary ? row_len(a.base_len, s, a.n_base) : 0u);
  }

  for (; s < a.n_series; s += stride) {
    const uint32_t nc = row_len(a.cur_len, s, a.n_cur);
    const uint32_t nb = canary ? row_len(a.base_len, s, a.n_base) : 0u;
    const uint64_t sn = s + stride;

    // prefetch the next series into the other stage (its previous contents were
    // written through the generic proxy: fence before the async proxy reuses it)
    if (a.n_stages == 2 && sn < a.n_series) {
      fm_fence_proxy_async();
      __syncwarp();
      if (lane == 0)
        issue_stage(a, w, st ^ 1u, sn, row_len(a.cur_len, sn, a.n_cur),
                    canary ? row_len(a.base_len, sn, a.n_base) : 0u);
    }
    fm_mbar_wait(&w.bar[st], phase[st]);
    phase[st] ^= 1u;

    float* cur = w.cur[st];
    float* base = w.base[st];
    const uint32_t ty = a.metric_type ? a.metric_type[s] : 0u;
    const fm_type_params tp = a.P->types[ty < a.P->n_types ? ty : 0u];

    fm_series_rec rec;
    fm_rec_init(rec);
    fm_rank_aux ax = {0, 0, 0, 0, 0, 0u, 0u};
    uint32_t status = 0;

    // ---- a2: model window ------------------------------------------------
    bool model_ok;
    const bool holt_winters = g.ma_mode == FM_MODEL_HOLT_WINTERS;
    FmHw hw = {0.0, nullptr, 0u, 0u};
    if (holt_winters) {
      double* season = a.hw_scratch ? a.hw_scratch + ((uint64_t)blockIdx.x * wpc + warp) * g.season_len : nullptr;
      if (a.hist) model_ok = fm_model_hw(g, a.hist + s * a.hist_pitch, row_len(a.hist_len, s, a.n_hist), lane, season, rec.mean, rec.sigma, hw);
      else model_ok = fm_model_hw(g, canary ? base : nullptr, nb, lane, season, rec.mean, rec.sigma, hw);
    } else if (a.hist) {
      const uint32_t nh = row_len(a.hist_len, s, a.n_hist);
      model_ok = fm_model(g, a.hist + s * a.hist_pitch, nh, true, lane, rec.mean, rec.sigma);
    } else {
      model_ok = fm_model(g, canary ? base : nullptr, nb, false, lane, rec.mean, rec.sigma);
    }
    if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
    if (nc == 0) status |= FM_ST_NO_CURRENT;

    // ---- a4..a6: pairwise tests ------------------------------------------
    bool differ = false;
    if (canary && g.pairwise_mode != FM_PW_NONE && nc > 0 && nb > 0) {
      FmTest tests[3];
      tests[0] = tests[1] = tests[2] = FmTest{false, false, 0.0, 1.0};
      const bool want_mw = fm_mode_has(g.pairwise_mode, FM_PW_MANN_WHITE) && nc >= g.min_mw && nb >= g.min_mw;
      const bool want_kw = fm_mode_has(g.pairwise_mode, FM_PW_KRUSKAL) && nc >= g.min_kw && nb >= g.min_kw;
      const bool want_wx = fm_mode_has(g.pairwise_mode, FM_PW_WILCOXON);

      // Wilcoxon first: it needs the baseline row in time order
      if (want_wx) {
        const uint32_t n = min(nc, nb);
        const uint32_t oc = g.pair_rule == FM_PAIR_TAIL ? nc - n : 0u;
        const uint32_t ob = g.pair_rule == FM_PAIR_TAIL ? nb - n : 0u;
        const uint32_t n2 = fm_pow2_ceil(n, 64u);
        uint32_t nz = 0;
        for (uint32_t i = lane; i < n2; i += 32) {
          double d = 0.0, k = __longlong_as_double(0x7ff0000000000000ll);
          if (i < n) {
            d = __dsub_rn((double)cur[oc + i], (double)base[ob + i]);
            k = fabs(d);
            nz += (d == 0.0);
            w.wd[i] = d;
          }
          w.wk[i] = k;
        }
        nz = fm_warp_sum(nz);
        const uint32_t np = n - nz;
        ax.n_wx = np;
        __syncwarp();
        if (np >= g.min_wx && np > 0) {
          fm_warp_bitonic<double>(w.wk, n2, lane);
          long long wp = 0, wm = 0;
          for (uint32_t i = lane; i < n; i += 32) {
            const double d = w.wd[i];
            if (d != 0.0) {
              const double k = fabs(d);
              const uint32_t lo = fm_lower_bound<double>(w.wk, n, k) - nz;
              const uint32_t hi = fm_upper_bound<double>(w.wk, n, k) - nz;
              const long long r2 = (long long)(lo + hi + 1u);
              if (d > 0.0) wp += r2; else wm += r2;
            }
          }
          long long T = fm_tie_term_sorted<double>(w.wk, nz, n, lane);
          wp = fm_warp_sum(wp); wm = fm_warp_sum(wm); T = fm_warp_sum(T);
          ax.w2_plus = wp; ax.w2_minus = wm; ax.t_wx = T;
          tests[1] = fm_wilcoxon(wp, wm, T, np, g.cc_wx);
          rec.w = tests[1].stat; rec.p_w = tests[1].p; status |= FM_ST_WX_RUN;
          if (tests[1].degenerate) status |= FM_ST_DEGENERATE;
        }
        __syncwarp();
      }

      if (want_mw || want_kw) {
        // sorted copy of the current row, and the baseline row sorted in place;
        // -0 is folded onto +0 so that equal values are equal keys
        const uint32_t c2 = fm_pow2_ceil(nc, 64u), b2 = fm_pow2_ceil(nb, 64u);
        for (uint32_t i = lane; i < c2; i += 32) w.cs[i] = i < nc ? cur[i] + 0.0f : FM_INF;
        for (uint32_t i = lane; i < b2; i += 32) base[i] = i < nb ? base[i] + 0.0f : FM_INF;
        __syncwarp();
        fm_warp_bitonic<float>(w.cs, c2, lane);
        fm_warp_bitonic<float>(base, b2, lane);

        // every distinct current value: its run in cur (p..p+tc) and in base (lb..lb+tb)
        long long r2c = 0, T = 0;
        for (uint32_t p = lane; p < nc; p += 32) {
          const float x = w.cs[p];
          if (p == 0 || w.cs[p - 1] != x) {
            const uint32_t tc = fm_upper_bound<float>(w.cs, nc, x) - p;
            const uint32_t lb = fm_lower_bound<float>(base, nb, x);
            const uint32_t tb = fm_upper_bound<float>(base, nb, x) - lb;
            r2c += (long long)tc * (long long)(2u * (lb + p) + tb + tc + 1u);
            T += fm_tie(tc + tb) - fm_tie(tb);
          }
        }
        T += fm_tie_term_sorted<float>(base, 0u, nb, lane);
        r2c = fm_warp_sum(r2c); T = fm_warp_sum(T);
        ax.r2_cur = r2c; ax.t_mw = T;
        if (want_mw) {
          tests[0] = fm_mann_whitney(r2c, T, nc, nb, g.cc_mw);
          rec.u1 = tests[0].stat; rec.p_u = tests[0].p; status |= FM_ST_MW_RUN;
          if (tests[0].degenerate) status |= FM_ST_DEGENERATE;
        }
        if (want_kw) {
          const double N = (double)nc + (double)nb;
          const double rx = 0.5 * (double)r2c;
          const double ry = 0.5 * (double)((long long)(nc + nb) * ((long long)(nc + nb) + 1) - r2c);
          const double sres = rx * rx / (double)nc + ry * ry / (double)nb;
          tests[2] = fm_kruskal(sres, N, T, 2u);
          rec.h = tests[2].stat; rec.p_h = tests[2].p; status |= FM_ST_KW_RUN;
          if (tests[2].degenerate) status |= FM_ST_DEGENERATE;
        }
      }
      bool skipped = false;
      differ = fm_decide(g.pairwise_mode, g.alpha, tests, 3, skipped);
      if (skipped) status |= FM_ST_PAIRWISE_SKIPPED;
      if (differ) status |= FM_ST_DIFFER;
    }

    // ---- a7, a8 -----------------------------------------------------------
    if (model_ok && nc > 0) {
      if (holt_winters) fm_threshold_hw(g, tp, differ, cur, nc, lane, hw, rec);
      else fm_threshold(g, tp, differ, cur, nc, lane, rec);
    }
    if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT)) rec.verdict = FM_UNKNOWN;
    rec.status = (uint8_t)status;

    if (lane == 0) {
      fm_write_record(a.out, s, a.record_kind, rec);
      if (a.aux) a.aux[s] = ax;
      if (a.api && a.api_id) fm_rollup_add(a.api, a.api_id[s], a.n_apis, rec.verdict);
    }
    __syncwarp();

    if (a.n_stages == 2) {
      st ^= 1u;
    } else if (sn < a.n_series) {
      fm_fence_proxy_async();
      __syncwarp();
      if (lane == 0)
        issue_stage(a, w, 0, sn, row_len(a.cur_len, sn, a.n_cur),
                    canary ? row_len(a.base_len, sn, a.n_base) : 0u);
    }
  }
}

// ---------------------------------------------------------------------------
// Rollover strategy: Kruskal-Wallis over the pod groups of a service
// (BASELINE.json configs[3]); one warp per service.
__global__ void __launch_bounds__(512) k_score_groups(const FmGArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int wpc = blockDim.x >> 5;
  unsigned char* p = smem + (size_t)warp * a.warp_bytes;
  float* raw = reinterpret_cast<float*>(p);
  float* pool = raw + a.raw_cap;
  const fm_global_params& g = a.P->g;

  const uint64_t n_iter = a.list ? (uint64_t)*a.list_n : a.n_services;
  for (uint64_t it = (uint64_t)blockIdx.x * wpc + warp; it < n_iter; it += (uint64_t)gridDim.x * wpc) {
    const uint64_t v = a.list ? (uint64_t)a.list[it] : it;
    const uint32_t g0 = a.group_offsets[v], g1 = a.group_offsets[v + 1];
    const uint32_t K = g1 - g0;
    const uint32_t ty = a.metric_type ? a.metric_type[v] : 0u;
    const fm_type_params tp = a.P->types[ty < a.P->n_types ? ty : 0u];

    double mean = 0.0, sigma = 0.0;
    bool model_ok = false;
    if (a.hist) {
      const uint32_t nh = row_len(a.hist_len, v, a.n_hist);
      model_ok = fm_model(g, a.hist + v * a.hist_pitch, nh, true, lane, mean, sigma);
    }

    FmTest kw = {false, false, 0.0, 1.0};
    uint32_t status_kw = 0;
    bool differ = false;
    if (g.pairwise_mode == FM_PW_FRIEDMAN) {
      uint32_t n = K ? row_len(a.cur_len, g0, a.n_cur) : 0u;
      bool eligible = K >= 3 && K <= FM_MAX_GROUPS && n >= g.min_kw && n > 0;
      for (uint32_t k = 1; eligible && k < K; ++k)
        if (row_len(a.cur_len, g0 + k, a.n_cur) != n) eligible = false;
      if (eligible) {
        kw = fm_friedman(a.cur + (uint64_t)g0 * a.cur_pitch, a.cur_pitch, K, n, lane);
        status_kw |= FM_ST_KW_RUN;
        if (kw.degenerate) status_kw |= FM_ST_DEGENERATE;
        else differ = kw.p < g.alpha;
      }
      if (!(status_kw & FM_ST_KW_RUN) || kw.degenerate) status_kw |= FM_ST_PAIRWISE_SKIPPED;
      if (differ) status_kw |= FM_ST_DIFFER;
    }
    if (g.pairwise_mode != FM_PW_NONE && fm_mode_has(g.pairwise_mode, FM_PW_KRUSKAL)) {
      bool eligible = K >= 2 && K <= FM_MAX_GROUPS;
      uint32_t N = 0;
      for (uint32_t k = 0; eligible && k < K; ++k) {
        const uint32_t n = row_len(a.cur_len, g0 + k, a.n_cur);
        if (n < g.min_kw || n == 0) eligible = false;
        N += n;
      }
      if (eligible && N <= a.raw_cap) {
        // pooled sample in group order (raw) and sorted (pool)
        const uint32_t n2 = fm_pow2_ceil(N, 64u);
        uint32_t o = 0;
        for (uint32_t k = 0; k < K; ++k) {
          const uint32_t n = row_len(a.cur_len, g0 + k, a.n_cur);
          const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
          for (uint32_t i = lane; i < n; i += 32) { const float x = row[i] + 0.0f; raw[o + i] = x; pool[o + i] = x; }
          o += n;
        }
        for (uint32_t i = N + lane; i < n2; i += 32) pool[i] = FM_INF;
        __syncwarp();
        fm_warp_bitonic<float>(pool, n2, lane);
        double sres = 0.0;
        o = 0;
        for (uint32_t k = 0; k < K; ++k) {
          const uint32_t n = row_len(a.cur_len, g0 + k, a.n_cur);
          long long r2 = 0;
          for (uint32_t i = lane; i < n; i += 32) {
            const float x = raw[o + i];
            r2 += (long long)(fm_lower_bound<float>(pool, N, x) + fm_upper_bound<float>(pool, N, x) + 1u);
          }
          r2 = fm_warp_sum(r2);
          const double R = 0.5 * (double)r2;
          sres += R * R / (double)n;
          o += n;
        }
        long long T = fm_warp_sum(fm_tie_term_sorted<float>(pool, 0u, N, lane));
        kw = fm_kruskal(sres, (double)N, T, K);
        status_kw |= FM_ST_KW_RUN;
        if (kw.degenerate) status_kw |= FM_ST_DEGENERATE;
        else differ = kw.p < g.alpha;
        __syncwarp();
      }
      if (!(status_kw & FM_ST_KW_RUN) || kw.degenerate) status_kw |= FM_ST_PAIRWISE_SKIPPED;
      if (differ) status_kw |= FM_ST_DIFFER;
    }

    for (uint32_t k = 0; k < K; ++k) {
      fm_series_rec rec;
      fm_rec_init(rec);
      rec.h = kw.stat; rec.p_h = kw.p; rec.mean = mean; rec.sigma = sigma;
      uint32_t status = status_kw;
      const uint32_t nc = row_len(a.cur_len, g0 + k, a.n_cur);
      if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
      if (nc == 0) status |= FM_ST_NO_CURRENT;
      if (model_ok && nc > 0)
        fm_threshold(g, tp, differ, a.cur + (uint64_t)(g0 + k) * a.cur_pitch, nc, lane, rec);
      if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT)) rec.verdict = FM_UNKNOWN;
      rec.status = (uint8_t)status;
      if (lane == 0) {
        a.out[g0 + k] = rec;
        if (a.svc) fm_rollup_add(a.svc, (uint32_t)v, 0xFFFFFFFFu, rec.verdict);
      }
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// a9: severity from the accumulated child counters, then service level.
__global__ void k_rollup_severity(fm_rollup_rec* r, uint64_t n, const uint32_t* parent,
                                  fm_rollup_rec* up, uint32_t n_up) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  fm_rollup_rec x = r[i];
  const uint32_t sev = x.n_unhealthy > 0 ? FM_UNHEALTHY
                     : (x.n_children > 0 && x.n_unknown == x.n_children) ? FM_UNKNOWN : FM_HEALTHY;
  r[i].severity = (uint8_t)sev;
  if (up && parent) fm_rollup_add(up, parent[i], n_up, sev);
}

// ---------------------------------------------------------------------------
// Parity aid: 2*midrank of every point of cur U base for ONE series.
__global__ void __launch_bounds__(32) k_rank_dump(const float* cur, uint32_t nc, const float* base,
                                                  uint32_t nb, float* sorted, uint32_t cap,
                                                  uint32_t* r2) {
  const int lane = threadIdx.x;
  const uint32_t N = nc + nb;
  for (uint32_t i = lane; i < cap; i += 32)
    sorted[i] = i < nc ? cur[i] + 0.0f : (i < N ? base[i - nc] + 0.0f : FM_INF);
  __syncwarp();
  // global-memory bitonic by one warp: only used on small parity cases
  fm_warp_bitonic<float>(sorted, cap, lane);
  for (uint32_t i = lane; i < N; i += 32) {
    const float x = (i < nc ? cur[i] : base[i - nc]) + 0.0f;
    r2[i] = fm_lower_bound<float>(sorted, N, x) + fm_upper_bound<float>(sorted, N, x) + 1u;
  }
}

// ---------------------------------------------------------------------------
// A.2 trailing-window mean / sigma series for ONE row by pivoted prefix sums.
// One CTA; p1/p2 are global scratch of n doubles each.
__global__ void __launch_bounds__(1024) k_ma_window(const float* x, uint32_t n, uint32_t w,
                                                    uint32_t ddof, double* p1, double* p2,
                                                    double* mean_out, double* sigma_out) {
  __shared__ double c1[1024], c2[1024];
  const uint32_t t = threadIdx.x, nt = blockDim.x;
  const uint32_t chunk = (n + nt - 1) / nt;
  const uint32_t lo = min(t * chunk, n), hi = min(lo + chunk, n);
  const double pivot = n ? (double)x[0] : 0.0;
  double s1 = 0.0, s2 = 0.0;
  for (uint32_t i = lo; i < hi; ++i) { const double d = (double)x[i] - pivot; s1 += d; s2 = fma(d, d, s2); }
  c1[t] = s1; c2[t] = s2;
  __syncthreads();
  if (t == 0) {
    double a1 = 0.0, a2 = 0.0;
    for (uint32_t k = 0; k < nt; ++k) { const double b1 = c1[k], b2 = c2[k]; c1[k] = a1; c2[k] = a2; a1 += b1; a2 += b2; }
  }
  __syncthreads();
  s1 = c1[t]; s2 = c2[t];
  for (uint32_t i = lo; i < hi; ++i) { const double d = (double)x[i] - pivot; s1 += d; s2 = fma(d, d, s2); p1[i] = s1; p2[i] = s2; }
  __threadfence_block();
  __syncthreads();
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  for (uint32_t i = t; i < n; i += nt) {
    if (w == 0 || i + 1 < w) { mean_out[i] = nan; sigma_out[i] = nan; continue; }
    const double q1 = p1[i] - (i >= w ? p1[i - w] : 0.0);
    const double q2 = p2[i] - (i >= w ? p2[i - w] : 0.0);
    const double md = q1 / (double)w;
    double m2 = q2 - (double)w * md * md;
    if (m2 < 0.0) m2 = 0.0;
    mean_out[i] = pivot + md;
    sigma_out[i] = w > ddof ? sqrt(m2 / (double)(w - ddof)) : 0.0;
  }
}

// ---------------------------------------------------------------------------
// Synthetic rows (include/fm_synth.h): one thread per sample, rows coalesced.
__global__ void k_synth_rows(fm_synth_desc d, float* out, uint32_t n, uint32_t pitch, uint32_t kind) {
  const uint64_t row = blockIdx.y + (uint64_t)blockIdx.z * gridDim.y;
  if (row >= d.n_series) return;
  const uint64_t id = d.row_ids ? d.row_ids[row] : d.first_series + row;
  const fm_synth_series sp = fm_synth_series_params(d.seed, d.config, id, d.n_cur);
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < pitch; t += gridDim.x * blockDim.x)
    out[row * pitch + t] = t < n ? fm_synth_sample(d.seed, &sp, id, kind, t) : 0.0f;
}
// rows resident in the cache -> the chunk's row matrix: one CTA per row, 128-bit copies; rows
// without a slot were uploaded into place by the host
__global__ void __launch_bounds__(128) k_gather_rows(const float4* __restrict__ cache, uint32_t pitch4,
                                                     const uint32_t* __restrict__ slot, float4* __restrict__ dst) {
  const uint32_t sl = slot[blockIdx.x];
  if (sl == 0xFFFFFFFFu) return;
  const float4* src = cache + (uint64_t)sl * pitch4;
  float4* d = dst + (uint64_t)blockIdx.x * pitch4;
  for (uint32_t i = threadIdx.x; i < pitch4; i += 128) d[i] = fm_ldg_stream(src + i);
}

__global__ void k_synth_types(fm_synth_desc d, uint8_t* mt) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < d.n_series) mt[i] = (uint8_t)fm_synth_type(d.config, d.row_ids ? d.row_ids[i] : d.first_series + i);
}

}  // namespace

// ------------------------------------------------------------- launchers ---
cudaError_t fm_score_smem_attr(int bytes) {
  return cudaFuncSetAttribute(k_score_series, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
cudaError_t fm_groups_smem_attr(int bytes) {
  return cudaFuncSetAttribute(k_score_groups, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

int fm_launch_score(const FmKArgs& a, int warps_per_cta, int grid, cudaStream_t st) {
  k_score_series<<<grid, warps_per_cta * 32, (size_t)warps_per_cta * a.warp_bytes, st>>>(a);
  return 1;
}

int fm_launch_groups(const FmGArgs& a, int warps_per_cta, int grid, cudaStream_t st) {
  k_score_groups<<<grid, warps_per_cta * 32, (size_t)warps_per_cta * a.warp_bytes, st>>>(a);
  return 1;
}

int fm_launch_rollup_finalize(fm_rollup_rec* api, uint32_t n_apis, const uint32_t* api_service,
                              fm_rollup_rec* svc, uint32_t n_services, cudaStream_t st) {
  int n = 0;
  if (api && n_apis) {
    k_rollup_severity<<<(n_apis + 255) / 256, 256, 0, st>>>(api, n_apis, svc ? api_service : nullptr, svc, n_services);
    ++n;
  }
  if (svc && n_services) {
    k_rollup_severity<<<(n_services + 255) / 256, 256, 0, st>>>(svc, n_services, nullptr, nullptr, 0u);
    ++n;
  }
  return n;
}

int fm_launch_gather_rows(const float* cache, uint32_t pitch, const uint32_t* d_slot, float* dst, uint64_t n,
                          cudaStream_t st) {
  if (!n || !pitch) return 0;
  k_gather_rows<<<(unsigned)n, 128, 0, st>>>(reinterpret_cast<const float4*>(cache), pitch >> 2, d_slot,
                                             reinterpret_cast<float4*>(dst));
  return 1;
}

int fm_launch_rollup_finalize_one(fm_rollup_rec* recs, uint64_t n, cudaStream_t st) {
  if (!recs || !n) return 0;
  k_rollup_severity<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(recs, n, nullptr, nullptr, 0u);
  return 1;
}

int fm_launch_rank_dump(const float* d_cur, uint32_t n_cur, const float* d_base, uint32_t n_base,
                        float* d_sorted, uint32_t cap, uint32_t* d_r2, cudaStream_t st) {
  k_rank_dump<<<1, 32, 0, st>>>(d_cur, n_cur, d_base, n_base, d_sorted, cap, d_r2);
  return 1;
}

int fm_launch_ma_window(const float* d_x, uint32_t n, uint32_t w, uint32_t ddof, double* d_p1,
                        double* d_p2, double* d_mean, double* d_sigma, cudaStream_t st) {
  k_ma_window<<<1, 1024, 0, st>>>(d_x, n, w, ddof, d_p1, d_p2, d_mean, d_sigma);
  return 1;
}

// ---------------------------------------------------------------------------
// fm_ma_window_series_static: k_ma_window for callers that may not allocate
// (see include/foremast_cuda.h).  The two prefix-sum arrays that the ctx-based
// entry point takes from a cudaMalloc'ed arena live here, in a fixed array that
// is part of the library image on every device.
//
// g_ma_static below is a HOST-side handle in host code, not a device address:
// the kernel must be given the address cudaGetSymbolAddress returns for the
// device that is current, so that lookup is done once per device and cached.
static __device__ __align__(256) double g_ma_static[2 * FM_STATIC_MA_MAX_POINTS];

namespace {
constexpr int kMaxStaticDevices = 64;
std::mutex g_ma_static_mu;
double* g_ma_static_addr[kMaxStaticDevices] = {};
}  // namespace

extern "C" int32_t fm_ma_window_series_static(int32_t device, void* stream, const float* d_x,
                                              uint32_t n, uint32_t w, uint32_t ddof,
                                              double* d_mean_out, double* d_sigma_out) {
  if (device < 0 || device >= kMaxStaticDevices) {
    fm_set_thread_error("fm_ma_window_series_static: device %d out of range", device);
    return FM_E_INVALID;
  }
  if (n == 0) return FM_OK;
  if (!d_x || !d_mean_out || !d_sigma_out) {
    fm_set_thread_error("fm_ma_window_series_static: null device pointer");
    return FM_E_INVALID;
  }
  if (n > FM_STATIC_MA_MAX_POINTS) {
    fm_set_thread_error("fm_ma_window_series_static: n = %u exceeds the built-in scratch of %u points; "
                        "use fm_ma_window_series (which allocates) for longer rows",
                        n, FM_STATIC_MA_MAX_POINTS);
    return FM_E_UNSUPPORTED;
  }
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) {
    fm_set_thread_error("fm_ma_window_series_static: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    return FM_E_CUDA;
  }
  double* scratch;
  {
    // set-up, once per device, with that device current
    std::lock_guard<std::mutex> lk(g_ma_static_mu);
    if (!g_ma_static_addr[device]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_ma_static);
      if (e != cudaSuccess || !p) {
        fm_set_thread_error("fm_ma_window_series_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        return FM_E_CUDA;
      }
      g_ma_static_addr[device] = static_cast<double*>(p);
    }
    scratch = g_ma_static_addr[device];
  }
  // entries [0, n) of both halves are rewritten by the kernel before it reads them
  k_ma_window<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(
      d_x, n, w, ddof, scratch, scratch + FM_STATIC_MA_MAX_POINTS, d_mean_out, d_sigma_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    fm_set_thread_error("fm_ma_window_series_static: launch: %s", cudaGetErrorString(e));
    return FM_E_CUDA;
  }
  return FM_OK;
}

int fm_launch_synth(const fm_synth_desc& d, float* cur, float* base, float* hist, uint8_t* mt,
                    cudaStream_t st) {
  int n = 0;
  const unsigned gy = (unsigned)(d.n_series < 32768 ? d.n_series : 32768);
  const unsigned gz = (unsigned)((d.n_series + gy - 1) / (gy ? gy : 1));
  if (!d.n_series) return 0;
  if (cur && d.n_cur) { k_synth_rows<<<dim3((d.cur_pitch + 255) / 256, gy, gz), 256, 0, st>>>(d, cur, d.n_cur, d.cur_pitch, FM_ROW_CUR); ++n; }
  if (base && d.n_base) { k_synth_rows<<<dim3((d.base_pitch + 255) / 256, gy, gz), 256, 0, st>>>(d, base, d.n_base, d.base_pitch, FM_ROW_BASE); ++n; }
  if (hist && d.n_hist) { k_synth_rows<<<dim3((d.hist_pitch + 255) / 256, gy, gz), 256, 0, st>>>(d, hist, d.n_hist, d.hist_pitch, FM_ROW_HIST); ++n; }
  if (mt) { k_synth_types<<<(unsigned)((d.n_series + 255) / 256), 256, 0, st>>>(d, mt); ++n; }
  return n;
}
