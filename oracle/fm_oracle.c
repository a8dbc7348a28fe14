/*
 * fm_oracle.c — scalar fp64 CPU restatement of the scoring path.
 * TEST INFRASTRUCTURE ONLY; PARITY UNPINNED upstream — see fm_oracle.h.
 *
 * Section numbers (A.x) refer to DESIGN.md section 3 / SURVEY.md appendix A.
 */
#define _GNU_SOURCE
#include "fm_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/fm_synth.h"

/* ---------------------------------------------------------------- A.1 --- */
/* README.md:22 (ML_ALGORITHM = moving_average_all, foremast-brain.yaml:24-25):
 * mean and sigma of the whole model window, two passes in fp64. */
void fmo_moments(const float* x, uint32_t n, uint32_t ddof, double* mean, double* sigma) {
  if (n == 0) { *mean = 0.0; *sigma = 0.0; return; }
  double s = 0.0;
  for (uint32_t i = 0; i < n; ++i) s += (double)x[i];
  const double mu = s / (double)n;
  double m2 = 0.0;
  for (uint32_t i = 0; i < n; ++i) { const double d = (double)x[i] - mu; m2 += d * d; }
  *mean = mu;
  *sigma = (n > ddof) ? sqrt(m2 / (double)(n - ddof)) : 0.0;
}

/* ---------------------------------------------------------------- A.2 --- */
void fmo_ma_window_series(const float* x, uint32_t n, uint32_t w, uint32_t ddof,
                          double* mean_out, double* sigma_out) {
  for (uint32_t t = 0; t < n; ++t) {
    if (w == 0 || t + 1 < w) { mean_out[t] = NAN; sigma_out[t] = NAN; continue; }
    fmo_moments(x + (t + 1 - w), w, ddof, &mean_out[t], &sigma_out[t]);
  }
}

/* ---------------------------------------------------------------- A.3 --- */
typedef struct { double v; uint32_t i; } fmo_kv;

static int fmo_kv_cmp(const void* a, const void* b) {
  const fmo_kv* p = (const fmo_kv*)a; const fmo_kv* q = (const fmo_kv*)b;
  if (p->v < q->v) return -1;
  if (p->v > q->v) return 1;
  return (p->i > q->i) - (p->i < q->i);
}

/* bottom-up merge sort on (value, index); -0 == +0 by IEEE compare */
static void fmo_sort_kv(fmo_kv* a, fmo_kv* tmp, uint32_t n) {
  if (n < 64) { qsort(a, n, sizeof(fmo_kv), fmo_kv_cmp); return; }
  /* insertion-sort runs of 16, then merge */
  for (uint32_t lo = 0; lo < n; lo += 16) {
    const uint32_t hi = lo + 16 < n ? lo + 16 : n;
    for (uint32_t i = lo + 1; i < hi; ++i) {
      fmo_kv k = a[i]; uint32_t j = i;
      while (j > lo && fmo_kv_cmp(&a[j - 1], &k) > 0) { a[j] = a[j - 1]; --j; }
      a[j] = k;
    }
  }
  fmo_kv* src = a; fmo_kv* dst = tmp;
  for (uint32_t width = 16; width < n; width *= 2) {
    for (uint32_t lo = 0; lo < n; lo += 2 * width) {
      uint32_t mid = lo + width < n ? lo + width : n;
      uint32_t hi = lo + 2 * width < n ? lo + 2 * width : n;
      uint32_t i = lo, j = mid, k = lo;
      while (i < mid && j < hi) dst[k++] = (fmo_kv_cmp(&src[j], &src[i]) < 0) ? src[j++] : src[i++];
      while (i < mid) dst[k++] = src[i++];
      while (j < hi) dst[k++] = src[j++];
    }
    fmo_kv* t = src; src = dst; dst = t;
  }
  if (src != a) memcpy(a, src, (size_t)n * sizeof(fmo_kv));
}

typedef struct { fmo_kv* a; fmo_kv* tmp; double* v; uint32_t* r2; uint32_t cap; } fmo_ws;

static int fmo_ws_reserve(fmo_ws* w, uint32_t n) {
  if (n <= w->cap) return 0;
  free(w->a); free(w->tmp); free(w->v); free(w->r2);
  w->a = (fmo_kv*)malloc((size_t)n * sizeof(fmo_kv));
  w->tmp = (fmo_kv*)malloc((size_t)n * sizeof(fmo_kv));
  w->v = (double*)malloc((size_t)n * sizeof(double));
  w->r2 = (uint32_t*)malloc((size_t)n * sizeof(uint32_t));
  w->cap = n;
  return (w->a && w->tmp && w->v && w->r2) ? 0 : -1;
}
static void fmo_ws_free(fmo_ws* w) { free(w->a); free(w->tmp); free(w->v); free(w->r2); memset(w, 0, sizeof(*w)); }

/* Midranks: a maximal run of equal values at 1-based sorted positions i..j
 * gets rank (i+j)/2, stored as r2 = i+j; tie term = sum over runs of t^3 - t. */
static void fmo_midranks_ws(fmo_ws* w, const double* v, uint32_t n, uint32_t* r2, int64_t* tie_term) {
  int64_t T = 0;
  for (uint32_t i = 0; i < n; ++i) { w->a[i].v = v[i]; w->a[i].i = i; }
  fmo_sort_kv(w->a, w->tmp, n);
  uint32_t lo = 0;
  while (lo < n) {
    uint32_t hi = lo + 1;
    while (hi < n && w->a[hi].v == w->a[lo].v) ++hi;
    const uint32_t t = hi - lo;
    const uint32_t rr = (lo + 1) + hi; /* first + last 1-based position */
    for (uint32_t k = lo; k < hi; ++k) r2[w->a[k].i] = rr;
    T += (int64_t)t * t * t - (int64_t)t;
    lo = hi;
  }
  *tie_term = T;
}

void fmo_midranks(const double* v, uint32_t n, uint32_t* r2, int64_t* tie_term) {
  fmo_ws w; memset(&w, 0, sizeof(w));
  if (n == 0) { *tie_term = 0; return; }
  fmo_ws_reserve(&w, n);
  fmo_midranks_ws(&w, v, n, r2, tie_term);
  fmo_ws_free(&w);
}

void fmo_rank_dump(const float* cur, uint32_t n_cur, const float* base, uint32_t n_base,
                   uint32_t* r2_out) {
  const uint32_t n = n_cur + n_base;
  if (n == 0) return;
  double* v = (double*)malloc((size_t)n * sizeof(double));
  for (uint32_t i = 0; i < n_cur; ++i) v[i] = (double)cur[i];
  for (uint32_t i = 0; i < n_base; ++i) v[n_cur + i] = (double)base[i];
  int64_t T;
  fmo_midranks(v, n, r2_out, &T);
  free(v);
}

/* ---------------------------------------------------------------- A.5 --- */
/* Chi-square survival function for integer dof, closed form.
 * dof = 2m:   Q = exp(-x/2) * sum_{j<m} (x/2)^j / j!
 * dof = 2m+1: Q = erfc(sqrt(x/2)) + exp(-x/2) * sqrt(2x/pi) * sum_{j=1..m} x^(j-1)/(2j-1)!! */
double fmo_chi2_sf(double x, uint32_t dof) {
  if (!(x > 0.0) || dof == 0) return 1.0;
  const double hx = 0.5 * x;
  const double e = exp(-hx);
  if ((dof & 1u) == 0u) {
    const uint32_t m = dof / 2;
    double term = 1.0, sum = 1.0;
    for (uint32_t j = 1; j < m; ++j) { term *= hx / (double)j; sum += term; }
    return e * sum;
  }
  const uint32_t m = dof / 2;
  double q = erfc(sqrt(hx));
  if (m > 0) {
    double term = 1.0, sum = 1.0; /* j = 1: x^0 / 1!! */
    for (uint32_t j = 2; j <= m; ++j) { term *= x / (double)(2 * j - 1); sum += term; }
    q += e * sqrt(2.0 * x / 3.14159265358979323846) * sum;
  }
  return q;
}

/* ---------------------------------------------------------- A.4 helpers --- */
static int fmo_mode_has(uint32_t mode, uint32_t test) {
  return mode == test || mode == FM_PW_ALL || mode == FM_PW_ANY;
}

typedef struct { int run; int degenerate; double stat; double p; } fmo_test;

/* Mann-Whitney from the exact integers (design.md:89-92 lists the tests;
 * README.md:36 MIN_MANN_WHITE_DATA_POINTS). */
static fmo_test fmo_mann_whitney(int64_t r2_cur, int64_t T, uint32_t n1, uint32_t n2, uint32_t cc) {
  fmo_test t = {1, 0, 0.0, 1.0};
  const double N = (double)n1 + (double)n2;
  const double u1 = 0.5 * (double)(r2_cur - (int64_t)n1 * ((int64_t)n1 + 1));
  const double u2 = (double)n1 * (double)n2 - u1;
  const double u = u1 > u2 ? u1 : u2;
  const double mu = 0.5 * (double)n1 * (double)n2;
  const double var = ((double)n1 * (double)n2 / 12.0) * ((N + 1.0) - (double)T / (N * (N - 1.0)));
  t.stat = u1;
  if (!(var > 0.0)) { t.degenerate = 1; return t; }
  const double z = (u - mu - (cc ? 0.5 : 0.0)) / sqrt(var);
  const double p = erfc(z / 1.4142135623730951);
  t.p = p > 1.0 ? 1.0 : p;
  return t;
}

static fmo_test fmo_wilcoxon(int64_t w2p, int64_t w2m, int64_t T, uint32_t np, uint32_t cc) {
  fmo_test t = {1, 0, 0.0, 1.0};
  const double w = 0.5 * (double)(w2p < w2m ? w2p : w2m);
  const double n = (double)np;
  const double mu = n * (n + 1.0) / 4.0;
  const double var = n * (n + 1.0) * (2.0 * n + 1.0) / 24.0 - (double)T / 48.0;
  t.stat = w;
  if (!(var > 0.0)) { t.degenerate = 1; return t; }
  double num = w - mu;
  if (cc && num != 0.0) num += 0.5;
  const double z = num / sqrt(var);
  const double p = erfc(fabs(z) / 1.4142135623730951);
  t.p = p > 1.0 ? 1.0 : p;
  return t;
}

/* Kruskal-Wallis from group rank sums (2*R_g) and sizes. */
static fmo_test fmo_kruskal(const int64_t* r2_g, const uint32_t* n_g, uint32_t K, int64_t T) {
  fmo_test t = {1, 0, 0.0, 1.0};
  double N = 0.0;
  for (uint32_t g = 0; g < K; ++g) N += (double)n_g[g];
  double s = 0.0;
  for (uint32_t g = 0; g < K; ++g) {
    const double R = 0.5 * (double)r2_g[g];
    s += R * R / (double)n_g[g];
  }
  const double h0 = 12.0 / (N * (N + 1.0)) * s - 3.0 * (N + 1.0);
  const double c = 1.0 - (double)T / (N * N * N - N);
  if (!(c > 0.0)) { t.degenerate = 1; return t; }
  t.stat = h0 / c;
  t.p = fmo_chi2_sf(t.stat, K - 1);
  return t;
}

/* A.6: combine the evaluated, non-degenerate tests. */
static int fmo_decide(uint32_t mode, double alpha, const fmo_test* tests, int n_tests, int* skipped) {
  int n_e = 0, any = 0, all = 1;
  for (int i = 0; i < n_tests; ++i) {
    if (!tests[i].run || tests[i].degenerate) continue;
    ++n_e;
    const int d = tests[i].p < alpha;
    any |= d; all &= d;
  }
  *skipped = (n_e == 0);
  if (n_e == 0) return 0;
  return mode == FM_PW_ALL ? all : any; /* single-test modes have one candidate */
}

/* A.7 + A.8: bounds, anomalous points, verdict (design.md:35,43). */
static void fmo_threshold(const fm_global_params* g, const fm_type_params* tp, int differ,
                          const float* cur, uint32_t n_cur, fm_series_rec* rec) {
  const double k = differ ? tp->k_low : tp->k;
  const double up = rec->mean + k * rec->sigma;
  double lo = rec->mean - k * rec->sigma;
  if (lo < tp->min_lower_bound) lo = tp->min_lower_bound;
  rec->upper = up; rec->lower = lo;
  uint32_t n_anom = 0; uint16_t first = 0xFFFFu; uint64_t mask = 0;
  for (uint32_t i = 0; i < n_cur; ++i) {
    const double x = (double)cur[i];
    const int a = ((tp->bound_mask & 1u) && x > up) || ((tp->bound_mask & 2u) && x < lo);
    if (a) {
      if (n_anom == 0) first = (uint16_t)(i < 0xFFFFu ? i : 0xFFFEu);
      ++n_anom;
      if (i < 64) mask |= 1ull << i;
    }
  }
  rec->n_anom = n_anom; rec->first_idx = first; rec->anom_mask = mask;
  rec->verdict = (n_anom >= g->min_anomaly_points && n_anom > 0) ? FM_UNHEALTHY : FM_HEALTHY;
}

/* model window -> mean/sigma according to ma_mode; returns 0 when insufficient */
static int fmo_model(const fm_global_params* g, const float* x, uint32_t n, double* mean, double* sigma) {
  uint32_t need = g->min_hist_points > g->ddof + 1 ? g->min_hist_points : g->ddof + 1;
  *mean = 0.0; *sigma = 0.0;
  if (g->ma_mode == FM_MODEL_HOLT_WINTERS) return 0;   /* series batches only: fmo_model_hw */
  if (g->ma_mode == FM_MA_WINDOW) {
    const uint32_t w = g->ma_window;
    if (w > need) need = w;
    if (n < need || w == 0) return 0;
    fmo_moments(x + (n - w), w, g->ddof, mean, sigma);
    return 1;
  }
  if (n < need) return 0;
  if (g->ma_mode == FM_MODEL_EWMA || g->ma_mode == FM_MODEL_DOUBLE_EXP) {
    /* design.md:60-66 "Exponential Smoothing", "Double Exponential Smoothing":
     * one-step-ahead forecast as the centre, RMS forecast error as sigma. */
    const double a = g->smooth_alpha, b = g->smooth_beta;
    double level = (double)x[0], trend = 0.0, sse = 0.0;
    const int dbl = g->ma_mode == FM_MODEL_DOUBLE_EXP;
    if (dbl && n > 1) trend = (double)x[1] - (double)x[0];
    uint32_t m = 0;
    for (uint32_t t = 1; t < n; ++t) {
      const double f = level + (dbl ? trend : 0.0);
      const double e = (double)x[t] - f;
      sse += e * e; ++m;
      const double nl = f + a * e;
      if (dbl) trend = trend + b * (nl - level - trend);
      level = nl;
    }
    *mean = level + (dbl ? trend : 0.0);
    *sigma = (m > g->ddof) ? sqrt(sse / (double)(m - g->ddof)) : 0.0;
    return 1;
  }
  fmo_moments(x, n, g->ddof, mean, sigma);
  return 1;
}

/* A.2c additive Holt-Winters (docs/guides/design.md:60-66 lists it; the recurrence is the
 * textbook one, declared in DESIGN.md section 3): season of m points; level = mean of the first
 * season, trend = (mean of the second season - level) / m, season_i = x_i - level; from t = m on the
 * one-step-ahead forecast level + trend + season[t mod m], its error, and the three updates.
 * Returns 0 when the window is shorter than two seasons.  season: m doubles, left holding the
 * final seasonal state (the caller frees it). */
static int fmo_model_hw(const fm_global_params* g, const float* x, uint32_t n, double* season,
                        double* c1, double* sigma, double* trend_out) {
  const uint32_t m = g->season_len;
  const uint32_t need = g->min_hist_points > g->ddof + 1 ? g->min_hist_points : g->ddof + 1;
  *c1 = 0.0; *sigma = 0.0; *trend_out = 0.0;
  if (m < 2 || n < 2 * m || n < need) return 0;
  double sa = 0.0, sb = 0.0;
  for (uint32_t i = 0; i < m; ++i) { sa += (double)x[i]; sb += (double)x[m + i]; }
  double level = sa / (double)m;
  double trend = (sb / (double)m - level) / (double)m;
  for (uint32_t i = 0; i < m; ++i) season[i] = (double)x[i] - level;
  const double a = g->smooth_alpha, b = g->smooth_beta, c = g->smooth_gamma;
  double sse = 0.0;
  for (uint32_t t = m; t < n; ++t) {
    const uint32_t slot = t % m;
    const double xt = (double)x[t], si = season[slot];
    const double e = xt - (level + trend + si);
    sse += e * e;
    const double nl = a * (xt - si) + (1.0 - a) * (level + trend);
    trend = b * (nl - level) + (1.0 - b) * trend;
    season[slot] = c * (xt - nl) + (1.0 - c) * si;
    level = nl;
  }
  const uint32_t cnt = n - m;
  *trend_out = trend;
  *c1 = level + trend + season[n % m];
  *sigma = cnt > g->ddof ? sqrt(sse / (double)(cnt - g->ddof)) : 0.0;
  return 1;
}

/* A.7 + A.8 for the seasonal model: point i is compared after its own centre has been moved onto
 * the first one's: x_i - (centre_i - centre_0) against centre_0 +- k sigma. */
static void fmo_threshold_hw(const fm_global_params* g, const fm_type_params* tp, int differ,
                             const float* cur, uint32_t n_cur, const double* season, uint32_t n, uint32_t m,
                             double trend, fm_series_rec* rec) {
  const double k = differ ? tp->k_low : tp->k;
  const double up = rec->mean + k * rec->sigma;
  double lo = rec->mean - k * rec->sigma;
  if (lo < tp->min_lower_bound) lo = tp->min_lower_bound;
  rec->upper = up; rec->lower = lo;
  const double s0 = season[n % m];
  uint32_t n_anom = 0; uint16_t first = 0xFFFFu; uint64_t mask = 0;
  for (uint32_t i = 0; i < n_cur; ++i) {
    const double off = (double)i * trend + (season[(n + i) % m] - s0);
    const double x = (double)cur[i] - off;
    const int a = ((tp->bound_mask & 1u) && x > up) || ((tp->bound_mask & 2u) && x < lo);
    if (a) {
      if (n_anom == 0) first = (uint16_t)(i < 0xFFFFu ? i : 0xFFFEu);
      ++n_anom;
      if (i < 64) mask |= 1ull << i;
    }
  }
  rec->n_anom = n_anom; rec->first_idx = first; rec->anom_mask = mask;
  rec->verdict = (n_anom >= g->min_anomaly_points && n_anom > 0) ? FM_UNHEALTHY : FM_HEALTHY;
}

static void fmo_score_series_ws(fmo_ws* ws, const fm_global_params* g, const fm_type_params* tp,
                                uint32_t strategy, const float* cur, uint32_t n_cur,
                                const float* base, uint32_t n_base, const float* hist,
                                uint32_t n_hist, fm_series_rec* rec, fm_rank_aux* aux) {
  memset(rec, 0, sizeof(*rec));
  rec->p_u = rec->p_w = rec->p_h = 1.0;
  rec->first_idx = 0xFFFFu;
  fm_rank_aux ax; memset(&ax, 0, sizeof(ax));
  uint32_t status = 0;
  const int canary = (strategy == FM_STRAT_CANARY) && base != NULL;
  if (!canary) { base = NULL; n_base = 0; }

  /* a2: the model window is the historical row, or the baseline row when no
   * historical row is sent (n_hist' = 0 in config 3). */
  const float* mx = hist ? hist : base;
  const uint32_t mn = hist ? n_hist : n_base;
  const int holt_winters = g->ma_mode == FM_MODEL_HOLT_WINTERS;
  double* season = NULL; double hw_trend = 0.0;
  int model_ok;
  if (holt_winters) {
    season = (g->season_len >= 2 && g->season_len <= (1u << 20)) ? (double*)malloc((size_t)g->season_len * sizeof(double)) : NULL;
    model_ok = (mx && season) ? fmo_model_hw(g, mx, mn, season, &rec->mean, &rec->sigma, &hw_trend) : 0;
  } else {
    model_ok = mx ? fmo_model(g, mx, mn, &rec->mean, &rec->sigma) : 0;
  }
  if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
  if (n_cur == 0) status |= FM_ST_NO_CURRENT;

  /* a4..a6: pairwise same-distribution tests, canary only (design.md:35). */
  int differ = 0;
  if (canary && g->pairwise_mode != FM_PW_NONE && n_cur > 0 && n_base > 0) {
    fmo_test tests[3]; memset(tests, 0, sizeof(tests));
    const int want_mw = fmo_mode_has(g->pairwise_mode, FM_PW_MANN_WHITE) && n_cur >= g->min_mw && n_base >= g->min_mw;
    const int want_kw = fmo_mode_has(g->pairwise_mode, FM_PW_KRUSKAL) && n_cur >= g->min_kw && n_base >= g->min_kw;
    const int want_wx = fmo_mode_has(g->pairwise_mode, FM_PW_WILCOXON);
    if (want_mw || want_kw) {
      const uint32_t N = n_cur + n_base;
      fmo_ws_reserve(ws, N);
      for (uint32_t i = 0; i < n_cur; ++i) ws->v[i] = (double)cur[i];
      for (uint32_t i = 0; i < n_base; ++i) ws->v[n_cur + i] = (double)base[i];
      int64_t T;
      fmo_midranks_ws(ws, ws->v, N, ws->r2, &T);
      int64_t r2c = 0;
      for (uint32_t i = 0; i < n_cur; ++i) r2c += ws->r2[i];
      ax.r2_cur = r2c; ax.t_mw = T;
      if (want_mw) {
        tests[0] = fmo_mann_whitney(r2c, T, n_cur, n_base, g->cc_mw);
        rec->u1 = tests[0].stat; rec->p_u = tests[0].p; status |= FM_ST_MW_RUN;
        if (tests[0].degenerate) status |= FM_ST_DEGENERATE;
      }
      if (want_kw) {
        int64_t r2g[2]; uint32_t ng[2] = {n_cur, n_base};
        r2g[0] = r2c; r2g[1] = (int64_t)N * ((int64_t)N + 1) - r2c;
        tests[2] = fmo_kruskal(r2g, ng, 2, T);
        rec->h = tests[2].stat; rec->p_h = tests[2].p; status |= FM_ST_KW_RUN;
        if (tests[2].degenerate) status |= FM_ST_DEGENERATE;
      }
    }
    if (want_wx) {
      /* A.4 Wilcoxon: pair the first (HEAD) or last (TAIL) min(n) samples;
       * differences in fp64, zeros dropped (zero_method = wilcox). */
      const uint32_t n = n_cur < n_base ? n_cur : n_base;
      const uint32_t oc = g->pair_rule == FM_PAIR_TAIL ? n_cur - n : 0;
      const uint32_t ob = g->pair_rule == FM_PAIR_TAIL ? n_base - n : 0;
      fmo_ws_reserve(ws, n ? n : 1);
      uint32_t np = 0;
      /* ws->v holds |d| of the kept pairs; the sign goes to the tmp index slot */
      double* sgn = (double*)malloc((size_t)(n ? n : 1) * sizeof(double));
      for (uint32_t i = 0; i < n; ++i) {
        const double d = (double)cur[oc + i] - (double)base[ob + i];
        if (d != 0.0) { ws->v[np] = fabs(d); sgn[np] = d; ++np; }
      }
      ax.n_wx = np;
      if (np >= g->min_wx && np > 0) {
        int64_t T;
        fmo_midranks_ws(ws, ws->v, np, ws->r2, &T);
        int64_t wp = 0, wm = 0;
        for (uint32_t i = 0; i < np; ++i) { if (sgn[i] > 0.0) wp += ws->r2[i]; else wm += ws->r2[i]; }
        ax.w2_plus = wp; ax.w2_minus = wm; ax.t_wx = T;
        tests[1] = fmo_wilcoxon(wp, wm, T, np, g->cc_wx);
        rec->w = tests[1].stat; rec->p_w = tests[1].p; status |= FM_ST_WX_RUN;
        if (tests[1].degenerate) status |= FM_ST_DEGENERATE;
      }
      free(sgn);
    }
    int skipped = 0;
    differ = fmo_decide(g->pairwise_mode, g->alpha, tests, 3, &skipped);
    if (skipped) status |= FM_ST_PAIRWISE_SKIPPED;
    if (differ) status |= FM_ST_DIFFER;
  }

  /* a7, a8 */
  if (model_ok && n_cur > 0) {
    if (holt_winters) fmo_threshold_hw(g, tp, differ, cur, n_cur, season, mn, g->season_len, hw_trend, rec);
    else fmo_threshold(g, tp, differ, cur, n_cur, rec);
  }
  free(season);
  if ((status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT)) != 0) rec->verdict = FM_UNKNOWN;
  rec->status = (uint8_t)status;
  if (aux) *aux = ax;
}

void fmo_score_series(const fm_global_params* g, const fm_type_params* tp, uint32_t strategy,
                      const float* cur, uint32_t n_cur, const float* base, uint32_t n_base,
                      const float* hist, uint32_t n_hist, fm_series_rec* rec, fm_rank_aux* aux) {
  fmo_ws ws; memset(&ws, 0, sizeof(ws));
  fmo_score_series_ws(&ws, g, tp, strategy, cur, n_cur, base, n_base, hist, n_hist, rec, aux);
  fmo_ws_free(&ws);
}

static void fmo_compact(const fm_series_rec* r, fm_compact_rec* c) {
  c->anom_mask = r->anom_mask; c->n_anom = r->n_anom; c->verdict = r->verdict;
  c->status = r->status; c->first_idx = r->first_idx;
}

/* ---------------------------------------------------------------- A.9 --- */
void fmo_rollup(const uint8_t* verdict, const uint32_t* parent, uint64_t n, fm_rollup_rec* out,
                uint32_t n_parents) {
  memset(out, 0, (size_t)n_parents * sizeof(fm_rollup_rec));
  for (uint64_t i = 0; i < n; ++i) {
    fm_rollup_rec* r = &out[parent[i]];
    r->n_children++;
    if (verdict[i] == FM_UNHEALTHY) r->n_unhealthy++;
    if (verdict[i] == FM_UNKNOWN) r->n_unknown++;
  }
  for (uint32_t p = 0; p < n_parents; ++p) {
    fm_rollup_rec* r = &out[p];
    r->severity = r->n_unhealthy > 0 ? FM_UNHEALTHY
                : (r->n_children > 0 && r->n_unknown == r->n_children) ? FM_UNKNOWN : FM_HEALTHY;
  }
}

/* ------------------------------------------------------------ batch ----- */
typedef struct {
  const fm_global_params* g; const fm_type_params* table; uint32_t n_types;
  const fm_batch* b; void* out; fm_rank_aux* aux; uint64_t lo, hi;
} fmo_job;

static uint32_t fmo_len(const uint32_t* len, uint64_t s, uint32_t n) { return len ? (len[s] < n ? len[s] : n) : n; }

static void* fmo_batch_worker(void* arg) {
  fmo_job* j = (fmo_job*)arg; const fm_batch* b = j->b;
  fmo_ws ws; memset(&ws, 0, sizeof(ws));
  for (uint64_t s = j->lo; s < j->hi; ++s) {
    const uint32_t ty = b->metric_type ? b->metric_type[s] : 0;
    const fm_type_params* tp = &j->table[ty < j->n_types ? ty : 0];
    fm_series_rec rec;
    fmo_score_series_ws(&ws, j->g, tp, b->strategy,
                        b->cur ? b->cur + s * b->cur_pitch : NULL, b->cur ? fmo_len(b->cur_len, s, b->n_cur) : 0,
                        b->base ? b->base + s * b->base_pitch : NULL, b->base ? fmo_len(b->base_len, s, b->n_base) : 0,
                        b->hist ? b->hist + s * b->hist_pitch : NULL, b->hist ? fmo_len(b->hist_len, s, b->n_hist) : 0,
                        &rec, j->aux ? &j->aux[s] : NULL);
    if (b->record_kind == FM_REC_COMPACT) fmo_compact(&rec, &((fm_compact_rec*)j->out)[s]);
    else ((fm_series_rec*)j->out)[s] = rec;
  }
  fmo_ws_free(&ws);
  return NULL;
}

int32_t fmo_score_batch(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                        const fm_batch* b, void* out_records, fm_rank_aux* out_aux,
                        fm_rollup_rec* out_api, fm_rollup_rec* out_svc, uint32_t n_threads) {
  if (!g || !table || !b || !out_records || n_types == 0) return FM_E_INVALID;
  if (n_threads == 0) n_threads = 1;
  if (n_threads > 1024) n_threads = 1024;
  if ((uint64_t)n_threads > b->n_series) n_threads = b->n_series ? (uint32_t)b->n_series : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * n_threads);
  fmo_job* jobs = (fmo_job*)malloc(sizeof(fmo_job) * n_threads);
  for (uint32_t t = 0; t < n_threads; ++t) {
    jobs[t] = (fmo_job){g, table, n_types, b, out_records, out_aux,
                        b->n_series * t / n_threads, b->n_series * (t + 1) / n_threads};
    if (n_threads == 1) fmo_batch_worker(&jobs[t]);
    else pthread_create(&th[t], NULL, fmo_batch_worker, &jobs[t]);
  }
  if (n_threads > 1) for (uint32_t t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
  free(th); free(jobs);

  if (b->api_id && out_api && b->n_apis) {
    uint8_t* v = (uint8_t*)malloc(b->n_series ? b->n_series : 1);
    for (uint64_t s = 0; s < b->n_series; ++s)
      v[s] = b->record_kind == FM_REC_COMPACT ? ((fm_compact_rec*)out_records)[s].verdict
                                               : ((fm_series_rec*)out_records)[s].verdict;
    fmo_rollup(v, b->api_id, b->n_series, out_api, b->n_apis);
    free(v);
    if (b->api_service && out_svc && b->n_services) {
      uint8_t* av = (uint8_t*)malloc(b->n_apis);
      for (uint32_t a = 0; a < b->n_apis; ++a) av[a] = out_api[a].severity;
      fmo_rollup(av, b->api_service, b->n_apis, out_svc, b->n_services);
      free(av);
    }
  }
  return FM_OK;
}

/* ----------------------------------------------------------- groups ----- */
typedef struct {
  const fm_global_params* g; const fm_type_params* table; uint32_t n_types;
  const fm_group_batch* b; fm_series_rec* out; uint64_t lo, hi;
} fmo_gjob;

static void* fmo_group_worker(void* arg) {
  fmo_gjob* j = (fmo_gjob*)arg; const fm_group_batch* b = j->b; const fm_global_params* g = j->g;
  fmo_ws ws; memset(&ws, 0, sizeof(ws));
  for (uint64_t v = j->lo; v < j->hi; ++v) {
    const uint32_t g0 = b->group_offsets[v], g1 = b->group_offsets[v + 1];
    const uint32_t K = g1 - g0;
    const uint32_t ty = b->metric_type ? b->metric_type[v] : 0;
    const fm_type_params* tp = &j->table[ty < j->n_types ? ty : 0];
    double mean = 0.0, sigma = 0.0;
    const float* hx = b->hist ? b->hist + v * b->hist_pitch : NULL;
    const uint32_t hn = b->hist ? fmo_len(b->hist_len, v, b->n_hist) : 0;
    const int model_ok = hx ? fmo_model(g, hx, hn, &mean, &sigma) : 0;

    /* Kruskal-Wallis over the K groups (README.md:34,38). */
    fmo_test kw; memset(&kw, 0, sizeof(kw)); kw.p = 1.0;
    uint32_t status_kw = 0; int differ = 0;
    if (g->pairwise_mode == FM_PW_FRIEDMAN) {
      /* Friedman chi-square (docs/guides/design.md:92), DESIGN.md 3 (A.4b): the K groups as related
       * samples, block i = the i-th point of every group.  Needs K >= 3 groups of one common
       * length n >= MIN_KRUSKAL_DATA_POINTS. */
      uint32_t n = K ? fmo_len(b->cur_len, g0, b->n_cur) : 0;
      int eligible = K >= 3 && K <= FM_MAX_GROUPS && n >= g->min_kw && n > 0;
      for (uint32_t k = 1; eligible && k < K; ++k)
        if (fmo_len(b->cur_len, g0 + k, b->n_cur) != n) eligible = 0;
      if (eligible) {
        int64_t ssq = 0, ties = 0;            /* sum_j (2 R_j)^2 and sum over blocks of sum t(t^2-1) */
        int64_t r2[FM_MAX_GROUPS]; memset(r2, 0, sizeof(r2));
        for (uint32_t i = 0; i < n; ++i) {
          for (uint32_t a = 0; a < K; ++a) {
            const float xa = b->cur[(uint64_t)(g0 + a) * b->cur_pitch + i];
            int64_t lt = 0, eq = 0;
            for (uint32_t c = 0; c < K; ++c) {
              const float xc = b->cur[(uint64_t)(g0 + c) * b->cur_pitch + i];
              lt += xc < xa; eq += xc == xa;
            }
            r2[a] += 2 * lt + eq + 1;         /* twice the midrank inside the block */
            ties += eq * eq - 1;              /* every member of a tie group of t adds t^2 - 1 */
          }
        }
        for (uint32_t a = 0; a < K; ++a) ssq += r2[a] * r2[a];
        const double dk = (double)K, dn = (double)n;
        const double c = 1.0 - (double)ties / (dk * (dk * dk - 1.0) * dn);
        status_kw |= FM_ST_KW_RUN;
        kw.run = 1;
        if (!(c > 0.0)) { kw.degenerate = 1; status_kw |= FM_ST_DEGENERATE; }
        else {
          const double q0 = 12.0 / (dk * dn * (dk + 1.0)) * (0.25 * (double)ssq) - 3.0 * dn * (dk + 1.0);
          kw.stat = q0 / c;
          kw.p = fmo_chi2_sf(kw.stat, K - 1);
          differ = kw.p < g->alpha;
        }
      }
      if (!(status_kw & FM_ST_KW_RUN) || kw.degenerate) status_kw |= FM_ST_PAIRWISE_SKIPPED;
      if (differ) status_kw |= FM_ST_DIFFER;
    }
    if (fmo_mode_has(g->pairwise_mode, FM_PW_KRUSKAL) && g->pairwise_mode != FM_PW_NONE) {
      uint32_t ng[FM_MAX_GROUPS]; int64_t r2g[FM_MAX_GROUPS]; uint32_t N = 0; int eligible = K >= 2 && K <= FM_MAX_GROUPS;
      for (uint32_t k = 0; eligible && k < K; ++k) {
        ng[k] = fmo_len(b->cur_len, g0 + k, b->n_cur);
        if (ng[k] < g->min_kw || ng[k] == 0) eligible = 0;
        N += ng[k];
      }
      if (eligible) {
        fmo_ws_reserve(&ws, N);
        uint32_t o = 0;
        for (uint32_t k = 0; k < K; ++k) {
          const float* row = b->cur + (uint64_t)(g0 + k) * b->cur_pitch;
          for (uint32_t i = 0; i < ng[k]; ++i) ws.v[o++] = (double)row[i];
        }
        int64_t T;
        fmo_midranks_ws(&ws, ws.v, N, ws.r2, &T);
        o = 0;
        for (uint32_t k = 0; k < K; ++k) { int64_t s = 0; for (uint32_t i = 0; i < ng[k]; ++i) s += ws.r2[o++]; r2g[k] = s; }
        kw = fmo_kruskal(r2g, ng, K, T);
        status_kw |= FM_ST_KW_RUN;
        if (kw.degenerate) status_kw |= FM_ST_DEGENERATE;
        else differ = kw.p < g->alpha;
      }
      if (!(status_kw & FM_ST_KW_RUN) || kw.degenerate) status_kw |= FM_ST_PAIRWISE_SKIPPED;
      if (differ) status_kw |= FM_ST_DIFFER;
    }
    for (uint32_t k = 0; k < K; ++k) {
      fm_series_rec* rec = &j->out[g0 + k];
      memset(rec, 0, sizeof(*rec));
      rec->p_u = rec->p_w = 1.0; rec->h = kw.stat; rec->p_h = kw.p; rec->first_idx = 0xFFFFu;
      rec->mean = mean; rec->sigma = sigma;
      uint32_t status = status_kw;
      const uint32_t nc = fmo_len(b->cur_len, g0 + k, b->n_cur);
      if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
      if (nc == 0) status |= FM_ST_NO_CURRENT;
      if (model_ok && nc > 0) fmo_threshold(g, tp, differ, b->cur + (uint64_t)(g0 + k) * b->cur_pitch, nc, rec);
      if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT)) rec->verdict = FM_UNKNOWN;
      rec->status = (uint8_t)status;
    }
  }
  fmo_ws_free(&ws);
  return NULL;
}

int32_t fmo_score_groups(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                         const fm_group_batch* b, fm_series_rec* out_groups,
                         fm_rollup_rec* out_svc, uint32_t n_threads) {
  if (!g || !table || !b || !out_groups || !b->group_offsets || n_types == 0) return FM_E_INVALID;
  if (n_threads == 0) n_threads = 1;
  if ((uint64_t)n_threads > b->n_services) n_threads = b->n_services ? (uint32_t)b->n_services : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * n_threads);
  fmo_gjob* jobs = (fmo_gjob*)malloc(sizeof(fmo_gjob) * n_threads);
  for (uint32_t t = 0; t < n_threads; ++t) {
    jobs[t] = (fmo_gjob){g, table, n_types, b, out_groups,
                         b->n_services * t / n_threads, b->n_services * (t + 1) / n_threads};
    if (n_threads == 1) fmo_group_worker(&jobs[t]);
    else pthread_create(&th[t], NULL, fmo_group_worker, &jobs[t]);
  }
  if (n_threads > 1) for (uint32_t t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
  free(th); free(jobs);
  if (out_svc) {
    uint8_t* v = (uint8_t*)malloc(b->n_groups ? b->n_groups : 1);
    uint32_t* parent = (uint32_t*)malloc((b->n_groups ? b->n_groups : 1) * sizeof(uint32_t));
    for (uint64_t s = 0; s < b->n_services; ++s)
      for (uint32_t k = b->group_offsets[s]; k < b->group_offsets[s + 1]; ++k) { parent[k] = (uint32_t)s; v[k] = out_groups[k].verdict; }
    fmo_rollup(v, parent, b->n_groups, out_svc, (uint32_t)b->n_services);
    free(v); free(parent);
  }
  return FM_OK;
}

/* ------------------------------------------------- timed CPU baseline --- */
typedef struct {
  const fm_global_params* g; const fm_type_params* table; uint32_t n_types;
  const fm_synth_desc* d; uint32_t strategy; int hist_is_base; uint64_t lo, hi;
  double seconds; uint64_t checksum;
} fmo_bjob;

static double fmo_now(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec; }

static void* fmo_bench_worker(void* arg) {
  fmo_bjob* j = (fmo_bjob*)arg; const fm_synth_desc* d = j->d;
  fmo_ws ws; memset(&ws, 0, sizeof(ws));
  float* cur = (float*)malloc(sizeof(float) * (d->n_cur ? d->n_cur : 1));
  float* base = (float*)malloc(sizeof(float) * (d->n_base ? d->n_base : 1));
  float* hist = (float*)malloc(sizeof(float) * (d->n_hist ? d->n_hist : 1));
  double secs = 0.0; uint64_t sum = 0;
  for (uint64_t s = j->lo; s < j->hi; ++s) {
    const uint64_t id = d->first_series + s;
    const fm_synth_series sp = fm_synth_series_params(d->seed, d->config, id, d->n_cur);
    for (uint32_t t = 0; t < d->n_cur; ++t) cur[t] = fm_synth_sample(d->seed, &sp, id, FM_ROW_CUR, t);
    for (uint32_t t = 0; t < d->n_base; ++t) base[t] = fm_synth_sample(d->seed, &sp, id, FM_ROW_BASE, t);
    if (!j->hist_is_base) for (uint32_t t = 0; t < d->n_hist; ++t) hist[t] = fm_synth_sample(d->seed, &sp, id, FM_ROW_HIST, t);
    const fm_type_params* tp = &j->table[sp.type < j->n_types ? sp.type : 0];
    fm_series_rec rec;
    const double t0 = fmo_now();
    fmo_score_series_ws(&ws, j->g, tp, j->strategy, cur, d->n_cur,
                        d->n_base ? base : NULL, d->n_base,
                        (j->hist_is_base || d->n_hist == 0) ? NULL : hist, d->n_hist, &rec, NULL);
    secs += fmo_now() - t0;
    sum += rec.verdict + 3u * rec.n_anom + 7u * rec.status;
  }
  free(cur); free(base); free(hist); fmo_ws_free(&ws);
  j->seconds = secs; j->checksum = sum;
  return NULL;
}

int32_t fmo_bench_synth(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                        const fm_synth_desc* desc, uint32_t strategy, int hist_is_base,
                        uint32_t n_threads, double* seconds, uint64_t* checksum) {
  if (!g || !table || !desc || !seconds) return FM_E_INVALID;
  if (n_threads == 0) n_threads = 1;
  if ((uint64_t)n_threads > desc->n_series) n_threads = desc->n_series ? (uint32_t)desc->n_series : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * n_threads);
  fmo_bjob* jobs = (fmo_bjob*)malloc(sizeof(fmo_bjob) * n_threads);
  for (uint32_t t = 0; t < n_threads; ++t) {
    jobs[t] = (fmo_bjob){g, table, n_types, desc, strategy, hist_is_base,
                         desc->n_series * t / n_threads, desc->n_series * (t + 1) / n_threads, 0.0, 0};
    pthread_create(&th[t], NULL, fmo_bench_worker, &jobs[t]);
  }
  double mx = 0.0; uint64_t sum = 0;
  for (uint32_t t = 0; t < n_threads; ++t) {
    pthread_join(th[t], NULL);
    if (jobs[t].seconds > mx) mx = jobs[t].seconds;
    sum += jobs[t].checksum;
  }
  free(th); free(jobs);
  *seconds = mx; if (checksum) *checksum = sum;
  return FM_OK;
}

/* A.2d: two metrics judged jointly, "Bivariate Normal Distribution" (docs/guides/design.md:67-76).
 * Two-pass means and covariances (ddof of g), squared Mahalanobis distance of every current pair
 * against k^2. */
int32_t fmo_score_pairs(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                        const fm_pair_batch* b, fm_pair_rec* out) {
  uint32_t need = g->min_hist_points > g->ddof + 1 ? g->min_hist_points : g->ddof + 1;
  if (need < 3) need = 3;
  for (uint64_t s = 0; s < b->n_pairs; ++s) {
    fm_pair_rec r; memset(&r, 0, sizeof(r));
    r.first_idx = 0xFFFFu;
    const uint32_t nh = fmo_len(b->hist_len, s, b->n_hist), nc = fmo_len(b->cur_len, s, b->n_cur);
    const float* hx = b->hist_x + s * b->hist_pitch; const float* hy = b->hist_y + s * b->hist_pitch;
    uint32_t status = 0;
    if (nc == 0) status |= FM_ST_NO_CURRENT;
    int ok = nh >= need;
    double det = 0.0;
    if (ok) {
      double mx = 0.0, my = 0.0;
      for (uint32_t t = 0; t < nh; ++t) { mx += (double)hx[t]; my += (double)hy[t]; }
      mx /= (double)nh; my /= (double)nh;
      double vx = 0.0, vy = 0.0, cv = 0.0;
      for (uint32_t t = 0; t < nh; ++t) {
        const double dx = (double)hx[t] - mx, dy = (double)hy[t] - my;
        vx += dx * dx; vy += dy * dy; cv += dx * dy;
      }
      const double dd = (double)(nh - g->ddof);
      r.mean_x = mx; r.mean_y = my; r.var_x = vx / dd; r.var_y = vy / dd; r.cov_xy = cv / dd;
      det = r.var_x * r.var_y - r.cov_xy * r.cov_xy;
      if (!(det > 1e-9 * r.var_x * r.var_y)) { status |= FM_ST_DEGENERATE; ok = 0; }   /* 1 - rho^2 <= 1e-9 */
    } else {
      status |= FM_ST_HIST_INSUFFICIENT;
    }
    if (ok && nc > 0) {
      const uint32_t ty = b->metric_type ? b->metric_type[s] : 0;
      const double k = table[ty < n_types ? ty : 0].k, k2 = k * k;
      const float* cx = b->cur_x + s * b->cur_pitch; const float* cy = b->cur_y + s * b->cur_pitch;
      uint32_t n_anom = 0; uint64_t mask = 0; double dmax = 0.0;
      for (uint32_t i = 0; i < nc; ++i) {
        const double dx = (double)cx[i] - r.mean_x, dy = (double)cy[i] - r.mean_y;
        const double d2 = (r.var_y * dx * dx - 2.0 * r.cov_xy * dx * dy + r.var_x * dy * dy) / det;
        if (d2 > dmax) dmax = d2;
        if (d2 > k2) {
          if (n_anom == 0) r.first_idx = (uint16_t)(i < 0xFFFFu ? i : 0xFFFEu);
          ++n_anom;
          if (i < 64) mask |= 1ull << i;
        }
      }
      r.d2_max = dmax; r.n_anom = n_anom; r.anom_mask = mask;
      r.verdict = (n_anom >= g->min_anomaly_points && n_anom > 0) ? FM_UNHEALTHY : FM_HEALTHY;
    }
    if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT | FM_ST_DEGENERATE)) r.verdict = FM_UNKNOWN;
    r.status = (uint8_t)status;
    out[s] = r;
  }
  return 0;
}
