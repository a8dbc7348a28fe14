// fm_canary_common.cuh — pieces shared by the canary kernels (fm_canary_sort.cu, fm_canary_elect.cu):
// row staging through the TMA unit, the 64-key warp sorts, the Wilcoxon signed-rank statistic on
// <= 64 pairs, and the common epilogue (one erfc for three p-values, pairwise decision, thresholds,
// anomaly ballot, record write-out, fused roll-up).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_device.cuh"

namespace fmc {

constexpr uint32_t kCur = 64;          // floats reserved for the current row at the start of a stage

__device__ __forceinline__ uint32_t row_len(const uint32_t* len, uint64_t s, uint32_t n) {
  return len ? min(__ldg(len + s), n) : n;
}

// lane 0: arm the barrier and start the bulk copies of series s into `stage`
// (current row at [0, 64), baseline row at [64, ...)).
__device__ __forceinline__ void issue_rows(const FmKArgs& a, float* stage, uint64_t* bar, uint64_t s,
                                           uint32_t nc, uint32_t nb) {
  const uint32_t cb = ((nc + 3u) & ~3u) * 4u, bb = ((nb + 3u) & ~3u) * 4u;
  fm_mbar_expect_tx(bar, cb + bb);
  if (cb) fm_bulk_g2s(stage, a.cur + s * a.cur_pitch, cb, bar);
  if (bb) fm_bulk_g2s(stage + kCur, fm_base_row(a, s), bb, bar);
}

// ---- 64 keys across the warp (a = index lane, b = index lane + 32), ascending -------------
__device__ __forceinline__ void sort64(float& a, float& b, int lane) {
#pragma unroll
  for (int k = 2; k <= 64; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j == 32) {
        const float lo = fminf(a, b), hi = fmaxf(a, b);
        a = lo; b = hi;
      } else {
        const bool up_a = (lane & k) == 0, up_b = ((lane + 32) & k) == 0;
        const bool lower = (lane & j) == 0;
        const float oa = __shfl_xor_sync(FM_FULL, a, j), ob = __shfl_xor_sync(FM_FULL, b, j);
        a = (lower == up_a) ? fminf(a, oa) : fmaxf(a, oa);
        b = (lower == up_b) ? fminf(b, ob) : fmaxf(b, ob);
      }
    }
  }
}
// same network on 64-bit keys (|d| << 1 | sign: equal keys are identical, so the side that keeps
// the maximum may replace on equality and one comparison serves both roles)
__device__ __forceinline__ void cex64(unsigned long long& v, unsigned long long o, bool take_min) {
  const bool repl = (o < v) == take_min;
  v = repl ? o : v;
}
__device__ __forceinline__ void sort64_abs(unsigned long long& a, unsigned long long& b, int lane) {
#pragma unroll
  for (int k = 2; k <= 64; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j == 32) {
        const unsigned long long a0 = a, b0 = b;
        cex64(a, b0, true);
        cex64(b, a0, false);
      } else {
        const bool up_a = (lane & k) == 0, up_b = ((lane + 32) & k) == 0;
        const bool lower = (lane & j) == 0;
        const unsigned long long oa = __shfl_xor_sync(FM_FULL, a, j), ob = __shfl_xor_sync(FM_FULL, b, j);
        cex64(a, oa, lower == up_a);
        cex64(b, ob, lower == up_b);
      }
    }
  }
}

// For a sorted 64-key sequence (positions lane and lane + 32; `sa` / `sb` = my key differs from
// the previous position's): start of my tie run (lo) and the first run start after me (e).
__device__ __forceinline__ void runs64(bool sa, bool sb, int lane, uint32_t& lo_a, uint32_t& lo_b,
                                       uint32_t& e_a, uint32_t& e_b) {
  lo_a = sa ? (uint32_t)lane : 0u; lo_b = sb ? (uint32_t)lane + 32u : 0u;
  uint32_t hi_a = sa ? (uint32_t)lane : 64u, hi_b = sb ? (uint32_t)lane + 32u : 64u;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t ua = __shfl_up_sync(FM_FULL, lo_a, o), ub = __shfl_up_sync(FM_FULL, lo_b, o);
    const uint32_t da = __shfl_down_sync(FM_FULL, hi_a, o), db = __shfl_down_sync(FM_FULL, hi_b, o);
    if (lane >= o) { lo_a = max(lo_a, ua); lo_b = max(lo_b, ub); }
    if (lane + o < 32) { hi_a = min(hi_a, da); hi_b = min(hi_b, db); }
  }
  lo_b = max(lo_b, __shfl_sync(FM_FULL, lo_a, 31));
  const uint32_t first_b = __shfl_sync(FM_FULL, hi_b, 0);
  hi_a = min(hi_a, first_b);
  e_a = __shfl_down_sync(FM_FULL, hi_a, 1);
  e_b = __shfl_down_sync(FM_FULL, hi_b, 1);
  if (lane == 31) { e_a = first_b; e_b = 64u; }
}
// runs of a sorted float sequence held as (sa, sb)
__device__ __forceinline__ void runs64f(float sa, float sb, int lane, uint32_t& lo_a, uint32_t& lo_b,
                                        uint32_t& e_a, uint32_t& e_b) {
  const float qa = __shfl_up_sync(FM_FULL, sa, 1);
  const float last_sa = __shfl_sync(FM_FULL, sa, 31);
  float qb = __shfl_up_sync(FM_FULL, sb, 1);
  if (lane == 0) qb = last_sa;
  runs64(lane == 0 || qa != sa, qb != sb, lane, lo_a, lo_b, e_a, e_b);
}

struct Prep { bool run; bool degenerate; double stat; double arg; };   // p = min(1, erfc(arg))

// ---- Wilcoxon signed rank on the first / last min(n1, n2) pairs (A.4) ---------------------
// Phase 1 reads the rows (time order) from the stage; phase 2 needs no shared memory, so the
// caller may hand the stage back to the TMA unit in between.
struct WxKeys { unsigned long long ka, kb; uint32_t nz, n_pairs; };
__device__ __forceinline__ unsigned long long wx_key(double d) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(d);
  return (b << 1) | (b >> 63);
}

__device__ __forceinline__ WxKeys wilcoxon_load(const fm_global_params& g, uint32_t stage_s, uint32_t nc,
                                                uint32_t nb, int lane, uint32_t base_off = kCur) {
  // key = |d| as an ordered 64-bit integer shifted left by one, bit 0 = (d < 0); zero differences
  // and padding get the largest key so that they sort behind the n' used ones
  WxKeys k = {~0ull, ~0ull, 0u, min(nc, nb)};
  const uint32_t base_s = stage_s + base_off * 4u;     // floats between the two staged rows
  const uint32_t oc = g.pair_rule == FM_PAIR_TAIL ? nc - k.n_pairs : 0u;
  const uint32_t ob = g.pair_rule == FM_PAIR_TAIL ? nb - k.n_pairs : 0u;
  if ((uint32_t)lane < k.n_pairs) {
    const double d = __dsub_rn((double)fm_lds_f32(stage_s + 4u * (oc + lane)),
                               (double)fm_lds_f32(base_s + 4u * (ob + lane)));
    if (d != 0.0) k.ka = wx_key(d); else ++k.nz;
  }
  if ((uint32_t)lane + 32u < k.n_pairs) {
    const double d = __dsub_rn((double)fm_lds_f32(stage_s + 4u * (oc + lane + 32u)),
                               (double)fm_lds_f32(base_s + 4u * (ob + lane + 32u)));
    if (d != 0.0) k.kb = wx_key(d); else ++k.nz;
  }
  return k;
}

__device__ __forceinline__ Prep wilcoxon_stat(const fm_global_params& g, WxKeys k, int lane, fm_rank_aux& ax) {
  Prep t = {false, false, 0.0, 0.0};
  const uint32_t nz = fm_warp_sum(k.nz);
  const uint32_t np = k.n_pairs - nz;
  ax.n_wx = np;
  if (!(np >= g.min_wx && np > 0)) return t;
  unsigned long long ka = k.ka, kb = k.kb;
  sort64_abs(ka, kb, lane);
  const unsigned long long ma = ka >> 1, mb = kb >> 1;
  const unsigned long long pa = __shfl_up_sync(FM_FULL, ma, 1);
  const unsigned long long last_a = __shfl_sync(FM_FULL, ma, 31);
  unsigned long long pb = __shfl_up_sync(FM_FULL, mb, 1);
  if (lane == 0) pb = last_a;
  uint32_t lo_a, lo_b, e_a, e_b;
  runs64(lane == 0 || pa != ma, pb != mb, lane, lo_a, lo_b, e_a, e_b);
  long long wp = 0, wm = 0, T = 0;
  if ((uint32_t)lane < np) {
    const long long r2 = (long long)(lo_a + e_a + 1u), tt = (long long)(e_a - lo_a);
    if (ka & 1ull) wm += r2; else wp += r2;
    T += tt * tt - 1;
  }
  if ((uint32_t)lane + 32u < np) {
    const long long r2 = (long long)(lo_b + e_b + 1u), tt = (long long)(e_b - lo_b);
    if (kb & 1ull) wm += r2; else wp += r2;
    T += tt * tt - 1;
  }
  wp = fm_warp_sum(wp); wm = fm_warp_sum(wm); T = fm_warp_sum(T);
  ax.w2_plus = wp; ax.w2_minus = wm; ax.t_wx = T;
  t.run = true;
  const double wmin = 0.5 * (double)(wp < wm ? wp : wm);
  const double dn = (double)np;
  const double mu = dn * (dn + 1.0) / 4.0;
  const double var = dn * (dn + 1.0) * (2.0 * dn + 1.0) / 24.0 - (double)T / 48.0;
  t.stat = wmin;
  if (!(var > 0.0)) { t.degenerate = true; return t; }
  double num = wmin - mu;
  if (g.cc_wx && num != 0.0) num += 0.5;
  t.arg = fabs(num / sqrt(var)) / 1.4142135623730951;
  return t;
}

// Mann-Whitney and 2-group Kruskal-Wallis from the exact integers (A.4)
__device__ __forceinline__ void rank_tests(const fm_global_params& g, bool want_mw, bool want_kw, uint32_t nc,
                                           uint32_t nb, long long r2c, long long T, Prep& t_mw, Prep& t_kw) {
  const double N = (double)nc + (double)nb;
  if (want_mw) {
    t_mw.run = true;
    const double u1 = 0.5 * (double)(r2c - (long long)nc * ((long long)nc + 1));
    const double u2 = (double)nc * (double)nb - u1;
    const double u = u1 > u2 ? u1 : u2;
    const double mu = 0.5 * (double)nc * (double)nb;
    const double var = ((double)nc * (double)nb / 12.0) * ((N + 1.0) - (double)T / (N * (N - 1.0)));
    t_mw.stat = u1;
    if (!(var > 0.0)) t_mw.degenerate = true;
    else t_mw.arg = ((u - mu - (g.cc_mw ? 0.5 : 0.0)) / sqrt(var)) / 1.4142135623730951;
  }
  if (want_kw) {
    t_kw.run = true;
    const double rx = 0.5 * (double)r2c;
    const double ry = 0.5 * (double)((long long)(nc + nb) * ((long long)(nc + nb) + 1) - r2c);
    const double sres = rx * rx / (double)nc + ry * ry / (double)nb;
    const double h0 = 12.0 / (N * (N + 1.0)) * sres - 3.0 * (N + 1.0);
    const double c = 1.0 - (double)T / (N * N * N - N);
    if (!(c > 0.0)) t_kw.degenerate = true;
    else {
      t_kw.stat = h0 / c;
      t_kw.arg = t_kw.stat > 0.0 ? sqrt(0.5 * t_kw.stat) : -1.0;   // chi2 sf, 1 dof; H <= 0 -> 1
    }
  }
}

// a5..a9: p-values, decision, bounds, anomalous points (current row held in registers in time
// order: craw_a = point lane, craw_b = point lane + 32), verdict, record, roll-up.
__device__ __forceinline__ void finish_series(const FmKArgs& a, const fm_global_params& g,
                                              const fm_type_params& tp, uint64_t s, int lane, uint32_t nc,
                                              bool pair, const Prep& t_mw, const Prep& t_wx, const Prep& t_kw,
                                              bool model_ok, float craw_a, float craw_b, fm_series_rec& rec,
                                              const fm_rank_aux& ax, uint32_t status) {
  // one erfc for the three tests: lanes 0, 1, 2 evaluate their own argument
  const double arg = lane == 0 ? t_mw.arg : (lane == 1 ? t_wx.arg : t_kw.arg);
  double pe = erfc(arg);
  pe = pe > 1.0 ? 1.0 : pe;
  const double p_mw = __shfl_sync(FM_FULL, pe, 0), p_wx = __shfl_sync(FM_FULL, pe, 1),
               p_kw = __shfl_sync(FM_FULL, pe, 2);
  const double pv_mw = t_mw.degenerate ? 1.0 : p_mw, pv_wx = t_wx.degenerate ? 1.0 : p_wx,
               pv_kw = t_kw.degenerate ? 1.0 : p_kw;
  if (t_mw.run) { rec.u1 = t_mw.stat; rec.p_u = pv_mw; status |= FM_ST_MW_RUN; }
  if (t_wx.run) { rec.w = t_wx.stat; rec.p_w = pv_wx; status |= FM_ST_WX_RUN; }
  if (t_kw.run) { rec.h = t_kw.degenerate ? 0.0 : t_kw.stat; rec.p_h = pv_kw; status |= FM_ST_KW_RUN; }
  if ((t_mw.run && t_mw.degenerate) || (t_wx.run && t_wx.degenerate) || (t_kw.run && t_kw.degenerate))
    status |= FM_ST_DEGENERATE;

  bool differ = false;
  if (pair) {   // A.6 over the evaluated, non-degenerate tests
    const bool e0 = t_mw.run && !t_mw.degenerate, e1 = t_wx.run && !t_wx.degenerate,
               e2 = t_kw.run && !t_kw.degenerate;
    const bool d0 = pv_mw < g.alpha, d1 = pv_wx < g.alpha, d2 = pv_kw < g.alpha;
    const bool any = (e0 && d0) || (e1 && d1) || (e2 && d2);
    const bool all = (!e0 || d0) && (!e1 || d1) && (!e2 || d2);
    if (!(e0 || e1 || e2)) status |= FM_ST_PAIRWISE_SKIPPED;
    else differ = g.pairwise_mode == FM_PW_ALL ? all : any;
    if (differ) status |= FM_ST_DIFFER;
  }
  if (model_ok && nc > 0) {   // A.7 / A.8
    const double k = differ ? tp.k_low : tp.k;
    const double up = __dadd_rn(rec.mean, __dmul_rn(k, rec.sigma));
    double lo = __dsub_rn(rec.mean, __dmul_rn(k, rec.sigma));
    if (lo < tp.min_lower_bound) lo = tp.min_lower_bound;
    rec.upper = up; rec.lower = lo;
    const double xa = (double)craw_a, xb = (double)craw_b;
    const bool fa = (uint32_t)lane < nc && (((tp.bound_mask & 1u) && xa > up) || ((tp.bound_mask & 2u) && xa < lo));
    const bool fb = (uint32_t)lane + 32u < nc && (((tp.bound_mask & 1u) && xb > up) || ((tp.bound_mask & 2u) && xb < lo));
    const uint32_t ba = __ballot_sync(FM_FULL, fa), bb = __ballot_sync(FM_FULL, fb);
    const uint64_t mask = (uint64_t)ba | ((uint64_t)bb << 32);
    const uint32_t cnt = (uint32_t)__popc(ba) + (uint32_t)__popc(bb);
    rec.n_anom = cnt;
    rec.first_idx = (uint16_t)(mask ? (uint32_t)__ffsll((long long)mask) - 1u : 0xFFFFu);
    rec.anom_mask = mask;
    rec.verdict = (cnt >= g.min_anomaly_points && cnt > 0) ? FM_UNHEALTHY : FM_HEALTHY;
  }
  if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT)) rec.verdict = FM_UNKNOWN;
  rec.status = (uint8_t)status;
  if (lane == 0) {
    fm_write_record(a.out, s, a.record_kind, rec);
    if (a.aux) a.aux[s] = ax;
    if (a.api && a.api_id) fm_rollup_add(a.api, a.api_id[s], a.n_apis, rec.verdict);
  }
}

}  // namespace fmc
