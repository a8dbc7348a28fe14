// fm_canary_count.cu — canary hot path WITHOUT sorting the baseline (BASELINE.json configs[0..2]).
//
// One warp per series, persistent over the batch, rows staged by the TMA unit (cp.async.bulk +
// mbarrier) exactly as in fm_canary_sort.cu.  What differs is the rank stage:
//
//   2*R1  Only the <= 63 current points are sorted (two keys per lane, warp network).  Every
//         baseline point b, held in a register, is located in that sorted row by a branch-free
//         binary search in shared memory (5 probes after a first comparison against the middle
//         key, which every lane holds in a register), giving lbc(b) = #{c < b} and, from a
//         run-length stored beside the key it lands on, ubc(b) = #{c <= b}:
//             2*R1 = n1(n1+1) + 2 n1 n2 - sum_b (lbc(b) + ubc(b)).
//   T     The tie term needs the multiset of all n1 + n2 values, not their order.  It is built by
//         ELECTION in a shared-memory table: pass A, every key stores its bit pattern at
//         slot hash(key) (plain store, somebody wins); pass B, every key reads the slot back: my
//         own pattern there -> my class owns the slot, a returned atomic increment of the slot's
//         counter is my arrival number p and T += 3p(p+1) (sum_{p<t} 3p(p+1) = t^3 - t);
//         another pattern -> my whole class lost (equal values hash equally) and the key goes to
//         an overflow list, which is replayed with another hash multiplier until it is empty.
//         Classes move together, so each is counted exactly once.  Register slots that hold no
//         sample carry one NaN pattern; the class they form is subtracted at the end.
//   W     Wilcoxon as before (64 fp64 |differences| sorted across the warp), tie runs read off two
//         ballots instead of warp scans.
//   a5..  ONE fp64 division, ONE square root and ONE erfc per series: lanes 0..4 each bring their
//         statistic into the form num * sqrt(Q / P) (Mann-Whitney, Wilcoxon, Kruskal, sigma, mean)
//         and evaluate it side by side.
//
// The integers (2*R1, T, W+, W-) are bit-identical to the sorting kernels and to the oracle; the
// fp64 results agree to rounding (1e-15), far inside the 1e-6 contract.
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_canary_common.cuh"

namespace {
using namespace fmc;

constexpr uint32_t kNoKey = 0x7FC00001u;      // NaN pattern of register slots that hold no sample
constexpr uint32_t kSearchBytes = 272;        // 65 floats (one pad word between the halves) + pad
constexpr uint32_t kFinalBytes = 528;         // 65 x {key, run length << 16 | 2 * position} + pad
constexpr uint32_t kMaxRounds = 16;

__device__ __forceinline__ uint2 lds_v2(uint32_t a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts_v2(uint32_t a, uint32_t x, uint32_t y) {
  asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void sts_zero16(uint32_t a) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(a), "r"(0u) : "memory");
}
__device__ __forceinline__ uint32_t mult_of(uint32_t round) {
  return (0x9E3779B1u + round * 0x7F4A7C15u) | 1u;          // a different odd multiplier per round
}
__device__ __forceinline__ uint32_t slot_of(uint32_t bits, uint32_t mult, uint32_t m) {
  return __umulhi(bits * mult, m) << 2;                       // byte offset into a table of m slots
}

// Runs of a sorted 64-key sequence held as positions lane (a) and lane + 32 (b), from the ballots
// of "my key differs from the previous position's": start of my run (lo) and first position
// after it (e).  Bit 0 of ba is always set.
__device__ __forceinline__ void runs_from_ballots(uint32_t ba, uint32_t bb, int lane, uint32_t& lo_a,
                                                  uint32_t& e_a, uint32_t& lo_b, uint32_t& e_b) {
  const uint32_t le = (2u << lane) - 1u;                      // bits <= lane
  const uint32_t up_a = ba & ~le, up_b = bb & ~le, dn_b = bb & le;
  lo_a = 31u - (uint32_t)__clz((int)(ba & le));
  e_a = up_a ? (uint32_t)__ffs((int)up_a) - 1u : (bb ? 31u + (uint32_t)__ffs((int)bb) : 64u);
  lo_b = dn_b ? 63u - (uint32_t)__clz((int)dn_b) : 31u - (uint32_t)__clz((int)ba);
  e_b = up_b ? 31u + (uint32_t)__ffs((int)up_b) : 64u;
}

// Wilcoxon signed rank, integers only: n' = non-zero pairs, 2W+, 2W-, tie term.
struct WxInts { uint32_t np, wp2, wm2, t; bool run; };
__device__ __forceinline__ WxInts wilcoxon_ints(const fm_global_params& g, WxKeys k, int lane) {
  WxInts r = {0u, 0u, 0u, 0u, false};
  r.np = k.n_pairs - __reduce_add_sync(FM_FULL, k.nz);
  if (!(r.np >= g.min_wx && r.np > 0)) return r;
  r.run = true;
  unsigned long long ka = k.ka, kb = k.kb;
  sort64_abs(ka, kb, lane);
  const unsigned long long ma = ka >> 1, mb = kb >> 1;
  const unsigned long long pa = __shfl_up_sync(FM_FULL, ma, 1);
  const unsigned long long last_a = __shfl_sync(FM_FULL, ma, 31);
  unsigned long long pb = __shfl_up_sync(FM_FULL, mb, 1);
  if (lane == 0) pb = last_a;
  const uint32_t ba = __ballot_sync(FM_FULL, lane == 0 || pa != ma), bb = __ballot_sync(FM_FULL, pb != mb);
  uint32_t lo_a, e_a, lo_b, e_b;
  runs_from_ballots(ba, bb, lane, lo_a, e_a, lo_b, e_b);
  uint32_t wp = 0, wm = 0, T = 0;
  if ((uint32_t)lane < r.np) {
    const uint32_t r2 = lo_a + e_a + 1u, tt = e_a - lo_a;
    if (ka & 1ull) wm += r2; else wp += r2;
    T += tt * tt - 1u;
  }
  if ((uint32_t)lane + 32u < r.np) {
    const uint32_t r2 = lo_b + e_b + 1u, tt = e_b - lo_b;
    if (kb & 1ull) wm += r2; else wp += r2;
    T += tt * tt - 1u;
  }
  r.wp2 = __reduce_add_sync(FM_FULL, wp);
  r.wm2 = __reduce_add_sync(FM_FULL, wm);
  r.t = __reduce_add_sync(FM_FULL, T);
  return r;
}

struct Tie { uint32_t sq, lin; };   // sum p^2 and MINUS sum p over my keys (T = 3 (sq - lin))

// One search step: if (v < b) addr += imm — a predicated add, no select + add pair.
template <uint32_t IMM>
__device__ __forceinline__ void step_lt(uint32_t& addr, float v, float b) {
  asm("{\n\t.reg .pred p;\n\tsetp.lt.f32 p, %1, %2;\n\t@p add.u32 %0, %0, %3;\n\t}"
      : "+r"(addr) : "f"(v), "f"(b), "n"(IMM));
}

// Pass B, first half, for one key: v = what the slot holds.  My own pattern -> my class owns the
// slot and the returned decrement of its counter is MINUS my arrival number.  No predicates and no
// branches (ptxas turns a predicated ATOMS into a branch): everybody runs the atomic, with the
// comparison mask (all ones / zero) as the addend, so a key whose class lost adds nothing to the
// winner's counter; everybody stores its pattern at the tail of its overflow list (entry k of
// lane l at list + 128 k + 4 l: conflict-free) and only the losers advance the tail.
// Returns -(arrival number) mod 2^32 for a member, 0 otherwise.
__device__ __forceinline__ uint32_t elect_count(uint32_t v, uint32_t bits, uint32_t caddr, uint32_t& lptr) {
  uint32_t inc, c;
  asm("set.eq.u32.u32 %0, %1, %2;" : "=r"(inc) : "r"(v), "r"(bits));
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(c) : "r"(caddr), "r"(inc) : "memory");
  fm_sts_u32(lptr, bits);
  lptr += 128u & ~inc;
  return c & inc;
}

// G baseline keys at a time: position of each in the sorted current row (binary search, the G
// probe chains interleaved by hand so that their shared-memory latencies overlap) and their
// election pass B.
template <int G>
__device__ __forceinline__ void rank_group(const float* x, uint32_t search_lo, uint32_t search_hi, uint32_t final_k,
                                           float c31, uint32_t tbl_s, uint32_t c_off, uint32_t mult, uint32_t m,
                                           uint32_t& acc_all, uint32_t& acc_eq, Tie& tie, uint32_t& lptr) {
  uint32_t addr[G], h[G], vv[G];
#pragma unroll
  for (int k = 0; k < G; ++k) {
    addr[k] = (c31 < x[k]) ? search_hi : search_lo;
    h[k] = tbl_s + slot_of(__float_as_uint(x[k]), mult, m);
  }
#pragma unroll
  for (int k = 0; k < G; ++k) vv[k] = fm_lds_u32(h[k]);
  {
    float v[G];
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k] + 60u);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<64u>(addr[k], v[k], x[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k] + 28u);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<32u>(addr[k], v[k], x[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k] + 12u);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<16u>(addr[k], v[k], x[k]);
  }
  uint32_t c[G];
#pragma unroll
  for (int k = 0; k < G; ++k) c[k] = elect_count(vv[k], __float_as_uint(x[k]), h[k] + c_off, lptr);
  {
    float v[G];
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k] + 4u);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<8u>(addr[k], v[k], x[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<4u>(addr[k], v[k], x[k]);
  }
  uint2 w[G];
#pragma unroll
  for (int k = 0; k < G; ++k) w[k] = lds_v2(final_k + 2u * addr[k]);
#pragma unroll
  for (int k = 0; k < G; ++k) { tie.sq += c[k] * c[k]; tie.lin += c[k]; }
#pragma unroll
  for (int k = 0; k < G; ++k) {
    acc_all += w[k].y;
    if (__uint_as_float(w[k].x) == x[k]) acc_eq += w[k].y;
  }
}

// election pass B for G keys that are not searched (current row, overflow rounds); the G chains
// (slot load, atomic) are in flight together
template <int G>
__device__ __forceinline__ void classify(const uint32_t* bits, uint32_t tbl_s, uint32_t c_off, uint32_t mult,
                                         uint32_t m, Tie& t, uint32_t& lptr) {
  uint32_t h[G], v[G], c[G];
#pragma unroll
  for (int k = 0; k < G; ++k) h[k] = tbl_s + slot_of(bits[k], mult, m);
#pragma unroll
  for (int k = 0; k < G; ++k) v[k] = fm_lds_u32(h[k]);
#pragma unroll
  for (int k = 0; k < G; ++k) c[k] = elect_count(v[k], bits[k], h[k] + c_off, lptr);
#pragma unroll
  for (int k = 0; k < G; ++k) { t.sq += c[k] * c[k]; t.lin += c[k]; }
}

// ---- singleton filter (FM_COUNT_PREFILTER) ------------------------------------------------------
// Most keys of a continuous series are the only one of their value, and such a key adds nothing to
// the tie term.  Sweep 1 therefore does not elect anything: every key adds 1 to a 16-bit counter
// picked by a hash of its bits (fire-and-forget red; the whole table region serves as 4 x slots
// counters).  In sweep 2 a key whose counter reads 1 shares its hash, hence its value, with nobody
// and is done; the others (a third of a continuous row: hash neighbours, and every real tie) go to
// the lists, and only they run the election rounds.  Exact: a counter never under-counts, and
// 16 bits hold every possible class (a row has at most 2111 keys).
#ifndef FM_COUNT_PREFILTER
#define FM_COUNT_PREFILTER 1
#endif
constexpr uint32_t kCntMult = 0x85EBCA6Bu;
__device__ __forceinline__ uint32_t cnt_addr(uint32_t tbl_s, uint32_t bits, uint32_t m16, uint32_t& sh) {
  const uint32_t i = __umulhi(bits * kCntMult, m16);
  sh = (i & 1u) << 4;
  return tbl_s + ((i >> 1) << 2);
}
__device__ __forceinline__ void cnt_add(uint32_t tbl_s, uint32_t bits, uint32_t m16) {
  uint32_t sh;
  const uint32_t a = cnt_addr(tbl_s, bits, m16, sh);
  fm_reds_add(a, 1u << sh);
}
// list entry for a key unless it is alone on its counter (w = the counter word)
__device__ __forceinline__ void list_unless_single(uint32_t w, uint32_t sh, uint32_t bits, uint32_t& lptr) {
  const uint32_t many = ((w >> sh) & 0xFFFFu) > 1u ? 0xFFFFFFFFu : 0u;
  fm_sts_u32(lptr, bits);
  lptr += 128u & many;
}
// rank_group without the election: search + singleton test
template <int G>
__device__ __forceinline__ void rank_group_p(const float* x, uint32_t search_lo, uint32_t search_hi, uint32_t final_k,
                                             float c31, float c15, float c47, uint32_t tbl_s, uint32_t m16, uint32_t& acc_all,
                                             uint32_t& acc_eq, uint32_t& lptr) {
  uint32_t addr[G], h[G], sh[G], vv[G];
#pragma unroll
  for (int k = 0; k < G; ++k) {
    addr[k] = (c31 < x[k]) ? search_hi : search_lo;
    h[k] = cnt_addr(tbl_s, __float_as_uint(x[k]), m16, sh[k]);
  }
  float top[G];                                    // second level from registers: the key is c15 or c47
#pragma unroll
  for (int k = 0; k < G; ++k) top[k] = (c31 < x[k]) ? c47 : c15;
#pragma unroll
  for (int k = 0; k < G; ++k) vv[k] = fm_lds_u32(h[k]);
  {
    float v[G];
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<64u>(addr[k], top[k], x[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k] + 28u);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<32u>(addr[k], v[k], x[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k] + 12u);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<16u>(addr[k], v[k], x[k]);
  }
#pragma unroll
  for (int k = 0; k < G; ++k) list_unless_single(vv[k], sh[k], __float_as_uint(x[k]), lptr);
  {
    float v[G];
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k] + 4u);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<8u>(addr[k], v[k], x[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) v[k] = fm_lds_f32(addr[k]);
#pragma unroll
    for (int k = 0; k < G; ++k) step_lt<4u>(addr[k], v[k], x[k]);
  }
  uint2 w[G];
#pragma unroll
  for (int k = 0; k < G; ++k) w[k] = lds_v2(final_k + 2u * addr[k]);
#pragma unroll
  for (int k = 0; k < G; ++k) {
    acc_all += w[k].y;
    if (__uint_as_float(w[k].x) == x[k]) acc_eq += w[k].y;
  }
}

// vector row c of the staged baseline: lane l takes float4 number l + 32 c (conflict-free).  Returns
// the index of its first element; k = the four keys (-0 folded onto +0, missing points = the NaN
// pattern), q = the raw samples.
__device__ __forceinline__ uint32_t keys4(uint32_t base_s, int c, int lane, uint32_t nb, float nokey, float* k,
                                          float4& q) {
  const uint32_t e0 = ((uint32_t)lane + 32u * (uint32_t)c) << 2;
  q = fm_lds_v4(base_s + (e0 << 2));
  if (128u * (uint32_t)c + 127u < nb) {                     // warp-uniform: the whole vector row is valid
    k[0] = q.x + 0.0f; k[1] = q.y + 0.0f; k[2] = q.z + 0.0f; k[3] = q.w + 0.0f;
  } else {
    k[0] = e0 < nb ? q.x + 0.0f : nokey;      k[1] = e0 + 1u < nb ? q.y + 0.0f : nokey;
    k[2] = e0 + 2u < nb ? q.z + 0.0f : nokey; k[3] = e0 + 3u < nb ? q.w + 0.0f : nokey;
  }
  return e0;
}

// The window / smoothing / separate-history forms of the model are a call, not inline code: the
// series loop stays small enough for the instruction cache.
// a5..a9: every statistic as num * sqrt(Q / P), one lane each (lane 0 Mann-Whitney, 1 Wilcoxon,
// 2 Kruskal-Wallis, 3 sigma, 4 mean): one fp64 division, one square root and one erfc per series;
// then decision, thresholds, anomaly ballot, record, roll-up.  Whole warp.
__device__ __forceinline__ void finish_stats(const FmKArgs& a, const fm_global_params& g, const fm_type_params& tp,
                                             uint64_t s, int lane, uint32_t nc, uint32_t nb, bool pair, bool want_mw,
                                             bool want_kw, const WxInts& wx, bool fuse_model, bool model_ok, double s1,
                                             double s2, double pivot, long long r2c, long long Tt, float craw_a,
                                             float craw_b, bool va, bool vb, fm_series_rec& rec, fm_rank_aux& ax,
                                             uint32_t status) {
  // ---- a5: every statistic as num * sqrt(Q / P), one lane each ---------------------------------
  const double dn1 = (double)nc, dn2 = (double)nb, N = dn1 + dn2, dT = (double)Tt;
  const bool run_mw = want_mw, run_kw = want_kw, run_wx = wx.run;
  double Q = 0.0, P = 1.0, num = 0.0;
  double u1 = 0.0, wmin = 0.0;
  if (run_mw) {
    u1 = 0.5 * (double)(r2c - (long long)nc * ((long long)nc + 1));
    const double u2 = dn1 * dn2 - u1, u = u1 > u2 ? u1 : u2;
    if (lane == 0) {
      num = u - 0.5 * dn1 * dn2 - (g.cc_mw ? 0.5 : 0.0);
      P = dn1 * dn2 * ((N + 1.0) * N * (N - 1.0) - dT);          // 12 N (N-1) var
      Q = 6.0 * N * (N - 1.0);
    }
  }
  if (run_wx) {
    ax.w2_plus = wx.wp2; ax.w2_minus = wx.wm2; ax.t_wx = wx.t;
    const double dn = (double)wx.np;
    wmin = 0.5 * (double)min(wx.wp2, wx.wm2);
    if (lane == 1) {
      num = wmin - dn * (dn + 1.0) * 0.25;
      if (g.cc_wx && num != 0.0) num += 0.5;
      num = fabs(num);
      P = 2.0 * dn * (dn + 1.0) * (2.0 * dn + 1.0) - (double)wx.t;   // 48 var
      Q = 24.0;
    }
  }
  if (run_kw && lane == 2) {
    const long long Ni = (long long)nc + nb, r2y = Ni * (Ni + 1) - r2c;
    double dA;
    if (Ni <= 4096) {                                                // every product below 2^63
      const long long A = 3ll * (r2c * r2c * (long long)nb + r2y * r2y * (long long)nc) -
                          3ll * (Ni + 1) * (Ni + 1) * Ni * (long long)nc * (long long)nb;
      dA = (double)A;
    } else {                                                         // long baselines: 128-bit integers
      const __int128 A = (__int128)3 * ((__int128)(r2c * r2c) * nb + (__int128)r2y * r2y * nc) -
                         (__int128)3 * (Ni + 1) * (Ni + 1) * Ni * ((__int128)nc * nb);
      const bool neg = A < 0;
      const unsigned __int128 M = neg ? (unsigned __int128)(-A) : (unsigned __int128)A;
      dA = (double)(unsigned long long)(M >> 64) * 18446744073709551616.0 + (double)(unsigned long long)M;
      if (neg) dA = -dA;
    }
    num = 1.0;
    P = dn1 * dn2 * (N * N * N - N - dT);
    Q = 0.5 * dA * (N - 1.0);                                        // Q / P = H / 2
  }
  if (fuse_model && model_ok) {
    if (lane == 3) {
      const double m2n = s2 * dn2 - s1 * s1;
      Q = (nb > g.ddof && m2n > 0.0) ? m2n : 0.0;
      P = nb > g.ddof ? dn2 * (double)(nb - g.ddof) : 1.0;
    }
    if (lane == 4) { Q = s1; P = dn2; }
  }
  const bool deg = !(P > 0.0);                                        // zero variance: p = 1
  const double q = deg ? 0.0 : Q / P;
  const double rt = sqrt(q);
  double arg = num * rt;
  if (lane == 2) arg = q > 0.0 ? rt : -1.0;                          // chi2 sf, 1 dof; H <= 0 -> 1
  double pe = erfc(arg);
  pe = (pe > 1.0 || deg) ? 1.0 : pe;
  const uint32_t degb = __ballot_sync(FM_FULL, deg);
  const double p_mw = __shfl_sync(FM_FULL, pe, 0), p_wx = __shfl_sync(FM_FULL, pe, 1),
               p_kw = __shfl_sync(FM_FULL, pe, 2);
  const double h_kw = 2.0 * __shfl_sync(FM_FULL, q, 2);
  if (fuse_model && model_ok) {
    rec.sigma = __shfl_sync(FM_FULL, rt, 3);
    rec.mean = pivot + __shfl_sync(FM_FULL, q, 4);
  }
  const bool d_mw = run_mw && (degb & 1u), d_wx = run_wx && (degb & 2u), d_kw = run_kw && (degb & 4u);
  if (run_mw) { rec.u1 = u1; rec.p_u = p_mw; status |= FM_ST_MW_RUN; }
  if (run_wx) { rec.w = wmin; rec.p_w = p_wx; status |= FM_ST_WX_RUN; }
  if (run_kw) { rec.h = d_kw ? 0.0 : h_kw; rec.p_h = p_kw; status |= FM_ST_KW_RUN; }
  if (d_mw || d_wx || d_kw) status |= FM_ST_DEGENERATE;

  bool differ = false;
  if (pair) {   // A.6 over the evaluated, non-degenerate tests
    const bool e0 = run_mw && !d_mw, e1 = run_wx && !d_wx, e2 = run_kw && !d_kw;
    const bool f0 = p_mw < g.alpha, f1 = p_wx < g.alpha, f2 = p_kw < g.alpha;
    const bool any = (e0 && f0) || (e1 && f1) || (e2 && f2);
    const bool all = (!e0 || f0) && (!e1 || f1) && (!e2 || f2);
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
    const bool fa = va && (((tp.bound_mask & 1u) && xa > up) || ((tp.bound_mask & 2u) && xa < lo));
    const bool fb = vb && (((tp.bound_mask & 1u) && xb > up) || ((tp.bound_mask & 2u) && xb < lo));
    const uint32_t ma = __ballot_sync(FM_FULL, fa), mb = __ballot_sync(FM_FULL, fb);
    const uint64_t mask = (uint64_t)ma | ((uint64_t)mb << 32);
    const uint32_t cnt = (uint32_t)__popc(ma) + (uint32_t)__popc(mb);
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

// ---- heavy ties: the rank integers from the multiset alone -----------------------------------
// An error-count series (a quarter of a job's metrics) has a handful of distinct values.  When the
// sorted current row shows few of them, the series first tries this path instead of sweep 2: every
// key reads its slot (pass A has run) and, if its class owns it, adds itself to the slot's counter
// with a fire-and-forget red (baseline points in the low half, current points in the high half);
// no search, no returned atomic, no overflow list.  If every key was placed and the table holds at
// most 64 classes, the (value, counts) list goes into the consumed stage, lane l takes entries l
// and l + 32, and one loop over the list gives, per class j with t_j = c_j + b_j points and L_j
// points below it:  2 R1 = sum c_j (2 L_j + t_j + 1),  T = sum t_j^3 - t_j  -- the same integers as
// the search path.  Anything else (a lost election, more classes) clears the counters again and
// returns false: the series then takes sweep 2 as if nothing had happened (the values are intact).
constexpr uint32_t kTieClasses = 64u;
struct TieOut { long long r2c, T; bool ok; };
__device__ __noinline__ TieOut ranks_by_classes(uint32_t stage_s, uint32_t base_s, uint32_t tbl_s, uint32_t c_off,
                                                uint32_t m_slots, uint32_t NV, uint32_t NS, uint32_t nb, uint32_t hka,
                                                uint32_t hkb, bool va, bool vb, int lane) {
  TieOut out = {0, 0, false};
  const uint32_t mult0 = mult_of(0);
  uint32_t lost = 0;
  auto place = [&](uint32_t bits, uint32_t add, bool real) {
    const uint32_t h = tbl_s + slot_of(bits, mult0, m_slots);
    const bool won = fm_lds_u32(h) == bits;
    fm_reds_add(h + c_off, (won && real) ? add : 0u);
    lost |= (real && !won) ? 1u : 0u;
  };
  place(hka, 0x10000u, va);
  place(hkb, 0x10000u, vb);
#pragma unroll 1
  for (uint32_t c = 0; c < NV; ++c) {
    const uint32_t e0 = ((uint32_t)lane + 32u * c) << 2;
    const float4 q = fm_lds_v4(base_s + (e0 << 2));
    place(__float_as_uint(q.x + 0.0f), 1u, e0 < nb);
    place(__float_as_uint(q.y + 0.0f), 1u, e0 + 1u < nb);
    place(__float_as_uint(q.z + 0.0f), 1u, e0 + 2u < nb);
    place(__float_as_uint(q.w + 0.0f), 1u, e0 + 3u < nb);
  }
#pragma unroll 1
  for (uint32_t j = 0; j < NS; ++j) {
    const uint32_t e = 128u * NV + (uint32_t)lane + 32u * j;
    const float q = fm_lds_f32(base_s + (min(e, 128u * NV + 32u * NS - 1u) << 2));
    place(__float_as_uint(q + 0.0f), 1u, e < nb);
  }
  __syncwarp();
  // occupied slots -> dense list in the stage (every lane scans its share of the counters)
  uint32_t occ = 0;
  bool fail = __any_sync(FM_FULL, lost != 0u);
  if (!fail) {
    for (uint32_t o = (uint32_t)lane * 16u; o < c_off; o += 512u) {
      const float4 w = fm_lds_v4(tbl_s + c_off + o);
      occ += (__float_as_uint(w.x) ? 1u : 0u) + (__float_as_uint(w.y) ? 1u : 0u) + (__float_as_uint(w.z) ? 1u : 0u) +
             (__float_as_uint(w.w) ? 1u : 0u);
    }
    uint32_t incl = occ;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t u = __shfl_up_sync(FM_FULL, incl, d);
      if (lane >= d) incl += u;
    }
    const uint32_t D = __shfl_sync(FM_FULL, incl, 31);
    fail = D > kTieClasses;
    if (!fail) {
      uint32_t at = stage_s + (incl - occ) * 8u;
      for (uint32_t o = (uint32_t)lane * 16u; o < c_off; o += 512u) {
        const float4 w = fm_lds_v4(tbl_s + c_off + o);
        const uint32_t ww[4] = {__float_as_uint(w.x), __float_as_uint(w.y), __float_as_uint(w.z), __float_as_uint(w.w)};
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (ww[i]) { sts_v2(at, fm_lds_u32(tbl_s + o + 4u * i), ww[i]); at += 8u; }
      }
      __syncwarp();
      float v0 = 0.f, v1 = 0.f;
      uint32_t w0 = 0, w1 = 0;
      if ((uint32_t)lane < D) { const uint2 e = lds_v2(stage_s + 8u * lane); v0 = __uint_as_float(e.x); w0 = e.y; }
      if ((uint32_t)lane + 32u < D) { const uint2 e = lds_v2(stage_s + 8u * (lane + 32u)); v1 = __uint_as_float(e.x); w1 = e.y; }
      uint32_t l0 = 0, l1 = 0;
#pragma unroll 2
      for (uint32_t i = 0; i < D; ++i) {               // broadcast reads
        const uint2 e = lds_v2(stage_s + 8u * i);
        const uint32_t t = (e.y & 0xFFFFu) + (e.y >> 16);
        const float vi = __uint_as_float(e.x);
        l0 += (vi < v0) ? t : 0u;
        l1 += (vi < v1) ? t : 0u;
      }
      const uint32_t t0 = (w0 & 0xFFFFu) + (w0 >> 16), t1 = (w1 & 0xFFFFu) + (w1 >> 16);
      const uint32_t r2 = (w0 >> 16) * (2u * l0 + t0 + 1u) + (w1 >> 16) * (2u * l1 + t1 + 1u);
      const unsigned long long tt = (unsigned long long)t0 * t0 * t0 - t0 + (unsigned long long)t1 * t1 * t1 - t1;
      out.r2c = (long long)__reduce_add_sync(FM_FULL, r2);
      out.T = fm_warp_sum((long long)tt);
      out.ok = true;
    }
  }
  if (fail) {                                       // sweep 2 expects empty counters
    __syncwarp();
    for (uint32_t o = (uint32_t)lane * 16u; o < c_off; o += 512u) sts_zero16(tbl_s + c_off + o);
  }
  __syncwarp();
  return out;
}

struct ModelOut { double mean, sigma; bool ok; };
__device__ __noinline__ ModelOut model_call(const fm_global_params* g, const float* x, uint32_t n, bool vec_ok, int lane) {
  ModelOut o;
  o.ok = fm_model(*g, x, n, vec_ok, lane, o.mean, o.sigma);
  return o;
}

#ifndef FM_COUNT_MAXW
#define FM_COUNT_MAXW 16
#endif
__global__ void __launch_bounds__(32 * FM_COUNT_MAXW, 1) k_canary_count(const FmKArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  // tier: NV vector rows (4 keys per lane each) + NS scalar rows of the staged baseline
  const uint32_t NV = a.cnt_nv, NS = a.cnt_ns;
  const uint32_t kCap = 128u * NV + 32u * NS;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int wpc = blockDim.x >> 5;
  unsigned char* wp = smem + (size_t)warp * a.warp_bytes;
  uint64_t* bar = reinterpret_cast<uint64_t*>(wp);
  float* stage = reinterpret_cast<float*>(wp + 16);
  const uint32_t stage_s = fm_smem_u32(stage), base_s = stage_s + kCur * 4u;
  const uint32_t search_s = stage_s + (kCur + a.base_cap) * 4u;
  const uint32_t final_s = search_s + kSearchBytes;
  const uint32_t tbl_s = final_s + kFinalBytes;
  const uint32_t m_slots = a.log_m, c_off = 4u * a.log_m;     // a.log_m carries the slot count (a multiple of 4)
  const fm_global_params& g = a.P->g;

  if (lane == 0) { fm_mbar_init(bar, 1); fm_fence_mbar_init(); }
  __syncwarp();

  const uint64_t stride = (uint64_t)gridDim.x * wpc;
  uint64_t s = (uint64_t)blockIdx.x * wpc + warp;
  uint32_t phase = 0;
  if (s < a.n_series && lane == 0)
    issue_rows(a, stage, bar, s, row_len(a.cur_len, s, a.n_cur), row_len(a.base_len, s, a.n_base));

  // Every warp of the CTA makes the same number of trips and meets the others at a barrier at the
  // top of each: the warps of an SM then walk the (long) loop body together and share its
  // instruction fetches instead of each missing the instruction cache on its own.
  const uint64_t trips = (a.n_series + stride - 1) / stride;
  for (uint64_t trip = 0; trip < trips; ++trip, s += stride) {
    if (a.n_stages == 2u && ((uint32_t)trip & a.sync_mask) == 0u) asm volatile("bar.sync 0;" ::: "memory");
    if (s >= a.n_series) continue;
    const uint32_t nc = row_len(a.cur_len, s, a.n_cur);
    const uint32_t nb = row_len(a.base_len, s, a.n_base);
    const uint64_t sn = s + stride;
    if (sn < a.n_series) {     // pull the next rows into L2 while this series is ranked
      const uint32_t nbn = row_len(a.base_len, sn, a.n_base);
      const char* pb = reinterpret_cast<const char*>(fm_base_row(a, sn));
      for (uint32_t o = (uint32_t)lane * 128u; o < nbn * 4u; o += 4096u) fm_prefetch_l2(pb + o);
      if (lane < 2) fm_prefetch_l2(reinterpret_cast<const char*>(a.cur + sn * a.cur_pitch) + lane * 128);
    }
    fm_mbar_wait(bar, phase);
    phase ^= 1u;

    const uint32_t ty = a.metric_type ? a.metric_type[s] : 0u;
    const fm_type_params tp = a.P->types[ty < a.P->n_types ? ty : 0u];
    fm_series_rec rec;
    fm_rec_init(rec);
    fm_rank_aux ax = {0, 0, 0, 0, 0, 0u, 0u};
    uint32_t status = 0;
    if (nc == 0) status |= FM_ST_NO_CURRENT;

    const bool pair = g.pairwise_mode != FM_PW_NONE && nc > 0 && nb > 0;
    const bool want_mw = pair && fm_mode_has(g.pairwise_mode, FM_PW_MANN_WHITE) && nc >= g.min_mw && nb >= g.min_mw;
    const bool want_kw = pair && fm_mode_has(g.pairwise_mode, FM_PW_KRUSKAL) && nc >= g.min_kw && nb >= g.min_kw;
    const bool want_wx = pair && fm_mode_has(g.pairwise_mode, FM_PW_WILCOXON);
    const bool ranks = want_mw || want_kw;
    const bool fuse_model = (a.hist == nullptr) && g.ma_mode == FM_MA_ALL;

    // ---- current row (time order for the flags, canonical -0 -> +0 for the ranks) ------------
    const bool va = (uint32_t)lane < nc, vb = (uint32_t)lane + 32u < nc;
    const float craw_a = va ? fm_lds_f32(stage_s + 4u * lane) : 0.0f;
    const float craw_b = vb ? fm_lds_f32(stage_s + 4u * (lane + 32u)) : 0.0f;
    const float ca = va ? craw_a + 0.0f : CUDART_INF_F;
    const float cb = vb ? craw_b + 0.0f : CUDART_INF_F;

    // ---- Wilcoxon (pairs need both rows in time order) -----------------------------------------
    WxInts wx = {0u, 0u, 0u, 0u, false};
    if (want_wx) wx = wilcoxon_ints(g, wilcoxon_load(g, stage_s, nc, nb, lane), lane);
    ax.n_wx = wx.np;

    // ---- model on a window / smoothing recurrence reads the staged row in time order ----------
    bool model_ok = false;
    if (!a.hist && !fuse_model) {
      const ModelOut mo = model_call(&g, stage + kCur, nb, false, lane);
      model_ok = mo.ok; rec.mean = mo.mean; rec.sigma = mo.sigma;
    }

    // ---- sorted current row -> search table (before sweep 1: what the row looks like decides what
    // sweep 1 builds) ------------------------------------------------------------------------------
    uint32_t cur_runs = 65u;                              // distinct values of the current row (+1 for its padding)
    float c31 = 0.f, c15 = 0.f, c47 = 0.f;
    if (ranks) {
      float sa = ca, sb = cb;
      sort64(sa, sb, lane);
      {
        const float qa = __shfl_up_sync(FM_FULL, sa, 1);
        const float last_a = __shfl_sync(FM_FULL, sa, 31);
        float qb = __shfl_up_sync(FM_FULL, sb, 1);
        if (lane == 0) qb = last_a;
        const uint32_t ba = __ballot_sync(FM_FULL, lane == 0 || qa != sa), bb = __ballot_sync(FM_FULL, qb != sb);
        uint32_t lo_a, e_a, lo_b, e_b;
        runs_from_ballots(ba, bb, lane, lo_a, e_a, lo_b, e_b);
        cur_runs = (uint32_t)__popc(ba) + (uint32_t)__popc(bb);
        // run length counts real samples only (the +inf padding behind position nc is not a sample)
        const uint32_t rl_a = min(e_a, nc) - min(lo_a, nc), rl_b = min(e_b, nc) - min(lo_b, nc);
        fm_sts_f32(search_s + 4u * lane, sa);
        fm_sts_f32(search_s + 4u * (lane + 33u), sb);
        sts_v2(final_s + 8u * lane, __float_as_uint(sa), (rl_a << 16) | (2u * lane));
        sts_v2(final_s + 8u * (lane + 33u), __float_as_uint(sb), (rl_b << 16) | (2u * (lane + 32u)));
      }
      c31 = __shfl_sync(FM_FULL, sa, 31);
      c15 = __shfl_sync(FM_FULL, sa, 15);
      c47 = __shfl_sync(FM_FULL, sb, 15);
    }
    // few distinct current values, most of them repeated: probably an error-count series, whose rank
    // integers come from its classes (ranks_by_classes): sweep 1 then elects instead of counting
    const bool tie_cand = ranks && (a.tie_runs >= 65u || (cur_runs <= a.tie_runs && 2u * cur_runs <= nc));   // 65: every series (tests)

    // ---- sweep 1 over the staged baseline: moments, election pass A -------------------------------
    // (the keys are NOT kept: sweep 2 reads the stage again, so that a lane holds a few keys at a
    // time and both sweeps are short rolled loops that stay in the instruction cache)
    const double pivot = nb ? (double)fm_lds_f32(base_s) : 0.0;
    FmMom mom = {0.0, 0.0};
    const float nokey = __uint_as_float(kNoKey);
    const uint32_t mult0 = mult_of(0);
    const uint32_t hka = va ? __float_as_uint(ca) : kNoKey, hkb = vb ? __float_as_uint(cb) : kNoKey;
    const uint32_t nfull = min(NV, nb >> 7);                  // vector rows without a missing point
    const uint32_t m16 = 4u * m_slots;                        // 16-bit counters in the table region
    auto cnt_add_all = [&](uint32_t bits) { cnt_add(tbl_s, bits, 4u * m_slots); };
    auto pass_a = [&](uint32_t bits) {                        // sweep 1's part of the multiset
#if FM_COUNT_PREFILTER
      if (tie_cand) fm_sts_u32(tbl_s + slot_of(bits, mult0, m_slots), bits);
      else cnt_add(tbl_s, bits, m16);
#else
      fm_sts_u32(tbl_s + slot_of(bits, mult0, m_slots), bits);
#endif
    };
    if (ranks) {
#if FM_COUNT_PREFILTER
      for (uint32_t o = (uint32_t)lane * 16u; o < 2u * c_off; o += 512u) sts_zero16(tbl_s + o);
      __syncwarp();
#else
      for (uint32_t o = (uint32_t)lane * 16u; o < c_off; o += 512u) sts_zero16(tbl_s + c_off + o);
#endif
      pass_a(hka);
      pass_a(hkb);
    }
    if (ranks || fuse_model) {
      auto row1 = [&](uint32_t c) {
        const float4 q = fm_lds_v4(base_s + (((uint32_t)lane + 32u * c) << 4));
        if (fuse_model) {
          fm_mom_add(mom, q.x, pivot); fm_mom_add(mom, q.y, pivot);
          fm_mom_add(mom, q.z, pivot); fm_mom_add(mom, q.w, pivot);
        }
        if (ranks) {
          const uint32_t k0 = __float_as_uint(q.x + 0.0f), k1 = __float_as_uint(q.y + 0.0f),
                         k2 = __float_as_uint(q.z + 0.0f), k3 = __float_as_uint(q.w + 0.0f);
          pass_a(k0); pass_a(k1); pass_a(k2); pass_a(k3);
        }
      };
#pragma unroll 1
      for (uint32_t c = 0; c < nfull; ++c) row1(c);
#pragma unroll 1
      for (uint32_t c = nfull; c < NV; ++c) {                 // ragged rows: the tail of the row and what lies behind it
        float k[4];
        float4 q;
        const uint32_t e0 = keys4(base_s, c, lane, nb, nokey, k, q);
        if (fuse_model) {
          if (e0 < nb) fm_mom_add(mom, q.x, pivot);
          if (e0 + 1u < nb) fm_mom_add(mom, q.y, pivot);
          if (e0 + 2u < nb) fm_mom_add(mom, q.z, pivot);
          if (e0 + 3u < nb) fm_mom_add(mom, q.w, pivot);
        }
        if (ranks) {
#pragma unroll
          for (int i = 0; i < 4; ++i) pass_a(__float_as_uint(k[i]));
        }
      }
#pragma unroll 1
      for (uint32_t j = 0; j < NS; ++j) {
        const uint32_t e = 128u * NV + (uint32_t)lane + 32u * j;
        float v = nokey;
        if (e < nb) {
          const float q = fm_lds_f32(base_s + (e << 2));
          v = q + 0.0f;
          if (fuse_model) fm_mom_add(mom, q, pivot);
        }
        if (ranks) pass_a(__float_as_uint(v));
      }
    }
    if (a.hist) {
      const ModelOut mo = model_call(&g, fm_hist_row(a, s), row_len(a.hist_len, s, a.n_hist), true, lane);
      model_ok = mo.ok; rec.mean = mo.mean; rec.sigma = mo.sigma;
    } else if (fuse_model)
      model_ok = nb >= max(g.min_hist_points, g.ddof + 1u);
    if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
    double s1 = 0.0, s2 = 0.0;
    if (fuse_model && model_ok) { s1 = fm_warp_sum(mom.s1); s2 = fm_warp_sum(mom.s2); }

    long long r2c = 0, Tt = 0;
    if (ranks) {

      // ---- sweep 2: every key's place in the sorted current row + election pass B ---------------
      // The overflow lists grow into the stage BEHIND the read position: a lane has produced at most
      // one entry per key it has consumed, entry k of a lane sits in stage word 32 k + lane, and the
      // 2 + 4 c + 4 rows of 32 words behind vector row c have been read by every lane by then.
      __syncwarp();                                         // pass A and the search table are complete
      bool tie_done = false;
      if (tie_cand) {
        const TieOut to = ranks_by_classes(stage_s, base_s, tbl_s, c_off, m_slots, NV, NS, nb, hka, hkb, va, vb, lane);
        if (to.ok) { r2c = to.r2c; Tt = to.T; ax.r2_cur = r2c; ax.t_mw = Tt; tie_done = true; }
#if FM_COUNT_PREFILTER
        else {                                              // the class path used the table: count again
          for (uint32_t o = (uint32_t)lane * 16u; o < 2u * c_off; o += 512u) sts_zero16(tbl_s + o);
          __syncwarp();
          cnt_add_all(hka);
          cnt_add_all(hkb);
#pragma unroll 1
          for (uint32_t c = 0; c < NV; ++c) {
            float k[4];
            float4 q;
            keys4(base_s, c, lane, nb, nokey, k, q);
#pragma unroll
            for (int i = 0; i < 4; ++i) cnt_add_all(__float_as_uint(k[i]));
          }
#pragma unroll 1
          for (uint32_t j = 0; j < NS; ++j) {
            const uint32_t e = 128u * NV + (uint32_t)lane + 32u * j;
            cnt_add_all(__float_as_uint(e < nb ? fm_lds_f32(base_s + (e << 2)) + 0.0f : nokey));
          }
          __syncwarp();
        }
#endif
      }
      if (!tie_done) {
      Tie tie = {0u, 0u};
      const uint32_t list0 = stage_s + 4u * lane;
      uint32_t lptr = list0;
      uint32_t acc_all = 0, acc_eq = 0;
      const uint32_t final_k = final_s - 2u * search_s;     // entry address = final_k + 2 * probe address
#if FM_COUNT_PREFILTER
      {
        uint32_t sha, shb;
        const uint32_t aa = cnt_addr(tbl_s, hka, m16, sha), ab = cnt_addr(tbl_s, hkb, m16, shb);
        const uint32_t wa = fm_lds_u32(aa), wb = fm_lds_u32(ab);
        list_unless_single(wa, sha, hka, lptr);
        list_unless_single(wb, shb, hkb, lptr);
      }
#define FM_RANK_GROUP(G, K) rank_group_p<G>(K, search_s, search_s + 132u, final_k, c31, c15, c47, tbl_s, m16, acc_all, acc_eq, lptr)
#else
      {
        const uint32_t ck[2] = {hka, hkb};
        classify<2>(ck, tbl_s, c_off, mult0, m_slots, tie, lptr);
      }
#define FM_RANK_GROUP(G, K) rank_group<G>(K, search_s, search_s + 132u, final_k, c31, tbl_s, c_off, mult0, m_slots, acc_all, acc_eq, tie, lptr)
#endif
      uint32_t c = 0;
      auto rows2 = [&](uint32_t cc) {                         // two full vector rows: eight probe chains in flight
        const float4 q0 = fm_lds_v4(base_s + (((uint32_t)lane + 32u * cc) << 4));
        const float4 q1 = fm_lds_v4(base_s + (((uint32_t)lane + 32u * cc + 32u) << 4));
        const float k[8] = {q0.x + 0.0f, q0.y + 0.0f, q0.z + 0.0f, q0.w + 0.0f,
                            q1.x + 0.0f, q1.y + 0.0f, q1.z + 0.0f, q1.w + 0.0f};
        FM_RANK_GROUP(8, k);
      };
#pragma unroll 1
      for (; c + 2u <= nfull; c += 2u) rows2(c);
#pragma unroll 1
      for (; c < NV; ++c) {
        float k[4];
        float4 q;
        keys4(base_s, c, lane, nb, nokey, k, q);
        FM_RANK_GROUP(4, k);
      }
#pragma unroll 1
      for (uint32_t j = 0; j < NS; ++j) {
        const uint32_t e = 128u * NV + (uint32_t)lane + 32u * j;
        const float k = e < nb ? fm_lds_f32(base_s + (e << 2)) + 0.0f : nokey;
        FM_RANK_GROUP(1, &k);
      }

#undef FM_RANK_GROUP
      // ---- overflow rounds -------------------------------------------------------------------------
      uint32_t n_l = (lptr - list0) >> 7;
      unsigned long long t_extra = 0;
      for (uint32_t round = 1;; ++round) {
        if (__reduce_max_sync(FM_FULL, n_l) == 0u) break;
        __syncwarp();                                       // the previous pass B is complete
        if (round == kMaxRounds) {
          // Not reached in practice.  What is left are whole classes: count each key's class directly.
          fm_sts_u32(search_s + 4u * lane, n_l);
          __syncwarp();
          for (uint32_t k = 0; k < n_l; ++k) {
            const uint32_t mine = fm_lds_u32(list0 + 128u * k);
            uint32_t t = 0;
            for (uint32_t l2 = 0; l2 < 32u; ++l2) {
              const uint32_t n2 = fm_lds_u32(search_s + 4u * l2);
              for (uint32_t k2 = 0; k2 < n2; ++k2) t += fm_lds_u32(stage_s + 4u * l2 + 128u * k2) == mine;
            }
            t_extra += (unsigned long long)t * t - 1ull;
          }
          break;
        }
        for (uint32_t o = (uint32_t)lane * 16u; o < c_off; o += 512u) sts_zero16(tbl_s + c_off + o);
        const uint32_t mult = mult_of(round);
        {
          uint32_t k = 0;
          for (; k + 4u <= n_l; k += 4u) {
            uint32_t bits[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) bits[i] = fm_lds_u32(list0 + 128u * (k + i));
#pragma unroll
            for (int i = 0; i < 4; ++i) fm_sts_u32(tbl_s + slot_of(bits[i], mult, m_slots), bits[i]);
          }
          for (; k < n_l; ++k) {
            const uint32_t bits = fm_lds_u32(list0 + 128u * k);
            fm_sts_u32(tbl_s + slot_of(bits, mult, m_slots), bits);
          }
        }
        __syncwarp();
        uint32_t wptr = list0;
        {
          uint32_t k = 0;                                   // in place: wptr never passes the read position
          for (; k + 4u <= n_l; k += 4u) {
            uint32_t bits[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) bits[i] = fm_lds_u32(list0 + 128u * (k + i));
            classify<4>(bits, tbl_s, c_off, mult, m_slots, tie, wptr);
          }
          for (; k < n_l; ++k) {
            const uint32_t bits = fm_lds_u32(list0 + 128u * k);
            classify<1>(&bits, tbl_s, c_off, mult, m_slots, tie, wptr);
          }
        }
        n_l = (wptr - list0) >> 7;
      }

      // ---- integers ---------------------------------------------------------------------------------
      const uint32_t sum_lu = __reduce_add_sync(FM_FULL, (acc_all & 0xFFFFu) + (acc_eq >> 16));
      r2c = (long long)nc * (nc + 1) + 2ll * nc * nb - (long long)sum_lu;
      const unsigned long long tl = (unsigned long long)tie.sq + (0u - tie.lin);
      const unsigned long long t3 = (unsigned long long)__reduce_add_sync(FM_FULL, (uint32_t)(tl & 0xFFFFu)) +
                                    ((unsigned long long)__reduce_add_sync(FM_FULL, (uint32_t)(tl >> 16)) << 16);
      const unsigned long long ninv = (unsigned long long)(kCap - nb) + (64u - nc);
      Tt = (long long)(3ull * t3 + (unsigned long long)fm_warp_sum((long long)t_extra) - (ninv * ninv * ninv - ninv));
      ax.r2_cur = r2c; ax.t_mw = Tt;
      }
    }

    // the stage (rows, then lists) has been consumed: hand it to the TMA unit for the next series
    __syncwarp();
    if (sn < a.n_series) {
      fm_fence_proxy_async();
      if (lane == 0)
        issue_rows(a, stage, bar, sn, row_len(a.cur_len, sn, a.n_cur), row_len(a.base_len, sn, a.n_base));
    }

    finish_stats(a, g, tp, s, lane, nc, nb, pair, want_mw, want_kw, wx, fuse_model, model_ok, s1, s2, pivot, r2c, Tt,
                 craw_a, craw_b, va, vb, rec, ax, status);
    __syncwarp();
  }
}

// =================================================================================================
// k_canary_team — the same algorithm with SEVERAL warps per series (one CTA per series), for
// baselines above 2048 points (BASELINE.json configs[1] read literally: Mann-Whitney of 60 points
// against the 10080-point window).  A 10080-point row needs 40 KB of stage and a 56 KB multiset
// table: a warp on its own would own 100 KB and two warps would fill an SM.  A team of W = 8 warps
// shares them: 16 warps per SM, every lane still handles ~40 keys, and U stays a search per key, so
// the cost per key is that of the 1440-point kernel.
//   warp 0    sorts the current row and builds the search tables, then turns the integers into the
//             record (a5..a9);
//   warp 1    runs the Wilcoxon test and the separate-history model;
//   all       rank their share of the baseline's vector rows (warp w: rows [NV w / W, NV (w+1) / W));
//             the last warp also takes the scalar rows and the current points.
// CTA barriers: after pass A, around every overflow round, before the hand-over of the partial
// sums.  Each warp's overflow lists grow into ITS part of the stage behind its own read position.
//
// Stage (floats): [current 64][pad 32][baseline cap][pad 64]; warp 0's lists start at word 0, warp
// w's at the first word of its first vector row.
constexpr uint32_t kPairCurPad = 96;     // floats before the baseline row
constexpr uint32_t kPairTailPad = 64;    // floats behind it
constexpr uint32_t kTeamMaxWarps = 16;
constexpr uint32_t kXchgBytes = kTeamMaxWarps * 96;

struct PairXchg {                         // warp w -> warp 0 (one entry per warp)
  double s1, s2, mean, sigma;
  unsigned long long t3, t_extra;
  uint32_t sum_lu, np, wp2, wm2, twx, flags;   // flags: bit 0 Wilcoxon ran, bit 1 model ok
};

__device__ __forceinline__ void issue_rows_pair(const FmKArgs& a, float* stage, uint64_t* bar, uint64_t s,
                                                uint32_t nc, uint32_t nb) {
  const uint32_t cb = ((nc + 3u) & ~3u) * 4u, bb = ((nb + 3u) & ~3u) * 4u;
  fm_mbar_expect_tx(bar, cb + bb);
  if (cb) fm_bulk_g2s(stage, a.cur + s * a.cur_pitch, cb, bar);
  if (bb) fm_bulk_g2s(stage + kPairCurPad, fm_base_row(a, s), bb, bar);
}

__global__ void __launch_bounds__(512, 1) k_canary_team(const FmKArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t NV = a.cnt_nv, NS = a.cnt_ns;
  const uint32_t kCap = 128u * NV + 32u * NS;
  const int lane = threadIdx.x & 31;
  const uint32_t warp = threadIdx.x >> 5, W = blockDim.x >> 5;
  const bool w0 = warp == 0, w1 = warp == 1, wlast = warp == W - 1u;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  float* stage = reinterpret_cast<float*>(smem + 16);
  const uint32_t stage_s = fm_smem_u32(stage), base_s = stage_s + kPairCurPad * 4u;
  const uint32_t search_s = stage_s + (kPairCurPad + a.base_cap + kPairTailPad) * 4u;
  const uint32_t final_s = search_s + kSearchBytes;
  PairXchg* xchg = reinterpret_cast<PairXchg*>(smem + 16 + (kPairCurPad + a.base_cap + kPairTailPad) * 4u +
                                               kSearchBytes + kFinalBytes);
  const uint32_t tbl_s = final_s + kFinalBytes + kXchgBytes;
  const uint32_t m_slots = a.log_m, c_off = 4u * a.log_m;
  const fm_global_params& g = a.P->g;
  // my vector rows of the baseline; the last warp also takes the scalar rows and the current points
  const uint32_t r0 = NV * warp / W, r1 = NV * (warp + 1u) / W;
  const uint32_t list0 = (w0 ? stage_s : base_s + 512u * r0) + 4u * (uint32_t)lane;

  if (threadIdx.x == 0) { fm_mbar_init(bar, 1); fm_fence_mbar_init(); }
  __syncthreads();

  uint64_t s = blockIdx.x;
  uint32_t phase = 0;
  if (s < a.n_series && threadIdx.x == 0)
    issue_rows_pair(a, stage, bar, s, row_len(a.cur_len, s, a.n_cur), row_len(a.base_len, s, a.n_base));

  for (; s < a.n_series; s += gridDim.x) {
    const uint32_t nc = row_len(a.cur_len, s, a.n_cur);
    const uint32_t nb = row_len(a.base_len, s, a.n_base);
    const uint64_t sn = s + gridDim.x;
    if (sn < a.n_series) {     // pull the next rows into L2 while this series is ranked
      const uint32_t nbn = row_len(a.base_len, sn, a.n_base);
      const char* pb = reinterpret_cast<const char*>(fm_base_row(a, sn));
      for (uint32_t o = threadIdx.x * 128u; o < nbn * 4u; o += blockDim.x * 128u) fm_prefetch_l2(pb + o);
      if (threadIdx.x < 2) fm_prefetch_l2(reinterpret_cast<const char*>(a.cur + sn * a.cur_pitch) + threadIdx.x * 128);
    }
    fm_mbar_wait(bar, phase);
    phase ^= 1u;

    const bool pair = g.pairwise_mode != FM_PW_NONE && nc > 0 && nb > 0;
    const bool want_mw = pair && fm_mode_has(g.pairwise_mode, FM_PW_MANN_WHITE) && nc >= g.min_mw && nb >= g.min_mw;
    const bool want_kw = pair && fm_mode_has(g.pairwise_mode, FM_PW_KRUSKAL) && nc >= g.min_kw && nb >= g.min_kw;
    const bool want_wx = pair && fm_mode_has(g.pairwise_mode, FM_PW_WILCOXON);
    const bool ranks = want_mw || want_kw;
    const bool fuse_model = (a.hist == nullptr) && g.ma_mode == FM_MA_ALL;

    const bool va = (uint32_t)lane < nc, vb = (uint32_t)lane + 32u < nc;
    const float craw_a = va ? fm_lds_f32(stage_s + 4u * lane) : 0.0f;
    const float craw_b = vb ? fm_lds_f32(stage_s + 4u * (lane + 32u)) : 0.0f;
    const float ca = va ? craw_a + 0.0f : CUDART_INF_F;
    const float cb = vb ? craw_b + 0.0f : CUDART_INF_F;
    const uint32_t hka = va ? __float_as_uint(ca) : kNoKey, hkb = vb ? __float_as_uint(cb) : kNoKey;
    const uint32_t mult0 = mult_of(0);
    const float nokey = __uint_as_float(kNoKey);

    // ---- role phase: warp 0 sorts the current row, warp 1 runs Wilcoxon and the separate models ----
    WxInts wx = {0u, 0u, 0u, 0u, false};
    ModelOut mo = {0.0, 0.0, false};
    if (w0) {
      if (ranks) {
        float sa = ca, sb = cb;
        sort64(sa, sb, lane);
        const float qa = __shfl_up_sync(FM_FULL, sa, 1);
        const float last_a = __shfl_sync(FM_FULL, sa, 31);
        float qb = __shfl_up_sync(FM_FULL, sb, 1);
        if (lane == 0) qb = last_a;
        const uint32_t ba = __ballot_sync(FM_FULL, lane == 0 || qa != sa), bb = __ballot_sync(FM_FULL, qb != sb);
        uint32_t lo_a, e_a, lo_b, e_b;
        runs_from_ballots(ba, bb, lane, lo_a, e_a, lo_b, e_b);
        const uint32_t rl_a = min(e_a, nc) - min(lo_a, nc), rl_b = min(e_b, nc) - min(lo_b, nc);
        fm_sts_f32(search_s + 4u * lane, sa);
        fm_sts_f32(search_s + 4u * (lane + 33u), sb);
        sts_v2(final_s + 8u * lane, __float_as_uint(sa), (rl_a << 16) | (2u * lane));
        sts_v2(final_s + 8u * (lane + 33u), __float_as_uint(sb), (rl_b << 16) | (2u * (lane + 32u)));
      }
    } else if (w1) {
      if (want_wx) wx = wilcoxon_ints(g, wilcoxon_load(g, stage_s, nc, nb, lane, kPairCurPad), lane);
      if (a.hist) mo = model_call(&g, fm_hist_row(a, s), row_len(a.hist_len, s, a.n_hist), true, lane);
      else if (!fuse_model) mo = model_call(&g, stage + kPairCurPad, nb, false, lane);
    }

    // ---- sweep 1 over my rows: moments, election pass A --------------------------------------------
    const double pivot = nb ? (double)fm_lds_f32(base_s) : 0.0;
    FmMom mom = {0.0, 0.0};
    const uint32_t nfull = min(r1, max(r0, nb >> 7));        // my rows in [r0, nfull) have no missing point
    if (ranks) {
      for (uint32_t o = threadIdx.x * 16u; o < c_off; o += blockDim.x * 16u) sts_zero16(tbl_s + c_off + o);
      if (wlast) {
        fm_sts_u32(tbl_s + slot_of(hka, mult0, m_slots), hka);
        fm_sts_u32(tbl_s + slot_of(hkb, mult0, m_slots), hkb);
      }
    }
    if (ranks || fuse_model) {
#pragma unroll 1
      for (uint32_t c = r0; c < nfull; ++c) {
        const float4 q = fm_lds_v4(base_s + (((uint32_t)lane + 32u * c) << 4));
        if (fuse_model) {
          fm_mom_add(mom, q.x, pivot); fm_mom_add(mom, q.y, pivot);
          fm_mom_add(mom, q.z, pivot); fm_mom_add(mom, q.w, pivot);
        }
        if (ranks) {
          const uint32_t k0 = __float_as_uint(q.x + 0.0f), k1 = __float_as_uint(q.y + 0.0f),
                         k2 = __float_as_uint(q.z + 0.0f), k3 = __float_as_uint(q.w + 0.0f);
          fm_sts_u32(tbl_s + slot_of(k0, mult0, m_slots), k0);
          fm_sts_u32(tbl_s + slot_of(k1, mult0, m_slots), k1);
          fm_sts_u32(tbl_s + slot_of(k2, mult0, m_slots), k2);
          fm_sts_u32(tbl_s + slot_of(k3, mult0, m_slots), k3);
        }
      }
#pragma unroll 1
      for (uint32_t c = nfull; c < r1; ++c) {
        float k[4];
        float4 q;
        const uint32_t e0 = keys4(base_s, c, lane, nb, nokey, k, q);
        if (fuse_model) {
          if (e0 < nb) fm_mom_add(mom, q.x, pivot);
          if (e0 + 1u < nb) fm_mom_add(mom, q.y, pivot);
          if (e0 + 2u < nb) fm_mom_add(mom, q.z, pivot);
          if (e0 + 3u < nb) fm_mom_add(mom, q.w, pivot);
        }
        if (ranks) {
#pragma unroll
          for (int i = 0; i < 4; ++i) fm_sts_u32(tbl_s + slot_of(__float_as_uint(k[i]), mult0, m_slots), __float_as_uint(k[i]));
        }
      }
      if (wlast) {
#pragma unroll 1
        for (uint32_t j = 0; j < NS; ++j) {
          const uint32_t e = 128u * NV + (uint32_t)lane + 32u * j;
          float v = nokey;
          if (e < nb) {
            const float q = fm_lds_f32(base_s + (e << 2));
            v = q + 0.0f;
            if (fuse_model) fm_mom_add(mom, q, pivot);
          }
          if (ranks) fm_sts_u32(tbl_s + slot_of(__float_as_uint(v), mult0, m_slots), __float_as_uint(v));
        }
      }
    }
    __syncthreads();                                         // pass A, counters and search tables complete

    // ---- sweep 2 over my rows, overflow rounds ----------------------------------------------------------
    Tie tie = {0u, 0u};
    unsigned long long tsq = 0, tlin = 0;              // long rows: p^2 sums leave 32 bits, so flush per group
    auto flush = [&]() { tsq += tie.sq; tlin += 0u - tie.lin; tie.sq = 0u; tie.lin = 0u; };
    uint32_t acc_all = 0, acc_eq = 0;
    unsigned long long t_extra = 0;
    if (ranks) {
      const float c31 = fm_lds_f32(search_s + 124u);
      const uint32_t final_k = final_s - 2u * search_s;
      uint32_t lptr = list0;
      uint32_t c = r0;
#pragma unroll 1
      for (; c + 2u <= nfull; c += 2u) {
        const float4 q0 = fm_lds_v4(base_s + (((uint32_t)lane + 32u * c) << 4));
        const float4 q1 = fm_lds_v4(base_s + (((uint32_t)lane + 32u * c + 32u) << 4));
        const float k[8] = {q0.x + 0.0f, q0.y + 0.0f, q0.z + 0.0f, q0.w + 0.0f,
                            q1.x + 0.0f, q1.y + 0.0f, q1.z + 0.0f, q1.w + 0.0f};
        rank_group<8>(k, search_s, search_s + 132u, final_k, c31, tbl_s, c_off, mult0, m_slots, acc_all, acc_eq, tie, lptr);
        flush();
      }
#pragma unroll 1
      for (; c < r1; ++c) {
        float k[4];
        float4 q;
        keys4(base_s, c, lane, nb, nokey, k, q);
        rank_group<4>(k, search_s, search_s + 132u, final_k, c31, tbl_s, c_off, mult0, m_slots, acc_all, acc_eq, tie, lptr);
        flush();
      }
      if (wlast) {
#pragma unroll 1
        for (uint32_t j = 0; j < NS; ++j) {
          const uint32_t e = 128u * NV + (uint32_t)lane + 32u * j;
          const float k = e < nb ? fm_lds_f32(base_s + (e << 2)) + 0.0f : nokey;
          rank_group<1>(&k, search_s, search_s + 132u, final_k, c31, tbl_s, c_off, mult0, m_slots, acc_all, acc_eq, tie, lptr);
          flush();
        }
        const uint32_t ck[2] = {hka, hkb};                     // the current points last: their list rows are the tail pad
        classify<2>(ck, tbl_s, c_off, mult0, m_slots, tie, lptr);
        flush();
      }

      uint32_t n_l = (lptr - list0) >> 7;
      for (uint32_t round = 1;; ++round) {
        if (!__syncthreads_or(n_l > 0u)) break;                // also: everybody's previous pass B is complete
        if (round == kMaxRounds) {
          // Not reached in practice.  What is left are whole classes: count each key's class directly.
          // (the tables are dead by now: their first words take every thread's list length and base)
          fm_sts_u32(tbl_s + 4u * threadIdx.x, n_l);
          fm_sts_u32(tbl_s + 2048u + 4u * threadIdx.x, list0);
          __syncthreads();
          for (uint32_t k = 0; k < n_l; ++k) {
            const uint32_t mine = fm_lds_u32(list0 + 128u * k);
            uint32_t t = 0;
            for (uint32_t l2 = 0; l2 < blockDim.x; ++l2) {
              const uint32_t n2 = fm_lds_u32(tbl_s + 4u * l2), b2 = fm_lds_u32(tbl_s + 2048u + 4u * l2);
              for (uint32_t k2 = 0; k2 < n2; ++k2) t += fm_lds_u32(b2 + 128u * k2) == mine;
            }
            t_extra += (unsigned long long)t * t - 1ull;
          }
          break;
        }
        for (uint32_t o = threadIdx.x * 16u; o < c_off; o += blockDim.x * 16u) sts_zero16(tbl_s + c_off + o);
        const uint32_t mult = mult_of(round);
        {
          uint32_t k = 0;
          for (; k + 4u <= n_l; k += 4u) {
            uint32_t bits[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) bits[i] = fm_lds_u32(list0 + 128u * (k + i));
#pragma unroll
            for (int i = 0; i < 4; ++i) fm_sts_u32(tbl_s + slot_of(bits[i], mult, m_slots), bits[i]);
          }
          for (; k < n_l; ++k) {
            const uint32_t bits = fm_lds_u32(list0 + 128u * k);
            fm_sts_u32(tbl_s + slot_of(bits, mult, m_slots), bits);
          }
        }
        __syncthreads();
        uint32_t wptr = list0;
        {
          uint32_t k = 0;
          for (; k + 4u <= n_l; k += 4u) {
            uint32_t bits[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) bits[i] = fm_lds_u32(list0 + 128u * (k + i));
            classify<4>(bits, tbl_s, c_off, mult, m_slots, tie, wptr);
            flush();
          }
          for (; k < n_l; ++k) {
            const uint32_t bits = fm_lds_u32(list0 + 128u * k);
            classify<1>(&bits, tbl_s, c_off, mult, m_slots, tie, wptr);
            flush();
          }
        }
        n_l = (wptr - list0) >> 7;
      }
    }

    // ---- partial sums; warp 1 hands its own over -------------------------------------------------------
    uint32_t sum_lu = 0;
    unsigned long long t3 = 0, tx = 0;
    if (ranks) {
      sum_lu = __reduce_add_sync(FM_FULL, (acc_all & 0xFFFFu) + (acc_eq >> 16));
      t3 = (unsigned long long)fm_warp_sum((long long)(tsq + tlin));
      tx = (unsigned long long)fm_warp_sum((long long)t_extra);
    }
    double s1 = 0.0, s2 = 0.0;
    if (fuse_model) { s1 = fm_warp_sum(mom.s1); s2 = fm_warp_sum(mom.s2); }
    if (!w0 && lane == 0) {
      PairXchg& x = xchg[warp];
      x.s1 = s1; x.s2 = s2; x.mean = mo.mean; x.sigma = mo.sigma;
      x.t3 = t3; x.t_extra = tx; x.sum_lu = sum_lu;
      x.np = wx.np; x.wp2 = wx.wp2; x.wm2 = wx.wm2; x.twx = wx.t;
      x.flags = (wx.run ? 1u : 0u) | (mo.ok ? 2u : 0u);
    }
    __syncthreads();                                         // hand-over visible; nobody reads the stage any more
    if (sn < a.n_series && threadIdx.x == 0) {
      fm_fence_proxy_async();
      issue_rows_pair(a, stage, bar, sn, row_len(a.cur_len, sn, a.n_cur), row_len(a.base_len, sn, a.n_base));
    }
    if (!w0) continue;                                       // the others go on to the next series

    // ---- warp 0: a5..a9 -----------------------------------------------------------------------------------
    PairXchg x1 = xchg[1];                                   // warp 1: Wilcoxon and the model; all: partial sums
    for (uint32_t w = 2; w < W; ++w) {
      x1.s1 += xchg[w].s1; x1.s2 += xchg[w].s2; x1.t3 += xchg[w].t3; x1.t_extra += xchg[w].t_extra;
      x1.sum_lu += xchg[w].sum_lu;
    }
    const uint32_t ty = a.metric_type ? a.metric_type[s] : 0u;
    const fm_type_params tp = a.P->types[ty < a.P->n_types ? ty : 0u];
    fm_series_rec rec;
    fm_rec_init(rec);
    fm_rank_aux ax = {0, 0, 0, 0, 0, 0u, 0u};
    uint32_t status = 0;
    if (nc == 0) status |= FM_ST_NO_CURRENT;
    wx.np = x1.np; wx.wp2 = x1.wp2; wx.wm2 = x1.wm2; wx.t = x1.twx; wx.run = (x1.flags & 1u) != 0;
    ax.n_wx = wx.np;
    bool model_ok;
    if (fuse_model) {
      model_ok = nb >= max(g.min_hist_points, g.ddof + 1u);
      s1 += x1.s1; s2 += x1.s2;
    } else {
      model_ok = (x1.flags & 2u) != 0;
      rec.mean = x1.mean; rec.sigma = x1.sigma;
    }
    if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
    long long r2c = 0, Tt = 0;
    if (ranks) {
      r2c = (long long)nc * (nc + 1) + 2ll * nc * nb - (long long)(sum_lu + x1.sum_lu);
      const unsigned long long ninv = (unsigned long long)(kCap - nb) + (64u - nc);
      Tt = (long long)(3ull * (t3 + x1.t3) + tx + x1.t_extra - (ninv * ninv * ninv - ninv));
      ax.r2_cur = r2c; ax.t_mw = Tt;
    }
    finish_stats(a, g, tp, s, lane, nc, nb, pair, want_mw, want_kw, wx, fuse_model, model_ok, s1, s2, pivot, r2c, Tt,
                 craw_a, craw_b, va, vb, rec, ax, status);
  }
}

}  // namespace

// tiers (vector rows, scalar rows) -> capacity 128 NV + 32 NS baseline points.  The rows are walked
// by rolled loops, so one kernel serves every tier; a tier only sizes the stage and balances the
// keys over the lanes (1440 points = 11 vector rows + 1 scalar row = 45 keys in every lane).
#define FM_COUNT_TIERS(X) X(1, 0) X(2, 0) X(4, 0) X(8, 0) X(11, 1) X(12, 0) X(16, 0)

cudaError_t fm_canary_count_smem_attr(int bytes) {
  cudaError_t e = cudaFuncSetAttribute(k_canary_count, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_canary_team, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  return e;
}

// smallest tier that holds n_base points; 0 if none does.  Above 2048 points (k_canary_team) the
// tier is the row itself: whole vector rows plus the scalar rows of the remainder.
uint32_t fm_canary_count_tier(uint32_t n_base, int* nv, int* ns) {
#define X(NV, NS) if (n_base <= 128u * NV + 32u * NS) { *nv = NV; *ns = NS; return 128u * NV + 32u * NS; }
  FM_COUNT_TIERS(X)
#undef X
  if (n_base <= 16320u) {
    *nv = (int)(n_base / 128u);
    *ns = (int)((n_base % 128u + 31u) / 32u);
    return 128u * (uint32_t)*nv + 32u * (uint32_t)*ns;
  }
  return 0;
}

size_t fm_canary_count_fixed_bytes() { return 16 + kSearchBytes + kFinalBytes; }
int fm_canary_count_max_warps() { return FM_COUNT_MAXW; }

int fm_launch_canary_count(const FmKArgs& a, int warps_per_cta, int grid, cudaStream_t st) {
  const size_t smem = (size_t)warps_per_cta * a.warp_bytes;
  k_canary_count<<<grid, warps_per_cta * 32, smem, st>>>(a);
  return 1;
}

// k_canary_team: bytes per CTA beside the baseline capacity and the multiset table
size_t fm_canary_team_fixed_bytes() {
  return 16 + (size_t)(kPairCurPad + kPairTailPad) * 4 + kSearchBytes + kFinalBytes + kXchgBytes;
}

// a.warp_bytes = bytes of shared memory per CTA (= per series in flight)
int fm_launch_canary_team(const FmKArgs& a, int warps, int grid, cudaStream_t st) {
  k_canary_team<<<grid, warps * 32, a.warp_bytes, st>>>(a);
  return 1;
}
