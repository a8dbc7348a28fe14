// fm_device.cuh — device-side building blocks shared by the scoring kernels.
//
// Formula references (A.x) are DESIGN.md section 3; the reference text each one
// implements is cited there and in include/foremast_cuda.h.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "fm_internal.h"

#define FM_FULL 0xffffffffu

// ------------------------------------------------------------------ PTX ----
__device__ __forceinline__ uint32_t fm_smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void fm_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(fm_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fm_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fm_smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void fm_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "FM_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra FM_DONE_%=;\n"
      "bra FM_WAIT_%=;\n"
      "FM_DONE_%=:\n"
      "}\n" ::"r"(fm_smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D bulk copy global -> shared through the TMA unit; completion on `bar`.
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void fm_bulk_g2s(void* dst, const void* src, uint32_t bytes,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(fm_smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(fm_smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fm_fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void fm_fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// Explicit shared-space accesses on 32-bit shared addresses.  Pointers carved out of the dynamic
// shared array at run time reach ptxas as generic pointers (LD / ATOM.E.*.GPU, which for atomics
// is an order of magnitude slower than ATOMS); the hot loops therefore use these.
__device__ __forceinline__ uint32_t fm_lds_u32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ float fm_lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ float4 fm_lds_v4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(a)
               : "memory");
  return v;
}
__device__ __forceinline__ void fm_sts_v4(uint32_t a, float x, float y, float z, float w) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ void fm_sts_u32(uint32_t a, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void fm_sts_f32(uint32_t a, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
}
__device__ __forceinline__ uint32_t fm_atoms_cas(uint32_t a, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.shared.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "r"(a), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t fm_atoms_add(uint32_t a, uint32_t val) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(a), "r"(val) : "memory");
  return old;
}

__device__ __forceinline__ void fm_reds_add(uint32_t a, uint32_t val) {   // no return: fire and forget
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(val) : "memory");
}
__device__ __forceinline__ uint32_t fm_lds_u16(uint32_t a) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void fm_sts_u16(uint32_t a, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((uint16_t)v) : "memory");
}
// -1 when a < b else 0 (one FSET; three of them add up with one IADD3)
__device__ __forceinline__ int fm_set_lt(float a, float b) {
  int r;
  asm("set.lt.s32.f32 %0, %1, %2;" : "=r"(r) : "f"(a), "f"(b));
  return r;
}

__device__ __forceinline__ void fm_prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
// 1.0f when a == b else 0.0f (one FSET)
__device__ __forceinline__ float fm_setf_eq(float a, float b) {
  float r;
  asm("set.eq.f32.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
// a product the compiler may not turn into a select
__device__ __forceinline__ float fm_fmul(float a, float b) {
  float r;
  asm("mul.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

__device__ __forceinline__ float4 fm_ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// ------------------------------------------------------------- reductions --
__device__ __forceinline__ double fm_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FM_FULL, v, o);
  return v;
}
__device__ __forceinline__ long long fm_warp_sum(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FM_FULL, v, o);
  return v;
}
__device__ __forceinline__ uint32_t fm_warp_sum(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FM_FULL, v, o);
  return v;
}
__device__ __forceinline__ uint32_t fm_warp_min(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(FM_FULL, v, o));
  return v;
}

__device__ __forceinline__ uint32_t fm_pow2_ceil(uint32_t n, uint32_t lo) {
  uint32_t p = lo;
  while (p < n) p <<= 1;
  return p;
}

// t^3 - t
__device__ __forceinline__ long long fm_tie(uint32_t t) {
  const long long x = (long long)t;
  return x * x * x - x;
}

// ------------------------------------------------------------- moments -----
// Baseline / history row of series s: in the caller's matrix, or in place in the row cache.
__device__ __forceinline__ const float* fm_base_row(const FmKArgs& a, uint64_t s) {
  if (a.row_slot) {
    const uint32_t sl = __ldg(a.row_slot + s);
    if (sl != 0xFFFFFFFFu) return a.cache_base + (uint64_t)sl * a.base_pitch;
  }
  return a.base + s * a.base_pitch;
}
__device__ __forceinline__ const float* fm_hist_row(const FmKArgs& a, uint64_t s) {
  if (a.row_slot) {
    const uint32_t sl = __ldg(a.row_slot + s);
    if (sl != 0xFFFFFFFFu) return a.cache_hist + (uint64_t)sl * a.hist_pitch;
  }
  return a.hist + s * a.hist_pitch;
}

// A.1: mean / sigma of a window with a pivot (the first sample) subtracted
// before squaring so that sum(d^2) - sum(d)^2/n does not cancel
// (SURVEY.md 7.2-4).  Partial sums are per lane; fm_moments_finish reduces.
struct FmMom { double s1, s2; };

__device__ __forceinline__ void fm_mom_add(FmMom& m, float x, double pivot) {
  const double d = (double)x - pivot;
  m.s1 += d;
  m.s2 = fma(d, d, m.s2);
}

__device__ __forceinline__ void fm_moments_finish(FmMom m, double pivot, uint32_t n, uint32_t ddof,
                                                  double& mean, double& sigma) {
  const double s1 = fm_warp_sum(m.s1);
  const double s2 = fm_warp_sum(m.s2);
  const double dn = (double)n;
  mean = pivot + s1 / dn;
  double m2 = s2 - s1 * s1 / dn;
  if (m2 < 0.0) m2 = 0.0;
  sigma = (n > ddof) ? sqrt(m2 / (double)(n - ddof)) : 0.0;
}

// whole-warp moments over a GLOBAL row (16-byte aligned start), streaming loads
__device__ __forceinline__ void fm_moments_global(const float* __restrict__ x, uint32_t n,
                                                  uint32_t ddof, int lane, double& mean,
                                                  double& sigma) {
  FmMom m = {0.0, 0.0};
  const uint32_t n4 = n >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  // software pipelined: the next four 16-byte loads of a lane are in flight while the current
  // four are accumulated
  float4 q[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    q[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (lane + 32u * j < n4) q[j] = fm_ldg_stream(x4 + lane + 32u * j);
  }
  // pivot = x[0]: lane 0 holds it in its first vector (no load of its own in front of the stream)
  const float p0 = __shfl_sync(FM_FULL, q[0].x, 0);
  const double pivot = (double)(n4 ? p0 : __ldg(x));
  for (uint32_t i = lane; i < n4; i += 128) {
    float4 r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      r[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      const uint32_t v = i + 32u * (4 + j);
      if (v < n4) r[j] = fm_ldg_stream(x4 + v);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (i + 32u * j < n4) {
        fm_mom_add(m, q[j].x, pivot); fm_mom_add(m, q[j].y, pivot);
        fm_mom_add(m, q[j].z, pivot); fm_mom_add(m, q[j].w, pivot);
      }
      q[j] = r[j];
    }
  }
  const uint32_t t = (n4 << 2) + lane;
  if (t < n) fm_mom_add(m, __ldg(x + t), pivot);
  fm_moments_finish(m, pivot, n, ddof, mean, sigma);
}

// whole-warp moments over an arbitrarily aligned row (shared or global), scalar loads
__device__ __forceinline__ void fm_moments_any(const float* x, uint32_t n, uint32_t ddof, int lane,
                                               double& mean, double& sigma) {
  const double pivot = (double)x[0];
  FmMom m = {0.0, 0.0};
  for (uint32_t i = lane; i < n; i += 32) fm_mom_add(m, x[i], pivot);
  fm_moments_finish(m, pivot, n, ddof, mean, sigma);
}

// Exponential / double exponential smoothing (design.md:60-66), A.2: centre = one-step-ahead
// forecast after the last sample, sigma = RMS one-step forecast error.  The recurrence
//   level' = (1-a)(level + trend) + a x,   trend' = (1-ab) trend - ab level + ab x
// is linear and time-invariant in the state (level, trend), so the warp evaluates it as a scan:
// lane l owns a contiguous chunk of steps, pass 1 runs the chunk from the zero state (response r)
// and from the two unit states (the columns of M^len), a 5-step warp scan composes the affine maps
// (P, r) into every lane's true start state, pass 2 re-runs the chunk from that state and sums the
// squared forecast errors.  Same arithmetic as the sequential form up to fp64 rounding order.
struct FmAffine { double p00, p01, p10, p11, r0, r1; };
__device__ __forceinline__ FmAffine fm_affine_after(const FmAffine& first, const FmAffine& then) {
  // then o first
  FmAffine o;
  o.p00 = then.p00 * first.p00 + then.p01 * first.p10;
  o.p01 = then.p00 * first.p01 + then.p01 * first.p11;
  o.p10 = then.p10 * first.p00 + then.p11 * first.p10;
  o.p11 = then.p10 * first.p01 + then.p11 * first.p11;
  o.r0 = then.p00 * first.r0 + then.p01 * first.r1 + then.r0;
  o.r1 = then.p10 * first.r0 + then.p11 * first.r1 + then.r1;
  return o;
}
__device__ __forceinline__ void fm_model_smoothing(const fm_global_params& g, const float* x,
                                                   uint32_t n, int lane, double& mean,
                                                   double& sigma) {
  const bool dbl = g.ma_mode == FM_MODEL_DOUBLE_EXP;
  const double a = g.smooth_alpha, ab = dbl ? g.smooth_alpha * g.smooth_beta : 0.0;
  const double m00 = 1.0 - a, m01 = 1.0 - a, m10 = -ab, m11 = 1.0 - ab;
  const double level0 = (double)x[0], trend0 = (dbl && n > 1) ? (double)x[1] - (double)x[0] : 0.0;
  const uint32_t steps = n - 1u;                         // t = 1 .. n-1
  const uint32_t L = (steps + 31u) / 32u;
  const uint32_t t0 = 1u + (uint32_t)lane * L, t1 = min(n, t0 + L);
  // pass 1: zero-state response and M^len of my chunk
  FmAffine me = {1.0, 0.0, 0.0, 1.0, 0.0, 0.0};
  for (uint32_t t = t0; t < t1; ++t) {
    const double xt = (double)x[t];
    const double q00 = m00 * me.p00 + m01 * me.p10, q01 = m00 * me.p01 + m01 * me.p11;
    const double q10 = m10 * me.p00 + m11 * me.p10, q11 = m10 * me.p01 + m11 * me.p11;
    const double s0 = m00 * me.r0 + m01 * me.r1 + a * xt, s1 = m10 * me.r0 + m11 * me.r1 + ab * xt;
    me.p00 = q00; me.p01 = q01; me.p10 = q10; me.p11 = q11; me.r0 = s0; me.r1 = s1;
  }
  // inclusive scan of the affine maps over the lanes
  FmAffine inc = me;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    FmAffine up;
    up.p00 = __shfl_up_sync(FM_FULL, inc.p00, o); up.p01 = __shfl_up_sync(FM_FULL, inc.p01, o);
    up.p10 = __shfl_up_sync(FM_FULL, inc.p10, o); up.p11 = __shfl_up_sync(FM_FULL, inc.p11, o);
    up.r0 = __shfl_up_sync(FM_FULL, inc.r0, o); up.r1 = __shfl_up_sync(FM_FULL, inc.r1, o);
    if (lane >= o) inc = fm_affine_after(up, inc);
  }
  // state at the end of my chunk, and (from the lane before) at its start
  const double end_l = inc.p00 * level0 + inc.p01 * trend0 + inc.r0;
  const double end_t = inc.p10 * level0 + inc.p11 * trend0 + inc.r1;
  double level = __shfl_up_sync(FM_FULL, end_l, 1), trend = __shfl_up_sync(FM_FULL, end_t, 1);
  if (lane == 0) { level = level0; trend = trend0; }
  // pass 2: forecast errors of my chunk from its true start state
  double sse = 0.0;
  for (uint32_t t = t0; t < t1; ++t) {
    const double f = level + trend;
    const double e = (double)x[t] - f;
    sse = fma(e, e, sse);
    const double nl = f + a * e;
    trend = trend + (dbl ? g.smooth_beta : 0.0) * (nl - level - trend);
    level = nl;
  }
  sse = fm_warp_sum(sse);
  mean = __shfl_sync(FM_FULL, end_l, 31) + __shfl_sync(FM_FULL, end_t, 31);
  sigma = (steps > g.ddof) ? sqrt(sse / (double)(steps - g.ddof)) : 0.0;
}

// Additive Holt-Winters (design.md:60-66), A.2c: season of m points, level / trend / seasonal
// factors smooth_alpha / smooth_beta / smooth_gamma.  Start: level = mean of the first season,
// trend = (mean of the second season - level) / m, season_i = x_i - level.  From t = m on:
//   f = level + trend + season[t mod m]  (one-step-ahead forecast; its errors give sigma),
//   level' = a (x_t - season) + (1-a)(level + trend), trend' = b (level' - level) + (1-b) trend,
//   season[t mod m]' = c (x_t - level') + (1-c) season.
// The recurrence rewrites a different seasonal slot at every step, so it is not a fixed linear
// map that a scan could compose: lane 0 runs it, the seasonal state lives in `season` (m doubles
// of global scratch per warp), the two start sums and everything after the recurrence use the
// whole warp.  Centre of current point i (0-based): level + (i+1) trend + season[(n+i) mod m];
// `c1` is the centre of point 0, `hw` carries what fm_threshold_hw needs for the others.
struct FmHw { double trend; const double* season; uint32_t n, m; };

__device__ __forceinline__ bool fm_model_hw(const fm_global_params& g, const float* x, uint32_t n, int lane,
                                            double* season, double& c1, double& sigma, FmHw& hw) {
  const uint32_t m = g.season_len;
  const uint32_t need = max(g.min_hist_points, g.ddof + 1u);
  c1 = 0.0; sigma = 0.0;
  hw.trend = 0.0; hw.season = season; hw.n = n; hw.m = m;
  if (x == nullptr || season == nullptr || m < 2u || n < 2u * m || n < need) return false;
  double sa = 0.0, sb = 0.0;
  for (uint32_t i = lane; i < m; i += 32) { sa += (double)x[i]; sb += (double)x[m + i]; }
  sa = fm_warp_sum(sa); sb = fm_warp_sum(sb);
  double level = sa / (double)m;
  double trend = (sb / (double)m - level) / (double)m;
  for (uint32_t i = lane; i < m; i += 32) season[i] = (double)x[i] - level;
  __syncwarp();
  double sse = 0.0;
  if (lane == 0) {
    const double a = g.smooth_alpha, b = g.smooth_beta, c = g.smooth_gamma;
    uint32_t slot = 0;                                  // t mod m, t starts at m
    for (uint32_t t = m; t < n; ++t) {
      const double xt = (double)x[t], si = season[slot];
      const double e = xt - (level + trend + si);
      sse = fma(e, e, sse);
      const double nl = a * (xt - si) + (1.0 - a) * (level + trend);
      trend = b * (nl - level) + (1.0 - b) * trend;
      season[slot] = c * (xt - nl) + (1.0 - c) * si;
      level = nl;
      if (++slot == m) slot = 0;
    }
  }
  __syncwarp();
  level = __shfl_sync(FM_FULL, level, 0); trend = __shfl_sync(FM_FULL, trend, 0); sse = __shfl_sync(FM_FULL, sse, 0);
  const uint32_t cnt = n - m;
  hw.trend = trend;
  c1 = level + trend + season[n % m];
  sigma = cnt > g.ddof ? sqrt(sse / (double)(cnt - g.ddof)) : 0.0;
  return true;
}

// a2: model window -> (mean, sigma).  `x` global or shared; `vec_ok` = row is a
// 16-byte aligned global row.  Returns false when the window is too short.
__device__ __forceinline__ bool fm_model(const fm_global_params& g, const float* x, uint32_t n,
                                         bool vec_ok, int lane, double& mean, double& sigma) {
  uint32_t need = max(g.min_hist_points, g.ddof + 1u);
  mean = 0.0; sigma = 0.0;
  if (x == nullptr || g.ma_mode == FM_MODEL_HOLT_WINTERS) return false;   // Holt-Winters: fm_model_hw, general kernel
  if (g.ma_mode == FM_MA_WINDOW) {
    const uint32_t w = g.ma_window;
    need = max(need, w);
    if (n < need || w == 0) return false;
    fm_moments_any(x + (n - w), w, g.ddof, lane, mean, sigma);
    return true;
  }
  if (n < need) return false;
  if (g.ma_mode == FM_MODEL_EWMA || g.ma_mode == FM_MODEL_DOUBLE_EXP) {
    fm_model_smoothing(g, x, n, lane, mean, sigma);
    return true;
  }
  if (vec_ok) fm_moments_global(x, n, g.ddof, lane, mean, sigma);
  else fm_moments_any(x, n, g.ddof, lane, mean, sigma);
  return true;
}

// ------------------------------------------------------- sort and search ---
// Warp-level bitonic sort of n2 (power of two >= 64) keys in shared memory.
template <typename T>
__device__ __forceinline__ void fm_warp_bitonic(T* s, uint32_t n2, int lane) {
  const uint32_t half = n2 >> 1;
  for (uint32_t k = 2; k <= n2; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t i = lane; i < half; i += 32) {
        const uint32_t a = ((i & ~(j - 1)) << 1) | (i & (j - 1));
        const uint32_t b = a | j;
        const bool up = (a & k) == 0;
        const T x = s[a], y = s[b];
        if ((x > y) == up) { s[a] = y; s[b] = x; }
      }
      __syncwarp();
    }
  }
}

template <typename T>
__device__ __forceinline__ uint32_t fm_lower_bound(const T* a, uint32_t n, T x) {
  uint32_t lo = 0, len = n;
  while (len) {
    const uint32_t h = len >> 1;
    if (a[lo + h] < x) { lo += h + 1; len -= h + 1; } else len = h;
  }
  return lo;
}
template <typename T>
__device__ __forceinline__ uint32_t fm_upper_bound(const T* a, uint32_t n, T x) {
  uint32_t lo = 0, len = n;
  while (len) {
    const uint32_t h = len >> 1;
    if (!(x < a[lo + h])) { lo += h + 1; len -= h + 1; } else len = h;
  }
  return lo;
}

// A.3: sum over tie runs of (t^3 - t) of a sorted array (per-lane partial).
template <typename T>
__device__ __forceinline__ long long fm_tie_term_sorted(const T* a, uint32_t lo, uint32_t n,
                                                        int lane) {
  long long t = 0;
  for (uint32_t p = lo + lane; p < n; p += 32) {
    const T v = a[p];
    const bool start = (p == lo) || (a[p - 1] != v);
    if (start && p + 1 < n && a[p + 1] == v) t += fm_tie(fm_upper_bound(a, n, v) - p);
  }
  return t;
}

// ----------------------------------------------------------- p-values ------
struct FmTest { bool run; bool degenerate; double stat; double p; };

// A.4 Mann-Whitney (normal approximation, tie-corrected, continuity optional)
__device__ __forceinline__ FmTest fm_mann_whitney(long long r2_cur, long long T, uint32_t n1,
                                                  uint32_t n2, uint32_t cc) {
  FmTest t = {true, false, 0.0, 1.0};
  const double N = (double)n1 + (double)n2;
  const double u1 = 0.5 * (double)(r2_cur - (long long)n1 * ((long long)n1 + 1));
  const double u2 = (double)n1 * (double)n2 - u1;
  const double u = u1 > u2 ? u1 : u2;
  const double mu = 0.5 * (double)n1 * (double)n2;
  const double var = ((double)n1 * (double)n2 / 12.0) * ((N + 1.0) - (double)T / (N * (N - 1.0)));
  t.stat = u1;
  if (!(var > 0.0)) { t.degenerate = true; return t; }
  const double z = (u - mu - (cc ? 0.5 : 0.0)) / sqrt(var);
  const double p = erfc(z / 1.4142135623730951);
  t.p = p > 1.0 ? 1.0 : p;
  return t;
}

__device__ __forceinline__ FmTest fm_wilcoxon(long long w2p, long long w2m, long long T,
                                              uint32_t np, uint32_t cc) {
  FmTest t = {true, false, 0.0, 1.0};
  const double w = 0.5 * (double)(w2p < w2m ? w2p : w2m);
  const double n = (double)np;
  const double mu = n * (n + 1.0) / 4.0;
  const double var = n * (n + 1.0) * (2.0 * n + 1.0) / 24.0 - (double)T / 48.0;
  t.stat = w;
  if (!(var > 0.0)) { t.degenerate = true; return t; }
  double num = w - mu;
  if (cc && num != 0.0) num += 0.5;
  const double z = num / sqrt(var);
  const double p = erfc(fabs(z) / 1.4142135623730951);
  t.p = p > 1.0 ? 1.0 : p;
  return t;
}

// 1/j for the chi-square series (entry 0 unused)
static __constant__ double fm_inv_small[32] = {
    0.0, 1.0, 1.0 / 2, 1.0 / 3, 1.0 / 4, 1.0 / 5, 1.0 / 6, 1.0 / 7, 1.0 / 8, 1.0 / 9, 1.0 / 10, 1.0 / 11, 1.0 / 12, 1.0 / 13,
    1.0 / 14, 1.0 / 15, 1.0 / 16, 1.0 / 17, 1.0 / 18, 1.0 / 19, 1.0 / 20, 1.0 / 21, 1.0 / 22, 1.0 / 23, 1.0 / 24, 1.0 / 25,
    1.0 / 26, 1.0 / 27, 1.0 / 28, 1.0 / 29, 1.0 / 30, 1.0 / 31};
// A.5 chi-square survival, integer dof, closed form
__device__ __forceinline__ double fm_chi2_sf(double x, uint32_t dof) {
  if (!(x > 0.0) || dof == 0) return 1.0;
  const double hx = 0.5 * x;
  const double e = exp(-hx);
  const uint32_t m = dof >> 1;
  // (terms by a table of reciprocals: an fp64 division is ~30 instructions; dof <= FM_MAX_GROUPS - 1 = 15)
  if ((dof & 1u) == 0u) {
    double term = 1.0, sum = 1.0;
    for (uint32_t j = 1; j < m; ++j) { term *= hx * fm_inv_small[j & 31u]; sum += term; }
    return e * sum;
  }
  double q = erfc(sqrt(hx));
  if (m > 0) {
    double term = 1.0, sum = 1.0;
    for (uint32_t j = 2; j <= m; ++j) { term *= x * fm_inv_small[(2 * j - 1) & 31u]; sum += term; }
    q += e * sqrt(2.0 * x / 3.14159265358979323846) * sum;
  }
  return q;
}

// Kruskal-Wallis from H0 pieces: s = sum R_g^2 / n_g
__device__ __forceinline__ FmTest fm_kruskal(double s, double N, long long T, uint32_t K) {
  FmTest t = {true, false, 0.0, 1.0};
  const double h0 = 12.0 / (N * (N + 1.0)) * s - 3.0 * (N + 1.0);
  const double c = 1.0 - (double)T / (N * N * N - N);
  if (!(c > 0.0)) { t.degenerate = true; return t; }
  t.stat = h0 / c;
  t.p = fm_chi2_sf(t.stat, K - 1);
  return t;
}

// A.4b Friedman chi-square over K related samples of n points each (x[j * stride + i] = point i of
// group j, global or shared memory): block i = the i-th point of every group, lanes own blocks
// lane, lane + 32, ...; twice the midrank of a value inside its block and the tie term are exact
// integers, the statistic follows in fp64.  Whole warp; result uniform.
__device__ __forceinline__ FmTest fm_friedman(const float* x, uint64_t stride, uint32_t K, uint32_t n, int lane) {
  FmTest t = {true, false, 0.0, 1.0};
  long long ssq = 0, ties = 0;
  for (uint32_t j = 0; j < K; ++j) {
    long long r2 = 0;
    for (uint32_t i = lane; i < n; i += 32) {
      const float xj = x[j * stride + i];
      int lt = 0, eq = 0;
      for (uint32_t m = 0; m < K; ++m) {
        const float xm = x[m * stride + i];
        lt += xm < xj; eq += xm == xj;
      }
      r2 += 2 * lt + eq + 1;
      ties += (long long)eq * eq - 1;
    }
    r2 = fm_warp_sum(r2);
    ssq += r2 * r2;
  }
  ties = fm_warp_sum(ties);
  const double dk = (double)K, dn = (double)n;
  const double c = 1.0 - (double)ties / (dk * (dk * dk - 1.0) * dn);
  if (!(c > 0.0)) { t.degenerate = true; return t; }
  const double q0 = 12.0 / (dk * dn * (dk + 1.0)) * (0.25 * (double)ssq) - 3.0 * dn * (dk + 1.0);
  t.stat = q0 / c;
  t.p = fm_chi2_sf(t.stat, K - 1);
  return t;
}

__device__ __forceinline__ bool fm_mode_has(uint32_t mode, uint32_t test) {
  return mode == test || mode == FM_PW_ALL || mode == FM_PW_ANY;
}

// A.6
__device__ __forceinline__ bool fm_decide(uint32_t mode, double alpha, const FmTest* t, int n,
                                          bool& skipped) {
  int ne = 0; bool any = false, all = true;
  for (int i = 0; i < n; ++i) {
    if (!t[i].run || t[i].degenerate) continue;
    ++ne;
    const bool d = t[i].p < alpha;
    any |= d; all &= d;
  }
  skipped = (ne == 0);
  if (ne == 0) return false;
  return mode == FM_PW_ALL ? all : any;
}

// A.7 + A.8: bounds, anomalous points of the current row (shared or global),
// verdict.  Whole warp; fills the threshold part of `rec` (uniform).
__device__ __forceinline__ void fm_threshold(const fm_global_params& g, const fm_type_params& tp,
                                             bool differ, const float* cur, uint32_t n_cur,
                                             int lane, fm_series_rec& rec) {
  const double k = differ ? tp.k_low : tp.k;
  const double up = __dadd_rn(rec.mean, __dmul_rn(k, rec.sigma));
  double lo = __dsub_rn(rec.mean, __dmul_rn(k, rec.sigma));
  if (lo < tp.min_lower_bound) lo = tp.min_lower_bound;
  rec.upper = up; rec.lower = lo;
  uint32_t cnt = 0, first = 0xFFFFFFFFu;
  uint64_t mask = 0;
  for (uint32_t i0 = 0; i0 < n_cur; i0 += 32) {
    const uint32_t i = i0 + lane;
    bool a = false;
    if (i < n_cur) {
      const double x = (double)cur[i];
      a = ((tp.bound_mask & 1u) && x > up) || ((tp.bound_mask & 2u) && x < lo);
    }
    const uint32_t b = __ballot_sync(FM_FULL, a);
    if (b) {
      if (first == 0xFFFFFFFFu) first = i0 + (uint32_t)__ffs((int)b) - 1u;
      cnt += (uint32_t)__popc(b);
      if (i0 < 64) mask |= (uint64_t)b << i0;
    }
  }
  rec.n_anom = cnt;
  rec.first_idx = (uint16_t)(first == 0xFFFFFFFFu ? 0xFFFFu : (first < 0xFFFFu ? first : 0xFFFEu));
  rec.anom_mask = mask;
  rec.verdict = (cnt >= g.min_anomaly_points && cnt > 0) ? FM_UNHEALTHY : FM_HEALTHY;
}

// A.7 + A.8 for a seasonal model: current point i is compared after its own centre has been moved
// onto the first one's, x_i - (centre_i - centre_0) against centre_0 +- k sigma (rec.upper / lower
// are the bounds of point 0; point i's are those plus the same offset).
__device__ __forceinline__ void fm_threshold_hw(const fm_global_params& g, const fm_type_params& tp,
                                                bool differ, const float* cur, uint32_t n_cur, int lane,
                                                const FmHw& hw, fm_series_rec& rec) {
  const double k = differ ? tp.k_low : tp.k;
  const double up = __dadd_rn(rec.mean, __dmul_rn(k, rec.sigma));
  double lo = __dsub_rn(rec.mean, __dmul_rn(k, rec.sigma));
  if (lo < tp.min_lower_bound) lo = tp.min_lower_bound;
  rec.upper = up; rec.lower = lo;
  const double s0 = hw.season[hw.n % hw.m];
  uint32_t cnt = 0, first = 0xFFFFFFFFu;
  uint64_t mask = 0;
  for (uint32_t i0 = 0; i0 < n_cur; i0 += 32) {
    const uint32_t i = i0 + lane;
    bool a = false;
    if (i < n_cur) {
      const double off = __dadd_rn(__dmul_rn((double)i, hw.trend), __dsub_rn(hw.season[(hw.n + i) % hw.m], s0));
      const double x = __dsub_rn((double)cur[i], off);
      a = ((tp.bound_mask & 1u) && x > up) || ((tp.bound_mask & 2u) && x < lo);
    }
    const uint32_t b = __ballot_sync(FM_FULL, a);
    if (b) {
      if (first == 0xFFFFFFFFu) first = i0 + (uint32_t)__ffs((int)b) - 1u;
      cnt += (uint32_t)__popc(b);
      if (i0 < 64) mask |= (uint64_t)b << i0;
    }
  }
  rec.n_anom = cnt;
  rec.first_idx = (uint16_t)(first == 0xFFFFFFFFu ? 0xFFFFu : (first < 0xFFFFu ? first : 0xFFFEu));
  rec.anom_mask = mask;
  rec.verdict = (cnt >= g.min_anomaly_points && cnt > 0) ? FM_UNHEALTHY : FM_HEALTHY;
}

__device__ __forceinline__ void fm_rec_init(fm_series_rec& r) {
  r.mean = r.sigma = r.upper = r.lower = 0.0;
  r.u1 = 0.0; r.p_u = 1.0; r.w = 0.0; r.p_w = 1.0; r.h = 0.0; r.p_h = 1.0;
  r.n_anom = 0; r.first_idx = 0xFFFFu; r.verdict = FM_HEALTHY; r.status = 0; r.anom_mask = 0;
}

// One lane writes the record (full or compact).
__device__ __forceinline__ void fm_write_record(void* out, uint64_t s, uint32_t kind,
                                                const fm_series_rec& r) {
  if (kind == FM_REC_COMPACT) {
    fm_compact_rec c;
    c.anom_mask = r.anom_mask; c.n_anom = r.n_anom; c.verdict = r.verdict; c.status = r.status;
    c.first_idx = r.first_idx;
    reinterpret_cast<uint4*>(out)[s] = *reinterpret_cast<const uint4*>(&c);
  } else {
    reinterpret_cast<fm_series_rec*>(out)[s] = r;
  }
}

// Fused roll-up (a9): the per-API counters live in the output records.
__device__ __forceinline__ void fm_rollup_add(fm_rollup_rec* api, uint32_t id, uint32_t limit, uint32_t verdict) {
  if (id >= limit) return;          // an id outside the caller's table is dropped, never written
  atomicAdd(&api[id].n_children, 1u);
  if (verdict == FM_UNHEALTHY) atomicAdd(&api[id].n_unhealthy, 1u);
  if (verdict == FM_UNKNOWN) atomicAdd(&api[id].n_unknown, 1u);
}
