// fm_ma.cu — bound check without a baseline (rollingUpdate / continuous / hpa strategies and the
// window-length sweep of BASELINE.json configs[4]): historical model -> mean, sigma -> bounds ->
// anomalous current points -> verdict -> roll-up.  Pure streaming, HBM-bound.
//
// A group of G lanes owns one series (G = 2, 4, 8 or 16 by history length), so a warp works on
// 32/G series at a time: a 60-point history is two 128-bit loads per lane, and the per-series
// bookkeeping (reduction, fp64 square root, record) is shared by 32/G series per instruction.
// Rows are read once with coalesced ld.global.nc.L1::no_allocate.v4 (consecutive series are
// consecutive rows, so a warp's groups read one contiguous span); the history loop is software
// pipelined with eight 16-byte loads per lane in flight, and the current row is requested first
// so that its latency hides under the moments.  No shared memory, no staging: nothing is reused.
// Measured (1M series, 16-byte records, B200): 58 % of the measured HBM copy rate at 60 points,
// 78 % at 240, 95 % at 480, >100 % from 960 on (profiles/r01/sweep_c5.md).
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_device.cuh"

namespace {

__device__ __forceinline__ uint32_t row_len(const uint32_t* len, uint64_t s, uint32_t n) {
  return len ? min(__ldg(len + s), n) : n;
}

template <int G>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FM_FULL, v, o);
  return v;
}
template <int G>
__device__ __forceinline__ uint32_t group_or(uint32_t v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v |= __shfl_xor_sync(FM_FULL, v, o);
  return v;
}

__device__ __forceinline__ void mom4(FmMom& m, const float4& q, double pivot) {
  fm_mom_add(m, q.x, pivot); fm_mom_add(m, q.y, pivot); fm_mom_add(m, q.z, pivot); fm_mom_add(m, q.w, pivot);
}

#ifndef FM_MA_CTAS
#define FM_MA_CTAS 4
#endif
// resident CTAs per SM (the register budget follows: 4 -> 64 registers)
constexpr int ma_ctas(int g) { return g <= 2 ? 3 : FM_MA_CTAS; }

template <int G>
__global__ void __launch_bounds__(256, ma_ctas(G)) k_ma_only(const FmKArgs a) {
  constexpr int SPW = 32 / G;                    // series per warp
  const int lane = threadIdx.x & 31;
  const int gl = lane & (G - 1), grp = lane / G;
  const fm_global_params& g = a.P->g;
  const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const uint64_t w0 = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const uint32_t need = max(g.min_hist_points, g.ddof + 1u);
  const uint32_t n4max = a.n_hist >> 2, c4max = (a.n_cur + 3u) >> 2;

  for (uint64_t s0 = w0 * SPW; s0 < a.n_series; s0 += warps * SPW) {   // warp-uniform loop
    const uint64_t s = s0 + grp;
    const bool live = s < a.n_series;
    const uint32_t nh = live ? row_len(a.hist_len, s, a.n_hist) : 0u;
    const uint32_t nc = live ? row_len(a.cur_len, s, a.n_cur) : 0u;
    const float* hx = fm_hist_row(a, live ? s : 0);
    const float4* h4 = reinterpret_cast<const float4*>(hx);

    // current row: requested before the history so that its latency hides under the moments
    const float4* c4 = reinterpret_cast<const float4*>(a.cur + (live ? s : 0) * a.cur_pitch);
    constexpr int CV = (16 + G - 1) / G;          // float4 per lane that cover 64 current points
    float4 cq[CV];
#pragma unroll
    for (int j = 0; j < CV; ++j) {
      const uint32_t v = gl + j * G;
      cq[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (v < c4max && 4u * v < nc) cq[j] = fm_ldg_stream(c4 + v);
    }

    // ---- a2: pivoted moments of the history row ---------------------------------------------
    // software pipelined: the next four 16-byte loads of a lane are in flight while the current
    // four are accumulated (8 loads = 128 B per lane outstanding)
    FmMom m = {0.0, 0.0};
    const uint32_t n4 = nh >> 2;
    float4 q[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      q[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gl + j * G < n4) q[j] = fm_ldg_stream(h4 + gl + j * G);
    }
    // pivot = the row's first sample: it is in the first vector of the group's first lane (rows
    // shorter than one vector fetch it on their own)
    const float p0 = __shfl_sync(FM_FULL, q[0].x, lane & ~(G - 1));
    const double pivot = nh >= 4u ? (double)p0 : (nh ? (double)__ldg(hx) : 0.0);
#ifndef FM_MA_SHORT
#define FM_MA_SHORT 1
#endif
#ifndef FM_MA_WORDS
#define FM_MA_WORDS 1
#endif
    if (FM_MA_SHORT && n4max <= 4u * G) {          // kernel-uniform: the whole row is in q already
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (gl + j * G < n4) mom4(m, q[j], pivot);
    } else {
      for (uint32_t i = gl; i < n4max; i += 4 * G) {
        float4 r[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          r[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          const uint32_t v = i + (4 + j) * G;
          if (v < n4) r[j] = fm_ldg_stream(h4 + v);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (i + j * G < n4) mom4(m, q[j], pivot);
          q[j] = r[j];
        }
      }
    }
    for (uint32_t t = (n4 << 2) + gl; t < nh; t += G)   // up to three points after the last full vector
      fm_mom_add(m, __ldg(hx + t), pivot);
    const double s1 = group_sum<G>(m.s1), s2 = group_sum<G>(m.s2);
    const bool model_ok = nh >= need;
    // mean = pivot + s1 / n and sigma = sqrt((s2 n - s1^2) / (n (n - ddof))): ONE division per
    // series, the even lanes of a group taking the first quotient and the odd lanes the second
    double mean = 0.0, sigma = 0.0;
    {
      const double dn = nh ? (double)nh : 1.0;
      const bool odd = (gl & 1) != 0;
      const double m2n = s2 * dn - s1 * s1;
      const double num = odd ? (m2n > 0.0 ? m2n : 0.0) : s1;
      const double den = odd ? dn * (double)(nh > g.ddof ? nh - g.ddof : 1u) : dn;
      const double q = num / den;
      const double r = sqrt(odd ? q : 0.0);
      const int base_lane = lane & ~(G - 1);
      const double q0 = __shfl_sync(FM_FULL, q, base_lane), r1 = __shfl_sync(FM_FULL, r, base_lane + 1);
      if (model_ok) {
        mean = pivot + q0;
        sigma = nh > g.ddof ? r1 : 0.0;
      }
    }

    // ---- a3, a7: bounds and anomalous current points ------------------------------------------
    const uint32_t ty = (live && a.metric_type) ? a.metric_type[s] : 0u;
    const fm_type_params tp = a.P->types[ty < a.P->n_types ? ty : 0u];
    const double up = __dadd_rn(mean, __dmul_rn(tp.k, sigma));
    double lo = __dsub_rn(mean, __dmul_rn(tp.k, sigma));
    if (lo < tp.min_lower_bound) lo = tp.min_lower_bound;
    uint32_t mlo = 0, mhi = 0;                   // anomaly mask of the (<= 64) current points
    if (model_ok) {
      // fp32 x against an fp64 bound, exactly, without converting every x: x > up <=> x > rd(up)
      // (rd(up) is the largest float <= up, so the next float above it exceeds up), and
      // x < lo <=> x < ru(lo); NaN bounds compare false either way.  A side that is not checked
      // gets an infinite bound; the points behind the row's end are masked once at the end: no
      // branch and no index test per point.
      const float up_f = (tp.bound_mask & 1u) ? __double2float_rd(up) : CUDART_INF_F;
      const float lo_f = (tp.bound_mask & 2u) ? __double2float_ru(lo) : -CUDART_INF_F;
      // lane gl's vector j covers points 4 (gl + j G) ..: the part 4 G j of the bit position is a
      // compile-time constant (and stays inside one 32-bit word for G <= 8), the part 4 gl is one
      // shift of each word at the end
      uint32_t w[2] = {0u, 0u};
      unsigned long long m64 = 0;
#pragma unroll
      for (int j = 0; j < CV; ++j) {
        const uint32_t nib = ((cq[j].x > up_f || cq[j].x < lo_f) ? 1u : 0u) | ((cq[j].y > up_f || cq[j].y < lo_f) ? 2u : 0u) |
                             ((cq[j].z > up_f || cq[j].z < lo_f) ? 4u : 0u) | ((cq[j].w > up_f || cq[j].w < lo_f) ? 8u : 0u);
        if constexpr (FM_MA_WORDS && G <= 8) {
          if (4 * G * j < 64) w[(4 * G * j) >> 5] |= nib << ((4 * G * j) & 31);
        } else {
          const uint32_t v = gl + j * G;             // float4 number: points 4v .. 4v+3
          if (v < 16u) m64 |= (unsigned long long)nib << (4u * v);
        }
      }
      if constexpr (FM_MA_WORDS && G <= 8) m64 = (unsigned long long)(w[0] << (4 * gl)) | ((unsigned long long)(w[1] << (4 * gl)) << 32);
      if (nc < 64u) m64 &= (1ull << nc) - 1ull;
      mlo = (uint32_t)m64; mhi = (uint32_t)(m64 >> 32);
    }
    mlo = group_or<G>(mlo); mhi = group_or<G>(mhi);

    // ---- a8: verdict, record -------------------------------------------------------------------
    uint32_t verdict_out = 0xFFu;
    if (live && gl == 0) {
      uint32_t status = 0;
      if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
      if (nc == 0) status |= FM_ST_NO_CURRENT;
      const uint64_t mask = (uint64_t)mlo | ((uint64_t)mhi << 32);
      const uint32_t cnt = (uint32_t)__popc(mlo) + (uint32_t)__popc(mhi);
      uint32_t verdict = (cnt >= g.min_anomaly_points && cnt > 0) ? FM_UNHEALTHY : FM_HEALTHY;
      if (status) verdict = FM_UNKNOWN;
      const uint16_t first = (uint16_t)(mask ? (uint32_t)__ffsll((long long)mask) - 1u : 0xFFFFu);
      if (a.record_kind == FM_REC_COMPACT) {
        fm_compact_rec c;
        c.anom_mask = mask; c.n_anom = cnt; c.verdict = (uint8_t)verdict; c.status = (uint8_t)status;
        c.first_idx = first;
        reinterpret_cast<uint4*>(a.out)[s] = *reinterpret_cast<const uint4*>(&c);
      } else {
        fm_series_rec r;
        fm_rec_init(r);
        r.mean = mean; r.sigma = sigma;
        if (model_ok && nc > 0) { r.upper = up; r.lower = lo; }
        r.n_anom = cnt; r.first_idx = first; r.verdict = (uint8_t)verdict; r.status = (uint8_t)status;
        r.anom_mask = mask;
        reinterpret_cast<fm_series_rec*>(a.out)[s] = r;
      }
      if (a.aux) a.aux[s] = fm_rank_aux{0, 0, 0, 0, 0, 0u, 0u};
      verdict_out = verdict;
    }
    // ---- a9: roll-up, one add per API and counter among the series of this warp (the series of an
    // API are neighbours in a batch: 32 / G series usually name one or two APIs)
    if (a.api && a.api_id) {                       // kernel-uniform
      if (gl == 0) {
        constexpr uint32_t leaders = G == 32 ? 1u : (G == 16 ? 0x00010001u : (G == 8 ? 0x01010101u : (G == 4 ? 0x11111111u : 0x55555555u)));
        const uint32_t id = live ? a.api_id[s] : 0xFFFFFFFFu;
        const uint32_t peers = __match_any_sync(leaders, id);
        const uint32_t bad = __ballot_sync(leaders, verdict_out == FM_UNHEALTHY) & peers;
        const uint32_t unk = __ballot_sync(leaders, verdict_out == FM_UNKNOWN) & peers;
        if (id < a.n_apis && (peers & ((1u << lane) - 1u)) == 0u) {   // first lane of its API; ids outside the table are dropped
          atomicAdd(&a.api[id].n_children, (uint32_t)__popc(peers));
          if (bad) atomicAdd(&a.api[id].n_unhealthy, (uint32_t)__popc(bad));
          if (unk) atomicAdd(&a.api[id].n_unknown, (uint32_t)__popc(unk));
        }
      }
    }
  }
}

}  // namespace

int fm_ma_only_ctas(int group) { return ma_ctas(group); }

int fm_launch_ma_only(const FmKArgs& a, int group, int grid, cudaStream_t st) {
  if (group == 2) k_ma_only<2><<<grid, 256, 0, st>>>(a);
  else if (group == 4) k_ma_only<4><<<grid, 256, 0, st>>>(a);
  else if (group == 8) k_ma_only<8><<<grid, 256, 0, st>>>(a);
  else if (group == 16) k_ma_only<16><<<grid, 256, 0, st>>>(a);
  else k_ma_only<32><<<grid, 256, 0, st>>>(a);
  return 1;
}
