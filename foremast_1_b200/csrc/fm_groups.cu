// fm_groups.cu — rollover strategy (BASELINE.json configs[3]): Kruskal-Wallis over the pod groups of a
// service plus the bound check of every group against the service's pooled history, for the
// common shape: at most 8 groups of at most 64 points.  One warp per service, persistent grid.
//
// Group k is loaded into registers 2k and 2k+1 of every lane (points lane and lane + 32): no index
// arithmetic, no pooled copy in shared memory.  The pooled sample (<= 512 keys, 16 per lane) is sorted by the register network of
// fm_sortnet.cuh (odd-even merge sort per lane, sign-folded shuffle stages across lanes), the tie
// term comes from the run scan, the sorted sample is parked in 2 KB of shared memory and every
// point finds its midrank there with two branch-free searches (10 probes each):
// 2 R_k = sum over the group's points of (lower + upper + 1).
// Services with more than 8 groups are appended to a list that k_score_groups (the general
// shared-memory kernel, fm_kernels.cu) works off afterwards.
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_device.cuh"
#include "fm_sortnet.cuh"

#ifndef FM_GROUPS_SYNC
#define FM_GROUPS_SYNC 1
#endif

namespace {
using namespace fmc;

constexpr int kGroups = 8;        // groups held in registers
constexpr int kR = 2 * kGroups;   // keys per lane
constexpr uint32_t kSyncEvery = FM_GROUPS_SYNC;   // services between two CTA barriers (power of two)
constexpr bool kSyncTwice = false;  // second barrier between the history stream and the rank stage
constexpr int kWarps = 32;        // one CTA per SM: all its warps walk the code together

__device__ __forceinline__ uint32_t row_len(const uint32_t* len, uint64_t s, uint32_t n) {
  return len ? min(__ldg(len + s), n) : n;
}

__global__ void __launch_bounds__(kWarps * 32, 1) k_groups_sort(const FmGArgs a, uint32_t* defer_list, uint32_t* defer_n) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int wpc = blockDim.x >> 5;
  const uint32_t park_s = fm_smem_u32(smem) + (uint32_t)warp * (32u * kR * 4u);
  const fm_global_params& g = a.P->g;
  const bool want_kw = g.pairwise_mode != FM_PW_NONE && fm_mode_has(g.pairwise_mode, FM_PW_KRUSKAL);

  // The loop is CTA-uniform with one barrier per service and the per-group parts are rolled
  // loops (their values are re-read through L1): 24 warps per SM at unrelated places of a long
  // unrolled body starve on instruction fetch (measured: 17 warps stalled on it per issue).
  uint32_t round = 0;
  for (uint64_t v0 = (uint64_t)blockIdx.x * wpc; v0 < a.n_services; v0 += (uint64_t)gridDim.x * wpc, ++round) {
    if ((round & (kSyncEvery - 1u)) == 0u) __syncthreads();
    const uint64_t v = v0 + warp;
    uint32_t g0 = 0, K = 0;
    bool active = v < a.n_services;
    if (active) {
      g0 = a.group_offsets[v];
      K = a.group_offsets[v + 1] - g0;
      if (K > (uint32_t)kGroups) {
        if (lane == 0) defer_list[atomicAdd(defer_n, 1u)] = (uint32_t)v;
        active = false;
      }
    }
    fm_type_params tp = a.P->types[0];
    double mean = 0.0, sigma = 0.0;
    bool model_ok = false;
    if (active) {
      const uint32_t ty = a.metric_type ? a.metric_type[v] : 0u;
      tp = a.P->types[ty < a.P->n_types ? ty : 0u];

      // the rows of this warp's NEXT service into L2 while this one is computed: its first touches
      // would otherwise each wait for DRAM (a service is only ~7 KB)
      const uint64_t vn = v + (uint64_t)gridDim.x * wpc;
      if (vn < a.n_services) {
        if (a.hist) {
          const char* hp = reinterpret_cast<const char*>(a.hist + vn * a.hist_pitch);
          for (uint32_t o = (uint32_t)lane * 128u; o < a.n_hist * 4u; o += 32u * 128u) fm_prefetch_l2(hp + o);
        }
        const uint32_t gn0 = a.group_offsets[vn], gn1 = a.group_offsets[vn + 1];
        const uint32_t row_bytes = a.n_cur * 4u;
        const uint32_t lines = (row_bytes + 127u) / 128u;           // per group row
        for (uint32_t j = lane; j < (gn1 - gn0) * lines; j += 32u) {
          const uint32_t k = j / lines, l = j - k * lines;
          fm_prefetch_l2(reinterpret_cast<const char*>(a.cur + (uint64_t)(gn0 + k) * a.cur_pitch) + l * 128u);
        }
      }

      // ---- a2: model from the pooled history (streamed from global memory) --------------------
      if (a.hist) {
        const uint32_t nh = row_len(a.hist_len, v, a.n_hist);
        model_ok = fm_model(g, a.hist + v * a.hist_pitch, nh, true, lane, mean, sigma);
      }
    }
    if (kSyncTwice) __syncthreads();
    if (!active) continue;

    // ---- a4/a5: Kruskal-Wallis over the pooled sample ---------------------------------------------
    FmTest kw = {false, false, 0.0, 1.0};
    uint32_t status_kw = 0;
    bool differ = false;
    if (g.pairwise_mode == FM_PW_FRIEDMAN) {
      uint32_t n = K ? row_len(a.cur_len, g0, a.n_cur) : 0u;
      bool eligible = K >= 3 && n >= g.min_kw && n > 0;
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
    if (want_kw) {
      // groups into registers (the sort does not care which key sits where)
      float s[kR];
      uint32_t N = 0;
      bool eligible = K >= 2;
#pragma unroll
      for (int k = 0; k < kGroups; ++k) {
        s[2 * k] = CUDART_INF_F; s[2 * k + 1] = CUDART_INF_F;
        if ((uint32_t)k < K) {
          const uint32_t n = row_len(a.cur_len, g0 + k, a.n_cur);
          const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
          if ((uint32_t)lane < n) s[2 * k] = __ldg(row + lane) + 0.0f;            // -0 -> +0
          if ((uint32_t)lane + 32u < n) s[2 * k + 1] = __ldg(row + lane + 32) + 0.0f;
          N += n;
          if (n < g.min_kw || n == 0) eligible = false;
        }
      }
      if (eligible) {
        sort_rows<kR, kR>(s, lane, park_s);
        long long T = (long long)tie_term_and_park<kR>(s, N, lane, park_s);
        T = fm_warp_sum(T);
        double sres = 0.0;
#pragma unroll 1
        for (uint32_t k = 0; k < K; ++k) {
          const uint32_t n = row_len(a.cur_len, g0 + k, a.n_cur);
          const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
          long long r2 = 0;
#pragma unroll 1
          for (uint32_t i = lane; i < n; i += 32) {
            const float q = __ldg(row + i) + 0.0f;
            r2 += (long long)(parked_lower<kR>(park_s, N, q) + parked_upper<kR>(park_s, N, q) + 1u);
          }
          r2 = fm_warp_sum(r2);
          const double R = 0.5 * (double)r2;
          sres += R * R / (double)n;
        }
        kw = fm_kruskal(sres, (double)N, T, K);
        status_kw |= FM_ST_KW_RUN;
        if (kw.degenerate) status_kw |= FM_ST_DEGENERATE;
        else differ = kw.p < g.alpha;
        __syncwarp();
      }
      if (!(status_kw & FM_ST_KW_RUN) || kw.degenerate) status_kw |= FM_ST_PAIRWISE_SKIPPED;
      if (differ) status_kw |= FM_ST_DIFFER;
    }

    // ---- a3, a7, a8: bounds and verdict of every group, a9: service roll-up -------------------------
    const double kk = differ ? tp.k_low : tp.k;
    const double up = __dadd_rn(mean, __dmul_rn(kk, sigma));
    double lo = __dsub_rn(mean, __dmul_rn(kk, sigma));
    if (lo < tp.min_lower_bound) lo = tp.min_lower_bound;
    const bool chk_up = (tp.bound_mask & 1u) != 0, chk_lo = (tp.bound_mask & 2u) != 0;
#pragma unroll 1
    for (uint32_t k = 0; k < K; ++k) {
      fm_series_rec rec;
      fm_rec_init(rec);
      rec.h = kw.stat; rec.p_h = kw.p; rec.mean = mean; rec.sigma = sigma;
      uint32_t status = status_kw;
      const uint32_t nc = row_len(a.cur_len, g0 + k, a.n_cur);
      const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
      if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
      if (nc == 0) status |= FM_ST_NO_CURRENT;
      if (model_ok && nc > 0) {
        const double xa = (uint32_t)lane < nc ? (double)__ldg(row + lane) : 0.0;
        const double xb = (uint32_t)lane + 32u < nc ? (double)__ldg(row + lane + 32) : 0.0;
        const bool fa = (uint32_t)lane < nc && ((chk_up && xa > up) || (chk_lo && xa < lo));
        const bool fb = (uint32_t)lane + 32u < nc && ((chk_up && xb > up) || (chk_lo && xb < lo));
        const uint32_t ba = __ballot_sync(FM_FULL, fa), bb = __ballot_sync(FM_FULL, fb);
        const uint64_t mask = (uint64_t)ba | ((uint64_t)bb << 32);
        const uint32_t cnt = (uint32_t)__popc(ba) + (uint32_t)__popc(bb);
        rec.upper = up; rec.lower = lo;
        rec.n_anom = cnt;
        rec.first_idx = (uint16_t)(mask ? (uint32_t)__ffsll((long long)mask) - 1u : 0xFFFFu);
        rec.anom_mask = mask;
        rec.verdict = (cnt >= g.min_anomaly_points && cnt > 0) ? FM_UNHEALTHY : FM_HEALTHY;
      }
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

}  // namespace

int fm_launch_groups_sort(const FmGArgs& a, uint32_t* defer_list, uint32_t* defer_n, int grid, cudaStream_t st) {
  constexpr int smem = kWarps * 32 * kR * 4;
  if (smem > 48 * 1024 &&
      cudaFuncSetAttribute(k_groups_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
    return 0;   // the caller's cudaGetLastError() reports it
  k_groups_sort<<<grid, kWarps * 32, smem, st>>>(a, defer_list, defer_n);
  return 1;
}
