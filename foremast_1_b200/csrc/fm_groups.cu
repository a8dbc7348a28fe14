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
//
// Heavy ties first: the rollover metrics are error counts per scrape interval, a pooled sample of
// 180-480 points with a handful of distinct values.  Before sorting anything the warp builds the
// multiset of the sample in a 256-slot hash table in its parking buffer (election: every key stores
// its bits into its slot if it read it empty, whoever reads its own bits back owns the slot and
// counts itself with a fire-and-forget shared-memory add; the others try the next multiplier).  With
// D <= 64 distinct values, at most two per lane, "points below me" is a D-step loop over the (value, count)
// list, which gives 2 * midrank and t^3 - t per value, and every point fetches the midrank of its
// value from its slot: ~0.5k instructions where sort + park + two searches per point take 2.7k.
// A sample with more distinct values (or four failed rounds) takes the network: same integers.
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_device.cuh"
#include "fm_sortnet.cuh"

#ifndef FM_GROUPS_SYNC
#define FM_GROUPS_SYNC 2
#endif

namespace {
using namespace fmc;

constexpr int kGroups = 8;        // groups held in registers
constexpr int kR = 2 * kGroups;   // keys per lane
constexpr uint32_t kSyncEvery = FM_GROUPS_SYNC;   // services between two CTA barriers (power of two)
constexpr bool kSyncTwice = false;  // second barrier between the history stream and the rank stage
#ifndef FM_GROUPS_WARPS
#define FM_GROUPS_WARPS 20
#endif
#ifndef FM_GROUPS_EARLY
#define FM_GROUPS_EARLY 1
#endif
#ifndef FM_GROUPS_CTAS
#define FM_GROUPS_CTAS 1
#endif
constexpr int kCtas = FM_GROUPS_CTAS;     // CTAs per SM
constexpr int kWarps = FM_GROUPS_WARPS;   // one CTA per SM: all its warps walk the code together

__device__ __forceinline__ uint32_t row_len(const uint32_t* len, uint64_t s, uint32_t n) {
  return len ? min(__ldg(len + s), n) : n;
}

constexpr uint32_t kSlots = 256u;         // hash slots per warp: values at +0, counts / midranks at +4*kSlots
constexpr uint32_t kEmpty = 0xFFFFFFFFu;   // a NaN pattern: rows are NaN-free (foremast_cuda.h)
constexpr int kRounds = 4;
constexpr uint32_t kMaxDistinct = 64u;     // distinct values the multiset path ranks: two list entries per lane
static_assert(2u * kSlots * 4u <= 32u * kR * 4u, "the table lives in the warp's parking buffer");

__device__ __forceinline__ uint32_t slot_of(uint32_t tbl_s, float key, uint32_t mult) {
  return tbl_s + (((__float_as_uint(key) * mult) >> 24) << 2);
}

// 2 * (rank sum) of group k into lane k and the tie term, from the multiset of the pooled sample.
// s[2k], s[2k+1]: the points lane and lane + 32 of group k (-0 folded into +0), valid: bit r = s[r]
// is a point, otherwise a copy of some point of the group.  Destroys s.  False: more than 64 distinct values or unplaced keys; nothing returned.
__device__ __forceinline__ bool ranks_by_multiset(float (&s)[kR], const uint32_t valid, const uint32_t K, const int lane,
                                                  const uint32_t tbl_s, uint32_t& r2_mine, long long& T) {
  __syncwarp();
  {
    const uint32_t z = tbl_s + (uint32_t)lane * 32u;
    const float e = __uint_as_float(kEmpty);
    fm_sts_v4(z, e, e, e, e); fm_sts_v4(z + 16u, e, e, e, e);
    fm_sts_v4(z + 4u * kSlots, 0.f, 0.f, 0.f, 0.f); fm_sts_v4(z + 4u * kSlots + 16u, 0.f, 0.f, 0.f, 0.f);
  }
  __syncwarp();
  // round 0, straight line: the table is empty, so every key stores (a lane behind the end of its
  // group holds a copy of one of the group's points: its stores change nothing and it never counts)
  uint32_t pend = 0;
  {
    constexpr uint32_t mult = 0x9E3779B1u;
#pragma unroll
    for (int k = 0; k < kGroups; ++k) {
      if ((uint32_t)k < K) {                       // warp-uniform
        fm_sts_f32(slot_of(tbl_s, s[2 * k], mult), s[2 * k]);
        fm_sts_f32(slot_of(tbl_s, s[2 * k + 1], mult), s[2 * k + 1]);
      }
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kGroups; ++k) {
      if ((uint32_t)k < K) {
#pragma unroll
        for (int r = 2 * k; r < 2 * k + 2; ++r) {
          const uint32_t a = slot_of(tbl_s, s[r], mult);
          const bool won = fm_lds_u32(a) == __float_as_uint(s[r]);
          const bool real = (valid >> r) & 1u;
          if (won && real) fm_reds_add(a + 4u * kSlots, 1u);
          if (!won && real) pend |= 1u << r;
          s[r] = won ? __uint_as_float(a + 4u * kSlots) : s[r];
        }
      }
    }
    __syncwarp();
  }
#pragma unroll 1
  for (int round = 1; round < kRounds; ++round) {
    if (!__any_sync(FM_FULL, pend != 0u)) break;
    const uint32_t mult = 0x9E3779B1u + 0x3C6EF372u * (uint32_t)round;   // odd
#pragma unroll
    for (int k = 0; k < kGroups; ++k) {
      if ((uint32_t)k < K) {                       // warp-uniform
#pragma unroll
        for (int r = 2 * k; r < 2 * k + 2; ++r) {
          if (pend & (1u << r)) {
            const uint32_t a = slot_of(tbl_s, s[r], mult);
            if (fm_lds_u32(a) == kEmpty) fm_sts_f32(a, s[r]);
          }
        }
      }
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kGroups; ++k) {
      if ((uint32_t)k < K) {
#pragma unroll
        for (int r = 2 * k; r < 2 * k + 2; ++r) {
          if (pend & (1u << r)) {
            const uint32_t a = slot_of(tbl_s, s[r], mult);
            if (fm_lds_u32(a) == __float_as_uint(s[r])) {
              fm_reds_add(a + 4u * kSlots, 1u);
              s[r] = __uint_as_float(a + 4u * kSlots);   // from here on: where my value's midrank will be
              pend &= ~(1u << r);
            }
          }
        }
      }
    }
    __syncwarp();
  }
  if (__any_sync(FM_FULL, pend != 0u)) return false;

  // the occupied slots, 8 per lane, into a dense (value, count | slot) list
  const uint32_t mine_s = tbl_s + (uint32_t)lane * 32u;
  const float4 va = fm_lds_v4(mine_s), vb = fm_lds_v4(mine_s + 16u);
  const float4 ca = fm_lds_v4(mine_s + 4u * kSlots), cb = fm_lds_v4(mine_s + 4u * kSlots + 16u);
  const float vv[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
  const uint32_t cc[8] = {__float_as_uint(ca.x), __float_as_uint(ca.y), __float_as_uint(ca.z), __float_as_uint(ca.w),
                          __float_as_uint(cb.x), __float_as_uint(cb.y), __float_as_uint(cb.z), __float_as_uint(cb.w)};
  uint32_t occ = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) occ += cc[j] ? 1u : 0u;
  uint32_t incl = occ;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(FM_FULL, incl, o);
    if (lane >= o) incl += u;
  }
  const uint32_t D = __shfl_sync(FM_FULL, incl, 31);
  if (D > kMaxDistinct) return false;
  __syncwarp();                                    // every lane holds its slots: the value half becomes the list
  uint32_t at = tbl_s + (incl - occ) * 8u;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (cc[j]) {
      fm_sts_f32(at, vv[j]);
      fm_sts_u32(at + 4u, cc[j] | (((uint32_t)lane * 8u + (uint32_t)j) << 16));
      at += 8u;
    }
  }
  __syncwarp();
  // lane l owns list entries l and l + 32
  float dv = 0.f, dv2 = 0.f;
  uint32_t dc = 0, dslot = 0, dc2 = 0, dslot2 = 0;
  if ((uint32_t)lane < D) {
    dv = fm_lds_f32(tbl_s + (uint32_t)lane * 8u);
    const uint32_t w = fm_lds_u32(tbl_s + (uint32_t)lane * 8u + 4u);
    dc = w & 0xFFFFu; dslot = w >> 16;
  }
  if ((uint32_t)lane + 32u < D) {
    dv2 = fm_lds_f32(tbl_s + ((uint32_t)lane + 32u) * 8u);
    const uint32_t w = fm_lds_u32(tbl_s + ((uint32_t)lane + 32u) * 8u + 4u);
    dc2 = w & 0xFFFFu; dslot2 = w >> 16;
  }
  uint32_t less = 0, less2 = 0;
  if (D <= 32u) {                                  // warp-uniform
#pragma unroll 2
    for (uint32_t i = 0; i < D; ++i) {             // broadcast reads
      const float vi = fm_lds_f32(tbl_s + i * 8u);
      const uint32_t ci = fm_lds_u32(tbl_s + i * 8u + 4u) & 0xFFFFu;
      less += (vi < dv) ? ci : 0u;
    }
  } else {
#pragma unroll 2
    for (uint32_t i = 0; i < D; ++i) {
      const float vi = fm_lds_f32(tbl_s + i * 8u);
      const uint32_t ci = fm_lds_u32(tbl_s + i * 8u + 4u) & 0xFFFFu;
      less += (vi < dv) ? ci : 0u;
      less2 += (vi < dv2) ? ci : 0u;
    }
  }
  T = (long long)__reduce_add_sync(FM_FULL, dc * (dc * dc - 1u) + dc2 * (dc2 * dc2 - 1u));   // sum t^3 - t <= 512^3
  if ((uint32_t)lane < D) fm_sts_u32(tbl_s + 4u * kSlots + dslot * 4u, 2u * less + dc + 1u);
  if ((uint32_t)lane + 32u < D) fm_sts_u32(tbl_s + 4u * kSlots + dslot2 * 4u, 2u * less2 + dc2 + 1u);
  __syncwarp();
  r2_mine = 0;
#pragma unroll
  for (int k = 0; k < kGroups; ++k) {
    if ((uint32_t)k < K) {
      uint32_t m = 0;
      if (valid & (1u << (2 * k))) m = fm_lds_u32(__float_as_uint(s[2 * k]));
      if (valid & (2u << (2 * k))) m += fm_lds_u32(__float_as_uint(s[2 * k + 1]));
      m = __reduce_add_sync(FM_FULL, m);
      if (lane == k) r2_mine = m;
    }
  }
  __syncwarp();
  return true;
}

__global__ void __launch_bounds__(kWarps * 32, kCtas) k_groups_sort(const FmGArgs a, uint32_t* defer_list, uint32_t* defer_n,
                                                                const uint32_t sync_mask) {
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
  uint32_t ng0 = 0, ng1 = 0, nty = 0;              // group range and metric type of the service this warp takes next
  {
    const uint64_t v = (uint64_t)blockIdx.x * wpc + warp;
    if (v < a.n_services) {
      ng0 = a.group_offsets[v]; ng1 = a.group_offsets[v + 1];
      if (a.metric_type) nty = a.metric_type[v];
    }
  }
  for (uint64_t v0 = (uint64_t)blockIdx.x * wpc; v0 < a.n_services; v0 += (uint64_t)gridDim.x * wpc, ++round) {
    if ((round & sync_mask) == 0u) __syncthreads();
    const uint64_t v = v0 + warp;
    const uint32_t g0 = ng0, K = ng1 - ng0;
    bool active = v < a.n_services;
    const uint64_t vn = v + (uint64_t)gridDim.x * wpc;
    const bool more = vn < a.n_services;
    const uint32_t ty = nty;
    if (more) {                                    // used after the model: no wait
      ng0 = a.group_offsets[vn]; ng1 = a.group_offsets[vn + 1];
      if (a.metric_type) nty = a.metric_type[vn];
    }
    if (active && K > (uint32_t)kGroups) {
      if (lane == 0) defer_list[atomicAdd(defer_n, 1u)] = (uint32_t)v;
      active = false;
    }
    fm_type_params tp = a.P->types[0];
    double mean = 0.0, sigma = 0.0;
    bool model_ok = false;
    float s[kR];                                   // group k: points lane and lane + 32 in s[2k], s[2k+1]
    uint32_t N = 0, valid = 0, n_mine = 1;
    bool eligible = K >= 2;
    auto load_points = [&]() {
#pragma unroll
      for (int k = 0; k < kGroups; ++k) {
        s[2 * k] = CUDART_INF_F; s[2 * k + 1] = CUDART_INF_F;
        if ((uint32_t)k < K) {
          const uint32_t n = row_len(a.cur_len, g0 + k, a.n_cur);
          const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
          if (n) {                                  // behind the row's end: a copy of its last point
            s[2 * k] = __ldg(row + min((uint32_t)lane, n - 1u));
            s[2 * k + 1] = __ldg(row + min((uint32_t)lane + 32u, n - 1u));
          }
          if ((uint32_t)lane < n) valid |= 1u << (2 * k);
          if ((uint32_t)lane + 32u < n) valid |= 2u << (2 * k);
          N += n;
          if (lane == k) n_mine = n;
          if (n < g.min_kw || n == 0) eligible = false;
        }
      }
    };
    if (active) {
      tp = a.P->types[ty < a.P->n_types ? ty : 0u];

      // the rows of this warp's NEXT service into L2 while this one is computed: its first touches
      // would otherwise each wait for DRAM (a service is only ~7 KB)
      if (more && a.hist) {
        const char* hp = reinterpret_cast<const char*>(a.hist + vn * a.hist_pitch);
        for (uint32_t o = (uint32_t)lane * 128u; o < a.n_hist * 4u; o += 32u * 128u) fm_prefetch_l2(hp + o);
      }
      // this service's points: requested before the history stream, consumed after it
      if (FM_GROUPS_EARLY && want_kw) load_points();

      // ---- a2: model from the pooled history (streamed from global memory) --------------------
      if (a.hist) {
        const uint32_t nh = row_len(a.hist_len, v, a.n_hist);
        model_ok = fm_model(g, a.hist + v * a.hist_pitch, nh, true, lane, mean, sigma);
      }
      if (more) {
        const uint32_t row_bytes = a.n_cur * 4u;
        const uint32_t lines = (row_bytes + 127u) / 128u;           // per group row
        for (uint32_t j = lane; j < (ng1 - ng0) * lines; j += 32u) {
          const uint32_t k = j / lines, l = j - k * lines;
          fm_prefetch_l2(reinterpret_cast<const char*>(a.cur + (uint64_t)(ng0 + k) * a.cur_pitch) + l * 128u);
        }
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
      if (!FM_GROUPS_EARLY) load_points();
#pragma unroll
      for (int r = 0; r < kR; ++r) s[r] += 0.0f;  // -0 -> +0
      if (eligible) {
        uint32_t r2_mine = 0;                      // lane k: twice the rank sum of group k
        long long T = 0;
        if (!ranks_by_multiset(s, valid, K, lane, park_s, r2_mine, T)) {
#pragma unroll
          for (int k = 0; k < kGroups; ++k) {      // the keys again (L1)
            s[2 * k] = CUDART_INF_F; s[2 * k + 1] = CUDART_INF_F;
            if ((uint32_t)k < K) {
              const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
              if (valid & (1u << (2 * k))) s[2 * k] = __ldg(row + lane) + 0.0f;
              if (valid & (2u << (2 * k))) s[2 * k + 1] = __ldg(row + lane + 32) + 0.0f;
            }
          }
          sort_rows<kR, kR>(s, lane, park_s);
          T = (long long)tie_term_and_park<kR>(s, N, lane, park_s);
          T = fm_warp_sum(T);
#pragma unroll 1
          for (uint32_t k = 0; k < K; ++k) {
            const uint32_t n = row_len(a.cur_len, g0 + k, a.n_cur);
            const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
            uint32_t r2 = 0;
#pragma unroll 1
            for (uint32_t i = lane; i < n; i += 32) {
              const float q = __ldg(row + i) + 0.0f;
              r2 += parked_lower<kR>(park_s, N, q) + parked_upper<kR>(park_s, N, q) + 1u;
            }
            r2 = __reduce_add_sync(FM_FULL, r2);
            if ((uint32_t)lane == k) r2_mine = r2;
          }
          __syncwarp();
        }
        // sum R_k^2 / n_k: one division for all groups
        const double R = 0.5 * (double)r2_mine;
        const double sres = fm_warp_sum((uint32_t)lane < K ? R * R / (double)n_mine : 0.0);
        kw = fm_kruskal(sres, (double)N, T, K);
        status_kw |= FM_ST_KW_RUN;
        if (kw.degenerate) status_kw |= FM_ST_DEGENERATE;
        else differ = kw.p < g.alpha;
      }
      if (!(status_kw & FM_ST_KW_RUN) || kw.degenerate) status_kw |= FM_ST_PAIRWISE_SKIPPED;
      if (differ) status_kw |= FM_ST_DIFFER;
    }

    // ---- a3, a7, a8: bounds and verdict of every group, a9: service roll-up -------------------------
    const double kk = differ ? tp.k_low : tp.k;
    const double up = __dadd_rn(mean, __dmul_rn(kk, sigma));
    double lo = __dsub_rn(mean, __dmul_rn(kk, sigma));
    if (lo < tp.min_lower_bound) lo = tp.min_lower_bound;
    // fp32 points against fp64 bounds, exactly: x > up <=> x > rd(up), x < lo <=> x < ru(lo) (fm_ma.cu)
    const float up_f = (tp.bound_mask & 1u) ? __double2float_rd(up) : CUDART_INF_F;
    const float lo_f = (tp.bound_mask & 2u) ? __double2float_ru(lo) : -CUDART_INF_F;
    uint32_t ba_mine = 0, bb_mine = 0, nc_mine = 0;     // lane k: flags and length of group k
#pragma unroll 2
    for (uint32_t k = 0; k < K; ++k) {
      const uint32_t nc = row_len(a.cur_len, g0 + k, a.n_cur);
      const float* row = a.cur + (uint64_t)(g0 + k) * a.cur_pitch;
      const float xa = (uint32_t)lane < nc ? __ldg(row + lane) : 0.f;
      const float xb = (uint32_t)lane + 32u < nc ? __ldg(row + lane + 32) : 0.f;
      const bool fa = model_ok && (uint32_t)lane < nc && (xa > up_f || xa < lo_f);
      const bool fb = model_ok && (uint32_t)lane + 32u < nc && (xb > up_f || xb < lo_f);
      const uint32_t ba = __ballot_sync(FM_FULL, fa), bb = __ballot_sync(FM_FULL, fb);
      if ((uint32_t)lane == k) { ba_mine = ba; bb_mine = bb; nc_mine = nc; }
    }
    uint32_t verdict = 0xFFu;
    if ((uint32_t)lane < K) {                      // lane k writes the record of group k
      fm_series_rec rec;
      fm_rec_init(rec);
      rec.h = kw.stat; rec.p_h = kw.p; rec.mean = mean; rec.sigma = sigma;
      uint32_t status = status_kw;
      if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;
      if (nc_mine == 0) status |= FM_ST_NO_CURRENT;
      if (model_ok && nc_mine > 0) {
        const uint64_t mask = (uint64_t)ba_mine | ((uint64_t)bb_mine << 32);
        const uint32_t cnt = (uint32_t)__popc(ba_mine) + (uint32_t)__popc(bb_mine);
        rec.upper = up; rec.lower = lo;
        rec.n_anom = cnt;
        rec.first_idx = (uint16_t)(mask ? (uint32_t)__ffsll((long long)mask) - 1u : 0xFFFFu);
        rec.anom_mask = mask;
        rec.verdict = (cnt >= g.min_anomaly_points && cnt > 0) ? FM_UNHEALTHY : FM_HEALTHY;
      }
      if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT)) rec.verdict = FM_UNKNOWN;
      rec.status = (uint8_t)status;
      verdict = rec.verdict;
      a.out[g0 + lane] = rec;
    }
    if (a.svc) {                                    // one warp owns the service: one add per counter
      const uint32_t bad = __popc(__ballot_sync(FM_FULL, verdict == FM_UNHEALTHY));
      const uint32_t unk = __popc(__ballot_sync(FM_FULL, verdict == FM_UNKNOWN));
      if (lane == 0) {
        atomicAdd(&a.svc[v].n_children, K);
        if (bad) atomicAdd(&a.svc[v].n_unhealthy, bad);
        if (unk) atomicAdd(&a.svc[v].n_unknown, unk);
      }
    }
    __syncwarp();
  }
}

}  // namespace

void fm_groups_sort_shape(int* warps_per_cta, int* ctas_per_sm) { *warps_per_cta = kWarps; *ctas_per_sm = kCtas; }

int fm_launch_groups_sort(const FmGArgs& a, uint32_t* defer_list, uint32_t* defer_n, int grid, cudaStream_t st) {
  constexpr int smem = kWarps * 32 * kR * 4;
  if (smem > 48 * 1024 &&
      cudaFuncSetAttribute(k_groups_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
    return 0;   // the caller's cudaGetLastError() reports it
  // services between two CTA barriers (power of two); tuning switch, read once
  static const uint32_t sync_every = [] {
    const char* e = getenv("FOREMAST_CUDA_GROUPS_SYNC");
    const long v = e ? atol(e) : (long)kSyncEvery;
    return (v >= 1 && v <= 1024 && (v & (v - 1)) == 0) ? (uint32_t)v : kSyncEvery;
  }();
  k_groups_sort<<<grid, kWarps * 32, smem, st>>>(a, defer_list, defer_n, sync_every - 1u);
  return 1;
}
