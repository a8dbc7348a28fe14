// fm_canary_sort.cu — canary hot path, register-resident ranking (BASELINE.json configs[0..2]).
//
// One warp per series, persistent over the batch:
//   a1  the row pair (current <= 64 points, baseline <= 32*R points) arrives in the warp's shared
//       memory stage through the TMA unit (cp.async.bulk + mbarrier); as soon as the warp has
//       pulled the rows into registers (conflict-free 128-bit LDS) it re-arms the same stage for
//       its NEXT series, so the copy runs under the whole ranking of the current one.
//   a2  pivoted fp64 moments from the registers, warp-shuffle reduction.
//   a4  the baseline row is sorted IN REGISTERS by a network over the warp (fm_sortnet.cuh: odd-even
//       merge sort per lane, bitonic merge phases with sign-folded shuffles or through the parking
//       buffer; fm_sort48.cuh: the padding-free 1536-key network for 1025..1536 points).  Tie term
//       from the run lengths (lane-local scan + a 5-step segmented warp scan for runs that cross
//       lanes).  The sorted row is then parked in shared memory (conflict-free) and each current
//       point finds its lower / upper bound there: 2*R1 = n1(n1+1) + sum(2*lb + (ub - lb)).
//       Wilcoxon: 64 fp64 |differences| sorted across the warp, tie runs by warp scans.
//   a5..a9  one erfc for the three p-values (lanes 0..2), decision, thresholds, anomaly ballot,
//       fused roll-up atomics.
// All the work is regular (no data-dependent loops), so the cost per series does not depend on
// how many ties the data has.  Integers (rank sums, tie terms) are bit-identical to the
// sort-in-shared-memory path of fm_kernels.cu and to the oracle.
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_canary_common.cuh"
#include "fm_sortnet.cuh"
#include "fm_sort48.cuh"

namespace {
using namespace fmc;

template <int R, int RV>
__global__ void __launch_bounds__(256, 2) k_canary_sort(const FmKArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int wpc = blockDim.x >> 5;
  unsigned char* wp = smem + (size_t)warp * a.warp_bytes;
  uint64_t* bar = reinterpret_cast<uint64_t*>(wp);
  float* stage = reinterpret_cast<float*>(wp + 16);
  const uint32_t stage_s = fm_smem_u32(stage), base_s = stage_s + kCur * 4u;
  const uint32_t park_s = stage_s + (kCur + a.base_cap) * 4u;
  const fm_global_params& g = a.P->g;

  if (lane == 0) { fm_mbar_init(bar, 1); fm_fence_mbar_init(); }
  __syncwarp();

  const uint64_t stride = (uint64_t)gridDim.x * wpc;
  uint64_t s = (uint64_t)blockIdx.x * wpc + warp;
  uint32_t phase = 0;
  if (s < a.n_series && lane == 0)
    issue_rows(a, stage, bar, s, row_len(a.cur_len, s, a.n_cur), row_len(a.base_len, s, a.n_base));

  for (; s < a.n_series; s += stride) {
    const uint32_t nc = row_len(a.cur_len, s, a.n_cur);
    const uint32_t nb = row_len(a.base_len, s, a.n_base);
    const uint64_t sn = s + stride;
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

    // ---- rows into registers (everything that needs time order is read here) --------------
    const float craw_a = (uint32_t)lane < nc ? fm_lds_f32(stage_s + 4u * lane) : 0.0f;
    const float craw_b = (uint32_t)lane + 32u < nc ? fm_lds_f32(stage_s + 4u * (lane + 32u)) : 0.0f;
    const float ca = (uint32_t)lane < nc ? craw_a + 0.0f : CUDART_INF_F;
    const float cb = (uint32_t)lane + 32u < nc ? craw_b + 0.0f : CUDART_INF_F;

    // ---- Wilcoxon signed rank (A.4): pairs need the baseline in time order; done before the
    // baseline row occupies the registers
    Prep t_wx = {false, false, 0.0, 0.0};
    if (want_wx) t_wx = wilcoxon_stat(g, wilcoxon_load(g, stage_s, nc, nb, lane), lane, ax);

    // baseline: vector c of lane = float4 #(lane + 32 c); any assignment of points to lanes will do
    // for a sort, this one is bank-conflict free.  Missing points are +inf and sort to the end.
    constexpr bool k48 = (R == 64 && RV == 48);      // 1536-key network without the padding (fm_sort48.cuh)
    constexpr int NK = k48 ? 48 : R;
    float v[NK];
    const double pivot = nb ? (double)fm_lds_f32(base_s) : 0.0;
    FmMom mom = {0.0, 0.0};
    if (ranks || fuse_model) {
#pragma unroll
      for (int c = RV / 4; c < NK / 4; ++c)
        v[4 * c + 0] = v[4 * c + 1] = v[4 * c + 2] = v[4 * c + 3] = CUDART_INF_F;
#pragma unroll
      for (int c = 0; c < RV / 4; ++c) {
        const uint32_t e0 = ((uint32_t)lane + 32u * c) << 2;
        float4 q = make_float4(CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F);
        if (e0 < nb) q = fm_lds_v4(base_s + (e0 << 2));
        const bool full = 128u * c + 127u < nb;          // warp-uniform: the whole vector row is valid
        if (!full) {
          if (e0 + 1u >= nb) q.y = CUDART_INF_F;
          if (e0 + 2u >= nb) q.z = CUDART_INF_F;
          if (e0 + 3u >= nb) q.w = CUDART_INF_F;
        }
        v[4 * c + 0] = q.x + 0.0f; v[4 * c + 1] = q.y + 0.0f;
        v[4 * c + 2] = q.z + 0.0f; v[4 * c + 3] = q.w + 0.0f;
        if (fuse_model) {
          if (full) {
            fm_mom_add(mom, q.x, pivot); fm_mom_add(mom, q.y, pivot);
            fm_mom_add(mom, q.z, pivot); fm_mom_add(mom, q.w, pivot);
          } else {
            if (e0 < nb) fm_mom_add(mom, q.x, pivot);
            if (e0 + 1u < nb) fm_mom_add(mom, q.y, pivot);
            if (e0 + 2u < nb) fm_mom_add(mom, q.z, pivot);
            if (e0 + 3u < nb) fm_mom_add(mom, q.w, pivot);
          }
        }
      }
    }

    // ---- a2: model ---------------------------------------------------------------------------
    bool model_ok;
    if (a.hist) {
      model_ok = fm_model(g, a.hist + s * a.hist_pitch, row_len(a.hist_len, s, a.n_hist), true, lane,
                          rec.mean, rec.sigma);
    } else if (fuse_model) {
      model_ok = nb >= max(g.min_hist_points, g.ddof + 1u);
      if (model_ok) fm_moments_finish(mom, pivot, nb, g.ddof, rec.mean, rec.sigma);
    } else {
      model_ok = fm_model(g, stage + kCur, nb, false, lane, rec.mean, rec.sigma);
    }
    if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;

    // the stage has been consumed: let the TMA unit fill it with this warp's next series
    __syncwarp();
    if (sn < a.n_series) {
      fm_fence_proxy_async();
      if (lane == 0)
        issue_rows(a, stage, bar, sn, row_len(a.cur_len, sn, a.n_cur), row_len(a.base_len, sn, a.n_base));
    }

    // ---- Mann-Whitney / Kruskal: sort the baseline in registers -------------------------------
    Prep t_mw = {false, false, 0.0, 0.0}, t_kw = {false, false, 0.0, 0.0};
    if (ranks) {
      unsigned long long T;
      if constexpr (k48) {
        sort_rows48(v, lane, park_s);
        T = tie_term_and_park48(v, nb, lane, park_s);
      } else {
        sort_rows<R, RV>(v, lane, park_s);
        T = tie_term_and_park<R>(v, nb, lane, park_s);
      }
      auto lower = [&](float x) { if constexpr (k48) return parked48_bound<false>(park_s, nb, x); else return parked_lower<R>(park_s, nb, x); };
      auto upper = [&](float x) { if constexpr (k48) return parked48_bound<true>(park_s, nb, x); else return parked_upper<R>(park_s, nb, x); };

      // current row: sorted for its own tie positions, each point located in the parked baseline
      float sa = ca, sb = cb;
      sort64(sa, sb, lane);
      uint32_t lo_a, lo_b, e_a, e_b;
      runs64f(sa, sb, lane, lo_a, lo_b, e_a, e_b);
      long long r2c = 0;
      if ((uint32_t)lane < nc) {
        const uint32_t lb = lower(sa), tb = upper(sa) - lb;
        const unsigned long long c = tb + ((uint32_t)lane - lo_a);
        r2c += 2ll * lb + tb;
        T += 3ull * c * (c + 1ull);
      }
      if ((uint32_t)lane + 32u < nc) {
        const uint32_t lb = lower(sb), tb = upper(sb) - lb;
        const unsigned long long c = tb + ((uint32_t)lane + 32u - lo_b);
        r2c += 2ll * lb + tb;
        T += 3ull * c * (c + 1ull);
      }
      r2c = fm_warp_sum(r2c) + (long long)nc * (nc + 1);
      const long long Tt = fm_warp_sum((long long)T);
      ax.r2_cur = r2c; ax.t_mw = Tt;
      rank_tests(g, want_mw, want_kw, nc, nb, r2c, Tt, t_mw, t_kw);
    }
    finish_series(a, g, tp, s, lane, nc, pair, t_mw, t_wx, t_kw, model_ok, craw_a, craw_b, rec, ax, status);
    __syncwarp();
  }
}

}  // namespace

template <int R, int RV>
cudaError_t smem_attr(int bytes) {
  return cudaFuncSetAttribute(k_canary_sort<R, RV>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

cudaError_t fm_canary_sort_smem_attr(int bytes) {
  cudaError_t e = smem_attr<16, 12>(bytes);
  if (e == cudaSuccess) e = smem_attr<16, 16>(bytes);
  if (e == cudaSuccess) e = smem_attr<32, 24>(bytes);
  if (e == cudaSuccess) e = smem_attr<32, 32>(bytes);
  if (e == cudaSuccess) e = smem_attr<64, 48>(bytes);
  if (e == cudaSuccess) e = smem_attr<64, 64>(bytes);
  return e;
}

int fm_launch_canary_sort(const FmKArgs& a, int regs, int valid_regs, int warps_per_cta, int grid, cudaStream_t st) {
  const size_t smem = (size_t)warps_per_cta * a.warp_bytes;
  const dim3 block(warps_per_cta * 32);
  const bool part = valid_regs < regs;
  if (regs == 16) {
    if (part) k_canary_sort<16, 12><<<grid, block, smem, st>>>(a);
    else k_canary_sort<16, 16><<<grid, block, smem, st>>>(a);
  } else if (regs == 32) {
    if (part) k_canary_sort<32, 24><<<grid, block, smem, st>>>(a);
    else k_canary_sort<32, 32><<<grid, block, smem, st>>>(a);
  } else {
    if (part) k_canary_sort<64, 48><<<grid, block, smem, st>>>(a);
    else k_canary_sort<64, 64><<<grid, block, smem, st>>>(a);
  }
  return 1;
}
