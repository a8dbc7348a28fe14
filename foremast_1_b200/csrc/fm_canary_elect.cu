// fm_canary_elect.cu — canary hot path without sorting the baseline (BASELINE.json configs[0..2]):
// current row <= 64 points, baseline row <= 1984 points, ONE CTA OF FOUR WARPS PER SERIES
// (persistent grid).  Four warps share one staging buffer pair and one table, so an SM keeps
// 24-32 warps resident instead of the 7 a warp-per-series layout would get out of 227 KB.
//
// With x the current sample (n1 points, sorted in registers by warp 0) and y the baseline (n2):
//
//   2*R1 = n1(n1+1) + 2*n1*n2 - 2*P - E     P = sum_y #{x < y}   a 3-level 4-ary search of every
//                                               baseline point in the sorted current row (64 keys,
//                                               21 nodes of 3 keys in shared memory)
//                                            E = sum_x #{y == x}  from the multiset below
//   T    = sum over tie classes of t^3 - t   from the multiset below
//
// The multiset of all n1 + n2 values is built by ELECTION in a shared-memory table of 32-bit
// slots {members:16 | index:16} — no compare-and-swap, no probing loops:
//   pass A  every point stores its index at slot hash(value)            (plain STS, nothing waits)
//   pass B  every point reads the slot back: the index found is the class WINNER.  I am the
//           winner -> done.  The winner has my value -> I am a MEMBER: bump the slot's member
//           count; the count returned is my position p in the class and T += 3p(p+1)
//           (sum_{p<t} 3p(p+1) = t^3 - t).  Another value owns the slot -> my whole class lost
//           it (equal values hash equally) and I keep the point in my private overflow list.
//   The overflow lists (~18 % at load 0.37) are replayed against the same table with another
//   hash multiplier until they are empty; classes move together, so each is counted exactly once.
// Same-address atomics serialise, so a series whose current row is mostly ties first merges the
// equal values of a warp row with match.any (2 cycles per DISTINCT value) and bumps once per group.
//
// Every thread keeps its share of the baseline (NV float4 vectors) and their slots in registers
// from pass A to pass B.  The integers (2*R1, T) are bit-identical to the sorting paths and to
// the oracle (tests/test_gpu_parity.py runs every path).
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_canary_common.cuh"

namespace {
using namespace fmc;

constexpr uint32_t kThreads = 128;
constexpr uint32_t kTreeFloats = 88;   // 21 nodes x 4 floats, [84] = 64th sorted key
constexpr uint32_t kMaxRounds = 12;
constexpr uint32_t kDirect = 0x80000000u;

__device__ __forceinline__ uint32_t hash_mult(uint32_t round) {
  return (0x9E3779B1u + round * 0x7F4A7C15u) | 1u;        // a different odd multiplier per round
}
__device__ __forceinline__ uint32_t slot_of(float x, uint32_t mult, uint32_t shift) {
  return ((__float_as_uint(x) * mult) >> shift) << 2;     // byte offset into the table
}
__device__ __forceinline__ unsigned long long tie_gain(uint32_t c, uint32_t n) {
  const unsigned long long C = c, N = n;                  // sum_{p=c}^{c+n-1} 3p(p+1)
  return N * (3ull * C * (C + N) + N * N - 1ull);
}
__device__ __forceinline__ void cta_sync() { asm volatile("bar.sync 0;" ::: "memory"); }

// sorted index -> float offset in the 4-ary search tree
__device__ __forceinline__ uint32_t tree_slot(uint32_t i) {
  if ((i & 15u) == 15u) return i == 63u ? 84u : (i >> 4);                 // node 0
  if ((i & 3u) == 3u) return (1u + (i >> 4)) * 4u + ((i & 15u) >> 2);      // nodes 1..4
  return (5u + (i >> 2)) * 4u + (i & 3u);                                  // nodes 5..20
}
__device__ __forceinline__ uint32_t below3(const float4& n, float y) {     // #{keys of the node < y}
  return (n.x < y ? 1u : 0u) + (n.y < y ? 1u : 0u) + (n.z < y ? 1u : 0u);
}
// sum over four points of #{sorted current values < y}; the node loads of the four are in flight together
__device__ __forceinline__ uint32_t tree_lower4(uint32_t tree, const float4& n0, float s63, const float* y) {
  uint32_t m[4];
  float4 n[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) m[k] = below3(n0, y[k]);
#pragma unroll
  for (int k = 0; k < 4; ++k) n[k] = fm_lds_v4(tree + ((1u + m[k]) << 4));
#pragma unroll
  for (int k = 0; k < 4; ++k) m[k] = 4u * m[k] + below3(n[k], y[k]);
#pragma unroll
  for (int k = 0; k < 4; ++k) n[k] = fm_lds_v4(tree + ((5u + m[k]) << 4));
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t pos = 4u * m[k] + below3(n[k], y[k]);
    sum += pos + ((pos == 63u && s63 < y[k]) ? 1u : 0u);
  }
  return sum;
}

struct Elect {
  uint32_t tbl, stage, curslot, list;   // shared-space addresses; list = this thread's column
  bool merged;
  unsigned long long tsum;              // per-thread part of T
  uint32_t nov;                         // entries in this thread's overflow list
};

// A member of a class whose winner sits in `slot`: bump the count, account my tie position(s).
// In merged mode the members of one warp row with equal values are bumped once by their leader.
__device__ __forceinline__ void member(Elect& e, int lane, bool mem, float x, uint32_t slot) {
  if (!e.merged) {
    if (mem) {
      const uint32_t c = fm_atoms_add(slot, 0x10000u) >> 16;     // members counted before me
      e.tsum += 3ull * (c + 1u) * (c + 2u);
    }
  } else {
    const uint32_t mm = __ballot_sync(FM_FULL, mem);
    if (mem) {
      const uint32_t peers = __match_any_sync(mm, __float_as_uint(x));
      if (lane == __ffs((int)peers) - 1) {
        const uint32_t n = (uint32_t)__popc(peers);
        const uint32_t c = fm_atoms_add(slot, n << 16) >> 16;
        e.tsum += tie_gain(c + 1u, n);
      }
    }
  }
}
__device__ __forceinline__ void overflow(Elect& e, bool ovf, uint32_t idx) {
  if (ovf) { fm_sts_u16(e.list + e.nov * (kThreads * 2u), idx); ++e.nov; }
}

// Pass-B treatment of one point; w = the slot word read back.  Whole warps call (merged mode votes).
__device__ __forceinline__ void classify(Elect& e, int lane, bool ok, float x, uint32_t idx, uint32_t h, uint32_t w) {
  const uint32_t widx = w & 0xFFFFu;
  const bool need = ok && widx != idx;                     // somebody else won my slot
  float key = 0.0f;
  if (need) key = fm_lds_f32(e.stage + (widx << 2));
  const bool mem = need && key == x;                       // IEEE ==: -0 == +0
  const bool ovf = need && !mem;
  member(e, lane, mem, x, e.tbl + h);
  if (ok && !ovf && idx < kCur) fm_sts_u32(e.curslot + (idx << 2), e.tbl + h);   // a current point found its class
  overflow(e, ovf, idx);
}

// What warp 1 (Wilcoxon) and the four warps' partial sums hand to warp 0 through shared memory.
struct Handoff {
  double wx_stat, wx_arg;
  long long w2_plus, w2_minus, t_wx;
  uint32_t n_wx, wx_flags;            // bit 0 run, bit 1 degenerate
  double s1[4], s2[4];
  unsigned long long tsum[4];
  uint32_t psum[4];
  uint32_t merged, pad;
};

template <int NV>
__global__ void __launch_bounds__(kThreads) k_canary_elect(const FmKArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  // CTA layout: [2 mbarriers][stage 0][stage 1][tree][curslot 64 x u32][handoff x 2][lists][table]
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  const size_t stage_bytes = (size_t)(kCur + a.base_cap) * 4;
  float* stage_p[2] = {reinterpret_cast<float*>(smem + 16),
                       reinterpret_cast<float*>(smem + 16 + (a.n_stages == 2 ? stage_bytes : 0))};
  unsigned char* q = smem + 16 + a.n_stages * stage_bytes;
  const uint32_t tree_s = fm_smem_u32(q);           q += kTreeFloats * 4;
  const uint32_t curslot_s = fm_smem_u32(q);        q += kCur * 4;
  Handoff* hand = reinterpret_cast<Handoff*>(q);    q += 2 * sizeof(Handoff);
  const uint32_t list_s = fm_smem_u32(q) + tid * 2u; q += (size_t)(4 * NV + 2) * kThreads * 2;
  const uint32_t tbl_s = fm_smem_u32(q);
  const fm_global_params& g = a.P->g;
  const uint32_t shift = 32u - a.log_m;

  if (tid == 0) {
    for (uint32_t i = 0; i < a.n_stages; ++i) fm_mbar_init(&bar[i], 1);
    fm_fence_mbar_init();
  }
  __syncthreads();

  uint64_t s = blockIdx.x;
  uint32_t phase[2] = {0u, 0u};
  uint32_t st = 0, par = 0;
  if (s < a.n_series && tid == 0)
    issue_rows(a, stage_p[0], &bar[0], s, row_len(a.cur_len, s, a.n_cur), row_len(a.base_len, s, a.n_base));

  for (; s < a.n_series; s += gridDim.x, par ^= 1u) {
    const uint32_t nc = row_len(a.cur_len, s, a.n_cur);
    const uint32_t nb = row_len(a.base_len, s, a.n_base);
    const uint64_t sn = s + gridDim.x;
    // Everybody is past the previous series' last barrier, so the other stage is free again.
    if (a.n_stages == 2 && sn < a.n_series && tid == 0) {
      fm_fence_proxy_async();
      issue_rows(a, stage_p[st ^ 1u], &bar[st ^ 1u], sn, row_len(a.cur_len, sn, a.n_cur),
                 row_len(a.base_len, sn, a.n_base));
    }
    fm_mbar_wait(&bar[st], phase[st]);
    phase[st] ^= 1u;

    const float* stage = stage_p[st];
    const uint32_t stage_s = fm_smem_u32(stage), base_s = stage_s + kCur * 4u;
    Handoff& H = hand[par];

    const bool pair = g.pairwise_mode != FM_PW_NONE && nc > 0 && nb > 0;
    const bool want_mw = pair && fm_mode_has(g.pairwise_mode, FM_PW_MANN_WHITE) && nc >= g.min_mw && nb >= g.min_mw;
    const bool want_kw = pair && fm_mode_has(g.pairwise_mode, FM_PW_KRUSKAL) && nc >= g.min_kw && nb >= g.min_kw;
    const bool want_wx = pair && fm_mode_has(g.pairwise_mode, FM_PW_WILCOXON);
    const bool ranks = want_mw || want_kw;
    const bool fuse_model = (a.hist == nullptr) && g.ma_mode == FM_MA_ALL;
    const uint32_t mult0 = hash_mult(0);

    Elect e;
    e.tbl = tbl_s; e.stage = stage_s; e.curslot = curslot_s; e.list = list_s;
    e.merged = false; e.tsum = 0; e.nov = 0;

    // ---- phase 1 ---------------------------------------------------------------------------
    // warp 0: current row -> registers, sorted, search tree, ties inside the row, its pass A
    float craw_a = 0.f, craw_b = 0.f, ca = CUDART_INF_F, cb = CUDART_INF_F;
    uint32_t ec_sum = 0, hca = 0, hcb = 0;
    double mean_nf = 0.0, sigma_nf = 0.0;
    bool model_nf_ok = false;
    if (warp == 0) {
      if ((uint32_t)lane < nc) { craw_a = fm_lds_f32(stage_s + 4u * lane); ca = craw_a + 0.0f; }
      if ((uint32_t)lane + 32u < nc) { craw_b = fm_lds_f32(stage_s + 4u * (lane + 32u)); cb = craw_b + 0.0f; }
      if (ranks) {
        float sa = ca, sb = cb;
        sort64(sa, sb, lane);
        fm_sts_f32(tree_s + (tree_slot((uint32_t)lane) << 2), sa);
        fm_sts_f32(tree_s + (tree_slot((uint32_t)lane + 32u) << 2), sb);
        fm_sts_u32(curslot_s + 4u * lane, 0u);
        fm_sts_u32(curslot_s + 4u * (lane + 32u), 0u);
        uint32_t lo_a, lo_b, e_a, e_b;
        runs64f(sa, sb, lane, lo_a, lo_b, e_a, e_b);
        uint32_t starts = 0;
        if ((uint32_t)lane < nc) { ec_sum += min(e_a, nc) - lo_a; starts += lo_a == (uint32_t)lane; }
        if ((uint32_t)lane + 32u < nc) { ec_sum += min(e_b, nc) - lo_b; starts += lo_b == (uint32_t)lane + 32u; }
        const bool mg = fm_warp_sum(starts) * 3u <= nc;       // few distinct values: heavy ties
        if (lane == 0) H.merged = mg ? 1u : 0u;
        hca = slot_of(ca, mult0, shift); hcb = slot_of(cb, mult0, shift);
        if ((uint32_t)lane < nc) fm_sts_u32(tbl_s + hca, (uint32_t)lane);
        if ((uint32_t)lane + 32u < nc) fm_sts_u32(tbl_s + hcb, (uint32_t)lane + 32u);
      }
      if (!a.hist && !fuse_model)   // window / smoothing models need the row in time order
        model_nf_ok = fm_model(g, stage + kCur, nb, false, lane, mean_nf, sigma_nf);
    }
    // warp 1: Wilcoxon signed rank on the first / last min(n1, n2) pairs
    if (warp == 1) {
      fm_rank_aux axw = {0, 0, 0, 0, 0, 0u, 0u};
      Prep t = {false, false, 0.0, 0.0};
      if (want_wx) t = wilcoxon_stat(g, wilcoxon_load(g, stage_s, nc, nb, lane), lane, axw);
      if (lane == 0) {
        H.wx_stat = t.stat; H.wx_arg = t.arg; H.wx_flags = (t.run ? 1u : 0u) | (t.degenerate ? 2u : 0u);
        H.w2_plus = axw.w2_plus; H.w2_minus = axw.w2_minus; H.t_wx = axw.t_wx; H.n_wx = axw.n_wx;
      }
    }
    // everybody: my NV baseline vectors into registers, moments, pass A
    float x[4 * NV];
    uint32_t h[4 * NV];
    FmMom mom = {0.0, 0.0};
    const double pivot = nb ? (double)fm_lds_f32(base_s) : 0.0;
    const uint32_t nbv = (nb + 3u) >> 2;
#pragma unroll
    for (int c = 0; c < NV; ++c) {
      const uint32_t v = tid + kThreads * c;
      float4 qv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (v < nbv) qv = fm_lds_v4(base_s + (v << 4));
      x[4 * c + 0] = qv.x + 0.0f; x[4 * c + 1] = qv.y + 0.0f; x[4 * c + 2] = qv.z + 0.0f; x[4 * c + 3] = qv.w + 0.0f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t eidx = 4u * v + k;
        h[4 * c + k] = slot_of(x[4 * c + k], mult0, shift);
        if (eidx < nb) {
          if (fuse_model) fm_mom_add(mom, x[4 * c + k], pivot);
          if (ranks) fm_sts_u32(tbl_s + h[4 * c + k], kCur + eidx);
        }
      }
    }
    cta_sync();     // pass A complete, tree and flags published

    // ---- phase 2: pass B + search ------------------------------------------------------------
    uint32_t psum = 0;
    if (ranks) {
      e.merged = H.merged != 0u;
      const float4 n0 = fm_lds_v4(tree_s);
      const float s63 = fm_lds_f32(tree_s + 84u * 4u);
      uint32_t w[4 * NV];
#pragma unroll
      for (int i = 0; i < 4 * NV; ++i) w[i] = fm_lds_u32(tbl_s + h[i]);
      uint32_t wa = 0, wb = 0;
      if (warp == 0) { wa = fm_lds_u32(tbl_s + hca); wb = fm_lds_u32(tbl_s + hcb); }
#pragma unroll
      for (int c = 0; c < NV; ++c) {
        const uint32_t e0 = 4u * (tid + kThreads * c);
        float y[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) y[k] = (e0 + k < nb) ? x[4 * c + k] : -CUDART_INF_F;   // -inf counts nothing
        psum += tree_lower4(tree_s, n0, s63, y);
      }
#pragma unroll
      for (int i = 0; i < 4 * NV; ++i) {
        const uint32_t eidx = 4u * (tid + kThreads * (i >> 2)) + (i & 3);
        classify(e, lane, eidx < nb, x[i], kCur + eidx, h[i], w[i]);
      }
      if (warp == 0) {
        classify(e, lane, (uint32_t)lane < nc, ca, (uint32_t)lane, hca, wa);
        classify(e, lane, (uint32_t)lane + 32u < nc, cb, (uint32_t)lane + 32u, hcb, wb);
      }
    }

    // ---- phase 3: overflow rounds (classes that lost their slot) ------------------------------
    uint32_t tj = 0;
    bool done_a = !(warp == 0 && (uint32_t)lane < nc), done_b = !(warp == 0 && (uint32_t)lane + 32u < nc);
    auto collect = [&]() {   // current points whose class has been counted: read its total
      if (!done_a) {
        const uint32_t sl = fm_lds_u32(curslot_s + 4u * lane);
        if (sl) { tj += (sl & kDirect) ? (sl & 0xFFFFu) : 1u + (fm_lds_u32(sl) >> 16); done_a = true; }
      }
      if (!done_b) {
        const uint32_t sl = fm_lds_u32(curslot_s + 4u * (lane + 32u));
        if (sl) { tj += (sl & kDirect) ? (sl & 0xFFFFu) : 1u + (fm_lds_u32(sl) >> 16); done_b = true; }
      }
    };
    if (ranks) {
      int more = __syncthreads_or(e.nov > 0);        // also: pass B of everybody is complete
      collect();
      for (uint32_t round = 1; more && round < kMaxRounds; ++round) {
        const uint32_t n_in = e.nov, mult = hash_mult(round);
        cta_sync();                                   // collect() reads done before slots are reused
        for (uint32_t j = 0; j < n_in; ++j) {
          const uint32_t idx = fm_lds_u16(list_s + j * (kThreads * 2u));
          fm_sts_u32(tbl_s + slot_of(fm_lds_f32(stage_s + (idx << 2)) + 0.0f, mult, shift), idx);
        }
        cta_sync();
        e.nov = 0;
        // a warp walks its lists together so that merged mode can vote
        const uint32_t n_warp = __reduce_max_sync(FM_FULL, n_in);
        for (uint32_t j = 0; j < n_warp; ++j) {
          const bool ok = j < n_in;
          const uint32_t idx = ok ? fm_lds_u16(list_s + j * (kThreads * 2u)) : 0u;
          const float xv = fm_lds_f32(stage_s + (idx << 2)) + 0.0f;
          const uint32_t hh = slot_of(xv, mult, shift);
          classify(e, lane, ok, xv, idx, hh, fm_lds_u32(tbl_s + hh));   // e.nov <= j: in-place compaction
        }
        more = __syncthreads_or(e.nov > 0);
        collect();
      }
      if (more) {
        // Not reached in practice (12 multipliers).  The classes left are complete (equal values
        // always overflow together): thread 0 counts them pairwise, sum of t^2 - 1 over the points.
        __shared__ uint32_t left[kThreads];
        left[tid] = e.nov;
        cta_sync();
        if (tid == 0) {
          for (uint32_t t1 = 0; t1 < kThreads; ++t1)
            for (uint32_t j1 = 0; j1 < left[t1]; ++j1) {
              const uint32_t i1 = fm_lds_u16(list_s + t1 * 2u + j1 * (kThreads * 2u));
              const float x1 = fm_lds_f32(stage_s + (i1 << 2));
              uint32_t t = 0;
              for (uint32_t t2 = 0; t2 < kThreads; ++t2)
                for (uint32_t j2 = 0; j2 < left[t2]; ++j2)
                  t += fm_lds_f32(stage_s + (fm_lds_u16(list_s + t2 * 2u + j2 * (kThreads * 2u)) << 2)) == x1;
              e.tsum += (unsigned long long)t * t - 1ull;
              if (i1 < kCur) fm_sts_u32(curslot_s + (i1 << 2), kDirect | t);
            }
        }
        cta_sync();
        collect();
      }
    }

    // ---- phase 4: partial sums to warp 0 ---------------------------------------------------------
    {
      const double s1 = fm_warp_sum(mom.s1), s2 = fm_warp_sum(mom.s2);
      const long long ts = fm_warp_sum((long long)e.tsum);
      const uint32_t ps = fm_warp_sum(psum);
      if (lane == 0) { H.s1[warp] = s1; H.s2[warp] = s2; H.tsum[warp] = (unsigned long long)ts; H.psum[warp] = ps; }
    }
    cta_sync();

    // ---- phase 5: warp 0 turns the integers into the record ---------------------------------------
    if (warp == 0) {
      const uint32_t ty = a.metric_type ? a.metric_type[s] : 0u;
      const fm_type_params tp = a.P->types[ty < a.P->n_types ? ty : 0u];
      fm_series_rec rec;
      fm_rec_init(rec);
      fm_rank_aux ax = {0, 0, 0, 0, 0, 0u, 0u};
      uint32_t status = 0;
      if (nc == 0) status |= FM_ST_NO_CURRENT;
      bool model_ok;
      if (a.hist) {
        model_ok = fm_model(g, a.hist + s * a.hist_pitch, row_len(a.hist_len, s, a.n_hist), true, lane,
                            rec.mean, rec.sigma);
      } else if (fuse_model) {
        model_ok = nb >= max(g.min_hist_points, g.ddof + 1u);
        if (model_ok) {
          const double s1 = (H.s1[0] + H.s1[1]) + (H.s1[2] + H.s1[3]), s2 = (H.s2[0] + H.s2[1]) + (H.s2[2] + H.s2[3]);
          const double dn = (double)nb;
          rec.mean = pivot + s1 / dn;
          double m2 = s2 - s1 * s1 / dn;
          if (m2 < 0.0) m2 = 0.0;
          rec.sigma = (nb > g.ddof) ? sqrt(m2 / (double)(nb - g.ddof)) : 0.0;
        }
      } else {
        model_ok = model_nf_ok; rec.mean = mean_nf; rec.sigma = sigma_nf;
      }
      if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;

      Prep t_mw = {false, false, 0.0, 0.0}, t_kw = {false, false, 0.0, 0.0};
      Prep t_wx = {(H.wx_flags & 1u) != 0, (H.wx_flags & 2u) != 0, H.wx_stat, H.wx_arg};
      ax.n_wx = H.n_wx; ax.w2_plus = H.w2_plus; ax.w2_minus = H.w2_minus; ax.t_wx = H.t_wx;
      if (ranks) {
        const long long P = (long long)H.psum[0] + H.psum[1] + H.psum[2] + H.psum[3];
        const long long E = (long long)fm_warp_sum(tj) - (long long)fm_warp_sum(ec_sum);
        const long long T = (long long)(H.tsum[0] + H.tsum[1] + H.tsum[2] + H.tsum[3]);
        const long long r2c = (long long)nc * (nc + 1) + 2ll * nc * nb - 2ll * P - E;
        ax.r2_cur = r2c; ax.t_mw = T;
        rank_tests(g, want_mw, want_kw, nc, nb, r2c, T, t_mw, t_kw);
      }
      finish_series(a, g, tp, s, lane, nc, pair, t_mw, t_wx, t_kw, model_ok, craw_a, craw_b, rec, ax, status);
    }

    if (a.n_stages == 2) {
      st ^= 1u;
    } else {
      cta_sync();     // single stage: nobody may still read it
      if (sn < a.n_series && tid == 0) {
        fm_fence_proxy_async();
        issue_rows(a, stage_p[0], &bar[0], sn, row_len(a.cur_len, sn, a.n_cur), row_len(a.base_len, sn, a.n_base));
      }
    }
  }
}

}  // namespace

cudaError_t fm_canary_elect_smem_attr(int bytes) {
  cudaError_t e = cudaFuncSetAttribute(k_canary_elect<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_canary_elect<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_canary_elect<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_canary_elect<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  return e;
}

// nv = float4 vectors of the baseline per thread (1..4); a.warp_bytes = bytes of shared memory per CTA
int fm_launch_canary_elect(const FmKArgs& a, int nv, int grid, cudaStream_t st) {
  const size_t smem = a.warp_bytes;
  if (nv == 1) k_canary_elect<1><<<grid, kThreads, smem, st>>>(a);
  else if (nv == 2) k_canary_elect<2><<<grid, kThreads, smem, st>>>(a);
  else if (nv == 3) k_canary_elect<3><<<grid, kThreads, smem, st>>>(a);
  else k_canary_elect<4><<<grid, kThreads, smem, st>>>(a);
  return 1;
}

size_t fm_canary_elect_fixed_bytes() { return 16 + kTreeFloats * 4 + kCur * 4 + 2 * sizeof(Handoff); }
