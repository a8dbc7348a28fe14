// fm_canary.cu — the canary hot path (BASELINE.json configs[0..2]): a short current
// row (<= 64 points) against a baseline row (<= 1984 points), one warp per series.
//
// Nothing is sorted except the current row.  With x the current sample (n1 points,
// sorted in registers by a 64-key warp bitonic network) and y the baseline (n2):
//
//   2*R1 = n1(n1+1) + 2*n1*n2 - 2*P - E        P = sum_y #{x < y}     (4-ary search of
//                                               every baseline point in the sorted current row)
//                                               E = sum_x #{y == x}    (from the multiset below)
//   T    = sum over all n1+n2 points of 3c(c+1), c = how many equal values were counted
//          before this one  (sum_{c<t} 3c(c+1) = t^3 - t)
//
// The multiset is a per-warp shared-memory hash table of 32-bit slots
// {generation:5 | index of the first point with this value:11 | count:16}; a point claims
// a free slot with atomicCAS or bumps the count of its value's slot with atomicAdd, and the
// value returned is c.  Generations make clearing unnecessary (31 series per clear).
// Shared-memory atomics on spread addresses run at LDS rate on B200
// (profiles/microbench/README.md); sorting the baseline would cost ~10x more.
//
// Both integers are exactly the ones the sort-based path (fm_kernels.cu) and the oracle
// produce; tests/test_gpu_parity.py compares them bit for bit.
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "fm_device.cuh"

namespace {

constexpr uint32_t kCur = 64;          // floats reserved for the current row in a stage
constexpr uint32_t kTreeFloats = 88;   // 21 nodes x 4 floats, [84] = 64th sorted key
constexpr unsigned long long kAbsMask = 0x7FFFFFFFFFFFFFFFull;

struct FastSmem {
  uint64_t* bar;
  float* stage[2];   // [0,64) current row, [64, 64+base_cap) baseline row
  float* tree;
  uint32_t* tbl;
};

__device__ __forceinline__ FastSmem carve_fast(unsigned char* p, const FmKArgs& a) {
  FastSmem w;
  w.bar = reinterpret_cast<uint64_t*>(p);
  p += 16;
  const size_t stage_bytes = (size_t)(kCur + a.base_cap) * 4;
  w.stage[0] = reinterpret_cast<float*>(p);
  p += stage_bytes;
  if (a.n_stages == 2) { w.stage[1] = reinterpret_cast<float*>(p); p += stage_bytes; }
  else w.stage[1] = w.stage[0];
  w.tree = reinterpret_cast<float*>(p);
  p += kTreeFloats * 4;
  w.tbl = reinterpret_cast<uint32_t*>(p);
  return w;
}

__device__ __forceinline__ uint32_t row_len(const uint32_t* len, uint64_t s, uint32_t n) {
  return len ? min(__ldg(len + s), n) : n;
}

__device__ __forceinline__ void issue_fast(const FmKArgs& a, const FastSmem& w, uint32_t st,
                                           uint64_t s, uint32_t nc, uint32_t nb) {
  const uint32_t cb = ((nc + 3u) & ~3u) * 4u, bb = ((nb + 3u) & ~3u) * 4u;
  fm_mbar_expect_tx(&w.bar[st], cb + bb);
  if (cb) fm_bulk_g2s(w.stage[st], a.cur + s * a.cur_pitch, cb, &w.bar[st]);
  if (bb) fm_bulk_g2s(w.stage[st] + kCur, a.base + s * a.base_pitch, bb, &w.bar[st]);
}

// Count the value x (canonical: -0 folded to +0) of point `idx` in the multiset; returns how
// many equal values were counted before it and the slot it lives in.
// tbl / keys are 32-bit shared-space addresses; `slot` is the byte offset of the slot in the table.
__device__ __forceinline__ uint32_t ms_hash(float x, uint32_t shift) {
  return ((__float_as_uint(x) * 0x9E3779B1u) >> shift) << 2;
}
// v = the slot's contents as already loaded by the caller (lets four probes be in flight).
// Counts n points of value x at once (n > 1 when a warp has merged equal values of one row).
__device__ __forceinline__ uint32_t ms_insert(uint32_t tbl, uint32_t bmask, uint32_t gen, uint32_t keys,
                                              float x, uint32_t idx, uint32_t h, uint32_t v,
                                              uint32_t& slot, uint32_t n = 1u) {
  const uint32_t claim = (gen << 27) | (idx << 16) | n;
  for (;;) {
    if ((v >> 27) != gen) {                       // free: left over from an older series
      const uint32_t old = fm_atoms_cas(tbl + h, v, claim);
      if (old == v) { slot = h; return 0u; }
      v = old;                                    // another lane claimed it just now
    }
    if (fm_lds_f32(keys + ((v >> 14) & 0x1FFCu)) == x) {   // same value (IEEE ==: -0 == +0)
      slot = h;
      return fm_atoms_add(tbl + h, n) & 0xFFFFu;
    }
    h = (h + 4u) & bmask;                         // another value lives here: probe on
    v = fm_lds_u32(tbl + h);
  }
}

// t^3 - t grows by this when a tie run of length c gains n more points
__device__ __forceinline__ unsigned long long tie_gain(uint32_t c, uint32_t n) {
  const unsigned long long C = c, N = n;
  return N * (3ull * C * (C + N) + N * N - 1ull);
}

// Equal values of one row merged before they reach the table: same-address shared atomics
// serialise (a row of 32 equal values costs ~500 cycles), match.any costs 2 cycles per DISTINCT
// value, so this is used only for series whose current row shows heavy ties.  Returns c for the
// group leader; `n` = size of my group, `leader` = am I its lowest lane.  All lanes of `mask`
// must call.
__device__ __forceinline__ uint32_t ms_insert_merged(uint32_t mask, int lane, uint32_t tbl, uint32_t bmask,
                                                     uint32_t gen, uint32_t keys, float x, uint32_t idx,
                                                     uint32_t h, uint32_t& slot, uint32_t& n, bool& leader) {
  const uint32_t peers = __match_any_sync(mask, __float_as_uint(x));
  const int lead = __ffs((int)peers) - 1;
  n = (uint32_t)__popc(peers);
  leader = lane == lead;
  uint32_t c = 0;
  if (leader) c = ms_insert(tbl, bmask, gen, keys, x, idx, h, fm_lds_u32(tbl + h), slot, n);
  slot = __shfl_sync(peers, slot, lead);
  return c;
}

// 64 floats across the warp (a = index lane, b = index lane + 32), ascending.
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

// Same network on 64-bit keys compared with the top bit (the sign of the difference) masked.
__device__ __forceinline__ void cex64(unsigned long long& v, unsigned long long o, bool take_min) {
  const unsigned long long vm = v & kAbsMask, om = o & kAbsMask;
  const bool repl = take_min ? (om < vm) : (vm < om);
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

// sorted index -> float offset in the 4-ary search tree
__device__ __forceinline__ uint32_t tree_slot(uint32_t i) {
  if ((i & 15u) == 15u) return i == 63u ? 84u : (i >> 4);                 // node 0
  if ((i & 3u) == 3u) return (1u + (i >> 4)) * 4u + ((i & 15u) >> 2);      // nodes 1..4
  return (5u + (i >> 2)) * 4u + (i & 3u);                                  // nodes 5..20
}

// number of sorted current values strictly below y
// (four points at a time: the two dependent node loads of each are issued back to back)
__device__ __forceinline__ uint32_t cnt3(const float4& n, float y) {
  return (n.x < y ? 1u : 0u) + (n.y < y ? 1u : 0u) + (n.z < y ? 1u : 0u);
}
__device__ __forceinline__ uint32_t tree_lower4(uint32_t tree, const float4& n0, float s63,
                                                const float (&y)[4]) {
  uint32_t m[4];
  float4 n[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) m[k] = cnt3(n0, y[k]);
#pragma unroll
  for (int k = 0; k < 4; ++k) n[k] = fm_lds_v4(tree + ((1u + m[k]) << 4));
#pragma unroll
  for (int k = 0; k < 4; ++k) m[k] = 4u * m[k] + cnt3(n[k], y[k]);
#pragma unroll
  for (int k = 0; k < 4; ++k) n[k] = fm_lds_v4(tree + ((5u + m[k]) << 4));
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t pos = 4u * m[k] + cnt3(n[k], y[k]);
    sum += pos + ((pos == 63u && s63 < y[k]) ? 1u : 0u);
  }
  return sum;
}

struct Prep { bool run; bool degenerate; double stat; double arg; };   // p = min(1, erfc(arg))

__global__ void __launch_bounds__(256) k_canary_fast(const FmKArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int wpc = blockDim.x >> 5;
  const FastSmem w = carve_fast(smem + (size_t)warp * a.warp_bytes, a);
  const fm_global_params& g = a.P->g;
  const uint32_t M = 1u << a.log_m, bmask = (M << 2) - 1u, hshift = 32u - a.log_m;
  const uint32_t tbl_s = fm_smem_u32(w.tbl), tree_s = fm_smem_u32(w.tree);

  if (lane == 0) {
    for (uint32_t s = 0; s < a.n_stages; ++s) fm_mbar_init(&w.bar[s], 1);
    fm_fence_mbar_init();
  }
  for (uint32_t i = lane; i < M; i += 32) fm_sts_u32(tbl_s + (i << 2), 0u);   // generation 0 = never valid
  __syncwarp();

  const uint64_t stride = (uint64_t)gridDim.x * wpc;
  uint64_t s = (uint64_t)blockIdx.x * wpc + warp;
  uint32_t phase[2] = {0u, 0u};
  uint32_t st = 0, gen = 0;

  if (s < a.n_series && lane == 0)
    issue_fast(a, w, 0, s, row_len(a.cur_len, s, a.n_cur), row_len(a.base_len, s, a.n_base));

  for (; s < a.n_series; s += stride) {
    const uint32_t nc = row_len(a.cur_len, s, a.n_cur);
    const uint32_t nb = row_len(a.base_len, s, a.n_base);
    const uint64_t sn = s + stride;
    if (a.n_stages == 2 && sn < a.n_series) {
      // the other stage was only READ through the generic proxy; order those reads before the refill
      fm_fence_proxy_async();
      __syncwarp();
      if (lane == 0)
        issue_fast(a, w, st ^ 1u, sn, row_len(a.cur_len, sn, a.n_cur), row_len(a.base_len, sn, a.n_base));
    }
    fm_mbar_wait(&w.bar[st], phase[st]);
    phase[st] ^= 1u;

    const float* stage = w.stage[st];
    const float* cur = stage;
    const float* base = stage + kCur;
    const uint32_t stage_s = fm_smem_u32(stage);      // shared-space addresses for the hot loops
    const uint32_t base_s = stage_s + kCur * 4u;
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
    // the model is fused into the baseline pass when it is the plain mean / sigma of that row
    const bool fuse_model = (a.hist == nullptr) && g.ma_mode == FM_MA_ALL;

    // current row in registers, -0 folded, padded with +inf
    const float ca = (uint32_t)lane < nc ? fm_lds_f32(stage_s + 4u * lane) + 0.0f : CUDART_INF_F;
    const float cb = (uint32_t)lane + 32u < nc ? fm_lds_f32(stage_s + 4u * (lane + 32u)) + 0.0f : CUDART_INF_F;

    // ---- Wilcoxon signed rank on the first / last min(n1, n2) pairs (A.4) -------------
    Prep t_wx = {false, false, 0.0, 0.0};
    if (want_wx) {
      const uint32_t n = min(nc, nb);
      const uint32_t oc = g.pair_rule == FM_PAIR_TAIL ? nc - n : 0u;
      const uint32_t ob = g.pair_rule == FM_PAIR_TAIL ? nb - n : 0u;
      // key = |d| as ordered 64-bit integer, bit 63 = (d < 0); zero differences and padding
      // get the largest key so that they sort behind the n' used ones
      unsigned long long ka = kAbsMask, kb = kAbsMask;
      uint32_t nz = 0;
      if ((uint32_t)lane < n) {
        const double d = __dsub_rn((double)fm_lds_f32(stage_s + 4u * (oc + lane)),
                                   (double)fm_lds_f32(base_s + 4u * (ob + lane)));
        if (d != 0.0) ka = (unsigned long long)__double_as_longlong(d); else ++nz;
      }
      if ((uint32_t)lane + 32u < n) {
        const double d = __dsub_rn((double)fm_lds_f32(stage_s + 4u * (oc + lane + 32u)),
                                   (double)fm_lds_f32(base_s + 4u * (ob + lane + 32u)));
        if (d != 0.0) kb = (unsigned long long)__double_as_longlong(d); else ++nz;
      }
      nz = fm_warp_sum(nz);
      const uint32_t np = n - nz;
      ax.n_wx = np;
      if (np >= g.min_wx && np > 0) {
        sort64_abs(ka, kb, lane);
        // tie runs: lo = first position of my run, hi = one past its last position
        const unsigned long long ma = ka & kAbsMask, mb = kb & kAbsMask;
        const unsigned long long pa = __shfl_up_sync(FM_FULL, ma, 1);
        const unsigned long long last_a = __shfl_sync(FM_FULL, ma, 31);
        unsigned long long pb = __shfl_up_sync(FM_FULL, mb, 1);
        if (lane == 0) pb = last_a;
        const bool sa = lane == 0 || pa != ma, sb = pb != mb;
        uint32_t lo_a = sa ? (uint32_t)lane : 0u, lo_b = sb ? (uint32_t)lane + 32u : 0u;
        uint32_t hi_a = sa ? (uint32_t)lane : 64u, hi_b = sb ? (uint32_t)lane + 32u : 64u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t ua = __shfl_up_sync(FM_FULL, lo_a, o), ub = __shfl_up_sync(FM_FULL, lo_b, o);
          const uint32_t da = __shfl_down_sync(FM_FULL, hi_a, o), db = __shfl_down_sync(FM_FULL, hi_b, o);
          if (lane >= o) { lo_a = max(lo_a, ua); lo_b = max(lo_b, ub); }
          if (lane + o < 32) { hi_a = min(hi_a, da); hi_b = min(hi_b, db); }
        }
        lo_b = max(lo_b, __shfl_sync(FM_FULL, lo_a, 31));
        const uint32_t first_b = __shfl_sync(FM_FULL, hi_b, 0);     // first run start at or after 32
        hi_a = min(hi_a, first_b);
        // exclusive: the first start strictly after my position
        uint32_t ea = __shfl_down_sync(FM_FULL, hi_a, 1), eb = __shfl_down_sync(FM_FULL, hi_b, 1);
        if (lane == 31) { ea = first_b; eb = 64u; }
        long long wp = 0, wm = 0, T = 0;
        if ((uint32_t)lane < np) {
          const long long r2 = (long long)(lo_a + ea + 1u), t = (long long)(ea - lo_a);
          if (ka >> 63) wm += r2; else wp += r2;
          T += t * t - 1;
        }
        if ((uint32_t)lane + 32u < np) {
          const long long r2 = (long long)(lo_b + eb + 1u), t = (long long)(eb - lo_b);
          if (kb >> 63) wm += r2; else wp += r2;
          T += t * t - 1;
        }
        wp = fm_warp_sum(wp); wm = fm_warp_sum(wm); T = fm_warp_sum(T);
        ax.w2_plus = wp; ax.w2_minus = wm; ax.t_wx = T;
        t_wx.run = true;
        const double wmin = 0.5 * (double)(wp < wm ? wp : wm);
        const double dn = (double)np;
        const double mu = dn * (dn + 1.0) / 4.0;
        const double var = dn * (dn + 1.0) * (2.0 * dn + 1.0) / 24.0 - (double)T / 48.0;
        t_wx.stat = wmin;
        if (!(var > 0.0)) t_wx.degenerate = true;
        else {
          double num = wmin - mu;
          if (g.cc_wx && num != 0.0) num += 0.5;
          t_wx.arg = fabs(num / sqrt(var)) / 1.4142135623730951;
        }
      }
    }

    // ---- sorted current row, search tree, multiset of the current row --------------------
    uint32_t slot_a = 0, slot_b = 0, ec_sum = 0;
    unsigned long long tsum = 0;
    float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f);
    float s63 = CUDART_INF_F;
    bool merged = false;                        // warp-uniform: merge equal values of a row first
    if (ranks) {
      if (++gen == 32u) {                       // 5-bit generation wrapped: clear the table
        __syncwarp();
        for (uint32_t i = lane; i < M; i += 32) fm_sts_u32(tbl_s + (i << 2), 0u);
        gen = 1u;
        __syncwarp();
      }
      float sa = ca, sb = cb;
      sort64(sa, sb, lane);
      fm_sts_f32(tree_s + (tree_slot((uint32_t)lane) << 2), sa);
      fm_sts_f32(tree_s + (tree_slot((uint32_t)lane + 32u) << 2), sb);
      // distinct current values (adjacent compare on the sorted row): few of them = heavy ties
      float pa = __shfl_up_sync(FM_FULL, sa, 1), pb = __shfl_up_sync(FM_FULL, sb, 1);
      const float last_a = __shfl_sync(FM_FULL, sa, 31);
      if (lane == 0) pb = last_a;
      const uint32_t distinct =
          (uint32_t)__popc(__ballot_sync(FM_FULL, (uint32_t)lane < nc && (lane == 0 || pa != sa))) +
          (uint32_t)__popc(__ballot_sync(FM_FULL, (uint32_t)lane + 32u < nc && pb != sb));
      merged = distinct * 3u <= nc;
      __syncwarp();
      n0 = fm_lds_v4(tree_s);
      s63 = fm_lds_f32(tree_s + 84u * 4u);

      // the current row goes into the multiset first: c then counts ties inside the row
      const bool va = (uint32_t)lane < nc, vb = (uint32_t)lane + 32u < nc;
      const uint32_t ha = ms_hash(ca, hshift), hb = ms_hash(cb, hshift);
      if (merged) {
        uint32_t n; bool lead;
        const uint32_t ma = __ballot_sync(FM_FULL, va), mb = __ballot_sync(FM_FULL, vb);
        if (va) {
          const uint32_t c = ms_insert_merged(ma, lane, tbl_s, bmask, gen, stage_s, ca, (uint32_t)lane, ha, slot_a, n, lead);
          if (lead) { ec_sum += n * (2u * c + n); tsum += tie_gain(c, n); }
        }
        if (vb) {
          const uint32_t c = ms_insert_merged(mb, lane, tbl_s, bmask, gen, stage_s, cb, (uint32_t)lane + 32u, hb, slot_b, n, lead);
          if (lead) { ec_sum += n * (2u * c + n); tsum += tie_gain(c, n); }
        }
      } else {
        if (va) {
          const uint32_t c = ms_insert(tbl_s, bmask, gen, stage_s, ca, (uint32_t)lane, ha, fm_lds_u32(tbl_s + ha), slot_a);
          ec_sum += 2u * c + 1u; tsum += 3ull * c * (c + 1u);
        }
        if (vb) {
          const uint32_t c = ms_insert(tbl_s, bmask, gen, stage_s, cb, (uint32_t)lane + 32u, hb, fm_lds_u32(tbl_s + hb), slot_b);
          ec_sum += 2u * c + 1u; tsum += 3ull * c * (c + 1u);
        }
      }
    }

    // ---- one pass over the baseline row: moments, multiset, search -----------------------
    FmMom mom = {0.0, 0.0};
    const double pivot = nb ? (double)fm_lds_f32(base_s) : 0.0;
    uint32_t psum = 0;
    if (ranks || fuse_model) {
      const uint32_t nv = (nb + 3u) >> 2;
      for (uint32_t v0 = 0; v0 < nv; v0 += 32) {           // warp-uniform trip count
        const uint32_t v = v0 + lane;
        const bool in = v < nv;
        const float4 q = fm_lds_v4(base_s + ((in ? v : 0u) << 4));
        const uint32_t e0 = v << 2;
        const bool ok[4] = {in, in && e0 + 1u < nb, in && e0 + 2u < nb, in && e0 + 3u < nb};
        // points past the end of a ragged row count nothing: -inf is below every current value
        const float y[4] = {ok[0] ? q.x + 0.0f : -CUDART_INF_F, ok[1] ? q.y + 0.0f : -CUDART_INF_F,
                            ok[2] ? q.z + 0.0f : -CUDART_INF_F, ok[3] ? q.w + 0.0f : -CUDART_INF_F};
        if (fuse_model) {
#pragma unroll
          for (int k = 0; k < 4; ++k) if (ok[k]) fm_mom_add(mom, y[k], pivot);
        }
        if (ranks) {
          uint32_t h[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) h[k] = ms_hash(y[k], hshift);
          if (!merged) {
            const uint32_t t0 = fm_lds_u32(tbl_s + h[0]);
            psum += tree_lower4(tree_s, n0, s63, y);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (ok[k]) {
                uint32_t sl;
                const uint32_t c = ms_insert(tbl_s, bmask, gen, stage_s, y[k], kCur + e0 + k, h[k],
                                             k == 0 ? t0 : fm_lds_u32(tbl_s + h[k]), sl);
                tsum += 3ull * c * (c + 1u);
              }
            }
          } else {
            psum += tree_lower4(tree_s, n0, s63, y);
            uint32_t leaders = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint32_t mk = __ballot_sync(FM_FULL, ok[k]);
              if (ok[k]) {
                uint32_t sl, n; bool lead;
                const uint32_t c = ms_insert_merged(mk, lane, tbl_s, bmask, gen, stage_s, y[k], kCur + e0 + k,
                                                    h[k], sl, n, lead);
                if (lead) { tsum += tie_gain(c, n); ++leaders; }
              }
            }
            // a row with many distinct values makes match.any the slow part: stop merging
            if (__reduce_add_sync(FM_FULL, leaders) > 48u) merged = false;
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
      model_ok = fm_model(g, base, nb, false, lane, rec.mean, rec.sigma);
    }
    if (!model_ok) status |= FM_ST_HIST_INSUFFICIENT;

    // ---- a4/a5: statistics from the exact integers -------------------------------------------
    Prep t_mw = {false, false, 0.0, 0.0}, t_kw = {false, false, 0.0, 0.0};
    if (ranks) {
      __syncwarp();
      uint32_t tj = 0;                          // combined multiplicities of the current values
      if ((uint32_t)lane < nc) tj += fm_lds_u32(tbl_s + slot_a) & 0xFFFFu;
      if ((uint32_t)lane + 32u < nc) tj += fm_lds_u32(tbl_s + slot_b) & 0xFFFFu;
      const long long P = (long long)fm_warp_sum(psum);
      const long long E = (long long)fm_warp_sum(tj) - (long long)fm_warp_sum(ec_sum);
      const long long T = fm_warp_sum((long long)tsum);
      const long long r2c = (long long)nc * (nc + 1) + 2ll * nc * nb - 2ll * P - E;
      ax.r2_cur = r2c; ax.t_mw = T;
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

    // ---- a6..a8 ----------------------------------------------------------------------------------
    bool differ = false;
    if (pair) {
      // A.6 over the evaluated, non-degenerate tests
      const bool e0 = t_mw.run && !t_mw.degenerate, e1 = t_wx.run && !t_wx.degenerate,
                 e2 = t_kw.run && !t_kw.degenerate;
      const bool d0 = pv_mw < g.alpha, d1 = pv_wx < g.alpha, d2 = pv_kw < g.alpha;
      const bool any = (e0 && d0) || (e1 && d1) || (e2 && d2);
      const bool all = (!e0 || d0) && (!e1 || d1) && (!e2 || d2);
      if (!(e0 || e1 || e2)) status |= FM_ST_PAIRWISE_SKIPPED;
      else differ = g.pairwise_mode == FM_PW_ALL ? all : any;
      if (differ) status |= FM_ST_DIFFER;
    }
    if (model_ok && nc > 0) fm_threshold(g, tp, differ, cur, nc, lane, rec);
    if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT)) rec.verdict = FM_UNKNOWN;
    rec.status = (uint8_t)status;

    if (lane == 0) {
      fm_write_record(a.out, s, a.record_kind, rec);
      if (a.aux) a.aux[s] = ax;
      if (a.api && a.api_id) fm_rollup_add(a.api, a.api_id[s], rec.verdict);
    }
    __syncwarp();

    if (a.n_stages == 2) {
      st ^= 1u;
    } else if (sn < a.n_series) {
      fm_fence_proxy_async();
      __syncwarp();
      if (lane == 0)
        issue_fast(a, w, 0, sn, row_len(a.cur_len, sn, a.n_cur), row_len(a.base_len, sn, a.n_base));
    }
  }
}

}  // namespace

cudaError_t fm_canary_smem_attr(int bytes) {
  return cudaFuncSetAttribute(k_canary_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

int fm_launch_canary_fast(const FmKArgs& a, int warps_per_cta, int grid, cudaStream_t st) {
  k_canary_fast<<<grid, warps_per_cta * 32, (size_t)warps_per_cta * a.warp_bytes, st>>>(a);
  return 1;
}
