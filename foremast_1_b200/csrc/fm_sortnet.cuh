// fm_sortnet.cuh — register-resident bitonic sort of 32*R fp32 keys across a warp, and what the
// rank kernels do with the sorted row: tie term, parking in shared memory, bound searches.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "fm_device.cuh"

namespace fmc {

// ---- 32*R keys, R per lane, blocked: key index = R*lane + r --------------------------------
// Phase 1 sorts the R keys of every lane with Batcher's odd-even merge sort (543 comparators for
// R = 64 against 672 of the bitonic network); registers >= RV are known to hold +inf padding and
// the comparators that touch them are dropped at compile time (384 left for RV = 48).
// The merge phases are bitonic in the all-ascending form: phase K first compares i with
// i ^ (K-1) ("flip"), then i with i ^ J for J = K/4 .. 1; the lower index always keeps the
// minimum, so every lane-local stage is a fixed pair of FMNMX on fixed registers.
// Stages that cross lanes need min in one lane and max in its partner; ptxas turns a per-lane
// choice into TWO predicated FMNMX per key (the ALU pipe is the bound of this kernel), so those
// stages run on sign-folded keys instead: a lane whose direction bit is set holds -key, receives
// its partner's folded key p and keeps min(w, -p) = -max(key, partner key) -- one FMNMX with a
// free operand negation for both roles.  Moving between the conventions of two stages is one
// FMUL by +-1 per key (exact; it issues on the otherwise idle FMA pipe).
// Template recursion keeps every register index a compile-time constant (a rolled loop would
// push v[] to local memory).
template <int R, int RV, int I, int J>
__device__ __forceinline__ void sort_ce(float (&v)[R]) {
  if constexpr (J < RV) {
    const float x = v[I], y = v[J];
    v[I] = fminf(x, y);
    v[J] = fmaxf(x, y);
  }
}
template <int R, int RV, int I, int END, int DIST, int STEP>
struct OemRow {
  static __device__ __forceinline__ void run(float (&v)[R]) {
    if constexpr (I + DIST < END) {
      sort_ce<R, RV, I, I + DIST>(v);
      OemRow<R, RV, I + STEP, END, DIST, STEP>::run(v);
    }
  }
};
template <int R, int RV, int LO, int N, int DIST>
struct OemMerge {
  static __device__ __forceinline__ void run(float (&v)[R]) {
    if constexpr (2 * DIST < N) {
      OemMerge<R, RV, LO, N, 2 * DIST>::run(v);
      OemMerge<R, RV, LO + DIST, N, 2 * DIST>::run(v);
      OemRow<R, RV, LO + DIST, LO + N, DIST, 2 * DIST>::run(v);
    } else {
      sort_ce<R, RV, LO, LO + DIST>(v);
    }
  }
};
template <int R, int RV, int LO, int N>
struct OemSort {
  static __device__ __forceinline__ void run(float (&v)[R]) {
    if constexpr (N > 1 && LO < RV) {
      OemSort<R, RV, LO, N / 2>::run(v);
      OemSort<R, RV, LO + N / 2, N / 2>::run(v);
      OemMerge<R, RV, LO, N, 1>::run(v);
    }
  }
};

__device__ __forceinline__ float sort_sign(int lane, int bit) { return (lane & bit) ? -1.0f : 1.0f; }

// lane-crossing stages of phase K (K > R): direction bits K/2R, K/4R, .., 1
template <int R, int BIT>
struct SortCross {
  static __device__ __forceinline__ void run(float (&v)[R], int lane) {
    if constexpr (BIT >= 1) {
      const float c = sort_sign(lane, BIT * 2) * sort_sign(lane, BIT);   // previous -> this convention
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float w = v[r] * c;
        v[r] = fminf(w, -__shfl_xor_sync(FM_FULL, w, BIT));
      }
      SortCross<R, BIT / 2>::run(v, lane);
    } else {
      const float c = sort_sign(lane, 1);                                // back to plain keys
#pragma unroll
      for (int r = 0; r < R; ++r) v[r] = fm_fmul(v[r], c);                // (a plain product becomes an FSEL)
    }
  }
};
template <int R, int K>
__device__ __forceinline__ void sort_flip(float (&v)[R], int lane) {
  if constexpr (K > R) {
    // partner: lane ^ (K/R - 1), register R-1-r.  Pairs (r, R-1-r) are exchanged together so that
    // both shuffles read values from before this stage.
    const float c = sort_sign(lane, K / (2 * R));
#pragma unroll
    for (int r = 0; r < R / 2; ++r) {
      const float wa = v[r] * c, wb = v[R - 1 - r] * c;
      const float o1 = __shfl_xor_sync(FM_FULL, wb, K / R - 1);
      const float o2 = __shfl_xor_sync(FM_FULL, wa, K / R - 1);
      v[r] = fminf(wa, -o1);
      v[R - 1 - r] = fminf(wb, -o2);
    }
    SortCross<R, K / (4 * R)>::run(v, lane);
  } else {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if ((r & (K / 2)) == 0) {
        const float x = v[r], y = v[r ^ (K - 1)];
        v[r] = fminf(x, y);
        v[r ^ (K - 1)] = fmaxf(x, y);
      }
    }
  }
}
template <int R, int J>
__device__ __forceinline__ void sort_disperse(float (&v)[R]) {
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if ((r & J) == 0) {
      const float x = v[r], y = v[r | J];
      v[r] = fminf(x, y);
      v[r | J] = fmaxf(x, y);
    }
  }
}
template <int R, int J>
struct SortMerge {                       // lane-local strides J, J/2, .., 1
  static __device__ __forceinline__ void run(float (&v)[R]) {
    sort_disperse<R, J>(v);
    SortMerge<R, J / 2>::run(v);
  }
};
template <int R>
struct SortMerge<R, 0> {
  static __device__ __forceinline__ void run(float (&)[R]) {}
};
// ---- lane-crossing strides through shared memory (R >= 32, phases K >= kSmemFactor*R) ---------
// A phase with k lane-crossing strides costs 64(3k+1) issue slots per lane as shuffles.  From k = 4
// on, the row goes through the warp's parking buffer twice instead: written blocked, read
// STRIPED (key 32*q + lane in register q), where every stride >= 32 is a lane-local FMNMX pair,
// then written striped and read blocked for the strides below 32.  The reversal of the upper
// half of each K-block (the "flip") is folded into the striped read.  The buffer holds lane block
// l at words (R+1)*l .. (R+1)*l + R-1: the odd pitch makes the blocked accesses conflict-free
// (bank = l + r), the striped ones touch 32 consecutive words.
// (A rolled loop over the phases -- classic alternating directions on sign-folded keys, ~1200
// instructions of loop body -- was measured 40 % slower: the back edges defeat the sequential
// instruction prefetch that this mostly straight-line kernel lives on.)
#ifndef FM_SORT_SMEM_FACTOR
#define FM_SORT_SMEM_FACTOR 16
#endif
template <int R>
__device__ __forceinline__ constexpr uint32_t xp_word(int q) {          // word of key 32*q (lane 0)
  return (uint32_t)((R + 1) * ((32 * q) / R) + (32 * q) % R);
}
template <int R, int K>
__device__ __forceinline__ void sort_phase_smem(float (&v)[R], int lane, uint32_t buf_s) {
  const uint32_t blk = buf_s + 4u * (uint32_t)(R + 1) * (uint32_t)lane;
  const uint32_t fwd = buf_s + 4u * (uint32_t)lane, rev = buf_s + 4u * (uint32_t)(lane ^ 31);
  __syncwarp();
#pragma unroll
  for (int r = 0; r < R; ++r) fm_sts_f32(blk + 4u * r, v[r]);
  __syncwarp();
#pragma unroll
  for (int q = 0; q < R; ++q) {
    if (q & (K / 64)) v[q] = fm_lds_f32(rev + 4u * xp_word<R>(q ^ (K / 64 - 1)));   // upper half: reversed
    else v[q] = fm_lds_f32(fwd + 4u * xp_word<R>(q));
  }
  SortMerge<R, K / 64>::run(v);              // strides K/2 .. 32 = register distances K/64 .. 1
  __syncwarp();
#pragma unroll
  for (int q = 0; q < R; ++q) fm_sts_f32(fwd + 4u * xp_word<R>(q), v[q]);
  __syncwarp();
#pragma unroll
  for (int r = 0; r < R; ++r) v[r] = fm_lds_f32(blk + 4u * r);
  SortMerge<R, 16>::run(v);                  // strides 16 .. 1
}

template <int R, int K, bool Done>
struct SortPhase {
  static __device__ __forceinline__ void run(float (&v)[R], int lane, uint32_t buf_s) {
    if constexpr (R >= 32 && K >= FM_SORT_SMEM_FACTOR * R) {
      sort_phase_smem<R, K>(v, lane, buf_s);
    } else {
      sort_flip<R, K>(v, lane);          // K > R: includes the lane-crossing strides
      SortMerge<R, R / 2>::run(v);
    }
    SortPhase<R, K * 2, (K * 2 > 32 * R)>::run(v, lane, buf_s);
  }
};
template <int R, int K>
struct SortPhase<R, K, true> {
  static __device__ __forceinline__ void run(float (&)[R], int, uint32_t) {}
};
// RV: registers [RV, R) of every lane hold +inf on entry; buf_s: 32*(R+1) words of shared memory
template <int R, int RV = R>
__device__ __forceinline__ void sort_rows(float (&v)[R], int lane, uint32_t buf_s) {
  OemSort<R, RV, 0, R>::run(v);
  SortPhase<R, 2 * R, false>::run(v, lane, buf_s);
}

// transposed parking of the sorted row: key index i = R*lane + r lives at word r*32 + lane, every
// word of the 32*R block is written (padding = +inf), so the bound searches are branch-free: five
// steps over the lane minima (words 0..31), then log2(R) steps down that lane's column.
template <int R>
__device__ __forceinline__ uint32_t park_addr(uint32_t base_s, uint32_t i) {
  return base_s + (((i & (R - 1)) << 5) + (i / R)) * 4u;
}
template <int R, bool UPPER>
__device__ __forceinline__ uint32_t parked_bound(uint32_t base_s, uint32_t n, float x) {
  // pred(key) = key < x (lower bound) or key <= x (upper bound); count of keys with pred true
  auto pred = [x](float k) { return UPPER ? !(x < k) : (k < x); };
  const bool any = pred(fm_lds_f32(base_s));
  uint32_t addr = base_s;                      // word (0, B): B = last lane whose minimum passes
#pragma unroll
  for (int h = 16; h > 0; h >>= 1)
    if (pred(fm_lds_f32(addr + 4u * h))) addr += 4u * h;
  const uint32_t col = addr;
#pragma unroll
  for (int h = R / 2; h > 0; h >>= 1)          // last register of that lane that passes
    if (pred(fm_lds_f32(addr + 128u * h))) addr += 128u * h;
  const uint32_t cnt = ((col - base_s) >> 2) * R + ((addr - col) >> 7) + 1u;
  return any ? min(cnt, n) : 0u;
}
template <int R>
__device__ __forceinline__ uint32_t parked_lower(uint32_t base_s, uint32_t n, float x) {
  return parked_bound<R, false>(base_s, n, x);
}
template <int R>
__device__ __forceinline__ uint32_t parked_upper(uint32_t base_s, uint32_t n, float x) {
  return parked_bound<R, true>(base_s, n, x);
}


// Tie term sum(t^3 - t) of the first n keys of a row sorted by sort_rows (+inf padding sorts
// last), and the row parked at `park_s`.  Returns this lane's part; callers warp-sum it.
// The runs are counted over all 32*R keys without a validity test per key (one FSET on the ALU
// pipe, three FMA-pipe operations per key: c = position inside the run, exact in fp32); the run of
// +inf that the padding forms, together with any +inf in the data, is then replaced by the run
// the data alone would have.
template <int R>
__device__ __forceinline__ unsigned long long tie_term_and_park(const float (&v)[R], uint32_t n, int lane,
                                                                 uint32_t park_s) {
  const uint32_t i0 = (uint32_t)lane * R;
  float c = 0.0f, a1 = 0.0f, a2 = 0.0f;       // a1 + a2 = sum of c(c+1)
#pragma unroll
  for (int r = 1; r < R; ++r) {
    const float eq = fm_setf_eq(v[r], v[r - 1]);
    c = fmaf(c, eq, eq);                      // (c + 1) inside a run, 0 at its first key
    a1 += c;
    a2 = fmaf(c, c, a2);
  }
  const uint32_t run = (uint32_t)c + 1u;      // run at my last key
  unsigned long long T = 3ull * (uint32_t)(a1 + a2);
  // runs that continue from the previous lane: cin = points of my first value held by earlier lanes
  const float prev_last = __shfl_up_sync(FM_FULL, v[R - 1], 1);
  const uint32_t prev_tail = __shfl_up_sync(FM_FULL, run, 1);      // previous lane: run at its last key
  const bool prev_full = __shfl_up_sync(FM_FULL, (run == (uint32_t)R) ? 1 : 0, 1) != 0;
  const bool link = lane > 0 && prev_last == v[0];
  uint32_t cin = link ? prev_tail : 0u;
  bool cont = link && prev_full;               // previous lane is one single run: add its carry too
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t cu = __shfl_up_sync(FM_FULL, cin, o);
    const int bu = __shfl_up_sync(FM_FULL, cont ? 1 : 0, o);
    if (lane >= o && cont) { cin += cu; cont = bu != 0; }
  }
  // park the sorted row (transposed: conflict-free) and correct my head run for the carry
  __syncwarp();
#pragma unroll
  for (int r = 0; r < R; ++r) fm_sts_f32(park_s + ((uint32_t)(r * 32 + lane) << 2), v[r]);
  __syncwarp();
  if (cin > 0) {
    // what lies beyond this lane is corrected by the lanes that hold it
    const uint32_t hlen = min(parked_upper<R>(park_s, 32u * R, v[0]) - i0, (uint32_t)R);
    const unsigned long long h = hlen, cc = cin;
    T += h * (3ull * cc * (cc + h));         // tie_gain(cin, hlen) - tie_gain(0, hlen)
  }
  if (lane == 0) {
    const unsigned long long m = 32u * R - parked_lower<R>(park_s, 32u * R, CUDART_INF_F);  // keys equal to +inf
    const unsigned long long q = m - (32u * R - n);                                         // of them in the data
    T += (q * q * q - q) - (m * m * m - m);   // modular: the warp sum is what counts
  }
  return T;
}

}  // namespace fmc
