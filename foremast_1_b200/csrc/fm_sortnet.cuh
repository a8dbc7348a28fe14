// fm_sortnet.cuh — register-resident bitonic sort of 32*R fp32 keys across a warp, and what the
// rank kernels do with the sorted row: tie term, parking in shared memory, bound searches.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "fm_device.cuh"

namespace fmc {

// ---- 32*R keys, R per lane, blocked: key index = R*lane + r --------------------------------
// Bitonic sort in its all-ascending form: phase K first compares i with i ^ (K-1) ("flip"), then
// i with i ^ J for J = K/4 .. 1; the lower index always keeps the minimum, so every lane-local
// stage is a fixed pair of FMNMX on fixed registers (no per-lane direction to select on).
// Template recursion over (K, J) keeps every register index a compile-time constant (a rolled
// loop would push v[] to local memory).
template <int R, int K>
__device__ __forceinline__ void sort_flip(float (&v)[R], int lane) {
  if constexpr (K > R) {
    // partner: lane ^ (K/R - 1), register R-1-r.  Pairs (r, R-1-r) are exchanged together so that
    // both shuffles read values from before this stage.
    const bool take_min = ((lane * R) & (K / 2)) == 0;
#pragma unroll
    for (int r = 0; r < R / 2; ++r) {
      const float o1 = __shfl_xor_sync(FM_FULL, v[R - 1 - r], K / R - 1);
      const float o2 = __shfl_xor_sync(FM_FULL, v[r], K / R - 1);
      v[r] = take_min ? fminf(v[r], o1) : fmaxf(v[r], o1);
      v[R - 1 - r] = take_min ? fminf(v[R - 1 - r], o2) : fmaxf(v[R - 1 - r], o2);
    }
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
__device__ __forceinline__ void sort_disperse(float (&v)[R], int lane) {
  if constexpr (J >= R) {
    const bool take_min = ((lane * R) & J) == 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float o = __shfl_xor_sync(FM_FULL, v[r], J / R);
      v[r] = take_min ? fminf(v[r], o) : fmaxf(v[r], o);
    }
  } else {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if ((r & J) == 0) {
        const float x = v[r], y = v[r | J];
        v[r] = fminf(x, y);
        v[r | J] = fmaxf(x, y);
      }
    }
  }
}
template <int R, int J>
struct SortMerge {
  static __device__ __forceinline__ void run(float (&v)[R], int lane) {
    sort_disperse<R, J>(v, lane);
    SortMerge<R, J / 2>::run(v, lane);
  }
};
template <int R>
struct SortMerge<R, 0> {
  static __device__ __forceinline__ void run(float (&)[R], int) {}
};
template <int R, int K, bool Done>
struct SortPhase {
  static __device__ __forceinline__ void run(float (&v)[R], int lane) {
    sort_flip<R, K>(v, lane);
    SortMerge<R, K / 4>::run(v, lane);
    SortPhase<R, K * 2, (K * 2 > 32 * R)>::run(v, lane);
  }
};
template <int R, int K>
struct SortPhase<R, K, true> {
  static __device__ __forceinline__ void run(float (&)[R], int) {}
};
template <int R>
__device__ __forceinline__ void sort_rows(float (&v)[R], int lane) {
  SortPhase<R, 2, false>::run(v, lane);
}

// transposed parking of the sorted row: key index i = R*lane + r lives at word r*32 + lane
template <int R>
__device__ __forceinline__ uint32_t park_addr(uint32_t base_s, uint32_t i) {
  return base_s + (((i & (R - 1)) << 5) + (i / R)) * 4u;
}
template <int R>
__device__ __forceinline__ uint32_t parked_lower(uint32_t base_s, uint32_t n, float x) {
  uint32_t lo = 0, len = n;
  while (len) {
    const uint32_t h = len >> 1;
    if (fm_lds_f32(park_addr<R>(base_s, lo + h)) < x) { lo += h + 1; len -= h + 1; } else len = h;
  }
  return lo;
}
template <int R>
__device__ __forceinline__ uint32_t parked_upper(uint32_t base_s, uint32_t n, float x) {
  uint32_t lo = 0, len = n;
  while (len) {
    const uint32_t h = len >> 1;
    if (!(x < fm_lds_f32(park_addr<R>(base_s, lo + h)))) { lo += h + 1; len -= h + 1; } else len = h;
  }
  return lo;
}


// Tie term sum(t^3 - t) of the first n keys of a row sorted by sort_rows (+inf padding sorts
// last), and the row parked at `park_s`.  Returns this lane's part; callers warp-sum it.
template <int R>
__device__ __forceinline__ unsigned long long tie_term_and_park(const float (&v)[R], uint32_t n, int lane,
                                                                 uint32_t park_s) {
  // lane-local runs; position i = R*lane + r counts while i < n
  const uint32_t i0 = (uint32_t)lane * R;
  const uint32_t nvalid = n > i0 ? min(n - i0, (uint32_t)R) : 0u;
  uint32_t run = 1, acc = 0;                  // acc = sum of c(c+1), c = position inside the run
#pragma unroll
  for (int r = 1; r < R; ++r) {
    const bool eq = (v[r] == v[r - 1]) && ((uint32_t)r < nvalid);
    run = eq ? run + 1u : 1u;
    acc += eq ? (run - 1u) * run : 0u;
  }
  unsigned long long T = 3ull * acc;
  // runs that continue from the previous lane: cin = points of my first value held by earlier lanes
  float vlast = v[0];
#pragma unroll
  for (int r = 1; r < R; ++r) vlast = ((uint32_t)r < nvalid) ? v[r] : vlast;
  const float prev_last = __shfl_up_sync(FM_FULL, vlast, 1);
  const uint32_t prev_tail = __shfl_up_sync(FM_FULL, run, 1);      // previous lane: run at its last key
  const bool prev_full = __shfl_up_sync(FM_FULL, (run == (uint32_t)R) ? 1 : 0, 1) != 0;
  const bool link = lane > 0 && nvalid > 0 && prev_last == v[0];
  uint32_t cin = link ? prev_tail : 0u;
  bool cont = link && prev_full;               // previous lane is one single run: add its carry too
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t cu = __shfl_up_sync(FM_FULL, cin, o);
    const int bu = __shfl_up_sync(FM_FULL, cont ? 1 : 0, o);
    if (lane >= o && cont) { cin += cu; cont = bu != 0; }
  }
  // park the sorted row (transposed: conflict-free) and correct my head run for the carry
#pragma unroll
  for (int r = 0; r < R; ++r) fm_sts_f32(park_s + ((uint32_t)(r * 32 + lane) << 2), v[r]);
  __syncwarp();
  if (cin > 0) {
    // what lies beyond this lane is corrected by the lanes that hold it
    const uint32_t hlen = min(parked_upper<R>(park_s, n, v[0]) - i0, nvalid);
    const unsigned long long h = hlen, c = cin;
    T += h * (3ull * c * (c + h));           // tie_gain(cin, hlen) - tie_gain(0, hlen)
  }
  return T;
}

}  // namespace fmc
