// fm_sort48.cuh — the 1536-key warp sort (48 keys per lane) of k_canary_sort for baselines of
// 1025..1536 points (BASELINE.json configs[2]: 1440), with its tie term, parking and bound searches.
//
// The 2048-key network of fm_sortnet.cuh spends 440 of its 1440 merge comparators per lane on the
// +inf padding of a 1440-point row.  This is the same merge tree with those comparators removed;
// gen/gen_sort48.py explains why the rest is uniform over the lanes in two layouts of the compact
// key index C (striped: C = 32 q + lane; blocked-16: C = 16 (lane + 32 t) + r), simulates the
// whole network and writes the comparator lists (fm_sort48_net.cuh).
//   local      odd-even merge sort of the lane's 48 keys (384 comparators)
//   phase 128  merge-split with the neighbouring lane on sign-folded keys (one FMNMX per key),
//              pruned 64-slot merge in the lane (128 comparators); both lanes end ascending /
//              descending-negated and write their run to the buffer in compact order
//   phases 256..2048  striped rows from the buffer (upper run read lane-reversed), lane-local
//              comparators (56 / 80 / 104 / 128), rows back; blocked-16 read (128-bit), stride 16
//              with the neighbouring lane (sign-folded shuffle), strides 8..1 lane-local (96
//              comparators), written back (128-bit) for the next phase
// Buffer: 1536 words, 16-byte chunks XOR-swizzled by (row & 3): both access shapes conflict-free.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "fm_device.cuh"
#include "fm_sortnet.cuh"

namespace fmc {

#include "fm_sort48_net.cuh"

constexpr int kKeys48 = 48;
constexpr uint32_t kSort48Total = 32u * kKeys48;

struct Sort48Addr {
  uint32_t fwd[4], rev[4], blk[4];
};
__device__ __forceinline__ Sort48Addr sort48_addr(uint32_t buf_s, int lane) {
  Sort48Addr a;
  const uint32_t l = (uint32_t)lane, lr = 31u - l;
#pragma unroll
  for (uint32_t j = 0; j < 4; ++j) {
    a.fwd[j] = buf_s + 16u * ((l >> 2) ^ j) + 4u * (l & 3u);
    a.rev[j] = buf_s + 16u * ((lr >> 2) ^ j) + 4u * (lr & 3u);
    a.blk[j] = buf_s + 64u * l + 16u * (j ^ ((l >> 1) & 3u));
  }
  return a;
}

// strides 16 (neighbouring lane) and 8..1 on the blocked-16 layout, read from / written to the buffer
template <bool WRITE_BACK>
__device__ __forceinline__ void sort48_blocked(float (&k)[kKeys48], int lane, const uint32_t (&blk)[4]) {
#pragma unroll
  for (int t = 0; t < 3; ++t) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float4 q = fm_lds_v4(blk[c] + 2048u * t);
      k[16 * t + 4 * c + 0] = q.x; k[16 * t + 4 * c + 1] = q.y;
      k[16 * t + 4 * c + 2] = q.z; k[16 * t + 4 * c + 3] = q.w;
    }
  }
  const float sg = sort_sign(lane, 1);             // the odd lane keeps the maximum: it holds -key
#pragma unroll
  for (int r = 0; r < kKeys48; ++r) {
    const float w = k[r] * sg;
    k[r] = fm_fmul(fminf(w, -__shfl_xor_sync(FM_FULL, w, 1)), sg);
  }
#pragma unroll
  for (int d = 8; d > 0; d >>= 1) {
#pragma unroll
    for (int r = 0; r < kKeys48; ++r) {
      if ((r & d) == 0) {
        const float x = k[r], y = k[r | d];
        k[r] = fminf(x, y);
        k[r | d] = fmaxf(x, y);
      }
    }
  }
  if (WRITE_BACK) {
    __syncwarp();
#pragma unroll
    for (int t = 0; t < 3; ++t) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        fm_sts_v4(blk[c] + 2048u * t, k[16 * t + 4 * c], k[16 * t + 4 * c + 1], k[16 * t + 4 * c + 2],
                  k[16 * t + 4 * c + 3]);
    }
  }
}

// On return k[16 t + r] holds key 16 (lane + 32 t) + r of the ascending order; registers of
// missing points must hold +inf on entry.  buf_s: 6144 bytes of shared memory of this warp.
__device__ __forceinline__ void sort_rows48(float (&k)[kKeys48], int lane, uint32_t buf_s) {
  OemSort<kKeys48, kKeys48, 0, 64>::run(k);
  // phase 128: the even lane keeps the 48 smallest of the pair as -key, the odd lane the 48
  // largest; both sequences are then valleys and the same pruned merge sorts them ascending
  {
    const float sg = sort_sign(lane, 1);
#pragma unroll
    for (int r = 0; r < kKeys48 / 2; ++r) {
      const float wa = k[r] * sg, wb = k[kKeys48 - 1 - r] * sg;
      const float o1 = __shfl_xor_sync(FM_FULL, wb, 1);
      const float o2 = __shfl_xor_sync(FM_FULL, wa, 1);
      k[r] = fmaxf(-wa, o1);
      k[kKeys48 - 1 - r] = fmaxf(-wb, o2);
    }
    sort48_local_merge(k);
    // compact write: the even lane holds its run descending and negated
    const bool odd = (lane & 1) != 0;
    const int c0 = kKeys48 * lane + (odd ? 0 : kKeys48 - 1), step = odd ? 1 : -1;
    const float sv = odd ? 1.0f : -1.0f;
    __syncwarp();
#pragma unroll
    for (int r = 0; r < kKeys48; ++r) {
      const uint32_t C = (uint32_t)(c0 + step * r);
      fm_sts_f32(buf_s + 4u * (C ^ (((C >> 5) & 3u) << 2)), fm_fmul(k[r], sv));
    }
    __syncwarp();
  }
  const Sort48Addr a = sort48_addr(buf_s, lane);
  sort48_striped<256>(k, a.fwd, a.rev);
  __syncwarp();
  sort48_blocked<true>(k, lane, a.blk);
  __syncwarp();
  sort48_striped<512>(k, a.fwd, a.rev);
  __syncwarp();
  sort48_blocked<true>(k, lane, a.blk);
  __syncwarp();
  sort48_striped<1024>(k, a.fwd, a.rev);
  __syncwarp();
  sort48_blocked<true>(k, lane, a.blk);
  __syncwarp();
  sort48_striped<2048>(k, a.fwd, a.rev);
  __syncwarp();
  sort48_blocked<false>(k, lane, a.blk);
}

// ---- parked copy for the bound searches: key 16 (l + 32 t) + r at word (16 t + r) * 32 + l --------
// three levels, branch-free: which third (2 probes), which lane's block (5 probes over the block
// minima, contiguous words), which register (4 probes down the column)
template <bool UPPER>
__device__ __forceinline__ uint32_t parked48_bound(uint32_t base_s, uint32_t n, float x) {
  auto pred = [x](float key) { return UPPER ? !(x < key) : (key < x); };
  const bool any = pred(fm_lds_f32(base_s));
  const bool t1 = pred(fm_lds_f32(base_s + 2048u)), t2 = pred(fm_lds_f32(base_s + 4096u));
  const uint32_t third = base_s + (t2 ? 4096u : (t1 ? 2048u : 0u));
  uint32_t addr = third;
#pragma unroll
  for (int h = 16; h > 0; h >>= 1)
    if (pred(fm_lds_f32(addr + 4u * h))) addr += 4u * h;
  const uint32_t col = addr;
#pragma unroll
  for (int h = 8; h > 0; h >>= 1)
    if (pred(fm_lds_f32(addr + 128u * h))) addr += 128u * h;
  const uint32_t cnt = ((third - base_s) >> 2) + ((col - third) >> 2) * 16u + ((addr - col) >> 7) + 1u;
  return any ? min(cnt, n) : 0u;
}

// Tie term sum(t^3 - t) of the first n keys of the row sorted by sort_rows48, and the row parked at
// park_s (same buffer).  Same scheme as tie_term_and_park (fm_sortnet.cuh) on three thirds: the run
// that reaches the end of one third is carried into lane 0 of the next.
__device__ __forceinline__ unsigned long long tie_term_and_park48(const float (&k)[kKeys48], uint32_t n, int lane,
                                                                   uint32_t park_s) {
  __syncwarp();
#pragma unroll
  for (int q = 0; q < kKeys48; ++q) fm_sts_f32(park_s + 4u * (uint32_t)(q * 32 + lane), k[q]);
  __syncwarp();
  unsigned long long T = 0;
  float last_prev = 0.0f;          // last key of the previous third and the run that ends there
  uint32_t tail_prev = 0;
#pragma unroll
  for (int t = 0; t < 3; ++t) {
    float c = 0.0f, a1 = 0.0f, a2 = 0.0f;
#pragma unroll
    for (int r = 1; r < 16; ++r) {
      const float eq = fm_setf_eq(k[16 * t + r], k[16 * t + r - 1]);
      c = fmaf(c, eq, eq);
      a1 += c;
      a2 = fmaf(c, c, a2);
    }
    const uint32_t run = (uint32_t)c + 1u;
    T += 3ull * (uint32_t)(a1 + a2);
    float prev_last = __shfl_up_sync(FM_FULL, k[16 * t + 15], 1);
    uint32_t prev_tail = __shfl_up_sync(FM_FULL, run, 1);
    bool prev_full = __shfl_up_sync(FM_FULL, (run == 16u) ? 1 : 0, 1) != 0;
    if (lane == 0) { prev_last = last_prev; prev_tail = tail_prev; prev_full = false; }
    const bool link = (lane > 0 || t > 0) && prev_last == k[16 * t];
    uint32_t cin = link ? prev_tail : 0u;
    bool cont = link && prev_full;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t cu = __shfl_up_sync(FM_FULL, cin, o);
      const int bu = __shfl_up_sync(FM_FULL, cont ? 1 : 0, o);
      if (lane >= o && cont) { cin += cu; cont = bu != 0; }
    }
    if (cin > 0) {
      const uint32_t i0 = 16u * ((uint32_t)lane + 32u * t);
      const uint32_t hlen = min(parked48_bound<true>(park_s, kSort48Total, k[16 * t]) - i0, 16u);
      const unsigned long long h = hlen, cc = cin;
      T += h * (3ull * cc * (cc + h));
    }
    // what the next third's lane 0 continues: the whole run that ends at lane 31's last key
    const uint32_t tail31 = __shfl_sync(FM_FULL, run + ((run == 16u) ? cin : 0u), 31);
    last_prev = __shfl_sync(FM_FULL, k[16 * t + 15], 31);
    tail_prev = tail31;
  }
  if (lane == 0) {
    const unsigned long long m = kSort48Total - parked48_bound<false>(park_s, kSort48Total, CUDART_INF_F);
    const unsigned long long q = m - (kSort48Total - n);
    T += (q * q * q - q) - (m * m * m - m);
  }
  return T;
}

}  // namespace fmc
