#!/usr/bin/env python3
"""Generator and simulator of the 1536-key warp sorting network (48 keys per lane) used by
k_canary_sort for 1024 < n_base <= 1536 (fm_sort48.cuh includes the file this script writes).

The network is the 2048-key bitonic merge tree of fm_sortnet.cuh with every comparator that would
touch the +inf padding (16 of the 64 slots of a lane) removed: 1000 comparators per lane instead
of 1440 in the merge phases.  Because the padding always ends up in whole 32-key rows, the
surviving comparators are uniform over the lanes in two layouts of the COMPACT (padding-free)
key index C in [0, 1536):

  striped     register q of lane l holds key C = 32 q + l       strides >= 32 are lane-local
  blocked-16  register 16 t + r of lane l holds C = 16 (l + 32 t) + r   strides < 16 are lane-local,
                                                  stride 16 is the neighbouring lane

Phase 128 (two lanes) runs in registers (merge-split with the partner lane, then a pruned 64-key
merge per lane); phases 256..2048 go through the warp's shared-memory buffer twice each.  This
script (1) builds the comparator lists, (2) SIMULATES the complete network, layout changes
included, on random, tied and padded inputs and checks the result against numpy's sort, (3) checks
that the shared-memory swizzle is bank-conflict free for both access shapes, (4) writes
fm_sort48_net.cuh.  Run it after any change:  python gen_sort48.py
"""
import os
import sys

import numpy as np

LANES, R, N = 32, 48, 1536


# ----------------------------------------------------------------------------- comparator lists
def local_merge64_pruned():
    """Bitonic merge of 64 slots whose upper 16 hold +inf, ascending: pairs on registers 0..47."""
    out = []
    for d in (32, 16, 8, 4, 2, 1):
        for i in range(64):
            if (i & d) == 0 and (i | d) < R:
                # slots >= 48 are +inf; after stride 32 the real keys of the upper half sit in
                # 32..47, and stride 16 of that half only meets padding
                if d == 16 and i >= 32:
                    continue
                out.append((i, i | d))
    return out


def striped_phase(K):
    """Phase K >= 256 in the striped layout.  Returns (loads, ces, n_rows) where loads are
    (dst_row, src_row, reversed) and ces are (row_lo, row_hi): minimum stays in row_lo."""
    a = K // 256
    rows_blk = 6 * a
    loads, ces = [], []
    for blk in range(2048 // K):
        base = rows_blk * blk
        A = lambda rho: base + rho                  # noqa: E731  input run A, row rho (0..3a-1)
        B = lambda j: base + 3 * a + j              # noqa: E731  input run B, row j
        for rho in range(4 * a):
            lo = base + rho
            if rho < a:
                loads.append((lo, A(rho), False))
            elif rho < 3 * a:
                hi = base + (4 * a + rho if rho < 2 * a else rho + 2 * a)
                loads.append((lo, A(rho), False))
                loads.append((hi, B(4 * a - 1 - rho), True))
                ces.append((lo, hi))
            else:
                loads.append((lo, B(4 * a - 1 - rho), True))
        d = 2 * a                                    # L: 4a rows
        while d >= 1:
            for r in range(4 * a):
                if (r & d) == 0:
                    ces.append((base + r, base + r + d))
            d //= 2
        d = a                                        # H': 2a rows
        while d >= 1:
            for r in range(2 * a):
                if (r & d) == 0:
                    ces.append((base + 4 * a + r, base + 4 * a + r + d))
            d //= 2
    return loads, ces


# ----------------------------------------------------------------------------------- simulation
def ce(k, i, j):
    lo, hi = np.minimum(k[i], k[j]), np.maximum(k[i], k[j])
    k[i], k[j] = lo, hi


def blocked16_stages(buf):
    """Strides 16 (neighbouring lane) and 8..1 on the compact array held in shared memory."""
    k = np.empty((R, LANES), buf.dtype)
    lane = np.arange(LANES)
    for t in range(3):
        for r in range(16):
            k[16 * t + r] = buf[16 * (lane + 32 * t) + r]
    for q in range(R):                               # stride 16: even lane keeps the minimum
        p = k[q][lane ^ 1]
        k[q] = np.where((lane & 1) == 0, np.minimum(k[q], p), np.maximum(k[q], p))
    for d in (8, 4, 2, 1):
        for q in range(R):
            if (q & d) == 0:
                ce(k, q, q | d)
    return k


def simulate(keys):
    """keys: (LANES, R) array, any order per lane (+inf padding allowed).  Returns the compact
    sorted array as the kernel leaves it (blocked-16 registers written back in compact order)."""
    lane = np.arange(LANES)
    k = np.sort(keys, axis=1).T.copy()               # local sort (any correct network)
    # phase 128: merge-split with the partner lane on sign-folded keys, pruned 64-merge
    sig = np.where(lane & 1, -1.0, 1.0).astype(k.dtype)
    w = k * sig
    newk = np.empty_like(k)
    for r in range(R):
        p = w[R - 1 - r][lane ^ 1]
        newk[r] = np.maximum(-w[r], p)
    k = newk
    for i, j in local_merge64_pruned():
        ce(k, i, j)
    buf = np.empty(N, k.dtype)
    for r in range(R):                               # compact write: even lanes reversed and negated
        C = np.where(lane & 1, R * lane + r, R * lane + (R - 1 - r))
        buf[C] = np.where(lane & 1, k[r], -k[r])
    for K in (256, 512, 1024, 2048):
        loads, ces = striped_phase(K)
        k = np.full((R, LANES), np.nan, buf.dtype)
        for dst, src, rev in loads:
            k[dst] = buf[32 * src + (31 - lane if rev else lane)]
        assert not np.isnan(k).any()
        for i, j in ces:
            ce(k, i, j)
        for q in range(R):
            buf[32 * q + lane] = k[q]
        k = blocked16_stages(buf)
        for t in range(3):
            for r in range(16):
                buf[16 * (lane + 32 * t) + r] = k[16 * t + r]
    return buf


def check_network():
    rng = np.random.default_rng(7)
    for trial in range(60):
        if trial % 3 == 0:
            x = rng.normal(size=(LANES, R))
        elif trial % 3 == 1:
            x = rng.integers(0, 5 + trial, size=(LANES, R)).astype(np.float64)     # heavy ties
        else:
            x = rng.normal(size=(LANES, R))
            flat = x.reshape(-1)
            flat[rng.choice(N, rng.integers(1, 400), replace=False)] = np.inf      # padding anywhere
        got = simulate(x.astype(np.float32))
        want = np.sort(x.astype(np.float32).reshape(-1))
        assert np.array_equal(got, want), trial
    # 0-1 principle on structured inputs as well: every threshold of a few random permutations
    for trial in range(20):
        perm = rng.permutation(N).reshape(LANES, R)
        for thr in rng.integers(0, N, 12):
            x = (perm >= thr).astype(np.float32)
            assert np.array_equal(simulate(x), np.sort(x.reshape(-1)))


# ------------------------------------------------------------------------------ swizzle checks
def word(C):
    """Shared-memory word of compact key C: 16-byte chunks XOR-swizzled by the row (C >> 5)."""
    return C ^ (((C >> 5) & 3) << 2)


def check_banks():
    lane = np.arange(LANES)
    for q in range(R):                               # striped, 32-bit: one wavefront
        assert len(set(word(32 * q + lane) % 32)) == 32
    for t in range(3):                               # blocked-16, 128-bit: 8 lanes per wavefront
        for c in range(4):
            for qw in range(4):
                ls = lane[8 * qw: 8 * qw + 8]
                w = word(16 * (ls + 32 * t) + 4 * c)
                assert (w % 4 == 0).all() and len(set((w // 4) % 8)) == 8
    # address forms used by the kernel
    for q in range(R):
        j = q & 3
        assert np.array_equal(word(32 * q + lane), 32 * q + (4 * ((lane >> 2) ^ j) + (lane & 3)))
    for t in range(3):
        for c in range(4):
            assert np.array_equal(word(16 * (lane + 32 * t) + 4 * c),
                                  16 * lane + 4 * (c ^ ((lane >> 1) & 3)) + 512 * t)


# ------------------------------------------------------------------------------------ emission
def emit(path):
    L = []
    w = L.append
    w("// fm_sort48_net.cuh -- GENERATED by gen/gen_sort48.py (comparator lists of the 1536-key network;")
    w("// see that script for the construction and its simulation).  Do not edit.")
    w("#pragma once")
    w("")
    w("#define FM_CE48(a, b) { const float x_ = k[a], y_ = k[b]; k[a] = fminf(x_, y_); k[b] = fmaxf(x_, y_); }")
    w("")
    w("// pruned 64-slot bitonic merge of a lane's 48 keys (slots 48..63 = +inf), ascending")
    w("__device__ __forceinline__ void sort48_local_merge(float (&k)[48]) {")
    for i, j in local_merge64_pruned():
        w(f"  FM_CE48({i}, {j})")
    w("}")
    w("")
    w("// striped part of phase K: rows from the buffer (fwd[j] / rev[j]: this lane's byte address in a")
    w("// row whose index is j mod 4, forward / lane-reversed), comparators, rows back to the buffer")
    w("template <int K> __device__ __forceinline__ void sort48_striped(float (&k)[48], const uint32_t (&fwd)[4],")
    w("                                                                const uint32_t (&rev)[4]);")
    for K in (256, 512, 1024, 2048):
        loads, ces = striped_phase(K)
        w(f"template <> __device__ __forceinline__ void sort48_striped<{K}>(float (&k)[48], const uint32_t (&fwd)[4],")
        w("                                                                   const uint32_t (&rev)[4]) {")
        for dst, src, rev in loads:
            w(f"  k[{dst}] = fm_lds_f32({'rev' if rev else 'fwd'}[{src & 3}] + {128 * src}u);")
        for i, j in ces:
            w(f"  FM_CE48({i}, {j})")
        w("  __syncwarp();")
        for q in range(R):
            w(f"  fm_sts_f32(fwd[{q & 3}] + {128 * q}u, k[{q}]);")
        w("}")
    w("")
    w("#undef FM_CE48")
    with open(path, "w") as f:
        f.write("\n".join(L) + "\n")
    n_ce = {K: len(striped_phase(K)[1]) for K in (256, 512, 1024, 2048)}
    return n_ce


if __name__ == "__main__":
    check_banks()
    check_network()
    here = os.path.dirname(os.path.abspath(__file__))
    out = os.path.join(here, "..", "fm_sort48_net.cuh")
    n_ce = emit(out)
    print("network verified; comparators per lane: local merge", len(local_merge64_pruned()),
          "striped", n_ce, "->", os.path.normpath(out))
    sys.exit(0)
