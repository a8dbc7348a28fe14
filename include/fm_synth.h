/*
 * fm_synth.h — counter-based synthetic metric series, bit-identical on CPU and GPU.
 *
 * Integer-only core (splitmix64 + Irwin-Hall(4) normal surrogate + popcount
 * counts), one exact int->fp32 conversion and one power-of-two scale per
 * sample, so host and device produce the same bits and a 6 GB workload never
 * has to exist on the host: any row can be regenerated from (seed, series id,
 * row kind, t).
 *
 * Value families follow the recording rules of the reference
 * (deploy/foremast/2_barrelman/metrics-rules-default.yaml:12-84): error series
 * are small integer counts / 60 (heavy ties, many exact zeros), latency is
 * positive seconds, cpu is cores in (1,4), memory is bytes ~5e8..1.6e9 with
 * sigma/mean ~1e-3.  Metric-type ids are the fm_default_params table:
 * 1 error5xx, 2 error4xx, 3 latency, 4 cpu, 5 memory.
 *
 * Pure C99 / CUDA C++; no dependencies.
 */
#ifndef FM_SYNTH_H_
#define FM_SYNTH_H_

#include <stdint.h>

#if defined(__CUDACC__)
#define FM_HD __host__ __device__ __forceinline__
#else
#define FM_HD static inline
#endif

#define FM_ROW_CUR  0u
#define FM_ROW_BASE 1u
#define FM_ROW_HIST 2u

#define FM_SYNTH_SEED 0xF04E3A57ull

FM_HD uint64_t fm_splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

FM_HD uint32_t fm_popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)__popcll((unsigned long long)x);
#else
  return (uint32_t)__builtin_popcountll((unsigned long long)x);
#endif
}

/* Per-series parameters, derived once per row. */
typedef struct {
  uint32_t type;       /* metric type id                          */
  uint32_t is_count;   /* 1: integer count family                 */
  int64_t  base_q;     /* continuous: centre in integer units     */
  int64_t  scale_q;    /* continuous: spread in integer units     */
  int32_t  exp2;       /* value = q * 2^exp2                      */
  uint64_t count_mask; /* count family: popcount(h & mask)        */
  uint32_t burst_add;  /* integer units added inside the burst    */
  uint32_t burst_t0, burst_len;   /* burst window in the current row, len 0 = none */
  uint32_t shifted;    /* 1: current row carries a location shift */
  int64_t  shift_q;
} fm_synth_series;

/* Metric type of series `id` under BASELINE.json config `config` (1..5). */
FM_HD uint32_t fm_synth_type(uint32_t config, uint64_t id) {
  if (config == 4u) return 1u;                    /* integer error counts only */
  const uint32_t k = (uint32_t)(id & 3ull);       /* 5xx, latency, cpu, memory */
  return k == 0u ? 1u : (k == 1u ? 3u : (k == 2u ? 4u : 5u));
}

FM_HD fm_synth_series fm_synth_series_params(uint64_t seed, uint32_t config, uint64_t id,
                                             uint32_t n_cur) {
  fm_synth_series p;
  /* config 4: ids are (service << 4) | group; the family is per service */
  const uint64_t fam = (config == 4u) ? (id >> 4) : id;
  const uint64_t hf = fm_splitmix64(seed ^ (fam * 0xD1342543DE82EF95ull) ^ 0xFA3111ull);
  const uint64_t hs = fm_splitmix64(seed ^ (id * 0xD1342543DE82EF95ull) ^ 0xB0757ull);
  p.type = fm_synth_type(config, id);
  p.is_count = (p.type == 1u || p.type == 2u) ? 1u : 0u;
  p.base_q = 0; p.scale_q = 0; p.exp2 = 0; p.count_mask = 0; p.burst_add = 0; p.shift_q = 0;
  if (p.is_count) {
    const uint32_t cls = (uint32_t)(hf % 3ull);           /* lambda 0.5 / 2 / 8 */
    p.count_mask = cls == 0u ? 0x1ull : (cls == 1u ? 0xFull : 0xFFFFull);
    p.burst_add = cls == 0u ? 3u : (cls == 1u ? 6u : 12u); /* 6 sigma of binomial(m, 1/2) */
  } else if (p.type == 3u) {                               /* latency, seconds */
    p.base_q = (int64_t)(1ull << 21) + (int64_t)(hf % (1ull << 22));
    p.scale_q = p.base_q >> 3;
    p.exp2 = -23;
  } else if (p.type == 4u) {                               /* cpu, cores */
    p.base_q = (int64_t)(1ull << 20) + (int64_t)(hf % (3ull << 20));
    p.scale_q = p.base_q >> 4;
    p.exp2 = -20;
  } else {                                                 /* memory, bytes */
    p.base_q = (int64_t)(1ull << 22) + (int64_t)(hf % (1ull << 23));
    p.scale_q = p.base_q >> 9;
    p.exp2 = 7;
  }
  if (!p.is_count) {
    /* sigma of the Irwin-Hall term is scale_q * 37837 / 65536 */
    const int64_t sigma_q = (p.scale_q * 37837) >> 16;
    p.burst_add = (uint32_t)(6 * sigma_q);
    p.shift_q = sigma_q >> 1;
  }
  /* 1 % of series: a 3..5 point burst in the current row */
  p.burst_len = 0u; p.burst_t0 = 0u;
  if ((hs & 0xFFFFull) < 655ull && n_cur >= 8u) {
    p.burst_len = 3u + (uint32_t)((hs >> 32) % 3ull);
    p.burst_t0 = (uint32_t)((hs >> 16) & 0xFFFFull) % (n_cur - 5u);
  }
  /* 5 % of series: the current row is shifted by half a sigma */
  p.shifted = (((hs >> 48) & 0xFFFFull) < 3277ull) ? 1u : 0u;
  return p;
}

/* Sample t of row `kind` of series `id`. */
FM_HD float fm_synth_sample(uint64_t seed, const fm_synth_series* p, uint64_t id,
                            uint32_t kind, uint32_t t) {
  const uint64_t h = fm_splitmix64(seed ^ (id << 20) ^ ((uint64_t)kind << 18) ^ (uint64_t)t);
  const int in_burst = (kind == FM_ROW_CUR) && (t - p->burst_t0 < p->burst_len);
  if (p->is_count) {
    uint32_t q = fm_popc64(h & p->count_mask);
    if (kind == FM_ROW_CUR && p->shifted) q += (((h >> 32) & 3ull) == 0ull) ? 1u : 0u;
    if (in_burst) q += p->burst_add;
#if defined(__CUDA_ARCH__)
    return __fdiv_rn((float)q, 60.0f);
#else
    return (float)q / 60.0f;
#endif
  }
  const int64_t g = (int64_t)(h & 0xFFFFull) + (int64_t)((h >> 16) & 0xFFFFull) +
                    (int64_t)((h >> 32) & 0xFFFFull) + (int64_t)((h >> 48) & 0xFFFFull);
  /* (g - 131070) * scale_q / 65536 without shifting a negative number */
  int64_t q = p->base_q + (int64_t)(((uint64_t)g * (uint64_t)p->scale_q) >> 16) -
              ((131070ll * p->scale_q) >> 16);
  if (kind == FM_ROW_CUR && p->shifted) q += p->shift_q;
  if (in_burst) q += (int64_t)p->burst_add;
  const float f = (float)(int32_t)q;                 /* q < 2^31 by construction */
  /* exact power-of-two scale */
  union { uint32_t u; float f; } s;
  s.u = (uint32_t)(127 + p->exp2) << 23;
  return f * s.f;
}

#endif /* FM_SYNTH_H_ */
