/*
 * foremast_cuda.h — C ABI of libforemast_cuda.so
 *
 * The batched canary-vs-baseline anomaly-scoring path of Foremast's brain
 * (moving-average mean +- k*sigma bounds, Mann-Whitney / Wilcoxon /
 * Kruskal-Wallis rank tests, verdict and roll-up), as sm_100a CUDA behind a
 * plain C interface.  Paths below are relative to the reference tree.
 *
 * The reference has no in-process plugin API for this path: foremast-service
 * writes an Elasticsearch document (foremast-service/pkg/models/models.go:83-101,
 * pkg/search/elasticsearchstore.go:22-62) and the brain (a separate Python
 * program, foremast-brain/README.md:1-38, not in the tree) polls it, scores it
 * and writes status/reason/anomalyInfo back (models.go:136-138,
 * pkg/converter/converter.go:11-30).  Each entry point here therefore cites the
 * reference text that defines the behaviour it implements; the binding a
 * maintainer adds on the reference side (cgo / ctypes) is in INTEGRATION.md.
 *
 * Conventions
 *   - every function returns int32_t: FM_OK (0) or a negative FM_E_* code;
 *     fm_last_error(ctx) gives a ctx-owned message valid until the next call.
 *   - all structs are POD; structs that may grow start with `struct_size`.
 *   - rows are series-major fp32, time-ascending, NaN-free (gaps are compacted
 *     by the host), `pitch` is in ELEMENTS, must be a multiple of 4 and the
 *     matrix base pointer must be 16-byte aligned (rows are moved with 16-byte
 *     bulk copies).  Per-row valid lengths are optional (NULL = every row has
 *     n_* valid points).
 *   - one ctx <-> one GPU <-> one stream; calls on one ctx must be serialised
 *     by the caller, different ctxs are independent.
 *   - no CPU fallback: without a usable CUDA device every compute entry point
 *     fails with FM_E_NODEVICE.
 */
#ifndef FOREMAST_CUDA_H_
#define FOREMAST_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FM_VERSION_MAJOR 0
#define FM_VERSION_MINOR 1

/* ---- error codes ------------------------------------------------------- */
#define FM_OK              0
#define FM_E_INVALID      -1  /* bad argument / layout contract violated      */
#define FM_E_CUDA         -2  /* CUDA runtime error (see fm_last_error)        */
#define FM_E_NOMEM        -3  /* host or device allocation failed              */
#define FM_E_UNSUPPORTED  -4  /* shape outside the limits stated below         */
#define FM_E_NODEVICE     -5  /* no CUDA device / driver                       */

/* ---- verdicts (design.md:25,43; converter.go:17-22) -------------------- */
#define FM_HEALTHY   0u  /* -> "completed_health"   */
#define FM_UNHEALTHY 1u  /* -> "completed_unhealth" */
#define FM_UNKNOWN   2u  /* -> "completed_unknown"  */

/* ---- per-series status bits -------------------------------------------- */
#define FM_ST_HIST_INSUFFICIENT 0x01u /* fewer than min_hist_points in the model window */
#define FM_ST_NO_CURRENT        0x02u /* no current points ("unknown", design.md:25)   */
#define FM_ST_DIFFER            0x04u /* pairwise says distributions differ: k_low used */
#define FM_ST_PAIRWISE_SKIPPED  0x08u /* pairwise requested but no eligible test        */
#define FM_ST_MW_RUN            0x10u /* Mann-Whitney evaluated                         */
#define FM_ST_WX_RUN            0x20u /* Wilcoxon evaluated                             */
#define FM_ST_KW_RUN            0x40u /* Kruskal-Wallis evaluated                       */
#define FM_ST_DEGENERATE        0x80u /* an evaluated test had zero variance: its p = 1 */

/* ---- strategies (metricsquery.go:17-19; BASELINE.json spellings) ------- */
#define FM_STRAT_CANARY      0u /* cur + base (+ hist): pairwise tests + bounds       */
#define FM_STRAT_HISTORICAL  1u /* cur + hist only: rollingUpdate/continuous/hpa      */

/* ---- moving-average model (foremast-brain.yaml:24-25) ------------------ */
#define FM_MA_ALL     0u  /* moving_average_all: one mean/sigma over the whole window */
#define FM_MA_WINDOW  1u  /* trailing window of ma_window points                      */
#define FM_MODEL_EWMA        2u /* exponential smoothing (design.md:60-66)            */
#define FM_MODEL_DOUBLE_EXP  3u /* double exponential smoothing (Holt linear)         */
#define FM_MODEL_HOLT_WINTERS 4u /* additive Holt-Winters, season of season_len points
                                  (design.md:60-66); series batches only, scored by
                                  the general kernel                                  */

/* ---- pairwise mode (foremast-brain/README.md:34) ----------------------- */
#define FM_PW_NONE       0u
#define FM_PW_MANN_WHITE 1u
#define FM_PW_WILCOXON   2u
#define FM_PW_KRUSKAL    3u
#define FM_PW_ALL        4u
#define FM_PW_ANY        5u
#define FM_PW_FRIEDMAN   6u  /* group batches only: Friedman chi-square over the pod groups as related
                              * samples ("Fried manchi square (special case)", docs/guides/design.md:92);
                              * not part of ALL / ANY; statistic and p-value are reported in h / p_h   */

#define FM_PAIR_HEAD 0u
#define FM_PAIR_TAIL 1u

#define FM_REC_FULL    0u  /* fm_series_rec, 96 B   */
#define FM_REC_COMPACT 1u  /* fm_compact_rec, 16 B  */

/* limits of this build */
#define FM_MAX_TYPES      64u
#define FM_MAX_CUR        4096u   /* n_cur (canary / group length)                 */
#define FM_MAX_RANK_N     16384u  /* n_cur + n_base, and sum of group lengths      */
#define FM_MAX_GROUPS     16u     /* Kruskal-Wallis groups per service             */

typedef struct fm_ctx fm_ctx;

/* Per metric-type thresholds: threshold{i}/bound{i}/min_lower_bound{i} of
 * deploy/foremast/3_brain/foremast-brain.yaml:26-73.  bound_mask: 1 = upper,
 * 2 = lower, 3 = both (README.md:24). k_low is the "lower threshold" used when
 * baseline and current differ (design.md:35). */
typedef struct {
  double   k;
  double   k_low;
  double   min_lower_bound;
  uint32_t bound_mask;
  uint32_t reserved;
} fm_type_params;

/* Global knobs: ML_ALGORITHM, MIN_HISTORICAL_DATA_POINT_TO_MEASURE,
 * ML_PAIRWISE_ALGORITHM, ML_PAIRWISE_THRESHOLD, MIN_*_DATA_POINTS
 * (README.md:22-38; foremast-brain.yaml:74-79). */
typedef struct {
  uint32_t struct_size;
  uint32_t ddof;               /* 0 or 1                                   */
  uint32_t ma_mode;            /* FM_MA_* / FM_MODEL_*                     */
  uint32_t ma_window;          /* w for FM_MA_WINDOW                       */
  uint32_t min_hist_points;    /* default 10                               */
  uint32_t pairwise_mode;      /* FM_PW_*                                  */
  uint32_t min_mw, min_wx, min_kw;   /* 20 / 20 / 5                        */
  uint32_t cc_mw, cc_wx;       /* continuity correction on/off (1 / 0)     */
  uint32_t pair_rule;          /* FM_PAIR_HEAD / FM_PAIR_TAIL              */
  uint32_t min_anomaly_points; /* default 1 (fail fast, design.md:43)      */
  uint32_t season_len;         /* FM_MODEL_HOLT_WINTERS: points per season (>= 2) */
  double   alpha;              /* ML_PAIRWISE_THRESHOLD, default 0.05      */
  double   smooth_alpha;       /* level factor for FM_MODEL_EWMA/DOUBLE_EXP */
  double   smooth_beta;        /* trend factor for FM_MODEL_DOUBLE_EXP / HOLT_WINTERS */
  double   smooth_gamma;       /* seasonal factor for FM_MODEL_HOLT_WINTERS  */
} fm_global_params;

/* Full per-series result, 96 B.  U1/W are exact half-integers; p-values and
 * H are fp64.  Tests that were not evaluated report stat = 0, p = 1. */
typedef struct {
  double   mean, sigma, upper, lower;   /* model + bounds (gauges _upper/_lower, metrics.js:14-20) */
  double   u1, p_u;                     /* Mann-Whitney U of the current sample, two-sided p        */
  double   w,  p_w;                     /* Wilcoxon min(W+,W-), two-sided p                         */
  double   h,  p_h;                     /* Kruskal-Wallis H (tie corrected), chi2 p                 */
  uint32_t n_anom;                      /* anomalous current points                                 */
  uint16_t first_idx;                   /* first anomalous index, 0xFFFF if none                    */
  uint8_t  verdict;                     /* FM_HEALTHY / FM_UNHEALTHY / FM_UNKNOWN                   */
  uint8_t  status;                      /* FM_ST_* bits                                             */
  uint64_t anom_mask;                   /* bit i = current point i anomalous (first 64 points)      */
} fm_series_rec;

typedef struct {
  uint64_t anom_mask;
  uint32_t n_anom;
  uint8_t  verdict;
  uint8_t  status;
  uint16_t first_idx;
} fm_compact_rec;

/* Exact integer rank statistics, for parity checks (optional output). */
typedef struct {
  int64_t r2_cur;      /* 2 * (rank sum of the current sample in cur U base)      */
  int64_t t_mw;        /* sum (t^3 - t) over tie runs of cur U base               */
  int64_t w2_plus;     /* 2 * W+                                                   */
  int64_t w2_minus;    /* 2 * W-                                                   */
  int64_t t_wx;        /* sum (t^3 - t) over tie runs of the non-zero |d|          */
  uint32_t n_wx;       /* non-zero differences used by Wilcoxon                    */
  uint32_t reserved;
} fm_rank_aux;

/* Roll-up of child verdicts (README.md:26; rule A.9 in DESIGN.md). */
typedef struct {
  uint32_t n_children;
  uint32_t n_unhealthy;
  uint32_t n_unknown;
  uint8_t  severity;    /* FM_HEALTHY / FM_UNHEALTHY / FM_UNKNOWN */
  uint8_t  reserved[3];
} fm_rollup_rec;

/* A batch of independent series.  The same struct describes host rows
 * (fm_score_batch) and device-resident rows (fm_score_batch_device). */
typedef struct {
  uint32_t struct_size;
  uint32_t strategy;        /* FM_STRAT_*                                          */
  uint64_t n_series;
  const float*    cur;   const uint32_t* cur_len;   uint32_t n_cur;   uint32_t cur_pitch;
  const float*    base;  const uint32_t* base_len;  uint32_t n_base;  uint32_t base_pitch;
  const float*    hist;  const uint32_t* hist_len;  uint32_t n_hist;  uint32_t hist_pitch;
                            /* hist == NULL with a baseline: the model window is the baseline row */
  const uint8_t*  metric_type;  /* index into the fm_set_params table, NULL = type 0 everywhere  */
  const uint32_t* api_id;       /* non-decreasing, < n_apis; NULL = no roll-up                    */
  const uint32_t* api_service;  /* [n_apis] non-decreasing, < n_services                          */
  uint32_t n_apis;
  uint32_t n_services;
  uint32_t record_kind;     /* FM_REC_FULL / FM_REC_COMPACT                        */
  uint32_t reserved;
} fm_batch;

/* Rollover strategy: Kruskal-Wallis over the pod groups of each service
 * (BASELINE.json configs[3]; README.md:34 KRUSKAL).  Rows of `cur` are groups;
 * service v owns rows group_offsets[v] .. group_offsets[v+1]-1 (3..FM_MAX_GROUPS
 * typical); every group is also bound-checked against the service's pooled
 * historical row. */
typedef struct {
  uint32_t struct_size;
  uint32_t reserved;
  uint64_t n_services;
  uint64_t n_groups;
  const uint32_t* group_offsets;   /* [n_services + 1] */
  const float*    cur;   const uint32_t* cur_len;   uint32_t n_cur;   uint32_t cur_pitch;   /* [n_groups] rows   */
  const float*    hist;  const uint32_t* hist_len;  uint32_t n_hist;  uint32_t hist_pitch;  /* [n_services] rows */
  const uint8_t*  metric_type;     /* [n_services] */
} fm_group_batch;

typedef struct {
  float    kernel_ms;      /* device time of all launches of the call (CUDA events on the ctx stream) */
  uint32_t launches;       /* kernels launched by this call                                            */
  float    h2d_ms, d2h_ms; /* reserved (0)                                                             */
  uint64_t h2d_bytes, d2h_bytes; /* fm_score_batch / fm_score_groups: bytes copied each way           */
  float    main_kernel_ms; /* device time of the scoring kernel alone (roofline numerator's clock)    */
  uint32_t reserved;
} fm_timing;

/* Synthetic workload descriptor (generator: include/fm_synth.h). */
typedef struct {
  uint32_t struct_size;
  uint32_t config;         /* 1..5 = BASELINE.json configs, value families per metric type */
  uint64_t seed;
  uint64_t first_series;   /* global id of row 0 (sharding by series id)                    */
  uint64_t n_series;
  uint32_t n_cur, n_base, n_hist;
  uint32_t cur_pitch, base_pitch, hist_pitch;
  const uint64_t* row_ids; /* optional [n_series] global id per row (device pointer for the
                              device variant); NULL: id = first_series + row.  Config 4 uses
                              id = (service << 4) | group. */
} fm_synth_desc;

/* ------------------------------------------------------------------------ */

/* (major << 16) | minor. */
int32_t fm_version(void);

/* sizeof() of the ABI structs as compiled, for bindings to verify their mirror:
 * 0 fm_global_params, 1 fm_type_params, 2 fm_batch, 3 fm_group_batch,
 * 4 fm_series_rec, 5 fm_compact_rec, 6 fm_rank_aux, 7 fm_rollup_rec,
 * 8 fm_timing, 9 fm_synth_desc; 0 for an unknown id. */
uint32_t fm_sizeof(uint32_t which);

/* Number of usable CUDA devices (0 and FM_OK when none). */
int32_t fm_device_count(int32_t* out);

/* Bind a GPU, create the stream and the grow-only device arenas.  Replaces the
 * brain process start-up (one brain pod == one scoring worker,
 * docs/guides/design.md:37-41). */
int32_t fm_ctx_create(int32_t device, fm_ctx** out);
int32_t fm_ctx_destroy(fm_ctx* ctx);

/* Fill `g` / `table[0..5]` with the deployed defaults of
 * deploy/foremast/3_brain/foremast-brain.yaml:24-79 (type 0 = default,
 * 1 error5xx, 2 error4xx, 3 latency, 4 cpu, 5 memory).  *n_types in: capacity,
 * out: 6. */
int32_t fm_default_params(fm_global_params* g, fm_type_params* table, uint32_t* n_types);

/* Install parameters on a ctx (the env-var block of foremast-brain.yaml:21-81). */
int32_t fm_set_params(fm_ctx* ctx, const fm_global_params* g,
                      const fm_type_params* table, uint32_t n_types);

/* Pinned host memory for row matrices and result arrays. */
int32_t fm_alloc_pinned(size_t bytes, void** out);
int32_t fm_free_pinned(void* p);

/* Device memory helpers for callers that keep rows resident (bench, tests). */
int32_t fm_device_alloc(fm_ctx* ctx, size_t bytes, void** out);
int32_t fm_device_free(fm_ctx* ctx, void* p);
int32_t fm_memcpy_h2d(fm_ctx* ctx, void* dst, const void* src, size_t bytes);
int32_t fm_memcpy_d2h(fm_ctx* ctx, void* dst, const void* src, size_t bytes);

/* Score a batch whose rows are in HOST memory; results land in host arrays.
 * This is the call the brain's job loop makes per poll: historical model ->
 * pairwise test -> lowered threshold -> anomaly points -> verdict
 * (docs/guides/design.md:35,43; foremast-brain/README.md:5-10).  Synchronous.
 * out_records: n_series records of batch->record_kind.  out_aux, out_api,
 * out_svc may be NULL.  Host->device copies are chunked and overlapped with
 * scoring; `timing` may be NULL. */
int32_t fm_score_batch(fm_ctx* ctx, const fm_batch* batch, void* out_records,
                       fm_rank_aux* out_aux, fm_rollup_rec* out_api,
                       fm_rollup_rec* out_svc, fm_timing* timing);

/* Device-resident row cache: MAX_CACHE_SIZE, "max cached model size"
 * (foremast-brain/README.md:30).  A job is re-judged on every poll until its
 * endTime (docs/guides/design.md:43); its baseline and 7-day history do not
 * change between polls, only the current window does.
 *
 * fm_cache_configure  max_bytes of HBM for resident rows on this ctx (0 = off,
 *                     the default; any call drops what is resident).
 * fm_score_batch_cached  fm_score_batch with one 64-bit id per series (the
 *                     caller's choice, e.g. hash of job id and metric alias).
 *                     Baseline / history rows whose id is resident are NOT read
 *                     from the host (those host rows need not even be valid:
 *                     ask fm_cache_query first and skip the fetch); the others
 *                     are uploaded and become resident, least recently scored
 *                     rows making room.  Current rows, metric types and API ids
 *                     are uploaded every call.  A resident row keeps the length
 *                     it was stored with (pass base_len / hist_len on every call
 *                     if any row is ragged).  Results are identical to
 *                     fm_score_batch on the same rows.  Two row layouts (n_base,
 *                     base_pitch, n_hist, hist_pitch) are resident at a time — the
 *                     canary rows and the history-only rows of a poll — each within
 *                     max_bytes; a third layout replaces the one used longest ago.
 * fm_cache_query      resident_out[i] = 1 if row_ids[i] is resident.
 * fm_cache_get_stats  counters since fm_cache_configure. */
typedef struct {
  uint64_t max_bytes, row_bytes, capacity_rows, resident_rows;
  uint64_t hits, misses, evictions, uncached;   /* uncached: rows of a call larger than the cache */
} fm_cache_stats;
int32_t fm_cache_configure(fm_ctx* ctx, uint64_t max_bytes);
int32_t fm_score_batch_cached(fm_ctx* ctx, const fm_batch* batch, const uint64_t* row_ids,
                              void* out_records, fm_rank_aux* out_aux, fm_rollup_rec* out_api,
                              fm_rollup_rec* out_svc, fm_timing* timing);
int32_t fm_cache_query(fm_ctx* ctx, const uint64_t* row_ids, uint64_t n, uint8_t* resident_out);
int32_t fm_cache_get_stats(fm_ctx* ctx, fm_cache_stats* out);

/* Two metrics of one job judged jointly: "Two Metrics — Bivariate Normal
 * Distribution" (docs/guides/design.md:67-76).  From the historical pairs
 * (hist_x[t], hist_y[t]) the means and the 2 x 2 covariance matrix (ddof of the
 * global params); a current pair is anomalous when its squared Mahalanobis
 * distance exceeds k^2, k = the threshold of the pair's metric type (rule A.2d in
 * DESIGN.md).  Fewer than max(min_hist_points, ddof + 1, 3) pairs -> unknown
 * (FM_ST_HIST_INSUFFICIENT); a singular covariance (constant or collinear
 * history: det <= 1e-9 var_x var_y) -> unknown (FM_ST_DEGENERATE).  Rows: 16-byte aligned, pitch a
 * multiple of 4, the two rows of a pair share their lengths. */
typedef struct {
  uint32_t struct_size;
  uint32_t reserved;
  uint64_t n_pairs;
  const float* cur_x;  const float* cur_y;  const uint32_t* cur_len;  uint32_t n_cur;  uint32_t cur_pitch;
  const float* hist_x; const float* hist_y; const uint32_t* hist_len; uint32_t n_hist; uint32_t hist_pitch;
  const uint8_t* metric_type;     /* [n_pairs], NULL = type 0 */
} fm_pair_batch;

typedef struct {
  double   mean_x, mean_y, var_x, var_y, cov_xy;
  double   d2_max;          /* largest squared Mahalanobis distance of a current pair */
  uint32_t n_anom;
  uint16_t first_idx;       /* 0xFFFF if none */
  uint8_t  verdict, status;
  uint64_t anom_mask;       /* first 64 current pairs */
} fm_pair_rec;

int32_t fm_score_pairs(fm_ctx* ctx, const fm_pair_batch* batch, fm_pair_rec* out, fm_timing* timing);
int32_t fm_score_pairs_device(fm_ctx* ctx, const fm_pair_batch* batch, fm_pair_rec* d_out, fm_timing* timing);

/* Several GPUs from ONE process (one host thread and one ctx per
 * GPU, results gathered into one pinned array; the shared-nothing brain nodes of
 * docs/guides/design.md:37-41 collapsed into the GPUs of one box).  A pool owns a
 * ctx per device (devices == NULL / n_devices == 0: every visible GPU; the same
 * device may be listed twice).  fm_pool_score_batch cuts the batch into one
 * contiguous range of series per ctx — at service boundaries when api_id /
 * api_service are given, so every API and service is rolled up on one GPU —
 * scores the ranges concurrently (fm_score_batch, or fm_score_batch_cached when
 * row_ids != NULL) and writes every range's records straight into its slice of
 * the caller's arrays.  Results are byte-identical to one ctx scoring the whole
 * batch.  No collective and no peer access.  timing_per_device: NULL or
 * fm_pool_size() entries.  A pool must be used from one thread at a time. */
typedef struct fm_pool fm_pool;
int32_t fm_pool_create(const int32_t* devices, int32_t n_devices, fm_pool** out);
int32_t fm_pool_destroy(fm_pool* pool);
int32_t fm_pool_size(fm_pool* pool);
fm_ctx* fm_pool_ctx(fm_pool* pool, int32_t i);
int32_t fm_pool_set_params(fm_pool* pool, const fm_global_params* g, const fm_type_params* table, uint32_t n_types);
int32_t fm_pool_cache_configure(fm_pool* pool, uint64_t max_bytes_per_device);
int32_t fm_pool_score_batch(fm_pool* pool, const fm_batch* batch, const uint64_t* row_ids, void* out_records,
                            fm_rank_aux* out_aux, fm_rollup_rec* out_api, fm_rollup_rec* out_svc,
                            fm_timing* timing_per_device);
const char* fm_pool_last_error(fm_pool* pool);

/* Same, with every pointer in `batch` and every output a DEVICE pointer.
 * Asynchronous on the ctx stream unless `timing` is non-NULL (then it
 * synchronises to read the events). */
int32_t fm_score_batch_device(fm_ctx* ctx, const fm_batch* batch, void* d_out_records,
                              fm_rank_aux* d_out_aux, fm_rollup_rec* d_out_api,
                              fm_rollup_rec* d_out_svc, fm_timing* timing);

/* Rollover strategy (Kruskal-Wallis over pod groups).  out_group_records:
 * n_groups fm_series_rec (model/bounds/anomalies per group, H and p_h of its
 * service); out_svc: n_services roll-ups over the groups, may be NULL. */
int32_t fm_score_groups(fm_ctx* ctx, const fm_group_batch* batch,
                        fm_series_rec* out_group_records, fm_rollup_rec* out_svc,
                        fm_timing* timing);
int32_t fm_score_groups_device(fm_ctx* ctx, const fm_group_batch* batch,
                               fm_series_rec* d_out_group_records, fm_rollup_rec* d_out_svc,
                               fm_timing* timing);

/* Parity aid: 2 * midrank of every point of cur U base (cur first), one series,
 * host pointers.  r2_out has n_cur + n_base entries. */
int32_t fm_rank_dump(fm_ctx* ctx, const float* cur, uint32_t n_cur,
                     const float* base, uint32_t n_base, uint32_t* r2_out);

/* Trailing-window moving average series (gauges _upper/_lower over time,
 * foremast-browser/src/config/metrics.js:14-20): for one host row x[0..n),
 * writes mean_t and sigma_t for t >= w-1 (NaN before). */
int32_t fm_ma_window_series(fm_ctx* ctx, const float* x, uint32_t n, uint32_t w,
                            uint32_t ddof, double* mean_out, double* sigma_out);

/* The same trailing-window series for callers that must not allocate.
 * (No counterpart in the reference: it exists for hosts of this library.)
 *
 * fm_ma_window_series above stages its row and keeps two prefix-sum arrays in
 * device memory that the ctx obtains with cudaMalloc.  Some embedders cannot
 * afford that: a service that hands the whole GPU to one pool owner at start-up
 * and forbids driver allocations afterwards, or a polling loop with a latency
 * budget that cudaMalloc's device-wide synchronisation would break.  This entry
 * point needs no ctx and performs no allocation of any kind: the caller passes
 * DEVICE pointers for the row and both outputs, and the prefix sums live in a
 * fixed array compiled into the library (2 * FM_STATIC_MA_MAX_POINTS doubles =
 * 256 KiB, resident on every GPU that loads the library until the process exits).
 *
 *   - n > FM_STATIC_MA_MAX_POINTS is refused with FM_E_UNSUPPORTED (message via
 *     fm_last_error(NULL)); nothing is launched.
 *   - makes `device` current and queues one kernel on `stream` (a cudaStream_t
 *     passed as void*, NULL = the default stream); it does not wait for it.
 *   - the first call on each device looks up the device address of the built-in
 *     array (once per device); make that first call before starting any stream
 *     capture on that device.
 *   - every call on a device shares the one array.  It keeps the previous call's
 *     prefix sums; that is harmless because each call rewrites entries [0, n)
 *     before reading them and reads nothing else, so nothing has to be cleared.
 *     Two calls that overlap on the same GPU (different streams or threads)
 *     race on it: serialise them. */
#define FM_STATIC_MA_MAX_POINTS 16384u
int32_t fm_ma_window_series_static(int32_t device, void* stream, const float* d_x, uint32_t n,
                                   uint32_t w, uint32_t ddof, double* d_mean_out,
                                   double* d_sigma_out);

/* Synthetic rows, bit-identical on host and device (include/fm_synth.h).
 * The *_device variant writes device matrices, the *_host variant host ones. */
int32_t fm_synth_fill_device(fm_ctx* ctx, const fm_synth_desc* d, float* d_cur,
                             float* d_base, float* d_hist, uint8_t* d_metric_type);
int32_t fm_synth_fill_host(const fm_synth_desc* d, float* cur, float* base,
                           float* hist, uint8_t* metric_type, uint32_t n_threads);

/* Host only (no GPU, no ctx): decode a Prometheus query_range response body -- the step right
 * before the hot path: the brain fetches the URLs of currentConfig / baselineConfig /
 * historicalConfig (foremast-service/cmd/manager/main.go:28-74) and gets
 * {"status":"success","data":{"result":[{"metric":{..},"values":[[ts,"v"],..]},..]}}
 * (foremast-browser/src/App.js:278-292).  The finite samples of result series i are written to
 * ts_out / val_out [series_start[i], series_start[i+1]) in the order of the body; NaN / +-Inf
 * samples (Prometheus gaps) are dropped; values are np.float32(float(v)) bit for bit.
 * series_start needs cap_series >= number of series + 1 entries.  Returns the number of samples
 * written, FM_E_INVALID for a malformed body or status != "success", FM_E_NOMEM when a capacity
 * is too small (text: fm_last_error(NULL)). */
int64_t fm_decode_query_range(const char* body, uint64_t len, int64_t* ts_out, float* val_out,
                              uint64_t cap_samples, uint32_t* series_start, uint32_t cap_series,
                              uint32_t* n_series);

/* Block until all work queued on the ctx stream has finished. */
int32_t fm_ctx_sync(fm_ctx* ctx);

/* Message for the last failing call on this ctx (NULL ctx: last create error). */
const char* fm_last_error(fm_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* FOREMAST_CUDA_H_ */
