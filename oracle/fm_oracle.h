/*
 * fm_oracle.h — CPU restatement of Foremast's anomaly-scoring path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference leg may build, load or call it, and only as the checker or the
 * timed CPU baseline.  libforemast_cuda.so never links or calls this code.
 *
 * PARITY UNPINNED.  The brain's implementation is not in /root/reference
 * (foremast-brain/ holds only README.md; upstream is the Python program
 * github.com/intuit/foremast-brain, shipped as image foremast/foremast-brain:0.0.18,
 * deploy/foremast/3_brain/foremast-brain.yaml:19, CHANGELOG.md:12) and no
 * reference test pins any number of this path.  This file is a straight scalar
 * fp64 transcription of the documented behaviour:
 *   foremast-brain/README.md:5-10,20-38          criteria and knobs
 *   deploy/foremast/3_brain/foremast-brain.yaml:21-81   deployed values
 *   docs/guides/design.md:25,31-43,89-92         flow, fail-fast, unknown
 *   foremast-service/pkg/converter/converter.go:11-30   status strings
 * with the conventions the reference leaves open fixed in DESIGN.md section 3
 * (SURVEY.md appendix A).  It is pinned instead against hand-derived
 * known-answer vectors and scipy on float64 inputs (tests/test_oracle.py,
 * tests/golden/).
 */
#ifndef FM_ORACLE_H_
#define FM_ORACLE_H_

#include <stdint.h>
#include "../include/foremast_cuda.h" /* POD structs and constants only */

#ifdef __cplusplus
extern "C" {
#endif

/* A.1 two-pass mean / sigma (README.md:22 ML_ALGORITHM moving_average_all). */
void fmo_moments(const float* x, uint32_t n, uint32_t ddof, double* mean, double* sigma);

/* A.2 trailing-window series by direct two-pass per window; NaN before t = w-1. */
void fmo_ma_window_series(const float* x, uint32_t n, uint32_t w, uint32_t ddof,
                          double* mean_out, double* sigma_out);

/* A.3 2*midrank of every value (input order) and tie term sum(t^3 - t). */
void fmo_midranks(const double* v, uint32_t n, uint32_t* r2, int64_t* tie_term);

/* 2*midranks of cur U base (cur first). */
void fmo_rank_dump(const float* cur, uint32_t n_cur, const float* base, uint32_t n_base,
                   uint32_t* r2_out);

/* A.5 chi-square survival for integer dof, and 2*sf(z) of the normal. */
double fmo_chi2_sf(double x, uint32_t dof);

/* One series, all stages a2..a8 of SURVEY.md section 8. aux may be NULL. */
void fmo_score_series(const fm_global_params* g, const fm_type_params* tp, uint32_t strategy,
                      const float* cur, uint32_t n_cur, const float* base, uint32_t n_base,
                      const float* hist, uint32_t n_hist, fm_series_rec* rec, fm_rank_aux* aux);

/* Whole batch (host pointers), n_threads host threads, static partition by
 * series id.  out_records honours batch->record_kind.  Any output but
 * out_records may be NULL. */
int32_t fmo_score_batch(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                        const fm_batch* batch, void* out_records, fm_rank_aux* out_aux,
                        fm_rollup_rec* out_api, fm_rollup_rec* out_svc, uint32_t n_threads);

/* Rollover strategy: Kruskal-Wallis over pod groups + per-group bounds. */
int32_t fmo_score_groups(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                         const fm_group_batch* batch, fm_series_rec* out_groups,
                         fm_rollup_rec* out_svc, uint32_t n_threads);

/* A.9 roll-up of verdicts. */
void fmo_rollup(const uint8_t* verdict, const uint32_t* parent, uint64_t n, fm_rollup_rec* out,
                uint32_t n_parents);

/* Timed CPU baseline: regenerates `n_series` synthetic series of `desc`
 * (include/fm_synth.h) thread-locally and scores them; returns seconds spent
 * scoring (generation excluded) in *seconds, max over threads. */
int32_t fmo_bench_synth(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                        const fm_synth_desc* desc, uint32_t strategy, int hist_is_base,
                        uint32_t n_threads, double* seconds, uint64_t* checksum);

#ifdef __cplusplus
}
#endif
/* A.2d: bivariate normal over metric pairs (docs/guides/design.md:67-76). */
int32_t fmo_score_pairs(const fm_global_params* g, const fm_type_params* table, uint32_t n_types,
                        const fm_pair_batch* b, fm_pair_rec* out);

#endif
