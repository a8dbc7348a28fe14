// fm_internal.h — shared between the host API (fm_api.cu) and the kernels.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/foremast_cuda.h"

// Parameters as the kernels see them (one device copy per ctx).
struct FmParamsDev {
  fm_global_params g;
  uint32_t n_types;
  uint32_t pad;
  fm_type_params types[FM_MAX_TYPES];
};

// Arguments of the per-series scoring kernels.  Pointers are device pointers.
struct FmKArgs {
  const float* cur;  const uint32_t* cur_len;  uint32_t n_cur,  cur_pitch;
  const float* base; const uint32_t* base_len; uint32_t n_base, base_pitch;
  const float* hist; const uint32_t* hist_len; uint32_t n_hist, hist_pitch;
  const uint8_t* metric_type;
  const uint32_t* api_id;
  uint32_t n_apis;         // roll-up ids >= n_apis are dropped on the device
  uint64_t n_series;
  void* out;               // fm_series_rec[] or fm_compact_rec[]
  fm_rank_aux* aux;        // optional
  fm_rollup_rec* api;      // optional: roll-up counters are accumulated in place
  const FmParamsDev* P;
  uint32_t record_kind;
  uint32_t strategy;
  // per-warp shared-memory layout, in elements
  uint32_t cur_cap;        // floats, raw current row (multiple of 4)
  uint32_t cs_cap;         // floats, sorted current row (power of two >= 32)
  uint32_t base_cap;       // floats, baseline row, sorted in place (power of two >= 32, or 0)
  uint32_t wx_cap;         // doubles, Wilcoxon |d| keys (power of two >= 32, or 0)
  uint32_t wx_n;           // doubles, raw differences (max pairs)
  uint32_t n_stages;       // 1 or 2 staging buffers per warp
  uint32_t warp_bytes;     // bytes of shared memory per warp
  uint32_t log_m;          // k_canary_elect: log2 of the election table size; k_canary_count: its slot count
  uint32_t cnt_nv, cnt_ns; // k_canary_count: vector rows / scalar rows of the staged baseline per lane
  double* hw_scratch;      // k_score_series, Holt-Winters: season_len doubles per resident warp
  // rows resident in the ctx's row cache (fm_score_batch_cached), read in place by k_canary_count,
  // k_canary_team and k_ma_only: series s takes its baseline / history row from slot row_slot[s] of
  // the cache unless that is 0xFFFFFFFF (row not resident: it is in base / hist like any other)
  uint32_t sync_mask;      // k_canary_count: CTA barrier on trips with (trip & sync_mask) == 0
  uint32_t tie_runs;       // k_canary_count: current rows with at most this many distinct values try the class path (0: off)
  const uint32_t* row_slot;
  const float* cache_base;
  const float* cache_hist;
};

struct FmGArgs {
  const uint32_t* group_offsets;
  const float* cur;  const uint32_t* cur_len;  uint32_t n_cur,  cur_pitch;
  const float* hist; const uint32_t* hist_len; uint32_t n_hist, hist_pitch;
  const uint8_t* metric_type;
  uint64_t n_services;
  fm_series_rec* out;
  fm_rollup_rec* svc;      // optional
  const FmParamsDev* P;
  const uint32_t* list;    // optional: work off list[0 .. *list_n) instead of all services
  const uint32_t* list_n;
  uint32_t pool_cap;       // floats: pooled sample, sorted (power of two)
  uint32_t raw_cap;        // floats: pooled sample in group order
  uint32_t warp_bytes;
};

// launchers (fm_kernels.cu); each returns the number of kernels launched
int fm_launch_score(const FmKArgs& a, int warps_per_cta, int grid, cudaStream_t st);
int fm_launch_groups(const FmGArgs& a, int warps_per_cta, int grid, cudaStream_t st);
// fm_groups.cu: services of <= 8 groups x <= 64 points; the others are appended to defer_list
int fm_launch_groups_sort(const FmGArgs& a, uint32_t* defer_list, uint32_t* defer_n, int grid, cudaStream_t st);
void fm_groups_sort_shape(int* warps_per_cta, int* ctas_per_sm);
int fm_launch_canary_sort(const FmKArgs& a, int regs, int valid_regs, int warps_per_cta, int grid, cudaStream_t st);
cudaError_t fm_canary_sort_smem_attr(int bytes);
int fm_launch_canary_elect(const FmKArgs& a, int nv, int grid, cudaStream_t st);
// fm_canary_count.cu: tier = (vector rows, scalar rows) per lane; a.log_m carries the slot count
int fm_launch_canary_count(const FmKArgs& a, int warps_per_cta, int grid, cudaStream_t st);
int fm_canary_count_max_warps();   // warps per CTA the kernel is built for (its register budget)
uint32_t fm_canary_count_tier(uint32_t n_base, int* nv, int* ns);
size_t fm_canary_count_fixed_bytes();
size_t fm_canary_team_fixed_bytes();
int fm_launch_canary_team(const FmKArgs& a, int warps, int grid, cudaStream_t st);
cudaError_t fm_canary_count_smem_attr(int bytes);
size_t fm_canary_elect_fixed_bytes();
int fm_launch_ma_only(const FmKArgs& a, int group, int grid, cudaStream_t st);
int fm_ma_only_ctas(int group);   // resident CTAs per SM the kernel is built for
cudaError_t fm_canary_elect_smem_attr(int bytes);
int fm_launch_rollup_finalize(fm_rollup_rec* api, uint32_t n_apis, const uint32_t* api_service,
                              fm_rollup_rec* svc, uint32_t n_services, cudaStream_t st);
int fm_launch_rollup_finalize_one(fm_rollup_rec* recs, uint64_t n, cudaStream_t st);
// fm_kernels.cu: cache rows (slot per row, 0xFFFFFFFF = skip) -> contiguous row matrix
int fm_launch_gather_rows(const float* cache, uint32_t pitch, const uint32_t* d_slot, float* dst, uint64_t n,
                          cudaStream_t st);
int fm_launch_rank_dump(const float* d_cur, uint32_t n_cur, const float* d_base, uint32_t n_base,
                        float* d_sorted /* pow2 scratch */, uint32_t cap, uint32_t* d_r2,
                        cudaStream_t st);
int fm_launch_ma_window(const float* d_x, uint32_t n, uint32_t w, uint32_t ddof, double* d_p1,
                        double* d_p2, double* d_mean, double* d_sigma, cudaStream_t st);
int fm_launch_synth(const fm_synth_desc& d, float* cur, float* base, float* hist, uint8_t* mt,
                    cudaStream_t st);

// fm_pairs.cu
int fm_launch_pairs(const fm_pair_batch& b, const FmParamsDev* P, fm_pair_rec* out, int num_sms, cudaStream_t st);

// message for fm_last_error(NULL) of the calling thread (entry points that take no ctx)
void fm_set_thread_error(const char* fmt, ...);

cudaError_t fm_score_smem_attr(int bytes);
cudaError_t fm_groups_smem_attr(int bytes);
