// fm_pairs.cu — "Two Metrics: Bivariate Normal Distribution" (docs/guides/design.md:67-76): two
// metrics of one job judged jointly.  From the historical pairs (x_t, y_t) the means and the 2 x 2
// covariance matrix; a current pair is anomalous when its Mahalanobis distance from the mean
// exceeds the metric type's threshold k (the k-sigma ellipse; A.2d in DESIGN.md section 3).
//
// One warp per pair, pure streaming: both history rows are read once with 128-bit loads (eight in
// flight per lane), five pivoted fp64 sums per lane, one warp reduction; the current rows are <=
// 4096 points.  HBM-bound like k_ma_only; algorithmic bytes per pair = 8 (n_hist + n_cur) + 64.
#include <cstdint>
#include <cuda_runtime.h>

#include "fm_device.cuh"

namespace {

struct Mom2 { double sx, sy, sxx, syy, sxy; };

__device__ __forceinline__ void add2(Mom2& m, float x, float y, double px, double py) {
  const double dx = (double)x - px, dy = (double)y - py;
  m.sx += dx; m.sy += dy;
  m.sxx = fma(dx, dx, m.sxx); m.syy = fma(dy, dy, m.syy); m.sxy = fma(dx, dy, m.sxy);
}

__global__ void __launch_bounds__(256, 4) k_score_pairs(const fm_pair_batch b, const FmParamsDev* P, fm_pair_rec* out) {
  const int lane = threadIdx.x & 31;
  const fm_global_params& g = P->g;
  const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const uint32_t need = max(max(g.min_hist_points, g.ddof + 1u), 3u);
  for (uint64_t s = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < b.n_pairs; s += warps) {
    const uint32_t nh = b.hist_len ? min(__ldg(b.hist_len + s), b.n_hist) : b.n_hist;
    const uint32_t nc = b.cur_len ? min(__ldg(b.cur_len + s), b.n_cur) : b.n_cur;
    const float* hx = b.hist_x + s * b.hist_pitch;
    const float* hy = b.hist_y + s * b.hist_pitch;
    const double px = nh ? (double)__ldg(hx) : 0.0, py = nh ? (double)__ldg(hy) : 0.0;
    Mom2 m = {0.0, 0.0, 0.0, 0.0, 0.0};
    const uint32_t n4 = nh >> 2;
    const float4* x4 = reinterpret_cast<const float4*>(hx);
    const float4* y4 = reinterpret_cast<const float4*>(hy);
    for (uint32_t i0 = 0; i0 < n4; i0 += 128) {          // 4 vectors of each row per lane in flight
      float4 qx[4], qy[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t v = i0 + lane + 32u * j;
        qx[j] = qy[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (v < n4) { qx[j] = fm_ldg_stream(x4 + v); qy[j] = fm_ldg_stream(y4 + v); }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (i0 + lane + 32u * j < n4) {
          add2(m, qx[j].x, qy[j].x, px, py); add2(m, qx[j].y, qy[j].y, px, py);
          add2(m, qx[j].z, qy[j].z, px, py); add2(m, qx[j].w, qy[j].w, px, py);
        }
      }
    }
    for (uint32_t t = (n4 << 2) + lane; t < nh; t += 32) add2(m, __ldg(hx + t), __ldg(hy + t), px, py);
    const double sx = fm_warp_sum(m.sx), sy = fm_warp_sum(m.sy), sxx = fm_warp_sum(m.sxx),
                 syy = fm_warp_sum(m.syy), sxy = fm_warp_sum(m.sxy);

    fm_pair_rec r;
    r.mean_x = r.mean_y = r.var_x = r.var_y = r.cov_xy = r.d2_max = 0.0;
    r.n_anom = 0; r.first_idx = 0xFFFFu; r.verdict = FM_HEALTHY; r.status = 0; r.anom_mask = 0;
    uint32_t status = 0;
    if (nc == 0) status |= FM_ST_NO_CURRENT;
    bool ok = nh >= need;
    double det = 0.0;
    if (ok) {
      const double dn = (double)nh, dd = (double)(nh - g.ddof);
      r.mean_x = px + sx / dn; r.mean_y = py + sy / dn;
      r.var_x = fmax(sxx - sx * sx / dn, 0.0) / dd;
      r.var_y = fmax(syy - sy * sy / dn, 0.0) / dd;
      r.cov_xy = (sxy - sx * sy / dn) / dd;
      det = r.var_x * r.var_y - r.cov_xy * r.cov_xy;
      // no ellipse: a constant or (numerically) collinear history, 1 - rho^2 <= 1e-9
      if (!(det > 1e-9 * r.var_x * r.var_y)) { status |= FM_ST_DEGENERATE; ok = false; }
    } else {
      status |= FM_ST_HIST_INSUFFICIENT;
    }
    if (ok && nc > 0) {
      const uint32_t ty = b.metric_type ? b.metric_type[s] : 0u;
      const double k = P->types[ty < P->n_types ? ty : 0u].k;
      const double k2 = k * k;
      const float* cx = b.cur_x + s * b.cur_pitch;
      const float* cy = b.cur_y + s * b.cur_pitch;
      uint32_t cnt = 0, first = 0xFFFFFFFFu;
      uint64_t mask = 0;
      double dmax = 0.0;
      for (uint32_t i0 = 0; i0 < nc; i0 += 32) {
        const uint32_t i = i0 + lane;
        bool a = false;
        if (i < nc) {
          const double dx = (double)__ldg(cx + i) - r.mean_x, dy = (double)__ldg(cy + i) - r.mean_y;
          const double d2 = (r.var_y * dx * dx - 2.0 * r.cov_xy * dx * dy + r.var_x * dy * dy) / det;
          dmax = fmax(dmax, d2);
          a = d2 > k2;
        }
        const uint32_t bl = __ballot_sync(FM_FULL, a);
        if (bl) {
          if (first == 0xFFFFFFFFu) first = i0 + (uint32_t)__ffs((int)bl) - 1u;
          cnt += (uint32_t)__popc(bl);
          if (i0 < 64) mask |= (uint64_t)bl << i0;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dmax = fmax(dmax, __shfl_xor_sync(FM_FULL, dmax, o));
      r.d2_max = dmax;
      r.n_anom = cnt;
      r.first_idx = (uint16_t)(first == 0xFFFFFFFFu ? 0xFFFFu : (first < 0xFFFFu ? first : 0xFFFEu));
      r.anom_mask = mask;
      r.verdict = (cnt >= g.min_anomaly_points && cnt > 0) ? FM_UNHEALTHY : FM_HEALTHY;
    }
    if (status & (FM_ST_NO_CURRENT | FM_ST_HIST_INSUFFICIENT | FM_ST_DEGENERATE)) r.verdict = FM_UNKNOWN;
    r.status = (uint8_t)status;
    if (lane == 0) out[s] = r;
  }
}

}  // namespace

int fm_launch_pairs(const fm_pair_batch& b, const FmParamsDev* P, fm_pair_rec* out, int num_sms, cudaStream_t st) {
  if (!b.n_pairs) return 0;
  const uint64_t need = (b.n_pairs + 7) / 8;
  const int grid = (int)(need < (uint64_t)num_sms * 4 ? need : (uint64_t)num_sms * 4);
  k_score_pairs<<<grid, 256, 0, st>>>(b, P, out);
  return 1;
}
