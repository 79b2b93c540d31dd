// Next rows of the path: validation utility, aim portfolio, trading-speed matrices for the
// out-of-sample dates and the back-test recursion with its statistics.
//
//   pfml_validation_utility  obj = r~' b - 1/2 b' denom b per (date, lambda)             PFML_hp_reals.py:90-102
//   pfml_aim_weights         w_aim = signal_t[d][feat] @ beta                            PFML_aim_fun.py:150-160
//   pfml_mfunc_dates         m = m_func(...) for a batch of dates from the factor form    PFML_best_hps.py:186-190
//   pfml_backtest            w = m w_start + (I - m) w_aim, drift, statistics            PFML_best_hps.py:168-239
#include "kernels.cuh"
#include "pfml_b200.h"

namespace pfml {
int mtilde_factor_for_dates(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                            double* m_out, int* status, void* workspace, size_t workspace_bytes, cudaStream_t s);   // pipeline.cu
size_t mtilde_factor_workspace_bytes(const pfml_panel_t* panel, int B, int n_max);
}

using namespace pfml;

namespace {

__device__ __forceinline__ int feat_index(int k, int p, int p_max) {
  const int h = p >> 1;
  return k == 0 ? 0 : (k <= h ? k : (p_max >> 1) + (k - h));
}

__device__ __forceinline__ double block_sum(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double s = 0.0;
  for (int k = 0; k < nw; k++) s += red[k];     // fixed order: deterministic
  return s;
}

// one CTA per (lambda, date)
__global__ void __launch_bounds__(256)
utility_kernel(int P, int p, int n_l, const double* __restrict__ r_tilde, const double* __restrict__ denom,
               const int* __restrict__ date_year, const double* __restrict__ beta, double* __restrict__ obj) {
  extern __shared__ double sb[];           // beta [q]
  __shared__ double red[8];
  const int l = blockIdx.x, b = blockIdx.y, q = p + 1, p_max = P - 1;
  const int y = date_year[b];
  if (y < 0) {
    if (threadIdx.x == 0) obj[(long)b * n_l + l] = __longlong_as_double(0x7ff8000000000000LL);
    return;
  }
  const double* bt = beta + ((long)y * n_l + l) * q;
  for (int i = threadIdx.x; i < q; i += blockDim.x) sb[i] = bt[i];
  __syncthreads();
  const double* D = denom + (long)b * P * P;
  const double* rt = r_tilde + (long)b * P;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  double lin = 0.0, quad = 0.0;
  for (int i = warp; i < q; i += nw) {
    const int fi = feat_index(i, p, p_max);
    const double* row = D + (long)fi * P;
    double s = 0.0;
    for (int j = lane; j < q; j += 32) s += row[feat_index(j, p, p_max)] * sb[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) { quad += sb[i] * s; lin += rt[fi] * sb[i]; }
  }
  const double tl = block_sum(lin, red);
  const double tq = block_sum(quad, red);
  if (threadIdx.x == 0) obj[(long)b * n_l + l] = tl - 0.5 * tq;
}

// w[b][i] = sum_j S[b][i][j] beta_full[b][j]; one warp per row
__global__ void aim_kernel(int NP, int P, const double* __restrict__ S, const int* __restrict__ n_sel, const double* __restrict__ beta_full,
                           double* __restrict__ w_aim) {
  const int b = blockIdx.y, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= NP) return;
  double s = 0.0;
  if (warp < n_sel[b]) {
    const double* row = S + ((long)b * NP + warp) * P;
    const double* bt = beta_full + (long)b * P;
    for (int j = lane; j < P; j += 32) s += row[j] * bt[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  }
  if (lane == 0) w_aim[(long)b * NP + warp] = s;
}

// The back-test is a recursion over dates; one CTA walks the dates in order.  `carry` [N] holds, per stock of the
// panel, next month's starting weight (w (1 + tr_ld1) / (1 + mu_ld1)); stocks that were not held start at 0.
__global__ void __launch_bounds__(1024)
backtest_kernel(int B, int NP, int N, const double* __restrict__ m, const double* __restrict__ w_aim, const int* __restrict__ date_t,
                const int* __restrict__ sel, const int* __restrict__ n_sel, const double* __restrict__ me,
                const double* __restrict__ tr_ld1, const double* __restrict__ mu_ld1, const double* __restrict__ ret_ld1,
                const double* __restrict__ lam, const double* __restrict__ wealth, double* __restrict__ carry,
                double* __restrict__ w_start_out, double* __restrict__ w_out, double* __restrict__ stats) {
  extern __shared__ double sh[];           // ws [NP], wa [NP], w [NP]
  __shared__ double red[32];
  double* ws = sh;
  double* wa = sh + NP;
  double* wn = sh + 2 * NP;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
  for (int b = 0; b < B; b++) {
    const int t = date_t[b], n = n_sel[b];
    const int* sl = sel + (long)b * NP;
    const long row0 = (long)t * N;
    if (b == 0) {
      double part = 0.0;
      for (int i = tid; i < n; i += blockDim.x) part += me[row0 + sl[i]];
      const double tot = block_sum(part, red);
      for (int i = tid; i < n; i += blockDim.x) ws[i] = me[row0 + sl[i]] / tot;        // value weights
    } else {
      for (int i = tid; i < n; i += blockDim.x) ws[i] = carry[sl[i]];
    }
    for (int i = tid; i < n; i += blockDim.x) wa[i] = w_aim[(long)b * NP + i];
    __syncthreads();
    const double* M = m + (long)b * NP * NP;
    for (int i = warp; i < n; i += nw) {
      const double* mr = M + (long)i * NP;
      double s1 = 0.0, s2 = 0.0;
      for (int j = lane; j < n; j += 32) {
        const double mij = mr[j];
        s1 += mij * ws[j];
        s2 += ((i == j ? 1.0 : 0.0) - mij) * wa[j];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
      if (lane == 0) wn[i] = s1 + s2;
    }
    __syncthreads();
    // statistics (compute_stats, PFML_best_hps.py:220-239) and outputs
    double inv = 0, sh_ = 0, to = 0, r = 0, tc = 0;
    for (int i = tid; i < n; i += blockDim.x) {
      const double w = wn[i], d = w - ws[i];
      inv += fabs(w);
      if (w < 0) sh_ += fabs(w);
      to += fabs(d);
      r += w * ret_ld1[row0 + sl[i]];
      tc += lam[row0 + sl[i]] * (d * d);
      w_start_out[(long)b * NP + i] = ws[i];
      w_out[(long)b * NP + i] = w;
    }
    for (int i = n + tid; i < NP; i += blockDim.x) { w_start_out[(long)b * NP + i] = 0.0; w_out[(long)b * NP + i] = 0.0; }
    const double v0 = block_sum(inv, red), v1 = block_sum(sh_, red), v2 = block_sum(to, red), v3 = block_sum(r, red),
                 v4 = block_sum(tc, red);
    if (tid == 0) {
      double* o = stats + (long)b * 5;
      o[0] = v0; o[1] = v1; o[2] = v2; o[3] = v3; o[4] = (wealth[t] / 2.0) * v4;
    }
    // next month's starting weights
    for (int i = tid; i < N; i += blockDim.x) carry[i] = 0.0;
    __syncthreads();
    const double mu1 = mu_ld1[t];
    // a NaN drifted weight (missing tr_ld1 / mu_ld1) is not carried: the reference writes w_opt_ld1 into next month's
    // w_start and then sets every NaN w_start to 0 (PFML_best_hps.py:204-210)
    for (int i = tid; i < n; i += blockDim.x) {
      const double v = wn[i] * (1.0 + tr_ld1[row0 + sl[i]]) / (1.0 + mu1);
      carry[sl[i]] = isnan(v) ? 0.0 : v;
    }
    __syncthreads();
  }
}

}  // namespace

extern "C" {

int pfml_validation_utility(const double* r_tilde, const double* denom, int B, int p_max, const int* date_year, const double* beta,
                            int p, int n_l, double* obj, void* stream) {
  if (!r_tilde || !denom || !date_year || !beta || !obj || B <= 0 || p < 2 || (p & 1) || p > p_max || n_l <= 0) {
    set_error("pfml_validation_utility: bad argument");
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  utility_kernel<<<dim3(n_l, B), 256, (p + 1) * sizeof(double), s>>>(p_max + 1, p, n_l, r_tilde, denom, date_year, beta, obj);
  PFML_CHECK_LAUNCH();
  return 0;
}

int pfml_aim_weights(const double* signal_t, const int* n_sel, int B, int NP, int P, const double* beta_full, double* w_aim,
                     void* stream) {
  if (!signal_t || !n_sel || !beta_full || !w_aim || B <= 0 || NP <= 0 || P <= 0) {
    set_error("pfml_aim_weights: bad argument");
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  aim_kernel<<<dim3(ceil_div((long)NP * 32, 256), B), 256, 0, s>>>(NP, P, signal_t, n_sel, beta_full, w_aim);
  PFML_CHECK_LAUNCH();
  return 0;
}

size_t pfml_mfunc_dates_workspace_bytes(const pfml_panel_t* panel, int B, int n_max) {
  return mtilde_factor_workspace_bytes(panel, B, n_max);
}

int pfml_mfunc_dates(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel, double* m_out,
                     int* status, void* workspace, size_t workspace_bytes, void* stream) {
  return mtilde_factor_for_dates(panel, B, n_max, date_t, sel, n_sel, m_out, status, workspace, workspace_bytes, (cudaStream_t)stream);
}

int pfml_backtest(const pfml_panel_t* panel, const double* me, const double* tr_ld1, const double* mu_ld1, int B, int n_max,
                  const int* date_t, const int* sel, const int* n_sel, const double* m, const double* w_aim, double* carry,
                  double* w_start, double* w, double* stats, void* stream) {
  if (!panel || !me || !tr_ld1 || !mu_ld1 || !date_t || !sel || !n_sel || !m || !w_aim || !carry || !w_start || !w || !stats ||
      B <= 0 || n_max <= 0) {
    set_error("pfml_backtest: null/empty argument");
    return -1;
  }
  const int NP = pfml_padded_n(n_max);
  const size_t smem = 3 * (size_t)NP * sizeof(double);
  if (smem > 200 * 1024) {
    set_error("pfml_backtest: universe too large for the shared-memory weight vectors (n_max = %d)", n_max);
    return -2;
  }
  cudaStream_t s = (cudaStream_t)stream;
  PFML_CONFIG_ONCE(backtest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  backtest_kernel<<<1, 1024, smem, s>>>(B, NP, panel->N, m, w_aim, date_t, sel, n_sel, me, tr_ld1, mu_ld1, panel->ret_ld1,
                                        panel->lam, panel->wealth, carry, w_start, w, stats);
  PFML_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"
