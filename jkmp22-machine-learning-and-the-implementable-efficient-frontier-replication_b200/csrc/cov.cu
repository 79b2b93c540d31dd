// Barra / EWMA covariance estimator kernels ("Estimate Covariance Matrix.py").  All three
// are HBM/L2-bound streaming kernels: rows are read with the thread index on the contiguous factor axis, the
// small K x K algebra lives in shared memory.
//
//   pfml_xs_ols_daily   per-day cross-sectional OLS: coef = solve(X'X, X'y), residuals         :214-241
//   pfml_wcov_wcor      EWMA-weighted correlation (hl_cor) x variance (hl_var) per calc date   :297-335,
//                       General_functions.py:745-835 (R's cov.wt, center = TRUE, unbiased)
//   pfml_ewma_vol       per-stock EWMA volatility recurrence (the reference's numba kernel)    :345-386
#include "kernels.cuh"
#include "pfml_b200.h"

using namespace pfml;

namespace {

constexpr int KC = 32;   // factors padded (K <= 32)

// ---- a15: one CTA per day -------------------------------------------------------------------------
// Rows of day d are [day_ptr[d], day_ptr[d+1]).  X'X and X'y are accumulated from 64-row chunks staged in shared
// memory; the K x K system is solved by Cholesky.  A pivot that is zero relative to the largest diagonal entry
// (an industry without any stock that day gives an exactly zero row/column, the case in which the reference's
// np.linalg.solve raises and it falls back to pinv, :226-229) drops that factor: its coefficient is 0, which is
// what pinv returns for a zero row/column; status bit 0 records the fallback.
__global__ void __launch_bounds__(128)
xs_ols_kernel(int K, const double* __restrict__ X, const double* __restrict__ y, const long* __restrict__ day_ptr,
              double* __restrict__ coef, double* __restrict__ resid, int* __restrict__ status) {
  constexpr int CH = 64, S = KC + 1;
  __shared__ double A[KC][S];
  __shared__ double rhs[KC];
  __shared__ double Xc[CH][S];
  __shared__ double yc[CH];
  __shared__ double beta[KC];
  __shared__ int dropped[KC];
  const int d = blockIdx.x, tid = threadIdx.x;
  const long r0 = day_ptr[d], r1 = day_ptr[d + 1];
  double acc[8];
#pragma unroll
  for (int q = 0; q < 8; q++) acc[q] = 0.0;
  double accy = 0.0;
  for (long i0 = r0; i0 < r1; i0 += CH) {
    const int nr = (int)((r1 - i0) < CH ? (r1 - i0) : CH);
    __syncthreads();
    for (int e = tid; e < CH * K; e += 128) {
      const int r = e / K, c = e % K;
      Xc[r][c] = r < nr ? X[(i0 + r) * K + c] : 0.0;
    }
    if (tid < CH) yc[tid] = tid < nr ? y[i0 + tid] : 0.0;
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 8; q++) {
      const int e = tid + 128 * q, a = e / KC, c = e % KC;
      if (a < K && c <= a) {
        double s = 0.0;
        for (int i = 0; i < nr; i++) s += Xc[i][a] * Xc[i][c];
        acc[q] += s;
      }
    }
    if (tid < K) {
      double s = 0.0;
      for (int i = 0; i < nr; i++) s += Xc[i][tid] * yc[i];
      accy += s;
    }
  }
#pragma unroll
  for (int q = 0; q < 8; q++) {
    const int e = tid + 128 * q, a = e / KC, c = e % KC;
    if (a < K && c <= a) { A[a][c] = acc[q]; A[c][a] = acc[q]; }
  }
  if (tid < K) { rhs[tid] = accy; dropped[tid] = 0; }
  __syncthreads();
  if (tid < 32) {
    const int r = tid;
    double dmax = 0.0;
    for (int c = 0; c < K; c++) dmax = fmax(dmax, A[c][c]);
    bool any = false;
    for (int c = 0; c < K; c++) {
      const double piv = A[c][c];
      const bool ok = piv > 1e-13 * dmax;
      const double dd = ok ? sqrt(piv) : 1.0;
      __syncwarp();
      if (r == c) { A[c][c] = dd; dropped[c] = ok ? 0 : 1; }
      if (r > c && r < K) A[r][c] = ok ? A[r][c] / dd : 0.0;
      any |= !ok;
      __syncwarp();
      if (r > c && r < K)
        for (int cc = c + 1; cc <= r; cc++) A[r][cc] -= A[r][c] * A[cc][c];
      __syncwarp();
    }
    if (r == 0) {
      // forward then backward substitution (serial: K = 25)
      for (int i = 0; i < K; i++) {
        double s = dropped[i] ? 0.0 : rhs[i];
        for (int a = 0; a < i; a++) s -= A[i][a] * beta[a];
        beta[i] = dropped[i] ? 0.0 : s / A[i][i];
      }
      for (int i = K - 1; i >= 0; i--) {
        double s = beta[i];
        for (int a = i + 1; a < K; a++) s -= A[a][i] * beta[a];
        beta[i] = dropped[i] ? 0.0 : s / A[i][i];
      }
      if (any && status) atomicOr(status + d, 1);
    }
  }
  __syncthreads();
  if (tid < K) coef[(long)d * K + tid] = beta[tid];
  for (long i = r0 + tid; i < r1; i += 128) {
    const double* xr = X + i * K;
    double s = 0.0;
    for (int c = 0; c < K; c++) s += xr[c] * beta[c];
    resid[i] = y[i] - s;
  }
}

// ---- a16: one CTA per calc date -------------------------------------------------------------------
__global__ void __launch_bounds__(256)
wcov_wcor_kernel(int K, int obs, const double* __restrict__ F, const int* __restrict__ end_idx, const double* __restrict__ w_cor,
                 const double* __restrict__ w_var, double scale, double* __restrict__ out) {
  constexpr int S = KC + 1;
  __shared__ double mu_c[KC], mu_v[KC], cov[KC][S], var_v[KC];
  __shared__ double red[4][256 / 32];
  __shared__ double sums[4];     // sum w_cor, sum w_var, sum wn_cor^2, sum wn_var^2
  const int cd = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int e = end_idx[cd];
  const int t = (e + 1) < obs ? (e + 1) : obs;
  const long first = (long)e + 1 - t;
  const double* wc = w_cor + (obs - t);
  const double* wv = w_var + (obs - t);
  // weight sums (sequential-in-thread, then a fixed tree: deterministic)
  double s[4] = {0, 0, 0, 0};
  for (int i = tid; i < t; i += 256) { s[0] += wc[i]; s[1] += wv[i]; }
  for (int q = 0; q < 2; q++) {
    double v = s[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[q][warp] = v;
  }
  __syncthreads();
  if (tid < 2) { double v = 0; for (int k = 0; k < 8; k++) v += red[tid][k]; sums[tid] = v; }
  __syncthreads();
  const double swc = sums[0], swv = sums[1];
  for (int i = tid; i < t; i += 256) { const double a = wc[i] / swc, b = wv[i] / swv; s[2] += a * a; s[3] += b * b; }
  for (int q = 2; q < 4; q++) {
    double v = s[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[q][warp] = v;
  }
  __syncthreads();
  if (tid >= 2 && tid < 4) { double v = 0; for (int k = 0; k < 8; k++) v += red[tid][k]; sums[tid] = v; }
  // weighted means: warp w handles factors w, w+8, ...
  for (int c = warp; c < K; c += 8) {
    double a = 0, b = 0, na = 0, nb = 0;
    for (int i = lane; i < t; i += 32) {
      const double x = F[(first + i) * K + c];
      const double wa = wc[i] / swc, wb = wv[i] / swv;
      a += wa * x; b += wb * x; na += wa; nb += wb;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o);
      na += __shfl_xor_sync(0xffffffffu, na, o); nb += __shfl_xor_sync(0xffffffffu, nb, o);
    }
    if (lane == 0) { mu_c[c] = a / na; mu_v[c] = b / nb; }      // np.average divides by the weight sum
  }
  __syncthreads();
  const double den_c = 1.0 - sums[2], den_v = 1.0 - sums[3];
  // centred, weighted products: pair (a, c<=a) per warp-iteration
  for (int pr = warp; pr < K * (K + 1) / 2 + K; pr += 8) {
    double acc = 0.0;
    if (pr < K * (K + 1) / 2) {
      int a = 0, rem = pr;
      while (rem > a) { rem -= a + 1; a++; }
      const int c = rem;
      for (int i = lane; i < t; i += 32) {
        const double wn = wc[i] / swc;
        const double* row = F + (first + i) * K;
        acc += wn * (row[a] - mu_c[a]) * (row[c] - mu_c[c]);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) { cov[a][c] = acc / den_c; cov[c][a] = acc / den_c; }
    } else {
      const int a = pr - K * (K + 1) / 2;
      for (int i = lane; i < t; i += 32) {
        const double wn = wv[i] / swv;
        const double xc = F[(first + i) * K + a] - mu_v[a];
        acc += wn * xc * xc;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) var_v[a] = acc / den_v;
    }
  }
  __syncthreads();
  // cov_est = sd (cor) sd with unit-diagonal cor, times `scale` (x 21 at :476)
  for (int e2 = tid; e2 < K * K; e2 += 256) {
    const int a = e2 / K, c = e2 % K;
    const double cor = (a == c) ? 1.0 : cov[a][c] / (sqrt(cov[a][a]) * sqrt(cov[c][c]));
    out[(long)cd * K * K + e2] = (sqrt(var_v[a]) * cor * sqrt(var_v[c])) * scale;
  }
}

// ---- a17: one thread per stock series ---------------------------------------------------------------
__global__ void ewma_vol_kernel(int n_series, const double* __restrict__ x, const long* __restrict__ ptr, double lam, int start,
                                double* __restrict__ vol) {
  const int sidx = blockIdx.x * blockDim.x + threadIdx.x;
  if (sidx >= n_series) return;
  const long b0 = ptr[sidx], n = ptr[sidx + 1] - b0;
  const double* xs = x + b0;
  double* vs = vol + b0;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  double ss = 0.0;
  int cnt = 0;
  if (n > start) {
    for (int i = 0; i < start; i++) {
      const double v = xs[i];
      if (!isnan(v)) { ss = __dadd_rn(ss, __dmul_rn(v, v)); cnt++; }
    }
  }
  if (n <= start || cnt <= 1) {
    for (long i = 0; i < n; i++) vs[i] = nan;
    return;
  }
  for (int i = 0; i < start; i++) vs[i] = nan;
  double var = ss / (double)(cnt - 1);
  vs[start] = sqrt(var);
  const double oml = 1.0 - lam;
  for (long i = start + 1; i < n; i++) {
    const double xp = xs[i - 1];
    if (!isnan(xp)) var = __dadd_rn(__dmul_rn(lam, var), __dmul_rn(oml, __dmul_rn(xp, xp)));   // no FMA: numba does not fuse
    vs[i] = sqrt(var);
  }
}

}  // namespace

extern "C" {

int pfml_xs_ols_daily(const double* X, const double* y, const long* day_ptr, int n_days, int K, double* coef, double* resid,
                      int* status, void* stream) {
  if (!X || !y || !day_ptr || !coef || !resid || n_days <= 0 || K <= 0 || K > KC) {
    set_error("pfml_xs_ols_daily: null/empty argument or K > %d", KC);
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(T_COV, s);
  if (status) PFML_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(int) * n_days, s));
  xs_ols_kernel<<<n_days, 128, 0, s>>>(K, X, y, day_ptr, coef, resid, status);
  PFML_CHECK_LAUNCH();
  return 0;
}

int pfml_wcov_wcor(const double* fct_ret, const int* end_idx, int n_calc, int K, int obs, const double* w_cor,
                   const double* w_var, double scale, double* cov_out, void* stream) {
  if (!fct_ret || !end_idx || !w_cor || !w_var || !cov_out || n_calc <= 0 || K <= 0 || K > KC || obs <= 0) {
    set_error("pfml_wcov_wcor: null/empty argument or K > %d", KC);
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(T_COV, s);
  wcov_wcor_kernel<<<n_calc, 256, 0, s>>>(K, obs, fct_ret, end_idx, w_cor, w_var, scale, cov_out);
  PFML_CHECK_LAUNCH();
  return 0;
}

int pfml_ewma_vol(const double* x, const long* series_ptr, int n_series, double lam, int start, double* vol, void* stream) {
  if (!x || !series_ptr || !vol || n_series <= 0 || start < 1) {
    set_error("pfml_ewma_vol: null/empty argument");
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(T_COV, s);
  ewma_vol_kernel<<<ceil_div(n_series, 64), 64, 0, s>>>(n_series, x, series_ptr, lam, start, vol);
  PFML_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"
