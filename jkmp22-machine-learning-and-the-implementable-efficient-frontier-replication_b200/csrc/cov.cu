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
// Rows of day d are [day_ptr[d], day_ptr[d+1]).  The day's rows [x_1 .. x_K, y] are streamed through shared memory in
// chunks of 128 rows (coalesced flat copies of X, row stride padded to 28 doubles so that four consecutive columns are
// one 16-byte-aligned pair of loads).  The augmented Gram matrix [X y]'[X y] is formed in 4 x 4 register tiles: the 28
// lower-triangle tiles of the 7 x 7 tile grid go to 28 lanes of every warp, warp w takes the rows i = w (mod 4) of the
// chunk -- four 16-byte loads feed sixteen FMAs, where one accumulator per thread needed two loads per FMA and was bound
// by shared-memory bandwidth.  The four partial matrices are added in warp order (deterministic).  The K x K system is
// solved by a warp: Cholesky, then forward / backward substitution with one lane per unknown.  A pivot that is zero
// relative to the largest diagonal entry (an industry without any stock that day gives an exactly zero row / column, the
// case in which the reference's np.linalg.solve raises and it falls back to pinv, :226-229) drops that factor: its
// coefficient is 0, which is what pinv returns for a zero row / column; status bit 0 records the fallback.  The residual
// pass streams the rows a second time (L2 hits: a day is 100 KB) with one thread per row.
constexpr int KA = 28;   // padded width of an augmented row [x (K <= 27) | y | 0..]

__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(128)
xs_ols_kernel(int K, const double* __restrict__ X, const double* __restrict__ y, const long* __restrict__ day_ptr,
              double* __restrict__ coef, double* __restrict__ resid, int* __restrict__ status) {
  constexpr int CH = 64, S = KC + 1;
  __shared__ __align__(16) double Xc[2][CH * KA];  // two chunks of augmented rows: the next one loads while this one is used
  __shared__ double A[KC][S];
  __shared__ double rhs[KC], beta[KC];
  __shared__ int dropped[KC];
  const int d = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long r0 = day_ptr[d], r1 = day_ptr[d + 1];
  const int n_chunks = (int)((r1 - r0 + CH - 1) / CH);
  // lane -> lower-triangle tile (ta >= tc) of the 7 x 7 grid
  int ta = 0, tc = 0;
  {
    int l = lane;
    while (l > ta) { l -= ta + 1; ta++; }
    tc = l;
  }
  const bool active = lane < 28;
  double acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) acc[u][v] = 0.0;
  const float invK = 1.0f / (float)K;
  // asynchronous copies (8 bytes each, many in flight per thread, no registers held): chunk `ck` of the day into buffer `bf`
  auto load_chunk = [&](int ck, int bf, bool with_y) {
    const long i0 = r0 + (long)ck * CH;
    const int nr = (int)((r1 - i0) < CH ? (r1 - i0) : CH);
    const double* src = X + i0 * K;
    for (int e = tid; e < nr * K; e += 128) {
      const int r = (int)(((float)e + 0.5f) * invK), c = e - r * K;
      cp_async8(&Xc[bf][r * KA + c], src + e);
    }
    if (with_y && tid < nr) cp_async8(&Xc[bf][tid * KA + K], y + i0 + tid);
    cp_async_commit();
  };
  for (int e = tid; e < 2 * CH * KA; e += 128) Xc[0][e] = 0.0;   // the padding columns stay zero (the copies never touch them)
  __syncthreads();
  if (n_chunks > 0) load_chunk(0, 0, true);
  for (int ck = 0; ck < n_chunks; ck++) {
    const int bf = ck & 1;
    if (ck + 1 < n_chunks) { load_chunk(ck + 1, bf ^ 1, true); cp_async_wait<1>(); } else cp_async_wait<0>();
    __syncthreads();
    const long i0 = r0 + (long)ck * CH;
    const int nr = (int)((r1 - i0) < CH ? (r1 - i0) : CH);
    if (active) {
      for (int i = warp; i < nr; i += 4) {
        const double2* pa = reinterpret_cast<const double2*>(&Xc[bf][i * KA + 4 * ta]);
        const double2* pc = reinterpret_cast<const double2*>(&Xc[bf][i * KA + 4 * tc]);
        const double2 a01 = pa[0], a23 = pa[1], c01 = pc[0], c23 = pc[1];
        const double av[4] = {a01.x, a01.y, a23.x, a23.y}, cv[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
          for (int v = 0; v < 4; v++) acc[u][v] += av[u] * cv[v];
      }
    }
    __syncthreads();                                        // buffer bf may be overwritten by the load issued next iteration
  }
  // the residual pass re-reads the day: start its first chunk now, under the solve (the Gram partials go to buffer 1)
  static_assert(4 * KA * (KA + 1) <= 2 * CH * KA, "partial Gram matrices fit in the chunk buffers");
  double* Gp = &Xc[0][0];                                   // [4][KA][KA + 1] over both buffers
  auto Gat = [&](int w, int r, int c) -> double& { return Gp[(w * KA + r) * (KA + 1) + c]; };
  if (active) {
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
      for (int v = 0; v < 4; v++) Gat(warp, 4 * ta + u, 4 * tc + v) = acc[u][v];
  }
  if (tid < KC) dropped[tid] = 0;
  __syncthreads();
  for (int e = tid; e < (K + 1) * KA; e += 128) {          // rows 0 .. K of the augmented matrix, columns <= row
    const int a = e / KA, c = e % KA;
    if (c <= a && c < K) {
      const double v = ((Gat(0, a, c) + Gat(1, a, c)) + Gat(2, a, c)) + Gat(3, a, c);
      if (a < K) { A[a][c] = v; A[c][a] = v; }
      else rhs[c] = v;                                       // row K of [X y]'[X y] is X'y
    }
  }
  __syncthreads();
  for (int e = tid; e < 2 * CH * KA; e += 128) Xc[0][e] = 0.0;
  __syncthreads();
  if (n_chunks > 0) load_chunk(0, 0, false);
  if (tid < 32) {
    const int r = lane;
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
    // forward substitution L z = rhs, then backward L' beta = z: lane r owns unknown r
    const bool mine = r < K;
    const bool drop = mine && dropped[r];
    double sv = (mine && !drop) ? rhs[r] : 0.0;
    double z = 0.0;
    for (int i = 0; i < K; i++) {
      const double zi = __shfl_sync(0xffffffffu, (r == i) ? (drop ? 0.0 : sv / A[r][r]) : 0.0, i);
      if (r == i) z = zi;
      if (mine && r > i) sv -= A[r][i] * zi;
    }
    sv = z;
    double bt = 0.0;
    for (int i = K - 1; i >= 0; i--) {
      const double bi = __shfl_sync(0xffffffffu, (r == i) ? (drop ? 0.0 : sv / A[r][r]) : 0.0, i);
      if (r == i) bt = bi;
      if (mine && r < i) sv -= A[i][r] * bi;
    }
    if (mine) beta[r] = bt;
    if (r == 0 && any && status) atomicOr(status + d, 1);
  }
  __syncthreads();
  if (tid < K) coef[(long)d * K + tid] = beta[tid];
  for (int ck = 0; ck < n_chunks; ck++) {
    const int bf = ck & 1;
    if (ck + 1 < n_chunks) { load_chunk(ck + 1, bf ^ 1, false); cp_async_wait<1>(); } else cp_async_wait<0>();
    __syncthreads();
    const long i0 = r0 + (long)ck * CH;
    const int nr = (int)((r1 - i0) < CH ? (r1 - i0) : CH);
    if (tid < nr) {
      const double* xr = &Xc[bf][tid * KA];
      double sacc = 0.0;
      for (int c = 0; c < K; c++) sacc += xr[c] * beta[c];
      resid[i0 + tid] = y[i0 + tid] - sacc;
    }
    __syncthreads();
  }
}

// ---- a16: one CTA per calc date -------------------------------------------------------------------
// weighted_cor_wt (weights w_cor) and the diagonal of weighted_cov_wt (weights w_var) over the window of the calc date,
// General_functions.py:745-835, then sd cor sd ("Estimate Covariance Matrix.py":316-320).  The normalised weights
// w_i / sum(w) live in shared memory (2 x obs doubles, dynamic).  The window (L2-resident: the whole D x K matrix is
// 1 MB) is streamed twice through 128-row chunks in shared memory, row stride padded to 28: pass 1 the weighted means,
// pass 2 the centred weighted Gram matrix in 4 x 4 register tiles (28 lower-triangle tiles on 28 lanes of each of the
// 8 warps, warp w takes rows i = w (mod 8): four 16-byte loads feed 16 FMAs) plus the w_var-weighted variances on
// lanes 0 .. K-1.  Partial results of the 8 warps are added in warp order (deterministic).
__global__ void __launch_bounds__(256)
wcov_wcor_kernel(int K, int obs, const double* __restrict__ F, const int* __restrict__ end_idx, const double* __restrict__ w_cor,
                 const double* __restrict__ w_var, double scale, double* __restrict__ out) {
  constexpr int S = KC + 1, CH = 128;
  __shared__ __align__(16) double Xc[CH * KA];
  __shared__ double mu_c[KA], mu_v[KA], cov[KC][S], var_v[KC];
  __shared__ double red[4][8];
  __shared__ double part[8][KA][2];     // per-warp partial sums of pass 1, then of the variances
  __shared__ double sums[4];            // sum w_cor, sum w_var, sum wn_cor^2, sum wn_var^2
  extern __shared__ double s_wn[];      // [obs] normalised w_cor, [obs] normalised w_var, then the 8 partial Gram matrices
  double* wnc = s_wn;
  double* wnv = s_wn + obs;
  double* Gp = s_wn + 2 * obs;          // [8][KA][KA + 1]
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
  for (int i = tid; i < t; i += 256) { const double a = wc[i] / swc, b = wv[i] / swv; wnc[i] = a; wnv[i] = b; s[2] += a * a; s[3] += b * b; }
  for (int q = 2; q < 4; q++) {
    double v = s[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[q][warp] = v;
  }
  __syncthreads();
  if (tid >= 2 && tid < 4) { double v = 0; for (int k = 0; k < 8; k++) v += red[tid][k]; sums[tid] = v; }
  for (int i = tid; i < CH * KA; i += 256) Xc[i] = 0.0;      // padding columns stay zero
  __syncthreads();
  const float invK = 1.0f / (float)K;
  auto stage = [&](int i0, int nr) {                         // rows i0 .. i0 + nr - 1 of the window -> Xc (coalesced flat reads)
    const double* src = F + (first + i0) * K;
    for (int q = tid; q < nr * K; q += 256) {
      const int r = (int)(((float)q + 0.5f) * invK), c = q - r * K;
      Xc[r * KA + c] = src[q];
    }
  };
  // ---- pass 1: weighted means (np.average divides by the weight sum) -----------------------------------------
  double a1 = 0, b1 = 0, na = 0, nb = 0;
  for (int i0 = 0; i0 < t; i0 += CH) {
    const int nr = t - i0 < CH ? t - i0 : CH;
    __syncthreads();
    stage(i0, nr);
    __syncthreads();
    if (lane < KA)
      for (int i = warp; i < nr; i += 8) {
        const double x = Xc[i * KA + lane], wa = wnc[i0 + i], wb = wnv[i0 + i];
        a1 += wa * x; b1 += wb * x; na += wa; nb += wb;
      }
  }
  if (lane < KA) { part[warp][lane][0] = a1; part[warp][lane][1] = b1; }
  if (lane == 0) { red[0][warp] = na; red[1][warp] = nb; }
  __syncthreads();
  if (tid < KA) {
    double a = 0, b = 0, sa = 0, sb = 0;
    for (int w = 0; w < 8; w++) { a += part[w][tid][0]; b += part[w][tid][1]; sa += red[0][w]; sb += red[1][w]; }
    mu_c[tid] = a / sa; mu_v[tid] = b / sb;
  }
  __syncthreads();
  const double den_c = 1.0 - sums[2], den_v = 1.0 - sums[3];
  // ---- pass 2: centred weighted Gram matrix (w_cor) in 4 x 4 tiles, variances (w_var) on lanes < K -----------------
  int ta = 0, tc = 0;
  {
    int l = lane;
    while (l > ta) { l -= ta + 1; ta++; }
    tc = l;
  }
  const bool active = lane < 28;
  double acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) acc[u][v] = 0.0;
  double ma[4], mc[4];
#pragma unroll
  for (int u = 0; u < 4; u++) { ma[u] = active ? mu_c[4 * ta + u] : 0.0; mc[u] = active ? mu_c[4 * tc + u] : 0.0; }
  const double mv = lane < K ? mu_v[lane] : 0.0;
  double vacc = 0.0;
  for (int i0 = 0; i0 < t; i0 += CH) {
    const int nr = t - i0 < CH ? t - i0 : CH;
    __syncthreads();
    stage(i0, nr);
    __syncthreads();
    for (int i = warp; i < nr; i += 8) {
      const double wn = wnc[i0 + i];
      if (active) {
        const double2* pa = reinterpret_cast<const double2*>(Xc + i * KA + 4 * ta);
        const double2* pc = reinterpret_cast<const double2*>(Xc + i * KA + 4 * tc);
        const double2 a01 = pa[0], a23 = pa[1], c01 = pc[0], c23 = pc[1];
        const double av[4] = {wn * (a01.x - ma[0]), wn * (a01.y - ma[1]), wn * (a23.x - ma[2]), wn * (a23.y - ma[3])};
        const double cv[4] = {c01.x - mc[0], c01.y - mc[1], c23.x - mc[2], c23.y - mc[3]};
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
          for (int v = 0; v < 4; v++) acc[u][v] += av[u] * cv[v];
      }
      if (lane < K) {
        const double xc = Xc[i * KA + lane] - mv;
        vacc += wnv[i0 + i] * xc * xc;
      }
    }
  }
  auto Gat = [&](int w, int r, int c) -> double& { return Gp[(w * KA + r) * (KA + 1) + c]; };
  if (active) {
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
      for (int v = 0; v < 4; v++) Gat(warp, 4 * ta + u, 4 * tc + v) = acc[u][v];
  }
  if (lane < KA) part[warp][lane][0] = vacc;
  __syncthreads();
  for (int q = tid; q < K * KA; q += 256) {
    const int a = q / KA, c = q % KA;
    if (c <= a) {
      double v = 0.0;
      for (int w = 0; w < 8; w++) v += Gat(w, a, c);
      cov[a][c] = v / den_c; cov[c][a] = v / den_c;
    }
  }
  if (tid < K) {
    double v = 0.0;
    for (int w = 0; w < 8; w++) v += part[w][tid][0];
    var_v[tid] = v / den_v;
  }
  __syncthreads();
  // cov_est = sd (cor) sd with unit-diagonal cor, times `scale` (x 21 at :476)
  for (int e2 = tid; e2 < K * K; e2 += 256) {
    const int a = e2 / K, c = e2 % K;
    const double cor = (a == c) ? 1.0 : cov[a][c] / (sqrt(cov[a][a]) * sqrt(cov[c][c]));
    out[(long)cd * K * K + e2] = (sqrt(var_v[a]) * cor * sqrt(var_v[c])) * scale;
  }
}

// ---- a17: one warp per stock series -------------------------------------------------------------------
// The recurrence var_i = lam var_{i-1} + (1 - lam) x_{i-1}^2 (NaN x: carry) is a 5 000-step dependent chain per series and
// is evaluated in the reference's order with un-fused multiply and add (numba does not contract), so it cannot be split
// across lanes without changing the rounding.  What CAN be parallel is everything around it: a warp loads 32 consecutive
// observations with one coalesced request (the next 32 are already in flight), every lane squares and scales its own
// observation, then all lanes step through the 32 updates together -- operand j arrives by shuffle, lane j keeps var_j --
// and the 32 square roots are taken and stored coalesced.  The chain costs two dependent FP64 operations per step;
// nothing else is on the critical path.
__global__ void __launch_bounds__(128)
ewma_vol_kernel(int n_series, const double* __restrict__ x, const long* __restrict__ ptr, double lam, int start, double* __restrict__ vol) {
  const int sidx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (sidx >= n_series) return;
  const long b0 = ptr[sidx], n = ptr[sidx + 1] - b0;
  const double* xs = x + b0;
  double* vs = vol + b0;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  double ss = 0.0;
  int cnt = 0;
  if (n > start) {
    // initial variance: squares of the first `start` observations summed in order (every lane computes the same sum)
    for (int i0 = 0; i0 < start; i0 += 32) {
      const double v = (i0 + lane < start) ? xs[i0 + lane] : nan;
      const int m = start - i0 < 32 ? start - i0 : 32;
      for (int j = 0; j < m; j++) {
        const double vj = __shfl_sync(0xffffffffu, v, j);
        if (!isnan(vj)) { ss = __dadd_rn(ss, __dmul_rn(vj, vj)); cnt++; }
      }
    }
  }
  if (n <= start || cnt <= 1) {
    for (long i = lane; i < n; i += 32) vs[i] = nan;
    return;
  }
  for (int i = lane; i < start; i += 32) vs[i] = nan;
  double var = ss / (double)(cnt - 1);
  if (lane == 0) vs[start] = sqrt(var);
  const double oml = 1.0 - lam;
  // element i (start < i < n) needs x[i - 1]; blocks of 32 elements, the next block's observations prefetched
  long base = start + 1;
  double xp = (base + lane < n) ? xs[base + lane - 1] : nan;
  for (; base < n; base += 32) {
    const long nb = base + 32;
    const double xnext = (nb + lane < n) ? xs[nb + lane - 1] : nan;
    const unsigned miss = __ballot_sync(0xffffffffu, isnan(xp));
    const double t = __dmul_rn(oml, __dmul_rn(xp, xp));
    const int m = (int)(n - base < 32 ? n - base : 32);
    double mine = 0.0;
    if (m == 32) {
#pragma unroll
      for (int j = 0; j < 32; j++) {
        const double tj = __shfl_sync(0xffffffffu, t, j);
        if (!((miss >> j) & 1u)) var = __dadd_rn(__dmul_rn(lam, var), tj);      // no FMA: numba does not fuse
        if (lane == j) mine = var;
      }
    } else {
      for (int j = 0; j < m; j++) {
        const double tj = __shfl_sync(0xffffffffu, t, j);
        if (!((miss >> j) & 1u)) var = __dadd_rn(__dmul_rn(lam, var), tj);
        if (lane == j) mine = var;
      }
    }
    if (lane < m) vs[base + lane] = sqrt(mine);
    xp = xnext;
  }
}

}  // namespace

extern "C" {

int pfml_xs_ols_daily(const double* X, const double* y, const long* day_ptr, int n_days, int K, double* coef, double* resid,
                      int* status, void* stream) {
  if (!X || !y || !day_ptr || !coef || !resid || n_days <= 0 || K <= 0 || K > KA - 1) {
    set_error("pfml_xs_ols_daily: null/empty argument or K > %d", KA - 1);
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
  if (!fct_ret || !end_idx || !w_cor || !w_var || !cov_out || n_calc <= 0 || K <= 0 || K > KA || obs <= 0) {
    set_error("pfml_wcov_wcor: null/empty argument or K > %d", KA);
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(T_COV, s);
  const size_t smem = (2 * (size_t)obs + 8 * KA * (KA + 1)) * sizeof(double);
  constexpr size_t kMaxDyn = 176 * 1024;                  // + 41 KB of static shared memory stays below the 227 KB of a CTA
  if (smem > kMaxDyn) {
    set_error("pfml_wcov_wcor: window of %d observations exceeds the shared-memory weights (max 7900)", obs);
    return -2;
  }
  PFML_CONFIG_ONCE(wcov_wcor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDyn);
  wcov_wcor_kernel<<<n_calc, 256, smem, s>>>(K, obs, fct_ret, end_idx, w_cor, w_var, scale, cov_out);
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
  ewma_vol_kernel<<<ceil_div((long)n_series * 32, 128), 128, 0, s>>>(n_series, x, series_ptr, lam, start, vol);
  PFML_CHECK_LAUNCH();
  return 0;
}

}  // extern "C"
