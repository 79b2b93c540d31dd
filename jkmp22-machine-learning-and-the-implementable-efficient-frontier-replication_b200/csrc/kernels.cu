// Non-GEMM kernels of the PFML date-step: gathers, standardisation, element-wise matrix updates, the
// 64x64 base-block inversion of the recursive block inverse, reductions.  All FP64, all HBM/L2-bound:
// threads run along the contiguous (column) axis so every warp access is a full 256-byte row segment.
#include "kernels.cuh"

namespace pfml {

namespace {

__device__ __forceinline__ double nan_to_one(double v) { return isnan(v) ? 1.0 : v; }

// ------------------------------------------------------------------------------------------------
// per-date vectors (PFML_Input_Data.py:321-342, :404-405, :458; General_functions.py:948)
// ------------------------------------------------------------------------------------------------
__global__ void prep_vectors_kernel(PanelView pv, int NP, const int* __restrict__ date_t, const int* __restrict__ sel,
                                    const int* __restrict__ nsel, BatchVecs v) {
  const int b = blockIdx.x;
  const int t = date_t[b], n = nsel[b];
  const int nt = pv.lb + 2;
  const double w = pv.wealth[t];
  if (threadIdx.x == 0) {
    double* sc = v.scal + (long)b * NSCAL;
    sc[0] = w;
    sc[1] = 1.0 / w;
    sc[2] = 1.0 + pv.rf[t] + pv.mu;
    sc[3] = pv.gamma_rel;
  }
  for (int e = threadIdx.x; e < KP * KP; e += blockDim.x) {
    int a = e / KP, c = e % KP;
    v.Fg[(long)b * KP * KP + e] = (a < pv.K && c < pv.K) ? pv.gamma_rel * pv.fct_cov[((long)t * pv.K + a) * pv.K + c] : 0.0;
  }
  for (int i = threadIdx.x; i < NP; i += blockDim.x) {
    const long o = (long)b * NP + i;
    if (i < n) {
      const int s = sel[(long)b * NP + i];
      const long row = (long)t * pv.N + s;
      const double lam = pv.lam[row];
      const double isl = 1.0 / sqrt(lam);
      v.sl[o] = sqrt(lam);
      v.isl[o] = isl;
      for (int tau = 0; tau < nt; tau++) v.g[((long)b * nt + tau) * NP + i] = nan_to_one(pv.gt[(long)(t - tau) * pv.N + s]);
      v.g0raw[o] = pv.gt[row];
      v.rv[o] = isl * pv.ret_ld1[row];
      v.dvl[o] = pv.gamma_rel * pv.ivol[row] / lam;
    } else {
      v.sl[o] = 1.0;
      v.isl[o] = 1.0;
      for (int tau = 0; tau < nt; tau++) v.g[((long)b * nt + tau) * NP + i] = 1.0;
      v.g0raw[o] = 1.0;
      v.rv[o] = 0.0;
      v.dvl[o] = 0.0;
    }
  }
  // gathered loadings, L @ F and lambda^-1/2 L
  const double* F = pv.fct_cov + (long)t * pv.K * pv.K;
  for (int e = threadIdx.x; e < NP * KP; e += blockDim.x) {
    const int i = e / KP, c = e % KP;
    const long o = ((long)b * NP + i) * KP + c;
    double lg = 0.0, lf = 0.0, lt = 0.0;
    if (i < n && c < pv.K) {
      const int s = sel[(long)b * NP + i];
      const double* Lrow = pv.fct_load + ((long)t * pv.N + s) * pv.K;
      lg = Lrow[c];
      for (int a = 0; a < pv.K; a++) lf += Lrow[a] * F[(long)a * pv.K + c];
      lt = lg / sqrt(pv.lam[(long)t * pv.N + s]);
    }
    v.Lg[o] = lg;
    v.LF[o] = lf;
    v.Lt[o] = lt;
  }
}

// ------------------------------------------------------------------------------------------------
// x, sigma_gr, y of m_func (General_functions.py:944-955) from the factor form of Sigma (:869).
// Padded rows/cols (i >= n) are decoupled dummy stocks: x_ii = 1, sigma_gr_ii = 1, off-diagonals 0.
// ------------------------------------------------------------------------------------------------
__global__ void build_x_factor_kernel(PanelView pv, int NP, const int* __restrict__ date_t, const int* __restrict__ sel,
                                      const int* __restrict__ nsel, BatchVecs v, double* __restrict__ x,
                                      double* __restrict__ sgr, double* __restrict__ yv) {
  __shared__ double sLF[32][KP + 1];
  __shared__ double sL[32][KP + 1];
  const int b = blockIdx.z, i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
  const int n = nsel[b], t = date_t[b];
  const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
  for (int e = ty * 32 + tx; e < 32 * KP; e += 256) {
    int r = e / KP, c = e % KP;
    sLF[r][c] = v.LF[((long)b * NP + i0 + r) * KP + c];
    sL[r][c] = v.Lg[((long)b * NP + j0 + r) * KP + c];
  }
  __syncthreads();
  const double* sc = v.scal + (long)b * NSCAL;
  const double iw = sc[1], mubar = sc[2], gam = sc[3];
  const double cs = 1.0 / (mubar * mubar);
  const int j = j0 + tx;
  for (int r = ty; r < 32; r += 8) {
    const int i = i0 + r;
    double sig = 0.0;
#pragma unroll
    for (int c = 0; c < KP; c++) sig += sLF[r][c] * sL[tx][c];
    double xv, sv;
    if (i < n && j < n) {
      if (i == j) sig += pv.ivol[(long)t * pv.N + sel[(long)b * NP + i]];
      const double sg = sig * gam;
      xv = (iw * v.isl[(long)b * NP + i]) * sg * v.isl[(long)b * NP + j];
      sv = cs * (mubar * mubar + sg / gam);
    } else {
      xv = (i == j) ? 1.0 : 0.0;
      sv = (i == j) ? 1.0 : 0.0;
    }
    x[((long)b * NP + i) * NP + j] = xv;
    sgr[((long)b * NP + i) * NP + j] = sv;
    if (i == j) yv[(long)b * NP + i] = 1.0 + sv;
  }
}

__global__ void build_x_dense_kernel(int n, int NP, const double* __restrict__ sigma_gam, const double* __restrict__ lam,
                                     const double* __restrict__ w, const double* __restrict__ rf, double mu, double gam,
                                     double* __restrict__ x, double* __restrict__ sgr, double* __restrict__ yv,
                                     double* __restrict__ isl_o, double* __restrict__ sl_o) {
  const int b = blockIdx.z;
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= NP) return;
  const double mubar = 1.0 + rf[b] + mu;
  const double cs = 1.0 / (mubar * mubar);
  double xv, sv;
  if (i < n && j < n) {
    const double sg = sigma_gam[((long)b * n + i) * n + j];
    const double li = lam[(long)b * n + i], lj = lam[(long)b * n + j];
    xv = ((1.0 / w[b]) * (1.0 / sqrt(li))) * sg * (1.0 / sqrt(lj));
    sv = cs * (mubar * mubar + sg / gam);
  } else {
    xv = (i == j) ? 1.0 : 0.0;
    sv = (i == j) ? 1.0 : 0.0;
  }
  x[((long)b * NP + i) * NP + j] = xv;
  sgr[((long)b * NP + i) * NP + j] = sv;
  if (i == j) {
    yv[(long)b * NP + i] = 1.0 + sv;
    const double l = i < n ? lam[(long)b * n + i] : 1.0;
    isl_o[(long)b * NP + i] = 1.0 / sqrt(l);
    sl_o[(long)b * NP + i] = sqrt(l);
  }
}

// ------------------------------------------------------------------------------------------------
// Standardise + scale (PFML_Input_Data.py:357-388) for every (date, tau): de-mean each RFF column over the
// date-d universe (two passes: mean, then sum of squares of the centred column, quirk Q6), l2-normalise,
// multiply rows by 1/vol_scale, then by sqrt(lambda_i) (tilde domain) and, if `prescale`, for tau in 2..lb by
// g_{tau-1} (pre-scaled Horner addend of the DMMA path).  Output columns are in feat_all order: constant first (quirk Q3/Q4).
// One thread per output column, rows streamed; the slab stays L2-resident between the three passes.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
standardize_kernel(PanelView pv, int NP, int P, long ldP, const int* __restrict__ date_t, const int* __restrict__ sel,
                   const int* __restrict__ nsel, BatchVecs v, double* __restrict__ St, double* __restrict__ signal_t, int prescale,
                   double* __restrict__ signals_all) {
  const int b = blockIdx.z, tau = blockIdx.y;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;   // output column (0 = constant)
  const int n = nsel[b], t = date_t[b] - tau;
  const int nt = pv.lb + 2;
  extern __shared__ int s_sel[];
  for (int i = threadIdx.x; i < n; i += blockDim.x) s_sel[i] = sel[(long)b * NP + i];
  __syncthreads();
  if (j >= P) return;
  const double* __restrict__ X = pv.rff + (long)t * pv.N * pv.p + (j > 0 ? j - 1 : 0);
  double mean = 0.0, scale;
  if (j == 0) {
    scale = sqrt(1.0 / (double)n);                        // constant: not de-meaned, sum of squares = n
  } else {
    double sum = 0.0;
#pragma unroll 8
    for (int i = 0; i < n; i++) sum += X[(long)s_sel[i] * pv.p];
    mean = sum / (double)n;
    double ss = 0.0;
#pragma unroll 8
    for (int i = 0; i < n; i++) {
      double c = X[(long)s_sel[i] * pv.p] - mean;
      ss += c * c;
    }
    scale = sqrt(1.0 / ss);
  }
  const double* __restrict__ vs = pv.vol_scale + (long)t * pv.N;
  const double* __restrict__ slb = v.sl + (long)b * NP;
  const double* __restrict__ gprev = (prescale && tau >= 2 && tau <= pv.lb) ? v.g + ((long)b * nt + (tau - 1)) * NP : nullptr;
  double* __restrict__ out = St + (((long)b * nt + tau) * NP) * ldP + j;
  // signal_t: slab 0 only (what the reference stores, :485-489); signals_all: every slab [B][lb+2][NP][P] (stage entry point)
  double* __restrict__ sig = signals_all ? signals_all + (((long)b * nt + tau) * NP) * P + j
                                         : ((tau == 0 && signal_t) ? signal_t + (long)b * NP * P + j : nullptr);
#pragma unroll 4
  for (int i = 0; i < n; i++) {
    const int s = s_sel[i];
    const double c = (j == 0) ? 1.0 : X[(long)s * pv.p] - mean;
    const double sv = (1.0 / vs[s]) * (c * scale);
    if (sig) sig[(long)i * P] = sv;
    double tv = slb[i] * sv;
    if (gprev) tv *= gprev[i];
    out[(long)i * ldP] = tv;
  }
  for (int i = n; i < NP; i++) {
    out[(long)i * ldP] = 0.0;
    if (sig) sig[(long)i * P] = 0.0;
  }
}

// ------------------------------------------------------------------------------------------------
// element-wise helpers on [B][NP][NP]
// ------------------------------------------------------------------------------------------------
__global__ void copy2_kernel(long n, const double* __restrict__ src, double* __restrict__ d1, double* __restrict__ d2) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    double v = src[i];
    d1[i] = v;
    if (d2) d2[i] = v;
  }
}

__global__ void add_kernel(long n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ out) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = a[i] + b[i];
}

// const_l1~ = I + sum_{k=1}^{nk} B~_k + B~_lb  and  G_0 (I + sum_{k=1}^{nk} B~_k) from the stored chain matrices
struct ChainPtrs { const double* bk[16]; };
__global__ void chain_sum_kernel(int NP, int nk, ChainPtrs p, const double* __restrict__ b11, const double* __restrict__ g0, long sg,
                                 double* __restrict__ wk, double* __restrict__ cbs) {
  const int b = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NP) return;
  const long o = ((long)b * NP + r) * NP + c;
  double acc = (r == c) ? 1.0 : 0.0;
  for (int k = 0; k < nk; k++) acc += p.bk[k][o];
  wk[o] = acc + b11[o];
  cbs[o] = acc * g0[(long)b * sg + r];
}

// a9 finish for the upper-triangle products: denom = risk + tc, and risk, tc, denom mirrored into their lower triangles
// (32 x 32 tiles through shared memory so that both the reads and the transposed writes are coalesced).
//   tc_nan[date]    omega_chg held NaN (quirk Q8: missing growth rate at date d): every element of tc and denom is NaN, as
//                   in the reference's dense product
//   bad[date] & 16  an operand of the INT8 products of this date held NaN / Inf (it was sliced as zero): risk, tc and denom
//                   are all NaN, which is what the reference's FP64 arithmetic returns for such a date
// With `bucket` (expanding-window sums of PFML_Search_Coef.py:77-88, :112-121 fused in): blockIdx.z is a window k, the
// block walks the dates of that window in order and adds their in-order sum of denom to d_sum[k] (z = n_buckets takes the
// dates that belong to no window).  Without it blockIdx.z is the date.
__device__ __forceinline__ double qnan() { return __longlong_as_double(0x7ff8000000000000ll); }

template <bool BUCKETS>
__global__ void sym_finish_kernel(int B, int P, double* __restrict__ risk, double* __restrict__ tc, double* __restrict__ denom,
                                  const int* __restrict__ tc_nan, const int* __restrict__ bad, const int* __restrict__ bucket,
                                  int n_buckets, double* __restrict__ d_sum) {
  __shared__ double tr[32][33], tt[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bj < bi) return;
  const int tx = threadIdx.x, ty = threadIdx.y;
  double acc_u[4], acc_l[4];
  bool any = false;
  const int b_lo = BUCKETS ? 0 : blockIdx.z, b_hi = BUCKETS ? B : blockIdx.z + 1;
  for (int b = b_lo; b < b_hi; b++) {
    if (BUCKETS) {
      const int kb = bucket[b];
      if (((kb >= 0 && kb < n_buckets) ? kb : n_buckets) != (int)blockIdx.z) continue;     // uniform over the block
    }
    const long zo = (long)b * P * P;
    const bool all_nan = bad && (bad[b] & 16), t_nan = all_nan || (tc_nan && tc_nan[b]);
    for (int k = ty; k < 32; k += 8) {
      const int r = bi * 32 + k, c = bj * 32 + tx;
      const bool ok = r < P && c < P;
      tr[k][tx] = ok ? (all_nan ? qnan() : risk[zo + (long)r * P + c]) : 0.0;
      tt[k][tx] = ok ? (t_nan ? qnan() : tc[zo + (long)r * P + c]) : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int k = ty + 8 * q;
      // upper tile position (r, c): only elements on / above the diagonal are valid inputs
      int r = bi * 32 + k, c = bj * 32 + tx;
      if (r < P && c < P && c >= r) {
        const double d = tr[k][tx] + tt[k][tx];
        denom[zo + (long)r * P + c] = d;
        if (t_nan) tc[zo + (long)r * P + c] = tt[k][tx];
        if (all_nan) risk[zo + (long)r * P + c] = tr[k][tx];
        if (BUCKETS) acc_u[q] = any ? acc_u[q] + d : d;
      }
      // mirrored position (c', r') = (bj * 32 + k, bi * 32 + tx) takes the value of (bi * 32 + tx, bj * 32 + k)
      r = bj * 32 + k; c = bi * 32 + tx;
      if (r < P && c < P && r > c) {
        const double rv = tr[tx][k], tv = tt[tx][k];
        risk[zo + (long)r * P + c] = rv;
        tc[zo + (long)r * P + c] = tv;
        denom[zo + (long)r * P + c] = rv + tv;
        if (BUCKETS) acc_l[q] = any ? acc_l[q] + (rv + tv) : (rv + tv);
      }
    }
    any = true;
    if (BUCKETS) __syncthreads();
  }
  if (BUCKETS && any && (int)blockIdx.z < n_buckets) {
    double* out = d_sum + (long)blockIdx.z * P * P;
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int k = ty + 8 * q;
      int r = bi * 32 + k, c = bj * 32 + tx;
      if (r < P && c < P && c >= r) out[(long)r * P + c] += acc_u[q];
      r = bj * 32 + k; c = bi * 32 + tx;
      if (r < P && c < P && r > c) out[(long)r * P + c] += acc_l[q];
    }
  }
}

// r_tilde (and, off the symmetric INT8 path, the matrices) of a date whose operands held NaN / Inf -> NaN
__global__ void poison_kernel(int P, const int* __restrict__ bad, double* __restrict__ r_tilde, double* __restrict__ denom,
                              double* __restrict__ risk, double* __restrict__ tc) {
  const int b = blockIdx.x;
  if (!(bad[b] & 16)) return;
  for (int j = threadIdx.x; j < P; j += blockDim.x) r_tilde[(long)b * P + j] = qnan();
  for (long e = threadIdx.x; e < (long)P * P; e += blockDim.x) {
    if (denom) denom[(long)b * P * P + e] = qnan();
    if (risk) risk[(long)b * P * P + e] = qnan();
    if (tc) tc[(long)b * P * P + e] = qnan();
  }
}

// omega~ = Lambda^1/2 omega for rows < n, zero rows beyond (stage entry: a9 from a given omega)
__global__ void tilde_rows_kernel(int NP, int P, const double* __restrict__ in, const int* __restrict__ nsel, const double* __restrict__ sl,
                                  double* __restrict__ out, long ldo, long stride_o) {
  const int b = blockIdx.z, i = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= P) return;
  out[(long)b * stride_o + (long)i * ldo + j] = i < nsel[b] ? sl[(long)b * NP + i] * in[((long)b * NP + i) * P + j] : 0.0;
}

// partial Frobenius sums of M, Mi and M - I over a chunk of rows; deterministic two-stage reduction
__global__ void db_norms_partial_kernel(int NP, const double* __restrict__ M, const double* __restrict__ Mi, double* __restrict__ part) {
  const int b = blockIdx.y, ch = blockIdx.x;
  const int rows_per = (NP + NORM_CHUNKS - 1) / NORM_CHUNKS;
  const int r0 = ch * rows_per, r1 = min(NP, r0 + rows_per);
  double s0 = 0, s1 = 0, s2 = 0;
  for (int r = r0; r < r1; r++) {
    const double* m = M + ((long)b * NP + r) * NP;
    const double* mi = Mi + ((long)b * NP + r) * NP;
    for (int c = threadIdx.x; c < NP; c += blockDim.x) {
      double a = m[c], q = mi[c], d = a - (r == c ? 1.0 : 0.0);
      s0 += a * a;
      s1 += q * q;
      s2 += d * d;
    }
  }
  __shared__ double red[3][256];
  red[0][threadIdx.x] = s0; red[1][threadIdx.x] = s1; red[2][threadIdx.x] = s2;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      red[0][threadIdx.x] += red[0][threadIdx.x + o];
      red[1][threadIdx.x] += red[1][threadIdx.x + o];
      red[2][threadIdx.x] += red[2][threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x < 3) part[((long)b * NORM_CHUNKS + ch) * 3 + threadIdx.x] = red[threadIdx.x][0];
}

// Frobenius-norm scaling of the product-form Denman-Beavers iteration (Higham, Functions of Matrices, 6.29):
//   mu = (||M^-1||_F / ||M||_F)^(1/4), switched off (mu = 1) once ||M - I||_F < 1e-2.
// sc: 0 alpha = 0.5/mu, 1 beta = 0.5*mu, 2 mu^2, 3 1/mu^2, 4 ||M-I||_F (before this update)
__global__ void db_norms_final_kernel(int B, const double* __restrict__ part, double* __restrict__ sc) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double s0 = 0, s1 = 0, s2 = 0;
  for (int ch = 0; ch < NORM_CHUNKS; ch++) {
    s0 += part[((long)b * NORM_CHUNKS + ch) * 3 + 0];
    s1 += part[((long)b * NORM_CHUNKS + ch) * 3 + 1];
    s2 += part[((long)b * NORM_CHUNKS + ch) * 3 + 2];
  }
  const double res = sqrt(s2);
  double mu = 1.0;
  if (res >= 1e-2) mu = sqrt(sqrt(sqrt(s1) / sqrt(s0)));
  double* o = sc + (long)b * 8;
  o[0] = 0.5 / mu;
  o[1] = 0.5 * mu;
  o[2] = mu * mu;
  o[3] = 1.0 / (mu * mu);
  o[4] = res;
}

__global__ void db_update_M_kernel(int NP, double* __restrict__ M, const double* __restrict__ Mi, double* __restrict__ Wk,
                                   const double* __restrict__ sc) {
  const int b = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NP) return;
  const double mu2 = sc[(long)b * 8 + 2], imu2 = sc[(long)b * 8 + 3];
  const long o = ((long)b * NP + r) * NP + c;
  const double v = 0.5 * ((r == c ? 1.0 : 0.0) + 0.5 * (mu2 * M[o] + imu2 * Mi[o]));
  M[o] = v;
  Wk[o] = v;
}

// m_tilde_0 = 0.5 (sigma_hat - sqrtm(sigma_hat^2 - 4I)), sigma_hat = x + 2I  (General_functions.py:955-956)
__global__ void mt0_kernel(int NP, const double* __restrict__ x, const double* __restrict__ Y, double* __restrict__ mt) {
  const int b = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NP) return;
  const long o = ((long)b * NP + r) * NP + c;
  mt[o] = 0.5 * ((x[o] + (r == c ? 2.0 : 0.0)) - Y[o]);
}

// x + y - m_tilde o sigma_gr (Hadamard)  (General_functions.py:960)
__global__ void fp_build_kernel(int NP, const double* __restrict__ x, const double* __restrict__ yv, const double* __restrict__ mt,
                                const double* __restrict__ sgr, double* __restrict__ Wk) {
  const int b = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NP) return;
  const long o = ((long)b * NP + r) * NP + c;
  double v = x[o];
  if (r == c) v += yv[(long)b * NP + r];
  Wk[o] = v - mt[o] * sgr[o];
}

// B_1 = m_tilde G_1 (column scaling) and CB = I + B_1
__global__ void chain_init_kernel(int NP, const double* __restrict__ mt, const double* __restrict__ g1, long sg,
                                  double* __restrict__ Bt, double* __restrict__ CB) {
  const int b = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NP) return;
  const long o = ((long)b * NP + r) * NP + c;
  const double v = mt[o] * g1[(long)b * sg + c];
  Bt[o] = v;
  CB[o] = v + (r == c ? 1.0 : 0.0);
}

__global__ void row_scale_kernel(int rows, int cols, long ld, long stride, const double* __restrict__ in,
                                 const double* __restrict__ rsv, long srs, double* __restrict__ out, long ldo, long stride_o) {
  const int b = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  out[(long)b * stride_o + (long)r * ldo + c] = rsv[(long)b * srs + r] * in[(long)b * stride + (long)r * ld + c];
}

// m = Lambda^-1/2 m_tilde Lambda^1/2 (General_functions.py:962), compact n x n output
__global__ void m_from_mtilde_kernel(int n, int NP, const double* __restrict__ mt, const double* __restrict__ isl,
                                     const double* __restrict__ sl, double* __restrict__ m_out) {
  const int b = blockIdx.z, r = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  m_out[((long)b * n + r) * n + c] = isl[(long)b * NP + r] * mt[((long)b * NP + r) * NP + c] * sl[(long)b * NP + c];
}

// ------------------------------------------------------------------------------------------------
// Gauss-Jordan inverse of one NB x NB diagonal block per CTA (NB = 64 or 128), matrix held in REGISTERS.
// 256 threads in a 16 x 16 grid with a CYCLIC distribution: thread (ti, tj) owns rows ti + 16 r and columns
// tj + 16 c (r, c < TS = NB / 16).  Column k = 16 cb + kk then lives in register column cb of the threads with
// tj == kk, so an outer loop over cb unrolled at compile time makes every register index static (no selects,
// no local memory).  Per pivot step the owners publish column k and the pivot row through double-buffered
// shared memory and every thread applies the rank-1 update to its tile (TS^2 DFMAs).
// PIVOT = false (SPD blocks of m_tilde's fixed point): pivot k is row k, one barrier per step.
// PIVOT = true  (general blocks): implicit partial pivoting -- every warp finds the unused row with the largest
// |a[r][k]| from the published column, rows are never moved and the permutation is undone on output
// (Ainv[k][p_c] = stored[p_k][c]).
// ------------------------------------------------------------------------------------------------
template <int NB, bool PIVOT>
__global__ void __launch_bounds__(256)
gj_inverse_kernel(int NP, const double* __restrict__ src, double* __restrict__ dst, int off, int* __restrict__ status) {
  constexpr int TS = NB / 16;
  __shared__ double prow[2][NB];
  __shared__ double pcol[2][NB];
  __shared__ int s_piv[NB];                // pivot row of step k
  __shared__ int s_pinv[NB];               // step at which row r was the pivot
  __shared__ int s_used[NB];
  __shared__ int bad;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int ti = tid >> 4, tj = tid & 15, lane = tid & 31;
  const double* S = src + (long)b * NP * NP + (long)off * NP + off;
  double* Dp = dst + (long)b * NP * NP + (long)off * NP + off;
  if (tid == 0) bad = 0;
  if (PIVOT && tid < NB) s_used[tid] = 0;
  double a[TS][TS];
#pragma unroll
  for (int r = 0; r < TS; r++)
#pragma unroll
    for (int c = 0; c < TS; c++) a[r][c] = S[(long)(ti + 16 * r) * NP + tj + 16 * c];
  if (PIVOT) __syncthreads();
#pragma unroll
  for (int cb = 0; cb < TS; cb++) {
#pragma unroll 1
    for (int kk = 0; kk < 16; kk++) {
      const int k = 16 * cb + kk, buf = k & 1;
      if (tj == kk) {                      // owners of column k publish it and clear it in their tile
#pragma unroll
        for (int r = 0; r < TS; r++) {
          pcol[buf][ti + 16 * r] = a[r][cb];
          a[r][cb] = 0.0;
        }
      }
      int p = k;
      if (PIVOT) {
        __syncthreads();
        double best = -1.0;
        int bi = -1;
#pragma unroll
        for (int m = 0; m < NB / 32; m++) {
          const int i = lane + 32 * m;
          const double v = fabs(pcol[buf][i]);       // NaN never wins a comparison: an all-NaN column is flagged
          if (!s_used[i] && v > best) { best = v; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const double ob = __shfl_xor_sync(0xffffffffu, best, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (ob > best || (ob == best && oi >= 0 && (bi < 0 || oi < bi))) { best = ob; bi = oi; }
        }
        p = bi < 0 ? k : bi;                         // identical in every warp
        if (tid == 0) {
          s_piv[k] = p; s_pinv[p] = k;
          if (!(best > 0.0) || !isfinite(best)) bad = 1;
        }
      }
      const int rhi = PIVOT ? (p >> 4) : cb, rlo = PIVOT ? (p & 15) : kk;
      if (ti == rlo) {                     // owners of the pivot row publish it
#pragma unroll
        for (int r = 0; r < TS; r++)
          if (r == rhi) {
#pragma unroll
            for (int c = 0; c < TS; c++) prow[buf][tj + 16 * c] = a[r][c];
          }
      }
      __syncthreads();
      if (PIVOT && tid == 0) s_used[p] = 1;          // read again only after the next step's first barrier
      const double pv = pcol[buf][p];
      if (!PIVOT && tid == 0 && (!(fabs(pv) > 0.0) || !isfinite(pv))) bad = 1;
      const double ipv = 1.0 / pv;
      double q[TS], f[TS];
#pragma unroll
      for (int c = 0; c < TS; c++) q[c] = prow[buf][tj + 16 * c] * ipv;
      if (tj == kk) q[cb] = ipv;
#pragma unroll
      for (int r = 0; r < TS; r++) f[r] = pcol[buf][ti + 16 * r];
#pragma unroll
      for (int r = 0; r < TS; r++)
#pragma unroll
        for (int c = 0; c < TS; c++) a[r][c] -= f[r] * q[c];
      if (ti == rlo) {                     // the pivot row itself becomes q
#pragma unroll
        for (int r = 0; r < TS; r++)
          if (r == rhi) {
#pragma unroll
            for (int c = 0; c < TS; c++) a[r][c] = q[c];
          }
      }
      // double-buffered prow/pcol: the next step writes the other buffer, and its barrier orders this step's reads
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < TS; r++) {
    const int row = ti + 16 * r;
    const int orow = PIVOT ? s_pinv[row] : row;
#pragma unroll
    for (int c = 0; c < TS; c++) {
      const int col = tj + 16 * c;
      Dp[(long)orow * NP + (PIVOT ? s_piv[col] : col)] = a[r][c];
    }
  }
  if (tid == 0 && bad && status) atomicOr(status + b, 1);
}

// out[c] = sum_r A[r][c] v[r]   (r_tilde = omega^T r, PFML_Input_Data.py:470)
__global__ void gemv_t_kernel(int rows, int cols, const double* __restrict__ A, long ld, long sA, const double* __restrict__ v,
                              long sv, double* __restrict__ out, long so) {
  const int b = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  const double* a = A + (long)b * sA + c;
  const double* vv = v + (long)b * sv;
  double s = 0.0;
#pragma unroll 8
  for (int r = 0; r < rows; r++) s += a[(long)r * ld] * vv[r];
  out[(long)b * so + c] = s;
}

__global__ void check_status_kernel(int B, const double* __restrict__ sc, int* __restrict__ status) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double res = sc[(long)b * 8 + 4];
  if (!(res < 1e-9)) atomicOr(status + b, 2);   // Denman-Beavers did not converge (or NaN input)
}

// g_t = (1 + tr_ld0) / (1 + mu_ld0)  (PFML_Input_Data.py:404, :458); NaN propagates
__global__ void gt_kernel(long n, const double* __restrict__ tr, const double* __restrict__ mu, double* __restrict__ gt) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) gt[i] = (1.0 + tr[i]) / (1.0 + mu[i]);
}

// vol_scale = sqrt(diag(L F L' + diag(ivol))) (PFML_Input_Data.py:278-281) without forming Sigma; NaN if absent
__global__ void vol_scale_kernel(const double* __restrict__ L, const double* __restrict__ F, const double* __restrict__ ivol,
                                 const unsigned char* __restrict__ has, int N, int K, double* __restrict__ vs) {
  const int t = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  extern __shared__ double sF[];
  for (int e = threadIdx.x; e < K * K; e += blockDim.x) sF[e] = F[(long)t * K * K + e];
  __syncthreads();
  if (i >= N) return;
  const long row = (long)t * N + i;
  if (has && !has[row]) {
    vs[row] = __longlong_as_double(0x7ff8000000000000LL);
    return;
  }
  const double* l = L + row * K;
  double q = 0.0;
  for (int c = 0; c < K; c++) {
    double lf = 0.0;
    for (int a = 0; a < K; a++) lf += l[a] * sF[a * K + c];
    q += lf * l[c];
  }
  vs[row] = sqrt(q + ivol[row]);
}

// per-month median (skipna) of vol_scale, written into the NaN slots (PFML_Input_Data.py:303-305).
// One CTA per month; bitonic sort of the month's values in shared memory (NaN -> +inf).
__global__ void median_fill_kernel(int N, int NPOW, double* __restrict__ vs) {
  extern __shared__ double sv[];
  const int t = blockIdx.x;
  double* v = vs + (long)t * N;
  __shared__ int n_nan;
  if (threadIdx.x == 0) n_nan = 0;
  __syncthreads();
  int local = 0;
  for (int i = threadIdx.x; i < NPOW; i += blockDim.x) {
    double x = i < N ? v[i] : INFINITY;
    if (i < N && isnan(x)) { x = INFINITY; local++; }
    sv[i] = x;
  }
  if (local) atomicAdd(&n_nan, local);
  __syncthreads();
  if (n_nan == 0) return;
  for (int k = 2; k <= NPOW; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < NPOW; i += blockDim.x) {
        int ixj = i ^ j;
        if (ixj > i) {
          bool up = (i & k) == 0;
          double a = sv[i], b = sv[ixj];
          if ((a > b) == up) { sv[i] = b; sv[ixj] = a; }
        }
      }
      __syncthreads();
    }
  const int m = N - n_nan;
  double med = __longlong_as_double(0x7ff8000000000000LL);
  if (m > 0) med = (m & 1) ? sv[m / 2] : 0.5 * (sv[m / 2 - 1] + sv[m / 2]);
  for (int i = threadIdx.x; i < N; i += blockDim.x)
    if (isnan(v[i])) v[i] = med;
}

inline dim3 grid_rc(int cols, int rows, int B, int bt = 128) { return dim3(ceil_div(cols, bt), rows, B); }

}  // namespace

int prep_vectors(const PanelView& pv, int B, int NP, const int* date_t, const int* sel, const int* nsel, const BatchVecs& v,
                 cudaStream_t s) {
  prep_vectors_kernel<<<B, 256, 0, s>>>(pv, NP, date_t, sel, nsel, v);
  PFML_CHECK_LAUNCH();
  return 0;
}

int build_x_factor(const PanelView& pv, int B, int NP, const int* date_t, const int* sel, const int* nsel, const BatchVecs& v,
                   double* x, double* sgr, double* yv, cudaStream_t s) {
  dim3 grid(NP / 32, NP / 32, B), block(32, 8);
  build_x_factor_kernel<<<grid, block, 0, s>>>(pv, NP, date_t, sel, nsel, v, x, sgr, yv);
  PFML_CHECK_LAUNCH();
  return 0;
}

int build_x_dense(int B, int n, int NP, const double* sigma_gam, const double* lam, const double* w, const double* rf, double mu,
                  double gam, double* x, double* sgr, double* yv, double* isl, double* sl, cudaStream_t s) {
  build_x_dense_kernel<<<grid_rc(NP, NP, B), 128, 0, s>>>(n, NP, sigma_gam, lam, w, rf, mu, gam, x, sgr, yv, isl, sl);
  PFML_CHECK_LAUNCH();
  return 0;
}

int standardize(const PanelView& pv, int B, int NP, int P, long ldP, const int* date_t, const int* sel, const int* nsel,
                const BatchVecs& v, double* St, double* signal_t, bool prescale, int ntau, cudaStream_t s, double* signals_all) {
  dim3 grid(ceil_div(P, 128), ntau, B);
  standardize_kernel<<<grid, 128, NP * sizeof(int), s>>>(pv, NP, P, ldP, date_t, sel, nsel, v, St, signal_t, prescale ? 1 : 0, signals_all);
  PFML_CHECK_LAUNCH();
  return 0;
}

int copy2(int B, long n_elem, const double* src, double* d1, double* d2, cudaStream_t s) {
  long n = (long)B * n_elem;
  int blocks = (int)((n + 255) / 256 > 148 * 16 ? 148 * 16 : (n + 255) / 256);
  copy2_kernel<<<blocks, 256, 0, s>>>(n, src, d1, d2);
  PFML_CHECK_LAUNCH();
  return 0;
}

int chain_sum(int B, int NP, int nk, const double* const* bk, const double* b11, const double* g0, long sg, double* wk, double* cbs,
              cudaStream_t s) {
  if (nk > 16) return -1;
  ChainPtrs p{};
  for (int k = 0; k < nk; k++) p.bk[k] = bk[k];
  chain_sum_kernel<<<grid_rc(NP, NP, B), 128, 0, s>>>(NP, nk, p, b11, g0, sg, wk, cbs);
  PFML_CHECK_LAUNCH();
  return 0;
}

int sym_finish(int B, int P, double* risk, double* tc, double* denom, const int* tc_nan, const int* bad, const int* bucket,
               int n_buckets, double* d_sum, cudaStream_t s) {
  const int nb = ceil_div(P, 32);
  if (bucket && d_sum && n_buckets > 0)
    sym_finish_kernel<true><<<dim3(nb, nb, n_buckets + 1), dim3(32, 8), 0, s>>>(B, P, risk, tc, denom, tc_nan, bad, bucket, n_buckets, d_sum);
  else
    sym_finish_kernel<false><<<dim3(nb, nb, B), dim3(32, 8), 0, s>>>(B, P, risk, tc, denom, tc_nan, bad, nullptr, 0, nullptr);
  PFML_CHECK_LAUNCH();
  return 0;
}

int poison_bad_dates(int B, int P, const int* bad, double* r_tilde, double* denom, double* risk, double* tc, cudaStream_t s) {
  poison_kernel<<<B, 256, 0, s>>>(P, bad, r_tilde, denom, risk, tc);
  PFML_CHECK_LAUNCH();
  return 0;
}

int tilde_rows(int B, int NP, int P, const double* in, const int* nsel, const double* sl, double* out, long ldo, long stride_o,
               cudaStream_t s) {
  tilde_rows_kernel<<<grid_rc(P, NP, B), 128, 0, s>>>(NP, P, in, nsel, sl, out, ldo, stride_o);
  PFML_CHECK_LAUNCH();
  return 0;
}

int add_mat(int B, long n_elem, const double* a, const double* b, double* out, cudaStream_t s) {
  long n = (long)B * n_elem;
  int blocks = (int)((n + 255) / 256 > 148 * 16 ? 148 * 16 : (n + 255) / 256);
  add_kernel<<<blocks, 256, 0, s>>>(n, a, b, out);
  PFML_CHECK_LAUNCH();
  return 0;
}

int db_norms(int B, int NP, const double* M, const double* Mi, double* part, double* sc, cudaStream_t s) {
  db_norms_partial_kernel<<<dim3(NORM_CHUNKS, B), 256, 0, s>>>(NP, M, Mi, part);
  PFML_CHECK_LAUNCH();
  db_norms_final_kernel<<<ceil_div(B, 128), 128, 0, s>>>(B, part, sc);
  PFML_CHECK_LAUNCH();
  return 0;
}

int db_update_M(int B, int NP, double* M, const double* Mi, double* Wk, const double* sc, cudaStream_t s) {
  db_update_M_kernel<<<grid_rc(NP, NP, B), 128, 0, s>>>(NP, M, Mi, Wk, sc);
  PFML_CHECK_LAUNCH();
  return 0;
}

int mt0_from_sqrt(int B, int NP, const double* x, const double* Y, double* mt, cudaStream_t s) {
  mt0_kernel<<<grid_rc(NP, NP, B), 128, 0, s>>>(NP, x, Y, mt);
  PFML_CHECK_LAUNCH();
  return 0;
}

int fp_build(int B, int NP, const double* x, const double* yv, const double* mt, const double* sgr, double* Wk, cudaStream_t s) {
  fp_build_kernel<<<grid_rc(NP, NP, B), 128, 0, s>>>(NP, x, yv, mt, sgr, Wk);
  PFML_CHECK_LAUNCH();
  return 0;
}

int chain_init(int B, int NP, const double* mt, const double* g1, long sg, double* Bt, double* CB, cudaStream_t s) {
  chain_init_kernel<<<grid_rc(NP, NP, B), 128, 0, s>>>(NP, mt, g1, sg, Bt, CB);
  PFML_CHECK_LAUNCH();
  return 0;
}

int row_scale(int B, int rows, int cols, long ld, long stride, const double* in, const double* rsv, long srs, double* out,
              long ldo, long stride_o, cudaStream_t s) {
  row_scale_kernel<<<grid_rc(cols, rows, B), 128, 0, s>>>(rows, cols, ld, stride, in, rsv, srs, out, ldo, stride_o);
  PFML_CHECK_LAUNCH();
  return 0;
}

int m_from_mtilde(int B, int n, int NP, const double* mt, const double* isl, const double* sl, double* m_out, cudaStream_t s) {
  m_from_mtilde_kernel<<<grid_rc(n, n, B), 128, 0, s>>>(n, NP, mt, isl, sl, m_out);
  PFML_CHECK_LAUNCH();
  return 0;
}

int base_inverse(int B, int NP, const double* src, double* dst, int off, int size, bool spd, int* status, cudaStream_t s) {
  if (size == 64) {
    if (spd) gj_inverse_kernel<64, false><<<B, 256, 0, s>>>(NP, src, dst, off, status);
    else gj_inverse_kernel<64, true><<<B, 256, 0, s>>>(NP, src, dst, off, status);
  } else if (size == 128) {
    if (spd) gj_inverse_kernel<128, false><<<B, 256, 0, s>>>(NP, src, dst, off, status);
    else gj_inverse_kernel<128, true><<<B, 256, 0, s>>>(NP, src, dst, off, status);
  } else {
    set_error("base_inverse: unsupported block size %d", size);
    return -2;
  }
  PFML_CHECK_LAUNCH();
  return 0;
}

int gemv_t(int B, int rows, int cols, const double* A, long ld, long sA, const double* v, long sv, double* out, long so,
           cudaStream_t s) {
  gemv_t_kernel<<<dim3(ceil_div(cols, 128), B), 128, 0, s>>>(rows, cols, A, ld, sA, v, sv, out, so);
  PFML_CHECK_LAUNCH();
  return 0;
}

int gt_panel(long n, const double* tr_ld0, const double* mu_ld0, double* gt, cudaStream_t s) {
  gt_kernel<<<ceil_div(n, 256), 256, 0, s>>>(n, tr_ld0, mu_ld0, gt);
  PFML_CHECK_LAUNCH();
  return 0;
}

int vol_scale_panel(const double* fct_load, const double* fct_cov, const double* ivol, const unsigned char* has, int T, int N,
                    int K, double* vol_scale, cudaStream_t s) {
  vol_scale_kernel<<<dim3(ceil_div(N, 128), T), 128, K * K * sizeof(double), s>>>(fct_load, fct_cov, ivol, has, N, K, vol_scale);
  PFML_CHECK_LAUNCH();
  if (has) {
    int npow = 1;
    while (npow < N) npow <<= 1;
    if (npow > 4096 * 4) {
      set_error("pfml_vol_scale: median fill supports N <= 16384 stocks per month (got %d)", N);
      return -2;
    }
    size_t smem = (size_t)npow * sizeof(double);
    PFML_CONFIG_ONCE(median_fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8);
    median_fill_kernel<<<T, 512, smem, s>>>(N, npow, vol_scale);
    PFML_CHECK_LAUNCH();
  }
  return 0;
}

int check_finite_status(int B, const double* sc, int* status, cudaStream_t s) {
  check_status_kernel<<<ceil_div(B, 128), 128, 0, s>>>(B, sc, status);
  PFML_CHECK_LAUNCH();
  return 0;
}

}  // namespace pfml
