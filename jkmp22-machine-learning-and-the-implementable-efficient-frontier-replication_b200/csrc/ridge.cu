// PFML_Search_Coef.py on the device: expanding-window sums (a11) and the ridge grid (a12/a13).
//
//   bucket sums     denom_sum_fun / r_tilde sums per window        PFML_Search_Coef.py:37-46, :77-88, :112-121
//   cumulative      running sums over the year buckets             :117, :121
//   ridge grid      beta = solve(denom_sub / n + lambda I, r_sub / n) for every (year, p, lambda), :125-135,
//                   denom_sub = the pfml_feat_fun(p) principal sub-matrix (General_functions.py:837-844):
//                   index set {0} U {1..p/2} U {p_max/2+1 .. p_max/2+p/2} of the feat_all layout.
//
// The solve is the recursive block inverse of the (symmetric positive definite for lambda > 0) system matrix
// followed by one step of iterative refinement against the un-inverted matrix, which brings the result back to
// LU-solve quality; lambda = 0 systems are as ill-conditioned as the reference's (quirk Q14) and are reported,
// not hidden (status bit 0 on a zero pivot).
#include "kernels.cuh"
#include "pfml_b200.h"

namespace pfml {
int spd_inverse(int B, int NP, double* Wk, double* X, double* T1, int* status, cudaStream_t s);   // pipeline.cu
}

using namespace pfml;

namespace {
__global__ void bucket_accumulate_kernel(int B, long elems, const double* __restrict__ in, const int* __restrict__ bucket,
                                         int n_buckets, double* __restrict__ out);
}

namespace pfml {
int bucket_accumulate(int B, long elems, const double* in, const int* bucket, int n_buckets, double* out, cudaStream_t s) {
  bucket_accumulate_kernel<<<dim3(ceil_div(elems, 256), n_buckets), 256, 0, s>>>(B, elems, in, bucket, n_buckets, out);
  PFML_CHECK_LAUNCH();
  return 0;
}
}  // namespace pfml

namespace {

__device__ __forceinline__ int feat_index(int k, int p, int p_max) {
  // pfml_feat_fun(p) inside pfml_feat_fun(p_max): constant, cos 1..p/2, sin 1..p/2
  const int h = p >> 1;
  return k == 0 ? 0 : (k <= h ? k : (p_max >> 1) + (k - h));
}

// out[k] += (x_b0 + x_b1 + ...) over the dates b0 < b1 < ... of window k, one thread per (element, window): the reference's
// left-to-right reduce() (PFML_Search_Coef.py:37-46), formed from the first term, then added to the running sum
__global__ void bucket_accumulate_kernel(int B, long elems, const double* __restrict__ in, const int* __restrict__ bucket,
                                         int n_buckets, double* __restrict__ out) {
  const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y;
  if (e >= elems) return;
  bool any = false;
  double acc = 0.0;
  for (int b = 0; b < B; b++) {
    if (bucket[b] != k) continue;
    const double x = in[(long)b * elems + e];
    acc = any ? acc + x : x;
    any = true;
  }
  if (any) out[(long)k * elems + e] += acc;
}

__global__ void bucket_cumsum_kernel(int n_buckets, long elems, double* __restrict__ x) {
  const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= elems) return;
  double acc = x[e];
  for (int k = 1; k < n_buckets; k++) {
    acc += x[(long)k * elems + e];
    x[(long)k * elems + e] = acc;
  }
}

// System matrices of one p for all (year, lambda): Wk[y * n_l + l] = D_cum[y][idx, idx] / n_y + lambda_l I, padded
// (problem = flat (year, lambda) index prob0 + blockIdx: a rank of a sharded grid works on a contiguous range of them)
__global__ void ridge_build_kernel(int P, int p, int p_max, int NPp, int n_l, const double* __restrict__ d_cum,
                                   const double* __restrict__ n_cum, const double* __restrict__ l_vec, double* __restrict__ Wk, int prob0) {
  const int prob = blockIdx.z, r = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NPp) return;
  const int y = (prob0 + prob) / n_l, l = (prob0 + prob) % n_l;
  const int q = p + 1;
  double v;
  if (r < q && c < q) {
    v = d_cum[((long)y * P + feat_index(r, p, p_max)) * P + feat_index(c, p, p_max)] / n_cum[y];
    if (r == c) v += l_vec[l];
  } else {
    v = (r == c) ? 1.0 : 0.0;
  }
  Wk[((long)prob * NPp + r) * NPp + c] = v;
}

// beta = X rhs  (first pass)  or  beta += X (rhs - M beta)  (refinement), one CTA per problem
__global__ void __launch_bounds__(256)
ridge_apply_kernel(int P, int p, int p_max, int NPp, int n_l, const double* __restrict__ d_cum, const double* __restrict__ r_cum,
                   const double* __restrict__ n_cum, const double* __restrict__ l_vec, const double* __restrict__ X,
                   double* __restrict__ beta, int refine, int prob0) {
  extern __shared__ double sh[];          // rhs / residual [q], beta [q]
  const int prob = blockIdx.x, y = (prob0 + prob) / n_l, l = (prob0 + prob) % n_l, q = p + 1;
  double* res = sh;
  double* bt = sh + q;
  const double inv_n = 1.0 / n_cum[y];
  double* out = beta + (long)(prob0 + prob) * q;
  for (int i = threadIdx.x; i < q; i += blockDim.x) {
    res[i] = r_cum[(long)y * P + feat_index(i, p, p_max)] / n_cum[y];
    bt[i] = refine ? out[i] : 0.0;
  }
  __syncthreads();
  if (refine) {
    // residual = rhs - (D_sub / n + lambda I) beta, from the un-inverted matrix
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int i = warp; i < q; i += nw) {
      const double* drow = d_cum + ((long)y * P + feat_index(i, p, p_max)) * P;
      double s = 0.0;
      for (int j = lane; j < q; j += 32) s += (drow[feat_index(j, p, p_max)] / n_cum[y]) * bt[j];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) res[i] = res[i] - (s + l_vec[l] * bt[i]);
    }
    __syncthreads();
  }
  (void)inv_n;
  const double* Xp = X + (long)prob * NPp * NPp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int i = warp; i < q; i += nw) {
    const double* xr = Xp + (long)i * NPp;
    double s = 0.0;
    for (int j = lane; j < q; j += 32) s += xr[j] * res[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[i] = bt[i] + s;
  }
}

}  // namespace

namespace pfml {
// the grid restricted to the flat (year, lambda) problems [prob_lo, prob_hi) of every p; beta / status are indexed as in
// the full grid, entries of other problems are left untouched
int ridge_grid_range(const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max, const int* p_vec_host, int n_p,
                     const double* l_vec, int n_l, double* beta, int* status, void* workspace, size_t workspace_bytes, cudaStream_t s,
                     int prob_lo, int prob_hi) {
  if (!r_cum || !d_cum || !n_cum || !p_vec_host || !l_vec || !beta || !workspace || Y <= 0 || n_p <= 0 || n_l <= 0) {
    set_error("pfml_ridge_grid: null/empty argument");
    return -1;
  }
  if (workspace_bytes < pfml_ridge_grid_workspace_bytes(Y, p_max, n_l)) {
    set_error("pfml_ridge_grid: workspace too small");
    return -3;
  }
  const int nprob = Y * n_l;
  if (prob_lo < 0 || prob_hi > nprob || prob_lo > prob_hi) {
    set_error("pfml_ridge_grid: problem range [%d, %d) outside [0, %d)", prob_lo, prob_hi, nprob);
    return -2;
  }
  for (int ip = 0; ip < n_p; ip++) {
    const int p = p_vec_host[ip];
    if (p < 2 || (p & 1) || p > p_max) {
      set_error("pfml_ridge_grid: p = %d must be even and in [2, p_max = %d]", p, p_max);
      return -2;
    }
  }
  ProfScope ps(T_RIDGE, s);
  const int P = p_max + 1, mine = prob_hi - prob_lo;
  const size_t per = ((size_t)nprob * pfml_padded_n(P) * pfml_padded_n(P) * sizeof(double) + 255) / 256 * 256;
  double* Wk = (double*)workspace;
  double* X = (double*)((char*)workspace + per);
  double* T1 = (double*)((char*)workspace + 2 * per);
  long off = 0;
  for (int ip = 0; ip < n_p; ip++) {
    const int p = p_vec_host[ip], q = p + 1;
    if (mine > 0) {
      int* st = status ? status + (long)ip * nprob + prob_lo : nullptr;
      if (st) PFML_CUDA_OK(cudaMemsetAsync(st, 0, sizeof(int) * mine, s));
      const int NPp = pfml_padded_n(q);
      ridge_build_kernel<<<dim3(ceil_div(NPp, 128), NPp, mine), 128, 0, s>>>(P, p, p_max, NPp, n_l, d_cum, n_cum, l_vec, Wk, prob_lo);
      PFML_CHECK_LAUNCH();
      PFML_TRY(spd_inverse(mine, NPp, Wk, X, T1, st, s));
      for (int pass = 0; pass < 2; pass++) {
        ridge_apply_kernel<<<mine, 256, 2 * q * sizeof(double), s>>>(P, p, p_max, NPp, n_l, d_cum, r_cum, n_cum, l_vec, X, beta + off,
                                                                     pass, prob_lo);
        PFML_CHECK_LAUNCH();
      }
    }
    off += (long)nprob * q;
  }
  return 0;
}
}  // namespace pfml

extern "C" {

int pfml_bucket_accumulate(const double* r_tilde, const double* denom, const int* bucket, int B, int P, int n_buckets,
                           double* r_sum, double* d_sum, void* stream) {
  if (!r_tilde || !denom || !bucket || !r_sum || !d_sum || B <= 0 || P <= 0 || n_buckets <= 0) {
    set_error("pfml_bucket_accumulate: null/empty argument");
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(T_RIDGE, s);
  PFML_TRY(bucket_accumulate(B, P, r_tilde, bucket, n_buckets, r_sum, s));
  PFML_TRY(bucket_accumulate(B, (long)P * P, denom, bucket, n_buckets, d_sum, s));
  return 0;
}

int pfml_bucket_cumsum(double* r_sum, double* d_sum, int P, int n_buckets, void* stream) {
  if (!r_sum || !d_sum || P <= 0 || n_buckets <= 0) {
    set_error("pfml_bucket_cumsum: null/empty argument");
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  bucket_cumsum_kernel<<<ceil_div(P, 256), 256, 0, s>>>(n_buckets, P, r_sum);
  PFML_CHECK_LAUNCH();
  bucket_cumsum_kernel<<<ceil_div((long)P * P, 256), 256, 0, s>>>(n_buckets, (long)P * P, d_sum);
  PFML_CHECK_LAUNCH();
  return 0;
}

size_t pfml_ridge_grid_workspace_bytes(int Y, int p_max, int n_l) {
  if (Y <= 0 || p_max <= 0 || n_l <= 0) return 0;
  const size_t NPp = pfml_padded_n(p_max + 1);
  return 3 * (((size_t)Y * n_l * NPp * NPp * sizeof(double) + 255) / 256 * 256);
}

// rank r of `world` solves the flat (year, lambda) problems [lo, hi) of every p: contiguous in beta, balanced to +-1 problem
void pfml_ridge_shard_range(int nprob, int rank, int world, int* lo, int* hi) {
  if (world < 1) world = 1;
  if (rank < 0) rank = 0;
  *lo = (int)((long)nprob * rank / world);
  *hi = (int)((long)nprob * (rank + 1) / world);
}

int pfml_ridge_grid_range(const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max, const int* p_vec_host,
                          int n_p, const double* l_vec, int n_l, int prob_lo, int prob_hi, double* beta, int* status, void* workspace,
                          size_t workspace_bytes, void* stream) {
  return ridge_grid_range(r_cum, d_cum, n_cum, Y, p_max, p_vec_host, n_p, l_vec, n_l, beta, status, workspace, workspace_bytes,
                          (cudaStream_t)stream, prob_lo, prob_hi);
}

int pfml_ridge_grid(const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max, const int* p_vec_host,
                    int n_p, const double* l_vec, int n_l, double* beta, int* status, void* workspace, size_t workspace_bytes,
                    void* stream) {
  return ridge_grid_range(r_cum, d_cum, n_cum, Y, p_max, p_vec_host, n_p, l_vec, n_l, beta, status, workspace, workspace_bytes,
                          (cudaStream_t)stream, 0, Y * n_l);
}

}  // extern "C"
