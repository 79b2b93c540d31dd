// m_tilde_0 = 0.5 (sigma_hat - sqrtm(sigma_hat^2 - 4I)),  sigma_hat = x + 2I   (General_functions.py:955-956)
// without a matrix square root, for x in the factor form the date-step has:
//
//     x = E + U~ U~',   E = diag((gamma/w) ivol_i / lambda_i),   U~ = Lambda^-1/2 L R,   R R' = (gamma/w) F.
//
// g(xi) = 0.5 (xi + 2 - sqrt(xi^2 + 4 xi)) is the Stieltjes transform of the semicircle law on [0, 4]:
//
//     g(xi) = (1 / 2 pi) int_0^4 sqrt(u (4 - u)) / (xi + u) du ,
//
// so m_tilde_0 = g(x) = sum_j w_j (x + u_j I)^-1 for any quadrature (u_j, w_j) of that integral.  With
// u = 4 sigmoid(s) the integrand is analytic in the strip |Im s| < pi for EVERY xi > 0 and decays like
// exp(-1.5 |s|), so the plain trapezoid rule with h = 0.5 on s in [-36, 24] (121 nodes) is accurate to 5e-16
// absolute for all xi in [1e-8, 1e3] (tools/proto_stieltjes.py).  Each resolvent is a diagonal plus a rank-K
// term (Woodbury):
//
//     (E + u I + U~ U~')^-1 = D - (D U~) T^-1 (D U~)',   D = (E + u I)^-1,   T = I + U~' D U~  (K x K, >= I),
//
// hence  m_tilde_0 = diag(sum_j w_j D_j) - G G',  G = [ sqrt(w_j) D_j U~ L_j^-T ]_j  (n x 121 K),  T_j = L_j L_j'.
// The dense work is ONE symmetric DMMA contraction of depth 121 K instead of ten (inverse + GEMM) Denman-Beavers
// steps; the dense-Sigma entry point pfml_mfunc_batched keeps Denman-Beavers.
#include "kernels.cuh"

namespace pfml {

namespace {

__device__ __forceinline__ void sq_node(int j, double& u, double& w) {
  const double s = SQ_SMIN + SQ_H * j;
  const double sg = 1.0 / (1.0 + exp(-s)), sgm = 1.0 / (1.0 + exp(s));
  const double t = sg * sgm;
  u = 4.0 * sg;
  w = SQ_H * (8.0 / 3.14159265358979323846) * t * sqrt(t);
}

// Per date: R = chol((gamma/w) F) (semi-definite tolerant), U~ = (Lambda^-1/2 L) R, E.
__global__ void __launch_bounds__(256)
stj_prep_kernel(int NP, int K, const int* __restrict__ nsel, BatchVecs v, double* __restrict__ Ut, double* __restrict__ Ev) {
  __shared__ double R[KP][KP + 1];
  const int b = blockIdx.x, tid = threadIdx.x, n = nsel[b];
  const double iw = v.scal[(long)b * NSCAL + 1];
  for (int e = tid; e < KP * KP; e += blockDim.x) R[e / KP][e % KP] = v.Fg[(long)b * KP * KP + e] * iw;
  __syncthreads();
  if (tid < 32) {
    double dmax = 0.0;
    for (int c = 0; c < K; c++) dmax = fmax(dmax, R[c][c]);
    const int r = tid;
    for (int c = 0; c < K; c++) {
      const double piv = R[c][c];
      const bool ok = piv > 1e-14 * dmax;          // rank-deficient factor covariance: drop the direction
      const double d = ok ? sqrt(piv) : 0.0;
      __syncwarp();
      if (r == c) R[c][c] = d;
      if (r > c && r < K) R[r][c] = ok ? R[r][c] / d : 0.0;
      __syncwarp();
      if (r > c && r < K)
        for (int cc = c + 1; cc <= r; cc++) R[r][cc] -= R[r][c] * R[cc][c];
      __syncwarp();
    }
    for (int c = r + 1; c < KP; c++) R[r][c] = 0.0;   // strict upper triangle
  }
  __syncthreads();
  for (int e = tid; e < NP * KP; e += blockDim.x) {
    const int i = e / KP, c = e % KP;
    double acc = 0.0;
    if (i < n && c < K) {
      const double* lt = v.Lt + ((long)b * NP + i) * KP;
      for (int a = c; a < K; a++) acc += lt[a] * R[a][c];
    }
    Ut[((long)b * NP + i) * KP + c] = acc;
  }
  for (int i = tid; i < NP; i += blockDim.x) Ev[(long)b * NP + i] = (i < n) ? v.dvl[(long)b * NP + i] * iw : 1.0;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// One CTA (4 warps) per (node j, date b):
//   T_j = I + U~' D_j U~          32 x 32 x n contraction on the DMMA pipe (warp w owns rows 8w..8w+7 of T)
//   T_j = L L', L^-1               one warp, shared memory
//   G_j = (D_j U~) (sqrt(w_j) L^-T)  n x 32 x 32 on the DMMA pipe, written straight to the stack G
// U~ is streamed through shared memory in chunks of 64 rows (stride 36 doubles: conflict-free for both the
// k-outer fragment pattern of the first product and the k-inner pattern of the second).
__global__ void __launch_bounds__(128)
stj_node_kernel(int NP, int K, int KQ, long ldG, const double* __restrict__ Ut, const double* __restrict__ Ev, double* __restrict__ G) {
  constexpr int CH = 64, S = KP + 4;
  __shared__ double T[KP][S];
  __shared__ double Li[KP][S];
  __shared__ double Uc[CH][S];
  __shared__ double dc[CH];
  const int j = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  double u, w;
  sq_node(j, u, w);
  const double* U = Ut + (long)b * NP * KP;
  const double* E = Ev + (long)b * NP;
  // ---- T = I + sum_i d_i u_i u_i' -----------------------------------------------------------------
  double acc[4][2];
#pragma unroll
  for (int q = 0; q < 4; q++) acc[q][0] = acc[q][1] = 0.0;
  for (int i0 = 0; i0 < NP; i0 += CH) {
    __syncthreads();
    for (int e = tid; e < CH * KP; e += 128) Uc[e / KP][e % KP] = U[(long)(i0 + e / KP) * KP + e % KP];
    if (tid < CH) dc[tid] = 1.0 / (E[i0 + tid] + u);
    __syncthreads();
#pragma unroll 4
    for (int k0 = 0; k0 < CH; k0 += 4) {
      const double af = Uc[k0 + t][8 * warp + g] * dc[k0 + t];
#pragma unroll
      for (int q = 0; q < 4; q++) dmma884(acc[q][0], acc[q][1], af, Uc[k0 + t][8 * q + g]);
    }
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int r = 8 * warp + g, c = 8 * q + 2 * t;
    T[r][c] = acc[q][0] + (r == c ? 1.0 : 0.0);
    T[r][c + 1] = acc[q][1] + (r == c + 1 ? 1.0 : 0.0);
    Li[r][c] = 0.0;
    Li[r][c + 1] = 0.0;
  }
  __syncthreads();
  // ---- Cholesky T = L L' (lower, in place) and L^-1, one warp -------------------------------------
  if (tid < 32) {
    const int r = tid;
    for (int c = 0; c < K; c++) {
      const double d = sqrt(T[c][c]);            // T >= I: always positive definite
      __syncwarp();
      if (r == c) T[c][c] = d;
      if (r > c && r < K) T[r][c] /= d;
      __syncwarp();
      if (r > c && r < K)
        for (int cc = c + 1; cc <= r; cc++) T[r][cc] -= T[r][c] * T[cc][c];
      __syncwarp();
    }
    if (r < K) {                                 // column r of L^-1 by forward substitution
      const int c = r;
      Li[c][c] = 1.0 / T[c][c];
      for (int rr = c + 1; rr < K; rr++) {
        double s = 0.0;
        for (int a = c; a < rr; a++) s += T[rr][a] * Li[a][c];
        Li[rr][c] = -s / T[rr][rr];
      }
    }
  }
  __syncthreads();
  // ---- G_j = (D_j U~) (sqrt(w) L^-T): B fragments (k = a, n = c) = sqrt(w) Li[c][a], kept in registers ----
  const double sw = sqrt(w);
  double bf[8][4];
#pragma unroll
  for (int kk = 0; kk < 8; kk++)
#pragma unroll
    for (int q = 0; q < 4; q++) bf[kk][q] = sw * Li[8 * q + g][4 * kk + t];
  double* Gj = G + (long)b * NP * ldG + (long)j * KQ;
  for (int i0 = 0; i0 < NP; i0 += CH) {
    __syncthreads();
    for (int e = tid; e < CH * KP; e += 128) Uc[e / KP][e % KP] = U[(long)(i0 + e / KP) * KP + e % KP];
    if (tid < CH) dc[tid] = 1.0 / (E[i0 + tid] + u);
    __syncthreads();
#pragma unroll
    for (int mi = 0; mi < 2; mi++) {
      const int r = 16 * warp + 8 * mi + g;
      const double dr = dc[r];
      double c4[4][2];
#pragma unroll
      for (int q = 0; q < 4; q++) c4[q][0] = c4[q][1] = 0.0;
#pragma unroll
      for (int kk = 0; kk < 8; kk++) {
        const double af = Uc[r][4 * kk + t] * dr;
#pragma unroll
        for (int q = 0; q < 4; q++) dmma884(c4[q][0], c4[q][1], af, bf[kk][q]);
      }
      double* row = Gj + (long)(i0 + r) * ldG;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int c = 8 * q + 2 * t;
        if (c + 1 < KQ) *reinterpret_cast<double2*>(row + c) = make_double2(c4[q][0], c4[q][1]);
        else if (c < KQ) row[c] = c4[q][0];
      }
    }
  }
}

// diagonal term sum_j w_j / (E_i + u_j), added in place; guard on the quadrature's design range
__global__ void stj_diag_kernel(int NP, const int* __restrict__ nsel, const double* __restrict__ Ev, double* __restrict__ mt,
                                int* __restrict__ status) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= NP) return;
  const double e = Ev[(long)b * NP + i];
  double s = 0.0;
  for (int j = 0; j < SQ_NODES; j++) {
    double u, w;
    sq_node(j, u, w);
    s += w / (e + u);
  }
  mt[((long)b * NP + i) * NP + i] += s;
  if (status && i < nsel[b] && !(e >= 1e-8)) atomicOr(status + b, 4);
}

}  // namespace

long stj_ldg(int K) { return (long)SQ_NODES * ((K + 1) / 2 * 2); }

int mt0_stieltjes(int B, int NP, int K, const int* nsel, const BatchVecs& v, double* Ut, double* Ev, double* G, double* mt,
                  int* status, cudaStream_t s) {
  const int KQ = (K + 1) / 2 * 2;
  const long ldG = stj_ldg(K);
  stj_prep_kernel<<<B, 256, 0, s>>>(NP, K, nsel, v, Ut, Ev);
  PFML_CHECK_LAUNCH();
  stj_node_kernel<<<dim3(SQ_NODES, B), 128, 0, s>>>(NP, K, KQ, ldG, Ut, Ev, G);
  PFML_CHECK_LAUNCH();
  GemmP g;
  g.A = G; g.lda = ldG; g.sA = (long)NP * ldG; g.a_kinner = true;
  g.B = G; g.ldb = ldG; g.sB = (long)NP * ldG; g.b_kinner = true;
  g.D = mt; g.ldd = NP; g.sD = (long)NP * NP;
  g.M = NP; g.N = NP; g.K = (int)ldG; g.alpha = -1.0; g.sym = 1;
  PFML_TRY(gemm(g, B, s));
  stj_diag_kernel<<<dim3(ceil_div(NP, 128), B), 128, 0, s>>>(NP, nsel, Ev, mt, status);
  PFML_CHECK_LAUNCH();
  return 0;
}

}  // namespace pfml
