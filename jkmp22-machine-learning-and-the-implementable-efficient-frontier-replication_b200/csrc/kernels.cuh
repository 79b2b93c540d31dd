// Launchers of the non-GEMM kernels of the PFML date-step (kernels.cu).  Internal header.
#pragma once
#include "common.cuh"

namespace pfml {

// Device view of the resident panel (mirrors pfml_panel_t of include/pfml_b200.h).
struct PanelView {
  int T, N, F, p, K;
  const double* rff;        // [T*N][p]  [cos | sin]
  const double* vol_scale;  // [T*N]
  const double* gt;         // [T*N]  (1+tr_ld0)/(1+mu_ld0), NaN where missing
  const double* lam;        // [T*N]
  const double* ret_ld1;    // [T*N]
  const double* fct_load;   // [T*N][K]
  const double* fct_cov;    // [T][K][K]
  const double* ivol;       // [T*N]
  const double* wealth;     // [T]
  const double* rf;         // [T]
  double gamma_rel, mu;
  int lb, iterations;
};

constexpr int KP = 32;         // factor count padded (K <= 32)
constexpr int BASE = 64;       // padding granularity of n (blocks of the recursive inversion are multiples of it)
constexpr int BASE_MAX = 128;  // largest block inverted by one CTA (register-resident Gauss-Jordan)
constexpr int NSCAL = 8;       // per-date scalars: 0 gamma/w... see prep_vectors
constexpr int NORM_CHUNKS = 16;
// trapezoid quadrature of the semicircle Stieltjes integral (stieltjes.cu): s_j = SQ_SMIN + j * SQ_H
constexpr int SQ_NODES = 121;
constexpr double SQ_H = 0.5, SQ_SMIN = -36.0;

struct BatchVecs {             // per-date vectors, all [B][NP] unless noted
  double* sl;      // sqrt(lambda)            (pad 1)
  double* isl;     // lambda^-1/2             (pad 1)
  double* g;       // [B][lb+2][NP] g_tau with NaN->1 (pad 1)
  double* g0raw;   // g_0 without NaN->1      (pad 1)
  double* rv;      // lambda^-1/2 * ret_ld1   (pad 0)
  double* dvl;     // gamma * ivol / lambda   (pad 0)
  double* Lg;      // [B][NP][KP] gathered factor loadings (pad 0)
  double* Lt;      // [B][NP][KP] lambda^-1/2 * L
  double* LF;      // [B][NP][KP] L @ fct_cov
  double* Fg;      // [B][KP][KP] gamma * fct_cov, zero padded
  double* scal;    // [B][NSCAL]: 0: w, 1: 1/w, 2: 1+rf+mu, 3: gamma
};

int prep_vectors(const PanelView& pv, int B, int NP, const int* date_t, const int* sel, const int* nsel,
                 const BatchVecs& v, cudaStream_t s);
int build_x_factor(const PanelView& pv, int B, int NP, const int* date_t, const int* sel, const int* nsel,
                   const BatchVecs& v, double* x, double* sgr, double* yv, cudaStream_t s);
int build_x_dense(int B, int n, int NP, const double* sigma_gam, const double* lam, const double* w, const double* rf,
                  double mu, double gam, double* x, double* sgr, double* yv, double* isl, double* sl, cudaStream_t s);

int standardize(const PanelView& pv, int B, int NP, int P, long ldP, const int* date_t, const int* sel, const int* nsel,
                const BatchVecs& v, double* St /*[B][lb+2][NP][ldP]*/, double* signal_t /*[B][NP][P] or null*/, bool prescale /* g_{tau-1} row scale for tau in 2..lb */,
                int ntau /* slabs 0 .. ntau-1 are written */, cudaStream_t s,
                double* signals_all = nullptr /* [B][lb+2][NP][P]: diag(1/vol) s of every slab, un-tilded */);

// n x n helpers on [B][NP][NP] buffers
int copy2(int B, long n_elem, const double* src, double* d1, double* d2, cudaStream_t s);
int db_norms(int B, int NP, const double* M, const double* Mi, double* part, double* sc /*[B][8]*/, cudaStream_t s);
int db_update_M(int B, int NP, double* M, const double* Mi, double* Wk, const double* sc, cudaStream_t s);
int mt0_from_sqrt(int B, int NP, const double* x, const double* Y, double* mt, cudaStream_t s);
int fp_build(int B, int NP, const double* x, const double* yv, const double* mt, const double* sgr, double* Wk, cudaStream_t s);
int chain_init(int B, int NP, const double* mt, const double* g1, long sg, double* Bt, double* CB, cudaStream_t s);
int emu_standardize_slices(const PanelView& pv, int B, int NP, int P, int tau0, int ntau, const int* date_t, const int* sel,
                           const int* nsel, const BatchVecs& v, int8_t* out, double* scale_out, int* err, cudaStream_t s);
int add_mat(int B, long n_elem, const double* a, const double* b, double* out, cudaStream_t s);
// a9 finish (mirror, denom = risk + tc, NaN semantics); with `bucket` the per-window sums of denom are accumulated into d_sum
int sym_finish(int B, int P, double* risk, double* tc, double* denom, const int* tc_nan /* [B] or null */, const int* bad /* [B] or null */,
               const int* bucket /* [B] or null */, int n_buckets, double* d_sum /* [n_buckets][P][P] */, cudaStream_t s);
int poison_bad_dates(int B, int P, const int* bad, double* r_tilde, double* denom, double* risk, double* tc, cudaStream_t s);
int tilde_rows(int B, int NP, int P, const double* in, const int* nsel, const double* sl, double* out, long ldo, long stride_o,
               cudaStream_t s);
int ridge_grid_range(const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max, const int* p_vec_host, int n_p,
                     const double* l_vec, int n_l, double* beta, int* status, void* workspace, size_t workspace_bytes, cudaStream_t s,
                     int prob_lo, int prob_hi);
// out[bucket[b]] += in-order sum of in[b] (ridge.cu; PFML_Search_Coef.py:77-88, :112-121)
int bucket_accumulate(int B, long elems, const double* in, const int* bucket, int n_buckets, double* out, cudaStream_t s);
int chain_sum(int B, int NP, int nk, const double* const* bk, const double* b11, const double* g0, long sg, double* wk, double* cbs,
              cudaStream_t s);
int row_scale(int B, int rows, int cols, long ld, long stride, const double* in, const double* rsv, long srs, double* out,
              long ldo, long stride_o, cudaStream_t s);
int m_from_mtilde(int B, int n, int NP, const double* mt, const double* isl, const double* sl, double* m_out, cudaStream_t s);

int base_inverse(int B, int NP, const double* src, double* dst, int off, int size, bool spd, int* status, cudaStream_t s);
int gemv_t(int B, int rows, int cols, const double* A, long ld, long sA, const double* v, long sv, double* out, long so,
           cudaStream_t s);
int check_finite_status(int B, const double* sc, int* status, cudaStream_t s);

// m_tilde_0 by quadrature + Woodbury in the factor form (stieltjes.cu)
long stj_ldg(int K);
int mt0_stieltjes(int B, int NP, int K, const int* nsel, const BatchVecs& v, double* Ut, double* Ev, double* G, double* mt,
                  int* status, cudaStream_t s);

// panel-level helpers
int gt_panel(long n, const double* tr_ld0, const double* mu_ld0, double* gt, cudaStream_t s);
int vol_scale_panel(const double* fct_load, const double* fct_cov, const double* ivol, const unsigned char* has, int T, int N,
                    int K, double* vol_scale, cudaStream_t s);

}  // namespace pfml
