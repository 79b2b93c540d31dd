/* pfml_b200.h -- C ABI of libpfml_b200.so: the Portfolio-ML (PFML) hot path of the JKMP22 replication,
 * hand-written for NVIDIA B200 (sm_100a).
 *
 * The reference (brockpat/JKMP22-...-Replication) is pure Python and has no FFI; its boundary is the set of
 * importable functions and scripts listed below.  Each entry point here replaces the arithmetic of one of
 * them; the Python host mirror (package `..._b200`, alias `pfml_b200`) keeps the reference's names and
 * argument meaning and calls these through ctypes.
 *
 * Conventions
 *   - every tensor is caller-owned, contiguous, row-major IEEE FP64 in DEVICE memory unless the name says
 *     `_host`; index arrays are int32;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); calls are asynchronous on it;
 *   - no hidden allocation: the caller passes a workspace whose size the matching *_workspace_bytes query
 *     returns.  Exception: the *_host convenience entry points keep one grow-only device arena per GPU (sized by
 *     the largest call so far, reused by later calls, never freed before process exit) plus their own streams;
 *   - return value: 0 ok; <0 bad argument / CUDA error (text via pfml_last_error(), per thread); numerical
 *     failures are reported PER DATE in the `status` array:
 *       bit 0 (1)   singular pivot in a block inverse,
 *       bit 1 (2)   matrix square root did not converge / non-finite input of m_func,
 *       bit 2 (4)   an eigenvalue of x = Sigma~ / w below 1e-8: outside the design range of the quadrature that
 *                   replaces sqrtm (General_functions.py:956),
 *       bit 3 (8)   a barrier wait inside the INT8-emulated GEMM timed out (kernel-wide: set on every date of the call),
 *       bit 4 (16)  NaN / Inf reached an operand of that GEMM for THIS date: its r_tilde, denom, risk and tc are
 *                   returned as NaN (what the reference's FP64 arithmetic yields), the other dates of the batch are
 *                   unaffected and bit-identical to a batch without the bad date;
 *   - never throws, never exits.  Threading: calls on different streams / devices may run concurrently from
 *     different host threads; the launch counter, the emulation mode (read once at the start of each call) and the
 *     one-time kernel configuration are atomic.  The *_host entry points serialise per GPU (one arena per GPU),
 *     not across GPUs, and restore the caller's current device.  The profiling hooks (pfml_prof_*) are for a
 *     single measuring thread.
 */
#ifndef PFML_B200_H
#define PFML_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* pfml_version(void);
const char* pfml_last_error(void);

/* Device-resident panel for a block of consecutive months t = 0..T-1 (dense: row index = t*N + i). */
typedef struct {
  int T, N, F, p, K;        /* months, stocks, characteristics, RFF features (cos+sin), factors (K <= 32) */
  const double* rff;        /* [T*N][p]  [cos | sin], output of pfml_rff_sincos  (PFML_Input_Data.py:254) */
  const double* vol_scale;  /* [T*N]     sqrt(diag(Sigma)) with eom-median fill   (PFML_Input_Data.py:274-305) */
  const double* gt;         /* [T*N]     (1+tr_ld0)/(1+mu_ld0), NaN where missing (PFML_Input_Data.py:404,:458) */
  const double* lam;        /* [T*N]     Kyle's lambda                            (PFML_Input_Data.py:116-120) */
  const double* ret_ld1;    /* [T*N]     lead returns                             (PFML_Input_Data.py:330) */
  const double* fct_load;   /* [T*N][K]  Barra loadings                           (Barra_Cov.pkl 'fct_load') */
  const double* fct_cov;    /* [T][K][K] factor covariance                        (Barra_Cov.pkl 'fct_cov') */
  const double* ivol;       /* [T*N]     idiosyncratic variance                   (Barra_Cov.pkl 'ivol_vec') */
  const double* wealth;     /* [T]       (wealth_processed.csv) */
  const double* rf;         /* [T]       decimal risk-free rate (FF_RF_monthly.csv / 100) */
  double gamma_rel, mu;     /* pf_set['gamma_rel'], pf_set['mu']  (General_functions.py:103-108) */
  int lb, iterations;       /* pf_set['lb_hor'] = 11; iterations = 10 (PFML_Input_Data.py:215) */
} pfml_panel_t;

/* a1 -- random Fourier features.  Replaces rff(X, W=W) + the [cos | sin] concatenation,
 * PFML_Input_Data.py:159-185, :254.  X [rows][F], W [F][p_half] -> out [rows][2*p_half]. */
int pfml_rff_sincos(const double* X, const double* W, long rows, int F, int p_half, double* out, void* stream);

/* a2 -- vol_scale = sqrt(diag(L F L' + diag(ivol))) per (month, stock) without forming Sigma, NaN where
 * `has` is 0, then the per-month median fill.  Replaces PFML_Input_Data.py:274-305 (create_cov + diag,
 * General_functions.py:869).  has may be NULL (all present). */
int pfml_vol_scale(const double* fct_load, const double* fct_cov, const double* ivol, const unsigned char* has,
                   int T, int N, int K, double* vol_scale, void* stream);

/* g_t = (1 + tr_ld0) / (1 + mu_ld0) per (month, stock), NaN kept (PFML_Input_Data.py:404, :458). */
int pfml_growth_ratio(const double* tr_ld0, const double* mu_ld0, long n, double* gt, void* stream);

/* a3..a9 -- B date-steps of the loop PFML_Input_Data.py:318-491 (create_cov :333, create_lambda :336,
 * m_func :345, standardise/scale :357-388, gtm chains :400-429, omega :438-462, r_tilde/risk/tc/denom
 * :470-481), one per entry of date_t.  sel [B][NP] holds the stock indices (ascending ids) of each date's
 * universe, n_sel [B] their count, NP = pfml_padded_n(n_max).  Outputs are in feat_all order
 * (constant, cos 1..p/2, sin 1..p/2; PFML_Input_Data.py:221,:484,:489), P = p + 1:
 *   r_tilde [B][P], denom/risk/tc [B][P][P], signal_t [B][NP][P] (rows >= n_sel[b] are zero).
 * risk, tc, signal_t may be NULL.  status [B] int32 (may be NULL) receives numerical-failure bits.
 * The n x n and n x P products run as FP64-equivalent GEMMs on the INT8 tensor cores (see pfml_dgemm_emulated;
 * pfml_set_emulation(0) or PFML_EMULATE_FP64=0 selects the FP64 DMMA kernel for them). */
int pfml_padded_n(int n_max);
size_t pfml_date_steps_workspace_bytes(const pfml_panel_t* panel, int B, int n_max);
int pfml_date_steps(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                    double* r_tilde, double* denom, double* risk, double* tc, double* signal_t, int* status,
                    void* workspace, size_t workspace_bytes, void* stream);

/* a3..a9 + a11 -- the same date-steps with the expanding-window sums of PFML_Search_Coef.py:77-88, :112-121 fused in:
 * bucket [B] (device int32) is the window of each date (0 = burn-in, k = k-th hp year, outside [0, n_buckets): none);
 * r_sum [n_buckets][P] and d_sum [n_buckets][P][P] are ACCUMULATED into (zero them before the first call): every
 * window receives the in-order (left-to-right, like the reference's reduce(+), :37-46) sum of its dates of this
 * call.  On the INT8 path the sums of denom are formed in the a9 finish kernel that writes denom, so denom is not
 * read again.  Partial sums of different ranks are then combined with pfml_comm_allreduce_sums. */
int pfml_date_steps_bucketed(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                             const int* bucket, int n_buckets, double* r_sum, double* d_sum, double* r_tilde, double* denom,
                             double* risk, double* tc, double* signal_t, int* status, void* workspace, size_t workspace_bytes,
                             void* stream);

/* a3..a9 (+ a11) for G bandwidths g of the SAME dates in one call (PFML_Input_Data.py:225 loops g outside the date loop
 * and recomputes everything per g; quirk Q10: Sigma, Lambda, m_func, the gtm chains and the two `const` matrices of
 * :400-456 do not depend on g).  rff_g is a HOST array of G device panels [T*N][p] (one W per g; panel->rff is ignored),
 * the output arguments are HOST arrays of G device pointers (risk_g / tc_g / signal_t_g and their entries may be NULL;
 * bucket NULL = no window sums).  m, the chain, the sums of the chain and the two inverses are computed once; per g only
 * a6, the n x P sums, the two solve products and a9 run.  Every g's outputs are bit-identical to a single-g
 * pfml_date_steps call on that g's panel.  Shapes the shared path does not cover (FP64 DMMA mode, universes beyond the
 * INT8 depth limit) fall back to one full pass per g.  status [B] ORs the flags of all g.  Workspace as pfml_date_steps. */
int pfml_date_steps_multi_g(const pfml_panel_t* panel, int G, const double* const* rff_g, int B, int n_max, const int* date_t,
                            const int* sel, const int* n_sel, const int* bucket, int n_buckets, double* const* r_sum_g,
                            double* const* d_sum_g, double* const* r_tilde_g, double* const* denom_g, double* const* risk_g,
                            double* const* tc_g, double* const* signal_t_g, int* status, void* workspace, size_t workspace_bytes,
                            void* stream);

/* Stage-level entry points: the same kernels as pfml_date_steps, stopped / started at a stage boundary, so that
 * each stage can be called, parity-tested and profiled on its own.  All take the workspace of
 * pfml_date_steps_workspace_bytes and the same date_t / sel / n_sel arrays; matrices are [B][NP][.] with rows
 * >= n_sel[b] zero, columns in feat_all order.
 *   pfml_standardize_scale     a6: signals[tau] = diag(1/vol_scale) s for tau = 0..lb+1 (PFML_Input_Data.py:357-388)
 *                              -> signals [B][lb+2][NP][P]
 *   pfml_omega_batched         a3..a8: omega = solve(const, sum A_theta S_theta), omega_chg = omega - G_0 omega_l1
 *                              (PFML_Input_Data.py:400-462) -> omega, omega_chg [B][NP][P]; m (may be NULL) receives
 *                              m_func's result [B][NP][NP] (General_functions.py:919-962)
 *   pfml_rtilde_denom_batched  a9 from GIVEN omega / omega_chg [B][NP][P] (e.g. the oracle's): r_tilde = omega' r,
 *                              risk = gamma omega' Sigma omega, tc = w omega_chg' Lambda omega_chg, denom = risk + tc
 *                              (PFML_Input_Data.py:470-481); risk / tc may be NULL */
int pfml_standardize_scale(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                           double* signals, void* workspace, size_t workspace_bytes, void* stream);
int pfml_omega_batched(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                       double* omega, double* omega_chg, double* m, int* status, void* workspace, size_t workspace_bytes,
                       void* stream);
int pfml_rtilde_denom_batched(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                              const double* omega, const double* omega_chg, double* r_tilde, double* denom, double* risk,
                              double* tc, int* status, void* workspace, size_t workspace_bytes, void* stream);

/* a5 -- m_func(w, mu, rf, sigma_gam, gam, K_Lambda, iterations), General_functions.py:919-962, batched over
 * B problems of equal size n: sigma_gam [B][n][n], lam [B][n] (diagonal of K_Lambda), w [B], rf [B] -> m [B][n][n]. */
size_t pfml_mfunc_workspace_bytes(int B, int n);
int pfml_mfunc_batched(int B, int n, const double* sigma_gam, const double* lam, const double* w, const double* rf,
                       double mu, double gam, int iterations, double* m_out, int* status, void* workspace,
                       size_t workspace_bytes, void* stream);

/* a11 -- expanding-window sums of PFML_Search_Coef.py:37-46, :77-88, :112-121.  bucket[b] (device int32) is the
 * window of date b (0 = burn-in, k = k-th hp year, -1 = unused); r_sum [n_buckets][P] and d_sum [n_buckets][P][P]
 * are ACCUMULATED into (zero them first; partial sums of different ranks are then added with an all-reduce).
 * Sums run over dates in order, one thread per element: the reference's left-to-right reduce().
 * pfml_bucket_cumsum turns bucket sums into the running sums the year loop keeps (:117, :121). */
int pfml_bucket_accumulate(const double* r_tilde, const double* denom, const int* bucket, int B, int P, int n_buckets,
                           double* r_sum, double* d_sum, void* stream);
int pfml_bucket_cumsum(double* r_sum, double* d_sum, int P, int n_buckets, void* stream);

/* a12/a13 -- the ridge grid of PFML_Search_Coef.py:125-135: for every year y < Y, every p in p_vec (HOST array,
 * even, <= p_max) and every lambda in l_vec (device): beta = solve(D_sub / n_y + lambda I, r_sub / n_y), where
 * D = d_cum[y], r = r_cum[y], n_y = n_cum[y] (device, number of dates summed) and `sub` is the index set of
 * pfml_feat_fun(p) (General_functions.py:837-844) inside the p_max layout.  beta is packed per p:
 * [p][Y][n_l][p+1]; status [n_p][Y][n_l] (may be NULL). */
size_t pfml_ridge_grid_workspace_bytes(int Y, int p_max, int n_l);
int pfml_ridge_grid(const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max, const int* p_vec_host,
                    int n_p, const double* l_vec, int n_l, double* beta, int* status, void* workspace,
                    size_t workspace_bytes, void* stream);

/* e -- the exchange steps of the path over NCCL (one process per GPU; the reference is single-process and has no
 * counterpart: its year loop PFML_Search_Coef.py:102-137 runs all of this serially).
 *   communicator   rank 0 calls pfml_comm_unique_id (128 bytes) and hands the id to the other ranks by any host-side
 *                  means (torch.distributed broadcast, a file, MPI); every rank then calls pfml_comm_init with its
 *                  device.  The handle is opaque; NCCL is opened at run time from the copy already in the process.
 *   pfml_comm_allreduce_sums   in-place FP64 sum all-reduce of the per-window partial sums every rank accumulated over
 *                  ITS dates (pfml_date_steps_bucketed / pfml_bucket_accumulate): r_sum [n_buckets][P],
 *                  d_sum [n_buckets][P][P], counts [n_buckets] (dates per window, may be NULL)
 *                  (PFML_Search_Coef.py:77-88, :112-121).  comm NULL or a single rank: no-op.
 *   pfml_ridge_grid_sharded    pfml_ridge_grid with the (year, lambda) problems of every p dealt over the ranks in
 *                  contiguous balanced ranges (pfml_ridge_shard_range) and beta / status all-gathered, so every rank
 *                  returns the full grid (PFML_Search_Coef.py:125-135).  Same arguments and workspace as pfml_ridge_grid.
 *   pfml_ridge_grid_range      one rank's share without a communicator: the flat (year, lambda) problems
 *                  [prob_lo, prob_hi) of every p; the other entries of beta / status are left untouched (what
 *                  pfml_ridge_grid_sharded runs before its all-gather) */
int pfml_comm_unique_id(void* id128);
int pfml_comm_init(const void* id128, int rank, int world, int device, void** comm);
int pfml_comm_rank(void* comm);
int pfml_comm_world(void* comm);
int pfml_comm_nccl_version(void);
int pfml_comm_destroy(void* comm);
int pfml_comm_allreduce_sums(void* comm, double* r_sum, double* d_sum, double* counts, int P, int n_buckets, void* stream);
void pfml_ridge_shard_range(int nprob, int rank, int world, int* lo, int* hi);
int pfml_ridge_grid_range(const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max, const int* p_vec_host,
                          int n_p, const double* l_vec, int n_l, int prob_lo, int prob_hi, double* beta, int* status,
                          void* workspace, size_t workspace_bytes, void* stream);
int pfml_ridge_grid_sharded(void* comm, const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max,
                            const int* p_vec_host, int n_p, const double* l_vec, int n_l, double* beta, int* status,
                            void* workspace, size_t workspace_bytes, void* stream);

/* a15 -- daily cross-sectional OLS of "Estimate Covariance Matrix.py":214-241.  Rows are sorted by day; day d owns
 * rows [day_ptr[d], day_ptr[d+1]) of X [rows][K] and y [rows] (day_ptr: device int64 [n_days+1]).  Outputs
 * coef [n_days][K] (= fct_ret) and the residuals [rows].  A factor whose column is empty that day (exactly
 * singular X'X, where the reference falls back to pinv, :226-229) gets coefficient 0 and status bit 0. */
int pfml_xs_ols_daily(const double* X, const double* y, const long* day_ptr, int n_days, int K, double* coef,
                      double* resid, int* status, void* stream);

/* a16 -- EWMA factor covariance per calc date, "Estimate Covariance Matrix.py":297-335 with weighted_cor_wt /
 * weighted_cov_wt of General_functions.py:745-835: window = the last min(obs, end_idx[c]+1) rows of fct_ret
 * [D][K] ending at row end_idx[c]; weights = the tails of w_cor / w_var [obs] (:280-282);
 * cov_out[c] = scale * sd_var cor sd_var (scale = 21 for the monthly conversion at :476). */
int pfml_wcov_wcor(const double* fct_ret, const int* end_idx, int n_calc, int K, int obs, const double* w_cor,
                   const double* w_var, double scale, double* cov_out, void* stream);

/* a17 -- ewma_vol(x, lam, start) of "Estimate Covariance Matrix.py":345-386 for n_series residual series stored
 * back to back (series s = x[series_ptr[s] .. series_ptr[s+1]), device int64 pointers): NaN for the first `start`
 * entries, var_start = sum x^2 / (count - 1), var_i = lam var_{i-1} + (1 - lam) x_{i-1}^2 (NaN carries). */
int pfml_ewma_vol(const double* x, const long* series_ptr, int n_series, double lam, int start, double* vol, void* stream);

/* f1 -- validation utility of PFML_hp_reals.py:90-102: obj[b][l] = r~_sub' beta - 1/2 beta' denom_sub beta for one p,
 * with beta [Y][n_l][p+1] (the ridge grid's block of that p) and date_year[b] (device int32) the hp-year index whose
 * coefficients date b is scored with (-1: not scored, obj = NaN).  `sub` is the pfml_feat_fun(p) index set. */
int pfml_validation_utility(const double* r_tilde, const double* denom, int B, int p_max, const int* date_year,
                            const double* beta, int p, int n_l, double* obj, void* stream);

/* f2 -- aim portfolio of PFML_aim_fun.py:150-160: w_aim[b][i] = signal_t[b][i][:] . beta_full[b][:], beta_full [B][P]
 * being the chosen coefficients scattered into the feat_all layout (zeros outside pfml_feat_fun(p)). */
int pfml_aim_weights(const double* signal_t, const int* n_sel, int B, int NP, int P, const double* beta_full,
                     double* w_aim, void* stream);

/* f3 -- m = m_func(...) (General_functions.py:919-962) for B dates straight from the panel's factor form, as
 * PFML_best_hps.py:186-190 calls it once per out-of-sample month: m_out [B][NP][NP].  The panel needs every field
 * except rff / vol_scale. */
size_t pfml_mfunc_dates_workspace_bytes(const pfml_panel_t* panel, int B, int n_max);
int pfml_mfunc_dates(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                     double* m_out, int* status, void* workspace, size_t workspace_bytes, void* stream);

/* f3 -- the back-test recursion pfml_w (PFML_best_hps.py:168-216) and compute_stats (:220-239) over B consecutive
 * dates: value-weighted start from me [T*N] at the first date, w = m w_start + (I - m) w_aim, next start =
 * w (1 + tr_ld1) / (1 + mu_ld1) (tr_ld1 [T*N], mu_ld1 [T]); stocks entering later start at 0.  carry [N] is scratch.
 * Outputs w_start, w [B][NP] and stats [B][5] = inv, shorting, turnover, r, tc. */
int pfml_backtest(const pfml_panel_t* panel, const double* me, const double* tr_ld1, const double* mu_ld1, int B, int n_max,
                  const int* date_t, const int* sel, const int* n_sel, const double* m, const double* w_aim, double* carry,
                  double* w_start, double* w, double* stats, void* stream);

/* General batched FP64 GEMM on the DMMA pipe, D = alpha * op(A) op(B) + beta * D0 (the reference's `@`):
 * trans_a/trans_b = 0: stored [M][K] / [K][N]; 1: stored transposed.  Batch strides in elements. */
int pfml_dgemm_batched(int trans_a, int trans_b, int M, int N, int K, double alpha, const double* A, long lda, long sA,
                       const double* B, long ldb, long sB, double beta, double* D, long ldd, long sD, int batch,
                       void* stream);

/* The same product on the INT8 tensor cores: D = alpha * A B + beta * D with A [M][K], B [K][N] row-major, computed as
 * 28 exact int8 x int8 -> int32 products of seven balanced base-256 digits per operand (tcgen05.mma kind::i8,
 * accumulators in TMEM, operands staged by bulk-TMA copies) and recombined exactly in int64, then FP64: 54 bits
 * below the largest element of each row of A / column of B are kept, measured 3e-15 relative to the DMMA product
 * at K = 512 (no reference
 * counterpart; it is how the date-step runs its n x n x P products).  Needs K % 64 == 0.  status[0] gets bit 3 on an
 * internal time-out, bit 4 if an operand holds NaN/Inf (they cannot be sliced).
 * pfml_set_emulation(mode) selects how pfml_date_steps runs its products (default 1, or env PFML_EMULATE_FP64):
 *   0 = FP64 DMMA kernel; 1 = INT8 path with all 7 digits (1e-13 of the DMMA result); 2 = INT8 path with 6 digits for the
 *   products whose errors do not compound (n x P sums, solves, a9; the chain keeps 7): 6 % faster, 7e-12 of the DMMA
 *   result at N = 500 and 2e-11 at N = 3000 (tools/emu_modes.py). */
size_t pfml_dgemm_emulated_workspace_bytes(int M, int N, int K, int batch);
int pfml_dgemm_emulated(int M, int N, int K, double alpha, const double* A, long lda, long sA, const double* B, long ldb,
                        long sB, double beta, double* D, long ldd, long sD, int batch, int* status, void* workspace,
                        size_t workspace_bytes, void* stream);
void pfml_set_emulation(int mode);

/* Batched inverse of well-conditioned n x n matrices by recursive Schur-complement blocks (np.linalg.inv at
 * General_functions.py:960).  A [B][n][n] -> Ainv [B][n][n]. */
size_t pfml_inverse_workspace_bytes(int B, int n);
int pfml_inverse_batched(int B, int n, const double* A, double* Ainv, int* status, void* workspace,
                         size_t workspace_bytes, void* stream);

/* The same inverse for callers that may not allocate device memory (no counterpart in the reference: this
 * entry point exists only for embedding the library).  Some hosts cannot call cudaMalloc where they need the
 * inverse -- inside a latency-critical section, inside a plugin whose process-wide memory pool is owned by
 * someone else, or simply without a spare 3*B*NP*NP-double buffer for the workspace above.  Here the work
 * matrices are a fixed 12 MiB array inside the library image (resident on every GPU that loads the library,
 * until the process exits), so only inputs and outputs are passed.
 *   limit   B * NP * NP <= PFML_INVERSE_STATIC_MAX_ELEMS with NP = pfml_padded_n(n); beyond it the call returns
 *           -3 with a message and launches nothing;
 *   sharing one array per GPU, reused by every call and never cleared (no step depends on old contents); calls
 *           that can overlap on one GPU race -- keep them on one stream or serialise them;
 *   set-up  the first call on each GPU resolves the array's device address; make it before any stream capture.
 * Allocates nothing: no cudaMalloc, cudaFree or allocating library routine on this path. */
#define PFML_INVERSE_STATIC_MAX_ELEMS (1 << 19)
int pfml_inverse_batched_static(int B, int n, const double* A, double* Ainv, int* status, void* stream);

/* Whole block through HOST buffers (the end-to-end call a drop-in user makes): uploads the block's inputs,
 * runs a1 + a2 + B date-steps on the device, downloads the outputs.  feats [T*N][F], W [F][p/2], tr_ld0/
 * mu_ld0/lam/ret_ld1/ivol [T*N], fct_load [T*N][K], fct_cov [T][K][K], barra_has [T*N] (uint8, may be NULL),
 * wealth/rf [T]; date_t/sel/n_sel as above but in host memory.  Output pointers are host memory (pinned
 * memory gives asynchronous copies); risk, tc, signal_t may be NULL.  date_t (lb + 1 <= t < T), n_sel (1..n_max),
 * sel (0..N-1), K, p and lb are validated on the host before anything is uploaded (-2 with a message). */
typedef struct {
  int T, N, F, p, K;
  const double *feats, *W, *tr_ld0, *mu_ld0, *lam, *ret_ld1, *fct_load, *fct_cov, *ivol, *wealth, *rf;
  const unsigned char* barra_has;
  double gamma_rel, mu;
  int lb, iterations;
} pfml_host_block_t;
int pfml_date_steps_host(const pfml_host_block_t* blk, int B, int n_max, const int* date_t, const int* sel,
                         const int* n_sel, double* r_tilde, double* denom, double* risk, double* tc,
                         double* signal_t, int* status, int chunk, int device);

/* The same call with the symmetric outputs in PACKED form: risk, tc and denom are symmetric and denom = risk + tc
 * exactly (PFML_Input_Data.py:473-481), so the upper triangles of risk and tc (row-major, row r holds columns r..P-1,
 * P (P + 1) / 2 doubles each) carry everything the reference stores per date in a third of the bytes: risk_ut, tc_ut
 * [B][P (P + 1) / 2].  With the environment variable PFML_HOST_REBUILD=1 this is also what crosses PCIe inside
 * pfml_date_steps_host, which then rebuilds the full squares in the caller's buffers with host threads
 * (PFML_HOST_UNPACK_THREADS, default = cores per local rank, at most 16) while the GPU works on the later chunks; the
 * rebuilt squares are bit-identical to the device-resident ones (opt-in: the rebuild is host-memory bound).
 * pfml_unpack_upper / pfml_unpack_reals do that rebuild for one date on the calling thread (pure host code):
 * full [P][P] from ut; risk, tc and denom = risk + tc [P][P] from the two triangles. */
int pfml_date_steps_host_packed(const pfml_host_block_t* blk, int B, int n_max, const int* date_t, const int* sel,
                                const int* n_sel, double* r_tilde, double* risk_ut, double* tc_ut, double* signal_t,
                                int* status, int chunk, int device);
int pfml_unpack_upper(const double* ut, int P, double* full);
int pfml_unpack_reals(const double* risk_ut, const double* tc_ut, int P, double* risk, double* tc, double* denom);

/* f4 -- cross-sectional percentile ranks of the characteristics, Prepare_Data.py:338-345 (+ the 0.5 imputation of :368
 * when impute_half != 0): X [T][N][F] dense, NaN = missing value / absent row; out [T][N][F] (may alias X only if the
 * caller does not need X).  Per (month, characteristic): pandas' groupby('eom').rank(pct=True) -- average rank of ties
 * over the count of non-missing values --, exact zeros put back to 0, missing -> 0.5 or NaN.  Bit-identical to pandas
 * (integer ranks, one FP64 division).  N <= 4096 rows per month. */
int pfml_percentile_rank(const double* X, int T, int N, int F, int impute_half, double* out, void* stream);

/* Measured ceiling of the INT8 tensor pipe for one instruction shape (no reference counterpart; bench.py prints it beside
 * the achieved rate of the emulated GEMM): `iters` back-to-back tcgen05.mma.kind::i8 of M = 128, N = n, K = 32 per SM on
 * operands resident in shared memory (ts_form = 0) or with A in TMEM (ts_form = 1), accumulators in TMEM, nothing else.
 * tops = 1e-12 * int8 multiply-adds * 2 per second over all SMs; clk_per_mma (may be NULL) at the nominal maximum SM
 * clock.  Allocates one int on the device; synchronises. */
int pfml_int8_mma_probe(int n, int iters, int ts_form, double* tops, double* clk_per_mma);
/* The instruction stream of ONE 32-byte k slab of the emulated GEMM (7 accumulators of 64 columns, 28 digit products),
 * repeated `iters` times per SM on resident operands: mode 0 = the production mix (three tcgen05.cp copies of A digits, 18
 * products with A in TMEM, 10 with A in shared memory); mode 1 = A digits 0..3 moved to TMEM by tcgen05.st from four stager
 * warps, double-buffered (22 + 6 products, no copy in the tensor queue).  clk_per_slab at the nominal maximum SM clock. */
int pfml_int8_pipe_probe(int iters, int mode, double* clk_per_slab, double* tops);

/* Device-time accounting with CUDA events (used by bench.py for the live roofline): enable, run, collect.
 * pfml_prof_collect synchronises the device, fills ms/flops/launches [n_tags] and clears the log.
 * Tag 0 = all DMMA GEMM launches; tags 1.. = pipeline stages (see pfml_prof_tag_name). */
void pfml_prof_enable(int on);
int pfml_prof_ntags(void);
const char* pfml_prof_tag_name(int tag);
int pfml_prof_collect(double* ms, double* flops, double* launches);
long pfml_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* PFML_B200_H */
