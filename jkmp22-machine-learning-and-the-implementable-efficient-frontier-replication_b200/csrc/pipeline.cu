// Orchestration of the PFML date-step on one B200 and the C ABI (include/pfml_b200.h).
//
// Mathematical restatement used on the device (every identity is exact; only rounding order differs from the
// reference, checked against the oracle at <= 1e-8 relative, typically 1e-13):
//   * Lambda is a vector (quirk Q11); everything runs in the "tilde" domain  m = L^-1/2 m~ L^1/2,
//     omega = L^-1/2 omega~  (L = diag(lambda)), so the symmetric, well-conditioned m~ is the GEMM operand.
//   * sqrtm(sigma_hat^2 - 4I) (General_functions.py:956) = sqrt(x^2 + 4x): scaled product-form
//     Denman-Beavers iteration, GEMM + block inverse only.
//   * inv() (General_functions.py:960) and the two solves (PFML_Input_Data.py:455-456): recursive
//     Schur-complement block inverse, 64x64 Gauss-Jordan base blocks, everything else DMMA GEMMs.
//   * chains (PFML_Input_Data.py:419-453): shared prefix A_{theta+1} = (m G_0) B_theta and Horner's rule,
//     11 n^3 + 12 n^2 P GEMMs instead of 22 + 24.
//   * risk = gamma omega' Sigma omega from the factor form Sigma = L F L' + diag(ivol) as one stacked
//     (n+K)-deep contraction; tc = w omega_chg' Lambda omega_chg = w omega~_chg' omega~_chg; both symmetric
//     (lower tiles computed, mirrored), denom = risk + tc fused in the tc epilogue.
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "kernels.cuh"
#include "pfml_b200.h"

namespace pfml {

// ------------------------------------------------------------------------------------------------
// errors, launch counter, profiling
// ------------------------------------------------------------------------------------------------
std::atomic<long> g_launch_count{0};
static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

static const char* kTagNames[] = {"dmma_gemm", "standardize", "m_func", "chain_nxn", "horner_nxP", "solve", "denom",
                                  "rff", "ridge", "cov", "other", "emu_i8_gemm"};
constexpr int N_TAGS = 12;   // slot 0 = all GEMM launches (tag -1), slots 1.. = ProfTag + ... see slot_of
static int slot_of(int tag) {
  switch (tag) {
    case -1: return 0;
    case -2: return 11;
    case T_STD: return 1;
    case T_MFUNC: return 2;
    case T_CHAIN: return 3;
    case T_HORNER: return 4;
    case T_SOLVE: return 5;
    case T_DENOM: return 6;
    case T_RFF: return 7;
    case T_RIDGE: return 8;
    case T_COV: return 9;
    default: return 10;
  }
}
struct ProfRec { int slot; cudaEvent_t e0, e1; double flops; };
static std::atomic<bool> g_prof_on{false};
static std::vector<ProfRec> g_recs;
static std::vector<cudaEvent_t> g_open[N_TAGS];
static std::mutex g_prof_mu;

void prof_begin(int tag, cudaStream_t s) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaEvent_t e;
  cudaEventCreate(&e);
  cudaEventRecord(e, s);
  g_open[slot_of(tag)].push_back(e);
}
void prof_end(int tag, cudaStream_t s, double flops, double) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  int sl = slot_of(tag);
  if (g_open[sl].empty()) return;
  cudaEvent_t e0 = g_open[sl].back();
  g_open[sl].pop_back();
  cudaEvent_t e1;
  cudaEventCreate(&e1);
  cudaEventRecord(e1, s);
  g_recs.push_back({sl, e0, e1, flops});
}

// ------------------------------------------------------------------------------------------------
// workspace carving
// ------------------------------------------------------------------------------------------------
struct Carver {
  char* base; size_t off = 0, cap;
  Carver(void* p, size_t c) : base((char*)p), cap(c) {}
  template <class T> T* take(size_t n) {
    size_t bytes = (n * sizeof(T) + 255) / 256 * 256;
    T* r = base ? (T*)(base + off) : nullptr;
    off += bytes;
    return r;
  }
};

// ------------------------------------------------------------------------------------------------
// recursive block inverse (np.linalg.inv, General_functions.py:960; np.linalg.solve, PFML_Input_Data.py:455)
//   M = [A B; C D],  S = D - C A^-1 B,
//   M^-1 = [A^-1 + W S^-1 V, -W S^-1; -S^-1 V, S^-1],  W = A^-1 B, V = C A^-1.
// Wk (destroyed) holds M, X receives M^-1, T1 holds W and V.  All [B][NP][NP].
// ------------------------------------------------------------------------------------------------
struct InvCtx { int B, NP; double *Wk, *X, *T1; int* status; cudaStream_t s; bool spd = false; };

static int inv_rec(const InvCtx& c, int off, int size) {
  if (size <= BASE_MAX) return base_inverse(c.B, c.NP, c.Wk, c.X, off, size, c.spd, c.status, c.s);
  const int nb = size / BASE, h = ((nb + 1) / 2) * BASE, r = size - h;
  const long NP = c.NP, NN = NP * NP;
  auto at = [&](double* base, int r0, int c0) { return base + (long)r0 * NP + c0; };
  PFML_TRY(inv_rec(c, off, h));
  GemmP g;
  g.lda = g.ldb = g.ldd = g.lde = NP;
  g.sA = g.sB = g.sD = g.sE = NN;
  // W = X11 * B
  g.A = at(c.X, off, off); g.B = at(c.Wk, off, off + h); g.D = at(c.T1, off, off + h);
  g.M = h; g.N = r; g.K = h;
  PFML_TRY(gemm(g, c.B, c.s));
  if (c.spd) {
    // symmetric positive definite: C = B', V = W', X21 = X12'; S and X11 are symmetric (lower tiles + mirror)
    // S = D - B' * W   (in place over D)
    g.A = at(c.Wk, off, off + h); g.a_kinner = false; g.B = at(c.T1, off, off + h); g.D = at(c.Wk, off + h, off + h);
    g.E = g.D; g.alpha = -1.0; g.beta = 1.0; g.sym = 1;
    g.M = r; g.N = r; g.K = h;
    PFML_TRY(gemm(g, c.B, c.s));
    PFML_TRY(inv_rec(c, off + h, r));
    // X12 = -W * X22, mirrored into X21
    g.a_kinner = true; g.sym = 0; g.E = nullptr; g.beta = 0.0;
    g.A = at(c.T1, off, off + h); g.B = at(c.X, off + h, off + h); g.D = at(c.X, off, off + h);
    g.DT = at(c.X, off + h, off); g.ldt = NP;
    g.M = h; g.N = r; g.K = r;
    PFML_TRY(gemm(g, c.B, c.s));
    // X11 = X11 - X12 * W'
    g.DT = nullptr;
    g.A = at(c.X, off, off + h); g.B = at(c.T1, off, off + h); g.b_kinner = true; g.D = at(c.X, off, off);
    g.E = g.D; g.beta = 1.0; g.sym = 1;
    g.M = h; g.N = h; g.K = r;
    PFML_TRY(gemm(g, c.B, c.s));
    return 0;
  }
  // V = C * X11
  g.A = at(c.Wk, off + h, off); g.B = at(c.X, off, off); g.D = at(c.T1, off + h, off);
  g.M = r; g.N = h; g.K = h;
  PFML_TRY(gemm(g, c.B, c.s));
  // S = D - C * W   (in place over D)
  g.A = at(c.Wk, off + h, off); g.B = at(c.T1, off, off + h); g.D = at(c.Wk, off + h, off + h);
  g.E = g.D; g.alpha = -1.0; g.beta = 1.0;
  g.M = r; g.N = r; g.K = h;
  PFML_TRY(gemm(g, c.B, c.s));
  PFML_TRY(inv_rec(c, off + h, r));
  // X12 = -W * X22
  g.E = nullptr; g.beta = 0.0;
  g.A = at(c.T1, off, off + h); g.B = at(c.X, off + h, off + h); g.D = at(c.X, off, off + h);
  g.M = h; g.N = r; g.K = r;
  PFML_TRY(gemm(g, c.B, c.s));
  // X21 = -X22 * V
  g.A = at(c.X, off + h, off + h); g.B = at(c.T1, off + h, off); g.D = at(c.X, off + h, off);
  g.M = r; g.N = h; g.K = r;
  PFML_TRY(gemm(g, c.B, c.s));
  // X11 = X11 - X12 * V
  g.A = at(c.X, off, off + h); g.B = at(c.T1, off + h, off); g.D = at(c.X, off, off);
  g.E = g.D; g.beta = 1.0;
  g.M = h; g.N = h; g.K = r;
  PFML_TRY(gemm(g, c.B, c.s));
  return 0;
}

static int inverse(const InvCtx& c) { return inv_rec(c, 0, c.NP); }

// symmetric positive definite batch inverse for other translation units (ridge.cu)
int spd_inverse(int B, int NP, double* Wk, double* X, double* T1, int* status, cudaStream_t s) {
  InvCtx c{B, NP, Wk, X, T1, status, s, true};
  return inverse(c);
}

// ------------------------------------------------------------------------------------------------
// m_tilde: General_functions.py:955-960 given x, sigma_gr, y.  Result pointer returned in *mt_out
// (one of mt / Xinv, they ping-pong).
// ------------------------------------------------------------------------------------------------
constexpr int DB_ITERS = 10;

// n x n x P products of the date-step on the INT8 tensor cores (emu.cu) instead of the DMMA pipe
// The mode is process-wide but read ONCE per call (a call in flight keeps the mode it started with).
static std::atomic<int> g_use_emu{-1};
static int emu_mode() {
  int m = g_use_emu.load(std::memory_order_relaxed);
  if (m < 0) {
    const char* e = getenv("PFML_EMULATE_FP64");
    m = e ? atoi(e) : 1;
    if (m < 0 || m > 2) m = 1;
    g_use_emu.store(m, std::memory_order_relaxed);
  }
  return m;
}

// err[0] bits 0-3: a barrier wait of the INT8 GEMM timed out (any matrix may be affected: status bit 3 on every date);
// err[EMU_ERR_SLOT0 + b] bit 4: NaN / Inf reached an operand of date / matrix b (status bit 4 on that date only).
// single != 0: one status word for the whole batch (pfml_dgemm_emulated).
__global__ void fold_emu_error_kernel(int B, const int* __restrict__ err, int* __restrict__ status, int single) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int e = ((err[0] & 15) ? 8 : 0) | (err[EMU_ERR_SLOT0 + b] & 16);
  if (e) atomicOr(status + (single ? 0 : b), e);
}

struct MtBufs {
  double *x, *sgr, *yv, *Mdb, *Wk, *Xinv, *T1, *Ya, *Yb, *mt, *part, *sc;
};

static int fixed_point(int B, int NP, int iterations, const MtBufs& m, int* status, cudaStream_t s, double** mt_out);

static int mtilde(int B, int NP, int iterations, const MtBufs& m, int* status, cudaStream_t s, double** mt_out) {
  const long NN = (long)NP * NP;
  GemmP g;
  g.lda = g.ldb = g.ldd = g.lde = NP;
  g.sA = g.sB = g.sD = g.sE = NN;
  g.M = g.N = g.K = NP;
  // A = sigma_hat^2 - 4I = x^2 + 4x   (symmetric)
  g.A = m.x; g.B = m.x; g.E = m.x; g.beta = 4.0; g.D = m.Mdb; g.sym = 1;
  PFML_TRY(gemm(g, B, s));
  PFML_TRY(copy2(B, NN, m.Mdb, m.Ya, m.Wk, s));
  double *Ya = m.Ya, *Yb = m.Yb;
  InvCtx ic{B, NP, m.Wk, m.Xinv, m.T1, status, s, true};
  for (int k = 0; k < DB_ITERS; k++) {
    PFML_TRY(inverse(ic));
    PFML_TRY(db_norms(B, NP, m.Mdb, m.Xinv, m.part, m.sc, s));
    // Y <- 0.5 mu Y + (0.5/mu) Y M^-1
    GemmP y;
    y.lda = y.ldb = y.ldd = y.lde = NP;
    y.sA = y.sB = y.sD = y.sE = NN;
    y.M = y.N = y.K = NP;
    y.A = Ya; y.B = m.Xinv; y.E = Ya; y.D = Yb;
    y.alpha = 1.0; y.beta = 1.0; y.az = m.sc + 0; y.bz = m.sc + 1; y.s_z = 8;
    PFML_TRY(gemm(y, B, s));
    // M <- 0.5 (I + 0.5 (mu^2 M + M^-1 / mu^2)),  copy into Wk for the next inverse
    PFML_TRY(db_update_M(B, NP, m.Mdb, m.Xinv, m.Wk, m.sc, s));
    double* t = Ya; Ya = Yb; Yb = t;
  }
  if (status) PFML_TRY(check_finite_status(B, m.sc, status, s));
  PFML_TRY(mt0_from_sqrt(B, NP, m.x, Ya, m.mt, s));
  return fixed_point(B, NP, iterations, m, status, s, mt_out);
}

// the `iterations` steps m~ <- inv(x + y - m~ o sigma_gr) of General_functions.py:959-960, starting from m.mt
static int fixed_point(int B, int NP, int iterations, const MtBufs& m, int* status, cudaStream_t s, double** mt_out) {
  double *mt = m.mt, *other = m.Xinv;
  for (int it = 0; it < iterations; it++) {
    PFML_TRY(fp_build(B, NP, m.x, m.yv, mt, m.sgr, m.Wk, s));
    InvCtx fc{B, NP, m.Wk, other, m.T1, status, s, true};
    PFML_TRY(inverse(fc));
    double* t = mt; mt = other; other = t;
  }
  *mt_out = mt;
  return 0;
}

// ------------------------------------------------------------------------------------------------
// date-step workspace
// ------------------------------------------------------------------------------------------------
struct StepWs {
  int B, NP, P, nt; long ldP;
  BatchVecs v;
  MtBufs m;
  double *XC, *XCl1, *Ut, *Ev, *G;
  int8_t *eAs, *eBs; double *esa, *esb; int *eex, *eerr;   // INT8-emulated GEMM operands (emu.cu)
  double* chain;                                            // B~_2 .. B~_{lb-1} of the chain (B~_1 sits in Ya)
  int8_t* eMs; double* esm;                                 // B-type slices of m~ (kept through the chain / term loop)
  int* ezf;                                                 // per-date flag: omega_chg holds NaN (quirk Q8)
  int8_t* eSs; double* eSsb;                                // B-type slices of the signal slabs 2 .. lb+1 and their scales
  double *St, *Za, *Zb, *Nn, *Nl1, *OM, *BS, *OMC;
  double *risk_tmp, *tc_tmp;
  size_t bytes;
};

static StepWs carve_step(void* ws, size_t cap, int B, int NP, int P, int nt, int K, bool need_risk, bool need_tc) {
  StepWs w{};
  w.B = B; w.NP = NP; w.P = P; w.nt = nt;
  w.ldP = round_up(P, 8);
  Carver c(ws, cap);
  const size_t NN = (size_t)NP * NP, BNN = B * NN, nP = (size_t)NP * w.ldP, BnP = B * nP;
  w.v.sl = c.take<double>((size_t)B * NP);
  w.v.isl = c.take<double>((size_t)B * NP);
  w.v.g = c.take<double>((size_t)B * nt * NP);
  w.v.g0raw = c.take<double>((size_t)B * NP);
  w.v.rv = c.take<double>((size_t)B * NP);
  w.v.dvl = c.take<double>((size_t)B * NP);
  w.v.Lg = c.take<double>((size_t)B * NP * KP);
  w.v.Lt = c.take<double>((size_t)B * NP * KP);
  w.v.LF = c.take<double>((size_t)B * NP * KP);
  w.v.Fg = c.take<double>((size_t)B * KP * KP);
  w.v.scal = c.take<double>((size_t)B * NSCAL);
  w.m.x = c.take<double>(BNN); w.m.sgr = c.take<double>(BNN); w.m.yv = c.take<double>((size_t)B * NP);
  w.m.Mdb = c.take<double>(BNN); w.m.Wk = c.take<double>(BNN); w.m.Xinv = c.take<double>(BNN);
  w.m.T1 = c.take<double>(BNN); w.m.Ya = c.take<double>(BNN); w.m.Yb = c.take<double>(BNN);
  w.m.mt = c.take<double>(BNN);
  w.m.part = c.take<double>((size_t)B * NORM_CHUNKS * 3);
  w.m.sc = c.take<double>((size_t)B * 8);
  w.XC = c.take<double>(BNN); w.XCl1 = c.take<double>(BNN);
  w.Ut = c.take<double>((size_t)B * NP * KP);
  w.Ev = c.take<double>((size_t)B * NP);
  w.St = c.take<double>((size_t)nt * BnP);
  // the quadrature stack G [B][NP][121 K] is dead before the signal slabs are written: alias it when it fits
  const size_t g_elems = (size_t)B * NP * stj_ldg(K);
  w.G = (g_elems <= (size_t)nt * BnP) ? w.St : c.take<double>(g_elems);
  w.Za = c.take<double>(BnP); w.Zb = c.take<double>(BnP);
  w.Nn = c.take<double>(BnP); w.Nl1 = c.take<double>(BnP);
  w.OM = c.take<double>((size_t)B * (NP + KP) * w.ldP);
  w.BS = c.take<double>((size_t)B * (NP + KP) * w.ldP);
  w.OMC = c.take<double>(BnP);
  const int Pn = P > NP ? P : NP;          // B operands are n x P (Horner, solves) or n x n (chain); A operands n x n or P x n (a9)
  w.eAs = c.take<int8_t>(emu_a_slices_bytes(B, Pn, NP));
  w.eBs = c.take<int8_t>(emu_b_slices_bytes(B, NP, Pn));
  w.esa = c.take<double>((size_t)B * Pn); w.esb = c.take<double>((size_t)B * Pn);
  w.eex = c.take<int>((size_t)B * Pn); w.eerr = c.take<int>(EMU_ERR_SLOT0 + (size_t)B); w.ezf = c.take<int>((size_t)B);
  w.chain = nt >= 5 ? c.take<double>((size_t)(nt - 4) * BNN) : nullptr;   // lb - 2 matrices
  w.eMs = c.take<int8_t>(emu_b_slices_bytes(B, NP, NP)); w.esm = c.take<double>((size_t)B * NP);
  w.eSs = nt >= 3 ? c.take<int8_t>((size_t)(nt - 2) * emu_b_slices_bytes(B, NP, P)) : nullptr;
  w.eSsb = nt >= 3 ? c.take<double>((size_t)(nt - 2) * B * P) : nullptr;
  w.risk_tmp = need_risk ? c.take<double>((size_t)B * P * P) : nullptr;
  w.tc_tmp = need_tc ? c.take<double>((size_t)B * P * P) : nullptr;
  w.bytes = c.off;
  return w;
}

static PanelView view_of(const pfml_panel_t* p) {
  PanelView v;
  v.T = p->T; v.N = p->N; v.F = p->F; v.p = p->p; v.K = p->K;
  v.rff = p->rff; v.vol_scale = p->vol_scale; v.gt = p->gt; v.lam = p->lam; v.ret_ld1 = p->ret_ld1;
  v.fct_load = p->fct_load; v.fct_cov = p->fct_cov; v.ivol = p->ivol; v.wealth = p->wealth; v.rf = p->rf;
  v.gamma_rel = p->gamma_rel; v.mu = p->mu; v.lb = p->lb; v.iterations = p->iterations;
  return v;
}

// Optional parts of a date-step call: the fused expanding-window sums and the stage taps / injections behind the
// stage-level entry points (pfml_omega_batched, pfml_rtilde_denom_batched).
struct StepOpts {
  const int* bucket = nullptr; int n_buckets = 0; double* r_sum = nullptr; double* d_sum = nullptr;
  double *omega_out = nullptr, *omega_chg_out = nullptr, *m_out = nullptr;    // un-tilded, [B][NP][P] / [B][NP][NP]
  const double *omega_in = nullptr, *omega_chg_in = nullptr;                  // a9 from a given omega / omega_chg [B][NP][P]
  bool stop_after_omega = false;
  bool keep_status = false;                 // do not clear `status` first (a further pass over the same dates)
  // further bandwidths g = 1 .. n_extra of the SAME dates (quirk Q10: m, the chain and the two inverses do not depend on
  // g): their feature panels and outputs (entries of risk / tc / signal / sums may be null)
  int n_extra = 0;
  const double* const* x_rff = nullptr;
  double* const* x_r_tilde = nullptr; double* const* x_denom = nullptr; double* const* x_risk = nullptr; double* const* x_tc = nullptr;
  double* const* x_signal = nullptr; double* const* x_r_sum = nullptr; double* const* x_d_sum = nullptr;
};

static int date_steps_impl(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                           double* r_tilde, double* denom, double* risk, double* tc, double* signal_t, int* status,
                           void* workspace, size_t workspace_bytes, cudaStream_t s, const StepOpts& opt = StepOpts()) {
  if (!panel || B <= 0 || n_max <= 0 || !date_t || !sel || !n_sel || !workspace ||
      (!opt.stop_after_omega && (!r_tilde || !denom))) {
    set_error("pfml_date_steps: null/empty argument");
    return -1;
  }
  if (panel->K > KP || panel->K <= 0) {
    set_error("pfml_date_steps: K = %d factors unsupported (1..%d)", panel->K, KP);
    return -2;
  }
  if (panel->lb < 2 || panel->p < 2 || (panel->p & 1)) {
    set_error("pfml_date_steps: need lb >= 2 and even p >= 2 (lb=%d p=%d)", panel->lb, panel->p);
    return -2;
  }
  const PanelView pv = view_of(panel);
  const int NP = pfml_padded_n(n_max), P = pv.p + 1, nt = pv.lb + 2, lb = pv.lb;
  StepWs w = carve_step(workspace, workspace_bytes, B, NP, P, nt, pv.K, risk == nullptr, tc == nullptr);
  if (w.bytes > workspace_bytes) {
    set_error("pfml_date_steps: workspace too small (%zu < %zu bytes)", workspace_bytes, w.bytes);
    return -3;
  }
  if (!risk) risk = w.risk_tmp;
  if (!tc) tc = w.tc_tmp;
  const long NN = (long)NP * NP, ldP = w.ldP, nP = (long)NP * ldP, sStack = (long)(NP + KP) * ldP;
  if (status && !opt.keep_status) PFML_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(int) * B, s));

  // ---- a3/a4 gathers ------------------------------------------------------------------------------
  {
    ProfScope ps(T_STD, s);
    PFML_TRY(prep_vectors(pv, B, NP, date_t, sel, n_sel, w.v, s));
  }
  auto St = [&](int tau) { return w.St + (long)tau * NP * ldP; };   // + b * nt * nP per batch
  const long sSt = (long)nt * nP;
  auto gv = [&](int tau) { return w.v.g + (long)tau * NP; };        // + b * nt * NP per batch
  const long sg = (long)nt * NP;

  const int mode = emu_mode();
  const bool emu = mode != 0 && emu_shape_ok(NP, P, NP);
  const bool emu_nn = emu && emu_shape_ok(NP, NP, NP);
  int* const bad = w.eerr + EMU_ERR_SLOT0;                 // per-date non-finite flags of the INT8 path
  if (emu) PFML_CUDA_OK(cudaMemsetAsync(w.eerr, 0, sizeof(int) * (EMU_ERR_SLOT0 + (size_t)B), s));
  // direct form of the n x P sums (INT8 path): Z = sum_k B~_k S~_{k+1} term by term next to the chain products
  const bool direct = emu_nn && lb - 1 <= 16 && w.chain != nullptr;
  // mode 2 (direct form only): the products whose errors do not compound (n x P sums, solves, a9) use 6 of the 7
  // digits (21 products); the chain products, whose results are multiplied on, always use all 7
  EmuExtra xq;
  xq.digits = (mode == 2 && direct) ? 6 : 0;
  const double* emu_a_cur = nullptr;       // whose slices currently sit in w.eAs
  auto nxP = [&](const double* A, const double* X, long sX, double alpha, const double* rs, long s_rs, const double* E, long sE,
                 double* D, long sD) -> int {
    // one n x n x P product D = alpha * rs o (A X) + E on whichever tensor pipe is selected; A is sliced on demand
    if (emu) {
      if (emu_a_cur != A) {
        PFML_TRY(emu_slice_a(A, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
        emu_a_cur = A;
      }
      PFML_TRY(emu_slice_b(X, ldP, sX, NP, P, B, w.eBs, w.esb, w.eex, w.eerr, s));
      return emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, NP, P, NP, B, alpha, rs, s_rs, E ? 1.0 : 0.0, E, ldP, sE, D, ldP, sD, xq, w.eerr, s);
    }
    GemmP g;
    g.A = A; g.lda = NP; g.sA = NN; g.B = X; g.ldb = ldP; g.sB = sX; g.D = D; g.ldd = ldP; g.sD = sD;
    g.M = NP; g.N = P; g.K = NP; g.alpha = alpha; g.rs = rs; g.s_rs = s_rs;
    if (E) { g.E = E; g.lde = ldP; g.sE = sE; g.beta = 1.0; }
    return gemm(g, B, s);
  };
  double* mt = nullptr;
  const double* bk[16] = {};
  double* B11 = w.m.x;
  const size_t slabS = emu_b_slices_bytes(B, NP, P);
  if (opt.omega_in) {
    // ---- stage entry: a9 from a given omega / omega_chg (un-tilded): omega~ = Lambda^1/2 omega ----------
    if (!opt.omega_chg_in) { set_error("pfml_rtilde_denom_batched: omega_chg missing"); return -1; }
    PFML_TRY(tilde_rows(B, NP, P, opt.omega_in, n_sel, w.v.sl, w.OM, ldP, sStack, s));
    PFML_TRY(tilde_rows(B, NP, P, opt.omega_chg_in, n_sel, w.v.sl, w.OMC, ldP, nP, s));
  } else {
  // ---- a5 m_tilde: m~_0 by quadrature + Woodbury (stieltjes.cu), then the fixed-point inverses -----
  {
    ProfScope ps(T_MFUNC, s);
    PFML_TRY(build_x_factor(pv, B, NP, date_t, sel, n_sel, w.v, w.m.x, w.m.sgr, w.m.yv, s));
    PFML_TRY(mt0_stieltjes(B, NP, pv.K, n_sel, w.v, w.Ut, w.Ev, w.G, w.m.mt, status, s));
    PFML_TRY(fixed_point(B, NP, pv.iterations, w.m, status, s, &mt));
  }
  if (opt.m_out) PFML_TRY(m_from_mtilde(B, NP, NP, mt, w.v.isl, w.v.sl, opt.m_out, s));

  // ---- a6 standardise (after m~: the signal slabs reuse the quadrature stack's memory) -------------
  {
    ProfScope ps(T_STD, s);
    // direct form: only S~_0 and S~_1 are needed in FP64 (addends); the other slabs go straight to INT8 digits
    PFML_TRY(standardize(pv, B, NP, P, ldP, date_t, sel, n_sel, w.v, w.St, signal_t, !direct, direct ? 2 : nt, s));
    if (direct) PFML_TRY(emu_standardize_slices(pv, B, NP, P, 2, lb, date_t, sel, n_sel, w.v, w.eSs, w.eSsb, w.eerr, s));
  }
  // free n x n buffers after m_tilde: x, sgr, Mdb, Ya, Yb, Wk, T1 and whichever of {mt, Xinv} is not `mt`
  double* Bt_a = w.m.Ya; double* Bt_b = w.m.Yb; double* CB = w.m.Mdb; double* CBs = w.m.sgr;
  double* spare = (mt == w.m.mt) ? w.m.Xinv : w.m.mt;

  double* Z = nullptr;
  if (direct) {
    // ---- a7 + a8 together: for k = 1 .. lb-1 the slices of B~_k serve two products, the chain step
    //   B~_{k+1} = B~_k m~ G_{k+1} and the term Z += B~_k S~_{k+1} (Z starts as S~_1).  m~ is sliced once as a B operand.
    PFML_TRY(chain_init(B, NP, mt, gv(1), sg, Bt_a, CB, s));
    PFML_TRY(emu_slice_b(mt, NP, NN, NP, NP, B, w.eMs, w.esm, w.eex, w.eerr, s));
    bk[0] = Bt_a;
    Z = w.Za;
    for (int k = 1; k < lb; k++) {
      const bool last = (k == lb - 1);
      PFML_TRY(emu_slice_a(Bt_a, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
      {
        ProfScope ps(T_CHAIN, s);
        EmuExtra x;
        x.cs = gv(k + 1); x.s_cs = sg;
        double* out = last ? B11 : w.chain + (size_t)(k - 1) * B * NN;
        PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eMs, w.esm, NP, NP, NP, B, 1.0, nullptr, 0, 0.0, nullptr, 0, 0, out, NP, NN, x, w.eerr, s));
        if (!last) bk[k] = out;
        Bt_a = out;
      }
      {
        ProfScope ps(T_HORNER, s);
        const double* E = (k == 1) ? St(1) : Z;     // slab k + 1 was sliced by the standardise kernel
        PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eSs + (size_t)(k - 1) * slabS, w.eSsb + (size_t)(k - 1) * B * P, NP, P, NP, B, 1.0, nullptr,
                            0, 1.0, E, ldP, (k == 1) ? sSt : nP, Z, ldP, nP, xq, w.eerr, s));
      }
    }
    // NOTE: in iteration k the A slices are those of B~_k (sliced before the chain step overwrote Bt_a's role)
    {
      ProfScope ps(T_CHAIN, s);
      // const_l1~ = I + sum_k B~_k + B11 -> Wk (to be inverted);  const~ = I + m~ (G_0 CB) -> spare (to be inverted)
      PFML_TRY(chain_sum(B, NP, lb - 1, bk, B11, gv(0), sg, w.m.Wk, CBs, s));
      PFML_TRY(emu_slice_a(mt, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
      PFML_TRY(emu_slice_b(CBs, NP, NN, NP, NP, B, w.eBs, w.esb, w.eex, w.eerr, s));
      EmuExtra x;
      x.diag = 1.0;
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, NP, NP, NP, B, 1.0, nullptr, 0, 0.0, nullptr, 0, 0, spare, NP, NN, x, w.eerr, s));
    }
    {
      ProfScope ps(T_HORNER, s);
      // N = S~_0 + m~ G_0 Z  (the A slices of m~ are still in place);  N_l1 = Z + B11 S~_{lb+1}
      if (NP <= 1024) {                 // G_0 Z is only ever this product's B operand: the row scale rides in the slicing
        PFML_TRY(emu_slice_nmajor512(Z, ldP, nP, NP, P, B, w.eBs, w.esb, w.eex, nullptr, nullptr, w.eerr, s, gv(0), sg));
      } else {
        double* ZG = w.Zb;
        PFML_TRY(row_scale(B, NP, P, ldP, nP, Z, gv(0), sg, ZG, ldP, nP, s));
        PFML_TRY(emu_slice_b(ZG, ldP, nP, NP, P, B, w.eBs, w.esb, w.eex, w.eerr, s));
      }
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, NP, P, NP, B, 1.0, nullptr, 0, 1.0, St(0), ldP, sSt, w.Nn, ldP, nP, xq, w.eerr, s));
      PFML_TRY(emu_slice_a(B11, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eSs + (size_t)(lb - 1) * slabS, w.eSsb + (size_t)(lb - 1) * B * P, NP, P, NP, B, 1.0, nullptr, 0,
                          1.0, Z, ldP, nP, w.Nl1, ldP, nP, xq, w.eerr, s));
    }
  } else {
  // ---- a7 chain of n x n products (tilde domain) ------------------------------------------------
  //   B~_1 = m~ G_1, B~_{k+1} = B~_k m~ G_{k+1};  CB = sum_{k=0}^{lb-1} B~_k;  B11 = B~_lb
  {
    ProfScope ps(T_CHAIN, s);
    PFML_TRY(chain_init(B, NP, mt, gv(1), sg, Bt_a, CB, s));
    const bool keep_all = emu_nn && lb - 1 <= 16 && w.chain != nullptr;
    if (emu_nn) {
      // INT8 tensor-core path: m~ is the B operand of every product (sliced once), B~_k the A operand.  Every B~_k is
      // kept (w.chain) and the sums CB + B11 and G_0 CB are formed by one kernel at the end instead of a read-modify-
      // write of CB in every epilogue.
      PFML_TRY(emu_slice_b(mt, NP, NN, NP, NP, B, w.eBs, w.esb, w.eex, w.eerr, s));
      const double* bk[16];
      bk[0] = Bt_a;
      for (int k = 1; k < lb; k++) {
        const bool last = (k == lb - 1);
        PFML_TRY(emu_slice_a(Bt_a, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
        EmuExtra x;
        x.cs = gv(k + 1); x.s_cs = sg;
        double* out = last ? B11 : (keep_all ? w.chain + (size_t)(k - 1) * B * NN : Bt_b);
        if (!last && !keep_all) { x.D2 = CB; x.E2 = CB; x.ld2 = NP; x.s2 = NN; }
        PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, NP, NP, NP, B, 1.0, nullptr, 0, 0.0, nullptr, 0, 0, out, NP, NN, x, w.eerr, s));
        if (!last) bk[k] = out;
        if (keep_all) Bt_a = out; else { double* t = Bt_a; Bt_a = Bt_b; Bt_b = t; }
      }
      if (keep_all) PFML_TRY(chain_sum(B, NP, lb - 1, bk, B11, gv(0), sg, w.m.Wk, CBs, s));
    } else {
      GemmP g;
      g.lda = g.ldb = g.ldd = NP; g.sA = g.sB = g.sD = NN;
      g.M = g.N = g.K = NP;
      for (int k = 1; k < lb; k++) {
        const bool last = (k == lb - 1);
        g.A = Bt_a; g.B = mt; g.D = last ? B11 : Bt_b;
        g.cs = gv(k + 1); g.s_cs = sg;
        g.D2 = last ? nullptr : CB; g.E2 = last ? nullptr : CB;
        PFML_TRY(gemm(g, B, s));
        double* t = Bt_a; Bt_a = Bt_b; Bt_b = t;
      }
    }
    // const_l1~ = CB + B11 -> Wk (to be inverted);  const~ = I + m~ (G_0 CB) -> spare (to be inverted)
    if (!keep_all) {
      PFML_TRY(add_mat(B, NN, CB, B11, w.m.Wk, s));
      PFML_TRY(row_scale(B, NP, NP, NP, NN, CB, gv(0), sg, CBs, NP, NN, s));
    }
    if (emu_nn) {
      PFML_TRY(emu_slice_a(mt, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
      emu_a_cur = mt;                    // the Horner products below reuse these slices of m~
      PFML_TRY(emu_slice_b(CBs, NP, NN, NP, NP, B, w.eBs, w.esb, w.eex, w.eerr, s));
      EmuExtra x;
      x.diag = 1.0;
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, NP, NP, NP, B, 1.0, nullptr, 0, 0.0, nullptr, 0, 0, spare, NP, NN, x, w.eerr, s));
    } else {
      GemmP c;
      c.lda = c.ldb = c.ldd = NP; c.sA = c.sB = c.sD = NN;
      c.M = c.N = c.K = NP;
      c.A = mt; c.B = CBs; c.D = spare; c.diag = 1.0;
      PFML_TRY(gemm(c, B, s));
    }
  }

  // ---- a8 Horner accumulation of the n x P sums -------------------------------------------------
  //   Z'_j = G_j (S~_{j+1} + m~ Z'_{j+1}),  Z'_{lb-1} = G_{lb-1} S~_lb ;  Z = S~_1 + m~ Z'_1
  //   N = S~_0 + m~ G_0 Z ;  N_l1 = Z + B11 S~_{lb+1}
  {
    ProfScope ps(T_HORNER, s);
    const double* Zp = St(lb);         // already scaled by g_{lb-1} in the standardise kernel
    long sZp = sSt;
    double *za = w.Za, *zb = w.Zb;
    for (int j = lb - 2; j >= 0; j--) {
      PFML_TRY(nxP(mt, Zp, sZp, 1.0, (j >= 1) ? gv(j) : nullptr, sg, St(j + 1), sSt, za, nP));
      Zp = za; sZp = nP;
      double* t = za; za = zb; zb = t;
    }
    Z = (double*)Zp;                    // unscaled Z (j = 0 step)
    double* ZG = za;                    // the other ping-pong buffer
    PFML_TRY(row_scale(B, NP, P, ldP, nP, Z, gv(0), sg, ZG, ldP, nP, s));
    PFML_TRY(nxP(mt, ZG, nP, 1.0, nullptr, 0, St(0), sSt, w.Nn, nP));
    PFML_TRY(nxP(B11, St(lb + 1), sSt, 1.0, nullptr, 0, Z, nP, w.Nl1, nP));
  }
  }   // !direct

  // ---- a8 solves: omega~ = const~^-1 N, omega~_l1 = const~_l1^-1 N_l1, omega~_chg -----------------
  {
    ProfScope ps(T_SOLVE, s);
    // const~ sits in `spare`; move it to Wk after const_l1 has been inverted
    InvCtx i1{B, NP, w.m.Wk, w.XCl1, w.m.T1, status, s};
    PFML_TRY(inverse(i1));
    InvCtx i0{B, NP, spare, w.XC, w.m.T1, status, s};
    PFML_TRY(inverse(i0));
    PFML_TRY(nxP(w.XC, w.Nn, nP, 1.0, nullptr, 0, nullptr, 0, w.OM, sStack));
    // omega~_chg = omega~ - g0raw o (const_l1^-1 N_l1)   (quirk Q8: g0 without NaN->1)
    PFML_TRY(nxP(w.XCl1, w.Nl1, nP, -1.0, w.v.g0raw, NP, w.OM, sStack, w.OMC, nP));
  }
  }   // !omega_in
  // stage taps: omega = Lambda^-1/2 omega~ (PFML_Input_Data.py:455), omega_chg (:459)
  if (opt.omega_out) PFML_TRY(row_scale(B, NP, P, ldP, sStack, w.OM, w.v.isl, NP, opt.omega_out, P, (long)NP * P, s));
  if (opt.omega_chg_out) PFML_TRY(row_scale(B, NP, P, ldP, nP, w.OMC, w.v.isl, NP, opt.omega_chg_out, P, (long)NP * P, s));
  if (opt.stop_after_omega) {
    if (emu && status) {
      fold_emu_error_kernel<<<ceil_div(B, 128), 128, 0, s>>>(B, w.eerr, status, 0);
      PFML_CHECK_LAUNCH();
    }
    return 0;
  }

  // ---- a9 r_tilde, risk, tc, denom (from omega~ in w.OM, omega~_chg in w.OMC) ------------------------------
  auto a9_stage = [&](double* r_tilde, double* denom, double* risk, double* tc, double* r_sum, double* d_sum) -> int {
    ProfScope ps(T_DENOM, s);
    bool fused_sums = false;
    PFML_TRY(gemv_t(B, NP, P, w.OM, ldP, sStack, w.v.rv, NP, r_tilde, P, s));
    // Y = (L^-1/2 Lf)' omega~  -> rows NP.. of the OM stack
    GemmP y;
    y.a_kinner = false; y.A = w.v.Lt; y.lda = KP; y.sA = (long)NP * KP;
    y.B = w.OM; y.ldb = ldP; y.sB = sStack;
    y.D = w.OM + (long)NP * ldP; y.ldd = ldP; y.sD = sStack;
    y.M = KP; y.N = P; y.K = NP;
    PFML_TRY(gemm(y, B, s));
    // gamma F Y -> rows NP.. of the BS stack;  gamma (ivol/lambda) o omega~ -> rows 0..NP-1
    GemmP f;
    f.A = w.v.Fg; f.lda = KP; f.sA = KP * KP;
    f.B = w.OM + (long)NP * ldP; f.ldb = ldP; f.sB = sStack;
    f.D = w.BS + (long)NP * ldP; f.ldd = ldP; f.sD = sStack;
    f.M = KP; f.N = P; f.K = KP;
    PFML_TRY(gemm(f, B, s));
    PFML_TRY(row_scale(B, NP, P, ldP, sStack, w.OM, w.v.dvl, NP, w.BS, ldP, sStack, s));
    // risk = [omega~; Y]' [gamma (ivol/lambda) omega~; gamma F Y]
    const bool emu_pp = emu && NP <= 1024 && emu_shape_ok(P, P, NP);   // the one-pass transposed slicing holds K <= 1024 in registers
    if (emu_pp) {
      // factor part (K = KP) on the DMMA pipe, then the n-deep part on the INT8 tensor cores accumulates onto it;
      // symmetric: only the upper triangle is computed; sym_finish mirrors it and forms denom
      GemmP r;
      r.a_kinner = false; r.A = w.OM + (long)NP * ldP; r.lda = ldP; r.sA = sStack;
      r.B = w.BS + (long)NP * ldP; r.ldb = ldP; r.sB = sStack;
      r.D = risk; r.ldd = P; r.sD = (long)P * P;
      r.M = P; r.N = P; r.K = KP; r.sym = 1;
      PFML_TRY(gemm(r, B, s));
      PFML_TRY(emu_slice_nmajor512(w.OM, ldP, sStack, NP, P, B, nullptr, nullptr, nullptr, w.eAs, w.esa, w.eerr, s));
      PFML_TRY(emu_slice_nmajor512(w.BS, ldP, sStack, NP, P, B, w.eBs, w.esb, w.eex, nullptr, nullptr, w.eerr, s));
      EmuExtra xr;
      xr.sym = true; xr.digits = xq.digits;
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, P, P, NP, B, 1.0, nullptr, 0, 1.0, risk, P, (long)P * P, risk, P, (long)P * P, xr,
                          w.eerr, s));
      // tc = w omega~_chg' omega~_chg ; denom = risk + tc
      PFML_CUDA_OK(cudaMemsetAsync(w.ezf, 0, sizeof(int) * B, s));
      PFML_TRY(emu_slice_nmajor512(w.OMC, ldP, nP, NP, P, B, w.eBs, w.esb, w.eex, w.eAs, w.esa, w.eerr, s, nullptr, 0, w.ezf));
      EmuExtra xt;
      xt.sym = true; xt.az = w.v.scal + 0; xt.s_az = NSCAL; xt.digits = xq.digits;
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, P, P, NP, B, 1.0, nullptr, 0, 0.0, nullptr, 0, 0, tc, P, (long)P * P, xt, w.eerr, s));
      // denom = risk + tc; lower triangles = copies of the upper; the per-window sums of denom ride along
      PFML_TRY(sym_finish(B, P, risk, tc, denom, w.ezf, bad, opt.bucket, opt.n_buckets, d_sum, s));
      PFML_TRY(poison_bad_dates(B, P, bad, r_tilde, nullptr, nullptr, nullptr, s));
      fused_sums = true;
    } else {
    GemmP r;
    r.a_kinner = false; r.A = w.OM; r.lda = ldP; r.sA = sStack;
    r.B = w.BS; r.ldb = ldP; r.sB = sStack;
    r.D = risk; r.ldd = P; r.sD = (long)P * P;
    r.M = P; r.N = P; r.K = NP + KP; r.sym = 1;
    PFML_TRY(gemm(r, B, s));
    // tc = w omega~_chg' omega~_chg ; denom = risk + tc
    GemmP t;
    t.a_kinner = false; t.A = w.OMC; t.lda = ldP; t.sA = nP;
    t.B = w.OMC; t.ldb = ldP; t.sB = nP;
    t.D = tc; t.ldd = P; t.sD = (long)P * P;
    t.D2 = denom; t.E2 = risk;
    t.az = w.v.scal + 0; t.s_z = NSCAL;
    t.M = P; t.N = P; t.K = NP; t.sym = 1;
    PFML_TRY(gemm(t, B, s));
    if (emu) PFML_TRY(poison_bad_dates(B, P, bad, r_tilde, denom, risk, tc, s));
    }
    if (emu && status) {
      fold_emu_error_kernel<<<ceil_div(B, 128), 128, 0, s>>>(B, w.eerr, status, 0);
      PFML_CHECK_LAUNCH();
    }
    // ---- a11 expanding-window sums (PFML_Search_Coef.py:77-88, :112-121) of this batch, accumulated into r_sum / d_sum ----
    if (opt.bucket && opt.n_buckets > 0) {
      if (r_sum) PFML_TRY(bucket_accumulate(B, P, r_tilde, opt.bucket, opt.n_buckets, r_sum, s));
      if (d_sum && !fused_sums) PFML_TRY(bucket_accumulate(B, (long)P * P, denom, opt.bucket, opt.n_buckets, d_sum, s));
    }
    return 0;
  };
  PFML_TRY(a9_stage(r_tilde, denom, risk, tc, opt.r_sum, opt.d_sum));

  // ---- further bandwidths of the same dates (quirk Q10): only the g-dependent stages are repeated ----------------
  // a6 on that g's feature panel, the n x P sums (the A digits of every kept B~_k are cut again: 11 small launches), the two
  // solve products against the shared inverses, a9.  Same kernels on the same operands as a single-g call with that
  // panel, in a different launch order: the results are bit-identical to it.
  for (int g = 0; g < opt.n_extra; g++) {
    if (!direct || opt.omega_in) { set_error("pfml_date_steps_multi_g: shared path unavailable for this shape"); return -2; }
    if (emu) PFML_CUDA_OK(cudaMemsetAsync(bad, 0, sizeof(int) * B, s));      // the flags of the previous g are in `status` already
    PanelView pg = pv;
    pg.rff = opt.x_rff[g];
    double* sig_g = opt.x_signal ? opt.x_signal[g] : nullptr;
    double* risk_g = (opt.x_risk && opt.x_risk[g]) ? opt.x_risk[g] : w.risk_tmp;
    double* tc_g = (opt.x_tc && opt.x_tc[g]) ? opt.x_tc[g] : w.tc_tmp;
    {
      ProfScope ps(T_STD, s);
      PFML_TRY(standardize(pg, B, NP, P, ldP, date_t, sel, n_sel, w.v, w.St, sig_g, false, 2, s));
      PFML_TRY(emu_standardize_slices(pg, B, NP, P, 2, lb, date_t, sel, n_sel, w.v, w.eSs, w.eSsb, w.eerr, s));
    }
    double* Zg = w.Za;
    {
      ProfScope ps(T_HORNER, s);
      for (int k = 1; k < lb; k++) {
        PFML_TRY(emu_slice_a(bk[k - 1], NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
        const double* E = (k == 1) ? St(1) : Zg;
        PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eSs + (size_t)(k - 1) * slabS, w.eSsb + (size_t)(k - 1) * B * P, NP, P, NP, B, 1.0, nullptr,
                            0, 1.0, E, ldP, (k == 1) ? sSt : nP, Zg, ldP, nP, xq, w.eerr, s));
      }
      PFML_TRY(emu_slice_a(mt, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
      if (NP <= 1024) {
        PFML_TRY(emu_slice_nmajor512(Zg, ldP, nP, NP, P, B, w.eBs, w.esb, w.eex, nullptr, nullptr, w.eerr, s, gv(0), sg));
      } else {
        PFML_TRY(row_scale(B, NP, P, ldP, nP, Zg, gv(0), sg, w.Zb, ldP, nP, s));
        PFML_TRY(emu_slice_b(w.Zb, ldP, nP, NP, P, B, w.eBs, w.esb, w.eex, w.eerr, s));
      }
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eBs, w.esb, NP, P, NP, B, 1.0, nullptr, 0, 1.0, St(0), ldP, sSt, w.Nn, ldP, nP, xq, w.eerr, s));
      PFML_TRY(emu_slice_a(B11, NP, NN, NP, NP, B, w.eAs, w.esa, w.eerr, s));
      PFML_TRY(emu_mma_ex(w.eAs, w.esa, w.eSs + (size_t)(lb - 1) * slabS, w.eSsb + (size_t)(lb - 1) * B * P, NP, P, NP, B, 1.0, nullptr, 0,
                          1.0, Zg, ldP, nP, w.Nl1, ldP, nP, xq, w.eerr, s));
    }
    {
      ProfScope ps(T_SOLVE, s);
      emu_a_cur = nullptr;
      PFML_TRY(nxP(w.XC, w.Nn, nP, 1.0, nullptr, 0, nullptr, 0, w.OM, sStack));
      PFML_TRY(nxP(w.XCl1, w.Nl1, nP, -1.0, w.v.g0raw, NP, w.OM, sStack, w.OMC, nP));
    }
    PFML_TRY(a9_stage(opt.x_r_tilde[g], opt.x_denom[g], risk_g, tc_g, opt.x_r_sum ? opt.x_r_sum[g] : nullptr,
                      opt.x_d_sum ? opt.x_d_sum[g] : nullptr));
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------
// m = Lambda^-1/2 m~ Lambda^1/2 for a batch of dates from the factor form (stage a5 of the date-step alone);
// used by the back-test, which the reference feeds with one m_func call per month (PFML_best_hps.py:186-190).
// m_out is [B][NP][NP] (rows / columns >= n_sel[b] belong to the decoupled dummy stocks).
// ------------------------------------------------------------------------------------------------
struct MfWs { BatchVecs v; MtBufs m; double *Ut, *Ev, *G; size_t bytes; };

static MfWs carve_mf(void* ws, size_t cap, int B, int NP, int nt, int K) {
  MfWs w{};
  Carver c(ws, cap);
  const size_t BNN = (size_t)B * NP * NP;
  w.v.sl = c.take<double>((size_t)B * NP); w.v.isl = c.take<double>((size_t)B * NP);
  w.v.g = c.take<double>((size_t)B * nt * NP); w.v.g0raw = c.take<double>((size_t)B * NP);
  w.v.rv = c.take<double>((size_t)B * NP); w.v.dvl = c.take<double>((size_t)B * NP);
  w.v.Lg = c.take<double>((size_t)B * NP * KP); w.v.Lt = c.take<double>((size_t)B * NP * KP);
  w.v.LF = c.take<double>((size_t)B * NP * KP); w.v.Fg = c.take<double>((size_t)B * KP * KP);
  w.v.scal = c.take<double>((size_t)B * NSCAL);
  w.m.x = c.take<double>(BNN); w.m.sgr = c.take<double>(BNN); w.m.yv = c.take<double>((size_t)B * NP);
  w.m.Mdb = nullptr; w.m.Ya = nullptr; w.m.Yb = nullptr;
  w.m.Wk = c.take<double>(BNN); w.m.Xinv = c.take<double>(BNN); w.m.T1 = c.take<double>(BNN); w.m.mt = c.take<double>(BNN);
  w.m.part = c.take<double>((size_t)B * NORM_CHUNKS * 3); w.m.sc = c.take<double>((size_t)B * 8);
  w.Ut = c.take<double>((size_t)B * NP * KP); w.Ev = c.take<double>((size_t)B * NP);
  w.G = c.take<double>((size_t)B * NP * stj_ldg(K));
  w.bytes = c.off;
  return w;
}

size_t mtilde_factor_workspace_bytes(const pfml_panel_t* panel, int B, int n_max) {
  if (!panel || B <= 0 || n_max <= 0) return 0;
  return carve_mf(nullptr, 0, B, pfml_padded_n(n_max), panel->lb + 2, panel->K).bytes;
}

int mtilde_factor_for_dates(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                            double* m_out, int* status, void* workspace, size_t workspace_bytes, cudaStream_t s) {
  if (!panel || B <= 0 || n_max <= 0 || !date_t || !sel || !n_sel || !m_out || !workspace) {
    set_error("pfml_mfunc_dates: null/empty argument");
    return -1;
  }
  if (panel->K > KP || panel->K <= 0) {
    set_error("pfml_mfunc_dates: K = %d factors unsupported (1..%d)", panel->K, KP);
    return -2;
  }
  const PanelView pv = view_of(panel);
  const int NP = pfml_padded_n(n_max);
  MfWs w = carve_mf(workspace, workspace_bytes, B, NP, pv.lb + 2, pv.K);
  if (w.bytes > workspace_bytes) {
    set_error("pfml_mfunc_dates: workspace too small (%zu < %zu bytes)", workspace_bytes, w.bytes);
    return -3;
  }
  if (status) PFML_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(int) * B, s));
  ProfScope ps(T_MFUNC, s);
  PFML_TRY(prep_vectors(pv, B, NP, date_t, sel, n_sel, w.v, s));
  PFML_TRY(build_x_factor(pv, B, NP, date_t, sel, n_sel, w.v, w.m.x, w.m.sgr, w.m.yv, s));
  PFML_TRY(mt0_stieltjes(B, NP, pv.K, n_sel, w.v, w.Ut, w.Ev, w.G, w.m.mt, status, s));
  double* mt = nullptr;
  PFML_TRY(fixed_point(B, NP, pv.iterations, w.m, status, s, &mt));
  PFML_TRY(m_from_mtilde(B, NP, NP, mt, w.v.isl, w.v.sl, m_out, s));
  return 0;
}

}  // namespace pfml

// =================================================================================================
// C ABI
// =================================================================================================
using namespace pfml;

extern "C" {

const char* pfml_version(void) { return "pfml_b200 0.3 (sm_100a, FP64 DMMA + INT8 tcgen05 emulation, NCCL exchange)"; }
const char* pfml_last_error(void) { return last_error(); }

int pfml_padded_n(int n_max) { return (int)round_up(n_max < 1 ? 1 : n_max, BASE); }

size_t pfml_date_steps_workspace_bytes(const pfml_panel_t* panel, int B, int n_max) {
  if (!panel || B <= 0 || n_max <= 0) return 0;
  StepWs w = carve_step(nullptr, 0, B, pfml_padded_n(n_max), panel->p + 1, panel->lb + 2, panel->K, true, true);
  return w.bytes;
}

int pfml_date_steps(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                    double* r_tilde, double* denom, double* risk, double* tc, double* signal_t, int* status,
                    void* workspace, size_t workspace_bytes, void* stream) {
  return date_steps_impl(panel, B, n_max, date_t, sel, n_sel, r_tilde, denom, risk, tc, signal_t, status, workspace,
                         workspace_bytes, (cudaStream_t)stream);
}

int pfml_date_steps_bucketed(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                             const int* bucket, int n_buckets, double* r_sum, double* d_sum, double* r_tilde, double* denom,
                             double* risk, double* tc, double* signal_t, int* status, void* workspace, size_t workspace_bytes,
                             void* stream) {
  if (!bucket || n_buckets <= 0 || !r_sum || !d_sum) {
    set_error("pfml_date_steps_bucketed: null/empty window arguments");
    return -1;
  }
  StepOpts o;
  o.bucket = bucket; o.n_buckets = n_buckets; o.r_sum = r_sum; o.d_sum = d_sum;
  return date_steps_impl(panel, B, n_max, date_t, sel, n_sel, r_tilde, denom, risk, tc, signal_t, status, workspace, workspace_bytes,
                         (cudaStream_t)stream, o);
}

int pfml_date_steps_multi_g(const pfml_panel_t* panel, int G, const double* const* rff_g, int B, int n_max, const int* date_t,
                            const int* sel, const int* n_sel, const int* bucket, int n_buckets, double* const* r_sum_g,
                            double* const* d_sum_g, double* const* r_tilde_g, double* const* denom_g, double* const* risk_g,
                            double* const* tc_g, double* const* signal_t_g, int* status, void* workspace, size_t workspace_bytes,
                            void* stream) {
  if (!panel || G < 1 || G > 16 || !rff_g || !r_tilde_g || !denom_g) {
    set_error("pfml_date_steps_multi_g: null argument or G outside 1..16");
    return -1;
  }
  for (int g = 0; g < G; g++)
    if (!rff_g[g] || !r_tilde_g[g] || !denom_g[g]) { set_error("pfml_date_steps_multi_g: null panel / output for g = %d", g); return -1; }
  if (bucket && (n_buckets <= 0 || !r_sum_g || !d_sum_g)) { set_error("pfml_date_steps_multi_g: window sums requested without buffers"); return -1; }
  pfml_panel_t p0 = *panel;
  p0.rff = rff_g[0];
  const int NP = pfml_padded_n(n_max), P = panel->p + 1;
  // the shared path needs the INT8 direct form (padded universe a multiple of 64 up to the emulation's depth limit, lb <= 17)
  const bool shared = G > 1 && emu_mode() != 0 && emu_shape_ok(NP, P, NP) && emu_shape_ok(NP, NP, NP) && panel->lb - 1 <= 16 && panel->lb + 2 >= 5;
  StepOpts o;
  o.bucket = bucket; o.n_buckets = bucket ? n_buckets : 0;
  o.r_sum = bucket ? r_sum_g[0] : nullptr; o.d_sum = bucket ? d_sum_g[0] : nullptr;
  if (shared) {
    o.n_extra = G - 1;
    o.x_rff = rff_g + 1; o.x_r_tilde = r_tilde_g + 1; o.x_denom = denom_g + 1;
    o.x_risk = risk_g ? risk_g + 1 : nullptr; o.x_tc = tc_g ? tc_g + 1 : nullptr; o.x_signal = signal_t_g ? signal_t_g + 1 : nullptr;
    o.x_r_sum = bucket ? r_sum_g + 1 : nullptr; o.x_d_sum = bucket ? d_sum_g + 1 : nullptr;
  }
  PFML_TRY(date_steps_impl(&p0, B, n_max, date_t, sel, n_sel, r_tilde_g[0], denom_g[0], risk_g ? risk_g[0] : nullptr, tc_g ? tc_g[0] : nullptr,
                           signal_t_g ? signal_t_g[0] : nullptr, status, workspace, workspace_bytes, (cudaStream_t)stream, o));
  if (!shared) {
    // no sharing possible for this shape / mode: one full pass per further bandwidth (status flags accumulate by OR)
    for (int g = 1; g < G; g++) {
      pfml_panel_t pg = *panel;
      pg.rff = rff_g[g];
      StepOpts og;
      og.bucket = bucket; og.n_buckets = bucket ? n_buckets : 0;
      og.r_sum = bucket ? r_sum_g[g] : nullptr; og.d_sum = bucket ? d_sum_g[g] : nullptr;
      og.keep_status = true;
      PFML_TRY(date_steps_impl(&pg, B, n_max, date_t, sel, n_sel, r_tilde_g[g], denom_g[g], risk_g ? risk_g[g] : nullptr,
                               tc_g ? tc_g[g] : nullptr, signal_t_g ? signal_t_g[g] : nullptr, status, workspace, workspace_bytes,
                               (cudaStream_t)stream, og));
    }
  }
  return 0;
}

// ---- stage-level entry points: a6, a3..a8, a9 on their own -----------------------------------------
int pfml_standardize_scale(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                           double* signals, void* workspace, size_t workspace_bytes, void* stream) {
  if (!panel || B <= 0 || n_max <= 0 || !date_t || !sel || !n_sel || !signals || !workspace) {
    set_error("pfml_standardize_scale: null/empty argument");
    return -1;
  }
  if (panel->lb < 0 || panel->p < 2 || (panel->p & 1)) {
    set_error("pfml_standardize_scale: need lb >= 0 and even p >= 2 (lb=%d p=%d)", panel->lb, panel->p);
    return -2;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const PanelView pv = view_of(panel);
  const int NP = pfml_padded_n(n_max), P = pv.p + 1, nt = pv.lb + 2;
  StepWs w = carve_step(workspace, workspace_bytes, B, NP, P, nt, pv.K, true, true);
  if (w.bytes > workspace_bytes) {
    set_error("pfml_standardize_scale: workspace too small (%zu < %zu bytes; size it with pfml_date_steps_workspace_bytes)",
              workspace_bytes, w.bytes);
    return -3;
  }
  ProfScope ps(T_STD, s);
  PFML_TRY(prep_vectors(pv, B, NP, date_t, sel, n_sel, w.v, s));
  return standardize(pv, B, NP, P, w.ldP, date_t, sel, n_sel, w.v, w.St, nullptr, false, nt, s, signals);
}

int pfml_omega_batched(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                       double* omega, double* omega_chg, double* m, int* status, void* workspace, size_t workspace_bytes,
                       void* stream) {
  if (!omega || !omega_chg) {
    set_error("pfml_omega_batched: null output");
    return -1;
  }
  StepOpts o;
  o.omega_out = omega; o.omega_chg_out = omega_chg; o.m_out = m; o.stop_after_omega = true;
  return date_steps_impl(panel, B, n_max, date_t, sel, n_sel, nullptr, nullptr, nullptr, nullptr, nullptr, status, workspace,
                         workspace_bytes, (cudaStream_t)stream, o);
}

int pfml_rtilde_denom_batched(const pfml_panel_t* panel, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
                              const double* omega, const double* omega_chg, double* r_tilde, double* denom, double* risk,
                              double* tc, int* status, void* workspace, size_t workspace_bytes, void* stream) {
  if (!omega || !omega_chg) {
    set_error("pfml_rtilde_denom_batched: null input");
    return -1;
  }
  StepOpts o;
  o.omega_in = omega; o.omega_chg_in = omega_chg;
  return date_steps_impl(panel, B, n_max, date_t, sel, n_sel, r_tilde, denom, risk, tc, nullptr, status, workspace,
                         workspace_bytes, (cudaStream_t)stream, o);
}

// ---- m_func ----------------------------------------------------------------------------------------
static MtBufs carve_mt(Carver& c, int B, int NP, double** isl, double** sl) {
  MtBufs m{};
  const size_t BNN = (size_t)B * NP * NP;
  m.x = c.take<double>(BNN); m.sgr = c.take<double>(BNN); m.yv = c.take<double>((size_t)B * NP);
  m.Mdb = c.take<double>(BNN); m.Wk = c.take<double>(BNN); m.Xinv = c.take<double>(BNN);
  m.T1 = c.take<double>(BNN); m.Ya = c.take<double>(BNN); m.Yb = c.take<double>(BNN); m.mt = c.take<double>(BNN);
  m.part = c.take<double>((size_t)B * NORM_CHUNKS * 3);
  m.sc = c.take<double>((size_t)B * 8);
  *isl = c.take<double>((size_t)B * NP);
  *sl = c.take<double>((size_t)B * NP);
  return m;
}

size_t pfml_mfunc_workspace_bytes(int B, int n) {
  if (B <= 0 || n <= 0) return 0;
  Carver c(nullptr, 0);
  double *a, *b;
  carve_mt(c, B, pfml_padded_n(n), &a, &b);
  return c.off;
}

int pfml_mfunc_batched(int B, int n, const double* sigma_gam, const double* lam, const double* w, const double* rf, double mu,
                       double gam, int iterations, double* m_out, int* status, void* workspace, size_t workspace_bytes,
                       void* stream) {
  if (B <= 0 || n <= 0 || !sigma_gam || !lam || !w || !rf || !m_out || !workspace || iterations < 0) {
    set_error("pfml_mfunc_batched: null/empty argument");
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int NP = pfml_padded_n(n);
  Carver c(workspace, workspace_bytes);
  double *isl, *sl;
  MtBufs m = carve_mt(c, B, NP, &isl, &sl);
  if (c.off > workspace_bytes) {
    set_error("pfml_mfunc_batched: workspace too small (%zu < %zu bytes)", workspace_bytes, c.off);
    return -3;
  }
  if (status) PFML_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(int) * B, s));
  ProfScope ps(T_MFUNC, s);
  PFML_TRY(build_x_dense(B, n, NP, sigma_gam, lam, w, rf, mu, gam, m.x, m.sgr, m.yv, isl, sl, s));
  double* mt = nullptr;
  PFML_TRY(mtilde(B, NP, iterations, m, status, s, &mt));
  PFML_TRY(m_from_mtilde(B, n, NP, mt, isl, sl, m_out, s));
  return 0;
}

// ---- generic GEMM / inverse --------------------------------------------------------------------------
int pfml_dgemm_batched(int trans_a, int trans_b, int M, int N, int K, double alpha, const double* A, long lda, long sA,
                       const double* B, long ldb, long sB, double beta, double* D, long ldd, long sD, int batch, void* stream) {
  GemmP g;
  g.A = A; g.lda = lda; g.sA = sA; g.B = B; g.ldb = ldb; g.sB = sB; g.D = D; g.ldd = ldd; g.sD = sD;
  g.M = M; g.N = N; g.K = K; g.alpha = alpha;
  g.a_kinner = !trans_a; g.b_kinner = trans_b != 0;
  if (beta != 0.0) { g.E = D; g.lde = ldd; g.sE = sD; g.beta = beta; }
  return gemm(g, batch, (cudaStream_t)stream);
}

// ---- FP64 GEMM emulated on the INT8 tensor cores (emu.cu) -------------------------------------------------------
struct EmuWs { int8_t *As, *Bs; double *sa, *sb; int *ex, *err; size_t bytes; };
static EmuWs carve_emu(void* ws, size_t cap, int M, int N, int K, int batch) {
  EmuWs w{};
  Carver c(ws, cap);
  w.As = c.take<int8_t>(emu_a_slices_bytes(batch, M, K));
  w.Bs = c.take<int8_t>(emu_b_slices_bytes(batch, K, N));
  w.sa = c.take<double>((size_t)batch * M); w.sb = c.take<double>((size_t)batch * N);
  w.ex = c.take<int>((size_t)batch * N); w.err = c.take<int>(EMU_ERR_SLOT0 + (size_t)batch);
  w.bytes = c.off;
  return w;
}

size_t pfml_dgemm_emulated_workspace_bytes(int M, int N, int K, int batch) {
  if (!emu_shape_ok(M, N, K) || batch <= 0) return 0;
  return carve_emu(nullptr, 0, M, N, K, batch).bytes;
}

int pfml_dgemm_emulated(int M, int N, int K, double alpha, const double* A, long lda, long sA, const double* B, long ldb, long sB,
                        double beta, double* D, long ldd, long sD, int batch, int* status, void* workspace, size_t workspace_bytes,
                        void* stream) {
  if (!A || !B || !D || !workspace || batch <= 0) {
    set_error("pfml_dgemm_emulated: null/empty argument");
    return -1;
  }
  if (!emu_shape_ok(M, N, K)) {
    set_error("pfml_dgemm_emulated: needs K %% 64 == 0 and K <= 16640 (got M=%d N=%d K=%d)", M, N, K);
    return -2;
  }
  cudaStream_t s = (cudaStream_t)stream;
  EmuWs w = carve_emu(workspace, workspace_bytes, M, N, K, batch);
  if (w.bytes > workspace_bytes) {
    set_error("pfml_dgemm_emulated: workspace too small (%zu < %zu bytes)", workspace_bytes, w.bytes);
    return -3;
  }
  PFML_CUDA_OK(cudaMemsetAsync(w.err, 0, sizeof(int) * (EMU_ERR_SLOT0 + (size_t)batch), s));
  PFML_TRY(emu_slice_a(A, lda, sA, M, K, batch, w.As, w.sa, w.err, s));
  PFML_TRY(emu_slice_b(B, ldb, sB, K, N, batch, w.Bs, w.sb, w.ex, w.err, s));
  PFML_TRY(emu_mma(w.As, w.sa, w.Bs, w.sb, M, N, K, batch, alpha, nullptr, 0, beta, beta != 0.0 ? D : nullptr, ldd, sD, D, ldd, sD,
                   w.err, s));
  if (status) {
    PFML_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(int), s));
    fold_emu_error_kernel<<<ceil_div(batch, 128), 128, 0, s>>>(batch, w.err, status, 1);
    PFML_CHECK_LAUNCH();
  }
  return 0;
}

void pfml_set_emulation(int on) { g_use_emu.store((on < 0 || on > 2) ? 1 : on, std::memory_order_relaxed); }

namespace {
__global__ void pad_identity_kernel(int n, int NP, const double* __restrict__ A, double* __restrict__ Wk) {
  const int b = blockIdx.z, r = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NP) return;
  Wk[((long)b * NP + r) * NP + c] = (r < n && c < n) ? A[((long)b * n + r) * n + c] : (r == c ? 1.0 : 0.0);
}
__global__ void unpad_kernel(int n, int NP, const double* __restrict__ X, double* __restrict__ out) {
  const int b = blockIdx.z, r = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  out[((long)b * n + r) * n + c] = X[((long)b * NP + r) * NP + c];
}
}  // namespace

size_t pfml_inverse_workspace_bytes(int B, int n) {
  if (B <= 0 || n <= 0) return 0;
  const size_t NP = pfml_padded_n(n);
  return 3 * (((size_t)B * NP * NP * sizeof(double) + 255) / 256 * 256);
}

int pfml_inverse_batched(int B, int n, const double* A, double* Ainv, int* status, void* workspace, size_t workspace_bytes,
                         void* stream) {
  if (B <= 0 || n <= 0 || !A || !Ainv || !workspace) {
    set_error("pfml_inverse_batched: null/empty argument");
    return -1;
  }
  if (workspace_bytes < pfml_inverse_workspace_bytes(B, n)) {
    set_error("pfml_inverse_batched: workspace too small");
    return -3;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int NP = pfml_padded_n(n);
  Carver c(workspace, workspace_bytes);
  const size_t BNN = (size_t)B * NP * NP;
  double* Wk = c.take<double>(BNN);
  double* X = c.take<double>(BNN);
  double* T1 = c.take<double>(BNN);
  if (status) PFML_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(int) * B, s));
  pad_identity_kernel<<<dim3(ceil_div(NP, 128), NP, B), 128, 0, s>>>(n, NP, A, Wk);
  PFML_CHECK_LAUNCH();
  InvCtx ic{B, NP, Wk, X, T1, status, s};
  PFML_TRY(inverse(ic));
  unpad_kernel<<<dim3(ceil_div(n, 128), n, B), 128, 0, s>>>(n, NP, X, Ainv);
  PFML_CHECK_LAUNCH();
  return 0;
}

// ---- a1 / a2 ----------------------------------------------------------------------------------------
int pfml_rff_sincos(const double* X, const double* W, long rows, int F, int p_half, double* out, void* stream) {
  if (!X || !W || !out || rows <= 0 || F <= 0 || p_half <= 0) {
    set_error("pfml_rff_sincos: null/empty argument");
    return -1;
  }
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(T_RFF, s);
  const long chunk = 1L << 20;     // grid.y limit: 65535 row tiles
  for (long r0 = 0; r0 < rows; r0 += chunk) {
    GemmP g;
    g.A = X + r0 * F; g.lda = F; g.B = W; g.ldb = p_half; g.D = out + r0 * 2L * p_half; g.ldd = 2L * p_half;
    g.M = (int)((rows - r0) < chunk ? (rows - r0) : chunk); g.N = p_half; g.K = F;
    g.sincos = 1;
    PFML_TRY(gemm(g, 1, s));
  }
  return 0;
}

int pfml_growth_ratio(const double* tr_ld0, const double* mu_ld0, long n, double* gt, void* stream) {
  if (!tr_ld0 || !mu_ld0 || !gt || n <= 0) {
    set_error("pfml_growth_ratio: null/empty argument");
    return -1;
  }
  return gt_panel(n, tr_ld0, mu_ld0, gt, (cudaStream_t)stream);
}

int pfml_vol_scale(const double* fct_load, const double* fct_cov, const double* ivol, const unsigned char* has, int T, int N,
                   int K, double* vol_scale, void* stream) {
  if (!fct_load || !fct_cov || !ivol || !vol_scale || T <= 0 || N <= 0 || K <= 0) {
    set_error("pfml_vol_scale: null/empty argument");
    return -1;
  }
  return vol_scale_panel(fct_load, fct_cov, ivol, has, T, N, K, vol_scale, (cudaStream_t)stream);
}

// ---- profiling ---------------------------------------------------------------------------------------
void pfml_prof_enable(int on) { g_prof_on.store(on != 0); }
int pfml_prof_ntags(void) { return N_TAGS; }
const char* pfml_prof_tag_name(int tag) { return (tag >= 0 && tag < N_TAGS) ? kTagNames[tag] : "?"; }
long pfml_launch_count(void) { return g_launch_count.load(std::memory_order_relaxed); }

int pfml_prof_collect(double* ms, double* flops, double* launches) {
  PFML_CUDA_OK(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (int i = 0; i < N_TAGS; i++) { ms[i] = 0; flops[i] = 0; launches[i] = 0; }
  for (auto& r : g_recs) {
    float t = 0.f;
    cudaEventElapsedTime(&t, r.e0, r.e1);
    ms[r.slot] += t; flops[r.slot] += r.flops; launches[r.slot] += 1;
    cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
  }
  g_recs.clear();
  for (int i = 0; i < N_TAGS; i++) {
    for (auto e : g_open[i]) cudaEventDestroy(e);
    g_open[i].clear();
  }
  return 0;
}

}  // extern "C"
