// pfml_date_steps_host: the end-to-end entry point over HOST buffers (see include/pfml_b200.h).
//
// One call = one block of consecutive months: upload the block's raw inputs, build the panel-level
// quantities on the device (RFF features a1, vol_scale a2, g_t), then run the requested date-steps in
// chunks.  Chunk c computes on the compute stream into staging buffer c%2 while the copy stream drains
// chunk c-1 to the caller's host buffers, so PCIe D2H (29 MB per date at p = 1024) overlaps the DMMA work.
// Device memory comes from a grow-only per-device arena that is reused across calls.
#include <cstdlib>
#include <mutex>
#include <vector>

#include "kernels.cuh"
#include "pfml_b200.h"

using namespace pfml;

namespace {

constexpr int MAX_CS = 4;           // compute streams (chunks in flight)

struct Arena {
  void* base = nullptr;
  size_t cap = 0;
  cudaStream_t compute[MAX_CS] = {}, copy = nullptr;
  cudaEvent_t panel_ready = nullptr;
  cudaEvent_t computed[MAX_CS] = {}, drained[MAX_CS] = {};
  bool init = false;
};
Arena g_arena[64];
std::mutex g_arena_mu;

struct Bump {
  char* base; size_t off = 0;
  explicit Bump(void* b) : base((char*)b) {}
  template <class T> T* take(size_t n) {
    size_t bytes = (n * sizeof(T) + 255) / 256 * 256;
    T* r = base ? (T*)(base + off) : nullptr;
    off += bytes;
    return r;
  }
};

struct HostPlan {
  double *feats, *W, *tr, *mu, *lam, *ret, *fct_load, *fct_cov, *ivol, *wealth, *rf, *rff, *vol_scale, *gt;
  unsigned char* has;
  int *date_t, *sel, *n_sel, *status;
  double* stage[2];
  void* ws[MAX_CS];
  size_t ws_bytes, stage_doubles, total;
};

HostPlan plan(void* base, const pfml_host_block_t* b, int B, int NP, int chunk, bool want_risk, bool want_tc, bool want_sig,
              size_t ws_bytes, int n_cs) {
  HostPlan h{};
  Bump m(base);
  const size_t TN = (size_t)b->T * b->N, P = b->p + 1;
  h.feats = m.take<double>(TN * b->F);
  h.W = m.take<double>((size_t)b->F * (b->p / 2));
  h.tr = m.take<double>(TN); h.mu = m.take<double>(TN); h.lam = m.take<double>(TN); h.ret = m.take<double>(TN);
  h.fct_load = m.take<double>(TN * b->K);
  h.fct_cov = m.take<double>((size_t)b->T * b->K * b->K);
  h.ivol = m.take<double>(TN);
  h.wealth = m.take<double>(b->T); h.rf = m.take<double>(b->T);
  h.rff = m.take<double>(TN * b->p);
  h.vol_scale = m.take<double>(TN); h.gt = m.take<double>(TN);
  h.has = m.take<unsigned char>(TN);
  h.date_t = m.take<int>(B); h.sel = m.take<int>((size_t)B * NP); h.n_sel = m.take<int>(B); h.status = m.take<int>(B);
  // device copy of ALL outputs of the call (B dates): no staging buffer has to be recycled, so a chunk never waits
  // for an earlier chunk's download
  h.stage_doubles = (size_t)B * (P + P * P * (1 + (want_risk ? 1 : 0) + (want_tc ? 1 : 0)) + (want_sig ? (size_t)NP * P : 0));
  h.stage[0] = m.take<double>(h.stage_doubles);
  h.stage[1] = nullptr;
  h.ws_bytes = ws_bytes;
  for (int i = 0; i < n_cs; i++) h.ws[i] = m.take<char>(ws_bytes);
  h.total = m.off;
  return h;
}

}  // namespace

extern "C" int pfml_date_steps_host(const pfml_host_block_t* blk, int B, int n_max, const int* date_t, const int* sel,
                                    const int* n_sel, double* r_tilde, double* denom, double* risk, double* tc,
                                    double* signal_t, int* status, int chunk, int device) {
  if (!blk || B <= 0 || n_max <= 0 || !date_t || !sel || !n_sel || !r_tilde || !denom) {
    set_error("pfml_date_steps_host: null/empty argument");
    return -1;
  }
  if (device < 0 || device >= 64) {
    set_error("pfml_date_steps_host: bad device %d", device);
    return -1;
  }
  // Chunks: uniform, default ceil(B / 6).  Consecutive chunks alternate between two compute streams (each with its
  // own workspace and staging buffer): the latency-bound kernels of one chunk (block inverses, small GEMMs of the
  // recursion) overlap the DMMA-bound kernels of the other, and the copy stream drains finished chunks meanwhile.
  std::vector<int> sched;
  if (const char* e = getenv("PFML_HOST_SCHED")) {        // tuning aid: explicit comma-separated chunk sizes
    int left = B;
    for (const char* q = e; *q && left > 0;) {
      int v = atoi(q);
      if (v <= 0) break;
      v = v < left ? v : left;
      sched.push_back(v); left -= v;
      while (*q && *q != ',') q++;
      if (*q == ',') q++;
    }
    if (left > 0) sched.push_back(left);
  } else {
    // measured best on B200 (tools/e2e_chunk_sweep.py): B / 6 when every output is downloaded (29 MB per date: the copy
    // engine must start early and stay busy), B / 3 when only r_tilde and denom are (8 MB per date: larger chunks
    // compute more efficiently)
    const bool all_outputs = risk || tc || signal_t;
    if (chunk <= 0) chunk = B <= 40 ? B : (all_outputs ? (B + 5) / 6 : (B + 2) / 3);
    for (int b0 = 0; b0 < B; b0 += chunk) sched.push_back((B - b0) < chunk ? (B - b0) : chunk);
  }
  chunk = 0;
  for (int c : sched) chunk = c > chunk ? c : chunk;      // buffers are sized for the largest chunk
  std::lock_guard<std::mutex> lk(g_arena_mu);
  PFML_CUDA_OK(cudaSetDevice(device));
  Arena& a = g_arena[device];
  if (!a.init) {
    PFML_CUDA_OK(cudaStreamCreateWithFlags(&a.copy, cudaStreamNonBlocking));
    PFML_CUDA_OK(cudaEventCreateWithFlags(&a.panel_ready, cudaEventDisableTiming));
    for (int i = 0; i < MAX_CS; i++) {
      PFML_CUDA_OK(cudaStreamCreateWithFlags(&a.compute[i], cudaStreamNonBlocking));
      PFML_CUDA_OK(cudaEventCreateWithFlags(&a.computed[i], cudaEventDisableTiming));
      PFML_CUDA_OK(cudaEventCreateWithFlags(&a.drained[i], cudaEventDisableTiming));
    }
    a.init = true;
  }
  static const int env_cs = getenv("PFML_HOST_STREAMS") ? atoi(getenv("PFML_HOST_STREAMS")) : 2;   // tuning aid
  const int n_cs = env_cs < 1 ? 1 : (env_cs > MAX_CS ? MAX_CS : env_cs);
  const int NP = pfml_padded_n(n_max);
  const size_t TN = (size_t)blk->T * blk->N, P = blk->p + 1;
  pfml_panel_t pan{};
  pan.T = blk->T; pan.N = blk->N; pan.F = blk->F; pan.p = blk->p; pan.K = blk->K;
  pan.gamma_rel = blk->gamma_rel; pan.mu = blk->mu; pan.lb = blk->lb; pan.iterations = blk->iterations;
  const size_t ws_bytes = pfml_date_steps_workspace_bytes(&pan, chunk, n_max);
  HostPlan h = plan(nullptr, blk, B, NP, chunk, risk != nullptr, tc != nullptr, signal_t != nullptr, ws_bytes, n_cs);
  if (h.total > a.cap) {
    if (a.base) PFML_CUDA_OK(cudaFree(a.base));
    a.base = nullptr; a.cap = 0;
    PFML_CUDA_OK(cudaMalloc(&a.base, h.total));
    a.cap = h.total;
  }
  h = plan(a.base, blk, B, NP, chunk, risk != nullptr, tc != nullptr, signal_t != nullptr, ws_bytes, n_cs);
  cudaStream_t cs = a.compute[0], ds = a.copy;
  cudaStream_t* cstream = a.compute;

  // ---- upload the block ---------------------------------------------------------------------------
#define H2D(dst, src, count, type) PFML_CUDA_OK(cudaMemcpyAsync(dst, src, (size_t)(count) * sizeof(type), cudaMemcpyHostToDevice, cs))
  H2D(h.feats, blk->feats, TN * blk->F, double);
  H2D(h.W, blk->W, (size_t)blk->F * (blk->p / 2), double);
  H2D(h.tr, blk->tr_ld0, TN, double);
  H2D(h.mu, blk->mu_ld0, TN, double);
  H2D(h.lam, blk->lam, TN, double);
  H2D(h.ret, blk->ret_ld1, TN, double);
  H2D(h.fct_load, blk->fct_load, TN * blk->K, double);
  H2D(h.fct_cov, blk->fct_cov, (size_t)blk->T * blk->K * blk->K, double);
  H2D(h.ivol, blk->ivol, TN, double);
  H2D(h.wealth, blk->wealth, blk->T, double);
  H2D(h.rf, blk->rf, blk->T, double);
  if (blk->barra_has) H2D(h.has, blk->barra_has, TN, unsigned char);
  H2D(h.date_t, date_t, B, int);
  H2D(h.sel, sel, (size_t)B * NP, int);
  H2D(h.n_sel, n_sel, B, int);
#undef H2D

  // ---- panel-level device work: a1 RFF, a2 vol_scale, g_t -----------------------------------------
  PFML_TRY(pfml_rff_sincos(h.feats, h.W, (long)TN, blk->F, blk->p / 2, h.rff, cs));
  PFML_TRY(pfml_vol_scale(h.fct_load, h.fct_cov, h.ivol, blk->barra_has ? h.has : nullptr, blk->T, blk->N, blk->K, h.vol_scale, cs));
  PFML_TRY(pfml_growth_ratio(h.tr, h.mu, (long)TN, h.gt, cs));
  pan.rff = h.rff; pan.vol_scale = h.vol_scale; pan.gt = h.gt; pan.lam = h.lam; pan.ret_ld1 = h.ret;
  pan.fct_load = h.fct_load; pan.fct_cov = h.fct_cov; pan.ivol = h.ivol; pan.wealth = h.wealth; pan.rf = h.rf;

  PFML_CUDA_OK(cudaEventRecord(a.panel_ready, cs));
  for (int i = 1; i < n_cs; i++) PFML_CUDA_OK(cudaStreamWaitEvent(a.compute[i], a.panel_ready, 0));
  // ---- chunks: compute on two alternating streams, drain on `ds` -----------------------------------
  const int n_chunks = (int)sched.size();
  const bool trace = getenv("PFML_HOST_TRACE") != nullptr;
  std::vector<cudaEvent_t> tev;
  auto mark = [&](cudaStream_t st) { if (trace) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); tev.push_back(e); } };
  mark(cs);
  int b0 = 0;
  for (int c = 0; c < n_chunks; b0 += sched[c], c++) {
    const int nb = sched[c], sb = c % n_cs;
    double* st = h.stage[0];
    double* d_rt = st + (size_t)b0 * P;                     st += (size_t)B * P;
    double* d_den = st + (size_t)b0 * P * P;                st += (size_t)B * P * P;
    double* d_risk = nullptr; if (risk) { d_risk = st + (size_t)b0 * P * P; st += (size_t)B * P * P; }
    double* d_tc = nullptr;   if (tc) { d_tc = st + (size_t)b0 * P * P; st += (size_t)B * P * P; }
    double* d_sig = nullptr;  if (signal_t) { d_sig = st + (size_t)b0 * NP * P; st += (size_t)B * NP * P; }
    cudaStream_t ks = cstream[sb];
    PFML_TRY(pfml_date_steps(&pan, nb, n_max, h.date_t + b0, h.sel + (size_t)b0 * NP, h.n_sel + b0, d_rt, d_den, d_risk, d_tc,
                             d_sig, h.status + b0, h.ws[sb], h.ws_bytes, ks));
    PFML_CUDA_OK(cudaEventRecord(a.computed[sb], ks));
    mark(ks);
    PFML_CUDA_OK(cudaStreamWaitEvent(ds, a.computed[sb], 0));
#define D2H(dst, src, count) PFML_CUDA_OK(cudaMemcpyAsync(dst, src, (size_t)(count) * sizeof(double), cudaMemcpyDeviceToHost, ds))
    D2H(r_tilde + (size_t)b0 * P, d_rt, (size_t)nb * P);
    D2H(denom + (size_t)b0 * P * P, d_den, (size_t)nb * P * P);
    if (risk) D2H(risk + (size_t)b0 * P * P, d_risk, (size_t)nb * P * P);
    if (tc) D2H(tc + (size_t)b0 * P * P, d_tc, (size_t)nb * P * P);
    if (signal_t) D2H(signal_t + (size_t)b0 * NP * P, d_sig, (size_t)nb * NP * P);
#undef D2H
    PFML_CUDA_OK(cudaEventRecord(a.drained[sb], ds));
    mark(ds);
  }
  for (int i = 0; i < n_cs; i++) PFML_CUDA_OK(cudaStreamSynchronize(a.compute[i]));
  if (status) PFML_CUDA_OK(cudaMemcpyAsync(status, h.status, sizeof(int) * B, cudaMemcpyDeviceToHost, ds));
  PFML_CUDA_OK(cudaStreamSynchronize(ds));
  if (trace) {
    fprintf(stderr, "[pfml host trace] panel ready at 0 ms;");
    for (size_t i = 1; i < tev.size(); i++) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, tev[0], tev[i]);
      fprintf(stderr, " %s%zu %.1f", (i & 1) ? "computed" : "drained", (i - 1) / 2, ms);
    }
    fprintf(stderr, "\n");
    for (auto e : tev) cudaEventDestroy(e);
  }
  return 0;
}
