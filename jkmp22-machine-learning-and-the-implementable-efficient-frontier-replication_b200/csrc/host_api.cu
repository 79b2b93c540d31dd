// pfml_date_steps_host / pfml_date_steps_host_packed: the end-to-end entry points over HOST buffers (include/pfml_b200.h).
//
// One call = one block of consecutive months: upload the block's raw inputs, build the panel-level quantities on the
// device (RFF features a1, vol_scale a2, g_t), then run the requested date-steps in chunks.  Chunk c computes on a
// compute stream while the copy stream drains the finished chunks to the host.
//
// What crosses PCIe on the way back: risk, tc and denom are symmetric and denom = risk + tc exactly, so of the three
// P x P squares the reference stores per date (PFML_Input_Data.py:484-491) only the upper triangles of risk and tc are
// information.  A pack kernel gathers those two triangles (2 x P (P + 1) / 2 doubles instead of 3 P^2: 8.4 MB instead of
// 25.2 MB per date at P = 1025) and
//   * pfml_date_steps_host_packed hands them to the caller as they are (the layout of the on-disk tensor store), while
//   * pfml_date_steps_host downloads the full squares by default; with PFML_HOST_REBUILD=1 it receives the triangles and
//     rebuilds the three squares in the caller's buffers with a few host threads, chunk by chunk, while the GPU is still
//     computing the later chunks (bit-identical to the squares the device holds: the mirror is a copy and denom is the
//     same IEEE addition the device performs).  Measured slower than the plain download on the 16-vCPU box, hence opt-in.
// Device memory comes from a grow-only arena per GPU that is reused across calls; the packed triangles of the full-
// square entry point land in a grow-only pinned host staging buffer, also per GPU.  Calls for the same GPU serialise
// on that GPU's arena; calls for different GPUs run concurrently.
#include <emmintrin.h>

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "kernels.cuh"
#include "pfml_b200.h"

using namespace pfml;

namespace {

constexpr int MAX_CS = 4;           // compute streams (chunks in flight)
constexpr int MAX_CHUNKS = 64;

struct Arena {
  std::mutex mu;
  void* base = nullptr;
  size_t cap = 0;
  void* pinned = nullptr;           // host staging of the packed triangles (full-square entry point)
  size_t pinned_cap = 0;
  cudaStream_t compute[MAX_CS] = {}, copy = nullptr;
  cudaEvent_t panel_ready = nullptr;
  cudaEvent_t computed[MAX_CS] = {}, drained[MAX_CHUNKS] = {};
  bool init = false;
};
Arena g_arena[64];

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

struct Want { bool risk, tc, sig; int n_ut; };   // n_ut packed triangles per date travel (0 = full squares, tuning aid)

struct HostPlan {
  double *feats, *W, *tr, *mu, *lam, *ret, *fct_load, *fct_cov, *ivol, *wealth, *rf, *rff, *vol_scale, *gt;
  unsigned char* has;
  int *date_t, *sel, *n_sel, *status;
  double *rt, *den, *risk, *tc, *sig, *ut;      // device copies of ALL outputs of the call (B dates) + packed triangles
  void* ws[MAX_CS];
  size_t ws_bytes, total;
};

HostPlan plan(void* base, const pfml_host_block_t* b, int B, int NP, const Want& w, size_t ws_bytes, int n_cs) {
  HostPlan h{};
  Bump m(base);
  const size_t TN = (size_t)b->T * b->N, P = b->p + 1, PT = P * (P + 1) / 2;
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
  // no staging buffer is recycled within a call, so a chunk never waits for an earlier chunk's download
  h.rt = m.take<double>((size_t)B * P);
  h.den = m.take<double>((size_t)B * P * P);
  h.risk = w.risk ? m.take<double>((size_t)B * P * P) : nullptr;
  h.tc = w.tc ? m.take<double>((size_t)B * P * P) : nullptr;
  h.sig = w.sig ? m.take<double>((size_t)B * NP * P) : nullptr;
  h.ut = w.n_ut ? m.take<double>((size_t)B * w.n_ut * PT) : nullptr;
  h.ws_bytes = ws_bytes;
  for (int i = 0; i < n_cs; i++) h.ws[i] = m.take<char>(ws_bytes);
  h.total = m.off;
  return h;
}

// upper triangle of src [nb][P][P], row-major, into dst [nb][stride]: row r starts at r P - r (r - 1) / 2
__global__ void pack_upper_kernel(int P, const double* __restrict__ src, double* __restrict__ dst, long dst_stride) {
  const int r = blockIdx.y, b = blockIdx.z;
  const int c = r + blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= P) return;
  const long off = (long)r * P - (long)r * (r - 1) / 2;
  dst[(long)b * dst_stride + off + (c - r)] = src[((long)b * P + r) * P + c];
}

// ---- host side: full squares from packed upper triangles ---------------------------------------------------------
// Lower triangle of A [P][P] from its upper triangle, in place: 64 x 64 blocks, each transposed with 2 x 2 SSE2 shuffles
// into an L1/L2-resident buffer and written out as 64 contiguous doubles per destination row.
void mirror_lower(int P, double* __restrict__ A) {
  constexpr int H = 64, W = 64;
  alignas(64) double tt[W][H];
  for (int r0 = 0; r0 < P; r0 += H) {
    const int r1 = r0 + H < P ? r0 + H : P, h = r1 - r0;
    for (int r = r0; r < r1; r++)                                    // the diagonal block
      for (int c = r + 1; c < r1; c++) A[(long)c * P + r] = A[(long)r * P + c];
    for (int c0 = r1; c0 < P; c0 += W) {
      const int c1 = c0 + W < P ? c0 + W : P, w = c1 - c0;
      const int w2 = w & ~1, h2 = h & ~1;
      for (int r = 0; r < h2; r += 2) {
        const double* s0 = A + (long)(r0 + r) * P + c0;
        const double* s1 = s0 + P;
        for (int c = 0; c < w2; c += 2) {
          const __m128d x0 = _mm_loadu_pd(s0 + c), x1 = _mm_loadu_pd(s1 + c);
          _mm_store_pd(&tt[c][r], _mm_unpacklo_pd(x0, x1));
          _mm_store_pd(&tt[c + 1][r], _mm_unpackhi_pd(x0, x1));
        }
        if (w2 < w) { tt[w2][r] = s0[w2]; tt[w2][r + 1] = s1[w2]; }
      }
      if (h2 < h) {
        const double* s0 = A + (long)(r0 + h2) * P + c0;
        for (int c = 0; c < w; c++) tt[c][h2] = s0[c];
      }
      for (int c = 0; c < w; c++) memcpy(A + (long)(c0 + c) * P + r0, tt[c], (size_t)h * sizeof(double));
    }
  }
}

// A (and Bm, and S = A + Bm) [P][P] from the packed upper triangles a (and b): the upper rows are contiguous copies
// (S: one IEEE addition per element, the same the device performs), the lower triangles are mirrors.
void unpack_date(int P, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ A, double* __restrict__ Bm,
                 double* __restrict__ S) {
  for (int r = 0; r < P; r++) {
    const long off = (long)r * P - (long)r * (r - 1) / 2;            // a[off + c - r] = element (r, c), c >= r
    const int len = P - r;
    memcpy(A + (long)r * P + r, a + off, (size_t)len * sizeof(double));
    if (b) {
      memcpy(Bm + (long)r * P + r, b + off, (size_t)len * sizeof(double));
      if (S) {
        double* __restrict__ sr = S + (long)r * P + r;
        const double* __restrict__ ar = a + off;
        const double* __restrict__ br = b + off;
        for (int c = 0; c < len; c++) sr[c] = ar[c] + br[c];
      }
    }
  }
  mirror_lower(P, A);
  if (b) {
    mirror_lower(P, Bm);
    if (S) mirror_lower(P, S);
  }
}

int unpack_threads() {
  if (const char* e = getenv("PFML_HOST_UNPACK_THREADS")) {
    const int v = atoi(e);
    if (v > 0) return v > 64 ? 64 : v;
  }
  int hw = (int)std::thread::hardware_concurrency();
  if (hw <= 0) hw = 8;
  int ranks = 1;                                          // torchrun exports the number of processes of this host
  if (const char* e = getenv("LOCAL_WORLD_SIZE")) ranks = atoi(e) > 0 ? atoi(e) : 1;
  int t = hw / ranks - 1;
  return t < 2 ? 2 : (t > 16 ? 16 : t);
}

// Host-side argument check before anything is uploaded (O(B NP)): bad dates or indices would otherwise be
// out-of-bounds reads on the device.
int validate(const char* fn, const pfml_host_block_t* blk, int B, int n_max, const int* date_t, const int* sel, const int* n_sel, int NP) {
  if (blk->T <= 0 || blk->N <= 0 || blk->F <= 0) { set_error("%s: empty block (T=%d N=%d F=%d)", fn, blk->T, blk->N, blk->F); return -2; }
  if (blk->K <= 0 || blk->K > KP) { set_error("%s: K = %d factors unsupported (1..%d)", fn, blk->K, KP); return -2; }
  if (blk->lb < 2 || blk->p < 2 || (blk->p & 1)) { set_error("%s: need lb >= 2 and even p >= 2 (lb=%d p=%d)", fn, blk->lb, blk->p); return -2; }
  if (blk->iterations < 0) { set_error("%s: iterations = %d", fn, blk->iterations); return -2; }
  if (!blk->feats || !blk->W || !blk->tr_ld0 || !blk->mu_ld0 || !blk->lam || !blk->ret_ld1 || !blk->fct_load || !blk->fct_cov ||
      !blk->ivol || !blk->wealth || !blk->rf) { set_error("%s: null input array in the block", fn); return -1; }
  if (n_max > blk->N) { set_error("%s: n_max = %d exceeds the block's N = %d stocks", fn, n_max, blk->N); return -2; }
  for (int b = 0; b < B; b++) {
    if (date_t[b] < blk->lb + 1 || date_t[b] >= blk->T) {
      set_error("%s: date_t[%d] = %d outside [lb + 1 = %d, T = %d): the look-back months must lie inside the block", fn, b, date_t[b],
                blk->lb + 1, blk->T);
      return -2;
    }
    if (n_sel[b] <= 0 || n_sel[b] > n_max) { set_error("%s: n_sel[%d] = %d outside (0, n_max = %d]", fn, b, n_sel[b], n_max); return -2; }
    const int* s = sel + (size_t)b * NP;
    for (int i = 0; i < n_sel[b]; i++)
      if (s[i] < 0 || s[i] >= blk->N) { set_error("%s: sel[%d][%d] = %d outside [0, N = %d)", fn, b, i, s[i], blk->N); return -2; }
  }
  return 0;
}

struct DeviceGuard {                  // the caller's current device is restored on every return path
  int prev = -1;
  DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct Outs {                         // caller's host buffers
  double *r_tilde, *denom, *risk, *tc, *signal_t;     // full squares (denom / risk / tc may be null in packed mode)
  double *risk_ut, *tc_ut;                            // packed mode
  int* status;
};

int run_chunks(Arena& a, const HostPlan& h, pfml_panel_t& pan, const std::vector<int>& sched, int n_cs, int B, int n_max, int NP,
               const Want& want, const Outs& o, bool packed_out, std::atomic<int>* ready) {
  const size_t P = pan.p + 1, PT = P * (P + 1) / 2;
  cudaStream_t ds = a.copy;
  const bool trace = getenv("PFML_HOST_TRACE") != nullptr;
  std::vector<cudaEvent_t> tev;
  auto mark = [&](cudaStream_t st) { if (trace) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); tev.push_back(e); } };
  mark(a.compute[0]);
  const int n_chunks = (int)sched.size();
  double* pin = (double*)a.pinned;
  int b0 = 0;
  for (int c = 0; c < n_chunks; b0 += sched[c], c++) {
    const int nb = sched[c], sb = c % n_cs;
    cudaStream_t ks = a.compute[sb];
    PFML_TRY(pfml_date_steps(&pan, nb, n_max, h.date_t + b0, h.sel + (size_t)b0 * NP, h.n_sel + b0, h.rt + (size_t)b0 * P,
                             h.den + (size_t)b0 * P * P, h.risk ? h.risk + (size_t)b0 * P * P : nullptr,
                             h.tc ? h.tc + (size_t)b0 * P * P : nullptr, h.sig ? h.sig + (size_t)b0 * NP * P : nullptr, h.status + b0,
                             h.ws[sb], h.ws_bytes, ks));
    if (want.n_ut) {
      // triangles that travel: [risk, tc] when both squares are wanted (denom = risk + tc is rebuilt), else [denom (, risk | tc)]
      const double* srcs[3] = {nullptr, nullptr, nullptr};
      int k = 0;
      if (want.risk && want.tc) { srcs[k++] = h.risk; srcs[k++] = h.tc; }
      else { srcs[k++] = h.den; if (want.risk) srcs[k++] = h.risk; if (want.tc) srcs[k++] = h.tc; }
      for (int q = 0; q < k; q++) {
        pack_upper_kernel<<<dim3(ceil_div((long)P, 256), (unsigned)P, nb), 256, 0, ks>>>((int)P, srcs[q] + (size_t)b0 * P * P,
                                                                                        h.ut + ((size_t)b0 * k + q) * PT, (long)k * PT);
        PFML_CHECK_LAUNCH();
      }
    }
    PFML_CUDA_OK(cudaEventRecord(a.computed[sb], ks));
    mark(ks);
    PFML_CUDA_OK(cudaStreamWaitEvent(ds, a.computed[sb], 0));
#define D2H(dst, src, count) PFML_CUDA_OK(cudaMemcpyAsync(dst, src, (size_t)(count) * sizeof(double), cudaMemcpyDeviceToHost, ds))
    D2H(o.r_tilde + (size_t)b0 * P, h.rt + (size_t)b0 * P, (size_t)nb * P);
    if (packed_out) {
      // date-major device layout [b][risk | tc] -> the caller's two arrays
      PFML_CUDA_OK(cudaMemcpy2DAsync(o.risk_ut + (size_t)b0 * PT, PT * 8, h.ut + (size_t)b0 * 2 * PT, 2 * PT * 8, PT * 8, nb,
                                     cudaMemcpyDeviceToHost, ds));
      PFML_CUDA_OK(cudaMemcpy2DAsync(o.tc_ut + (size_t)b0 * PT, PT * 8, h.ut + (size_t)b0 * 2 * PT + PT, 2 * PT * 8, PT * 8, nb,
                                     cudaMemcpyDeviceToHost, ds));
    } else if (want.n_ut) {
      D2H(pin + (size_t)b0 * want.n_ut * PT, h.ut + (size_t)b0 * want.n_ut * PT, (size_t)nb * want.n_ut * PT);
    } else {
      D2H(o.denom + (size_t)b0 * P * P, h.den + (size_t)b0 * P * P, (size_t)nb * P * P);
      if (o.risk) D2H(o.risk + (size_t)b0 * P * P, h.risk + (size_t)b0 * P * P, (size_t)nb * P * P);
      if (o.tc) D2H(o.tc + (size_t)b0 * P * P, h.tc + (size_t)b0 * P * P, (size_t)nb * P * P);
    }
    if (o.signal_t) D2H(o.signal_t + (size_t)b0 * NP * P, h.sig + (size_t)b0 * NP * P, (size_t)nb * NP * P);
#undef D2H
    PFML_CUDA_OK(cudaEventRecord(a.drained[c], ds));
    mark(ds);
  }
  // hand the drained chunks to the unpack threads as they arrive
  b0 = 0;
  for (int c = 0; c < n_chunks; b0 += sched[c], c++) {
    PFML_CUDA_OK(cudaEventSynchronize(a.drained[c]));
    if (ready) ready->store(b0 + sched[c], std::memory_order_release);
  }
  for (int i = 0; i < n_cs; i++) PFML_CUDA_OK(cudaStreamSynchronize(a.compute[i]));
  if (o.status) PFML_CUDA_OK(cudaMemcpyAsync(o.status, h.status, sizeof(int) * B, cudaMemcpyDeviceToHost, ds));
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

int host_impl(const char* fn, const pfml_host_block_t* blk, int B, int n_max, const int* date_t, const int* sel, const int* n_sel,
              const Outs& o, bool packed_out, int chunk, int device) {
  if (!blk || B <= 0 || n_max <= 0 || !date_t || !sel || !n_sel || !o.r_tilde || (packed_out ? (!o.risk_ut || !o.tc_ut) : !o.denom)) {
    set_error("%s: null/empty argument", fn);
    return -1;
  }
  if (device < 0 || device >= 64) {
    set_error("%s: bad device %d", fn, device);
    return -1;
  }
  const int NP = pfml_padded_n(n_max);
  PFML_TRY(validate(fn, blk, B, n_max, date_t, sel, n_sel, NP));
  Want want{};
  want.risk = packed_out || o.risk != nullptr;
  want.tc = packed_out || o.tc != nullptr;
  want.sig = o.signal_t != nullptr;
  // Full-square entry point: by default the squares themselves are downloaded.  PFML_HOST_REBUILD=1 sends the packed
  // triangles instead and rebuilds the squares with host threads -- a third of the PCIe bytes, but the transposing
  // rebuild is host-memory / TLB bound: on the 16-vCPU B200 box it measured 130 ms per 148 dates against 115 ms for the
  // plain download, so it is opt-in for hosts with more memory bandwidth per GPU.
  const char* rb = getenv("PFML_HOST_REBUILD");
  const bool rebuild_squares = rb && atoi(rb) != 0;
  want.n_ut = packed_out ? 2 : (!rebuild_squares ? 0 : ((want.risk && want.tc) ? 2 : 1 + (want.risk ? 1 : 0) + (want.tc ? 1 : 0)));

  // Chunks: uniform, default ceil(B / 6).  Consecutive chunks alternate between compute streams (each with its own
  // workspace): the latency-bound kernels of one chunk (block inverses, small GEMMs of the recursion) overlap the
  // tensor-bound kernels of the other, and the copy stream drains finished chunks meanwhile.
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
    // measured on B200 (tools/e2e_sched_sweep.py, tools/e2e_chunk_sweep.py): full squares (29 MB per date: the copy engine
    // must start early and stay busy) B / 6; r_tilde and denom only B / 3; packed triangles (12.6 MB per date: compute, not
    // PCIe, is the limit, so few large chunks, the last one smallest to keep the un-overlapped final download short):
    // 41 % / 34 % / 25 % of the dates
    const bool all_outputs = want.risk || want.tc || want.sig;
    if (chunk <= 0 && want.n_ut > 0 && B > 40) {
      const int c0 = (int)(0.405 * B + 0.5), c1 = (int)(0.34 * B + 0.5);
      sched = {c0, c1, B - c0 - c1};
    } else {
      if (chunk <= 0) chunk = B <= 40 ? B : (all_outputs ? (B + 5) / 6 : (B + 2) / 3);
      if ((B + chunk - 1) / chunk > MAX_CHUNKS) chunk = (B + MAX_CHUNKS - 1) / MAX_CHUNKS;
      for (int b0 = 0; b0 < B; b0 += chunk) sched.push_back((B - b0) < chunk ? (B - b0) : chunk);
    }
  }
  if ((int)sched.size() > MAX_CHUNKS) {
    set_error("%s: more than %d chunks", fn, MAX_CHUNKS);
    return -2;
  }
  chunk = 0;
  for (int c : sched) chunk = c > chunk ? c : chunk;      // buffers are sized for the largest chunk

  DeviceGuard guard;
  Arena& a = g_arena[device];
  std::lock_guard<std::mutex> lk(a.mu);                   // one call at a time per GPU; other GPUs are independent
  PFML_CUDA_OK(cudaSetDevice(device));
  if (!a.init) {
    PFML_CUDA_OK(cudaStreamCreateWithFlags(&a.copy, cudaStreamNonBlocking));
    PFML_CUDA_OK(cudaEventCreateWithFlags(&a.panel_ready, cudaEventDisableTiming));
    for (int i = 0; i < MAX_CS; i++) {
      PFML_CUDA_OK(cudaStreamCreateWithFlags(&a.compute[i], cudaStreamNonBlocking));
      PFML_CUDA_OK(cudaEventCreateWithFlags(&a.computed[i], cudaEventDisableTiming));
    }
    for (int i = 0; i < MAX_CHUNKS; i++) PFML_CUDA_OK(cudaEventCreateWithFlags(&a.drained[i], cudaEventDisableTiming));
    a.init = true;
  }
  static const int env_cs = getenv("PFML_HOST_STREAMS") ? atoi(getenv("PFML_HOST_STREAMS")) : 2;   // tuning aid
  const int n_cs = env_cs < 1 ? 1 : (env_cs > MAX_CS ? MAX_CS : env_cs);
  const size_t TN = (size_t)blk->T * blk->N, P = blk->p + 1, PT = P * (P + 1) / 2;
  pfml_panel_t pan{};
  pan.T = blk->T; pan.N = blk->N; pan.F = blk->F; pan.p = blk->p; pan.K = blk->K;
  pan.gamma_rel = blk->gamma_rel; pan.mu = blk->mu; pan.lb = blk->lb; pan.iterations = blk->iterations;
  const size_t ws_bytes = pfml_date_steps_workspace_bytes(&pan, chunk, n_max);
  HostPlan h = plan(nullptr, blk, B, NP, want, ws_bytes, n_cs);
  if (h.total > a.cap) {
    if (a.base) PFML_CUDA_OK(cudaFree(a.base));
    a.base = nullptr; a.cap = 0;
    PFML_CUDA_OK(cudaMalloc(&a.base, h.total));
    a.cap = h.total;
  }
  const bool rebuild = !packed_out && want.n_ut > 0;
  if (rebuild) {
    const size_t need = (size_t)B * want.n_ut * PT * sizeof(double);
    if (need > a.pinned_cap) {
      if (a.pinned) PFML_CUDA_OK(cudaFreeHost(a.pinned));
      a.pinned = nullptr; a.pinned_cap = 0;
      PFML_CUDA_OK(cudaHostAlloc(&a.pinned, need, cudaHostAllocDefault));
      a.pinned_cap = need;
    }
  }
  h = plan(a.base, blk, B, NP, want, ws_bytes, n_cs);
  cudaStream_t cs = a.compute[0];

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
  int rc = pfml_rff_sincos(h.feats, h.W, (long)TN, blk->F, blk->p / 2, h.rff, cs);
  if (rc == 0) rc = pfml_vol_scale(h.fct_load, h.fct_cov, h.ivol, blk->barra_has ? h.has : nullptr, blk->T, blk->N, blk->K, h.vol_scale, cs);
  if (rc == 0) rc = pfml_growth_ratio(h.tr, h.mu, (long)TN, h.gt, cs);
  pan.rff = h.rff; pan.vol_scale = h.vol_scale; pan.gt = h.gt; pan.lam = h.lam; pan.ret_ld1 = h.ret;
  pan.fct_load = h.fct_load; pan.fct_cov = h.fct_cov; pan.ivol = h.ivol; pan.wealth = h.wealth; pan.rf = h.rf;
  if (rc == 0 && cudaEventRecord(a.panel_ready, cs) != cudaSuccess) { set_error("%s: cudaEventRecord failed", fn); rc = -100; }
  for (int i = 1; i < n_cs && rc == 0; i++)
    if (cudaStreamWaitEvent(a.compute[i], a.panel_ready, 0) != cudaSuccess) { set_error("%s: cudaStreamWaitEvent failed", fn); rc = -100; }

  // ---- host threads that rebuild the full squares from the drained triangles ------------------------
  std::atomic<int> ready{0}, next{0};
  std::atomic<bool> abort_flag{false};
  std::vector<std::thread> pool;
  if (rc == 0 && rebuild) {
    const int nthreads = unpack_threads();
    const double* pin = (const double*)a.pinned;
    const int k = want.n_ut;
    const bool both = want.risk && want.tc;
    for (int t = 0; t < nthreads; t++)
      pool.emplace_back([&, pin, k, both] {
        for (;;) {
          const int b = next.fetch_add(1, std::memory_order_relaxed);
          if (b >= B) return;
          while (ready.load(std::memory_order_acquire) <= b) {
            if (abort_flag.load(std::memory_order_relaxed)) return;
            std::this_thread::yield();
          }
          const double* src = pin + (size_t)b * k * PT;
          const size_t sq = (size_t)b * P * P;
          if (both) {
            unpack_date((int)P, src, src + PT, o.risk + sq, o.tc + sq, o.denom + sq);
          } else {
            unpack_date((int)P, src, nullptr, o.denom + sq, nullptr, nullptr);
            if (want.risk) unpack_date((int)P, src + PT, nullptr, o.risk + sq, nullptr, nullptr);
            if (want.tc) unpack_date((int)P, src + PT, nullptr, o.tc + sq, nullptr, nullptr);
          }
        }
      });
  }
  // ---- chunks: compute on alternating streams, drain on the copy stream ------------------------------
  if (rc == 0) rc = run_chunks(a, h, pan, sched, n_cs, B, n_max, NP, want, o, packed_out, rebuild ? &ready : nullptr);
  if (rc != 0) {
    // nothing that still uses the arena may stay in flight once the lock is released
    abort_flag.store(true);
    for (int i = 0; i < n_cs; i++) cudaStreamSynchronize(a.compute[i]);
    cudaStreamSynchronize(a.copy);
  }
  for (auto& th : pool) th.join();
  return rc;
}

}  // namespace

extern "C" int pfml_date_steps_host(const pfml_host_block_t* blk, int B, int n_max, const int* date_t, const int* sel,
                                    const int* n_sel, double* r_tilde, double* denom, double* risk, double* tc,
                                    double* signal_t, int* status, int chunk, int device) {
  Outs o{r_tilde, denom, risk, tc, signal_t, nullptr, nullptr, status};
  return host_impl("pfml_date_steps_host", blk, B, n_max, date_t, sel, n_sel, o, false, chunk, device);
}

extern "C" int pfml_date_steps_host_packed(const pfml_host_block_t* blk, int B, int n_max, const int* date_t, const int* sel,
                                           const int* n_sel, double* r_tilde, double* risk_ut, double* tc_ut, double* signal_t,
                                           int* status, int chunk, int device) {
  Outs o{r_tilde, nullptr, nullptr, nullptr, signal_t, risk_ut, tc_ut, status};
  return host_impl("pfml_date_steps_host_packed", blk, B, n_max, date_t, sel, n_sel, o, true, chunk, device);
}

// Host helpers for the packed form (no device work): full P x P square(s) from upper triangle(s).
extern "C" int pfml_unpack_upper(const double* ut, int P, double* full) {
  if (!ut || !full || P <= 0) {
    set_error("pfml_unpack_upper: null/empty argument");
    return -1;
  }
  unpack_date(P, ut, nullptr, full, nullptr, nullptr);
  return 0;
}

extern "C" int pfml_unpack_reals(const double* risk_ut, const double* tc_ut, int P, double* risk, double* tc, double* denom) {
  if (!risk_ut || !tc_ut || !risk || !tc || !denom || P <= 0) {
    set_error("pfml_unpack_reals: null/empty argument");
    return -1;
  }
  unpack_date(P, risk_ut, tc_ut, risk, tc, denom);
  return 0;
}
