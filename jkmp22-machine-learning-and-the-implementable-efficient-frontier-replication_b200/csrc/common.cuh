// Shared declarations for the PFML sm_100a kernels (internal; the public C ABI is include/pfml_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <mutex>

namespace pfml {

// ---- error plumbing: the C ABI never throws; the first failure is latched per thread -------------
void set_error(const char* fmt, ...);
const char* last_error();

#define PFML_CUDA_OK(expr)                                                                   \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      pfml::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return -100;                                                                           \
    }                                                                                        \
  } while (0)

extern std::atomic<long> g_launch_count;   // kernels launched by this library (bench.py's gpu_launches)

#define PFML_CHECK_LAUNCH()                                                                  \
  do {                                                                                       \
    pfml::g_launch_count.fetch_add(1, std::memory_order_relaxed);                            \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess) {                                                                 \
      pfml::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return -101;                                                                           \
    }                                                                                        \
  } while (0)

// One-time (per device) kernel attribute.  cudaFuncSetAttribute is idempotent, so two threads racing through the first
// call both set the same value; the flag itself is atomic.
#define PFML_CONFIG_ONCE(kern, attr, value)                                                  \
  do {                                                                                       \
    static std::atomic<bool> _cfg[64];                                                       \
    int _dev = 0;                                                                            \
    PFML_CUDA_OK(cudaGetDevice(&_dev));                                                      \
    if (!_cfg[_dev & 63].load(std::memory_order_acquire)) {                                  \
      PFML_CUDA_OK(cudaFuncSetAttribute(kern, attr, value));                                 \
      _cfg[_dev & 63].store(true, std::memory_order_release);                                \
    }                                                                                        \
  } while (0)

#define PFML_TRY(expr)                                                                       \
  do {                                                                                       \
    int _rc = (expr);                                                                        \
    if (_rc != 0) return _rc;                                                                \
  } while (0)

// ---- batched FP64 GEMM on the DMMA pipe -----------------------------------------------------------
//   D[z] = alpha * az[z] * rs[r] * cs[c] * (opA[z] x opB[z]) + beta * bz[z] * E[z] + diag * I
//   opA[m][k] = a_kinner ? A[m*lda + k] : A[k*lda + m];  opB[k][n] = b_kinner ? B[n*ldb + k] : B[k*ldb + n]
//   sym != 0: M == N, only tiles on/below the diagonal are computed and the strict upper triangle is
//   written as the mirror image of the lower one (exactly symmetric output).
//   D2 (optional): D2 = D + E2 (same leading dimension / batch stride as D).
struct GemmP {
  const double* A = nullptr; long lda = 0; long sA = 0;
  const double* B = nullptr; long ldb = 0; long sB = 0;
  double* D = nullptr;       long ldd = 0; long sD = 0;
  const double* E = nullptr; long lde = 0; long sE = 0;
  double* D2 = nullptr; const double* E2 = nullptr;
  double* DT = nullptr; long ldt = 0;   // optional transposed copy DT[c][r] = D[r][c] (batch stride sD)
  int M = 0, N = 0, K = 0;
  bool a_kinner = true, b_kinner = false;
  double alpha = 1.0, beta = 0.0, diag = 0.0;
  const double* az = nullptr;  // per-batch multiplier of alpha (device, stride 1)
  const double* bz = nullptr;  // per-batch multiplier of beta
  long s_z = 1;                // element stride of az / bz between batches
  const double* rs = nullptr; long s_rs = 0;  // row scale (length M) per batch
  const double* cs = nullptr; long s_cs = 0;  // column scale (length N) per batch
  int sym = 0;
  int sincos = 0;              // epilogue: D[r][c] = cos(v), D[r][c + N] = sin(v)  (random Fourier features)
};

int gemm(const GemmP& p, int batch, cudaStream_t stream);
double gemm_flops(const GemmP& p, int batch);

// Error words of the INT8 path (`err`, device int array of EMU_ERR_SLOT0 + batch entries, zeroed by the caller):
//   err[0]                  bits 0-3: a barrier wait inside emu_gemm_kernel timed out (kernel-wide, not attributable to a matrix)
//   err[EMU_ERR_SLOT0 + z]  bit 4: matrix z of the batch held NaN / Inf in an operand (it cannot be sliced)
constexpr int EMU_ERR_SLOT0 = 64;

// ---- FP64 GEMM emulated on the INT8 tensor cores (emu.cu): tcgen05.mma kind::i8 + TMEM + bulk-TMA staging ----------
size_t emu_a_slices_bytes(int batch, int M, int K);
size_t emu_b_slices_bytes(int batch, int K, int N);
bool emu_shape_ok(int M, int N, int K);
int emu_slice_a(const double* A, long lda, long sA, int M, int K, int batch, int8_t* out, double* scale, int* err, cudaStream_t s);
int emu_slice_b(const double* B, long ldb, long sB, int K, int N, int batch, int8_t* out, double* scale, int* ex, int* err,
                cudaStream_t s);
int emu_mma(const int8_t* As, const double* sa, const int8_t* Bs, const double* sb, int M, int N, int K, int batch, double alpha,
            const double* rs, long s_rs, double beta, const double* E, long lde, long sE, double* D, long ldd, long sD, int* err,
            cudaStream_t s);
// optional epilogue extras of emu_mma: o = cs[col] * o (+ diag on the diagonal), and a second output D2 = E2 + o
struct EmuExtra {
  const double* cs = nullptr; long s_cs = 0;
  double diag = 0.0;
  double* D2 = nullptr; const double* E2 = nullptr; long ld2 = 0, s2 = 0;
  const double* az = nullptr; long s_az = 0;   // per-matrix factor on alpha
  bool sym = false;                            // symmetric result: only the upper triangle is written, the caller mirrors it
  int digits = 0;                              // leading base-256 digits of both operands to use (0 = all 7; 6 = 21 products)
};
int emu_slice_nmajor512(const double* X, long ld, long sX, int K, int C, int batch, int8_t* outB, double* sb, int* ex, int8_t* outA,
                        double* sa, int* err, cudaStream_t s, const double* row_scale = nullptr, long s_row_scale = 0,
                        int* zflag = nullptr /* [batch]: non-finite operand flagged per matrix instead of through err */);
int emu_mma_ex(const int8_t* As, const double* sa, const int8_t* Bs, const double* sb, int M, int N, int K, int batch, double alpha,
               const double* rs, long s_rs, double beta, const double* E, long lde, long sE, double* D, long ldd, long sD,
               const EmuExtra& x, int* err, cudaStream_t s);

// profiling categories (device-time accounting with CUDA events, see prof.cu)
enum ProfTag { T_OTHER = 0, T_STD = 1, T_MFUNC = 2, T_CHAIN = 3, T_HORNER = 4, T_SOLVE = 5, T_DENOM = 6,
               T_RFF = 7, T_RIDGE = 8, T_COV = 9, T_NTAGS = 10 };
void prof_begin(int tag, cudaStream_t s);
void prof_end(int tag, cudaStream_t s, double flops, double bytes);

struct ProfScope {
  int tag; cudaStream_t s; double fl, by;
  ProfScope(int t, cudaStream_t st, double flops = 0, double bytes = 0) : tag(t), s(st), fl(flops), by(bytes) { prof_begin(tag, s); }
  ~ProfScope() { prof_end(tag, s, fl, by); }
};

static inline int ceil_div(long a, long b) { return (int)((a + b - 1) / b); }
static inline long round_up(long a, long b) { return (a + b - 1) / b * b; }

}  // namespace pfml
