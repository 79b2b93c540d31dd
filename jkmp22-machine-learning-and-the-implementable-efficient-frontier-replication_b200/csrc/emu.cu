// FP64 GEMM emulated on the INT8 tensor cores of sm_100a: tcgen05.mma kind::i8, accumulators in TMEM, operands
// staged by bulk TMA copies (cp.async.bulk + mbarrier), Ozaki-style error-free slicing.
//
//   a_ik = 2^(ea_i + 2) sum_{t<S} da_t 256^-(t+1),   b_kj = 2^(eb_j + 2) sum_{t<S} db_t 256^-(t+1),   d in [-128, 127]
//   (balanced base-256 digits of round(x 2^(8 S - 2 - e)); S = 7 slices carry 54 bits)
//   C_ij = 2^(ea_i + eb_j - 12) sum_{d<S} 256^-d X_d,   X_d = sum_{i+j=d} da_i . db_j   (EXACT int32: K 128^2 S < 2^31)
//
// Only the slice pairs with i + j < S are multiplied (28 products for S = 7); the dropped ones are below
// 2^-58 of |a|_rowmax |b|_colmax per term and carry no systematic sign, so the result is as accurate as a native FP64
// GEMM (measured against the DMMA kernel in tests/test_emulated_gemm_gpu.py).  Why it pays: the DMMA pipe peaks at
// 37 TFLOP/s, the INT8 pipe at 4.5 POP/s; 28 INT8 products per FP64 product still leave > 60 FP64-equivalent TFLOP/s.
//
// Layout: slices are written to global memory already in the shared-memory order the UMMA descriptors expect
// (K-major, no swizzle, 8 rows x 16 B core matrices), one contiguous block per (row block, 64-byte k chunk, all
// slices), so a pipeline stage is ONE bulk copy per operand.  One CTA per SM walks the output tiles (128 x 64);
// the S accumulators X_d occupy S x 64 of the 512 TMEM columns.  Warp 0: TMA producer; warp 1: MMA issuer; warp 2:
// TMEM allocator; warps 4-19: epilogue (TMEM -> registers, exact int64 recombination, scaling through the exponent
// field, fused row / column scale, addend, identity, second output; stores transposed through shared memory).
#include <cstdlib>

#include "kernels.cuh"

namespace pfml {

namespace {

#ifndef PFML_EMU_SLICES
#define PFML_EMU_SLICES 7
#endif
constexpr int ES = PFML_EMU_SLICES;                          // slices (8: 56 bits; 7: 49 bits)
constexpr int EBM = 128, EBN = 64, EKC = 64;                 // CTA tile, k chunk (bytes) per stage
constexpr int ELBO = 128, ESBO = (EKC / 16) * 128;           // canonical layout strides (bytes)
constexpr int EA_STAGE = ES * EBM * EKC, EB_STAGE = ES * EBN * EKC;   // 64 KB, 32 KB
constexpr int ESTAGES = 2;
constexpr double kBScale = 1.0 / 4096.0;                     // a b = 2^(ea + eb - 12) sum_d acc_d 2^(-8 d): folded into the B scales

__host__ __device__ inline int canon(int r, int kk) { return (r % 8) * 16 + (r / 8) * ESBO + (kk / 16) * ELBO + kk % 16; }

// 16 values -> ES balanced base-256 digits each: F = round(x 2^(8 ES - 2 - e)) as a signed integer (|x| < 2^e, so
// |F| <= 2^(8 ES - 2)), F = sum_t d_t 256^(ES-1-t) with d_t in [-128, 127] taken from the bottom (the top digit stays
// within +-65).  Every slice carries 8 bits, 8 ES - 2 bits in all (54 for ES = 7: the whole FP64 significand of the
// largest element of the row / column); the digit signs are not tied to the sign of x, so the products left out
// (i + j >= ES) have no systematic sign.  Zeros, denormals and values below 2^(e - 8 ES + 1) give all-zero slices.
__device__ __forceinline__ void slice16(const double (&x)[16], int e, int8_t* dst, long slice_stride, int8_t* dst2 = nullptr,
                                        long slice_stride2 = 0) {
  // With the bias Bz = sum_t 0x80 256^t: the bytes g_t of F + Bz satisfy F = sum_t (g_t - 128) 256^t, and g_t - 128 as an
  // int8 is g_t ^ 0x80 -- all ES balanced digits of an element come from one 64-bit add and one xor.
  constexpr unsigned long long Bz = 0x8080808080808080ull >> (8 * (8 - ES));
  unsigned lo[16], hi[16];
  // F = round(x 2^(8 ES - 2 - e)): the scale is an exact power of two built in the exponent field, the product is exact and
  // the hardware conversion rounds to nearest (ties to even).  A scale outside the normal range (a row / column whose
  // largest element is below 2^-968, or a bogus exponent) gives all-zero digits; NaN / Inf have no digits either (they
  // are flagged per matrix by the callers).
  const int se = 1023 + (8 * ES - 2) - e;
  const double sc = (se >= 1 && se <= 2046) ? __hiloint2double(se << 20, 0) : 0.0;
#pragma unroll
  for (int k = 0; k < 16; k++) {
    const double xs = x[k] * sc;
    const long long fv = (fabs(xs) <= 9007199254740992.0 * 4.0) ? __double2ll_rn(xs) : 0ll;     // |F| <= 2^(8 ES - 2) for finite in-range values
    const unsigned long long g = ((unsigned long long)fv + Bz) ^ Bz;
    lo[k] = (unsigned)g; hi[k] = (unsigned)(g >> 32);
  }
#pragma unroll
  for (int t = 0; t < ES; t++) {
    const int p = ES - 1 - t;                              // byte position of digit t (most significant first)
    int4 v;
    unsigned w[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const unsigned a0 = p < 4 ? lo[4 * q] : hi[4 * q], a1 = p < 4 ? lo[4 * q + 1] : hi[4 * q + 1];
      const unsigned a2 = p < 4 ? lo[4 * q + 2] : hi[4 * q + 2], a3 = p < 4 ? lo[4 * q + 3] : hi[4 * q + 3];
      const unsigned sel = (unsigned)(p & 3) | ((unsigned)((p & 3) + 4) << 4);      // byte p of a, byte p of b
      const unsigned x01 = __byte_perm(a0, a1, sel), x23 = __byte_perm(a2, a3, sel);
      w[q] = __byte_perm(x01, x23, 0x5410);
    }
    v.x = (int)w[0]; v.y = (int)w[1]; v.z = (int)w[2]; v.w = (int)w[3];
    if (dst) *reinterpret_cast<int4*>(dst + t * slice_stride) = v;
    if (dst2) *reinterpret_cast<int4*>(dst2 + t * slice_stride2) = v;
  }
}

// A-type operand: X [R][K] with K contiguous (ld), scaled per row, blocks of BR rows (128 for A, 64 for B^T operands).
// Non-finite or zero rows get exponent 0.  Rows >= R inside the last block are zero-filled.
template <int BR>
__global__ void slice_kmajor_kernel(const double* __restrict__ X, long ld, long sX, int R, int K, int8_t* __restrict__ out,
                                    double* __restrict__ scale, double extra, int* __restrict__ err) {
  const int z = blockIdx.y, row = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32, lane = threadIdx.x & 31;
  const int RB = (R + BR - 1) / BR, KB = K / EKC;
  if (row >= RB * BR) return;
  constexpr long STAGE = (long)ES * BR * EKC;
  int8_t* base = out + (((long)z * RB + row / BR) * KB) * STAGE;
  const double* a = X + (long)z * sX + (long)row * ld;
  int e = 0;
  if (row < R) {
    double mx = 0.0;
    bool bad = false;
    for (int k = lane; k < K; k += 32) { const double v = fabs(a[k]); bad |= !(v <= 1.7976931348623157e308); mx = fmax(mx, v); }
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (bad) atomicOr(err + EMU_ERR_SLOT0 + z, 16);   // NaN / Inf cannot be sliced: reported per matrix through the status array
    e = (mx > 0.0 && isfinite(mx)) ? ilogb(mx) + 1 : 0;
    if (lane == 0) scale[(long)z * R + row] = scalbn(extra, e);
  }
  for (int k0 = lane * 16; k0 < K; k0 += 32 * 16) {
    double x[16];
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = row < R ? a[k0 + k] : 0.0;
    slice16(x, e, base + (long)(k0 / EKC) * STAGE + canon(row % BR, k0 % EKC), (long)BR * EKC);
  }
}

// The same for K <= 1024 (K % 64 == 0) in one pass over the data.  A CTA owns 8 rows: warp w takes the 64-byte k chunks w
// and w + 8 of all of them, lane = (row r, 16-element group c), so that one store instruction of a warp writes four
// whole 8 x 16 B core matrices (512 contiguous bytes) of the operand layout; the row maxima meet in shared memory.
constexpr int ONEPASS_KMAX = 1024;
template <int BR>
__global__ void __launch_bounds__(256)
slice_kmajor512_kernel(const double* __restrict__ X, long ld, long sX, int R, int K, int8_t* __restrict__ out,
                       double* __restrict__ scale, double extra, int* __restrict__ err) {
  static_assert(EKC == 64, "one warp per 64-byte k chunk");
  __shared__ double s_mx[8][8];
  const int z = blockIdx.y, w = threadIdx.x >> 5, lane = threadIdx.x & 31, r = lane & 7, c = lane >> 3;
  const int row = blockIdx.x * 8 + r, KB = K / EKC;
  const int RB = (R + BR - 1) / BR;
  constexpr long STAGE = (long)ES * BR * EKC;
  double x[2][16];
  double mx = 0.0;
  bool bad = false;
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int ch = w + 8 * h;
    if (row < R && ch < KB) {
      const double2* a = reinterpret_cast<const double2*>(X + (long)z * sX + (long)row * ld + ch * EKC + c * 16);
#pragma unroll
      for (int k = 0; k < 8; k++) { const double2 v = a[k]; x[h][2 * k] = v.x; x[h][2 * k + 1] = v.y; }
#pragma unroll
      for (int k = 0; k < 16; k++) { const double v = fabs(x[h][k]); bad |= !(v <= 1.7976931348623157e308); mx = fmax(mx, v); }
    } else {
#pragma unroll
      for (int k = 0; k < 16; k++) x[h][k] = 0.0;
    }
  }
  if (bad) atomicOr(err + EMU_ERR_SLOT0 + z, 16);
  mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
  mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
  if (c == 0) s_mx[w][r] = mx;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < 8; q++) mx = fmax(mx, s_mx[q][r]);
  const int e = (mx > 0.0 && isfinite(mx)) ? ilogb(mx) + 1 : 0;
  if (w == 0 && c == 0 && row < R) scale[(long)z * R + row] = scalbn(extra, e);
  if (row >= RB * BR) return;
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int ch = w + 8 * h;
    if (ch < KB) slice16(x[h], e, out + (((long)z * RB + row / BR) * KB + ch) * STAGE + canon(row % BR, c * 16), (long)BR * EKC);
  }
}

// B-type operand: X [K][C] with C contiguous (ld), scaled per column, blocks of 64 columns.
__global__ void col_exponent_kernel(const double* __restrict__ X, long ld, long sX, int K, int C, int* __restrict__ ex,
                                    double* __restrict__ scale, double extra, int* __restrict__ err) {
  const int z = blockIdx.y, n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= C) return;
  const double* b = X + (long)z * sX + n;
  double mx = 0.0;
  bool bad = false;
  for (int k = 0; k < K; k++) { const double v = fabs(b[(long)k * ld]); bad |= !(v <= 1.7976931348623157e308); mx = fmax(mx, v); }
  if (bad) atomicOr(err + EMU_ERR_SLOT0 + z, 16);
  const int e = (mx > 0.0 && isfinite(mx)) ? ilogb(mx) + 1 : 0;
  ex[(long)z * C + n] = e;
  scale[(long)z * C + n] = scalbn(extra, e);
}

__global__ void __launch_bounds__(256)
slice_nmajor_kernel(const double* __restrict__ X, long ld, long sX, int K, int C, int CB, const int* __restrict__ ex,
                    int8_t* __restrict__ out) {
  // one CTA per (column block of 64, k chunk of 64): thread = (column, 16 consecutive k); reads coalesced along columns
  const int z = blockIdx.z, cb = blockIdx.x, kb = blockIdx.y, KB = K / EKC;
  const int nn = threadIdx.x & 63, kq = threadIdx.x >> 6, n = cb * EBN + nn, k0 = kb * EKC + kq * 16;
  int8_t* base = out + (((long)z * CB + cb) * KB + kb) * (long)EB_STAGE;
  double x[16];
  const int e = n < C ? ex[(long)z * C + n] : 0;
#pragma unroll
  for (int k = 0; k < 16; k++) x[k] = n < C ? X[(long)z * sX + (long)(k0 + k) * ld + n] : 0.0;
  slice16(x, e, base + canon(nn, kq * 16), (long)EBN * EKC);
}

// N-major operand X [512][C] in one pass: a CTA owns 32 columns x all 512 k (thread = column, 32 consecutive k held in
// registers), the column maxima go through shared memory, no separate exponent kernel.  Writes the B-type slices
// (64-column blocks, `out`) and / or the A-type slices of X' (128-row blocks, `outA`: X' as the left operand).
constexpr int SNC = 16;   // columns per CTA of the one-pass N-major kernels: 256 threads at K = 512, two CTAs per SM overlap their load and store phases
__global__ void __launch_bounds__(SNC * 32)
slice_nmajor512_kernel(const double* __restrict__ X, long ld, long sX, int K, int C, int CB, int* __restrict__ ex, double* __restrict__ scale,
                       double extra, int8_t* __restrict__ out, int RBA, double* __restrict__ scaleA, int8_t* __restrict__ outA,
                       int* __restrict__ err, const double* __restrict__ rsc, long s_rsc, int* __restrict__ zflag) {
  const int KB = K / EKC, NKQ = K / 32;                  // blockDim.x = SNC * NKQ (K <= 1024: up to 512 threads)
  __shared__ double red[32][SNC + 1];
  const int z = blockIdx.y, lane = threadIdx.x % SNC, kq = threadIdx.x / SNC;
  const int n = blockIdx.x * SNC + lane, cb = n / EBN, nn = n % EBN;
  double x[32];
  double mx = 0.0;
  bool bad = false;
  if (n < C) {
    const double* b = X + (long)z * sX + (long)(kq * 32) * ld + n;
#pragma unroll
    for (int j = 0; j < 32; j++) x[j] = b[(long)j * ld];
    if (rsc) {                                             // optional row scale (the k index of the product) fused into the slicing
      const double* r = rsc + (long)z * s_rsc + kq * 32;
#pragma unroll
      for (int j = 0; j < 32; j++) x[j] *= r[j];
    }
#pragma unroll
    for (int j = 0; j < 32; j++) { const double v = fabs(x[j]); bad |= !(v <= 1.7976931348623157e308); mx = fmax(mx, v); }
  } else {
#pragma unroll
    for (int j = 0; j < 32; j++) x[j] = 0.0;
  }
  // a non-finite operand is an error (status bit 4) unless the caller asked for a per-matrix flag instead (zflag: the
  // reference's own NaN semantics, quirk Q8, are then reproduced by the caller)
  if (bad) { if (zflag) atomicOr(zflag + z, 1); else atomicOr(err + EMU_ERR_SLOT0 + z, 16); }
  red[kq][lane] = mx;
  __syncthreads();
  for (int j = 0; j < NKQ; j++) mx = fmax(mx, red[j][lane]);
  const int e = (mx > 0.0 && isfinite(mx)) ? ilogb(mx) + 1 : 0;
  if (kq == 0 && n < C) {
    if (out) { ex[(long)z * C + n] = e; scale[(long)z * C + n] = scalbn(extra, e); }
    if (outA) scaleA[(long)z * C + n] = scalbn(1.0, e);
  }
  int8_t* base = (out && cb < CB) ? out + (((long)z * CB + cb) * KB + kq / 2) * (long)EB_STAGE : nullptr;
  int8_t* baseA = (outA && n / EBM < RBA) ? outA + (((long)z * RBA + n / EBM) * KB + kq / 2) * (long)EA_STAGE : nullptr;
#pragma unroll
  for (int h = 0; h < 2; h++) {
    double y[16];
#pragma unroll
    for (int j = 0; j < 16; j++) y[j] = x[h * 16 + j];
    const int kk = (kq & 1) * 32 + h * 16;
    slice16(y, e, base ? base + canon(nn, kk) : nullptr, (long)EBN * EKC, baseA ? baseA + canon(n % EBM, kk) : nullptr, (long)EBM * EKC);
  }
}

// Standardise + scale (the arithmetic of standardize_kernel in kernels.cu, PFML_Input_Data.py:357-388) for the slabs that
// are only ever B operands of the INT8 products: the standardised, tilde-scaled column goes straight into its
// base-256 digits in the B-type layout, no FP64 copy is written.  A CTA owns 64 output columns (one B-type block);
// thread = (column, row quarter): it owns the 16-row groups ib = rq, rq + 4, ... in all three passes (mean; centred sum
// of squares + largest scaled deviation; digits), partial sums meet in shared memory in a fixed order.
constexpr int SSQ = 4;   // row quarters
__global__ void __launch_bounds__(64 * SSQ)
standardize_slice_kernel(PanelView pv, int NP, int P, int tau0, const int* __restrict__ date_t, const int* __restrict__ sel,
                         const int* __restrict__ nsel, BatchVecs v, int8_t* __restrict__ out, long slab_bytes,
                         double* __restrict__ scale_out, int* __restrict__ err) {
  const int b = blockIdx.z, tau = tau0 + blockIdx.y;
  const int cj = threadIdx.x & 63, rq = threadIdx.x >> 6;
  const int j = blockIdx.x * EBN + cj;                   // output column (0 = constant)
  const int n = nsel[b], t = date_t[b] - tau;
  const int CB = (P + EBN - 1) / EBN, KB = NP / EKC, NG = NP / 16;
  // per-row factors once per CTA: gather index, 1 / vol_scale and sqrt(lambda)
  extern __shared__ double s_dyn[];
  double* s_inv = s_dyn;
  double* s_sl = s_dyn + NP;
  double* s_red = s_dyn + 2 * NP;                        // [SSQ][64]
  int* s_sel = reinterpret_cast<int*>(s_dyn + 2 * NP + SSQ * 64);
  const double* __restrict__ vs = pv.vol_scale + (long)t * pv.N;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int si = sel[(long)b * NP + i];
    s_sel[i] = si;
    s_inv[i] = 1.0 / vs[si];
    s_sl[i] = v.sl[(long)b * NP + i];
  }
  __syncthreads();
  const bool real = j < P;
  const double* __restrict__ X = pv.rff + (long)t * pv.N * pv.p + (j > 0 ? j - 1 : 0);
  double mean = 0.0, scale = 0.0, mx = 0.0;
  bool bad = false;
  // pass 1: mean
  double part = 0.0;
  if (real && j > 0)
    for (int ib = rq; ib < NG; ib += SSQ) {
      const int i1 = min(ib * 16 + 16, n);
      for (int i = ib * 16; i < i1; i++) part += X[(long)s_sel[i] * pv.p];
    }
  s_red[rq * 64 + cj] = part;
  __syncthreads();
  if (real && j > 0) {
    double sum = 0.0;
#pragma unroll
    for (int q = 0; q < SSQ; q++) sum += s_red[q * 64 + cj];
    mean = sum / (double)n;
  }
  __syncthreads();
  // pass 2: centred sum of squares and the largest |sqrt(lambda) / vol_scale * c|
  part = 0.0;
  if (real)
    for (int ib = rq; ib < NG; ib += SSQ) {
      const int i1 = min(ib * 16 + 16, n);
      for (int i = ib * 16; i < i1; i++) {
        const double c = (j == 0) ? 1.0 : X[(long)s_sel[i] * pv.p] - mean;
        part += c * c;
        const double a = fabs(s_sl[i] * (s_inv[i] * c));
        bad |= !(a <= 1.7976931348623157e308);
        mx = fmax(mx, a);
      }
    }
  s_red[rq * 64 + cj] = part;
  __syncthreads();
  double ss = 0.0;
#pragma unroll
  for (int q = 0; q < SSQ; q++) ss += s_red[q * 64 + cj];
  __syncthreads();
  s_red[rq * 64 + cj] = mx;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < SSQ; q++) mx = fmax(mx, s_red[q * 64 + cj]);
  if (real) {
    scale = (j == 0) ? sqrt(1.0 / (double)n) : sqrt(1.0 / ss);   // constant: not de-meaned, sum of squares = n
    mx *= scale;                                                  // largest |value| up to rounding (digits tolerate 1.98 * 2^e)
    bad |= !(mx <= 1.7976931348623157e308);
  }
  if (bad) atomicOr(err + EMU_ERR_SLOT0 + b, 16);          // NaN / Inf cannot be sliced: reported per date through the status array
  const int e = (mx > 0.0 && isfinite(mx)) ? ilogb(mx) + 1 : 0;
  if (real && rq == 0) scale_out[((long)blockIdx.y * gridDim.z + b) * P + j] = scalbn(kBScale, e);
  // pass 3: digits
  int8_t* base = out + (long)blockIdx.y * slab_bytes + (((long)b * CB + blockIdx.x) * KB) * (long)EB_STAGE;
  for (int ib = rq; ib < NG; ib += SSQ) {
    const int i0 = ib * 16;
    double x[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
      const int i = i0 + k;
      double tv = 0.0;
      if (real && i < n) {
        const double c = (j == 0) ? 1.0 : X[(long)s_sel[i] * pv.p] - mean;
        tv = s_sl[i] * (s_inv[i] * (c * scale));
      }
      x[k] = tv;
    }
    slice16(x, e, base + (long)(i0 / EKC) * EB_STAGE + canon(cj, i0 % EKC), (long)EBN * EKC);
  }
}

// The same for NP <= 1024 in ONE pass over the gathered panel rows: a CTA owns SNC output columns, thread = (column,
// group of 32 rows) keeps its values in registers from the statistics to the digits (like slice_nmajor512_kernel).
__global__ void __launch_bounds__(SNC * 32)
standardize_slice512_kernel(PanelView pv, int NP, int P, int tau0, const int* __restrict__ date_t, const int* __restrict__ sel,
                            const int* __restrict__ nsel, BatchVecs v, int8_t* __restrict__ out, long slab_bytes,
                            double* __restrict__ scale_out, int* __restrict__ err) {
  // blockIdx.z walks the diagonals b - tau = const: with consecutive month-ends the (date, lag) pairs of one diagonal
  // read the same panel month, so launching them together keeps that month in L2 (DRAM reads 4.9 GB -> ~1 GB)
  const int ntau = gridDim.y, nB = gridDim.z - ntau + 1;
  const int b = (int)blockIdx.z + (int)blockIdx.y - (ntau - 1), tau = tau0 + blockIdx.y;
  if (b < 0 || b >= nB) return;
  const int lane = threadIdx.x % SNC, rg = threadIdx.x / SNC;    // rows rg * 32 .. rg * 32 + 31
  const int j = blockIdx.x * SNC + lane;                           // output column (0 = constant)
  const int n = nsel[b], t = date_t[b] - tau;
  const int CB = (P + EBN - 1) / EBN, KB = NP / EKC;
  // row factor sqrt(lambda_i) / vol_scale_i once per CTA (the FP64 kernel applies the two factors one after the other:
  // one rounding more, far below the digit resolution); non-finite factors are caught here, non-finite signals by the
  // sum of squares below
  __shared__ double s_rf[ONEPASS_KMAX], s_red[32][SNC + 1];     // blockDim.x = SNC * (NP / 32)
  __shared__ int s_sel[ONEPASS_KMAX];
  const int NRG = NP / 32;
  const double* __restrict__ vs = pv.vol_scale + (long)t * pv.N;
  bool bad = false;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int si = sel[(long)b * NP + i];
    s_sel[i] = si;
    const double rf = v.sl[(long)b * NP + i] * (1.0 / vs[si]);
    s_rf[i] = rf;
    bad |= !(fabs(rf) <= 1.7976931348623157e308);
  }
  __syncthreads();
  const bool real = j < P;
  const double* __restrict__ X = pv.rff + (long)t * pv.N * pv.p + (j > 0 ? j - 1 : 0);
  const int i0 = rg * 32;
  double x[32];
#pragma unroll
  for (int k = 0; k < 32; k++) {
    const int i = i0 + k;
    x[k] = (real && i < n) ? (j == 0 ? 1.0 : X[(long)s_sel[i] * pv.p]) : 0.0;
  }
  // mean (the constant column is not de-meaned)
  double part = 0.0;
#pragma unroll
  for (int k = 0; k < 32; k++) part += x[k];
  s_red[rg][lane] = part;
  __syncthreads();
  double mean = 0.0;
  if (j > 0) {
    for (int q = 0; q < NRG; q++) mean += s_red[q][lane];
    mean = mean / (double)n;
  }
  __syncthreads();
  // centred values, sum of squares, largest |sqrt(lambda) / vol_scale * c|
  double mx = 0.0;
  part = 0.0;
#pragma unroll
  for (int k = 0; k < 32; k++) {
    const int i = i0 + k;
    const double c = (real && i < n) ? x[k] - mean : 0.0;
    part += c * c;
    const double y = (i < n ? s_rf[i] : 0.0) * c;          // scaled below by the column's 1 / norm
    x[k] = y;
    mx = fmax(mx, fabs(y));
  }
  s_red[rg][lane] = part;
  __syncthreads();
  double ss = 0.0;
  for (int q = 0; q < NRG; q++) ss += s_red[q][lane];
  __syncthreads();
  s_red[rg][lane] = mx;
  __syncthreads();
  for (int q = 0; q < NRG; q++) mx = fmax(mx, s_red[q][lane]);
  const double scale = real ? ((j == 0) ? sqrt(1.0 / (double)n) : sqrt(1.0 / ss)) : 0.0;
  mx *= scale;                                               // largest |value| up to rounding (digits tolerate 1.98 * 2^e)
  bad |= real && !(ss <= 1.7976931348623157e308 && mx <= 1.7976931348623157e308);
  if (bad) atomicOr(err + EMU_ERR_SLOT0 + b, 16);
  const int e = (mx > 0.0 && isfinite(mx)) ? ilogb(mx) + 1 : 0;
  if (real && rg == 0) scale_out[((long)blockIdx.y * nB + b) * P + j] = scalbn(kBScale, e);
  if (i0 >= NP || j / EBN >= CB) return;
  int8_t* base = out + (long)blockIdx.y * slab_bytes + (((long)b * CB + j / EBN) * KB + i0 / EKC) * (long)EB_STAGE;
#pragma unroll
  for (int h = 0; h < 2; h++) {
    double y[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
      y[k] = x[h * 16 + k] * scale;
    }
    slice16(y, e, base + canon(j % EBN, (i0 + h * 16) % EKC), (long)EBN * EKC);
  }
}

}  // namespace

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  // start address >> 4 | leading byte offset >> 4 << 16 | stride byte offset >> 4 << 32 | version 1 << 46; no swizzle
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((ELBO >> 4) & 0x3FFF) << 16) | ((uint64_t)((ESBO >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46);
}
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {     // bounded: a bug must not hang the GPU
  uint32_t done = 0;
  for (int it = 0; it < (1 << 26) && !done; it++)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done != 0;
}
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
               ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_i8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}\n"
               ::"r"(d_tmem), "r"(a_tmem), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void utccp_128x32(uint32_t d_tmem, uint64_t src_desc) {   // 128 rows x 32 bytes, smem -> TMEM
  asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(d_tmem), "l"(src_desc) : "memory");
}
__device__ __forceinline__ bool elect_one() {      // one lane of a converged warp; lets ptxas issue the tensor ops without a per-lane loop
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, %1;\n\t@px mov.s32 %0, 1;\n\t}\n" : "+r"(pred) : "r"(0xFFFFFFFFu));
  return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}


#ifndef PFML_EMU_TS_SLICES
#define PFML_EMU_TS_SLICES 3
#endif
constexpr int ETS = PFML_EMU_TS_SLICES;                      // leading A slices held in TMEM (TS-form products)
constexpr int EACOL = ES * EBN;                               // their TMEM columns (8 per slice), after the accumulators
static_assert(EACOL + ETS * 8 <= 512 && ETS <= ES, "TMEM: accumulators + A slices");

struct EmuP {
  const int8_t* As; const int8_t* Bs; const double* sa; const double* sb;   // slices and 2^e scales (sb includes 2^-12)
  double* D; long ldd, sD;
  const double* E; long lde, sE;
  const double* rs; long s_rs;
  double alpha, beta;
  int M, N, K, MB, NB, batch;
  const double* cs; long s_cs; double diag;
  const double* az; long s_az;              // per-matrix factor on alpha (batch stride s_az)
  int sym;                                  // symmetric result: only the upper triangle is produced (the caller mirrors it)
  double* D2; const double* E2; long ld2, s2;
  int q;                                   // fixed-point position of the recombined sum (see the epilogue)
  int* err;
};

constexpr int EPI_G = 4;                                      // column groups of the epilogue -> 4 * EPI_G epilogue warps
constexpr int EMU_THREADS = 128 + 128 * EPI_G;

// ND <= ES: number of leading digits of both operands that are used (ND (ND + 1) / 2 products).  The operand layout
// always holds ES digits per stage; a balanced-digit number cut after ND digits is that number rounded to ND digits.
template <int ND>
__global__ void __launch_bounds__(EMU_THREADS, 1) emu_gemm_kernel(const EmuP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + ESTAGES * EA_STAGE;
  __shared__ __align__(8) uint64_t full_bar[ESTAGES], empty_bar[ESTAGES], acc_full, acc_empty;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int KB = p.K / EKC;
  const long n_tiles = (long)p.batch * p.MB * p.NB;
  if (tid == 0) {
    for (int s = 0; s < ESTAGES; s++) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full_bar[s])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty_bar[s])));
    }
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&acc_full)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&acc_empty)), "r"(128 * EPI_G));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));   // ES * 64 <= 512
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tbase = tmem_base;

  if (warp == 0) {
    // ===== producer: one bulk copy per operand per stage; runs ahead into the next tile while the epilogue works =====
    long it = 0;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      const int nb = (int)(t % p.NB), mb = (int)((t / p.NB) % p.MB), z = (int)(t / ((long)p.NB * p.MB));
      if (p.sym && nb * EBN + EBN - 1 < mb * EBM) continue;          // tile strictly below the diagonal
      const int8_t* a_src = p.As + (((long)z * p.MB + mb) * KB) * (long)EA_STAGE;
      const int8_t* b_src = p.Bs + (((long)z * p.NB + nb) * KB) * (long)EB_STAGE;
      for (int kc = 0; kc < KB; kc++, it++) {
        const int s = (int)(it % ESTAGES);
        if (!mbar_wait(smem_u32(&empty_bar[s]), (uint32_t)(((it / ESTAGES) & 1) ^ 1))) { atomicOr(p.err, 1); return; }
        if (elect_one()) {
          const uint32_t fb = smem_u32(&full_bar[s]);
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(ND * (EBM + EBN) * EKC) : "memory");
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(smem_u32(sA + s * EA_STAGE)), "l"(a_src + (long)kc * EA_STAGE), "r"(ND * EBM * EKC), "r"(fb) : "memory");
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(smem_u32(sB + s * EB_STAGE)), "l"(b_src + (long)kc * EB_STAGE), "r"(ND * EBN * EKC), "r"(fb) : "memory");
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: ES (ES + 1) / 2 digit pairs x 2 k-slabs per stage, accumulator d = i + j at TMEM column d * 64 =====
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(EBN >> 3) << 17) | ((uint32_t)(EBM >> 4) << 24);
    long it = 0;
    int tl = 0;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      if (p.sym && (int)(t % p.NB) * EBN + EBN - 1 < (int)((t / p.NB) % p.MB) * EBM) continue;
      if (!mbar_wait(smem_u32(&acc_empty), (uint32_t)((tl & 1) ^ 1))) { atomicOr(p.err, 4); return; }   // epilogue drained TMEM
      asm volatile("tcgen05.fence::after_thread_sync;");
      for (int kc = 0; kc < KB; kc++, it++) {
        const int s = (int)(it % ESTAGES);
        if (!mbar_wait(smem_u32(&full_bar[s]), (uint32_t)((it / ESTAGES) & 1))) { atomicOr(p.err, 2); return; }
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint64_t da0 = make_desc(smem_u32(sA + s * EA_STAGE)), db0 = make_desc(smem_u32(sB + s * EB_STAGE));
        if (elect_one()) {
        // Per 32-byte k slab.  The first ETS slices of A (used by 7, 6, 5 products) are first copied to TMEM
        // (tcgen05.cp, executes in issue order with the products) and their products read only B from shared memory;
        // the other slices are read from shared memory by each product.  A copy costs about as much as 4 products'
        // worth of A reads, so it pays only for the most reused slices (all-TMEM and all-shared measured equal).
        if (kc == 0) {
#pragma unroll
          for (int kk = 0; kk < EKC / 32; kk++) {
#pragma unroll
            for (int i = 0; i < (ETS < ND ? ETS : ND); i++) utccp_128x32(tbase + EACOL + i * 8, da0 + ((i * (EBM * EKC) + kk * 2 * ELBO) >> 4));
#pragma unroll
            for (int i = 0; i < ND; i++)
#pragma unroll
              for (int j = 0; j < ND - i; j++) {
                const uint32_t acc = (kk > 0 || i > 0) ? 1u : 0u;
                if (i < ETS) umma_i8_ts(tbase + (i + j) * EBN, tbase + EACOL + i * 8, db0 + ((j * (EBN * EKC) + kk * 2 * ELBO) >> 4), idesc, acc);
                else umma_i8(tbase + (i + j) * EBN, da0 + ((i * (EBM * EKC) + kk * 2 * ELBO) >> 4),
                             db0 + ((j * (EBN * EKC) + kk * 2 * ELBO) >> 4), idesc, acc);
              }
          }
        } else {
#pragma unroll
          for (int kk = 0; kk < EKC / 32; kk++) {
#pragma unroll
            for (int i = 0; i < (ETS < ND ? ETS : ND); i++) utccp_128x32(tbase + EACOL + i * 8, da0 + ((i * (EBM * EKC) + kk * 2 * ELBO) >> 4));
#pragma unroll
            for (int i = 0; i < ND; i++)
#pragma unroll
              for (int j = 0; j < ND - i; j++) {
                if (i < ETS) umma_i8_ts(tbase + (i + j) * EBN, tbase + EACOL + i * 8, db0 + ((j * (EBN * EKC) + kk * 2 * ELBO) >> 4), idesc, 1u);
                else umma_i8(tbase + (i + j) * EBN, da0 + ((i * (EBM * EKC) + kk * 2 * ELBO) >> 4),
                             db0 + ((j * (EBN * EKC) + kk * 2 * ELBO) >> 4), idesc, 1u);
              }
          }
        }
        umma_commit(smem_u32(&empty_bar[s]));
        }
        __syncwarp();
      }
      if (elect_one()) umma_commit(smem_u32(&acc_full));
      __syncwarp();
      tl++;
    }
  } else if (warp >= 4) {
    // ===== epilogue: EPI_WARPS warps, warp = (lane quarter q of TMEM, column group g of EBN / EPI_G columns) =====
    // Many warps because this part is latency-bound (about 35 dependent integer / FP64 instructions per element): with one
    // warp per scheduler it took longer than the products of the tile.
    const int q = warp & 3, g = (warp - 4) >> 2;
    double* stg = reinterpret_cast<double*>(smem + ESTAGES * (EA_STAGE + EB_STAGE)) + (warp - 4) * (32 * 9);   // 32 x 8 (+1 pad)
    constexpr int GC = EBN / EPI_G;                 // columns per warp
    const int shl = p.q - 24, shr = 8 * (ND - 1) - p.q;
    const long long rnd = shr > 0 ? (1ll << (shr - 1)) : 0;   // round the dropped low bits to nearest
    int tl = 0;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      const int nb = (int)(t % p.NB), mb = (int)((t / p.NB) % p.MB), z = (int)(t / ((long)p.NB * p.MB));
      if (p.sym && nb * EBN + EBN - 1 < mb * EBM) continue;
      const int row = mb * EBM + q * 32 + lane;
      const bool rok = row < p.M;
      // the operand scales are powers of two: they go into the exponent field instead of FP64 multiplies
      const int ea = rok ? (int)((__double_as_longlong(p.sa[(long)z * p.M + row]) >> 52) & 0x7ff) - 1023 - p.q : 0;
      double rsv = rok ? p.alpha : 0.0;
      if (p.az) rsv *= p.az[(long)z * p.s_az];
      if (p.rs && rok) rsv *= p.rs[(long)z * p.s_rs + row];
      const int colbase = nb * EBN + g * GC;
      const double* sbz = p.sb + (long)z * p.N + colbase;
      if (!mbar_wait(smem_u32(&acc_full), (uint32_t)(tl & 1))) { atomicOr(p.err, 8); return; }
      asm volatile("tcgen05.fence::after_thread_sync;");
      // phase 1 (the issuer waits for it): drain this warp's columns from TMEM into fixed-point sums, hand TMEM back
      long long S[GC];
#pragma unroll
      for (int c0 = 0; c0 < GC; c0 += 8) {
        uint32_t v[ND][8];
#pragma unroll
        for (int d = 0; d < ND; d++)
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                       : "=r"(v[d][0]), "=r"(v[d][1]), "=r"(v[d][2]), "=r"(v[d][3]), "=r"(v[d][4]), "=r"(v[d][5]), "=r"(v[d][6]), "=r"(v[d][7])
                       : "r"(tbase + ((uint32_t)(q * 32) << 16) + d * EBN + g * GC + c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
        if (c0 + 8 >= GC) {
          asm volatile("tcgen05.fence::before_thread_sync;");
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&acc_empty)) : "memory");
        }
#pragma unroll
        for (int c = 0; c < 8; c++) {
          long long hi = 0, lo = 0;
#pragma unroll
          for (int d = 0; d < ND; d++) {
            if (d < 4) hi = hi * 256 + (long long)(int32_t)v[d][c];
            else lo = lo * 256 + (long long)(int32_t)v[d][c];
          }
          // S = V 2^q with V = sum_d acc_d 2^(-8 d): the bits of the low slices below 2^-q are dropped (q is chosen from K
          // so that S fits 63 bits; the dropped part is < 2^(-q-12) of max|a| max|b|, below the slicing error)
          S[c0 + c] = (hi << shl) + ((lo + rnd) >> shr);
        }
      }
      // phase 2 (overlaps the next tile's products): scale, then store through a per-warp transpose in shared memory so
      // that one global instruction covers 4 rows x 64 contiguous bytes instead of 32 rows x 8 bytes
      // One FP64 instruction per element: the power-of-two scales are applied in the exponent field and alpha rs[row]
      // rides along in the pad column of the staging tile.  (ncu attributes most stall samples of this loop to the
      // DFMA / DMUL; removing them altogether did not change the kernel time -- the products, not this loop, pace it.)
      const int rr = lane >> 3, cc = lane & 7;
      stg[lane * 9 + 8] = rsv;
#pragma unroll
      for (int c0 = 0; c0 < GC; c0 += 8) {
#pragma unroll
        for (int c = 0; c < 8; c++) {
          const double x0 = (double)S[c0 + c];
          const int col = colbase + c0 + c;
          const int e = ea + (col < p.N ? (int)((__double_as_longlong(sbz[c0 + c]) >> 52) & 0x7ff) - 1023 : 0);
          const int h0 = __double2hiint(x0), nf = ((h0 >> 20) & 0x7ff) + e;
          // exponent out of range: flush to zero / overflow to infinity (never seen with the panel's magnitudes)
          const int h1 = (S[c0 + c] == 0 || nf <= 0) ? 0 : (nf >= 2047 ? ((h0 & 0x80000000) | 0x7ff00000) : h0 + (e << 20));
          const int l1 = (S[c0 + c] == 0 || nf <= 0 || nf >= 2047) ? 0 : __double2loint(x0);
          stg[lane * 9 + c] = __hiloint2double(h1, l1);
        }
        __syncwarp();
        const int col = colbase + c0 + cc;
        if (col < p.N) {
          const double csv = p.cs ? p.cs[(long)z * p.s_cs + col] : 1.0;
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const int r = mb * EBM + q * 32 + i * 4 + rr;
            if (r < p.M && !(p.sym && r > col)) {           // symmetric: only the upper triangle is produced, then mirrored
              const double x = stg[(i * 4 + rr) * 9 + cc], rv = stg[(i * 4 + rr) * 9 + 8];
              double o;
              if (!p.E) o = rv * x;
              else if (p.beta == 1.0) o = fma(rv, x, p.E[(long)z * p.sE + (long)r * p.lde + col]);
              else o = fma(rv, x, p.beta * p.E[(long)z * p.sE + (long)r * p.lde + col]);
              if (p.cs) o *= csv;
              if (p.diag != 0.0 && r == col) o += p.diag;
              p.D[(long)z * p.sD + (long)r * p.ldd + col] = o;
              if (p.D2) {
                const long o2 = (long)z * p.s2 + (long)r * p.ld2 + col;
                p.D2[o2] = p.E2[o2] + o;
              }
            }
          }
        }
        __syncwarp();
      }
      tl++;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase));
}

}  // namespace

// ---- host API (declared in common.cuh) ---------------------------------------------------------------------
size_t emu_a_slices_bytes(int batch, int M, int K) { return (size_t)batch * ceil_div(M, EBM) * (K / EKC) * (size_t)EA_STAGE; }
size_t emu_b_slices_bytes(int batch, int K, int N) { return (size_t)batch * ceil_div(N, EBN) * (K / EKC) * (size_t)EB_STAGE; }

bool emu_shape_ok(int M, int N, int K) { return M > 0 && N > 0 && K >= EKC && K % EKC == 0 && (long)K * 128 * 128 * ES < (1L << 31); }

// A [M][K] (K contiguous) -> slices + row scales 2^ea
int emu_slice_a(const double* A, long lda, long sA, int M, int K, int batch, int8_t* out, double* scale, int* err, cudaStream_t s) {
  const int rows = ceil_div(M, EBM) * EBM;
  if (K <= ONEPASS_KMAX && (lda % 2) == 0 && (sA % 2) == 0 && ((uintptr_t)A % 16) == 0)
    slice_kmajor512_kernel<EBM><<<dim3(ceil_div(rows, 8), batch), 256, 0, s>>>(A, lda, sA, M, K, out, scale, 1.0, err);
  else
    slice_kmajor_kernel<EBM><<<dim3(ceil_div(rows, 8), batch), 256, 0, s>>>(A, lda, sA, M, K, out, scale, 1.0, err);
  PFML_CHECK_LAUNCH();
  return 0;
}

// B [K][N] (N contiguous) -> slices + column scales 2^(eb - 14); ex [batch][N] is int scratch
int emu_slice_b(const double* B, long ldb, long sB, int K, int N, int batch, int8_t* out, double* scale, int* ex, int* err,
                cudaStream_t s) {
  if (K <= ONEPASS_KMAX) {
    // 64-column blocks are written by several CTAs of SNC columns; the padded part of the last block is zero-filled too
    slice_nmajor512_kernel<<<dim3((EBN / SNC) * ceil_div(N, EBN), batch), SNC * (K / 32), 0, s>>>(B, ldb, sB, K, N, ceil_div(N, EBN), ex, scale, kBScale,
                                                                             out, 0, nullptr, nullptr, err, nullptr, 0, nullptr);
    PFML_CHECK_LAUNCH();
    return 0;
  }
  col_exponent_kernel<<<dim3(ceil_div(N, 128), batch), 128, 0, s>>>(B, ldb, sB, K, N, ex, scale, kBScale, err);
  PFML_CHECK_LAUNCH();
  slice_nmajor_kernel<<<dim3(ceil_div(N, EBN), K / EKC, batch), 256, 0, s>>>(B, ldb, sB, K, N, ceil_div(N, EBN), ex, out);
  PFML_CHECK_LAUNCH();
  return 0;
}

// X [K][C] (C contiguous, K <= 1024) -> B-type slices of X (outB, may be null) and / or A-type slices of X' (outA, may be null)
int emu_slice_nmajor512(const double* X, long ld, long sX, int K, int C, int batch, int8_t* outB, double* sb, int* ex, int8_t* outA,
                        double* sa, int* err, cudaStream_t s, const double* row_scale, long s_row_scale, int* zflag) {
  if (K > ONEPASS_KMAX || K % EKC != 0) {
    set_error("emu_slice_nmajor512: K = %d (needs K %% 64 == 0 and K <= %d)", K, ONEPASS_KMAX);
    return -2;
  }
  const int CB = ceil_div(C, EBN), RBA = ceil_div(C, EBM);
  const int blocks = outA ? (EBM / SNC) * RBA : (EBN / SNC) * CB;   // SNC columns per CTA; the padded tail of the last block is zero-filled
  slice_nmajor512_kernel<<<dim3(blocks, batch), SNC * (K / 32), 0, s>>>(X, ld, sX, K, C, CB, ex, sb, kBScale, outB, RBA, sa, outA, err, row_scale,
                                                                        s_row_scale, zflag);
  PFML_CHECK_LAUNCH();
  return 0;
}

// slabs tau0 .. tau0 + ntau - 1 of the standardised signals as B-type slices (out: ntau blocks of emu_b_slices_bytes(B, NP, P),
// scale_out [ntau][B][P])
int emu_standardize_slices(const PanelView& pv, int B, int NP, int P, int tau0, int ntau, const int* date_t, const int* sel,
                           const int* nsel, const BatchVecs& v, int8_t* out, double* scale_out, int* err, cudaStream_t s) {
  const int CB = ceil_div(P, EBN);
  dim3 grid(CB, ntau, B);
  if (NP <= ONEPASS_KMAX) {
    standardize_slice512_kernel<<<dim3((EBN / SNC) * CB, ntau, B + ntau - 1), SNC * (NP / 32), 0, s>>>(pv, NP, P, tau0, date_t, sel, nsel, v, out,
                                                                      (long)emu_b_slices_bytes(B, NP, P), scale_out, err);
    PFML_CHECK_LAUNCH();
    return 0;
  }
  const size_t smem = (size_t)NP * (2 * sizeof(double) + sizeof(int)) + SSQ * 64 * sizeof(double);
  if (smem > 200 * 1024) {
    set_error("emu_standardize_slices: universe of %d stocks needs %zu bytes of shared memory", NP, smem);
    return -2;
  }
  if (smem > 48 * 1024)
    PFML_CUDA_OK(cudaFuncSetAttribute(standardize_slice_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  standardize_slice_kernel<<<grid, 64 * SSQ, smem, s>>>(pv, NP, P, tau0, date_t, sel, nsel, v, out,
                                                        (long)emu_b_slices_bytes(B, NP, P), scale_out, err);
  PFML_CHECK_LAUNCH();
  return 0;
}

// D = alpha * rs[r] * (A B) + beta * E from pre-sliced operands
int emu_mma(const int8_t* As, const double* sa, const int8_t* Bs, const double* sb, int M, int N, int K, int batch, double alpha,
            const double* rs, long s_rs, double beta, const double* E, long lde, long sE, double* D, long ldd, long sD, int* err,
            cudaStream_t s) {
  return emu_mma_ex(As, sa, Bs, sb, M, N, K, batch, alpha, rs, s_rs, beta, E, lde, sE, D, ldd, sD, EmuExtra(), err, s);
}

int emu_mma_ex(const int8_t* As, const double* sa, const int8_t* Bs, const double* sb, int M, int N, int K, int batch, double alpha,
               const double* rs, long s_rs, double beta, const double* E, long lde, long sE, double* D, long ldd, long sD,
               const EmuExtra& x, int* err, cudaStream_t s) {
  if (!emu_shape_ok(M, N, K)) {
    set_error("emu_mma: unsupported shape M=%d N=%d K=%d", M, N, K);
    return -2;
  }
  EmuP p;
  p.As = As; p.Bs = Bs; p.sa = sa; p.sb = sb; p.D = D; p.ldd = ldd; p.sD = sD; p.E = E; p.lde = lde; p.sE = sE;
  p.rs = rs; p.s_rs = s_rs; p.alpha = alpha; p.beta = beta; p.M = M; p.N = N; p.K = K;
  p.MB = ceil_div(M, EBM); p.NB = ceil_div(N, EBN); p.batch = batch; p.err = err;
  p.cs = x.cs; p.s_cs = x.s_cs; p.diag = x.diag; p.az = x.az; p.s_az = x.s_az; p.sym = x.sym ? 1 : 0; p.D2 = x.D2; p.E2 = x.E2; p.ld2 = x.ld2; p.s2 = x.s2;
  // |V| = |sum_d acc_d 2^(-8 d)| < 1.01 K 128^2 < 2^vb: the fixed-point position q leaves one spare bit below the sign
  int vb = 0;
  while ((1.01 * K * 16384.0) >= (double)(1ll << vb)) vb++;
  p.q = 62 - vb;
  const int digits = (x.digits >= 4 && x.digits < ES) ? x.digits : ES;
  if (digits != ES && digits != 6) {
    set_error("emu_mma: %d digits not built (6 or %d)", digits, ES);
    return -2;
  }
  if (p.q > 8 * (digits - 1)) p.q = 8 * (digits - 1);
  constexpr size_t smem = (size_t)ESTAGES * (EA_STAGE + EB_STAGE) + (size_t)(4 * EPI_G) * 32 * 9 * sizeof(double);
  static std::once_flag once[64];
  static int n_sm[64] = {};
  static cudaError_t cfg_rc[64] = {};
  int dev = 0;
  PFML_CUDA_OK(cudaGetDevice(&dev));
  std::call_once(once[dev & 63], [&] {       // per device, safe under concurrent first calls
    cudaError_t rc = cudaFuncSetAttribute(emu_gemm_kernel<ES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (rc == cudaSuccess) rc = cudaFuncSetAttribute(emu_gemm_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (rc == cudaSuccess) rc = cudaDeviceGetAttribute(&n_sm[dev & 63], cudaDevAttrMultiProcessorCount, dev);
    cfg_rc[dev & 63] = rc;
  });
  PFML_CUDA_OK(cfg_rc[dev & 63]);
  const long tiles = (long)batch * p.MB * p.NB;
  const long want = n_sm[dev & 63];                       // persistent: one CTA per SM (it owns the SM's whole TMEM)
  const int grid = (int)(tiles < want ? tiles : want);
  prof_begin(-2, s);
  if (digits == ES) emu_gemm_kernel<ES><<<grid, EMU_THREADS, smem, s>>>(p);
  else emu_gemm_kernel<6><<<grid, EMU_THREADS, smem, s>>>(p);
  PFML_CHECK_LAUNCH();
  prof_end(-2, s, 2.0 * M * (double)N * K * batch, 0.0);
  return 0;
}

}  // namespace pfml
