// Batched FP64 GEMM for sm_100a on the DMMA pipe.
//
// tcgen05.mma has no f64 kind, so on Blackwell the FP64 tensor path is the warp-level
// mma.sync.m8n8k4.f64 (SASS: DMMA.8x8x4; the m16n8k{4,8,16} PTX shapes lower to the same instruction on
// sm_100a).  Measured issue-bound peak on this pool's B200: 37.05 TFLOP/s (= 148 SMs x 64 FMA/clk x
// 1.965 GHz); cuBLAS DGEMM 8192^3 reaches 35.4.  One DMMA.8x8x4 needs only two 8-byte operand loads per
// thread for 8 FMAs and the pipe retires one instruction per 16 clk per SM sub-partition, so operand
// delivery is cheap: a 64x64x16 CTA tile needs 16 B/clk/SM from L2, far below the 40 B/clk/SM share of
// the L2 fabric.  The kernel is therefore a classic multi-stage cp.async (LDGSTS) pipeline into padded
// shared-memory tiles (conflict-free for the (groupID, threadID_in_group) fragment pattern), with the
// diagonal scalings / addends / symmetric mirroring the PFML chain needs fused into the epilogue.
#include "common.cuh"

namespace pfml {

namespace {

constexpr int PAD = 4;       // stride == 4 (mod 16) doubles -> the 16 lanes of a half-warp hit 16 distinct bank pairs

__device__ __forceinline__ void cp_async16(void* smem, const void* g, bool pred) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  int sz = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(g), "r"(sz));
}
__device__ __forceinline__ void cp_async8(void* smem, const void* g, bool pred) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  int sz = pred ? 8 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(s), "l"(g), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// Load one operand tile into shared memory.
//   KIN  (k contiguous in global): tile is [ROWS][BK] with row stride BK+PAD; element (r,k) = G[(r0+r)*ld + k0+k]
//   !KIN (k is the row in global):  tile is [BK][ROWS] with row stride ROWS+PAD; element (k,r) = G[(k0+k)*ld + r0+r]
// Out-of-range elements are zero-filled (cp.async src-size 0).  VEC: 16-byte copies (needs even ld / offsets
// and 16-byte aligned base), otherwise 8-byte copies.
template <int ROWS, int BK, bool KIN, bool VEC, int NT>
__device__ __forceinline__ void load_tile(double* s, const double* __restrict__ G, long ld, int r0, int k0, int R, int K, int tid) {
  if (KIN) {
    constexpr int S = BK + PAD;
    if (VEC) {
      constexpr int CH = ROWS * (BK / 2);
#pragma unroll
      for (int c = tid; c < CH; c += NT) {
        int r = c / (BK / 2), kc = (c % (BK / 2)) * 2;
        bool ok = (r0 + r < R) && (k0 + kc < K);
        const double* src = ok ? G + (long)(r0 + r) * ld + k0 + kc : G;
        cp_async16(s + r * S + kc, src, ok);
      }
    } else {
      constexpr int CH = ROWS * BK;
#pragma unroll
      for (int c = tid; c < CH; c += NT) {
        int r = c / BK, kc = c % BK;
        bool ok = (r0 + r < R) && (k0 + kc < K);
        const double* src = ok ? G + (long)(r0 + r) * ld + k0 + kc : G;
        cp_async8(s + r * S + kc, src, ok);
      }
    }
  } else {
    constexpr int S = ROWS + PAD;
    if (VEC) {
      constexpr int CH = BK * (ROWS / 2);
#pragma unroll
      for (int c = tid; c < CH; c += NT) {
        int k = c / (ROWS / 2), rc = (c % (ROWS / 2)) * 2;
        bool ok = (k0 + k < K) && (r0 + rc < R);
        const double* src = ok ? G + (long)(k0 + k) * ld + r0 + rc : G;
        cp_async16(s + k * S + rc, src, ok);
      }
    } else {
      constexpr int CH = BK * ROWS;
#pragma unroll
      for (int c = tid; c < CH; c += NT) {
        int k = c / ROWS, rc = c % ROWS;
        bool ok = (k0 + k < K) && (r0 + rc < R);
        const double* src = ok ? G + (long)(k0 + k) * ld + r0 + rc : G;
        cp_async8(s + k * S + rc, src, ok);
      }
    }
  }
}

template <int ROWS, int BK, bool KIN>
struct TileSize {
  static constexpr int value = KIN ? ROWS * (BK + PAD) : BK * (ROWS + PAD);
};

template <int BM, int BN, int WM, int WN, int BK, int STAGES, int MINB, bool AKIN, bool BKIN, bool VEC>
__global__ void __launch_bounds__((BM / WM) * (BN / WN) * 32, MINB)
gemm_kernel(const GemmP p) {
  constexpr int NWM = BM / WM, NWN = BN / WN;
  constexpr int NT = NWM * NWN * 32;
  constexpr int MT = WM / 8, NTL = WN / 8;            // 8x8 DMMA tiles per warp
  constexpr int A_SZ = TileSize<BM, BK, AKIN>::value;
  constexpr int B_SZ = TileSize<BN, BK, BKIN>::value;
  constexpr int SA = AKIN ? BK + PAD : BM + PAD;
  constexpr int SB = BKIN ? BK + PAD : BN + PAD;

  const int tile_m = blockIdx.y, tile_n = blockIdx.x, z = blockIdx.z;
  if (p.sym && tile_n * BN > tile_m * BM + BM - 1) return;

  extern __shared__ __align__(16) double smem[];
  double* As = smem;
  double* Bs = smem + STAGES * A_SZ;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  const int wm0 = (warp / NWN) * WM, wn0 = (warp % NWN) * WN;
  const int m0 = tile_m * BM, n0 = tile_n * BN;

  const double* __restrict__ A = p.A + (long)z * p.sA;
  const double* __restrict__ B = p.B + (long)z * p.sB;

  double acc[MT][NTL][2];
#pragma unroll
  for (int i = 0; i < MT; i++)
#pragma unroll
    for (int j = 0; j < NTL; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
  // 8x8 tiles of this warp that hold at least one valid output (edge CTAs skip the DMMA work of the rest,
  // e.g. the 9th column tile of P = 1025 holds a single valid column)
  const int imax = min(MT, max(0, (p.M - m0 - wm0 + 7) >> 3));
  const int jmax = min(NTL, max(0, (p.N - n0 - wn0 + 7) >> 3));
  const bool full = (imax == MT) && (jmax == NTL);
  const bool active = (imax > 0) && (jmax > 0);

  const int KT = (p.K + BK - 1) / BK;
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < KT) {
      load_tile<BM, BK, AKIN, VEC, NT>(As + s * A_SZ, A, p.lda, m0, s * BK, p.M, p.K, tid);
      load_tile<BN, BK, BKIN, VEC, NT>(Bs + s * B_SZ, B, p.ldb, n0, s * BK, p.N, p.K, tid);
    }
    cp_async_commit();
  }

  for (int kt = 0; kt < KT; kt++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    {
      int nk = kt + STAGES - 1;
      if (nk < KT) {
        int s = nk % STAGES;
        load_tile<BM, BK, AKIN, VEC, NT>(As + s * A_SZ, A, p.lda, m0, nk * BK, p.M, p.K, tid);
        load_tile<BN, BK, BKIN, VEC, NT>(Bs + s * B_SZ, B, p.ldb, n0, nk * BK, p.N, p.K, tid);
      }
      cp_async_commit();
    }
    const double* as = As + (kt % STAGES) * A_SZ;
    const double* bs = Bs + (kt % STAGES) * B_SZ;
    if (full) {
#pragma unroll
      for (int kk = 0; kk < BK; kk += 4) {
        double af[MT], bf[NTL];
#pragma unroll
        for (int i = 0; i < MT; i++) af[i] = AKIN ? as[(wm0 + 8 * i + g) * SA + kk + t] : as[(kk + t) * SA + wm0 + 8 * i + g];
#pragma unroll
        for (int j = 0; j < NTL; j++) bf[j] = BKIN ? bs[(wn0 + 8 * j + g) * SB + kk + t] : bs[(kk + t) * SB + wn0 + 8 * j + g];
#pragma unroll
        for (int i = 0; i < MT; i++)
#pragma unroll
          for (int j = 0; j < NTL; j++) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
      }
    } else if (active) {
#pragma unroll
      for (int kk = 0; kk < BK; kk += 4) {
        double af[MT], bf[NTL];
#pragma unroll
        for (int i = 0; i < MT; i++) af[i] = AKIN ? as[(wm0 + 8 * i + g) * SA + kk + t] : as[(kk + t) * SA + wm0 + 8 * i + g];
#pragma unroll
        for (int j = 0; j < NTL; j++) bf[j] = BKIN ? bs[(wn0 + 8 * j + g) * SB + kk + t] : bs[(kk + t) * SB + wn0 + 8 * j + g];
#pragma unroll
        for (int i = 0; i < MT; i++)
#pragma unroll
          for (int j = 0; j < NTL; j++)
            if (i < imax && j < jmax) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
      }
    }
  }
  cp_async_wait<0>();

  // ---- epilogue -----------------------------------------------------------------------------------
  double alpha = p.alpha, beta = p.beta;
  if (p.az) alpha *= p.az[(long)z * p.s_z];
  if (p.bz) beta *= p.bz[(long)z * p.s_z];
  double* D = p.D + (long)z * p.sD;   // may alias E / E2 (element-wise in-place updates)
  const double* E = p.E ? p.E + (long)z * p.sE : nullptr;
  double* D2 = p.D2 ? p.D2 + (long)z * p.sD : nullptr;
  double* DT = p.DT ? p.DT + (long)z * p.sD : nullptr;
  const double* E2 = p.E2 ? p.E2 + (long)z * p.sD : nullptr;
  const double* rs = p.rs ? p.rs + (long)z * p.s_rs : nullptr;
  const double* cs = p.cs ? p.cs + (long)z * p.s_cs : nullptr;
#pragma unroll
  for (int i = 0; i < MT; i++) {
    const int r = m0 + wm0 + 8 * i + g;
    if (r >= p.M) continue;
    const double rsv = rs ? alpha * rs[r] : alpha;
#pragma unroll
    for (int j = 0; j < NTL; j++) {
      const int c = n0 + wn0 + 8 * j + 2 * t;
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int cc = c + e;
        if (cc >= p.N) continue;
        if (p.sym && cc > r) continue;
        double v = acc[i][j][e] * rsv;
        if (cs) v *= cs[cc];
        if (E) v += beta * E[(long)r * p.lde + cc];
        if (r == cc) v += p.diag;
        if (p.sincos) {
          double sv, cv;
          sincos(v, &sv, &cv);
          D[(long)r * p.ldd + cc] = cv;
          D[(long)r * p.ldd + cc + p.N] = sv;
          continue;
        }
        D[(long)r * p.ldd + cc] = v;
        if (p.sym && cc != r) D[(long)cc * p.ldd + r] = v;
        if (DT) DT[(long)cc * p.ldt + r] = v;
        if (D2) {
          double v2 = v + E2[(long)r * p.ldd + cc];
          D2[(long)r * p.ldd + cc] = v2;
          if (p.sym && cc != r) D2[(long)cc * p.ldd + r] = v2;
        }
      }
    }
  }
}

template <int BM, int BN, int WM, int WN, int BK, int STAGES, int MINB, bool AKIN, bool BKIN, bool VEC>
int launch(const GemmP& p, int batch, cudaStream_t stream) {
  constexpr int NT = (BM / WM) * (BN / WN) * 32;
  constexpr size_t smem = (size_t)STAGES * (TileSize<BM, BK, AKIN>::value + TileSize<BN, BK, BKIN>::value) * sizeof(double);
  auto kern = gemm_kernel<BM, BN, WM, WN, BK, STAGES, MINB, AKIN, BKIN, VEC>;
  PFML_CONFIG_ONCE(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   // per instantiation and device
  dim3 grid(ceil_div(p.N, BN), ceil_div(p.M, BM), batch);
  kern<<<grid, NT, smem, stream>>>(p);
  PFML_CHECK_LAUNCH();
  return 0;
}

template <int BM, int BN, int WM, int WN, int BK, int STAGES, int MINB>
int dispatch_layout(const GemmP& p, int batch, cudaStream_t stream, bool vec) {
#define PFML_GEMM_CASE(AK, BK_)                                                                 \
  if (p.a_kinner == AK && p.b_kinner == BK_) {                                                  \
    return vec ? launch<BM, BN, WM, WN, BK, STAGES, MINB, AK, BK_, true>(p, batch, stream)      \
               : launch<BM, BN, WM, WN, BK, STAGES, MINB, AK, BK_, false>(p, batch, stream);    \
  }
  PFML_GEMM_CASE(true, false)
  PFML_GEMM_CASE(false, false)
  PFML_GEMM_CASE(true, true)
  PFML_GEMM_CASE(false, true)
#undef PFML_GEMM_CASE
  return -1;
}

inline bool aligned16(const void* ptr) { return ((uintptr_t)ptr & 15) == 0; }

}  // namespace

double gemm_flops(const GemmP& p, int batch) {
  double f = 2.0 * p.M * (double)p.N * p.K * batch;
  return p.sym ? 0.5 * f : f;
}

int gemm(const GemmP& p, int batch, cudaStream_t stream) {
  if (p.M <= 0 || p.N <= 0 || batch <= 0) return 0;
  if (p.K < 0 || !p.A || !p.B || !p.D) {
    set_error("gemm: bad arguments");
    return -2;
  }
  if (p.sym && p.M != p.N) {
    set_error("gemm: sym requires a square output");
    return -2;
  }
  bool vec = aligned16(p.A) && aligned16(p.B) && (p.lda % 2 == 0) && (p.ldb % 2 == 0) && (p.sA % 2 == 0) &&
             (p.sB % 2 == 0) && (p.a_kinner ? (p.K % 2 == 0) : true) && (p.b_kinner ? (p.K % 2 == 0) : true);
  // K-outer operands copy pairs along M/N: an odd edge only reads one element past the logical row inside
  // the (even) leading dimension, and that lane's output column is never stored.
  prof_begin(-1, stream);
  int rc;
  // One configuration for every shape: 64x64x16 CTA tile, 4 warps x (32x32), 3 cp.async stages, 3 CTAs per SM.
  // Measured on B200 for the Horner step (148 x [512 x 1025 x 512]): 30.5 TFLOP/s, against 24.2 for
  // 128x128 / 8 warps / 1 CTA per SM and 27.9 for 128x64 / 2 CTAs per SM -- independent CTAs hide each other's
  // per-k-tile barrier, which a single big CTA cannot (profiles/r01_gemm_tile_sweep.md).
  // Small outputs (the 128- and 256-blocks of the inverse recursion): 4 CTAs per SM, so that batch x 4 or batch x 16
  // tiles of a 148-matrix batch fill whole waves (592 resident CTAs) instead of 1.33 / 5.33 waves of 444.
  if (p.M <= 256 && p.N <= 256) rc = dispatch_layout<64, 64, 32, 32, 16, 3, 4>(p, batch, stream, vec);
  else rc = dispatch_layout<64, 64, 32, 32, 16, 3, 3>(p, batch, stream, vec);
  prof_end(-1, stream, gemm_flops(p, batch), 0.0);
  return rc;
}

}  // namespace pfml
