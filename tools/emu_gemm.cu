// Prototype: FP64 GEMM emulated on the INT8 tensor cores (tcgen05.mma kind::i8, accumulators in TMEM), Ozaki-style.
//
//   C[M x N] = A[M x K] B[K x N]   with  a_ik = 2^ea_i sum_t qa_t 2^(-7(t+1)),  b_kj = 2^eb_j sum_t qb_t 2^(-7(t+1)),
//   qa_t, qb_t in [-127, 127] (S = 8 slices = 56 bits), and only the slice pairs with i + j <= S-1 are multiplied:
//   C_ij = 2^(ea_i + eb_j) sum_d 2^(-7(d+2)) [ sum_{i+j=d} qa_i . qb_j ]   -- every bracket is an EXACT int32 sum
//   (K 127^2 (d+1) <= 512 * 16129 * 8 < 2^26), so the only errors are the dropped pairs (~2^-44 |a|max |b|max K).
//
// Slices are stored in global memory already in the shared-memory order the UMMA descriptors expect (K-major, no
// swizzle, 8 x 16 B core matrices), one contiguous block per (row block, 64-byte k chunk), so that a pipeline stage
// is ONE bulk copy (cp.async.bulk, the 1-D TMA) per operand.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o emu_gemm emu_gemm.cu && ./emu_gemm [batch]
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

constexpr int S = 8;            // slices
constexpr int W = 7;            // bits per slice
constexpr int BM = 128, BN = 64, KC = 64;                  // CTA tile, k chunk (bytes) per stage
constexpr int LBO = 128, SBO = (KC / 16) * 128;            // canonical layout strides (bytes)
constexpr int A_STAGE = S * BM * KC, B_STAGE = S * BN * KC;  // 64 KB, 32 KB
constexpr int STAGES = 2;

__host__ __device__ inline int canon(int r, int kk) { return (r % 8) * 16 + (r / 8) * SBO + (kk / 16) * LBO + kk % 16; }

// ---- slicing ---------------------------------------------------------------------------------------
// A [M][K] row-major (lda), scale per row.  Output As[z][m/BM][k/KC][slice][canon(128 x KC)], ea[z][M].
__device__ __forceinline__ void slice16(const double (&x)[16], int8_t* dst, long slice_stride) {
  double r[16];
#pragma unroll
  for (int k = 0; k < 16; k++) r[k] = x[k];
#pragma unroll
  for (int t = 0; t < S; t++) {
    union { int8_t b[16]; int4 v; } u;
#pragma unroll
    for (int k = 0; k < 16; k++) {
      r[k] *= 128.0;
      const double qd = trunc(r[k]);
      r[k] -= qd;
      u.b[k] = (int8_t)(int)qd;
    }
    *reinterpret_cast<int4*>(dst + t * slice_stride) = u.v;
  }
}

__global__ void slice_rows(const double* __restrict__ A, long lda, long sA, int M, int K, int8_t* __restrict__ As, double* __restrict__ ea) {
  // one warp per row: row maximum, then every lane slices 16 consecutive k (one 16-byte core-matrix row per slice)
  const int z = blockIdx.y, row = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32, lane = threadIdx.x & 31;
  if (row >= M) return;
  const double* a = A + (long)z * sA + (long)row * lda;
  double mx = 0.0;
  for (int k = lane; k < K; k += 32) mx = fmax(mx, fabs(a[k]));
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  const int e = mx > 0.0 ? ilogb(mx) + 1 : 0;
  if (lane == 0) ea[(long)z * M + row] = scalbn(1.0, e);
  const int MB = M / BM, KB = K / KC;
  int8_t* base = As + (((long)z * MB + row / BM) * KB) * (long)A_STAGE;
  for (int k0 = lane * 16; k0 < K; k0 += 32 * 16) {
    double x[16];
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = scalbn(a[k0 + k], -e);
    slice16(x, base + (long)(k0 / KC) * A_STAGE + canon(row % BM, k0 % KC), (long)BM * KC);
  }
}

// B [K][N] row-major (ldb), scale per column.  Output Bs[z][n/BN][k/KC][slice][canon(64 x KC)] (n is the "row").
__global__ void col_exponents(const double* __restrict__ B, long ldb, long sB, int K, int N, int* __restrict__ eb, double* __restrict__ ebs) {
  const int z = blockIdx.y, n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const double* b = B + (long)z * sB + n;
  double mx = 0.0;
  for (int k = 0; k < K; k++) mx = fmax(mx, fabs(b[(long)k * ldb]));
  const int e = mx > 0.0 ? ilogb(mx) + 1 : 0;
  eb[(long)z * N + n] = e;
  ebs[(long)z * N + n] = scalbn(1.0, e - 14);      // includes the 2^-14 of the two leading slice scalings
}

__global__ void slice_cols(const double* __restrict__ B, long ldb, long sB, int K, int N, int NB, const int* __restrict__ eb,
                           int8_t* __restrict__ Bs) {
  // one CTA (256 threads) per (n block of 64, k chunk of 64): thread = (n, 16 consecutive k); reads are coalesced along n
  const int z = blockIdx.z, nb = blockIdx.x, kb = blockIdx.y, KB = K / KC;
  const int nn = threadIdx.x & 63, kq = threadIdx.x >> 6, n = nb * BN + nn, k0 = kb * KC + kq * 16;
  int8_t* base = Bs + (((long)z * NB + nb) * KB + kb) * (long)B_STAGE;
  double x[16];
  const int e = n < N ? eb[(long)z * N + n] : 0;
#pragma unroll
  for (int k = 0; k < 16; k++) x[k] = n < N ? scalbn(B[(long)z * sB + (long)(k0 + k) * ldb + n], -e) : 0.0;
  slice16(x, base + canon(nn, kq * 16), (long)BN * KC);
}

// ---- tcgen05 helpers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((LBO >> 4) & 0x3FFF) << 16) | ((uint64_t)((SBO >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46);
}
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {     // bounded: a bug must not hang the GPU
  uint32_t done = 0;
  for (int it = 0; it < (1 << 24) && !done; it++)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done != 0;
}
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
               ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// ---- the emulated GEMM: one 128 x 64 output tile per CTA ------------------------------------------------
// grid (N/64 tiles, M/128 tiles, batch); 256 threads: warp 0 producer, warp 1 MMA issuer, warp 2 TMEM allocator,
// warps 4-7 epilogue (TMEM lane quarter = warp % 4).
__global__ void __launch_bounds__(256, 1)
emu_gemm_kernel(const int8_t* __restrict__ As, const int8_t* __restrict__ Bs, const double* __restrict__ ea, const double* __restrict__ eb,
                double* __restrict__ C, long ldc, long sC, int M, int N, int K, int* __restrict__ err) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                              // [STAGES][A_STAGE]
  uint8_t* sB = smem + STAGES * A_STAGE;           // [STAGES][B_STAGE]
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], acc_bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nb = blockIdx.x, mb = blockIdx.y, z = blockIdx.z;
  const int KB = K / KC, MB = M / BM, NB = gridDim.x;
  if (tid == 0) {
    for (int s = 0; s < STAGES; s++) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full_bar[s])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty_bar[s])));
    }
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&acc_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tbase = tmem_base;

  if (warp == 0 && lane == 0) {
    // ===== producer: one bulk copy per operand per stage =====
    const int8_t* a_src = As + (((long)z * MB + mb) * KB) * (long)A_STAGE;
    const int8_t* b_src = Bs + (((long)z * NB + nb) * KB) * (long)B_STAGE;
    for (int kc = 0; kc < KB; kc++) {
      const int s = kc % STAGES;
      if (!mbar_wait(smem_u32(&empty_bar[s]), ((kc / STAGES) & 1) ^ 1)) { atomicExch(err, 1); break; }
      const uint32_t fb = smem_u32(&full_bar[s]);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(A_STAGE + B_STAGE) : "memory");
#ifdef NO_LOAD
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(0) : "memory"); continue;
#endif
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(sA + s * A_STAGE)), "l"(a_src + (long)kc * A_STAGE), "r"(A_STAGE), "r"(fb) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(sB + s * B_STAGE)), "l"(b_src + (long)kc * B_STAGE), "r"(B_STAGE), "r"(fb) : "memory");
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    bool ok = true;
    for (int kc = 0; kc < KB && ok; kc++) {
      const int s = kc % STAGES;
      if (!mbar_wait(smem_u32(&full_bar[s]), (kc / STAGES) & 1)) { atomicExch(err, 2); ok = false; break; }
      asm volatile("tcgen05.fence::after_thread_sync;");
      // descriptors differ only in the 14-bit start-address field: add (byte offset >> 4) to a per-stage base
      const uint64_t da0 = make_desc(smem_u32(sA + s * A_STAGE)), db0 = make_desc(smem_u32(sB + s * B_STAGE));
      if (kc == 0) {
#pragma unroll
        for (int i = 0; i < S; i++)
#pragma unroll
          for (int j = 0; j < S - i; j++)
#pragma unroll
            for (int kk = 0; kk < KC / 32; kk++)
              umma_i8(tbase + (i + j) * BN, da0 + ((i * (BM * KC) + kk * 2 * LBO) >> 4), db0 + ((j * (BN * KC) + kk * 2 * LBO) >> 4),
                      idesc, (kk > 0 || i > 0) ? 1u : 0u);
      } else {
#pragma unroll
        for (int i = 0; i < S; i++)
#pragma unroll
          for (int j = 0; j < S - i; j++)
#pragma unroll
            for (int kk = 0; kk < KC / 32; kk++)
              umma_i8(tbase + (i + j) * BN, da0 + ((i * (BM * KC) + kk * 2 * LBO) >> 4), db0 + ((j * (BN * KC) + kk * 2 * LBO) >> 4),
                      idesc, 1u);
      }
      umma_commit(smem_u32(&empty_bar[s]));         // smem stage free once these MMAs have read it
    }
    if (ok) umma_commit(smem_u32(&acc_bar));
  } else if (warp >= 4) {
    // ===== epilogue: combine the 8 exact int32 sums in FP64 =====
    if (!mbar_wait(smem_u32(&acc_bar), 0)) { atomicExch(err, 3); }
    else {
      asm volatile("tcgen05.fence::after_thread_sync;");
      const int q = warp & 3, row = mb * BM + q * 32 + lane;
      const double rs = ea[(long)z * M + row];
      double* crow = C + (long)z * sC + (long)row * ldc + (long)nb * BN;
#ifndef NO_EPI
      // X = sum_d X_d 2^(-7 d):  hi = sum_{d<4} X_d 2^(7 (3-d)),  lo = sum_{d>=4} X_d 2^(7 (7-d))  (exact in int64),
      // X = (hi + lo 2^-28) 2^-21  -> two int64->double conversions per element instead of eight
      for (int c0 = 0; c0 < BN; c0 += 16) {
        long long hi[16], lo[16];
#pragma unroll
        for (int c = 0; c < 16; c++) hi[c] = lo[c] = 0;
        uint32_t v[S][16];
#pragma unroll
        for (int d = 0; d < S; d++)
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                       : "=r"(v[d][0]), "=r"(v[d][1]), "=r"(v[d][2]), "=r"(v[d][3]), "=r"(v[d][4]), "=r"(v[d][5]), "=r"(v[d][6]), "=r"(v[d][7]),
                         "=r"(v[d][8]), "=r"(v[d][9]), "=r"(v[d][10]), "=r"(v[d][11]), "=r"(v[d][12]), "=r"(v[d][13]), "=r"(v[d][14]), "=r"(v[d][15])
                       : "r"(tbase + ((uint32_t)(q * 32) << 16) + d * BN + c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
#pragma unroll
        for (int d = 0; d < S; d++)
#pragma unroll
          for (int c = 0; c < 16; c++) {
            if (d < 4) hi[c] = hi[c] * 128 + (long long)(int32_t)v[d][c];
            else lo[c] = lo[c] * 128 + (long long)(int32_t)v[d][c];
          }
#pragma unroll
        for (int c = 0; c < 16; c++) {
          const int col = nb * BN + c0 + c;
          const double x = ((double)hi[c] + (double)lo[c] * 3.725290298461914e-09) * 4.76837158203125e-07;   // 2^-28, 2^-21
          if (col < N) crow[c0 + c] = rs * (x * eb[(long)z * N + col]);
        }
      }
#endif
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase));
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

int main(int argc, char** argv) {
  const int batch = argc > 1 ? atoi(argv[1]) : 4;
  const int M = 512, K = 512, N = 1025, ldb = 1032, NB = (N + BN - 1) / BN, MB = M / BM, KB = K / KC;
  std::vector<double> A((size_t)M * K), B((size_t)K * ldb, 0.0);
  srand(7);
  auto rnd = []() { return (rand() / (double)RAND_MAX) * 2.0 - 1.0; };
  for (int i = 0; i < M; i++) { const double sc = pow(10.0, -3.0 * (i % 5)); for (int k = 0; k < K; k++) A[(size_t)i * K + k] = rnd() * sc; }
  for (int k = 0; k < K; k++) for (int j = 0; j < N; j++) B[(size_t)k * ldb + j] = rnd() * pow(10.0, -2.0 * (j % 4)) * (1.0 + 0.001 * k);
  double *dA, *dB, *dC; int8_t *dAs, *dBs; int *deb, *derr; double *dea, *debs;
  const size_t nA = (size_t)M * K, nB = (size_t)K * ldb, nC = (size_t)M * ldb;
  CK(cudaMalloc(&dA, nA * 8 * batch)); CK(cudaMalloc(&dB, nB * 8 * batch)); CK(cudaMalloc(&dC, nC * 8 * batch));
  CK(cudaMalloc(&dAs, (size_t)batch * MB * KB * A_STAGE)); CK(cudaMalloc(&dBs, (size_t)batch * NB * KB * B_STAGE));
  CK(cudaMalloc(&dea, (size_t)batch * M * 8)); CK(cudaMalloc(&deb, (size_t)batch * N * 4)); CK(cudaMalloc(&debs, (size_t)batch * N * 8)); CK(cudaMalloc(&derr, 4));
  for (int z = 0; z < batch; z++) {
    CK(cudaMemcpy(dA + z * nA, A.data(), nA * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB + z * nB, B.data(), nB * 8, cudaMemcpyHostToDevice));
  }
  CK(cudaMemset(derr, 0, 4)); CK(cudaMemset(dC, 0, nC * 8 * batch));
  const size_t smem = (size_t)STAGES * (A_STAGE + B_STAGE);
  CK(cudaFuncSetAttribute(emu_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1, e2; cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
  float ms_slice = 0, ms_gemm = 0;
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0);
    slice_rows<<<dim3(M / 8, batch), 256>>>(dA, K, nA, M, K, dAs, dea);
    col_exponents<<<dim3((N + 127) / 128, batch), 128>>>(dB, ldb, nB, K, N, deb, debs);
    slice_cols<<<dim3(NB, KB, batch), 256>>>(dB, ldb, nB, K, N, NB, deb, dBs);
    cudaEventRecord(e1);
    emu_gemm_kernel<<<dim3(NB, MB, batch), 256, smem>>>(dAs, dBs, dea, debs, dC, ldb, nC, M, N, K, derr);
    cudaEventRecord(e2);
    CK(cudaDeviceSynchronize());
    cudaEventElapsedTime(&ms_slice, e0, e1); cudaEventElapsedTime(&ms_gemm, e1, e2);
  }
  int herr = 0; CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost));
  printf("batch %d: slicing %.3f ms, emulated GEMM %.3f ms = %.1f FP64-equivalent TFLOP/s (with slicing %.1f); err flag %d\n", batch,
         ms_slice, ms_gemm, 2.0 * batch * M * N * K / ms_gemm * 1e-9, 2.0 * batch * M * N * K / (ms_gemm + ms_slice) * 1e-9, herr);
  std::vector<double> C(nC);
  CK(cudaMemcpy(C.data(), dC + (size_t)(batch - 1) * nC, nC * 8, cudaMemcpyDeviceToHost));
  double worst = 0, worst_norm = 0;
  for (int i = 0; i < M; i += 7)
    for (int j = 0; j < N; j += 3) {
      double s = 0, sa = 0;
      for (int k = 0; k < K; k++) { s += A[(size_t)i * K + k] * B[(size_t)k * ldb + j]; sa += fabs(A[(size_t)i * K + k] * B[(size_t)k * ldb + j]); }
      const double got = C[(size_t)i * ldb + j];
      worst = fmax(worst, fabs(got - s) / fmax(fabs(s), 1e-300));
      worst_norm = fmax(worst_norm, fabs(got - s) / sa);
    }
  printf("max elementwise rel err %.3e, max err relative to sum|a||b| %.3e\n", worst, worst_norm);
  return herr ? 2 : 0;
}
