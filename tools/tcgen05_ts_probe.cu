// Probe: tcgen05.mma kind::i8 (M=128, N=64, K=32) with the A operand in TMEM (copied there from shared memory by
// tcgen05.cp 128x256b) and B in shared memory; accumulator in TMEM.  Checks the product against the CPU and dumps
// the TMEM image of A on a mismatch.   nvcc -gencode arch=compute_100a,code=sm_100a -o tcgen05_ts_probe ...
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int M = 128, N = 64, K = 32;
constexpr int LBO = 128, SBO = 256;     // bytes: k-chunk (16 B) stride, 8-row group stride

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((LBO >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((SBO >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;               // descriptor version for sm_100
  return d;                             // layout_type = 0 (no swizzle), base_offset = 0
}

__global__ void __launch_bounds__(128) probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B, int32_t* __restrict__ D, uint32_t* __restrict__ Aimg) {
  __shared__ __align__(1024) int8_t sA[M * K];
  __shared__ __align__(1024) int8_t sB[N * K];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // canonical layout: byte(r, k) = (r % 8) * 16 + (r / 8) * SBO + (k / 16) * LBO + k % 16
  for (int e = tid; e < M * K; e += 128) { int r = e / K, k = e % K; sA[(r % 8) * 16 + (r / 8) * SBO + (k / 16) * LBO + k % 16] = A[e]; }
  for (int e = tid; e < N * K; e += 128) { int r = e / K, k = e % K; sB[(r % 8) * 16 + (r / 8) * SBO + (k / 16) * LBO + k % 16] = B[e]; }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");      // generic-proxy smem writes -> visible to the async (tensor) proxy
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t taddr = tmem_base;
  if (tid == 0) {
    const uint64_t da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    (void)da;
    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(taddr + 64), "l"(make_desc(smem_u32(sA))));
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(taddr), "r"(taddr + 64), "l"(db), "r"(idesc), "r"(0u));
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)));
  }
  // wait for the MMA (bounded spin: a wrong descriptor must not hang the GPU)
  uint32_t done = 0;
  for (int it = 0; it < (1 << 22) && !done; it++) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u));
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  if (done) {
    uint32_t v[64];
    const uint32_t a = taddr + ((uint32_t)(warp * 32) << 16);
#pragma unroll
    for (int c = 0; c < 64; c += 8) {
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(v[c]), "=r"(v[c + 1]), "=r"(v[c + 2]), "=r"(v[c + 3]), "=r"(v[c + 4]), "=r"(v[c + 5]), "=r"(v[c + 6]), "=r"(v[c + 7])
                   : "r"(a + c));
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;");
    for (int c = 0; c < 64; c++) D[(warp * 32 + lane) * N + c] = (int32_t)v[c];
    uint32_t w[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "r"(a + 64));
    asm volatile("tcgen05.wait::ld.sync.aligned;");
    for (int c = 0; c < 8; c++) Aimg[(warp * 32 + lane) * 8 + c] = w[c];
  } else if (tid == 0) {
    D[0] = -123456789;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(taddr));
}

int main() {
  std::vector<int8_t> A(M * K), B(N * K);
  srand(1);
  for (auto& x : A) x = (int8_t)(rand() % 129 - 64);
  for (auto& x : B) x = (int8_t)(rand() % 129 - 64);
  int8_t *dA, *dB; int32_t* dD; uint32_t* dI; cudaMalloc(&dI, M * 32);
  cudaMalloc(&dA, A.size()); cudaMalloc(&dB, B.size()); cudaMalloc(&dD, M * N * 4);
  cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, M * N * 4);
  probe<<<1, 128>>>(dA, dB, dD, dI);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<int32_t> D(M * N);
  cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
  if (D[0] == -123456789) { printf("mbarrier wait timed out\n"); return 2; }
  int bad = 0;
  for (int i = 0; i < M; i++)
    for (int j = 0; j < N; j++) {
      int32_t s = 0;
      for (int k = 0; k < K; k++) s += (int32_t)A[i * K + k] * (int32_t)B[j * K + k];
      if (s != D[i * N + j]) { if (bad < 5) printf("mismatch (%d,%d): got %d want %d\n", i, j, D[i * N + j], s); bad++; }
    }
  std::vector<uint32_t> I(M * 8);
  cudaMemcpy(I.data(), dI, M * 32, cudaMemcpyDeviceToHost);
  int badimg = 0;
  for (int r = 0; r < M; r++)
    for (int k = 0; k < K; k++) {
      const int8_t got = (int8_t)((I[r * 8 + k / 4] >> (8 * (k % 4))) & 0xff);
      if (got != A[r * K + k]) { if (badimg < 8) printf("A image (%d,%d): got %d want %d\n", r, k, got, A[r * K + k]); badimg++; }
    }
  printf("tcgen05.cp image: %d mismatches of %d\n", badimg, M * K);
  printf("tcgen05 i8 TS probe: %d mismatches of %d\n", bad, M * N);
  return bad ? 3 : 0;
}
