// Measured ceilings for the roofline of the INT8-emulated FP64 GEMM (bench.py prints them next to the achieved rate):
// pfml_int8_mma_probe times back-to-back tcgen05.mma.kind::i8 instructions (M = 128, N = n, K = 32) on every SM, operands
// resident in shared memory (or A in TMEM), accumulators in TMEM, no data movement -- the issue rate of the tensor pipe
// for that instruction shape, i.e. the speed-of-light of any kernel built from it.
#include "kernels.cuh"
#include "pfml_b200.h"

using namespace pfml;

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int PM = 128, PK = 32;          // rows, bytes of k per instruction
constexpr int P_LBO = 128, P_SBO = 256;   // canonical no-swizzle K-major layout of a 32-byte k slab

__device__ __forceinline__ bool probe_elect() {     // one lane of a converged warp (ptxas then emits straight-line UTCIMMA issue)
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, %1;\n\t@px mov.s32 %0, 1;\n\t}\n" : "+r"(pred) : "r"(0xFFFFFFFFu));
  return pred != 0;
}

__device__ __forceinline__ uint64_t probe_desc(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((P_LBO >> 4) & 0x3FFF) << 16) | ((uint64_t)((P_SBO >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46);
}

// n_acc accumulators of n columns are cycled so that consecutive instructions do not serialise on one accumulator;
// n_a / n_b distinct operand slabs are cycled so that the operand reads are real shared-memory traffic
__global__ void __launch_bounds__(128, 1) int8_probe_kernel(int n, int iters, int ts, int n_acc, int* __restrict__ ok) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  constexpr int NA = 4, NB = 4;
  uint8_t* sA = smem;                       // NA slabs of 128 x 32 B
  uint8_t* sB = smem + NA * PM * PK;        // NB slabs of n x 32 B
  for (int e = tid; e < NA * PM * PK + NB * n * PK; e += blockDim.x) smem[e] = (uint8_t)((e * 37 + 11) & 3);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tb = tmem_base;
  if (warp == 1) {
   if (probe_elect()) {
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(PM >> 4) << 24);
    const uint32_t a_tmem = tb + n_acc * n;                 // TS form: A slab (8 columns) behind the accumulators
    if (ts) {
      asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(a_tmem), "l"(probe_desc(smem_u32(sA))) : "memory");
    }
    // descriptors and accumulator addresses are loop-invariant: the loop body is 8 instructions and nothing else
    uint64_t da[4], db[4];
    uint32_t dacc[8];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      da[u] = probe_desc(smem_u32(sA + u * PM * PK));
      db[u] = probe_desc(smem_u32(sB + u * n * PK));
    }
#pragma unroll
    for (int u = 0; u < 8; u++) dacc[u] = tb + (u % n_acc) * n;
    for (int it = 0; it < iters; it += 8) {
      const uint32_t acc = it > 0 ? 1u : 0u;
#pragma unroll
      for (int u = 0; u < 8; u++) {
        if (ts)
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}\n"
                       ::"r"(dacc[u]), "r"(a_tmem), "l"(db[u & 3]), "r"(idesc), "r"(acc) : "memory");
        else
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
                       ::"r"(dacc[u]), "l"(da[(u + (u >> 2)) & 3]), "l"(db[u & 3]), "r"(idesc), "r"(acc) : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t done = 0;
    for (long spin = 0; spin < (1L << 30) && !done; spin++)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    if (!done) atomicExch(ok, 0);
   }
   __syncwarp();
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tb));
}


// ---- pipeline probe: one 32-byte k slab of the emulated GEMM per iteration (7 accumulators of 64 columns, 28 digit products) ----
//   mode 0  the production mix: the MMA lane copies A digits 0..2 to TMEM with tcgen05.cp, then 18 products with A in TMEM
//           and 10 with A in shared memory
//   mode 1  four stager warps (one per TMEM lane quarter) move A digits 0..3 from shared memory to TMEM with tcgen05.st,
//           double-buffered and one slab ahead of the MMA lane: 22 products with A in TMEM, 6 from shared memory, no copy
//           in the tensor queue
__device__ __forceinline__ bool probe_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (long spin = 0; spin < (1L << 28) && !done; spin++)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done != 0;
}

__global__ void __launch_bounds__(256, 1) int8_pipe_probe_kernel(int iters, int mode, int* __restrict__ ok) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t a_ready[2], a_free[2], done_bar;
  __shared__ uint32_t tmem_base;
  constexpr int ND = 7, NB = 64;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint8_t* sA = smem;                          // ND digit slabs of 128 x 32 B
  uint8_t* sB = smem + ND * PM * PK;           // ND digit slabs of 64 x 32 B
  for (int e = tid; e < ND * (PM + NB) * PK; e += blockDim.x) smem[e] = (uint8_t)((e * 37 + 11) & 3);
  if (tid == 0) {
    for (int b = 0; b < 2; b++) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 4;" ::"r"(smem_u32(&a_ready[b])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&a_free[b])));
    }
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&done_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tb = tmem_base;
  const uint32_t abase = tb + ND * NB;           // A region: 64 columns behind the accumulators
  const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NB >> 3) << 17) | ((uint32_t)(PM >> 4) << 24);
  if (warp == 1) {
    if (probe_elect()) {
      uint64_t da[ND], db[ND];
#pragma unroll
      for (int i = 0; i < ND; i++) { da[i] = probe_desc(smem_u32(sA + i * PM * PK)); db[i] = probe_desc(smem_u32(sB + i * NB * PK)); }
      for (int it = 0; it < iters; it++) {
        const uint32_t acc = it > 0 ? 1u : 0u;
        const int buf = it & 1;
        constexpr int NTS0 = 3, NTS1 = 4;
        if (mode == 0) {
#pragma unroll
          for (int i = 0; i < NTS0; i++)
            asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(abase + i * 8), "l"(da[i]) : "memory");
#pragma unroll
          for (int i = 0; i < ND; i++)
#pragma unroll
            for (int j = 0; j < ND - i; j++) {
              const uint32_t a2 = (i > 0) ? 1u : acc;
              if (i < NTS0)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}\n"
                             ::"r"(tb + (i + j) * NB), "r"(abase + i * 8), "l"(db[j]), "r"(idesc), "r"(a2) : "memory");
              else
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
                             ::"r"(tb + (i + j) * NB), "l"(da[i]), "l"(db[j]), "r"(idesc), "r"(a2) : "memory");
            }
        } else {
          if (!probe_wait(smem_u32(&a_ready[buf]), (uint32_t)((it >> 1) & 1))) { atomicExch(ok, 0); break; }
          asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll
          for (int i = 0; i < ND; i++)
#pragma unroll
            for (int j = 0; j < ND - i; j++) {
              const uint32_t a2 = (i > 0) ? 1u : acc;
              if (i < NTS1)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}\n"
                             ::"r"(tb + (i + j) * NB), "r"(abase + buf * 32 + i * 8), "l"(db[j]), "r"(idesc), "r"(a2) : "memory");
              else
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
                             ::"r"(tb + (i + j) * NB), "l"(da[i]), "l"(db[j]), "r"(idesc), "r"(a2) : "memory");
            }
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&a_free[buf])) : "memory");
        }
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&done_bar)) : "memory");
      if (!probe_wait(smem_u32(&done_bar), 0u)) atomicExch(ok, 0);
    }
    __syncwarp();
  } else if (warp >= 4 && mode == 1) {
    // stagers: warp 4 + q owns TMEM lanes 32 q .. 32 q + 31 (row = lane quarter * 32 + lane)
    const int q = warp & 3, row = q * 32 + lane;
    for (int it = 0; it < iters; it++) {
      const int buf = it & 1;
      if (!probe_wait(smem_u32(&a_free[buf]), (uint32_t)(((it >> 1) & 1) ^ 1))) { atomicExch(ok, 0); break; }
      asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll
      for (int i = 0; i < 4; i++) {
        // row `row` of digit slab i in the canonical layout: two 16-byte chunks (k 0..15, k 16..31)
        const uint8_t* p = sA + i * PM * PK + (row & 7) * 16 + (row >> 3) * P_SBO;
        const uint4 lo = *reinterpret_cast<const uint4*>(p), hi = *reinterpret_cast<const uint4*>(p + P_LBO);
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                     ::"r"(abase + ((uint32_t)(q * 32) << 16) + buf * 32 + i * 8), "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w), "r"(hi.x), "r"(hi.y),
                       "r"(hi.z), "r"(hi.w) : "memory");
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      asm volatile("tcgen05.fence::before_thread_sync;");
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&a_ready[buf])) : "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tb));
}

}  // namespace

extern "C" int pfml_int8_mma_probe(int n, int iters, int ts_form, double* tops, double* clk_per_mma) {
  if (!tops || n < 16 || n > 256 || (n % 16) || iters < 16 || (iters % 8)) {
    set_error("pfml_int8_mma_probe: need 16 <= n <= 256, n %% 16 == 0, iters >= 16");
    return -1;
  }
  int dev = 0, n_sm = 0, khz = 0;
  PFML_CUDA_OK(cudaGetDevice(&dev));
  PFML_CUDA_OK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  PFML_CUDA_OK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  // accumulators cycled by consecutive instructions (an instruction that accumulates into a tile still in flight waits for
  // it): as many as fit beside the TS-form A slab, at most 8; ts_form bits 8.. override the count (dependency-latency probe)
  int n_acc = (ts_form & 1) ? 496 / n : 512 / n;
  if (n_acc > 8) n_acc = 8;
  if ((ts_form >> 8) > 0 && (ts_form >> 8) <= n_acc) n_acc = ts_form >> 8;
  if (n_acc < 1) n_acc = 1;
  ts_form &= 1;
  const size_t smem = 4 * PM * PK + 4 * (size_t)n * PK;
  int* ok = nullptr;
  PFML_CUDA_OK(cudaMalloc(&ok, sizeof(int)));
  const int one = 1;
  PFML_CUDA_OK(cudaMemcpy(ok, &one, sizeof(int), cudaMemcpyHostToDevice));
  PFML_CUDA_OK(cudaFuncSetAttribute(int8_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  PFML_CUDA_OK(cudaEventCreate(&e0));
  PFML_CUDA_OK(cudaEventCreate(&e1));
  int8_probe_kernel<<<n_sm, 128, smem>>>(n, iters, ts_form, n_acc, ok);        // warm-up
  float best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    PFML_CUDA_OK(cudaEventRecord(e0));
    int8_probe_kernel<<<n_sm, 128, smem>>>(n, iters, ts_form, n_acc, ok);
    PFML_CUDA_OK(cudaEventRecord(e1));
    PFML_CUDA_OK(cudaEventSynchronize(e1));
    float ms = 0.f;
    PFML_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    best = ms < best ? ms : best;
  }
  PFML_CHECK_LAUNCH();
  int okh = 0;
  PFML_CUDA_OK(cudaMemcpy(&okh, ok, sizeof(int), cudaMemcpyDeviceToHost));
  cudaFree(ok);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (!okh) {
    set_error("pfml_int8_mma_probe: barrier wait timed out");
    return -102;
  }
  const double ops = 2.0 * PM * n * PK * (double)iters * n_sm;
  *tops = ops / (best * 1e-3) * 1e-12;
  if (clk_per_mma) *clk_per_mma = best * 1e-3 * (khz * 1e3) / iters;        // at the nominal maximum SM clock
  return 0;
}


// clk (nominal) per 32-byte k slab of the emulated GEMM's instruction stream: mode 0 = production mix, mode 1 = A digits staged
// by tcgen05.st from four stager warps (see the kernel)
extern "C" int pfml_int8_pipe_probe(int iters, int mode, double* clk_per_slab, double* tops) {
  if (!clk_per_slab || iters < 8 || (mode != 0 && mode != 1)) {
    set_error("pfml_int8_pipe_probe: bad argument");
    return -1;
  }
  int dev = 0, n_sm = 0, khz = 0;
  PFML_CUDA_OK(cudaGetDevice(&dev));
  PFML_CUDA_OK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  PFML_CUDA_OK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  const size_t smem = 7 * (size_t)(PM + 64) * PK;
  int* ok = nullptr;
  PFML_CUDA_OK(cudaMalloc(&ok, sizeof(int)));
  const int one = 1;
  PFML_CUDA_OK(cudaMemcpy(ok, &one, sizeof(int), cudaMemcpyHostToDevice));
  PFML_CUDA_OK(cudaFuncSetAttribute(int8_pipe_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  PFML_CUDA_OK(cudaEventCreate(&e0));
  PFML_CUDA_OK(cudaEventCreate(&e1));
  int8_pipe_probe_kernel<<<n_sm, 256, smem>>>(iters, mode, ok);
  float best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    PFML_CUDA_OK(cudaEventRecord(e0));
    int8_pipe_probe_kernel<<<n_sm, 256, smem>>>(iters, mode, ok);
    PFML_CUDA_OK(cudaEventRecord(e1));
    PFML_CUDA_OK(cudaEventSynchronize(e1));
    float ms = 0.f;
    PFML_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    best = ms < best ? ms : best;
  }
  PFML_CHECK_LAUNCH();
  int okh = 0;
  PFML_CUDA_OK(cudaMemcpy(&okh, ok, sizeof(int), cudaMemcpyDeviceToHost));
  cudaFree(ok);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (!okh) {
    set_error("pfml_int8_pipe_probe: barrier wait timed out");
    return -102;
  }
  *clk_per_slab = best * 1e-3 * (khz * 1e3) / iters;
  if (tops) *tops = 28.0 * 2.0 * PM * 64 * PK * (double)iters * n_sm / (best * 1e-3) * 1e-12;
  return 0;
}
