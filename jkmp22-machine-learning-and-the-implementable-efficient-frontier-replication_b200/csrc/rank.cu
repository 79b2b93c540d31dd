// f4 -- cross-sectional percentile ranks of the characteristics (Prepare_Data.py:338-345, :368): within every month, each
// characteristic is replaced by pandas' rank(pct=True) -- average rank of ties divided by the number of non-missing
// values --, exact zeros are put back to 0, and (optionally) missing values are imputed with 0.5.
//
// Dense layout X [T][N][F] (NaN = missing value or absent row).  One CTA per (month, characteristic): the column is
// gathered into shared memory, sorted with its row indices by a bitonic network (NaN last), tie groups get
// (first + last) / 2 -- the value pandas' sum-of-ranks / group-size produces, both are exact in FP64 --, then one FP64
// division by the count: bit-identical to pandas.  Sort-bound integer / compare work on 4 bytes of index and 8 of key per
// element; the panel is read and written once (uncoalesced column gathers: 32 B sectors of which 8 B are used).
#include "kernels.cuh"
#include "pfml_b200.h"

using namespace pfml;

namespace {

constexpr int RANK_MAX_N = 4096;

__global__ void __launch_bounds__(512)
pct_rank_kernel(int N, int F, int NPOW, const double* __restrict__ X, double* __restrict__ out, int impute) {
  extern __shared__ double s_key[];                 // [NPOW] keys, then [NPOW] int indices
  int* s_idx = reinterpret_cast<int*>(s_key + NPOW);
  __shared__ int s_cnt;
  const int t = blockIdx.y, f = blockIdx.x, tid = threadIdx.x;
  const double* col = X + ((long)t * N) * F + f;
  double* ocol = out + ((long)t * N) * F + f;
  if (tid == 0) s_cnt = 0;
  __syncthreads();
  int local = 0;
  for (int i = tid; i < NPOW; i += blockDim.x) {
    double v = i < N ? col[(long)i * F] : __longlong_as_double(0x7ff8000000000000ll);
    const bool miss = v != v;
    s_key[i] = miss ? __longlong_as_double(0x7ff0000000000000ll) : v;       // missing sorts last (+inf key, flagged by index)
    s_idx[i] = miss ? -1 - i : i;
    local += miss ? 0 : 1;
  }
  atomicAdd(&s_cnt, local);
  __syncthreads();
  // bitonic sort ascending by (key, missing last): a real +inf sorts before the missing entries through the index sign
  for (int k = 2; k <= NPOW; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < NPOW; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const double a = s_key[i], b = s_key[l];
          const int ia = s_idx[i], ib = s_idx[l];
          const bool up = (i & k) == 0;
          // order: non-missing by key; missing (index < 0) after everything
          const bool a_gt_b = (ia < 0 && ib >= 0) || ((ia < 0) == (ib < 0) && a > b);
          const bool b_gt_a = (ib < 0 && ia >= 0) || ((ia < 0) == (ib < 0) && b > a);
          if (up ? a_gt_b : b_gt_a) { s_key[i] = b; s_key[l] = a; s_idx[i] = ib; s_idx[l] = ia; }
        }
      }
      __syncthreads();
    }
  const int cnt = s_cnt;
  for (int i = tid; i < NPOW; i += blockDim.x) {
    const int id = s_idx[i];
    if (id < 0) {
      const int row = -1 - id;
      if (row < N) ocol[(long)row * F] = impute ? 0.5 : __longlong_as_double(0x7ff8000000000000ll);
      continue;
    }
    const double v = s_key[i];
    if (v == 0.0) { ocol[(long)id * F] = 0.0; continue; }                      // exact zeros are restored (:343-345)
    int lo = i, hi = i;                                                         // tie group [lo, hi] (positions among the cnt sorted values)
    while (lo > 0 && s_key[lo - 1] == v) lo--;
    while (hi + 1 < cnt && s_key[hi + 1] == v) hi++;
    const double avg = ((double)(lo + 1) + (double)(hi + 1)) * 0.5;             // average rank, 1-based
    ocol[(long)id * F] = avg / (double)cnt;
  }
}

}  // namespace

extern "C" int pfml_percentile_rank(const double* X, int T, int N, int F, int impute_half, double* out, void* stream) {
  if (!X || !out || T <= 0 || N <= 0 || F <= 0) {
    set_error("pfml_percentile_rank: null/empty argument");
    return -1;
  }
  if (N > RANK_MAX_N) {
    set_error("pfml_percentile_rank: %d rows per month exceed the shared-memory sort (max %d)", N, RANK_MAX_N);
    return -2;
  }
  int npow = 2;
  while (npow < N) npow <<= 1;
  const size_t smem = (size_t)npow * (sizeof(double) + sizeof(int));
  PFML_CONFIG_ONCE(pct_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RANK_MAX_N * 12);
  cudaStream_t s = (cudaStream_t)stream;
  for (int t0 = 0; t0 < T; t0 += 65535) {
    const int nt = T - t0 < 65535 ? T - t0 : 65535;
    pct_rank_kernel<<<dim3(F, nt), npow >= 512 ? 512 : (npow < 32 ? 32 : npow), smem, s>>>(N, F, npow, X + (long)t0 * N * F,
                                                                                         out + (long)t0 * N * F, impute_half);
    PFML_CHECK_LAUNCH();
  }
  return 0;
}
