// pfml_inverse_batched_static: the batched inverse for callers that are not allowed to allocate device memory.
//
// Who needs this: code that runs where cudaMalloc is off limits -- a host thread inside a latency-critical
// section (cudaMalloc takes driver locks and may synchronise the device), a plugin loaded into a process whose
// memory pool belongs to somebody else (an allocation behind the pool's back fragments it or breaks its
// accounting), or a caller that simply has no spare buffer of 3 x B x NP x NP doubles to hand over as the
// workspace that pfml_inverse_batched wants.  For them the three work matrices live in one array that is part
// of the library image itself.
//
// What that costs: the array below occupies 12 MiB on EVERY GPU that loads the library, from load until the
// process exits, whether the entry point is ever called or not.  That is why the limit is small.
//
// Rules of use
//   * Size limit: B * NP * NP <= PFML_INVERSE_STATIC_MAX_ELEMS (NP = pfml_padded_n(n)); e.g. 2 matrices of
//     n <= 512, 8 of n <= 256, 32 of n <= 128.  Larger requests fail with -3 and an explanatory message; nothing
//     is launched.
//   * One array per GPU, shared by every call: two calls that may overlap on the same GPU (different streams,
//     different host threads) race on it.  Serialise them yourself, e.g. keep them on one stream.
//   * The array keeps whatever the previous call left in it.  Nothing here relies on old contents or on zeros:
//     every element a call reads is written earlier in the same call (the padded copy of A, then W/V/S blocks
//     before their use).  The only thing cleared is the caller's `status` array (cudaMemsetAsync on the stream).
//   * The first call on each GPU looks up the device address of the array (once per device, with that device
//     current).  Make that first call before starting any stream capture on that GPU.
//   * No cudaMalloc / cudaFree / allocating library routine is called on this path.
#include <mutex>

#include "common.cuh"
#include "pfml_b200.h"

namespace {

constexpr size_t kMaxElems = PFML_INVERSE_STATIC_MAX_ELEMS;

// Work matrices Wk | X | T1 of the recursive block inverse, each kMaxElems doubles.  In HOST code the name
// `g_inverse_scratch` is only a handle for the runtime (it is not a device address and must never be passed to
// a kernel); cudaGetSymbolAddress below turns it into the real per-device pointer.
static __device__ __align__(256) double g_inverse_scratch[3 * kMaxElems];

constexpr int kMaxDevices = 64;
double* g_scratch_ptr[kMaxDevices] = {};
std::mutex g_setup_mu;

// Once per device, with that device current: resolve the device address of the static array.
int scratch_for_current_device(double** out) {
  int dev = -1;
  PFML_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    pfml::set_error("pfml_inverse_batched_static: device ordinal %d out of range", dev);
    return -1;
  }
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (!g_scratch_ptr[dev]) {
    double* p = nullptr;
    PFML_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_inverse_scratch));
    if (!p) {
      pfml::set_error("pfml_inverse_batched_static: cudaGetSymbolAddress returned a null address");
      return -100;
    }
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return 0;
}

}  // namespace

extern "C" int pfml_inverse_batched_static(int B, int n, const double* A, double* Ainv, int* status, void* stream) {
  if (B <= 0 || n <= 0 || !A || !Ainv) {
    pfml::set_error("pfml_inverse_batched_static: null/empty argument");
    return -1;
  }
  const size_t NP = (size_t)pfml_padded_n(n);
  const size_t need = (size_t)B * NP * NP;
  if (need > kMaxElems) {
    pfml::set_error("pfml_inverse_batched_static: B * NP * NP = %zu exceeds the static scratch limit of %zu doubles "
                    "(B = %d, n = %d, NP = %zu); use pfml_inverse_batched with a caller-provided workspace",
                    need, kMaxElems, B, n, NP);
    return -3;
  }
  double* scratch = nullptr;
  PFML_TRY(scratch_for_current_device(&scratch));
  // Same kernels as pfml_inverse_batched; the workspace is the library's own array instead of the caller's.
  return pfml_inverse_batched(B, n, A, Ainv, status, scratch, 3 * kMaxElems * sizeof(double), stream);
}
