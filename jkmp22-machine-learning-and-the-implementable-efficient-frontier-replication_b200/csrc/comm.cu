// The exchange steps of the path over NCCL (NVLink 5 / NVSwitch), behind the C ABI:
//
//   pfml_comm_allreduce_sums   FP64 sum all-reduce of the per-window partial sums  [n_buckets] x (P + P^2 + 1)
//                              (PFML_Search_Coef.py:77-88, :112-121: each rank summed its own dates)
//   pfml_ridge_grid_sharded    the ridge problems of PFML_Search_Coef.py:125-135 dealt over the ranks as contiguous,
//                              balanced ranges of the flat (year, lambda) index of every p, followed by an all-gather of
//                              beta (variable counts: one grouped ncclBroadcast per (rank, p) segment)
//
// These are the only collectives of the path: date-steps are independent (PFML_Input_Data.py:318 carries no state).  They
// move ~250 MB once per g for a 360-month panel (about a millisecond over NVSwitch) against seconds of date-steps, so
// they are plain NCCL calls on the caller's stream; no compute kernel waits on them tile by tile.
//
// NCCL is not linked: the library is opened at run time, preferring the copy already loaded into the process (the one
// torch.distributed uses) so that one process never holds two NCCL versions.  Only the types come from <nccl.h>.
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>
#include <mutex>

#include "kernels.cuh"
#include "pfml_b200.h"

using namespace pfml;

namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  const char* (*GetErrorString)(ncclResult_t);
  ncclResult_t (*GetVersion)(int*);
  bool ok = false;
  char why[256] = "";
};

NcclApi g_nccl;
std::once_flag g_nccl_once;

const NcclApi& nccl() {
  std::call_once(g_nccl_once, [] {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);     // already in the process (torch's)?
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
      snprintf(g_nccl.why, sizeof(g_nccl.why), "cannot open libnccl.so.2: %s", dlerror());
      return;
    }
#define SYM(field, name)                                                                  \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));                \
  if (!g_nccl.field) { snprintf(g_nccl.why, sizeof(g_nccl.why), "libnccl lacks %s", name); return; }
    SYM(GetUniqueId, "ncclGetUniqueId")
    SYM(CommInitRank, "ncclCommInitRank")
    SYM(CommDestroy, "ncclCommDestroy")
    SYM(AllReduce, "ncclAllReduce")
    SYM(Broadcast, "ncclBroadcast")
    SYM(GroupStart, "ncclGroupStart")
    SYM(GroupEnd, "ncclGroupEnd")
    SYM(GetErrorString, "ncclGetErrorString")
    SYM(GetVersion, "ncclGetVersion")
#undef SYM
    g_nccl.ok = true;
  });
  return g_nccl;
}

struct Comm {
  ncclComm_t nc = nullptr;
  int rank = 0, world = 1, device = 0;
};

#define PFML_NCCL_OK(expr)                                                                          \
  do {                                                                                              \
    ncclResult_t _r = (expr);                                                                       \
    if (_r != ncclSuccess) {                                                                        \
      set_error("%s failed: %s (%s:%d)", #expr, nccl().GetErrorString(_r), __FILE__, __LINE__);     \
      return -110;                                                                                  \
    }                                                                                               \
  } while (0)

int need_nccl(const char* fn) {
  if (!nccl().ok) {
    set_error("%s: NCCL unavailable (%s)", fn, nccl().why);
    return -111;
  }
  return 0;
}

}  // namespace

extern "C" {

int pfml_comm_unique_id(void* id128) {
  if (!id128) { set_error("pfml_comm_unique_id: null argument"); return -1; }
  PFML_TRY(need_nccl("pfml_comm_unique_id"));
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  PFML_NCCL_OK(nccl().GetUniqueId(reinterpret_cast<ncclUniqueId*>(id128)));
  return 0;
}

int pfml_comm_init(const void* id128, int rank, int world, int device, void** comm) {
  if (!id128 || !comm || world < 1 || rank < 0 || rank >= world || device < 0) {
    set_error("pfml_comm_init: bad argument (rank %d of %d, device %d)", rank, world, device);
    return -1;
  }
  PFML_TRY(need_nccl("pfml_comm_init"));
  struct Restore {                     // NCCL binds the communicator to the current device; the caller's device is put back
    int prev = -1;
    Restore() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
    ~Restore() { if (prev >= 0) cudaSetDevice(prev); }
  } restore;
  PFML_CUDA_OK(cudaSetDevice(device));
  Comm* c = new Comm;
  c->rank = rank; c->world = world; c->device = device;
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclResult_t r = nccl().CommInitRank(&c->nc, world, id, rank);
  if (r != ncclSuccess) {
    set_error("ncclCommInitRank failed: %s", nccl().GetErrorString(r));
    delete c;
    return -110;
  }
  *comm = c;
  return 0;
}

int pfml_comm_rank(void* comm) { return comm ? static_cast<Comm*>(comm)->rank : 0; }
int pfml_comm_world(void* comm) { return comm ? static_cast<Comm*>(comm)->world : 1; }

int pfml_comm_nccl_version(void) {
  int v = 0;
  if (nccl().ok && nccl().GetVersion(&v) == ncclSuccess) return v;
  return 0;
}

int pfml_comm_destroy(void* comm) {
  if (!comm) return 0;
  Comm* c = static_cast<Comm*>(comm);
  if (c->nc && nccl().ok) nccl().CommDestroy(c->nc);
  delete c;
  return 0;
}

int pfml_comm_allreduce_sums(void* comm, double* r_sum, double* d_sum, double* counts, int P, int n_buckets, void* stream) {
  if (!r_sum || !d_sum || P <= 0 || n_buckets <= 0) {
    set_error("pfml_comm_allreduce_sums: null/empty argument");
    return -1;
  }
  if (!comm || static_cast<Comm*>(comm)->world == 1) return 0;        // one rank: the partial sums are the sums
  PFML_TRY(need_nccl("pfml_comm_allreduce_sums"));
  Comm* c = static_cast<Comm*>(comm);
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(T_RIDGE, s);
  PFML_NCCL_OK(nccl().GroupStart());
  PFML_NCCL_OK(nccl().AllReduce(d_sum, d_sum, (size_t)n_buckets * P * P, ncclDouble, ncclSum, c->nc, s));
  PFML_NCCL_OK(nccl().AllReduce(r_sum, r_sum, (size_t)n_buckets * P, ncclDouble, ncclSum, c->nc, s));
  if (counts) PFML_NCCL_OK(nccl().AllReduce(counts, counts, (size_t)n_buckets, ncclDouble, ncclSum, c->nc, s));
  PFML_NCCL_OK(nccl().GroupEnd());
  return 0;
}

int pfml_ridge_grid_sharded(void* comm, const double* r_cum, const double* d_cum, const double* n_cum, int Y, int p_max,
                            const int* p_vec_host, int n_p, const double* l_vec, int n_l, double* beta, int* status, void* workspace,
                            size_t workspace_bytes, void* stream) {
  Comm* c = static_cast<Comm*>(comm);
  const int world = c ? c->world : 1, rank = c ? c->rank : 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int nprob = Y * n_l;
  int lo = 0, hi = nprob;
  pfml_ridge_shard_range(nprob, rank, world, &lo, &hi);
  PFML_TRY(ridge_grid_range(r_cum, d_cum, n_cum, Y, p_max, p_vec_host, n_p, l_vec, n_l, beta, status, workspace, workspace_bytes, s,
                            lo, hi));
  if (world == 1) return 0;
  PFML_TRY(need_nccl("pfml_ridge_grid_sharded"));
  // all-gather with variable counts: every (rank, p) segment of beta (and of status) is broadcast from its owner
  ProfScope ps(T_RIDGE, s);
  PFML_NCCL_OK(nccl().GroupStart());
  long off = 0;
  for (int ip = 0; ip < n_p; ip++) {
    const int q = p_vec_host[ip] + 1;
    for (int r = 0; r < world; r++) {
      int rlo, rhi;
      pfml_ridge_shard_range(nprob, r, world, &rlo, &rhi);
      if (rhi <= rlo) continue;
      double* seg = beta + off + (long)rlo * q;
      PFML_NCCL_OK(nccl().Broadcast(seg, seg, (size_t)(rhi - rlo) * q, ncclDouble, r, c->nc, s));
      if (status) {
        int* st = status + (long)ip * nprob + rlo;
        PFML_NCCL_OK(nccl().Broadcast(st, st, (size_t)(rhi - rlo), ncclInt32, r, c->nc, s));
      }
    }
    off += (long)nprob * q;
  }
  PFML_NCCL_OK(nccl().GroupEnd());
  return 0;
}

}  // extern "C"
