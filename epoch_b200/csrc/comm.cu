// comm.cu -- inter-GPU plumbing: NCCL point-to-point over NVLink 5 / NVSwitch.
// Replaces the MPI_SENDRECV call sites of boundary.F90 (do_field_mpi_with_lengths
// :354-449, particle_periodic_bcs :976-1072) and partlist_sendrecv
// (housekeeping/partlist.F90:808-862).  libnccl is resolved at run time (dlopen)
// so that a single-GPU host does not need it and so that a process which already
// loaded torch's bundled NCCL shares that copy.
#include "ctx.cuh"
#include <dlfcn.h>
#include <cstring>
#include <cstdlib>

namespace eb {

// minimal NCCL ABI (stable since 2.7): opaque comm, 128-byte unique id
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclFloat64 = 8 };   // ncclDataType_t value of ncclDouble

struct Nccl {
  void* h = nullptr;
  ncclComm_t comm = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static Nccl* load_nccl(std::string& err) {
  static Nccl lib;
  static bool tried = false;
  if (!tried) {
    tried = true;
    const char* names[] = {getenv("EB200_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      if (!n) continue;
      lib.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (lib.h) break;
    }
    if (lib.h) {
#define SYM(field, name) *(void**)(&lib.field) = dlsym(lib.h, name)
      SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank");
      SYM(CommDestroy, "ncclCommDestroy"); SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv");
      SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd");
      SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    }
  }
  if (!lib.h || !lib.GetUniqueId || !lib.CommInitRank || !lib.Send || !lib.Recv || !lib.GroupStart || !lib.GroupEnd) {
    err = "libnccl.so.2 could not be loaded (set EB200_NCCL_LIB)";
    return nullptr;
  }
  return &lib;
}

int32_t nccl_unique_id(void* out128) {
  std::string err;
  Nccl* l = load_nccl(err);
  if (!l) return fail(nullptr, EB200_E_NCCL, err);
  ncclUniqueId id;
  if (l->GetUniqueId(&id) != 0) return fail(nullptr, EB200_E_NCCL, "ncclGetUniqueId failed");
  std::memcpy(out128, &id, 128);
  return 0;
}

int32_t comm_init(eb200_ctx* c, const void* idbytes) {
  if (c->nproc <= 1) return 0;
  if (!idbytes) return fail(c, EB200_E_ARG, "nccl_unique_id required when nproc > 1");
  std::string err;
  Nccl* l = load_nccl(err);
  if (!l) return fail(c, EB200_E_NCCL, err);
  c->nccl = new Nccl(*l);
  ncclUniqueId id;
  std::memcpy(&id, idbytes, 128);
  ncclResult_t r = c->nccl->CommInitRank(&c->nccl->comm, c->nproc, id, c->rank);
  if (r != 0) return fail(c, EB200_E_NCCL, std::string("ncclCommInitRank: ") + (l->GetErrorString ? l->GetErrorString(r) : "error"));
  // connect the links to the Cartesian neighbours now: NCCL sets a peer's channels up at the first send / recv
  // (a few tenths of a second per peer), which the first step would otherwise pay (1.0 s at N = 4, round 2)
  double* tmp = nullptr;
  EB_CUDA(c, cudaMalloc(&tmp, 32 * sizeof(double)));
  EB_CUDA(c, cudaMemsetAsync(tmp, 0, 32 * sizeof(double), c->stream));
  for (int d = 0; d < c->ndims; ++d) {
    if (c->nprocdir[d] <= 1) continue;
    const int32_t rc = comm_exchange2(c, tmp, c->nbr_lo[d], 1, tmp + 8, c->nbr_hi[d], 1, tmp + 16, c->nbr_hi[d], 1, tmp + 24, c->nbr_lo[d], 1);
    if (rc != 0) { cudaFree(tmp); return rc; }
  }
  cudaError_t e = cudaStreamSynchronize(c->stream);
  cudaFree(tmp);
  if (e != cudaSuccess) return fail(c, EB200_E_CUDA, std::string("NCCL warm-up exchange: ") + cudaGetErrorString(e));
  return 0;
}

void comm_destroy(eb200_ctx* c) {
  if (c->nccl) {
    if (c->nccl->comm && c->nccl->CommDestroy) c->nccl->CommDestroy(c->nccl->comm);
    delete c->nccl;
    c->nccl = nullptr;
  }
}

int32_t comm_sendrecv(eb200_ctx* c, const double* send, int dst, double* recv, int src, size_t count) {
  return comm_sendrecv_v(c, send, dst, count, recv, src, count);
}

int32_t comm_sendrecv_v(eb200_ctx* c, const double* send, int dst, size_t nsend, double* recv, int src, size_t nrecv) {
  if (dst < 0 && src < 0) return 0;
  if (!c->nccl) return fail(c, EB200_E_STATE, "NCCL communicator not initialised");
  Nccl* l = c->nccl;
  ncclResult_t r = l->GroupStart();
  if (r == 0 && dst >= 0) r = l->Send(send, nsend, ncclFloat64, dst, l->comm, c->stream);
  if (r == 0 && src >= 0) r = l->Recv(recv, nrecv, ncclFloat64, src, l->comm, c->stream);
  ncclResult_t r2 = l->GroupEnd();
  if (r != 0 || r2 != 0)
    return fail(c, EB200_E_NCCL, std::string("nccl send/recv: ") + (l->GetErrorString ? l->GetErrorString(r ? r : r2) : "error"));
  return 0;
}

int32_t comm_exchange2(eb200_ctx* c, const double* a, int da, size_t na, const double* b, int db, size_t nb,
                       double* ra, int sa, size_t nra, double* rb, int sb, size_t nrb) {
  if (da < 0 && db < 0 && sa < 0 && sb < 0) return 0;
  if (!c->nccl) return fail(c, EB200_E_STATE, "NCCL communicator not initialised");
  Nccl* l = c->nccl;
  // (sends and receives between one pair of ranks match in issue order: a's partner is ra of the neighbour)
  ncclResult_t r = l->GroupStart();
  if (r == 0 && da >= 0) r = l->Send(a, na, ncclFloat64, da, l->comm, c->stream);
  if (r == 0 && db >= 0) r = l->Send(b, nb, ncclFloat64, db, l->comm, c->stream);
  if (r == 0 && sa >= 0) r = l->Recv(ra, nra, ncclFloat64, sa, l->comm, c->stream);
  if (r == 0 && sb >= 0) r = l->Recv(rb, nrb, ncclFloat64, sb, l->comm, c->stream);
  ncclResult_t r2 = l->GroupEnd();
  if (r != 0 || r2 != 0)
    return fail(c, EB200_E_NCCL, std::string("nccl exchange: ") + (l->GetErrorString ? l->GetErrorString(r ? r : r2) : "error"));
  return 0;
}

// every rank sends cnt_s[r] doubles at send + off_s[r] to rank r and receives cnt_r[r] doubles at
// recv + off_r[r] from it, in one NCCL group (balance.F90's MPI_ALLTOALLW of redistribute_field_*);
// the part a rank keeps is a device-to-device copy
int32_t comm_alltoallv(eb200_ctx* c, const double* send, const size_t* off_s, const size_t* cnt_s, double* recv,
                       const size_t* off_r, const size_t* cnt_r) {
  const int me = c->rank;
  if (cnt_s[me] != cnt_r[me]) return fail(c, EB200_E_STATE, "alltoallv: self counts differ");
  if (cnt_s[me])
    EB_CUDA(c, cudaMemcpyAsync(recv + off_r[me], send + off_s[me], cnt_s[me] * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  if (c->nproc <= 1) return 0;
  if (!c->nccl) return fail(c, EB200_E_STATE, "NCCL communicator not initialised");
  Nccl* l = c->nccl;
  ncclResult_t r = l->GroupStart();
  for (int p = 0; p < c->nproc && r == 0; ++p) {
    if (p == me) continue;
    if (cnt_s[p]) r = l->Send(send + off_s[p], cnt_s[p], ncclFloat64, p, l->comm, c->stream);
    if (r == 0 && cnt_r[p]) r = l->Recv(recv + off_r[p], cnt_r[p], ncclFloat64, p, l->comm, c->stream);
  }
  ncclResult_t r2 = l->GroupEnd();
  if (r != 0 || r2 != 0)
    return fail(c, EB200_E_NCCL, std::string("nccl alltoallv: ") + (l->GetErrorString ? l->GetErrorString(r ? r : r2) : "error"));
  return 0;
}

}  // namespace eb
