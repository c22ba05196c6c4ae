// np_comm.cu — rank-to-rank plumbing over NCCL (NVLink 5 / NVSwitch): the halo exchange that replaces
// FabArray::FillBoundary's MPI point-to-point, and the allreduce / allgather that replace
// ParallelAllReduce ("distributed communication backend").
//
// NCCL is bound at run time (dlopen), not at link time: a host process that already carries an NCCL
// (PyTorch bundles its own libnccl.so.2) must keep exactly that one, and a single-GPU caller needs none.
#include "np_solver.h"
#include <dlfcn.h>

namespace {
// The slice of the NCCL ABI this library uses (nccl.h, stable since 2.7).
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;                       // ncclSuccess == 0
enum { ncclSuccess_ = 0 };
enum { ncclSum_ = 0, ncclMax_ = 2 };            // ncclRedOp_t
enum { ncclFloat64_ = 8 };                      // ncclDataType_t

struct NcclApi {
    void* so = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    bool ok = false;
};

NcclApi& nccl()
{
    static NcclApi api;
    if (api.so) return api;
    // 1) whatever NCCL the process already loaded (PyTorch's), 2) the default search path
    void* so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!so) so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!so) so = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!so) return api;
    api.so = so;
#define NP_SYM(field, name) api.field = (decltype(api.field))dlsym(so, name)
    NP_SYM(GetUniqueId, "ncclGetUniqueId"); NP_SYM(CommInitRank, "ncclCommInitRank"); NP_SYM(CommDestroy, "ncclCommDestroy");
    NP_SYM(GroupStart, "ncclGroupStart"); NP_SYM(GroupEnd, "ncclGroupEnd"); NP_SYM(Send, "ncclSend"); NP_SYM(Recv, "ncclRecv");
    NP_SYM(AllReduce, "ncclAllReduce"); NP_SYM(AllGather, "ncclAllGather");
#undef NP_SYM
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.GroupStart && api.GroupEnd && api.Send && api.Recv
          && api.AllReduce && api.AllGather;
    return api;
}

int nccl_sendrecv(void* ctx, const double* send_lo, long long nsend_lo, double* recv_lo, long long nrecv_lo,
                  const double* send_hi, long long nsend_hi, double* recv_hi, long long nrecv_hi,
                  int lo_rank, int hi_rank, void* stream)
{
    NcclApi& N = nccl();
    ncclComm_t comm = (ncclComm_t)ctx; cudaStream_t st = (cudaStream_t)stream;
    // Order matters when lo_rank == hi_rank (two ranks, periodic): NCCL matches the sends and receives of one
    // peer pair in issue order, so both sides send [down, up] and receive [from above, from below].
    ncclResult_t rc = N.GroupStart();
    if (lo_rank >= 0 && rc == ncclSuccess_) rc = N.Send(send_lo, (size_t)nsend_lo, ncclFloat64_, lo_rank, comm, st);
    if (hi_rank >= 0 && rc == ncclSuccess_) rc = N.Send(send_hi, (size_t)nsend_hi, ncclFloat64_, hi_rank, comm, st);
    if (hi_rank >= 0 && rc == ncclSuccess_) rc = N.Recv(recv_hi, (size_t)nrecv_hi, ncclFloat64_, hi_rank, comm, st);
    if (lo_rank >= 0 && rc == ncclSuccess_) rc = N.Recv(recv_lo, (size_t)nrecv_lo, ncclFloat64_, lo_rank, comm, st);
    ncclResult_t rc2 = N.GroupEnd();
    return (rc == ncclSuccess_ && rc2 == ncclSuccess_) ? 0 : 1;
}
int nccl_allreduce(void* ctx, double* buf, int n, int op, void* stream)
{
    return nccl().AllReduce(buf, buf, (size_t)n, ncclFloat64_, op == 1 ? ncclMax_ : ncclSum_, (ncclComm_t)ctx, (cudaStream_t)stream) == ncclSuccess_ ? 0 : 1;
}
int nccl_allgather(void* ctx, const double* send, double* recv, long long cnt, void* stream)
{
    return nccl().AllGather(send, recv, (size_t)cnt, ncclFloat64_, (ncclComm_t)ctx, (cudaStream_t)stream) == ncclSuccess_ ? 0 : 1;
}
} // namespace

void np_comm_fill_nccl(NpComm& c, void* nccl_comm)
{
    if (!nccl_comm || !nccl().ok) return;
    c.nccl = nccl_comm; c.ctx = nccl_comm;
    c.sendrecv = nccl_sendrecv; c.allreduce = nccl_allreduce; c.allgather = nccl_allgather;
}

extern "C" {
int np_nccl_unique_id(void* id128)
{
    if (!nccl().ok) return NP_ERR_NCCL;
    ncclUniqueId id;
    if (nccl().GetUniqueId(&id) != ncclSuccess_) return NP_ERR_NCCL;
    memcpy(id128, &id, 128);
    return NP_OK;
}
int np_nccl_comm_create(void** comm, const void* id128, int rank, int nranks)
{
    if (!nccl().ok) return NP_ERR_NCCL;
    ncclUniqueId id; memcpy(&id, id128, 128);
    ncclComm_t c;
    if (nccl().CommInitRank(&c, nranks, id, rank) != ncclSuccess_) return NP_ERR_NCCL;
    *comm = (void*)c;
    return NP_OK;
}
int np_nccl_comm_destroy(void* comm)
{
    if (!nccl().ok) return NP_ERR_NCCL;
    return nccl().CommDestroy((ncclComm_t)comm) == ncclSuccess_ ? NP_OK : NP_ERR_NCCL;
}
}
