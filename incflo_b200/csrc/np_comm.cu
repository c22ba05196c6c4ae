// np_comm.cu — rank-to-rank plumbing over NCCL (NVLink 5 / NVSwitch): the halo exchange that replaces
// FabArray::FillBoundary's MPI point-to-point, and the allreduce / allgather that replace
// ParallelAllReduce ("distributed communication backend").
#include "np_solver.h"
#include <nccl.h>

static int nccl_sendrecv(void* ctx, const double* send_lo, long long nsend_lo, double* recv_lo, long long nrecv_lo,
                         const double* send_hi, long long nsend_hi, double* recv_hi, long long nrecv_hi,
                         int lo_rank, int hi_rank, void* stream)
{
    ncclComm_t comm = (ncclComm_t)ctx; cudaStream_t st = (cudaStream_t)stream;
    ncclResult_t rc = ncclGroupStart();
    if (lo_rank >= 0) {
        if (rc == ncclSuccess) rc = ncclSend(send_lo, (size_t)nsend_lo, ncclDouble, lo_rank, comm, st);
        if (rc == ncclSuccess) rc = ncclRecv(recv_lo, (size_t)nrecv_lo, ncclDouble, lo_rank, comm, st);
    }
    if (hi_rank >= 0) {
        if (rc == ncclSuccess) rc = ncclSend(send_hi, (size_t)nsend_hi, ncclDouble, hi_rank, comm, st);
        if (rc == ncclSuccess) rc = ncclRecv(recv_hi, (size_t)nrecv_hi, ncclDouble, hi_rank, comm, st);
    }
    ncclResult_t rc2 = ncclGroupEnd();
    return (rc == ncclSuccess && rc2 == ncclSuccess) ? 0 : 1;
}
static int nccl_allreduce(void* ctx, double* buf, int n, int op, void* stream)
{
    return ncclAllReduce(buf, buf, (size_t)n, ncclDouble, op == 1 ? ncclMax : ncclSum, (ncclComm_t)ctx, (cudaStream_t)stream) == ncclSuccess ? 0 : 1;
}
static int nccl_allgather(void* ctx, const double* send, double* recv, long long cnt, void* stream)
{
    return ncclAllGather(send, recv, (size_t)cnt, ncclDouble, (ncclComm_t)ctx, (cudaStream_t)stream) == ncclSuccess ? 0 : 1;
}

void np_comm_fill_nccl(NpComm& c, void* nccl_comm)
{
    if (!nccl_comm) return;
    c.nccl = nccl_comm; c.ctx = nccl_comm;
    c.sendrecv = nccl_sendrecv; c.allreduce = nccl_allreduce; c.allgather = nccl_allgather;
}

extern "C" {
int np_nccl_unique_id(void* id128)
{
    ncclUniqueId id;
    if (ncclGetUniqueId(&id) != ncclSuccess) return NP_ERR_NCCL;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    memcpy(id128, &id, 128);
    return NP_OK;
}
int np_nccl_comm_create(void** comm, const void* id128, int rank, int nranks)
{
    ncclUniqueId id; memcpy(&id, id128, 128);
    ncclComm_t c;
    if (ncclCommInitRank(&c, nranks, id, rank) != ncclSuccess) return NP_ERR_NCCL;
    *comm = (void*)c;
    return NP_OK;
}
int np_nccl_comm_destroy(void* comm) { return ncclCommDestroy((ncclComm_t)comm) == ncclSuccess ? NP_OK : NP_ERR_NCCL; }
}
