// TEST INFRASTRUCTURE ONLY.  Communication backend of the emulation build: the three collectives are
// callbacks installed by the test (Python, over torch.distributed/gloo), so the N>1 host logic of the
// solver (slab halos, replicated coarse levels, reductions) runs on CPU with world_size 2.
#include "np_solver.h"

struct EmuCb {
    int (*sendrecv)(const double*, long long, double*, long long, const double*, long long, double*, long long, int, int);
    int (*allreduce)(double*, int, int);
    int (*allgather)(const double*, double*, long long);
};
static EmuCb g_cb = {nullptr, nullptr, nullptr};

static int e_sendrecv(void*, const double* a, long long na, double* b, long long nb, const double* c, long long nc, double* d, long long nd,
                      int lo, int hi, void*) { return g_cb.sendrecv(a, na, b, nb, c, nc, d, nd, lo, hi); }
static int e_allreduce(void*, double* buf, int n, int op, void*) { return g_cb.allreduce(buf, n, op); }
static int e_allgather(void*, const double* s, double* r, long long cnt, void*) { return g_cb.allgather(s, r, cnt); }

void np_comm_fill_nccl(NpComm& c, void*)
{
    if (!g_cb.sendrecv) return;
    c.sendrecv = e_sendrecv; c.allreduce = e_allreduce; c.allgather = e_allgather;
}

extern "C" {
void np_emu_set_comm(void* sendrecv, void* allreduce, void* allgather)
{
    g_cb.sendrecv = (decltype(g_cb.sendrecv))sendrecv;
    g_cb.allreduce = (decltype(g_cb.allreduce))allreduce;
    g_cb.allgather = (decltype(g_cb.allgather))allgather;
}
long long np_emu_launch_count() { return np_emu::st().launches; }
int np_nccl_unique_id(void*) { return NP_ERR_NCCL; }
int np_nccl_comm_create(void**, const void*, int, int) { return NP_ERR_NCCL; }
int np_nccl_comm_destroy(void*) { return NP_ERR_NCCL; }
}
