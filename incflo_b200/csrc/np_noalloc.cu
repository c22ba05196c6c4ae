// np_noalloc.cu — np_apply_noalloc: the operator hook y = A x (K2, k_stencil) for callers that must not allocate device
// memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// np_apply stages x and y through two temporaries in the internal padded nodal layout, which it cudaMallocs and frees on
// every call.  Here those two temporaries live in ONE file-scope device array of fixed size (NP_NOALLOC_MAX_DOUBLES
// doubles = 8 MiB, held by every GPU that loads the library until the process exits).  The array's name is a host-side
// stand-in in host code, so its device address is looked up once per device with cudaGetSymbolAddress (that device
// current) and that pointer is what the kernels get.  Every call shares the array and finds the previous call's contents
// in it: both halves are cleared first (ghost nodes that no boundary fill writes must read as zero), and two calls that
// overlap on the GPU (different streams / threads / handles on one device) race.
#include "np_solver.h"
#include "np_kernels.cuh"
#include "../../include/nodalproj.h"

static __device__ __align__(256) double g_np_noalloc_scratch[NP_NOALLOC_MAX_DOUBLES];

namespace {
// device address of the scratch array on the current device; looked up once per device.  Returns null on failure.
double* noalloc_scratch_ptr()
{
#ifdef NP_EMU
    return g_np_noalloc_scratch;               // the emulation has one address space
#else
    enum { MAXDEV = 64 };
    static double* ptr[MAXDEV] = {nullptr};
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAXDEV) { cudaGetLastError(); return nullptr; }
    if (!ptr[dev]) {
        double* p = nullptr;
        if (cudaGetSymbolAddress((void**)&p, g_np_noalloc_scratch) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        ptr[dev] = p;
    }
    return ptr[dev];
#endif
}
} // namespace

extern "C" int np_apply_noalloc(np_handle* h, int l, const np_fab* x, np_fab* y)
{
    if (!h || l < 0 || l >= (int)h->S.lev.size() || !x || !y || !x->p || !y->p || x->ngrow < 0 || y->ngrow < 0) return NP_ERR_INVALID;
    NpSolver& S = h->S;
    if (!S.have_sigma) return NP_ERR_STATE;
    // device fabs only (a host fab would need a staging allocation), one rank or a replicated level (no communicator calls)
    if (!np_rt::is_device_ptr(x->p) || !np_rt::is_device_ptr(y->p)) return NP_ERR_INVALID;
    if (S.comm.nranks > 1 && !S.lev[l].replicated) return NP_ERR_INVALID;
    const LevelDev& d = S.lev[l].d;
    if (2 * d.ntot > (long long)NP_NOALLOC_MAX_DOUBLES) return NP_ERR_TOO_LARGE;
    double* base = noalloc_scratch_ptr();
    if (!base) return NP_ERR_CUDA;
    double* tx = base;
    double* ty = base + d.ntot;                          // ntot is even (rows are padded to 16 bytes): ty stays 16-byte aligned
    np_rt::memset_async(base, 0, (size_t)(2 * d.ntot) * sizeof(double), S.stream);
    FabView vx = S.view_nodal(*x, l), vy = S.view_nodal(*y, l);
    NP_LAUNCH(false, (k_copy_nodal<false>), grid_nodes(d), NP_BLK, 0, S.stream, d, tx, vx, 0);
    S.apply(l, ty, tx);                                  // ghost fill of tx + k_stencil; launches only
    NP_LAUNCH(false, (k_copy_nodal<true>), grid_nodes(d), NP_BLK, 0, S.stream, d, ty, vy, 0);
    np_rt::stream_sync(S.stream);
    return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
}
