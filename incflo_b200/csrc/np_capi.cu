// np_capi.cu — the extern "C" boundary declared in include/nodalproj.h.
#include "np_solver.h"
#include "np_kernels.cuh"
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

// ------------------------------------------------------------------------------------------------
// Handle cache.  The reference builds a fresh NodalProjector inside every ApplyProjection
// (/root/reference/src/projection/incflo_apply_nodal_projection.cpp:134, 148-151), i.e. np_create / np_destroy run every
// time step.  np_create allocates ~9 GB at 512^3, opens CUDA-IPC mappings and all-gathers handles; none of that
// depends on anything but (geometry, BoxArray, DistributionMapping, rank layout, communicator, options).  np_destroy
// therefore parks the handle (one slot per process) and np_create hands it out again when the key matches exactly; the
// handle is reset to the state a fresh one has, so the API stays stateless.  A different key evicts (really destroys)
// the parked handle.  np_cache_clear() / NP_HANDLE_CACHE=0 release / disable it.  With several ranks create / destroy /
// clear are collective, as they already were (peer-memory mappings of the neighbours stay valid only while every rank
// keeps its handle).
// ------------------------------------------------------------------------------------------------
namespace {
struct NpCacheKey {
    np_geom g; std::vector<np_box> boxes; std::vector<int> owner; int rank, nranks; void* comm; np_options o; int device;
    // field by field: the structs have padding holes that a caller's stack copy leaves undefined
    static bool same(const np_geom& a, const np_geom& b) {
        for (int d = 0; d < 3; ++d)
            if (a.ncell[d] != b.ncell[d] || a.prob_lo[d] != b.prob_lo[d] || a.prob_hi[d] != b.prob_hi[d] || a.periodic[d] != b.periodic[d]) return false;
        return true;
    }
    static bool same(const np_options& a, const np_options& b) {
        return a.max_coarsening_level == b.max_coarsening_level && a.maxiter == b.maxiter && a.bottom_maxiter == b.bottom_maxiter
            && a.num_pre_smooth == b.num_pre_smooth && a.num_post_smooth == b.num_post_smooth && a.num_final_smooth == b.num_final_smooth
            && a.bottom_solver == b.bottom_solver && a.verbose == b.verbose && a.bottom_verbose == b.bottom_verbose
            && a.bottom_rtol == b.bottom_rtol && a.bottom_atol == b.bottom_atol && a.smooth_zchunk == b.smooth_zchunk
            && a.interp_sigma_weighted == b.interp_sigma_weighted && a.bottom_normalize == b.bottom_normalize
            && a.agg_min_local_nz == b.agg_min_local_nz && a.use_graph == b.use_graph && a.const_sigma_fastpath == b.const_sigma_fastpath
            && a.profile_kernels == b.profile_kernels && a.overlap_halo == b.overlap_halo && a.p2p_halo == b.p2p_halo;
    }
    bool operator==(const NpCacheKey& b) const {
        if (!same(g, b.g) || !same(o, b.o) || boxes.size() != b.boxes.size() || owner != b.owner) return false;
        for (size_t q = 0; q < boxes.size(); ++q)
            for (int d = 0; d < 3; ++d) if (boxes[q].lo[d] != b.boxes[q].lo[d] || boxes[q].hi[d] != b.boxes[q].hi[d]) return false;
        return rank == b.rank && nranks == b.nranks && comm == b.comm && device == b.device;
    }
};
std::mutex g_cache_mu;
np_handle* g_cached = nullptr;
bool cache_enabled() { const char* e = std::getenv("NP_HANDLE_CACHE"); return !(e && e[0] == '0'); }
int current_device()
{
#ifdef NP_EMU
    return 0;
#else
    int d = -1; if (cudaGetDevice(&d) != cudaSuccess) { cudaGetLastError(); return -1; } return d;
#endif
}
void really_destroy(np_handle* h) { if (!h) return; h->S.destroy(); delete (NpCacheKey*)h->cache_key; delete h; }
} // namespace




extern "C" {

void np_default_options(np_options* o)
{
    o->max_coarsening_level = 30; o->maxiter = 100; o->bottom_maxiter = 100;
    o->num_pre_smooth = 2; o->num_post_smooth = 2; o->num_final_smooth = 8;
    o->bottom_solver = NP_BOTTOM_BICGCG; o->verbose = 0; o->bottom_verbose = 0;
    o->bottom_rtol = 1e-4; o->bottom_atol = -1.0;
    o->smooth_zchunk = 0; o->interp_sigma_weighted = 1; o->bottom_normalize = 1;
    o->agg_min_local_nz = 64; o->use_graph = 1; o->const_sigma_fastpath = 1; o->profile_kernels = 0; o->overlap_halo = 0; o->p2p_halo = 1;
}

int np_create(np_handle** h, const np_geom* geom, int nboxes_global, const np_box* all_boxes,
              const int* owner_rank, int my_rank, int nranks, void* nccl_comm_or_null, const np_options* opt)
{
    if (!h || !geom || !all_boxes || !owner_rank || nranks < 1 || my_rank < 0 || my_rank >= nranks) return NP_ERR_INVALID;
    np_options o; if (opt) o = *opt; else np_default_options(&o);
    NpCacheKey* key = new (std::nothrow) NpCacheKey();
    if (!key) return NP_ERR_INVALID;
    key->g = *geom;
    key->boxes.assign(all_boxes, all_boxes + nboxes_global); key->owner.assign(owner_rank, owner_rank + nboxes_global);
    key->rank = my_rank; key->nranks = nranks; key->comm = nccl_comm_or_null; key->o = o; key->device = current_device();
    if (cache_enabled()) {
        std::lock_guard<std::mutex> lk(g_cache_mu);
        if (g_cached && *(NpCacheKey*)g_cached->cache_key == *key) {
            np_handle* q = g_cached; g_cached = nullptr;
            delete key;
            q->S.reset_for_reuse();
            *h = q;
            return NP_OK;
        }
    }
    np_handle* p = new (std::nothrow) np_handle();
    if (!p) { delete key; return NP_ERR_INVALID; }
    p->cache_key = key;
    NpComm c;
    if (nranks > 1) {
        np_comm_fill_nccl(c, nccl_comm_or_null);
        if (!c.sendrecv) { delete key; delete p; return NP_ERR_NCCL; }
    }
    int rc = p->S.create(geom, nboxes_global, all_boxes, owner_rank, my_rank, nranks, c, &o);
    if (rc) { really_destroy(p); return rc; }
    *h = p;
    return NP_OK;
}

int np_set_bc(np_handle* h, const int lo[3], const int hi[3]) { return h ? h->S.set_bc(lo, hi) : NP_ERR_INVALID; }

int np_project(np_handle* h, int nlocal, np_fab* vel, const np_fab* sigma, double rtol, double atol, void* cuda_stream, np_stats* out)
{
    if (!h || nlocal != (int)h->S.lbox.size() || !vel || !sigma || !vel->p || !sigma->p) return NP_ERR_INVALID;
    h->S.stream = (np_stream_t)cuda_stream;
    int rc = nlocal == 1 ? h->S.project(vel, sigma, rtol, atol) : h->S.project_boxes(nlocal, vel, sigma, rtol, atol);
    if (out) *out = h->S.stats;
    return rc;
}

int np_apply_projection(np_handle* h, np_fab* vel, const np_fab* vel_old, const np_fab* density, np_fab* gp, np_fab* p,
                        const np_apply_args* args, double rtol, double atol, void* cuda_stream, np_stats* out)
{
    if (!h || !vel || !density || !gp || !p || !args || !vel->p || !density->p || !gp->p || !p->p) return NP_ERR_INVALID;
    if (h->S.lbox.size() != 1) return NP_ERR_INVALID;      // one box per rank only (several boxes: np_project)
    h->S.stream = (np_stream_t)cuda_stream;
    int rc = h->S.apply_projection(vel, vel_old, density, gp, p, args, rtol, atol);
    if (out) *out = h->S.stats;
    return rc;
}

int np_last_stats(np_handle* h, np_stats* out) { if (!h || !out) return NP_ERR_INVALID; *out = h->S.stats; return NP_OK; }

int np_get_phi(np_handle* h, int local_box, np_fab* out)
{
    if (!h || local_box < 0 || local_box >= (int)h->S.lbox.size() || !out) return NP_ERR_INVALID;
    if (!h->S.have_result) return NP_ERR_STATE;
    out->p = h->S.lbox.size() == 1 ? h->S.phi_out : h->S.mb_phi[local_box]; out->ngrow = 1; out->ncomp = 1;
    return NP_OK;
}
int np_get_gradphi(np_handle* h, int local_box, np_fab* out)
{
    if (!h || local_box < 0 || local_box >= (int)h->S.lbox.size() || !out) return NP_ERR_INVALID;
    if (!h->S.have_result) return NP_ERR_STATE;
    out->p = h->S.lbox.size() == 1 ? h->S.gphi : h->S.mb_gphi[local_box]; out->ngrow = 1; out->ncomp = 3;
    return NP_OK;
}

// copy internal results into a caller fab (host or device, any ngrow): interior (+ min(ngrow,1) ghosts).
// Host destination: a fab with ngrow == 1 has exactly the internal layout -> one device-to-host copy; any other
// ngrow is repacked on the device into the persistent staging buffer first.
static int copy_out(NpSolver& S, int local_box, np_fab* dst, const double* src_fab1, bool nodal, int ncomp)
{
    struct { int n[3]; } d;      // cells of the box the fab belongs to
    for (int q = 0; q < 3; ++q) d.n[q] = S.lbox[local_box].hi[q] - S.lbox[local_box].lo[q] + 1;
    FabView v; v.p = dst->p; v.ng = dst->ngrow;
    v.sy = d.n[0] + (nodal ? 1 : 0) + 2 * dst->ngrow; v.sz = v.sy * (d.n[1] + (nodal ? 1 : 0) + 2 * dst->ngrow);
    v.sn = v.sz * (d.n[2] + (nodal ? 1 : 0) + 2 * dst->ngrow);
    const long long nelem = v.sn * ncomp;
    bool host = false;
#ifndef NP_EMU
    host = !np_rt::is_device_ptr(dst->p);
#endif
    if (host && dst->ngrow == 1) {
        np_rt::copy_async(dst->p, src_fab1, (size_t)nelem * sizeof(double), S.stream);
        np_rt::stream_sync(S.stream);
        return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
    }
    double* dev = dst->p;
    if (host && !(dev = S.stage_get(2, nelem))) return NP_ERR_CUDA;
    v.p = dev;
    const int g = dst->ngrow > 0 ? 1 : 0;
    const int off = nodal ? 1 : 0;
    FabView sv; sv.p = const_cast<double*>(src_fab1); sv.ng = 1; sv.sy = d.n[0] + off + 2; sv.sz = sv.sy * (d.n[1] + off + 2); sv.sn = sv.sz * (d.n[2] + off + 2);
    NP_LAUNCH(false, k_copy_cell, dim3((d.n[0] + off + 2*g + 31) / 32, (d.n[1] + off + 2*g + 7) / 8, d.n[2] + off + 2*g), NP_BLK, 0, S.stream,
              d.n[0] + off, d.n[1] + off, v, sv, ncomp, g);
    if (host) np_rt::copy_async(dst->p, dev, (size_t)nelem * sizeof(double), S.stream);
    np_rt::stream_sync(S.stream);
    return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
}
int np_copy_phi(np_handle* h, int local_box, np_fab* dst)
{
    if (!h || local_box < 0 || local_box >= (int)h->S.lbox.size() || !dst || !dst->p) return NP_ERR_INVALID;
    if (!h->S.have_result) return NP_ERR_STATE;
    return copy_out(h->S, local_box, dst, h->S.lbox.size() == 1 ? h->S.phi_out : h->S.mb_phi[local_box], true, 1);
}
int np_copy_gradphi(np_handle* h, int local_box, np_fab* dst)
{
    if (!h || local_box < 0 || local_box >= (int)h->S.lbox.size() || !dst || !dst->p || dst->ncomp < 3) return NP_ERR_INVALID;
    if (!h->S.have_result) return NP_ERR_STATE;
    return copy_out(h->S, local_box, dst, h->S.lbox.size() == 1 ? h->S.gphi : h->S.mb_gphi[local_box], false, 3);
}

int np_is_device_pointer(const void* p)
{
#ifdef NP_EMU
    (void)p; return 0;
#else
    return np_rt::is_device_ptr(p);
#endif
}

void np_destroy(np_handle* h)
{
    if (!h) return;
    np_handle* evict = h;
    if (cache_enabled() && h->cache_key && !h->S.err) {
        np_rt::stream_sync(h->S.stream);
        std::lock_guard<std::mutex> lk(g_cache_mu);
        evict = g_cached; g_cached = h;
    }
    really_destroy(evict);
}

int np_cache_clear(void)
{
    np_handle* q;
    { std::lock_guard<std::mutex> lk(g_cache_mu); q = g_cached; g_cached = nullptr; }
    really_destroy(q);
    return NP_OK;
}

const char* np_strerror(int code)
{
    switch (code) {
    case NP_OK: return "ok";
    case NP_ERR_INVALID: return "nodalproj: invalid argument or unsupported box layout";
    case NP_ERR_NO_DEVICE: return "nodalproj: no CUDA device (there is no CPU fallback)";
    case NP_ERR_CUDA: return "nodalproj: CUDA error";
    case NP_ERR_NCCL: return "nodalproj: NCCL / communication error";
    case NP_ERR_NOT_CONVERGED: return "MLMG failed to converge";
    case NP_ERR_DIVERGED: return "MLMG failing so lets stop here";
    case NP_ERR_STATE: return "nodalproj: call order";
    case NP_ERR_TOO_LARGE: return "nodalproj: level too large for the fixed no-allocation scratch (np_apply_noalloc: 2 padded nodal arrays must fit NP_NOALLOC_MAX_DOUBLES = 2^20 doubles)";
    default: return "nodalproj: unknown error";
    }
}

int np_num_levels(np_handle* h) { return h ? (int)h->S.lev.size() : 0; }
int np_level_ncell(np_handle* h, int l, int n[3])
{
    if (!h || l < 0 || l >= (int)h->S.lev.size()) return NP_ERR_INVALID;
    for (int d = 0; d < 3; ++d) n[d] = h->S.lev[l].d.n[d];
    return NP_OK;
}

} // extern "C"
