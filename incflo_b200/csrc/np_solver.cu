// np_solver.cu — host driver: hierarchy, V-cycle (amrex::MLMG::solve / mgVcycle / bottomSolve restated as
// launch sequences), projection (amrex::NodalProjector::project).  Reference call site:
// /root/reference/src/projection/incflo_apply_nodal_projection.cpp:148-166.
#include "np_solver.h"
#include "np_kernels.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>

static inline int shell_blocks(long long n) { long long b = (n + 255) / 256; return (int)std::min<long long>(std::max<long long>(b, 1), 4096); }

static int alloc_d(double** p, long long n) {
    void* q = nullptr;
    if (np_rt::dev_malloc(&q, (size_t)n * sizeof(double))) return NP_ERR_CUDA;
    *p = (double*)q;
    return 0;
}

// written planes per smoother z-chunk: chunks of about `target` planes, never a sliver
static int pick_zchunk(int nnodes_z, int target) {
    int nch = std::max(1, (nnodes_z + target / 2) / target);
    return (nnodes_z + nch - 1) / nch;
}
// Smoother z-chunk of a level.  Small levels (one wave of 2 x 148 CTAs holds them with chunks of <= 32 planes): as many
// chunks as fit the wave, down to 4 planes each, because a CTA marches through its planes one after the other and
// the chunk length is the launch latency.  Large levels: enough chunks to give the 148 SMs about four CTAs each,
// 8..32 planes per chunk.  The result is rounded up to an even number: chunk interfaces (frozen during a smooth) then
// coincide with coarse-grid planes, which measurably improves the V-cycle contraction.
// Part of the smoother's definition; the oracle restates it.
static int gs_zchunk(int nx_nodes, int ny_nodes, int nz_nodes) {
    const int ncols = ((nx_nodes + GS_WX - 1) / GS_WX) * ((ny_nodes + GS_WY - 1) / GS_WY);
    const int nch1 = std::min(296 / ncols, std::max(1, nz_nodes / 4));
    int zc = 0;
    if (nch1 >= 1 && (nz_nodes + nch1 - 1) / nch1 <= 32) zc = (nz_nodes + nch1 - 1) / nch1;
    else {
        const int nch_want = (592 + ncols - 1) / ncols;
        const int target = std::min(32, std::max(8, nz_nodes / nch_want));
        zc = pick_zchunk(nz_nodes, target);
    }
    return zc + (zc & 1);
}

static void fill_coef(LevelDev& d) {
    const double fx = d.dxinv[0]*d.dxinv[0]/36.0, fy = d.dxinv[1]*d.dxinv[1]/36.0, fz = d.dxinv[2]*d.dxinv[2]/36.0;
    d.c.fxyz = fx+fy+fz;
    d.c.fmx2y2z = -fx+2*fy+2*fz; d.c.f2xmy2z = 2*fx-fy+2*fz; d.c.f2x2ymz = 2*fx+2*fy-fz;
    d.c.f4xm2ym2z = 4*fx-2*fy-2*fz; d.c.fm2x4ym2z = -2*fx+4*fy-2*fz; d.c.fm2xm2y4z = -2*fx-2*fy+4*fz;
}

static void set_dims(LevelDev& d, int n0, int n1, int n2) {
    d.n[0] = n0; d.n[1] = n1; d.n[2] = n2;
    d.nsy = (n0 + 3 + 1) & ~1LL; d.nsz = d.nsy * (n1 + 3); d.ntot = d.nsz * (n2 + 3);
    d.csy = (n0 + 2 + 1) & ~1LL; d.csz = d.csy * (n1 + 2); d.ctot = d.csz * (n2 + 2);
}

void NpSolver::zero(double* p, long long n) { np_rt::memset_async(p, 0, (size_t)n * sizeof(double), stream); }

int NpSolver::create(const np_geom* g, int nboxes, const np_box* boxes, const int* owner, int my_rank, int nranks,
                     const NpComm& c, const np_options* o)
{
    if (np_rt::init()) return NP_ERR_NO_DEVICE;
    if (const char* e = std::getenv("NP_FUSED_RR")) fused_rr = std::atoi(e) != 0;        // experiment knob: 0 = residual and restriction as two kernels
    if (const char* e = std::getenv("NP_GS_SMALL")) gs_small = std::atoi(e);
    if (const char* e = std::getenv("NP_GS_SMALL_VARIANT")) gs_small_variant = std::atoi(e);
    if (const char* e = std::getenv("NP_INTERP_MARCH")) interp_march = std::atoi(e) != 0;
    if (const char* e = std::getenv("NP_GS_VARIANT")) gs_variant_c = std::atoi(e);        // experiment knobs: which smoother kernel
    if (const char* e = std::getenv("NP_GSV_VARIANT")) gs_variant_v = std::atoi(e);
    geom = *g; opt = *o; comm = c; comm.rank = my_rank; comm.nranks = nranks;
    // ---- box layout: the boxes of rank r tile the r-th of nranks equal z-slabs (one box per rank: used as it is;
    //      several: gathered into the slab by np_project, see project_boxes) ----
    if (nboxes < nranks) return NP_ERR_INVALID;
    const int nzl = g->ncell[2] / nranks;
    if (nzl * nranks != g->ncell[2]) return NP_ERR_INVALID;
    {
        std::vector<long long> vol(nranks, 0);
        for (int b = 0; b < nboxes; ++b) {
            const np_box& bx = boxes[b];
            const int r = owner[b];
            if (r < 0 || r >= nranks) return NP_ERR_INVALID;
            for (int d = 0; d < 3; ++d) if (bx.lo[d] > bx.hi[d]) return NP_ERR_INVALID;
            if (bx.lo[0] < 0 || bx.lo[1] < 0 || bx.hi[0] >= g->ncell[0] || bx.hi[1] >= g->ncell[1]) return NP_ERR_INVALID;
            if (bx.lo[2] < r * nzl || bx.hi[2] >= (r + 1) * nzl) return NP_ERR_INVALID;       // inside its owner's slab
            vol[r] += (long long)(bx.hi[0] - bx.lo[0] + 1) * (bx.hi[1] - bx.lo[1] + 1) * (bx.hi[2] - bx.lo[2] + 1);
        }
        // inside the slab + total volume == slab volume + pairwise disjoint (checked for this rank's boxes) <=> exact tiling
        for (int r = 0; r < nranks; ++r) if (vol[r] != (long long)g->ncell[0] * g->ncell[1] * nzl) return NP_ERR_INVALID;
        lbox.clear();
        for (int b = 0; b < nboxes; ++b) if (owner[b] == my_rank) {
            np_box q = boxes[b]; q.lo[2] -= my_rank * nzl; q.hi[2] -= my_rank * nzl; lbox.push_back(q);
        }
        for (size_t a = 0; a < lbox.size(); ++a) for (size_t b = a + 1; b < lbox.size(); ++b) {
            bool ov = true;
            for (int d = 0; d < 3; ++d) if (lbox[a].hi[d] < lbox[b].lo[d] || lbox[b].hi[d] < lbox[a].lo[d]) ov = false;
            if (ov) return NP_ERR_INVALID;
        }
    }
    nloc[0] = g->ncell[0]; nloc[1] = g->ncell[1]; nloc[2] = nzl; koff = my_rank * nzl;
    for (int d = 0; d < 3; ++d) { bclo[d] = bchi[d] = g->periodic[d] ? NP_BC_PERIODIC : NP_BC_NEUMANN; }

    // ---- MG levels (MLLinOp::defineGrids: coarsen by 2 while every box stays coarsenable, min width 2) ----
    lev.clear();
    int n[3] = {nloc[0], nloc[1], nloc[2]};
    int nglob2 = g->ncell[2];
    double dxinv[3];
    for (int d = 0; d < 3; ++d) dxinv[d] = g->ncell[d] / (g->prob_hi[d] - g->prob_lo[d]);
    int ko = koff;
    bool replicated = false;
    for (;;) {
        NpLevel L;
        set_dims(L.d, n[0], n[1], n[2]);
        for (int d = 0; d < 3; ++d) L.d.dxinv[d] = dxinv[d];
        fill_coef(L.d);
        L.d.const_sigma = 0; L.d.sigma0 = 1.0;
        L.koff = ko; L.nglob2 = nglob2; L.replicated = replicated;
        L.zc = opt.smooth_zchunk > 0 ? opt.smooth_zchunk : gs_zchunk(n[0] + 1, n[1] + 1, n[2] + 1);
        lev.push_back(L);
        if ((int)lev.size() > opt.max_coarsening_level) break;
        // multi-GPU: below agg_min_local_nz the next level is replicated on every rank (whole domain)
        int nn[3] = {n[0], n[1], n[2]};
        bool to_replica = false;
        if (nranks > 1 && !replicated && n[2] / 2 < opt.agg_min_local_nz) { to_replica = true; nn[2] = nglob2; }
        bool ok = true;
        for (int d = 0; d < 3; ++d) if (nn[d] % 2 || nn[d] < 4) ok = false;
        if (!replicated && nranks > 1 && !to_replica && (n[2] % 2 || n[2] < 4)) ok = false;
        if (!ok) break;
        if (to_replica) { replicated = true; agg_lev = (int)lev.size(); ko = 0; }
        n[0] = nn[0] / 2; n[1] = nn[1] / 2; n[2] = nn[2] / 2; nglob2 /= 2; if (!to_replica) ko /= 2;
        for (int d = 0; d < 3; ++d) dxinv[d] *= 0.5;
    }
    // ---- allocations ----
    if (np_rt::last_error()) {}
    for (auto& L : lev) {
        if (alloc_d(&L.sigma, L.d.ctot) || alloc_d(&L.res, L.d.ntot) || alloc_d(&L.cor, L.d.ntot)
            || alloc_d(&L.cor_alt, L.d.ntot) || alloc_d(&L.rescor, L.d.ntot)) return NP_ERR_CUDA;
        zero(L.res, L.d.ntot); zero(L.cor, L.d.ntot); zero(L.cor_alt, L.d.ntot); zero(L.rescor, L.d.ntot);
        zero(L.sigma, L.d.ctot);
    }
    const LevelDev& d0 = lev[0].d;
    if (alloc_d(&sol, d0.ntot) || alloc_d(&rhs, d0.ntot)) return NP_ERR_CUDA;
    zero(sol, d0.ntot); zero(rhs, d0.ntot);
    const long long gcell = (long long)(d0.n[0]+2) * (d0.n[1]+2) * (d0.n[2]+2) * 3;
    const long long pnod = (long long)(d0.n[0]+3) * (d0.n[1]+3) * (d0.n[2]+3);
    if (alloc_d(&gphi, gcell) || alloc_d(&phi_out, pnod)) return NP_ERR_CUDA;
    zero(gphi, gcell); zero(phi_out, pnod);
    npart = nblocks(grid_nodes(d0)) + 16;
    if (alloc_d(&part0, npart) || alloc_d(&part1, npart) || alloc_d(&scal, 64)) return NP_ERR_CUDA;
    void* q = nullptr;
    if (np_rt::host_malloc(&q, 64 * sizeof(double))) return NP_ERR_CUDA;
    h_scal = (double*)q;
    if (np_rt::dev_malloc(&q, 8 * sizeof(int))) return NP_ERR_CUDA;
    bot_out = (int*)q;
    if (np_rt::host_malloc(&q, 8 * sizeof(int))) return NP_ERR_CUDA;
    h_bot_out = (int*)q;
    const LevelDev& db = lev.back().d;
    for (int w = 0; w < 7; ++w) { if (alloc_d(&bw[w], db.ntot)) return NP_ERR_CUDA; zero(bw[w], db.ntot); }
    if (agg_lev >= 0) {
        // all-gather staging for the first replicated level: one local coarse slab per rank
        LevelDev dl = lev[agg_lev - 1].d;
        set_dims(dl, dl.n[0] / 2, dl.n[1] / 2, dl.n[2] / 2);
        const long long m = std::max(dl.ntot, dl.ctot);
        if (alloc_d(&agg_send, m) || alloc_d(&agg_recv, m * nranks)) return NP_ERR_CUDA;
        zero(agg_send, m); zero(agg_recv, m * nranks);
    }
    for (int e = 0; e < 4; ++e) np_rt::event_create(&ev[e]);
    if (nranks > 1 && opt.overlap_halo) {
        have_comm_stream = np_rt::stream_create_high_priority(&comm_stream) == 0;
        np_rt::event_create(&ev_halo[0]); np_rt::event_create(&ev_halo[1]);
    }
    if (opt.profile_kernels) { prof_ev.resize(2048); for (auto& e : prof_ev) np_rt::event_create(&e); }
    for (auto& L : lev) setup_level_bc(L);
    if (nranks > 1 && opt.p2p_halo) setup_peer_halo();
#ifndef NP_EMU
    cudaFuncSetAttribute(k_gs2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GS_SMEM_BYTES);
    cudaFuncSetAttribute(k_gs2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GS_SMEM_BYTES);
    cudaFuncSetAttribute(k_bottom_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NP_BOTTOM_SMEM_MAX);
    cudaFuncSetAttribute(k_gs_small<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GSS_SMEM_BYTES(GSS_MAX_ZC));
    cudaFuncSetAttribute(k_gs_small<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GSS_SMEM_BYTES(GSS_MAX_ZC));
#define NP_GS3_ATTR(BW, BH, CS, MINB) cudaFuncSetAttribute(k_gs3<BW, BH, CS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G3<BW, BH>::SMEM_BYTES)
    NP_GS3_ATTR(2, 8, true, 2); NP_GS3_ATTR(2, 4, true, 1); NP_GS3_ATTR(4, 8, true, 2); NP_GS3_ATTR(4, 4, true, 2);
    NP_GS3_ATTR(2, 8, false, 1); NP_GS3_ATTR(2, 4, false, 1); NP_GS3_ATTR(4, 4, false, 1); NP_GS3_ATTR(2, 8, false, 2);
#undef NP_GS3_ATTR
    cudaFuncSetAttribute(k_resid_restrict<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RR_SMEM_BYTES);
    cudaFuncSetAttribute(k_resid_restrict<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RR_SMEM_BYTES);
    cudaFuncSetAttribute(k_stencil<0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ST_SMEM_BYTES);
    cudaFuncSetAttribute(k_stencil<1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ST_SMEM_BYTES);
    cudaFuncSetAttribute(k_stencil<1, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ST_SMEM_BYTES);
    cudaFuncSetAttribute(k_stencil<0, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ST_SMEM_BYTES);
    cudaFuncSetAttribute(k_stencil<1, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ST_SMEM_BYTES);
    cudaFuncSetAttribute(k_stencil<1, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ST_SMEM_BYTES);
#endif
    np_rt::stream_sync(stream);
    return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
}

void NpSolver::destroy()
{
    np_rt::stream_sync(stream);
    for (auto& L : lev) { np_rt::dev_free(L.sigma); np_rt::dev_free(L.res); np_rt::dev_free(L.cor); np_rt::dev_free(L.cor_alt); np_rt::dev_free(L.rescor); }
    lev.clear();
    double* ds[] = {sol, rhs, gphi, phi_out, part0, part1, scal, xbuf_lo, xbuf_hi, agg_send, agg_recv, stage[0], stage[1], stage[2], mb_vel, mb_sig};
    for (double* p : ds) if (p) np_rt::dev_free(p);
    for (double* p : mb_phi) if (p) np_rt::dev_free(p);
    for (double* p : mb_gphi) if (p) np_rt::dev_free(p);
    mb_phi.clear(); mb_gphi.clear(); mb_vel = mb_sig = nullptr;
    for (int w = 0; w < 7; ++w) if (bw[w]) np_rt::dev_free(bw[w]);
    if (h_scal) np_rt::host_free(h_scal);
    if (bot_out) np_rt::dev_free(bot_out);
    if (h_bot_out) np_rt::host_free(h_bot_out);
    for (int e = 0; e < 4; ++e) if (ev[e]) np_rt::event_destroy(ev[e]);
    for (void* m : peer.opened) np_rt::ipc_close(m);
    peer.opened.clear(); peer.ok = false;
    if (peer.flags) { np_rt::dev_free(peer.flags); peer.flags = nullptr; }
    if (peer.counter) { np_rt::dev_free(peer.counter); peer.counter = nullptr; }
    if (have_comm_stream) { np_rt::stream_sync(comm_stream); np_rt::stream_destroy(comm_stream); have_comm_stream = false; }
    for (int e = 0; e < 2; ++e) if (ev_halo[e]) { np_rt::event_destroy(ev_halo[e]); ev_halo[e] = 0; }
    for (auto& e : prof_ev) np_rt::event_destroy(e);
    prof_ev.clear();
    drop_graph();
    for (auto& e : gprof_ev) np_rt::event_destroy(e);
    gprof_ev.clear();
    if (have_cap_stream) { np_rt::stream_destroy(cap_stream); have_cap_stream = false; }
}

// A cached handle is handed out again by np_create (np_capi.cu): every array is cleared and every flag reset, so the handle
// is in exactly the state create() leaves, without the cudaMalloc / IPC / all-gather work (cudaMemsetAsync of ~9 GB at
// 512^3 takes a few milliseconds, the allocations it replaces a few hundred).
void NpSolver::reset_for_reuse()
{
    stream = 0;
    for (int d = 0; d < 3; ++d) { last_bclo[d] = bclo[d]; last_bchi[d] = bchi[d]; }
    reuse_pending = true;            // see ensure_clean()
    for (auto& L : lev) { L.cor_ghost_clean = false; L.d.const_sigma = 0; L.d.sigma0 = 1.0; }
    for (int d = 0; d < 3; ++d) bclo[d] = bchi[d] = geom.periodic[d] ? NP_BC_PERIODIC : NP_BC_NEUMANN;
    singular = true;
    for (auto& L : lev) setup_level_bc(L);
    have_sigma = have_result = false; bicg_failed_switch = false;
    stats = np_stats{}; err = 0; launches = 0; divu_npart = 0; red_n = 0; prof_n = 0; lvl0_off = nullptr; wtot = 1.0;
    graph_off = false; g_ms_smooth = 0.0; g_n_smooth = 0;      // (the captured graph stays: its key decides whether it still fits)
}

// First use of a handle that came back from the cache.  Every valid node / cell of every array is written before it is
// read, and so is every ghost value that a boundary fill or halo exchange produces; the only memory whose content
// survives from the previous use and can be seen again are ghost values next to faces whose boundary-condition kind
// changed (they are "left alone" under one kind and filled under another, e.g. the ghost cells of grad(phi) handed back to
// the caller).  With the same BCs as last time - every time step of an incflo run - nothing needs clearing; otherwise
// all arrays are cleared as create() does.
void NpSolver::ensure_clean()
{
    if (!reuse_pending) return;
    reuse_pending = false;
    bool same = true;
    for (int d = 0; d < 3; ++d) if (bclo[d] != last_bclo[d] || bchi[d] != last_bchi[d]) same = false;
    if (same) return;
    for (auto& L : lev) { zero(L.res, L.d.ntot); zero(L.cor, L.d.ntot); zero(L.cor_alt, L.d.ntot); zero(L.rescor, L.d.ntot); zero(L.sigma, L.d.ctot); }
    const LevelDev& d0 = lev[0].d;
    zero(sol, d0.ntot); zero(rhs, d0.ntot);
    zero(gphi, (long long)(d0.n[0]+2) * (d0.n[1]+2) * (d0.n[2]+2) * 3);
    zero(phi_out, (long long)(d0.n[0]+3) * (d0.n[1]+3) * (d0.n[2]+3));
    for (int w = 0; w < 7; ++w) zero(bw[w], lev.back().d.ntot);
}

// Per-level face types and ghost maps from the domain BCs and this rank's position in the slab stack.
void NpSolver::setup_level_bc(NpLevel& L)
{
    LevelDev& d = L.d;
    const bool sharded = comm.nranks > 1 && !L.replicated;
    for (int dim = 0; dim < 3; ++dim) {
        bool lo_dom = true, hi_dom = true;
        if (dim == 2 && sharded) { lo_dom = (L.koff == 0); hi_dom = (L.koff + d.n[2] == L.nglob2); }
        const int blo = bclo[dim], bhi = bchi[dim];
        d.face[dim]     = (lo_dom && blo != NP_BC_PERIODIC) ? (blo == NP_BC_DIRICHLET ? NP_FACE_DIRICHLET : NP_FACE_NEUMANN) : NP_FACE_NONE;
        d.face[dim + 3] = (hi_dom && bhi != NP_BC_PERIODIC) ? (bhi == NP_BC_DIRICHLET ? NP_FACE_DIRICHLET : NP_FACE_NEUMANN) : NP_FACE_NONE;
        const bool remote = (dim == 2 && sharded);
        d.map_lo[dim] = d.face[dim] == NP_FACE_NEUMANN ? NP_MAP_MIRROR
                      : (blo == NP_BC_PERIODIC && !remote ? NP_MAP_PERIODIC : NP_MAP_NONE);
        d.map_hi[dim] = d.face[dim + 3] == NP_FACE_NEUMANN ? NP_MAP_MIRROR
                      : (bhi == NP_BC_PERIODIC && !remote ? NP_MAP_PERIODIC : NP_MAP_NONE);
        // the high face is a duplicate when it is periodic or shared with the next rank
        const bool hi_dup = (bhi == NP_BC_PERIODIC) || (remote && !hi_dom);
        d.owned_hi[dim] = hi_dup ? d.n[dim] - 1 : d.n[dim];
    }
}

int NpSolver::set_bc(const int lo[3], const int hi[3])
{
    for (int d = 0; d < 3; ++d) {
        if ((lo[d] == NP_BC_PERIODIC) != (hi[d] == NP_BC_PERIODIC)) return NP_ERR_INVALID;
        if ((lo[d] == NP_BC_PERIODIC) != (geom.periodic[d] != 0)) return NP_ERR_INVALID;
        if (lo[d] < 0 || lo[d] > 2 || hi[d] < 0 || hi[d] > 2) return NP_ERR_INVALID;
        bclo[d] = lo[d]; bchi[d] = hi[d];
    }
    singular = true;
    for (int d = 0; d < 3; ++d) if (bclo[d] == NP_BC_DIRICHLET || bchi[d] == NP_BC_DIRICHLET) singular = false;
    for (auto& L : lev) setup_level_bc(L);
    return NP_OK;
}

// ------------------------------------------------------------------------------------------------
// Building blocks
// ------------------------------------------------------------------------------------------------
static inline bool is_sharded(const NpSolver& S, const NpLevel& L) { return S.comm.nranks > 1 && !L.replicated; }

// ring neighbours of this rank for a z-periodic / non-periodic slab stack
static void z_neighbours(const NpSolver& S, int& lo, int& hi) {
    const int r = S.comm.rank, R = S.comm.nranks;
    const bool per = S.bclo[2] == NP_BC_PERIODIC;
    lo = r > 0 ? r - 1 : (per ? R - 1 : -1);
    hi = r < R - 1 ? r + 1 : (per ? 0 : -1);
}

// Peer-memory halo exchange, set up once per handle (collective): every rank exports the IPC handles of the arrays that
// are ever exchanged (cor, cor_alt, rescor of the sharded levels, sol) and of its flag words, all ranks gather them, and
// each rank maps its two z-neighbours'.  Used only if EVERY rank succeeded; otherwise the NCCL path stays.
int NpSolver::setup_peer_halo()
{
    if (!comm.allgather || !comm.allreduce) return 1;
    for (auto& L : lev) if (!L.replicated) { peer.mine.push_back(L.cor); peer.mine.push_back(L.cor_alt); peer.mine.push_back(L.rescor); }
    peer.mine.push_back(sol);
    void* q = nullptr;
    if (np_rt::dev_malloc(&q, 8 * sizeof(unsigned long long))) return 1;
    peer.flags = (unsigned long long*)q; np_rt::memset_async(peer.flags, 0, 8 * sizeof(unsigned long long), stream);
    if (np_rt::dev_malloc(&q, 4 * sizeof(unsigned int))) return 1;
    peer.counter = (unsigned int*)q; np_rt::memset_async(peer.counter, 0, 4 * sizeof(unsigned int), stream);
    const int na = (int)peer.mine.size() + 1;                       // + the flag words
    std::vector<double> hsend((size_t)na * 8, 0.0), hrecv((size_t)na * 8 * comm.nranks, 0.0);
    int good = 1;
    for (int a = 0; a < na; ++a) {
        void* p = a < na - 1 ? (void*)peer.mine[a] : (void*)peer.flags;
        if (np_rt::ipc_export(p, &hsend[(size_t)a * 8])) good = 0;
    }
    double *dsend = nullptr, *drecv = nullptr;
    if (alloc_d(&dsend, (long long)na * 8) || alloc_d(&drecv, (long long)na * 8 * comm.nranks)) return 1;
    np_rt::copy_async(dsend, hsend.data(), hsend.size() * sizeof(double), stream);
    if (comm.allgather(comm.ctx, dsend, drecv, (long long)na * 8, (void*)stream)) { err = NP_ERR_NCCL; good = 0; }
    np_rt::copy_async(hrecv.data(), drecv, hrecv.size() * sizeof(double), stream);
    np_rt::stream_sync(stream);
    int lo, hi; z_neighbours(*this, lo, hi);
    peer.lo.assign(na - 1, nullptr); peer.hi.assign(na - 1, nullptr);
    auto open_rank = [&](int r, std::vector<double*>& arr, unsigned long long** fl) {
        for (int a = 0; a < na && good; ++a) {
            void* m = nullptr;
            if (np_rt::ipc_open(&hrecv[((size_t)r * na + a) * 8], &m)) { good = 0; break; }
            peer.opened.push_back(m);
            if (a < na - 1) arr[a] = (double*)m; else *fl = (unsigned long long*)m;
        }
    };
    if (good && lo >= 0) open_rank(lo, peer.lo, &peer.lo_flags);
    if (good && hi >= 0) {
        if (hi == lo) { peer.hi = peer.lo; peer.hi_flags = peer.lo_flags; }      // two ranks, periodic: one neighbour on both sides
        else open_rank(hi, peer.hi, &peer.hi_flags);
    }
    // all or nothing: one max-allreduce of "somebody failed"
    h_scal[16] = good ? 0.0 : 1.0;
    np_rt::copy_async(scal + 16, h_scal + 16, sizeof(double), stream);
    if (comm.allreduce(comm.ctx, scal + 16, 1, 1, (void*)stream)) good = 0;
    np_rt::copy_async(h_scal + 16, scal + 16, sizeof(double), stream);
    np_rt::stream_sync(stream);
    np_rt::dev_free(dsend); np_rt::dev_free(drecv);
    peer.ok = good && h_scal[16] == 0.0;
    if (opt.verbose >= 1 && comm.rank == 0) std::printf("nodalproj: halo exchange over %s\n", peer.ok ? "peer memory (CUDA IPC, NVLink)" : "NCCL send/recv");
    return peer.ok ? 0 : 1;
}

// one exchange of the z-boundary planes of x (level l) through peer memory; false: x is not registered, use NCCL
bool NpSolver::halo_p2p(int l, double* x, np_stream_t on)
{
    if (!peer.ok) return false;
    int id = -1;
    for (int a = 0; a < (int)peer.mine.size(); ++a) if (peer.mine[a] == x) { id = a; break; }
    if (id < 0) return false;
    const LevelDev& d = lev[l].d;
    const long long pl = d.nsz;
    int lo, hi; z_neighbours(*this, lo, hi);
    HaloP2P a{};
    // to the lower rank: my planes 0,1 -> its n2, n2+1; to the upper rank: my plane n2-1 -> its -1 (all slabs are equal)
    a.send_lo = x + pl * 1; a.n_lo = 2 * pl; a.dst_lo = lo >= 0 ? peer.lo[id] + pl * (long long)(d.n[2] + 1) : nullptr;
    a.send_hi = x + pl * (long long)d.n[2]; a.n_hi = pl; a.dst_hi = hi >= 0 ? peer.hi[id] + 0 : nullptr;
    a.mine = peer.flags; a.lo_flags = lo >= 0 ? peer.lo_flags : nullptr; a.hi_flags = hi >= 0 ? peer.hi_flags : nullptr;
    a.epoch_word = peer.flags + 4; a.counter = peer.counter;        // words 0..3: the flags the neighbours write; word 4: my exchange count
    const long long chunks = (3 * pl / 2 + 255) / 256;                          // 16-byte stores per thread pass
    const int nb = (int)std::min<long long>(64, std::max<long long>(1, chunks / 8));
    NP_LAUNCH(false, k_halo_p2p, dim3(nb), dim3(256), 0, on, a); ++launches;
    return true;
}

// FillBoundary + nodalSync + applyBC on one nodal array
void NpSolver::fill(int l, double* x)
{
    NpLevel& L = lev[l];
    const LevelDev& d = L.d;
    const long long nshell = 3LL * ((long long)(d.n[1]+3)*(d.n[2]+3) + (long long)(d.n[0]+3)*(d.n[2]+3) + (long long)(d.n[0]+3)*(d.n[1]+3));
    NP_LAUNCH(false, k_fill_boundary, dim3(shell_blocks(nshell)), dim3(256), 0, stream, d, x); ++launches;
    if (is_sharded(*this, L)) {
        if (halo_p2p(l, x, stream)) return;
        int lo, hi; z_neighbours(*this, lo, hi);
        const long long pl = d.nsz;
        // to the lower rank: my planes 0,1 (its n2, n2+1); to the upper rank: my plane n2-1 (its -1)
        int rc = comm.sendrecv(comm.ctx, x + pl * 1, 2 * pl, x + 0, pl,
                               x + pl * (long long)d.n[2], pl, x + pl * (long long)(d.n[2] + 1), 2 * pl, lo, hi, (void*)stream);
        if (rc) err = NP_ERR_NCCL;
    }
}

// stencil z-chunk: enough CTAs for ~6 per SM, 4..64 output planes each (two extra planes are staged per chunk)
static int st_zchunk(const LevelDev& d) {
    const int ncols = ((d.n[0] + GS_TX) / GS_TX) * ((d.n[1] + GS_TY) / GS_TY);
    const int nch_want = (888 + ncols - 1) / ncols;
    const int nn = d.n[2] + 1;
    return std::min(64, std::max(4, (nn + nch_want - 1) / nch_want));
}
static StCoef st_coef(const LevelDev& d) {
    const double sg = d.sigma0;
    StCoef c;
    c.d = -32.0 * d.c.fxyz * sg;
    c.w_fx = 4.0 * d.c.f4xm2ym2z * sg; c.w_fy = 4.0 * d.c.fm2x4ym2z * sg; c.w_fz = 4.0 * d.c.fm2xm2y4z * sg;
    c.w_exy = 2.0 * d.c.f2x2ymz * sg; c.w_exz = 2.0 * d.c.f2xmy2z * sg; c.w_eyz = 2.0 * d.c.fmx2y2z * sg;
    c.w_c = d.c.fxyz * sg;
    return c;
}

void NpSolver::apply(int l, double* y, double* x)
{
    fill(l, x);
    const LevelDev& d = lev[l].d;
    const int zc = st_zchunk(d);
    const dim3 grid((d.n[0] + GS_TX) / GS_TX, (d.n[1] + GS_TY) / GS_TY, (d.n[2] + zc) / zc);
    const double* nul = nullptr;
    if (d.const_sigma) NP_LAUNCH(true, (k_stencil<0, false, true>), grid, dim3(256), ST_SMEM_BYTES, stream, d, st_coef(d), y, (const double*)x, nul, nul, part0, part1, zc, nul);
    else               NP_LAUNCH(true, (k_stencil<0, false, false>), grid, dim3(256), ST_SMEM_BYTES, stream, d, st_coef(d), y, (const double*)x, nul, nul, part0, part1, zc, (const double*)lev[l].sigma);
    ++launches;
}

// y = (b - *boff) - A x.  with_red leaves per-CTA partials in part0 (max |y|) and part1 (dot-mask weighted
// sum of y) for finish_norm: the inf-norm and the next solvability offset come out of the residual pass.
void NpSolver::residual(int l, double* y, double* x, const double* b, bool with_red, const double* boff)
{
    if (!(x == lev[l].cor && lev[l].cor_ghost_clean)) fill(l, x);      // a smooth leaves the correction with clean ghosts
    const LevelDev& d = lev[l].d;
    const int zc = st_zchunk(d);
    const dim3 grid((d.n[0] + GS_TX) / GS_TX, (d.n[1] + GS_TY) / GS_TY, (d.n[2] + zc) / zc);
    const StCoef c = st_coef(d);
    const double* sg = lev[l].sigma;
    if (d.const_sigma) {
        if (with_red) NP_LAUNCH(true, (k_stencil<1, true, true>), grid, dim3(256), ST_SMEM_BYTES, stream, d, c, y, (const double*)x, b, boff, part0, part1, zc, sg);
        else          NP_LAUNCH(true, (k_stencil<1, false, true>), grid, dim3(256), ST_SMEM_BYTES, stream, d, c, y, (const double*)x, b, boff, part0, part1, zc, sg);
    } else {
        if (with_red) NP_LAUNCH(true, (k_stencil<1, true, false>), grid, dim3(256), ST_SMEM_BYTES, stream, d, c, y, (const double*)x, b, boff, part0, part1, zc, sg);
        else          NP_LAUNCH(true, (k_stencil<1, false, false>), grid, dim3(256), ST_SMEM_BYTES, stream, d, c, y, (const double*)x, b, boff, part0, part1, zc, sg);
    }
    red_n = nblocks(grid);
    ++launches;
}

// fold the partials left by a reducing kernel on level l: returns the global inf-norm; with want_off the
// solvability offset sum(w*y)/sum(w) of the same array is left on the device in scal[6]
double NpSolver::finish_norm(int l, bool want_off)
{
    const bool sh = is_sharded(*this, lev[l]);
    NP_LAUNCH(true, k_reduce_norm_off, dim3(1), dim3(1024), 0, stream, (const double*)part0, (const double*)part1, red_n, scal, wtot, want_off ? 1 : 0); ++launches;
    if (sh) {
        if (comm.allreduce(comm.ctx, scal, 1, 1, (void*)stream)) err = NP_ERR_NCCL;
        if (want_off) {
            if (comm.allreduce(comm.ctx, scal + 2, 1, 0, (void*)stream)) err = NP_ERR_NCCL;
            NP_LAUNCH(false, k_set_offset, dim3(1), dim3(32), 0, stream, scal, wtot); ++launches;
        }
    }
    np_rt::copy_async(h_scal, scal, sizeof(double), stream);
    if (in_capture) return 0.0;              // read after the graph has run
    np_rt::stream_sync(stream);
    return h_scal[0];
}

double NpSolver::norminf(int l, const double* x)
{
    const LevelDev& d = lev[l].d;
    NP_LAUNCH(true, k_norminf, grid_nodes(d), NP_BLK, 0, stream, d, x, part0); ++launches;
    red_n = nblocks(grid_nodes(d));
    return finish_norm(l, false);
}

// MLMG::makeSolvable on level l: x -= (sum w x)/(sum w) over owned nodes, scalars never leave the device
void NpSolver::make_solvable(int l, double* x)
{
    const LevelDev& d = lev[l].d;
    const long long nb = nblocks(grid_nodes(d));
    NP_LAUNCH(true, k_wdot, grid_nodes(d), NP_BLK, 0, stream, d, (const double*)x, (const double*)nullptr, part0, part1); ++launches;
    NP_LAUNCH(true, (k_reduce_final<false>), dim3(1), dim3(256), 0, stream, (const double*)part0, nb, scal, 2); ++launches;
    NP_LAUNCH(true, (k_reduce_final<false>), dim3(1), dim3(256), 0, stream, (const double*)part1, nb, scal, 3); ++launches;
    if (is_sharded(*this, lev[l])) { if (comm.allreduce(comm.ctx, scal + 2, 2, 0, (void*)stream)) err = NP_ERR_NCCL; }
    NP_LAUNCH(false, k_sub_offset, grid_nodes(d), NP_BLK, 0, stream, d, x, (const double*)(scal + 2)); ++launches;
}

// one Fsmooth (two GS sweeps) on lev[l].cor with rhs lev[l].res (- *lvl0_off on level 0), then sync + ghost
// fill of the result.  zero_in: the input is known to be zero and is not read.  add_to: the result is not
// stored as the correction but added to `add_to` in place (MLMG's sol += cor folded into the last smooth).
void NpSolver::smooth(int l, bool zero_in, double* add_to)
{
    NpLevel& L = lev[l];
    const LevelDev& d = L.d;
    if (!zero_in && !L.cor_ghost_clean) fill(l, L.cor);
    const dim3 grid((d.n[0] + 1 + GS_WX - 1) / GS_WX, (d.n[1] + 1 + GS_WY - 1) / GS_WY, (d.n[2] + 1 + L.zc - 1) / L.zc);
    const size_t smem = GS_SMEM_BYTES;
    // profile_kernels: event pairs around the level-0 launches; inside a graph capture they become event-record nodes
    const bool gprof = (l == 0 && opt.profile_kernels && in_capture && itg.n_prof + 1 <= (int)gprof_ev.size() / 2);
    const bool prof = (l == 0 && opt.profile_kernels && !in_capture && prof_n + 2 <= (int)prof_ev.size());
    if (prof) np_rt::event_record(prof_ev[prof_n], stream);
    if (gprof) np_rt::event_record_in_graph(gprof_ev[2 * itg.n_prof], stream);
    GsArgs ga{};
    ga.xout = add_to ? add_to : L.cor_alt; ga.xin = L.cor; ga.rhs = L.res; ga.sig = L.sigma;
    ga.boff = (l == 0) ? lvl0_off : nullptr; ga.addto = add_to; ga.zc = L.zc; ga.zero_in = zero_in ? 1 : 0;
    ga.zmap0 = 0; ga.zmapS = 1;
    if (d.const_sigma) {
        // constant sigma: the 26 neighbours fall into 7 weight classes; everything is pre-divided by the diagonal
        const double sg = d.sigma0, dinv = 1.0 / (-32.0 * d.c.fxyz * sg);
        GsCoef& gc = ga.c;
        gc.dinv = dinv;
        gc.a_fx = 4.0 * d.c.f4xm2ym2z * sg * dinv; gc.a_fy = 4.0 * d.c.fm2x4ym2z * sg * dinv; gc.a_fz = 4.0 * d.c.fm2xm2y4z * sg * dinv;
        gc.a_exy = 2.0 * d.c.f2x2ymz * sg * dinv; gc.a_exz = 2.0 * d.c.f2xmy2z * sg * dinv; gc.a_eyz = 2.0 * d.c.fmx2y2z * sg * dinv;
        gc.a_c = d.c.fxyz * sg * dinv;
    }
#define NP_GS3(BW, BH, CS, MINB) NP_LAUNCH(true, (k_gs3<BW, BH, CS, MINB>), g, dim3(G3<BW, BH>::CTA), (G3<BW, BH>::SMEM_BYTES), stream, d, ga)
    auto launch = [&](int nz) {
        const dim3 g(grid.x, grid.y, nz);
        // a launch of at most one CTA per SM is a chain of plane steps with nothing to hide behind: the 2 x 2-block kernel
        if (gs_small && ga.zc <= GSS_MAX_ZC && (long long)g.x * g.y * g.z <= (long long)gs_small * np_rt::sm_count()) {
            if (d.const_sigma) NP_LAUNCH(true, (k_gs_small<true>), g, dim3(GSS_CTA), GSS_SMEM_BYTES(ga.zc), stream, d, ga);
            else               NP_LAUNCH(true, (k_gs_small<false>), g, dim3(GSS_CTA), GSS_SMEM_BYTES(ga.zc), stream, d, ga);
            ++launches;
            return;
        }
        const bool small_launch = (long long)g.x * g.y * g.z <= (long long)np_rt::sm_count();
        if (d.const_sigma) {
            switch (small_launch && gs_small_variant ? gs_small_variant : gs_variant_c) {
            case 1: NP_GS3(2, 8, true, 2); break;
            case 2: NP_GS3(2, 4, true, 1); break;
            case 3: NP_GS3(4, 8, true, 2); break;
            case 4: NP_GS3(4, 4, true, 2); break;
            default: NP_LAUNCH(true, (k_gs2<true>), g, dim3(256), smem, stream, d, ga); break;
            }
        } else {
            switch (small_launch && gs_small_variant ? gs_small_variant : gs_variant_v) {
            case 1: NP_GS3(2, 8, false, 1); break;
            case 2: NP_GS3(2, 4, false, 1); break;
            case 3: NP_GS3(4, 4, false, 1); break;
            case 4: NP_GS3(2, 8, false, 2); break;
            default: NP_LAUNCH(true, (k_gs2<false>), g, dim3(256), smem, stream, d, ga); break;
            }
        }
        ++launches;
    };
#undef NP_GS3
    // Sharded level: relax the two end z-chunks first and ship their planes to the neighbours on the comm stream while
    // the interior chunks are relaxed (FillBoundary overlapped with the interior work).  The end chunks write the
    // planes that are sent (k = 0, 1, n2-1) as long as a chunk has two planes, which gs_zchunk guarantees.
    const int nch = (int)grid.z;
    // (only where the interior launch still fills the GPU several times over: on a one-wave level the split would double the latency)
    const bool split = !add_to && have_comm_stream && is_sharded(*this, L) && nch >= 3 && L.zc >= 2
                    && (opt.overlap_halo < 0 || (long long)grid.x * grid.y * (nch - 2) >= 4LL * 2 * np_rt::sm_count())
                    && (long long)(nch - 1) * L.zc <= d.n[2] - 1;          // the last chunk holds plane n2-1
    if (split) {
        ga.zmap0 = 0; ga.zmapS = nch - 1; launch(2);
        np_rt::event_record(ev_halo[0], stream);
        np_rt::stream_wait_event(comm_stream, ev_halo[0]);
        double* xo = L.cor_alt;
        const long long per = 3LL * (d.n[1] + 3) + 3LL * (d.n[0] + 3);
        NP_LAUNCH(false, k_fill_xy_planes, dim3(shell_blocks(per * 2)), dim3(256), 0, comm_stream, d, xo, 0, 2); ++launches;
        NP_LAUNCH(false, k_fill_xy_planes, dim3(shell_blocks(per)), dim3(256), 0, comm_stream, d, xo, d.n[2] - 1, 1); ++launches;
        if (!halo_p2p(l, xo, comm_stream)) {          // peer-memory stores beside the interior chunks; NCCL if peer memory is unavailable
            int lo, hi; z_neighbours(*this, lo, hi);
            const long long pl = d.nsz;
            if (comm.sendrecv(comm.ctx, xo + pl * 1, 2 * pl, xo + 0, pl, xo + pl * (long long)d.n[2], pl,
                              xo + pl * (long long)(d.n[2] + 1), 2 * pl, lo, hi, (void*)comm_stream)) err = NP_ERR_NCCL;
        }
        np_rt::event_record(ev_halo[1], comm_stream);
        ga.zmap0 = 1; ga.zmapS = 1; launch(nch - 2);
        if (prof) { np_rt::event_record(prof_ev[prof_n + 1], stream); prof_n += 2; }
    if (gprof) { np_rt::event_record_in_graph(gprof_ev[2 * itg.n_prof + 1], stream); ++itg.n_prof; }
        std::swap(L.cor, L.cor_alt);
        np_rt::stream_wait_event(stream, ev_halo[1]);
        const long long nshell = 3LL * ((long long)(d.n[1]+3)*(d.n[2]+3) + (long long)(d.n[0]+3)*(d.n[2]+3) + (long long)(d.n[0]+3)*(d.n[1]+3));
        NP_LAUNCH(false, k_fill_boundary, dim3(shell_blocks(nshell)), dim3(256), 0, stream, d, L.cor); ++launches;
        L.cor_ghost_clean = true;
        return;
    }
    launch(nch);
    if (prof) { np_rt::event_record(prof_ev[prof_n + 1], stream); prof_n += 2; }
    if (gprof) { np_rt::event_record_in_graph(gprof_ev[2 * itg.n_prof + 1], stream); ++itg.n_prof; }
    if (add_to) { L.cor_ghost_clean = false; return; }
    std::swap(L.cor, L.cor_alt);
    fill(l, L.cor);                 // nodalSync (+ the ghosts the next operation needs)
    L.cor_ghost_clean = true;
}

// res[lc] = R(fine), fine on level lc-1 (its ghosts are filled here: restriction's applyBC)
void NpSolver::restrict_to(int lc, double* crse, double* fine)
{
    fill(lc - 1, fine);
    const LevelDev& dc = lev[lc].d; const LevelDev& df = lev[lc - 1].d;
    if (lev[lc].replicated && !lev[lc - 1].replicated) { restrict_to_replica(lc, crse, fine); return; }
    NP_LAUNCH(false, k_restrict, grid_nodes(dc), NP_BLK, 0, stream, dc, df, crse, (const double*)fine); ++launches;
}

void NpSolver::interp_add(int lf, double* fine, const double* crse)
{
    const LevelDev& df = lev[lf].d; const LevelDev& dc = lev[lf + 1].d;
    const int koffc = (lev[lf + 1].replicated && !lev[lf].replicated) ? lev[lf].koff / 2 : 0;
    if (opt.interp_sigma_weighted && !df.const_sigma && interp_march) {
        // z-marching form: 31 x 7 coarse cells per CTA and plane, a run of coarse planes per CTA (about six CTAs per SM in all)
        const int nc0 = df.n[0] / 2, nc1 = df.n[1] / 2, nc2 = df.n[2] / 2;
        const int tx = (nc0 + IW_TX - 1) / IW_TX, ty = (nc1 + 1 + IW_TY - 1) / IW_TY;
        const int nch_want = (6 * np_rt::sm_count() + tx * ty - 1) / (tx * ty);
        const int kcz = std::min(32, std::max(std::min(4, nc2), (nc2 + nch_want - 1) / nch_want));
        NP_LAUNCH(true, k_interp_add_w2, dim3(tx, ty, (nc2 + kcz - 1) / kcz), dim3(256), 0, stream, df, dc, fine, crse, (const double*)lev[lf].sigma, koffc, kcz);
    } else if (opt.interp_sigma_weighted && !df.const_sigma)
        NP_LAUNCH(false, k_interp_add_w, dim3((df.n[0] / 2 + 1 + 31) / 32, (df.n[1] / 2 + 1 + 3) / 4, df.n[2] / 2 + 1), dim3(32, 4, 1), 0, stream, df, dc, fine, crse, (const double*)lev[lf].sigma, koffc);
    else
        NP_LAUNCH(false, k_interp_add_tri, dim3((df.n[0] / 2 + 1 + 31) / 32, (df.n[1] / 2 + 1 + 7) / 8, df.n[2] / 2 + 1), NP_BLK, 0, stream, df, dc, fine, crse, koffc);
    ++launches;
}

// MLMG::bottomSolve: x = cor[bottom] <- solve(res[bottom]).  Returns the number of Krylov iterations.
int NpSolver::bottom_solve()
{
    const int l = (int)lev.size() - 1;
    NpLevel& L = lev[l];
    int iters = 0;
    auto run = [&](int type) -> int {
        BottomArgs a;
        a.L = L.d; a.x = L.cor; a.b = L.res;
        a.r = bw[0]; a.rh = bw[1]; a.p = bw[2]; a.v = bw[3]; a.s = bw[4]; a.t = bw[5]; a.bb = bw[6];
        a.sig = L.sigma; a.maxiter = opt.bottom_maxiter; a.eps_rel = opt.bottom_rtol; a.eps_abs = opt.bottom_atol;
        a.singular = singular ? 1 : 0; a.normalize = opt.bottom_normalize; a.type = type; a.out = bot_out;
        const long long nv = (long long)(L.d.n[0] + 1) * (L.d.n[1] + 1) * (L.d.n[2] + 1);
        int nth = (int)std::min<long long>(1024, std::max<long long>(64, ((nv + 31) / 32) * 32));
        const size_t bsm = (size_t)8 * L.d.ntot * sizeof(double);
        a.in_smem = bsm <= NP_BOTTOM_SMEM_MAX ? 1 : 0;
        NP_LAUNCH(true, k_bottom_solve, dim3(1), dim3(nth), a.in_smem ? bsm : (size_t)0, stream, a); ++launches;
        np_rt::copy_async(h_bot_out, bot_out, 2 * sizeof(int), stream);
        if (in_capture) return 0;            // the status is looked at after the graph has run (iterate_graph)
        np_rt::stream_sync(stream);
        iters += h_bot_out[1];
        return h_bot_out[0];
    };
    if (opt.bottom_solver == NP_BOTTOM_SMOOTHER) {
        for (int i = 0; i < opt.num_final_smooth; ++i) smooth(l, i == 0, nullptr);
        return 0;
    }
    int type = (opt.bottom_solver == NP_BOTTOM_CG || (opt.bottom_solver == NP_BOTTOM_BICGCG && bicg_failed_switch)) ? 1 : 0;
    int ret = run(type);
    if (ret != 0 && ret != 9) {
        if (opt.bottom_verbose) std::printf("MLMG: Bottom solve failed (ret = %d)\n", ret);
        if (opt.bottom_solver == NP_BOTTOM_BICGCG) {
            const int other = 1 - type;
            ret = run(other);
            if (ret == 0) bicg_failed_switch = (other == 1);
        }
        if (ret != 0) zero(L.cor, L.d.ntot);
    }
    L.cor_ghost_clean = false;
    if (ret != 0) for (int i = 0; i < opt.num_final_smooth; ++i) smooth(l, false, nullptr);
    return iters;
}

// MLMG::mgVcycle(0,0) on res[0] (- *lvl0_off): leaves the correction in cor[0], or, with add_to_sol, adds it
// to sol inside the last post-smooth (MLMG::oneIter's sol += cor)
void NpSolver::vcycle(bool add_to_sol)
{
    np_rt::Range nvtx_vc("MLMG::mgVcycle");
    const int nl = (int)lev.size();
    for (int l = 0; l < nl - 1; ++l) {
        NpLevel& L = lev[l];
        if (opt.num_pre_smooth == 0) { zero(L.cor, L.d.ntot); L.cor_ghost_clean = true; }
        for (int i = 0; i < opt.num_pre_smooth; ++i) smooth(l, i == 0, nullptr);      // cor = 0 is implied by zero_in
        if (fused_rr) resid_restrict(l, l == 0 ? lvl0_off : nullptr);              // computeResOfCorrection + restriction, one pass
        else {
            residual(l, L.rescor, L.cor, L.res, false, l == 0 ? lvl0_off : nullptr);
            restrict_to(l + 1, lev[l + 1].res, L.rescor);
        }
    }
    stats.bottom_iters_total += bottom_solve();
    for (int l = nl - 2; l >= 0; --l) {
        NpLevel& L = lev[l];
        interp_add(l, L.cor, lev[l + 1].cor);                       // addInterpCorrection
        L.cor_ghost_clean = false;
        for (int i = 0; i < opt.num_post_smooth; ++i)
            smooth(l, false, (add_to_sol && l == 0 && i == opt.num_post_smooth - 1) ? sol : nullptr);
    }
    if (nl == 1 && add_to_sol) { NP_LAUNCH(false, k_add, grid_nodes(lev[0].d), NP_BLK, 0, stream, lev[0].d, sol, (const double*)lev[0].cor); ++launches; }
    else if (add_to_sol && opt.num_post_smooth == 0) { NP_LAUNCH(false, k_add, grid_nodes(lev[0].d), NP_BLK, 0, stream, lev[0].d, sol, (const double*)lev[0].cor); ++launches; }
}

void NpSolver::drop_graph()
{
    if (itg.exec) np_rt::graph_destroy(itg.exec);
    itg.exec = nullptr; itg.key.clear(); itg.launches = 0; itg.n_prof = 0;
}

// One MLMG iteration (mgVcycle with sol += cor, the new residual, its norm and the next solvability offset) replayed from a
// CUDA graph; captured on first use and whenever anything baked into it has changed.  lvl0_off is set by the caller.
int NpSolver::iterate_graph(const double* rhs_off, double* resnorm)
{
    if (!opt.use_graph || graph_off || lev.size() < 2 || have_comm_stream) return 1;
    // several ranks: the exchanges inside the graph are the same launches in the same order as on the plain path (the halo
    // kernel counts its exchanges on the device), so ranks may even mix the two; the first iteration of a handle runs plain so
    // that NCCL has set up every collective it needs before one is captured
    if (comm.nranks > 1 && !graph_warm) { graph_warm = true; return 1; }
    std::vector<unsigned char> key;
    auto put = [&](const void* p, size_t n) { key.insert(key.end(), (const unsigned char*)p, (const unsigned char*)p + n); };
    for (const NpLevel& L : lev) {
        put(&L.d, sizeof(LevelDev)); put(&L.sigma, sizeof(double*)); put(&L.res, sizeof(double*)); put(&L.cor, sizeof(double*));
        put(&L.cor_alt, sizeof(double*)); put(&L.rescor, sizeof(double*)); put(&L.zc, sizeof(int));
    }
    put(&sol, sizeof(double*)); put(&rhs, sizeof(double*)); put(&rhs_off, sizeof(double*)); put(&lvl0_off, sizeof(double*));
    put(&opt, sizeof(np_options)); put(&wtot, sizeof(double));
    const int sw[8] = {singular ? 1 : 0, bicg_failed_switch ? 1 : 0, fused_rr ? 1 : 0, interp_march ? 1 : 0, gs_variant_c, gs_variant_v, gs_small, peer.ok ? 1 : 0};
    put(sw, sizeof(sw));
    if (!itg.exec || key != itg.key) {
        drop_graph();
        if (!have_cap_stream) { if (np_rt::stream_create(&cap_stream)) { graph_off = true; return 1; } have_cap_stream = true; }
        if (opt.profile_kernels && gprof_ev.empty()) { gprof_ev.resize(2 * 16); for (auto& e : gprof_ev) np_rt::event_create(&e); }
        // the smoother swaps cor / cor_alt; a V-cycle does not depend on what they held before, so the host state is put
        // back after the capture and every replay starts from the same assignment
        std::vector<double*> save;
        for (NpLevel& L : lev) { save.push_back(L.cor); save.push_back(L.cor_alt); L.cor_ghost_clean = false; }
        const long long l0 = launches;
        const np_stream_t user = stream;
        const int err0 = err;
        bool ok = np_rt::capture_begin(cap_stream) == 0;
        if (ok) {
            stream = cap_stream; in_capture = true;
            vcycle(true);
            residual(0, lev[0].res, sol, rhs, true, rhs_off);
            finish_norm(0, singular);
            in_capture = false; stream = user;
            ok = np_rt::capture_end(cap_stream, &itg.exec) == 0 && !err;
        }
        itg.launches = launches - l0; launches = l0;
        for (size_t q = 0; q < lev.size(); ++q) { lev[q].cor = save[2 * q]; lev[q].cor_alt = save[2 * q + 1]; lev[q].cor_ghost_clean = false; }
        if (!ok) { drop_graph(); np_rt::last_error(); err = err0; graph_off = true; return 1; }
        itg.key = key; ++stats.graph_captures;
    }
    if (np_rt::graph_launch(itg.exec, stream)) { np_rt::last_error(); drop_graph(); graph_off = true; return 1; }
    launches += itg.launches; ++stats.graph_replays;
    np_rt::stream_sync(stream);
    for (int q = 0; q < itg.n_prof; ++q) { g_ms_smooth += np_rt::event_ms(gprof_ev[2 * q], gprof_ev[2 * q + 1]); ++g_n_smooth; }
    for (NpLevel& L : lev) L.cor_ghost_clean = false;
    if (opt.bottom_solver != NP_BOTTOM_SMOOTHER) {
        if (h_bot_out[0] != 0 && h_bot_out[0] != 9) return 2;
        stats.bottom_iters_total += h_bot_out[1];
    }
    *resnorm = h_scal[0];
    return 0;
}

// MLMG::solve with the zero initial guess NodalProjector uses
int NpSolver::solve(double rtol, double atol)
{
    np_rt::Range nvtx_solve("MLMG::solve");
    NpLevel& L0 = lev[0];
    const LevelDev& d0 = L0.d;
    zero(sol, d0.ntot);
    bicg_failed_switch = false;
    stats.bottom_iters_total = 0; stats.n_vcycles = 0; stats.status = NP_OK; stats.n_levels = (int)lev.size();
    stats.graph_replays = 0; stats.graph_captures = 0;
    for (int i = 0; i < 128; ++i) stats.res_history[i] = 0.0;
    wtot = 1.0;
    for (int dd = 0; dd < 3; ++dd) wtot *= (double)geom.ncell[dd];   // sum of the dot-mask weights (periodic / Neumann dims)
    lvl0_off = nullptr;
    // MLMG::makeSolvable(rhs): the offset sum(w*rhs)/sum(w) stays a device scalar (scal[7]) that the level-0
    // residual subtracts on the fly; its partial sums came out of the divergence kernel
    const double* rhs_off = nullptr;
    if (singular) {
        if (divu_npart > 0) {
            NP_LAUNCH(true, k_reduce_sum_off, dim3(1), dim3(1024), 0, stream, (const double*)part1, divu_npart, scal, 8, 7, wtot); ++launches;
            if (is_sharded(*this, L0)) {
                if (comm.allreduce(comm.ctx, scal + 8, 1, 0, (void*)stream)) err = NP_ERR_NCCL;
                NP_LAUNCH(false, k_div_scalar, dim3(1), dim3(32), 0, stream, scal, 8, 7, wtot); ++launches;
            }
            rhs_off = scal + 7;
            divu_npart = 0;
        } else make_solvable(0, rhs);
    }
    // zero initial guess: the first residual IS the (solvable) rhs, so one pass gives both norms
    residual(0, L0.res, sol, rhs, true, rhs_off);
    const double resnorm0 = finish_norm(0, singular);
    const double rhsnorm0 = resnorm0;
    const double max_norm = rhsnorm0 >= resnorm0 ? rhsnorm0 : resnorm0;
    const double res_target = std::max(atol, std::max(rtol, 1e-16) * max_norm);
    stats.rhs_norm0 = rhsnorm0; stats.res_norm0 = resnorm0; stats.res_norm_final = resnorm0;
    if (!(resnorm0 == resnorm0) || std::isinf(resnorm0)) { stats.status = NP_ERR_DIVERGED; return stats.status; }     // NaN / Inf in u* or sigma
    if (resnorm0 <= res_target) return NP_OK;
    bool converged = false;
    for (int iter = 0; iter < opt.maxiter; ++iter) {
        // MLMG makes the residual solvable before every V-cycle; here the offset (scal[6], left by finish_norm)
        // is subtracted on the fly by the level-0 consumers of res instead of in a separate pass
        np_rt::Range nvtx_iter("MLMG::oneIter");
        lvl0_off = singular ? scal + 6 : nullptr;
        double rn = 0.0;
        const int grc = iterate_graph(rhs_off, &rn);
        if (grc == 2) {
            // a bottom solve inside the graph asked for MLMG's fallback (other Krylov method, final smooths), which the graph
            // does not hold: start again kernel by kernel (rhs is still intact; its offset is removed in place this time)
            lvl0_off = nullptr; graph_off = true;
            return solve(rtol, atol);
        }
        if (grc != 0) {
            vcycle(true);
            lvl0_off = nullptr;
            residual(0, L0.res, sol, rhs, true, rhs_off);
            rn = finish_norm(0, singular);
        }
        lvl0_off = nullptr;
        if (iter < 128) stats.res_history[iter] = rn;
        stats.n_vcycles = iter + 1; stats.res_norm_final = rn;
        if (opt.verbose >= 2 && comm.rank == 0) std::printf("MLMG: Iteration %3d Fine resid/bnorm = %.6e\n", iter + 1, rn / max_norm);
        if (err) { stats.status = err; return err; }
        if (rn <= res_target) { converged = true; break; }
        if (!(rn == rn) || rn > 1e20 * max_norm) { stats.status = NP_ERR_DIVERGED; return stats.status; }
    }
    if (!converged) stats.status = NP_ERR_NOT_CONVERGED;
    return stats.status;
}

// ------------------------------------------------------------------------------------------------
// Multi-GPU: the first replicated level is assembled from every rank's local coarse slab
// ------------------------------------------------------------------------------------------------
static LevelDev local_coarse_of(const NpLevel& F)
{
    LevelDev dl = F.d;
    set_dims(dl, F.d.n[0] / 2, F.d.n[1] / 2, F.d.n[2] / 2);
    return dl;
}
void NpSolver::restrict_to_replica(int lc, double* crse, const double* fine)
{
    const LevelDev dl = local_coarse_of(lev[lc - 1]);
    const LevelDev& dc = lev[lc].d;
    NP_LAUNCH(false, k_restrict, grid_nodes(dl), NP_BLK, 0, stream, dl, lev[lc - 1].d, agg_send, fine); ++launches;
    if (comm.allgather(comm.ctx, agg_send, agg_recv, dl.ntot, (void*)stream)) err = NP_ERR_NCCL;
    NP_LAUNCH(false, k_assemble_nodal, grid_nodes(dc), NP_BLK, 0, stream, dc, dl, crse, (const double*)agg_recv, comm.nranks); ++launches;
}
// MLMG::computeResOfCorrection + MLNodeLaplacian::restriction as one pass over level l (k_resid_restrict): the residual of the
// correction is restricted on the fly and stored on the boundary shell only; the coarse nodes on the faces of the box, which
// need ghost residuals (periodic images, mirrors, the neighbour rank's planes), come from that shell after the usual ghost
// fill (k_restrict_faces).  Result: lev[l + 1].res (through the all-gather when level l + 1 is the first replicated one).
void NpSolver::resid_restrict(int l, const double* boff)
{
    NpLevel& F = lev[l]; NpLevel& Cn = lev[l + 1];
    const LevelDev& df = F.d;
    if (!F.cor_ghost_clean) fill(l, F.cor);
    const bool to_replica = Cn.replicated && !F.replicated;
    const LevelDev dc = to_replica ? local_coarse_of(F) : Cn.d;
    RrArgs rr; rr.Lc = dc; rr.crse = to_replica ? agg_send : Cn.res;
    const int tx = std::max((df.n[0] - 1 + RR_SX - 1) / RR_SX, (dc.n[0] - 2) / 31 + 1);       // every fine node and every interior coarse node
    const int ty = std::max((df.n[1] - 1 + RR_SY - 1) / RR_SY, (dc.n[1] - 2) / 15 + 1);
    const int nck = std::max(dc.n[2] - 1, 1);                                                 // interior coarse planes
    const int nch_want = (888 + tx * ty - 1) / (tx * ty);
    rr.kcz = std::min(32, std::max(2, (nck + nch_want - 1) / nch_want));
    const dim3 grid(std::max(tx, 1), std::max(ty, 1), (nck + rr.kcz - 1) / rr.kcz);
    const StCoef c = st_coef(df);
    if (df.const_sigma) NP_LAUNCH(true, (k_resid_restrict<true>), grid, dim3(256), RR_SMEM_BYTES, stream, df, c, F.rescor, (const double*)F.cor, (const double*)F.res, boff, (const double*)F.sigma, rr);
    else                NP_LAUNCH(true, (k_resid_restrict<false>), grid, dim3(256), RR_SMEM_BYTES, stream, df, c, F.rescor, (const double*)F.cor, (const double*)F.res, boff, (const double*)F.sigma, rr);
    ++launches;
    fill(l, F.rescor);                     // ghosts of the shell: periodic images, mirrors, the neighbours' planes
    const long long nface = 2LL * ((long long)(dc.n[1] + 1) * (dc.n[2] + 1) + (long long)(dc.n[0] + 1) * (dc.n[2] + 1) + (long long)(dc.n[0] + 1) * (dc.n[1] + 1));
    NP_LAUNCH(false, k_restrict_faces, dim3(shell_blocks(nface)), dim3(256), 0, stream, dc, df, rr.crse, (const double*)F.rescor); ++launches;
    if (to_replica) {
        if (comm.allgather(comm.ctx, agg_send, agg_recv, dc.ntot, (void*)stream)) err = NP_ERR_NCCL;
        NP_LAUNCH(false, k_assemble_nodal, grid_nodes(Cn.d), NP_BLK, 0, stream, Cn.d, dc, Cn.res, (const double*)agg_recv, comm.nranks); ++launches;
    }
}
void NpSolver::avgdown_sigma_to_replica(int lc)
{
    const LevelDev dl = local_coarse_of(lev[lc - 1]);
    const LevelDev& dc = lev[lc].d;
    NP_LAUNCH(false, k_avgdown_sigma, grid_cells(dl), NP_BLK, 0, stream, dl, lev[lc - 1].d, agg_send, (const double*)lev[lc - 1].sigma); ++launches;
    if (comm.allgather(comm.ctx, agg_send, agg_recv, dl.ctot, (void*)stream)) err = NP_ERR_NCCL;
    NP_LAUNCH(false, k_assemble_cell, grid_cells(dc), NP_BLK, 0, stream, dc, dl, lev[lc].sigma, (const double*)agg_recv); ++launches;
}

// ------------------------------------------------------------------------------------------------
// sigma on every level (MLNodeLaplacian::setSigma + averageDownCoeffs + FillBoundaryCoeff)
// ------------------------------------------------------------------------------------------------
static void fill_sigma_level(NpSolver& S, int l)
{
    NpLevel& L = S.lev[l];
    const LevelDev& d = L.d;
    const bool sh = is_sharded(S, L);
    const int per0 = S.bclo[0] == NP_BC_PERIODIC, per1 = S.bclo[1] == NP_BC_PERIODIC, per2 = S.bclo[2] == NP_BC_PERIODIC;
    int zlo_local = 1, zhi_local = 1, p2 = per2;
    if (sh) {
        p2 = 0;
        zlo_local = (!per2 && L.koff == 0);
        zhi_local = (!per2 && L.koff + d.n[2] == L.nglob2);
    }
    const long long nshell = 2LL * ((long long)(d.n[1]+2)*(d.n[2]+2) + (long long)(d.n[0]+2)*(d.n[2]+2) + (long long)(d.n[0]+2)*(d.n[1]+2));
    NP_LAUNCH(false, k_fill_sigma, dim3(shell_blocks(nshell)), dim3(256), 0, S.stream, d, L.sigma, per0, per1, p2, zlo_local, zhi_local); ++S.launches;
    if (sh) {
        int lo, hi; z_neighbours(S, lo, hi);
        const long long pl = d.csz;
        if (S.comm.sendrecv(S.comm.ctx, L.sigma + pl * 1, pl, L.sigma + 0, pl,
                            L.sigma + pl * (long long)d.n[2], pl, L.sigma + pl * (long long)(d.n[2] + 1), pl, lo, hi, (void*)S.stream)) S.err = NP_ERR_NCCL;
    }
}

int NpSolver::set_sigma(const FabView& s)
{
    ensure_clean();
    const LevelDev& d0 = lev[0].d;
    dim3 gmm = grid_cells(d0); gmm.z = std::min<unsigned>(gmm.z, 16u);
    NP_LAUNCH(true, k_minmax_cell, gmm, NP_BLK, 0, stream, d0.n[0], d0.n[1], d0.n[2], s, part0, part1); ++launches;
    return finish_sigma(nblocks(gmm), &s);
}

// second half of setSigma: min / max partials are in part0 / part1.  Uniform sigma selects the constant-coefficient
// kernels (nothing is stored); otherwise level 0 is filled from `src` (null: its interior is already in place), its
// ghosts are set and the coarse levels are averaged down (averageDownCoeffs + FillBoundaryCoeff).
int NpSolver::finish_sigma(long long nbm, const FabView* src)
{
    const LevelDev& d0 = lev[0].d;
    NP_LAUNCH(true, k_reduce_minmax, dim3(1), dim3(256), 0, stream, (const double*)part0, (const double*)part1, nbm, scal + 4); ++launches;
    NP_LAUNCH(false, k_negate_one, dim3(1), dim3(32), 0, stream, scal + 4); ++launches;   // (-min, max): one max-allreduce
    if (comm.nranks > 1) { if (comm.allreduce(comm.ctx, scal + 4, 2, 1, (void*)stream)) err = NP_ERR_NCCL; }
    np_rt::copy_async(h_scal + 4, scal + 4, 2 * sizeof(double), stream);
    np_rt::stream_sync(stream);
    const double smin = -h_scal[4], smax = h_scal[5];
    const bool is_const = opt.const_sigma_fastpath && (smin == smax);
    stats.sigma_is_const = is_const ? 1 : 0;
    for (auto& L : lev) { L.d.const_sigma = is_const ? 1 : 0; L.d.sigma0 = smin; }
    if (!is_const) {
        if (src) { NP_LAUNCH(false, k_copy_sigma_plain, grid_cells(d0), NP_BLK, 0, stream, d0, lev[0].sigma, *src); ++launches; }
        fill_sigma_level(*this, 0);
        for (int l = 1; l < (int)lev.size(); ++l) {
            if (lev[l].replicated && !lev[l - 1].replicated) avgdown_sigma_to_replica(l);
            else { NP_LAUNCH(false, k_avgdown_sigma, grid_cells(lev[l].d), NP_BLK, 0, stream, lev[l].d, lev[l - 1].d, lev[l].sigma, (const double*)lev[l - 1].sigma); ++launches; }
            fill_sigma_level(*this, l);
        }
    }
    have_sigma = true;
    return err ? err : NP_OK;
}

FabView NpSolver::view_nodal(const np_fab& f, int l) const
{
    const LevelDev& d = lev[l].d;
    FabView v; v.p = f.p; v.ng = f.ngrow;
    v.sy = d.n[0] + 1 + 2 * f.ngrow; v.sz = v.sy * (d.n[1] + 1 + 2 * f.ngrow); v.sn = v.sz * (d.n[2] + 1 + 2 * f.ngrow);
    return v;
}
FabView NpSolver::view_cell(const np_fab& f, int l, int ncomp) const
{
    const LevelDev& d = lev[l].d;
    FabView v; v.p = f.p; v.ng = f.ngrow;
    v.sy = d.n[0] + 2 * f.ngrow; v.sz = v.sy * (d.n[1] + 2 * f.ngrow); v.sn = v.sz * (d.n[2] + 2 * f.ngrow);
    (void)ncomp;
    return v;
}

// ------------------------------------------------------------------------------------------------
// Host-pointer staging: a caller built without CUDA hands host fabs; they are copied to the device
// (cudaMemcpyAsync, pinned or pageable) and back.
// ------------------------------------------------------------------------------------------------
int NpSolver::import_fab(const np_fab& f, long long nelem, double** dev, bool* staged)
{
    if (np_rt::is_device_ptr(f.p)) { *dev = f.p; *staged = false; return NP_OK; }
#ifdef NP_EMU
    *dev = f.p; *staged = false; return NP_OK;
#else
    if (alloc_d(dev, nelem)) return NP_ERR_CUDA;
    np_rt::copy_async(*dev, f.p, (size_t)nelem * sizeof(double), stream);
    *staged = true;
    return NP_OK;
#endif
}
// persistent (grow-only) device staging for np_project / np_copy_* with host fabs: slot 0 vel, 1 sigma, 2 outputs
double* NpSolver::stage_get(int slot, long long n)
{
    if (stage_n[slot] < n) {
        if (stage[slot]) { np_rt::stream_sync(stream); np_rt::dev_free(stage[slot]); stage[slot] = nullptr; stage_n[slot] = 0; }
        if (alloc_d(&stage[slot], n)) return nullptr;
        stage_n[slot] = n;
    }
    return stage[slot];
}
int NpSolver::export_fab(const np_fab& f, long long nelem, double* dev, bool staged)
{
    if (!staged) return NP_OK;
    np_rt::copy_async(f.p, dev, (size_t)nelem * sizeof(double), stream);
    np_rt::stream_sync(stream);
    np_rt::dev_free(dev);
    return NP_OK;
}

// periodic / rank-neighbour ghost layer of a 3-component cell fab (vel.FillBoundary, fluxes.FillBoundary)
void np_fill_cell3(NpSolver& S, FabView u)
{
    NpLevel& L = S.lev[0];
    const LevelDev& d = L.d;
    const bool sh = is_sharded(S, L);
    const int per0 = S.bclo[0] == NP_BC_PERIODIC, per1 = S.bclo[1] == NP_BC_PERIODIC, per2 = S.bclo[2] == NP_BC_PERIODIC;
    const long long nshell = 2LL * ((long long)(d.n[1]+2)*(d.n[2]+2) + (long long)(d.n[0]+2)*(d.n[2]+2) + (long long)(d.n[0]+2)*(d.n[1]+2));
    NP_LAUNCH(false, k_fill_vel, dim3(shell_blocks(nshell)), dim3(256), 0, S.stream, d, u, per0, per1, sh ? 0 : per2); ++S.launches;
    if (sh) {
        int lo, hi; z_neighbours(S, lo, hi);
        // whole z-layers of the fab (all ghosts in x,y included) are contiguous per component
        for (int n = 0; n < 3; ++n) {
            double* base = u.p + u.sn * n;
            const long long pl = u.sz;
            const int g = u.ng;
            if (S.comm.sendrecv(S.comm.ctx, base + pl * g, pl, base + pl * (g - 1), pl,
                                base + pl * (long long)(g + d.n[2] - 1), pl, base + pl * (long long)(g + d.n[2]), pl,
                                lo, hi, (void*)S.stream)) S.err = NP_ERR_NCCL;
        }
    }
}

// amrex::NodalProjector::project(rtol, atol)
// device-timed phases and launch counts of the call that just ended (events ev[0..2] were recorded on the way)
void NpSolver::finish_stats()
{
    np_rt::event_record(ev[3], stream);
    // peer-memory halos: did a wait for a neighbour give up (k_halo_p2p, np_wait_flag)?  The word stays set: the handle is dead.
    if (peer.ok) np_rt::copy_async(h_scal + 18, peer.flags + 5, sizeof(unsigned long long), stream);
    np_rt::stream_sync(stream);
    if (peer.ok) {
        unsigned long long w; memcpy(&w, h_scal + 18, sizeof(w));
        if (w != 0ull) { if (!err) err = NP_ERR_NCCL; if (opt.verbose >= 1) fprintf(stderr, "[nodalproj] rank %d: a neighbour never reached halo exchange %llu\n", comm.rank, w); }
    }
    stats.ms_setup = np_rt::event_ms(ev[0], ev[1]);
    stats.ms_solve = np_rt::event_ms(ev[1], ev[2]);
    stats.ms_finish = np_rt::event_ms(ev[2], ev[3]);
    stats.ms_total = np_rt::event_ms(ev[0], ev[3]);
    stats.kernel_launches = launches;
    stats.n_smooth0 = prof_n / 2 + g_n_smooth; stats.ms_smooth0 = g_ms_smooth;
    for (int e = 0; e < prof_n; e += 2) stats.ms_smooth0 += np_rt::event_ms(prof_ev[e], prof_ev[e + 1]);
    if (opt.verbose >= 4) {     // per-launch times of the level-0 smoother (first V-cycle: zero-input, plain, plain, add-to)
        fprintf(stderr, "[nodalproj] level-0 smoother launches (ms):");
        for (int e = 0; e < prof_n && e < 32; e += 2) fprintf(stderr, " %.3f", np_rt::event_ms(prof_ev[e], prof_ev[e + 1]));
        fprintf(stderr, "\n");
    }
}

int NpSolver::project(np_fab* vel, const np_fab* sigma, double rtol, double atol)
{
    np_rt::Range nvtx_proj("NodalProjector::project");
    if (vel->ngrow < 1 || vel->ncomp < 3 || sigma->ncomp < 1) return NP_ERR_INVALID;
    err = 0; launches = 0; have_result = false; prof_n = 0; g_ms_smooth = 0.0; g_n_smooth = 0;
    ensure_clean();
    const LevelDev& d0 = lev[0].d;
    np_rt::event_record(ev[0], stream);
    // ---- inputs ----
    FabView uv = view_cell(*vel, 0, 3), sv = view_cell(*sigma, 0, 1);
    const long long nvel = uv.sn * 3, nsig = sv.sn;
    double *dvel = vel->p, *dsig = sigma->p; bool st_vel = false, st_sig = false;
#ifndef NP_EMU
    if (!np_rt::is_device_ptr(vel->p)) {        // host fabs (CPU build of the caller): stage through persistent device buffers
        if (!(dvel = stage_get(0, nvel))) return NP_ERR_CUDA;
        np_rt::copy_async(dvel, vel->p, (size_t)nvel * sizeof(double), stream); st_vel = true;
    }
    if (!np_rt::is_device_ptr(sigma->p)) {
        if (!(dsig = stage_get(1, nsig))) return NP_ERR_CUDA;
        np_rt::copy_async(dsig, sigma->p, (size_t)nsig * sizeof(double), stream); st_sig = true;
    }
#endif
    uv.p = dvel; sv.p = dsig;
    int rc = set_sigma(sv);
    if (rc) return rc;
    // ---- rhs = div(u*) ----
    np_fill_cell3(*this, uv);
    {   // rhs = div(u*) and, for singular problems, the partial sums of its solvability offset in the same pass
        dim3 gd = grid_nodes(d0); gd.z = std::min<unsigned>(gd.z, 16u);     // 16 runs of node planes per column
        if (singular) { NP_LAUNCH(true, (k_divu<true>), gd, NP_BLK, 0, stream, d0, rhs, uv, part1); divu_npart = nblocks(gd); }
        else          { NP_LAUNCH(true, (k_divu<false>), gd, NP_BLK, 0, stream, d0, rhs, uv, part1); divu_npart = 0; }
        ++launches;
    }
    np_rt::event_record(ev[1], stream);
    // ---- solve ----
    rc = solve(rtol, atol);
    np_rt::event_record(ev[2], stream);
    // ---- u = u* - sigma grad(phi); gradphi; phi ----
    FabView gv; gv.p = gphi; gv.ng = 1; gv.sy = d0.n[0] + 2; gv.sz = gv.sy * (d0.n[1] + 2); gv.sn = gv.sz * (d0.n[2] + 2);
    NP_LAUNCH(false, k_mknewu, grid_cells(d0), NP_BLK, 0, stream, d0, uv, gv, (const double*)sol, (const double*)lev[0].sigma); ++launches;
    np_fill_cell3(*this, gv);
    fill(0, sol);
    FabView pv; pv.p = phi_out; pv.ng = 1; pv.sy = d0.n[0] + 3; pv.sz = pv.sy * (d0.n[1] + 3); pv.sn = pv.sz * (d0.n[2] + 3);
    NP_LAUNCH(false, (k_copy_nodal<true>), dim3((d0.n[0] + 3 + 31) / 32, (d0.n[1] + 3 + 7) / 8, d0.n[2] + 3), NP_BLK, 0, stream, d0, sol, pv, 1); ++launches;
    if (st_vel) np_rt::copy_async(vel->p, dvel, (size_t)nvel * sizeof(double), stream);
    (void)st_sig;
    finish_stats();
    if (np_rt::last_error()) { stats.status = NP_ERR_CUDA; return NP_ERR_CUDA; }
    if (err) { stats.status = err; return err; }
    have_result = true;
    return rc;
}

// np_project with several boxes per rank.  AMReX hands the projector a MultiFab whose boxes (max_grid_size, default 32
// cells in 3D) are spread over the ranks; here the boxes of a rank must tile its z-slab.  Each box's u* (valid cells and
// the physical ghost cells outside the domain, which carry 0 or the inflow velocity) and sigma are gathered into
// slab-shaped fabs, NodalProjector::project runs on the slab as for a one-box layout, and u, phi (nodal, 1 ghost) and
// grad(phi) (cell, 3 comp, 1 ghost) are scattered back box by box; the ghost layers of the per-box results are the
// neighbouring boxes' values, i.e. what FillBoundary leaves there in the reference.
void NpSolver::box_copy(const FabView& dst, const int dorg[3], const FabView& src, const int sorg[3], const int rlo[3],
                        const int rhi[3], int ncomp, const np_box* guard_valid)
{
    BoxCopy a;
    for (int d = 0; d < 3; ++d) {
        a.rlo[d] = rlo[d]; a.rhi[d] = rhi[d]; a.dorg[d] = dorg[d]; a.sorg[d] = sorg[d];
        a.vlo[d] = guard_valid ? guard_valid->lo[d] : 0; a.vhi[d] = guard_valid ? guard_valid->hi[d] : 0;
        a.dlo[d] = 0; a.dhi[d] = geom.ncell[d] - 1;
    }
    a.dlo[2] = -koff; a.dhi[2] = geom.ncell[2] - 1 - koff;       // the domain in slab-local z
    a.guard = guard_valid ? 1 : 0;
    if (rhi[0] < rlo[0] || rhi[1] < rlo[1] || rhi[2] < rlo[2]) return;
    NP_LAUNCH(false, k_box_copy, dim3((rhi[0] - rlo[0] + 32) / 32, (rhi[1] - rlo[1] + 8) / 8, rhi[2] - rlo[2] + 1), NP_BLK, 0, stream,
              dst, src, a, ncomp); ++launches;
}
static FabView box_view(double* p, const np_box& b, int ng, int nodal)
{
    FabView v; v.p = p; v.ng = ng;
    v.sy = b.hi[0] - b.lo[0] + 1 + nodal + 2 * ng;
    v.sz = v.sy * (b.hi[1] - b.lo[1] + 1 + nodal + 2 * ng);
    v.sn = v.sz * (b.hi[2] - b.lo[2] + 1 + nodal + 2 * ng);
    return v;
}
int NpSolver::project_boxes(int nlocal, np_fab* vel, const np_fab* sigma, double rtol, double atol)
{
    if (nlocal != (int)lbox.size()) return NP_ERR_INVALID;
    const LevelDev& d0 = lev[0].d;
    const int zero3[3] = {0, 0, 0};
    np_box slab; for (int d = 0; d < 3; ++d) { slab.lo[d] = 0; slab.hi[d] = d0.n[d] - 1; }
    const FabView us = box_view(nullptr, slab, 1, 0), ss = box_view(nullptr, slab, 0, 0);
    if (!mb_vel) {
        if (alloc_d(&mb_vel, us.sn * 3) || alloc_d(&mb_sig, ss.sn)) return NP_ERR_CUDA;
        mb_phi.assign(nlocal, nullptr); mb_gphi.assign(nlocal, nullptr);
        for (int b = 0; b < nlocal; ++b)
            if (alloc_d(&mb_phi[b], box_view(nullptr, lbox[b], 1, 1).sn) || alloc_d(&mb_gphi[b], box_view(nullptr, lbox[b], 1, 0).sn * 3)) return NP_ERR_CUDA;
    }
    bool host = false;
#ifndef NP_EMU
    host = !np_rt::is_device_ptr(vel[0].p);
#endif
    // host callers: every box's vel is kept in a persistent device staging area (slot 0) until it is copied back
    mb_stage_off.assign(nlocal + 1, 0);
    long long smax = 0;
    for (int b = 0; b < nlocal; ++b) {
        if (!vel[b].p || !sigma[b].p || vel[b].ngrow < 1 || vel[b].ncomp < 3 || sigma[b].ncomp < 1) return NP_ERR_INVALID;
        mb_stage_off[b + 1] = mb_stage_off[b] + box_view(nullptr, lbox[b], vel[b].ngrow, 0).sn * 3;
        smax = std::max(smax, box_view(nullptr, lbox[b], sigma[b].ngrow, 0).sn);
    }
    double *stv = nullptr, *sts = nullptr;
    if (host && (!(stv = stage_get(0, mb_stage_off[nlocal])) || !(sts = stage_get(1, smax)))) return NP_ERR_CUDA;
    FabView usv = us; usv.p = mb_vel;
    FabView ssv = ss; ssv.p = mb_sig;
    zero(mb_vel, us.sn * 3);
    long long gather_launches = 0;
    for (int b = 0; b < nlocal; ++b) {
        const np_box& bx = lbox[b];
        FabView uv = box_view(vel[b].p, bx, vel[b].ngrow, 0), sv = box_view(sigma[b].p, bx, sigma[b].ngrow, 0);
        if (host) {
            uv.p = stv + mb_stage_off[b]; sv.p = sts;
            np_rt::copy_async(uv.p, vel[b].p, (size_t)(uv.sn * 3) * sizeof(double), stream);
            np_rt::copy_async(sv.p, sigma[b].p, (size_t)sv.sn * sizeof(double), stream);
        }
        int rlo[3], rhi[3];
        for (int d = 0; d < 3; ++d) { rlo[d] = bx.lo[d] - 1; rhi[d] = bx.hi[d] + 1; }
        box_copy(usv, zero3, uv, bx.lo, rlo, rhi, 3, &bx);
        box_copy(ssv, zero3, sv, bx.lo, bx.lo, bx.hi, 1, nullptr);
        gather_launches += 2;
    }
    np_fab vslab{mb_vel, 1, 3}, sslab{mb_sig, 0, 1};
    const int rc = project(&vslab, &sslab, rtol, atol);
    if (rc && !have_result) return rc;
    // ---- scatter: u (valid cells), phi and grad(phi) with their ghost layers ----
    FabView pv; pv.p = phi_out; pv.ng = 1; pv.sy = d0.n[0] + 3; pv.sz = pv.sy * (d0.n[1] + 3); pv.sn = pv.sz * (d0.n[2] + 3);
    FabView gv = us; gv.p = gphi;
    for (int b = 0; b < nlocal; ++b) {
        const np_box& bx = lbox[b];
        FabView uv = box_view(host ? stv + mb_stage_off[b] : vel[b].p, bx, vel[b].ngrow, 0);
        box_copy(uv, bx.lo, usv, zero3, bx.lo, bx.hi, 3, nullptr);
        if (host) np_rt::copy_async(vel[b].p, uv.p, (size_t)(uv.sn * 3) * sizeof(double), stream);
        int rlo[3], rhi[3], nhi[3];
        for (int d = 0; d < 3; ++d) { rlo[d] = bx.lo[d] - 1; rhi[d] = bx.hi[d] + 1; nhi[d] = bx.hi[d] + 2; }
        box_copy(box_view(mb_phi[b], bx, 1, 1), bx.lo, pv, zero3, rlo, nhi, 1, nullptr);
        box_copy(box_view(mb_gphi[b], bx, 1, 0), bx.lo, gv, zero3, rlo, rhi, 3, nullptr);
    }
    np_rt::stream_sync(stream);
    stats.kernel_launches = launches + gather_launches;
    if (np_rt::last_error()) { stats.status = NP_ERR_CUDA; return NP_ERR_CUDA; }
    return rc;
}

// incflo::ApplyProjection for one level in one call: the pointwise passes the caller runs before
// and after NodalProjector::project (/root/reference/src/projection/incflo_apply_nodal_projection.cpp:79-131, :156-203)
// are folded into the solver's first and last kernels; sigma = scaling/rho is never materialised as a caller fab.
int NpSolver::apply_projection(np_fab* vel, const np_fab* vel_old, const np_fab* rho, np_fab* gp, np_fab* p,
                               const np_apply_args* a, double rtol, double atol)
{
    const bool use_old = a->incremental || a->proj_for_small_dt;
    if (vel->ngrow < 1 || vel->ncomp < 3 || rho->ncomp < 1 || gp->ncomp < 3 || p->ncomp < 1) return NP_ERR_INVALID;
    if (use_old && (!vel_old || !vel_old->p || vel_old->ncomp < 3)) return NP_ERR_INVALID;
    // device fabs only (a CUDA build of the caller); a CPU caller keeps its own passes and uses np_project
    if (!np_rt::is_device_ptr(vel->p) || !np_rt::is_device_ptr(rho->p) || !np_rt::is_device_ptr(gp->p) || !np_rt::is_device_ptr(p->p)
        || (use_old && !np_rt::is_device_ptr(vel_old->p))) return NP_ERR_INVALID;
    err = 0; launches = 0; have_result = false; prof_n = 0; g_ms_smooth = 0.0; g_n_smooth = 0;
    ensure_clean();
    const LevelDev& d0 = lev[0].d;
    np_rt::event_record(ev[0], stream);
    FabView uv = view_cell(*vel, 0, 3), rv = view_cell(*rho, 0, 1), gv = view_cell(*gp, 0, 3), pv = view_nodal(*p, 0);
    FabView ov = use_old ? view_cell(*vel_old, 0, 3) : uv;
    // ---- before: u += gp*scaling/rho, u -= u_old, sigma = scaling/rho, setBndry(0) ----
    ApPre pre; pre.scaling = a->scaling_factor; pre.add_gp = a->incremental ? 0 : 1; pre.sub_old = use_old ? 1 : 0;
    NP_LAUNCH(true, k_apply_pre, grid_cells(d0), NP_BLK, 0, stream, d0, uv, ov, rv, gv, lev[0].sigma, pre, part0, part1); ++launches;
    if (a->zero_vel_ghosts) {
        const int g = uv.ng;
        NP_LAUNCH(false, k_zero_cell_ghosts, dim3((d0.n[0] + 2 * g + 31) / 32, (d0.n[1] + 2 * g + 7) / 8, d0.n[2] + 2 * g), NP_BLK, 0, stream,
                  d0.n[0], d0.n[1], d0.n[2], uv); ++launches;
    }
    int rc = finish_sigma(nblocks(grid_cells(d0)), nullptr);
    if (rc) return rc;
    // ---- rhs = div(u*) ----
    np_fill_cell3(*this, uv);
    {
        dim3 gd = grid_nodes(d0); gd.z = std::min<unsigned>(gd.z, 16u);     // 16 runs of node planes per column
        if (singular) { NP_LAUNCH(true, (k_divu<true>), gd, NP_BLK, 0, stream, d0, rhs, uv, part1); divu_npart = nblocks(gd); }
        else          { NP_LAUNCH(true, (k_divu<false>), gd, NP_BLK, 0, stream, d0, rhs, uv, part1); divu_npart = 0; }
        ++launches;
    }
    np_rt::event_record(ev[1], stream);
    rc = solve(rtol, atol);
    np_rt::event_record(ev[2], stream);
    // ---- after: u -= sigma grad(phi) (+ u_old), gp (+)= grad(phi), p (+)= phi ----
    ApPost post; post.add_old = use_old ? 1 : 0; post.accumulate = a->incremental ? 1 : 0;
    NP_LAUNCH(false, k_apply_post, grid_cells(d0), NP_BLK, 0, stream, d0, uv, ov, gv, (const double*)sol, (const double*)lev[0].sigma, post); ++launches;
    fill(0, sol);
    NP_LAUNCH(false, k_apply_post_p, grid_nodes(d0), NP_BLK, 0, stream, d0, (const double*)sol, pv, post.accumulate); ++launches;
    finish_stats();
    if (np_rt::last_error()) { stats.status = NP_ERR_CUDA; return NP_ERR_CUDA; }
    if (err) { stats.status = err; return err; }
    return rc;
}

