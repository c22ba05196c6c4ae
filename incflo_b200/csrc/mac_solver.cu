// mac_solver.cu — host side of the MAC projection (amrex::MacProjector::project -> amrex::MLMG::solve on
// amrex::MLABecLaplacian) and its C ABI (include/macproj.h).  All arithmetic is in mac_kernels.cuh.
#include "np_solver.h"
#include "mac_kernels.cuh"
#include "../../include/macproj.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <new>
#include <vector>

enum { MP_RESTART = -77 };        // MacSolver::solve: a bottom solve inside the iteration graph needed MLMG's fallback path

struct MacLev {
    MacLevel d{};
    double *bx = nullptr, *by = nullptr, *bz = nullptr;
    double *res = nullptr, *cor = nullptr, *rescor = nullptr;   // every cell array has one spare plane below and above (ghost planes
                                                                 // of a slab: k = -1 and k = n2)
    double* a = nullptr;          // MLABecLaplacian's a coefficient (cell array without ghost planes), allocated by set_acoeffs
    bool replicated = false;      // multi-GPU: this level holds the whole domain on every rank
    MacBeta beta() const { return MacBeta{bx, by, bz}; }
};

struct MacSolver {
    mp_options opt{};
    np_geom geom{};
    int bclo[3] = {0, 0, 0}, bchi[3] = {0, 0, 0};
    NpComm comm;                  // z-slab sharding: halo planes, reductions, the gather onto the replicated coarse levels
    int agg_lev = -1;             // first replicated level, -1: none (one rank)
    double *agg_send = nullptr, *agg_recv = nullptr;
    std::vector<MacLev> lev;
    double *sol = nullptr, *rhs = nullptr;
    double *part0 = nullptr, *part1 = nullptr; long long npart = 0;
    double *scal = nullptr, *h_scal = nullptr;
    int *bot_out = nullptr, *h_bot_out = nullptr;
    double* bw[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    double* stage[4] = {nullptr, nullptr, nullptr, nullptr}; long long stage_n[4] = {0, 0, 0, 0};
    double* wrapbuf = nullptr; long long wrapbuf_n = 0;
    std::vector<np_box> lbox;                            // the caller's boxes (one rank: they tile the domain)
    double* mb_f[3] = {nullptr, nullptr, nullptr}; double* mb_phi = nullptr;   // several boxes: domain-shaped gather buffers
    std::vector<np_event_t> prof_ev; int prof_n = 0;     // profile_kernels: start/stop pairs around the level-0 smooths
    np_stream_t stream = 0;
    np_event_t ev[4] = {0, 0, 0, 0};
    np_stats stats{};
    bool singular = true, have_beta = false, have_result = false, bicg_failed_switch = false;
    long long launches = 0;
    int bottom_iters = 0;
    int err = 0;
    double alpha = 0.0, bscal = 1.0;      // MLABecLaplacian::setScalars (the MAC projection: 0, 1)
    bool have_a = false;
    // mp_options.use_graph: one MLMG iteration as a CUDA graph, keyed on what it bakes in (as NpSolver::iterate_graph); one rank
    struct IterGraph { np_rt::np_graph_t exec = nullptr; std::vector<unsigned char> key; long long launches = 0; int n_prof = 0; } itg;
    np_stream_t cap_stream = 0; bool have_cap_stream = false, in_capture = false, graph_off = false, graph_warm = false;
    std::vector<np_event_t> gprof_ev; double g_ms_smooth = 0.0; int g_n_smooth = 0;
    void drop_graph() { if (itg.exec) np_rt::graph_destroy(itg.exec); itg = IterGraph{}; }

    static dim3 blk() { return dim3(32, 8, 1); }
    static dim3 grid_cells(const MacLevel& L) { return dim3((L.n[0] + 31) / 32, (L.n[1] + 7) / 8, L.n[2]); }
    static dim3 grid_pairs(const MacLevel& L) { return dim3(((L.n[0] + 1) / 2 + 31) / 32, (L.n[1] + 7) / 8, L.n[2]); }
    static dim3 grid_adotx(const MacLevel& L) { return dim3(((L.n[0] + 1) / 2 + 31) / 32, (L.n[1] + 7) / 8, (L.n[2] + MAC_KZ - 1) / MAC_KZ); }
    static long long nblk(dim3 g) { return (long long)g.x * g.y * g.z; }
    static int alloc_d(double** p, long long n) { void* q = nullptr; if (np_rt::dev_malloc(&q, (size_t)n * sizeof(double))) return 1; *p = (double*)q; return 0; }
    void zero(double* p, long long n) { np_rt::memset_async(p, 0, (size_t)n * sizeof(double), stream); }
    // cell array of a level with two ghost planes on either side in z (planes -2, -1 and n2, n2+1 of a slab); the returned
    // pointer is plane 0
    int alloc_cell(double** p, const MacLevel& d)
    {
        double* q = nullptr;
        if (alloc_d(&q, d.ntot + 4 * d.sxy)) return 1;
        np_rt::memset_async(q, 0, (size_t)(d.ntot + 4 * d.sxy) * sizeof(double), stream);
        *p = q + 2 * d.sxy;
        return 0;
    }
    static void free_cell(double* p, const MacLevel& d) { if (p) np_rt::dev_free(p - 2 * d.sxy); }
    bool sharded(int l) const { return comm.nranks > 1 && !lev[l].replicated; }
    void z_neighbours(int& lo, int& hi) const
    {
        const int r = comm.rank, R = comm.nranks;
        const bool per = bclo[2] == NP_BC_PERIODIC;
        lo = r > 0 ? r - 1 : (per ? R - 1 : -1);
        hi = r < R - 1 ? r + 1 : (per ? 0 : -1);
    }
    // Peer-memory halo exchange (the nodal solver's: NpSolver::setup_peer_halo, k_halo_p2p): every rank exports the CUDA-IPC handles
    // of the cell arrays that are exchanged inside a solve (cor, rescor, res of the sharded levels, sol) and of its flag words,
    // all ranks gather them, each rank maps its two z-neighbours'.  Used only if EVERY rank succeeded; else grouped ncclSend/ncclRecv.
    NpSolver::PeerHalo peer;
    std::vector<long long> peer_off;         // offset of a registered array from the base of its allocation (its two low ghost planes)
    int setup_peer()
    {
        if (!comm.allgather || !comm.allreduce) return 1;
        auto reg = [&](double* p, const MacLevel& d) { peer.mine.push_back(p); peer_off.push_back(2 * d.sxy); };
        for (auto& L : lev) if (!L.replicated) { reg(L.cor, L.d); reg(L.rescor, L.d); reg(L.res, L.d); }
        reg(sol, lev[0].d);
        void* q = nullptr;
        if (np_rt::dev_malloc(&q, 8 * sizeof(unsigned long long))) return 1;
        peer.flags = (unsigned long long*)q; np_rt::memset_async(peer.flags, 0, 8 * sizeof(unsigned long long), stream);
        if (np_rt::dev_malloc(&q, 4 * sizeof(unsigned int))) return 1;
        peer.counter = (unsigned int*)q; np_rt::memset_async(peer.counter, 0, 4 * sizeof(unsigned int), stream);
        const int na = (int)peer.mine.size() + 1;                       // + the flag words
        std::vector<double> hsend((size_t)na * 8, 0.0), hrecv((size_t)na * 8 * comm.nranks, 0.0);
        int good = 1;
        for (int a = 0; a < na; ++a) {
            void* p = a < na - 1 ? (void*)(peer.mine[a] - peer_off[a]) : (void*)peer.flags;
            if (np_rt::ipc_export(p, &hsend[(size_t)a * 8])) good = 0;
        }
        double *dsend = nullptr, *drecv = nullptr;
        if (alloc_d(&dsend, (long long)na * 8) || alloc_d(&drecv, (long long)na * 8 * comm.nranks)) return 1;
        np_rt::copy_async(dsend, hsend.data(), hsend.size() * sizeof(double), stream);
        if (comm.allgather(comm.ctx, dsend, drecv, (long long)na * 8, (void*)stream)) { err = NP_ERR_NCCL; good = 0; }
        np_rt::copy_async(hrecv.data(), drecv, hrecv.size() * sizeof(double), stream);
        np_rt::stream_sync(stream);
        int lo, hi; z_neighbours(lo, hi);
        peer.lo.assign(na - 1, nullptr); peer.hi.assign(na - 1, nullptr);
        auto open_rank = [&](int r, std::vector<double*>& arr, unsigned long long** fl) {
            for (int a = 0; a < na && good; ++a) {
                void* m = nullptr;
                if (np_rt::ipc_open(&hrecv[((size_t)r * na + a) * 8], &m)) { good = 0; break; }
                peer.opened.push_back(m);
                if (a < na - 1) arr[a] = (double*)m + peer_off[a]; else *fl = (unsigned long long*)m;
            }
        };
        if (good && lo >= 0) open_rank(lo, peer.lo, &peer.lo_flags);
        if (good && hi >= 0) {
            if (hi == lo) { peer.hi = peer.lo; peer.hi_flags = peer.lo_flags; }      // two ranks, periodic: one neighbour on both sides
            else open_rank(hi, peer.hi, &peer.hi_flags);
        }
        h_scal[40] = good ? 0.0 : 1.0;                                  // all or nothing: one max-allreduce of "somebody failed"
        np_rt::copy_async(scal + 40, h_scal + 40, sizeof(double), stream);
        if (comm.allreduce(comm.ctx, scal + 40, 1, 1, (void*)stream)) good = 0;
        np_rt::copy_async(h_scal + 40, scal + 40, sizeof(double), stream);
        np_rt::stream_sync(stream);
        np_rt::dev_free(dsend); np_rt::dev_free(drecv);
        peer.ok = good && h_scal[40] == 0.0;
        np_rt::last_error();
        return peer.ok ? 0 : 1;
    }
    bool halo_p2p(const MacLevel& d, double* x, int depth)
    {
        if (!peer.ok) return false;
        int id = -1;
        for (int a = 0; a < (int)peer.mine.size(); ++a) if (peer.mine[a] == x) { id = a; break; }
        if (id < 0) return false;
        const long long m = d.sxy * depth;
        int lo, hi; z_neighbours(lo, hi);
        HaloP2P a{};
        // to the lower rank: my planes 0 .. depth-1 -> its planes n2 ..; to the upper rank: my planes n2-depth .. n2-1 -> its planes -depth .. -1
        a.send_lo = x; a.n_lo = m; a.dst_lo = lo >= 0 ? peer.lo[id] + d.sxy * (long long)d.n[2] : nullptr;
        a.send_hi = x + d.sxy * (long long)d.n[2] - m; a.n_hi = m; a.dst_hi = hi >= 0 ? peer.hi[id] - m : nullptr;
        a.mine = peer.flags; a.lo_flags = lo >= 0 ? peer.lo_flags : nullptr; a.hi_flags = hi >= 0 ? peer.hi_flags : nullptr;
        a.epoch_word = peer.flags + 4; a.counter = peer.counter;
        const long long chunks = (m + 255) / 256;
        const int nb = (int)std::min<long long>(64, std::max<long long>(1, chunks / 8));
        NP_LAUNCH(false, k_halo_p2p, dim3(nb), dim3(256), 0, stream, a); ++launches;
        return true;
    }
    // ghost planes of a cell array of a sharded level, `depth` (1 or 2) planes per side: my planes 0.. -> the lower rank's planes
    // n2.., my planes ..n2-1 -> the upper rank's planes ..-1
    void halo(int l, double* x, int depth = 1)
    {
        if (!sharded(l)) return;
        const MacLevel& d = lev[l].d;
        if (halo_p2p(d, x, depth)) return;
        const long long m = d.sxy * depth;
        int lo, hi; z_neighbours(lo, hi);
        if (comm.sendrecv(comm.ctx, x, m, x - m, m, x + d.sxy * (long long)d.n[2] - m, m, x + d.sxy * (long long)d.n[2], m,
                          lo, hi, (void*)stream)) err = NP_ERR_NCCL;
    }

    void setup_bc()
    {
        singular = true;
        for (int d = 0; d < 3; ++d) if (bclo[d] == NP_BC_DIRICHLET || bchi[d] == NP_BC_DIRICHLET) singular = false;
        if (alpha != 0.0 && have_a) singular = false;
        for (auto& L : lev) for (int d = 0; d < 3; ++d) {
            L.d.bclo[d] = bclo[d]; L.d.bchi[d] = bchi[d];
            if (d == 2 && comm.nranks > 1 && !L.replicated) {      // slab faces shared with a neighbouring rank (periodic z: all of them)
                const bool per = bclo[2] == NP_BC_PERIODIC;
                if (per || comm.rank > 0) L.d.bclo[2] = MAC_REMOTE;
                if (per || comm.rank < comm.nranks - 1) L.d.bchi[2] = MAC_REMOTE;
            }
            const bool hi_order = std::min(L.d.n[d] + 1, opt.maxorder) >= 3;     // mllinop_apply_bc: NX = min(blen+1, maxorder)
            L.d.dc0[d] = hi_order ? -2.0 : -1.0; L.d.dc1[d] = hi_order ? 1.0 / 3.0 : 0.0;
            auto fac = [&](int t) { return (t == NP_BC_PERIODIC || t == MAC_REMOTE) ? 0.0 : t == NP_BC_NEUMANN ? 1.0 : L.d.dc0[d]; };
            L.d.flo[d] = fac(L.d.bclo[d]); L.d.fhi[d] = fac(L.d.bchi[d]);
        }
    }

    int create(const np_geom* g, int nboxes, const np_box* boxes, const int* owner, int my_rank, int nranks, const NpComm& c, const mp_options* o)
    {
        if (np_rt::init()) return NP_ERR_NO_DEVICE;
        geom = *g; opt = *o; comm = c; comm.rank = my_rank; comm.nranks = nranks;
        if (const char* e = std::getenv("NP_MAC_FUSED_MIN_N")) opt.fused_rb_min_n = std::atoi(e);      // experiment knob (A/B of the one-pass smoother)
        if (nranks < 1 || my_rank < 0 || my_rank >= nranks || nboxes < 1) return NP_ERR_INVALID;
        for (int d = 0; d < 3; ++d) if (g->ncell[d] < 2) return NP_ERR_INVALID;
        int nloc[3] = {g->ncell[0], g->ncell[1], g->ncell[2]};
        int koff = 0;
        if (nranks > 1) {
            // one box per rank: equal z-slabs ordered by rank (the nodal solver's layout)
            const int nzl = g->ncell[2] / nranks;
            if (nboxes != nranks || nzl * nranks != g->ncell[2] || (nzl & 1)) return NP_ERR_INVALID;
            for (int r = 0; r < nranks; ++r) {
                const np_box& b = boxes[r];
                if (owner[r] != r || b.lo[0] != 0 || b.lo[1] != 0 || b.hi[0] != g->ncell[0] - 1 || b.hi[1] != g->ncell[1] - 1
                    || b.lo[2] != r * nzl || b.hi[2] != (r + 1) * nzl - 1) return NP_ERR_INVALID;
            }
            nloc[2] = nzl; koff = my_rank * nzl;
            np_box mine = boxes[my_rank]; mine.lo[2] = 0; mine.hi[2] = nzl - 1;
            lbox.assign(1, mine);
        } else {   // one rank: the boxes must tile the domain: inside it, pairwise disjoint, volumes adding up
            long long vol = 0;
            lbox.assign(boxes, boxes + nboxes);
            for (int b = 0; b < nboxes; ++b) {
                if (owner[b] != 0) return NP_ERR_INVALID;
                long long v = 1;
                for (int d = 0; d < 3; ++d) {
                    if (boxes[b].lo[d] < 0 || boxes[b].hi[d] >= g->ncell[d] || boxes[b].lo[d] > boxes[b].hi[d]) return NP_ERR_INVALID;
                    v *= boxes[b].hi[d] - boxes[b].lo[d] + 1;
                }
                vol += v;
                for (int a = 0; a < b; ++a) {
                    bool ov = true;
                    for (int d = 0; d < 3; ++d) if (boxes[a].hi[d] < boxes[b].lo[d] || boxes[b].hi[d] < boxes[a].lo[d]) ov = false;
                    if (ov) return NP_ERR_INVALID;
                }
            }
            if (vol != (long long)g->ncell[0] * g->ncell[1] * g->ncell[2]) return NP_ERR_INVALID;
        }
        for (int d = 0; d < 3; ++d) bclo[d] = bchi[d] = g->periodic[d] ? NP_BC_PERIODIC : NP_BC_NEUMANN;
        // ---- MG levels (MLLinOp::defineGrids on the GLOBAL box: coarsen by 2 while it stays coarsenable, min width 2).  A slab is
        //      coarsened locally while it keeps at least agg_min_local_nz planes; below that the level is replicated on every
        //      rank (one all-gather on the way down, nothing on the way up), so the hierarchy is the one-box hierarchy ----
        int n[3] = {nloc[0], nloc[1], nloc[2]};
        int nglob2 = g->ncell[2];
        double dxinv[3];
        for (int d = 0; d < 3; ++d) dxinv[d] = g->ncell[d] / (g->prob_hi[d] - g->prob_lo[d]);
        bool replicated = false;
        int ko = koff;
        for (;;) {
            MacLev L;
            for (int d = 0; d < 3; ++d) { L.d.n[d] = n[d]; L.d.dxinv[d] = dxinv[d]; L.d.dh[d] = dxinv[d] * dxinv[d]; }
            L.d.sx = (n[0] + 1) & ~1LL; L.d.sxy = L.d.sx * n[1]; L.d.ntot = L.d.sxy * n[2] + 2;
            L.d.fxs = (n[0] + 2) & ~1LL;
            L.d.const_beta = 0; L.d.beta0 = 1.0; L.d.koff = ko; L.d.kcoff = 0;
            L.replicated = replicated;
            lev.push_back(L);
            if ((int)lev.size() > opt.max_coarsening_level) break;
            const int gn[3] = {n[0], n[1], replicated || nranks == 1 ? n[2] : nglob2};
            bool ok = true;
            for (int d = 0; d < 3; ++d) if (gn[d] % 2 || gn[d] < 4) ok = false;
            if (!ok) break;
            bool to_replica = false;
            if (nranks > 1 && !replicated) {
                if (n[2] % 2) return NP_ERR_INVALID;            // a slab that cannot be halved cannot hand its part of the coarse level over
                if (n[2] / 2 < opt.agg_min_local_nz || (n[2] / 2) % 2) to_replica = true;
            }
            if (to_replica) {
                replicated = true; agg_lev = (int)lev.size();
                lev.back().d.kcoff = ko / 2;                     // parents of this slab start at plane koff/2 of the replicated level
                n[2] = nglob2 / 2; ko = 0;
            } else { n[2] /= 2; ko /= 2; }
            n[0] /= 2; n[1] /= 2; nglob2 /= 2;
            for (int d = 0; d < 3; ++d) dxinv[d] *= 0.5;
        }
        if (nranks > 1 && !lev.back().replicated) return NP_ERR_INVALID;   // the bottom solver needs the whole coarsest level
        for (auto& L : lev) {
            const MacLevel& d = L.d;
            if (alloc_cell(&L.res, d) || alloc_cell(&L.cor, d) || alloc_cell(&L.rescor, d)) return NP_ERR_CUDA;
            // face coefficients: one spare plane below and above (ghost planes of a slab for the one-pass smoother)
            const long long px_ = d.fxs * d.n[1], py_ = d.sx * (d.n[1] + 1), pz_ = d.sxy;
            if (alloc_d(&L.bx, px_ * (d.n[2] + 2) + 2) || alloc_d(&L.by, py_ * (d.n[2] + 2) + 2) || alloc_d(&L.bz, pz_ * (d.n[2] + 3) + 2)) return NP_ERR_CUDA;
            zero(L.bx, px_ * (d.n[2] + 2) + 2); zero(L.by, py_ * (d.n[2] + 2) + 2); zero(L.bz, pz_ * (d.n[2] + 3) + 2);
            L.bx += px_; L.by += py_; L.bz += pz_;
        }
        const MacLevel& d0 = lev[0].d;
        if (alloc_cell(&sol, d0) || alloc_cell(&rhs, d0)) return NP_ERR_CUDA;
        dim3 gmax((d0.n[0] + 32) / 32, (d0.n[1] + 8) / 8, d0.n[2] + 1);
        npart = nblk(gmax) + 16;
        if (alloc_d(&part0, npart) || alloc_d(&part1, npart) || alloc_d(&scal, 64)) return NP_ERR_CUDA;
        zero(scal, 64);
        void* q = nullptr;
        if (np_rt::host_malloc(&q, 64 * sizeof(double))) return NP_ERR_CUDA;
        h_scal = (double*)q;
        if (np_rt::dev_malloc(&q, 8 * sizeof(int))) return NP_ERR_CUDA;
        bot_out = (int*)q;
        if (np_rt::host_malloc(&q, 8 * sizeof(int))) return NP_ERR_CUDA;
        h_bot_out = (int*)q;
        for (int w = 0; w < 6; ++w) { if (alloc_d(&bw[w], lev.back().d.ntot)) return NP_ERR_CUDA; zero(bw[w], lev.back().d.ntot); }
        if (agg_lev >= 0) {
            // staging for the gather onto the first replicated level: this rank's part of that level (cells, or one of its face
            // arrays with the extra row / plane), and the parts of all ranks
            const MacLevel& C = lev[agg_lev].d;
            const long long nzc = C.n[2] / nranks;
            const long long m = std::max(C.fxs, C.sx) * (long long)(C.n[1] + 1) * (nzc + 1) + 2;
            if (alloc_d(&agg_send, m) || alloc_d(&agg_recv, m * nranks)) return NP_ERR_CUDA;
            zero(agg_send, m); zero(agg_recv, m * nranks);
        }
        for (int e = 0; e < 4; ++e) np_rt::event_create(&ev[e]);
        if (opt.profile_kernels) { prof_ev.resize(4096); for (auto& e : prof_ev) np_rt::event_create(&e); }
        setup_bc();
        np_rt::stream_sync(stream);
        if (nranks > 1 && opt.p2p_halo) setup_peer();        // collective; falls back to NCCL when any rank cannot map its neighbours
        return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
    }
    void destroy()
    {
        np_rt::stream_sync(stream);
        if (!lev.empty()) { free_cell(sol, lev[0].d); free_cell(rhs, lev[0].d); }
        for (auto& L : lev) {
            if (L.bx) np_rt::dev_free(L.bx - L.d.fxs * L.d.n[1]);
            if (L.by) np_rt::dev_free(L.by - L.d.sx * (L.d.n[1] + 1));
            if (L.bz) np_rt::dev_free(L.bz - L.d.sxy);
            free_cell(L.res, L.d); free_cell(L.cor, L.d); free_cell(L.rescor, L.d);
            if (L.a) np_rt::dev_free(L.a);
        }
        lev.clear();
        double* a[] = {agg_send, agg_recv, mb_f[0], mb_f[1], mb_f[2], mb_phi, wrapbuf, part0, part1, scal, stage[0], stage[1], stage[2], stage[3], bw[0], bw[1], bw[2], bw[3], bw[4], bw[5]};
        for (double* p : a) if (p) np_rt::dev_free(p);
        if (h_scal) np_rt::host_free(h_scal);
        if (bot_out) np_rt::dev_free(bot_out);
        if (h_bot_out) np_rt::host_free(h_bot_out);
        for (int e = 0; e < 4; ++e) if (ev[e]) np_rt::event_destroy(ev[e]);
        for (auto& e : prof_ev) np_rt::event_destroy(e);
        prof_ev.clear();
        drop_graph();
        for (void* m : peer.opened) np_rt::ipc_close(m);
        peer.opened.clear(); peer.ok = false;
        if (peer.flags) { np_rt::dev_free(peer.flags); peer.flags = nullptr; }
        if (peer.counter) { np_rt::dev_free(peer.counter); peer.counter = nullptr; }
        for (auto& e : gprof_ev) np_rt::event_destroy(e);
        gprof_ev.clear();
        if (have_cap_stream) { np_rt::stream_destroy(cap_stream); have_cap_stream = false; }
    }

    // ---- caller fabs ----
    // view of a fab whose valid range starts at local index 0; nodal_dir = -1 (cell) or the direction a face fab is nodal in
    FabView view(const np_fab& f, int l, int nodal_dir) const
    {
        const MacLevel& d = lev[l].d;
        FabView v; v.p = f.p; v.ng = f.ngrow;
        v.sy = d.n[0] + (nodal_dir == 0) + 2 * f.ngrow;
        v.sz = v.sy * (d.n[1] + (nodal_dir == 1) + 2 * f.ngrow);
        v.sn = v.sz * (d.n[2] + (nodal_dir == 2) + 2 * f.ngrow);
        return v;
    }
    double* stage_get(int slot, long long n)
    {
        if (stage_n[slot] < n) {
            if (stage[slot]) { np_rt::stream_sync(stream); np_rt::dev_free(stage[slot]); stage[slot] = nullptr; stage_n[slot] = 0; }
            if (alloc_d(&stage[slot], n)) return nullptr;
            stage_n[slot] = n;
        }
        return stage[slot];
    }
    // device-side view of a caller fab: device pointers are used in place, host fabs are staged (copied in when `in`)
    bool import(const np_fab& f, int l, int nodal_dir, int slot, bool in, FabView* v, bool* staged)
    {
        *v = view(f, l, nodal_dir); *staged = false;
#ifndef NP_EMU
        if (!np_rt::is_device_ptr(f.p)) {
            double* dev = stage_get(slot, v->sn);
            if (!dev) return false;
            if (in) np_rt::copy_async(dev, f.p, (size_t)v->sn * sizeof(double), stream);
            v->p = dev; *staged = true;
        }
#endif
        return true;
    }
    void export_back(const np_fab& f, const FabView& v, bool staged)
    {
        if (staged) np_rt::copy_async(f.p, v.p, (size_t)v.sn * sizeof(double), stream);
    }

    // ---- building blocks ----
    void gsrb(int l, double* phi, const double* b, int colour, bool zero_in, double* add_to, const double* boff)
    {
        const MacLev& L = lev[l];
        // a level with an odd number of cells in a periodic direction (only the coarsest can be): the periodic image of a cell
        // has the cell's own colour, so images are read from a snapshot - the ghost cells applyBC filled before the sweep
        const double* xw = phi;
        bool odd_per = false;
        for (int d = 0; d < 3; ++d) if (bclo[d] == NP_BC_PERIODIC && (L.d.n[d] & 1)) odd_per = true;
        if (odd_per && !zero_in) {
            if (!wrapbuf || wrapbuf_n < L.d.ntot) { if (wrapbuf) { np_rt::stream_sync(stream); np_rt::dev_free(wrapbuf); } wrapbuf = nullptr; wrapbuf_n = 0; if (alloc_d(&wrapbuf, L.d.ntot)) { err = NP_ERR_CUDA; return; } wrapbuf_n = L.d.ntot; }
            np_rt::copy_async(wrapbuf, phi, (size_t)L.d.ntot * sizeof(double), stream);
            xw = wrapbuf;
        }
        if (L.d.const_beta) NP_LAUNCH(false, (k_mac_gsrb<true>), grid_pairs(L.d), blk(), 0, stream, L.d, L.beta(), phi, b, colour, zero_in ? 1 : 0, add_to, boff, xw);
        else NP_LAUNCH(false, (k_mac_gsrb<false>), grid_pairs(L.d), blk(), 0, stream, L.d, L.beta(), phi, b, colour, zero_in ? 1 : 0, add_to, boff, xw);
        ++launches;
    }
    // MLCellLinOp::smooth: red then black (the boundary conditions are evaluated inside the kernels)
    void smooth(int l, double* phi, const double* b, bool zero_in, double* add_to, const double* boff)
    {
        if (!zero_in) halo(l, phi);                // applyBC's FillBoundary across the slab faces, before each colour
        gsrb(l, phi, b, 0, zero_in, nullptr, boff);
        halo(l, phi);
        gsrb(l, phi, b, 1, false, add_to, boff);
    }
    // levels large enough to fill the GPU run both colours in one out-of-place pass (k_mac_rb2)
    bool fused_ok(int l) const
    {
        const MacLevel& d = lev[l].d;
        // alpha * a (the diffusion solves): the whole-tile kernel only, and only where it pays - measured at 512^3 / 256^3 with
        // Rayleigh-Taylor density: one pass instead of a launch per colour is 2.20 against 2.73 ms per smooth on a 512^3 level with
        // variable b (1.74 / 1.79 uniform) but 0.247 against 0.224 ms on a 256^3 level: the extra coefficient stream costs the
        // latency-bound one-pass kernel more than the second pass costs the streaming one there
        if (d.acoef && (!(d.n[0] % RB_TX == 0 && d.n[1] % RB_TY == 0) || d.n[0] < 2 * opt.fused_rb_min_n)) return false;
        if (opt.fused_rb_min_n <= 0 || d.n[0] < opt.fused_rb_min_n || (d.n[0] & 1)) return false;
        // a slab runs the one-pass smoother on two ghost planes per side (the reds of the planes next to the slab are recomputed
        // there, from the right-hand side and the coefficients of one ghost plane); only the hoisted-index kernel knows about
        // ghost planes, so other slabs exchange one plane before each colour and use one launch per colour
        if (sharded(l) && !(d.n[0] % RB_TX == 0 && d.n[1] % RB_TY == 0 && d.n[2] >= 4)) return false;
        if (d.const_beta && !opt.fused_rb_const) return false;
        for (int q = 0; q < 3; ++q) if (bclo[q] == NP_BC_PERIODIC && (d.n[q] & 1)) return false;
        return true;
    }
    // one smooth of lev[l].cor against lev[l].res; with crse: cor += interpolation(crse) first (addInterpCorrection).
    // The fused pass writes into lev[l].rescor and the two pointers are swapped, so lev[l].cor is always the current iterate.
    void smooth_level(int l, bool zero_in, const double* crse, double* add_to, const double* boff)
    {
        const bool prof = opt.profile_kernels && !in_capture && l == 0 && prof_n + 2 <= (int)prof_ev.size();
        const bool gprof = opt.profile_kernels && in_capture && l == 0 && itg.n_prof + 1 <= (int)gprof_ev.size() / 2;
        if (prof) np_rt::event_record(prof_ev[prof_n++], stream);
        if (gprof) np_rt::event_record_in_graph(gprof_ev[2 * itg.n_prof], stream);
        smooth_level_impl(l, zero_in, crse, add_to, boff);
        if (prof) np_rt::event_record(prof_ev[prof_n++], stream);
        if (gprof) { np_rt::event_record_in_graph(gprof_ev[2 * itg.n_prof + 1], stream); ++itg.n_prof; }
    }
    void smooth_level_impl(int l, bool zero_in, const double* crse, double* add_to, const double* boff)
    {
        MacLev& L = lev[l];
        if (!fused_ok(l)) {
            if (crse) interp_add(l, L.cor, crse);
            smooth(l, L.cor, L.res, zero_in, add_to, boff);
            return;
        }
        if (sharded(l)) {                        // interpolation as a pass of its own (its parents may live on another rank's
            if (crse) { interp_add(l, L.cor, crse); crse = nullptr; }      // planes), then both ghost planes of the iterate
            if (!zero_in) halo(l, L.cor, 2);
        }
        RbArgs A;
        A.out = L.rescor; A.in = L.cor; A.rhs = L.res; A.boff = boff; A.add_to = add_to; A.crse = crse; A.zero_in = zero_in ? 1 : 0;
        A.zc = std::min(32, std::max(4, L.d.n[2] / 8));
        const dim3 g((L.d.n[0] + RB_TX - 1) / RB_TX, (L.d.n[1] + RB_TY - 1) / RB_TY, (L.d.n[2] + A.zc - 1) / A.zc);
        const MacLevel& C = crse ? lev[l + 1].d : L.d;
        const bool whole_tiles = L.d.n[0] % RB_TX == 0 && L.d.n[1] % RB_TY == 0;      // hoisted-index variant
        if (whole_tiles && L.d.acoef) {          // MLABecLaplacian with alpha * a (the diffusion solves)
            if (L.d.const_beta) NP_LAUNCH(true, (k_mac_rb3<true, true>), g, dim3(256), 0, stream, L.d, C, L.beta(), A);
            else NP_LAUNCH(true, (k_mac_rb3<false, true>), g, dim3(256), 0, stream, L.d, C, L.beta(), A);
        } else if (whole_tiles) {
            if (L.d.const_beta) NP_LAUNCH(true, (k_mac_rb3<true>), g, dim3(256), 0, stream, L.d, C, L.beta(), A);
            else NP_LAUNCH(true, (k_mac_rb3<false>), g, dim3(256), 0, stream, L.d, C, L.beta(), A);
        } else if (L.d.const_beta) NP_LAUNCH(true, (k_mac_rb2<true>), g, dim3(256), 0, stream, L.d, C, L.beta(), A);
        else NP_LAUNCH(true, (k_mac_rb2<false>), g, dim3(256), 0, stream, L.d, C, L.beta(), A);
        ++launches;
        std::swap(L.cor, L.rescor);
    }
    template <int MODE, bool REDUCE> void adotx(int l, double* y, const double* x, const double* b, const double* boff)
    {
        const MacLev& L = lev[l];
        if (L.d.const_beta) NP_LAUNCH(REDUCE, (k_mac_adotx<MODE, REDUCE, true>), grid_adotx(L.d), blk(), 0, stream, L.d, L.beta(), y, x, b, boff, part0, part1);
        else NP_LAUNCH(REDUCE, (k_mac_adotx<MODE, REDUCE, false>), grid_adotx(L.d), blk(), 0, stream, L.d, L.beta(), y, x, b, boff, part0, part1);
        ++launches;
    }
    // computeResOfCorrection + restriction: lev[l+1].res = average_down(lev[l].res - off - A lev[l].cor).  From the last sharded
    // level onto the first replicated one: every rank restricts its slab into the staging buffer and one all-gather puts the
    // slabs together on every rank (planes are contiguous, slabs are stacked in rank order).
    void resid_restrict(int l, const double* boff)
    {
        const MacLev& L = lev[l]; const MacLev& C = lev[l + 1];
        halo(l, L.cor);
        MacLevel Cl = C.d;                       // shape of the part of the coarse level this rank produces
        double* out = C.res;
        const bool gather = sharded(l) && C.replicated;
        if (gather) { Cl.n[2] = L.d.n[2] / 2; out = agg_send; }
        if (L.d.const_beta) NP_LAUNCH(false, (k_mac_resid_restrict<true>), grid_cells(Cl), blk(), 0, stream, Cl, L.d, L.beta(), out, (const double*)L.cor, (const double*)L.res, boff);
        else NP_LAUNCH(false, (k_mac_resid_restrict<false>), grid_cells(Cl), blk(), 0, stream, Cl, L.d, L.beta(), out, (const double*)L.cor, (const double*)L.res, boff);
        ++launches;
        if (gather && comm.allgather(comm.ctx, agg_send, C.res, Cl.sxy * (long long)Cl.n[2], (void*)stream)) err = NP_ERR_NCCL;
        halo(l + 1, C.res);                      // the one-pass smoother relaxes the reds of the first ghost plane too
    }
    // fold the partials of the last reducing kernel over level-0 cells: scal[base] = max, scal[base+2] = sum, scal[base+6] = mean
    void reduce(long long n, bool want_max, bool want_sum, int base)
    {
        const MacLevel& d0 = lev[0].d;
        const double ncells = (double)geom.ncell[0] * geom.ncell[1] * geom.ncell[2];
        NP_LAUNCH(true, k_mac_reduce, dim3(1), dim3(1024), 0, stream, want_max ? (const double*)part0 : (const double*)nullptr,
                  want_sum ? (const double*)part1 : (const double*)nullptr, n, scal + base, ncells);
        ++launches;
        (void)d0;
        if (comm.nranks > 1) {                    // level 0 is sharded: max and sum over the ranks, then the mean of the whole domain
            if (want_max && comm.allreduce(comm.ctx, scal + base, 1, 1, (void*)stream)) err = NP_ERR_NCCL;
            if (want_sum) {
                if (comm.allreduce(comm.ctx, scal + base + 2, 1, 0, (void*)stream)) err = NP_ERR_NCCL;
                NP_LAUNCH(false, k_div_scalar, dim3(1), dim3(32), 0, stream, scal + base, 2, 6, ncells); ++launches;
            }
        }
    }
    double fetch(int slot)
    {
        np_rt::copy_async(h_scal, scal, 16 * sizeof(double), stream);
        if (in_capture) return 0.0;          // read after the graph has run
        np_rt::stream_sync(stream);
        return h_scal[slot];
    }
    void restrict_to(int lc, double* crse, const double* fine)
    {
        NP_LAUNCH(false, k_mac_restrict, grid_cells(lev[lc].d), blk(), 0, stream, lev[lc].d, lev[lc - 1].d, crse, fine); ++launches;
    }
    void interp_add(int lf, double* fine, const double* crse)
    {
        const MacLevel& F = lev[lf].d;
        NP_LAUNCH(false, k_mac_interp_add, dim3((lev[lf + 1].d.n[0] + 31) / 32, (F.n[1] + 7) / 8, F.n[2]), blk(), 0, stream, F, lev[lf + 1].d, fine, crse); ++launches;
    }
    int run_bottom(bool use_cg, int* iters)
    {
        MacLev& L = lev.back();
        MacBottomArgs A;
        A.L = L.d; A.B = L.beta(); A.x = L.cor; A.b = L.res;
        A.r = bw[0]; A.rh = bw[1]; A.p = bw[2]; A.v = bw[3]; A.s = bw[4]; A.t = bw[5];
        A.maxiter = opt.bottom_maxiter; A.use_cg = use_cg ? 1 : 0; A.singular = singular ? 1 : 0;
        A.eps_rel = opt.bottom_rtol; A.eps_abs = opt.bottom_atol; A.out = bot_out;
        const long long nc = (long long)L.d.n[0] * L.d.n[1] * L.d.n[2];
        const int nth = nc <= 32 ? 32 : nc <= 256 ? 256 : 1024;
        if (L.d.const_beta) NP_LAUNCH(true, (k_mac_bottom<true>), dim3(1), dim3(nth), 0, stream, A);
        else NP_LAUNCH(true, (k_mac_bottom<false>), dim3(1), dim3(nth), 0, stream, A);
        ++launches;
        np_rt::copy_async(h_bot_out, bot_out, 2 * sizeof(int), stream);
        if (in_capture) { *iters = 0; return 0; }      // looked at after the graph has run (iterate_graph)
        np_rt::stream_sync(stream);
        *iters = h_bot_out[1];
        return h_bot_out[0];
    }
    // MLMG::bottomSolve
    void bottom_solve()
    {
        const int l = (int)lev.size() - 1;
        MacLev& L = lev[l];
        if (opt.bottom_solver == NP_BOTTOM_SMOOTHER) {
            if (opt.num_final_smooth <= 0) zero(L.cor, L.d.ntot);
            for (int i = 0; i < opt.num_final_smooth; ++i) smooth(l, L.cor, L.res, i == 0, nullptr, nullptr);
            return;
        }
        const bool cg_first = opt.bottom_solver == NP_BOTTOM_CG || (opt.bottom_solver == NP_BOTTOM_BICGCG && bicg_failed_switch);
        int nit = 0;
        int ret = run_bottom(cg_first, &nit);
        bottom_iters += nit;
        if (opt.bottom_verbose) fprintf(stderr, "[macproj] bottom %s: ret %d, %d iterations\n", cg_first ? "cg" : "bicgstab", ret, nit);
        if (ret != 0 && ret != 9) {
            if (opt.bottom_solver == NP_BOTTOM_BICGCG) {
                ret = run_bottom(!cg_first, &nit); bottom_iters += nit;
                if (ret == 0) bicg_failed_switch = cg_first ? false : true;
            }
            if (ret != 0) {
                if (opt.num_final_smooth <= 0) zero(L.cor, L.d.ntot);
                for (int i = 0; i < opt.num_final_smooth; ++i) smooth(l, L.cor, L.res, i == 0, nullptr, nullptr);
            }
        }
    }
    // MLMG::mgVcycle on lev[*].res -> lev[*].cor; the last post-smooth of level 0 also does MLMG's sol += cor
    void vcycle(const double* off0)
    {
        const int nl = (int)lev.size();
        for (int l = 0; l < nl - 1; ++l) {
            MacLev& L = lev[l];
            const double* off = l == 0 ? off0 : nullptr;
            if (opt.num_pre_smooth <= 0) zero(L.cor, L.d.ntot);
            for (int i = 0; i < opt.num_pre_smooth; ++i) smooth_level(l, i == 0, nullptr, nullptr, off);
            resid_restrict(l, off);                                        // computeResOfCorrection + restriction, fused
        }
        bottom_solve();
        for (int l = nl - 2; l >= 0; --l) {
            MacLev& L = lev[l];
            const double* off = l == 0 ? off0 : nullptr;
            if (opt.num_post_smooth <= 0) interp_add(l, L.cor, lev[l + 1].cor);
            for (int i = 0; i < opt.num_post_smooth; ++i)                  // addInterpCorrection rides on the first post-smooth
                smooth_level(l, false, i == 0 ? lev[l + 1].cor : nullptr, (l == 0 && i == opt.num_post_smooth - 1) ? sol : nullptr, off);
            if (l == 0 && opt.num_post_smooth <= 0) add_cor_to_sol();
        }
        if (nl == 1) add_cor_to_sol();
    }
    void add_cor_to_sol()
    {
        // sol += cor as a pass of its own (only when no post-smooth can carry it)
        const MacLevel& d = lev[0].d;
        NP_LAUNCH(false, k_mac_axpy_cell, grid_cells(d), blk(), 0, stream, d, sol, (const double*)lev[0].cor); ++launches;
    }
    // one MLMG iteration (V-cycle with sol += cor, new residual, its norm and mean) replayed from a CUDA graph.
    // 0: done, *rn set; 1: no graph, run the plain path; 2: the bottom solve asked for MLMG's fallback path
    int iterate_graph(const double* off0, long long nb, double* rn)
    {
        if (!opt.use_graph || graph_off || lev.size() < 2) return 1;
        if (comm.nranks > 1 && !graph_warm) { graph_warm = true; return 1; }      // (as NpSolver::iterate_graph: NCCL set up before a capture)
        std::vector<unsigned char> key;
        auto put = [&](const void* p, size_t n) { key.insert(key.end(), (const unsigned char*)p, (const unsigned char*)p + n); };
        for (const MacLev& L : lev) {
            put(&L.d, sizeof(MacLevel)); put(&L.bx, sizeof(double*)); put(&L.by, sizeof(double*)); put(&L.bz, sizeof(double*));
            put(&L.res, sizeof(double*)); put(&L.cor, sizeof(double*)); put(&L.rescor, sizeof(double*)); put(&L.a, sizeof(double*));
        }
        put(&sol, sizeof(double*)); put(&rhs, sizeof(double*)); put(&off0, sizeof(double*)); put(&opt, sizeof(mp_options));
        put(&wrapbuf, sizeof(double*)); put(&wrapbuf_n, sizeof(long long));
        const double sc[2] = {alpha, bscal}; put(sc, sizeof(sc));
        const int sw[4] = {singular ? 1 : 0, bicg_failed_switch ? 1 : 0, have_a ? 1 : 0, peer.ok ? 1 : 0}; put(sw, sizeof(sw));
        if (!itg.exec || key != itg.key) {
            drop_graph();
            if (!have_cap_stream) { if (np_rt::stream_create(&cap_stream)) { graph_off = true; return 1; } have_cap_stream = true; }
            if (opt.profile_kernels && gprof_ev.empty()) { gprof_ev.resize(2 * 16); for (auto& e : gprof_ev) np_rt::event_create(&e); }
            std::vector<double*> save;
            for (MacLev& L : lev) { save.push_back(L.cor); save.push_back(L.rescor); }
            const long long l0 = launches; const np_stream_t user = stream; const int err0 = err; const int bi0 = bottom_iters;
            bool ok = np_rt::capture_begin(cap_stream) == 0;
            if (ok) {
                stream = cap_stream; in_capture = true;
                vcycle(off0);
                halo(0, sol);
                adotx<1, true>(0, lev[0].res, sol, rhs, nullptr);
                halo(0, lev[0].res);
                reduce(nb, true, singular, 0);
                fetch(0);
                in_capture = false; stream = user;
                ok = np_rt::capture_end(cap_stream, &itg.exec) == 0 && err == err0;
            }
            itg.launches = launches - l0; launches = l0; bottom_iters = bi0;
            for (size_t q = 0; q < lev.size(); ++q) { lev[q].cor = save[2 * q]; lev[q].rescor = save[2 * q + 1]; }
            if (!ok) { drop_graph(); np_rt::last_error(); err = err0; graph_off = true; return 1; }     // e.g. a scratch allocation inside the V-cycle
            itg.key = key; ++stats.graph_captures;
        }
        if (np_rt::graph_launch(itg.exec, stream)) { np_rt::last_error(); drop_graph(); graph_off = true; return 1; }
        launches += itg.launches; ++stats.graph_replays;
        np_rt::stream_sync(stream);
        for (int q = 0; q < itg.n_prof; ++q) { g_ms_smooth += np_rt::event_ms(gprof_ev[2 * q], gprof_ev[2 * q + 1]); ++g_n_smooth; }
        if (opt.bottom_solver != NP_BOTTOM_SMOOTHER) {
            if (h_bot_out[0] != 0 && h_bot_out[0] != 9) return 2;
            bottom_iters += h_bot_out[1];
        }
        *rn = h_scal[0];
        return 0;
    }
    // MLMG::solve; `sol` holds the initial guess
    int solve(double rtol, double atol, double rhsnorm0)
    {
        np_rt::Range nvtx_solve("MLMG::solve");
        const MacLevel& d0 = lev[0].d;
        const long long nb = nblk(grid_adotx(d0));
        bottom_iters = 0; bicg_failed_switch = false;
        halo(0, sol);
        adotx<1, true>(0, lev[0].res, sol, rhs, nullptr);
        halo(0, lev[0].res);
        reduce(nb, true, singular, 0);
        const double resnorm0 = fetch(0);
        const double max_norm = rhsnorm0 >= resnorm0 ? rhsnorm0 : resnorm0;
        const double res_target = std::max(atol, std::max(rtol, 1e-16) * max_norm);
        stats.rhs_norm0 = rhsnorm0; stats.res_norm0 = resnorm0; stats.n_vcycles = 0; stats.status = 0; stats.res_norm_final = resnorm0;
        stats.n_levels = (int)lev.size(); stats.bottom_iters_total = 0;
        // NaN / Inf in umac, beta or the initial guess: AMReX's MLMG fails; never report success on it
        if (!(resnorm0 == resnorm0) || !(rhsnorm0 == rhsnorm0) || std::isinf(resnorm0) || std::isinf(rhsnorm0)) { stats.status = NP_ERR_DIVERGED; return NP_ERR_DIVERGED; }
        if (!(resnorm0 > res_target)) return NP_OK;
        bool converged = false;
        int rc = NP_OK;
        for (int iter = 0; iter < opt.maxiter; ++iter) {
            double rn = 0.0;
            const int grc = iterate_graph(singular ? scal + 6 : nullptr, nb, &rn);
            if (grc == 2) { graph_off = true; return MP_RESTART; }      // the caller restores the initial guess and solves again, kernel by kernel
            if (grc != 0) {
            vcycle(singular ? scal + 6 : nullptr);       // makeSolvable(res) is applied by the consumers of res (lazy offset)
            halo(0, sol);
            adotx<1, true>(0, lev[0].res, sol, rhs, nullptr);
            halo(0, lev[0].res);
            reduce(nb, true, singular, 0);
            rn = fetch(0);
            }
            if (iter < 128) stats.res_history[iter] = rn;
            stats.n_vcycles = iter + 1; stats.res_norm_final = rn;
            if (opt.verbose >= 2) fprintf(stderr, "MLMG: Iteration %3d Fine resid/bnorm = %.6e\n", iter + 1, rn / max_norm);
            if (err) { rc = err; break; }                 // a failed exchange / collective: stop issuing them
            if (rn <= res_target) { converged = true; break; }
            if (!(rn <= 1e20 * max_norm)) { rc = NP_ERR_DIVERGED; break; }
        }
        if (!converged && rc == NP_OK) rc = NP_ERR_NOT_CONVERGED;
        stats.bottom_iters_total = bottom_iters;
        return rc;
    }

    int set_bc(const int lo[3], const int hi[3])
    {
        for (int d = 0; d < 3; ++d) {
            if (lo[d] < 0 || lo[d] > 2 || hi[d] < 0 || hi[d] > 2) return NP_ERR_INVALID;
            if ((lo[d] == NP_BC_PERIODIC) != (hi[d] == NP_BC_PERIODIC)) return NP_ERR_INVALID;
            if ((lo[d] == NP_BC_PERIODIC) != (geom.periodic[d] != 0)) return NP_ERR_INVALID;
            bclo[d] = lo[d]; bchi[d] = hi[d];
        }
        setup_bc();
        return NP_OK;
    }
    // MacProjector::updateBeta + MLABecLaplacian::averageDownCoeffs
    int update_beta(const np_fab beta[3])
    {
        MacLev& L0 = lev[0];
        double* dst[3] = {L0.bx, L0.by, L0.bz};
        for (int d = 0; d < 3; ++d) {
            if (!beta[d].p || beta[d].ncomp < 1) return NP_ERR_INVALID;
            FabView v; bool st;
            if (!import(beta[d], 0, d, d, true, &v, &st)) return NP_ERR_CUDA;
            dim3 g((L0.d.n[0] + (d == 0) + 31) / 32, (L0.d.n[1] + (d == 1) + 7) / 8, L0.d.n[2] + (d == 2));
            NP_LAUNCH(true, k_mac_copy_beta, g, blk(), 0, stream, L0.d, dst[d], v, d, part0, part1); ++launches;
            NP_LAUNCH(true, k_mac_reduce_minmax, dim3(1), dim3(1024), 0, stream, (const double*)part0, (const double*)part1, nblk(g), scal + 8, d == 0 ? 1 : 0); ++launches;
        }
        if (comm.nranks > 1) {                    // (-min, max) over the ranks in one max-allreduce
            NP_LAUNCH(false, k_negate_one, dim3(1), dim3(32), 0, stream, scal + 8); ++launches;
            if (comm.allreduce(comm.ctx, scal + 8, 2, 1, (void*)stream)) err = NP_ERR_NCCL;
            NP_LAUNCH(false, k_negate_one, dim3(1), dim3(32), 0, stream, scal + 8); ++launches;
        }
        fetch(0);
        const double bmin = h_scal[8], bmax = h_scal[9];
        const bool is_const = opt.const_beta_fastpath && bmin == bmax;
        for (auto& L : lev) { L.d.const_beta = is_const ? 1 : 0; L.d.beta0 = is_const ? bmin : 1.0; }
        if (!is_const) {
            for (size_t l = 1; l < lev.size(); ++l) {
                const MacLevel &C = lev[l].d, &F = lev[l - 1].d;
                const double* src[3] = {lev[l - 1].bx, lev[l - 1].by, lev[l - 1].bz};
                double* dd[3] = {lev[l].bx, lev[l].by, lev[l].bz};
                const bool gather = (int)l == agg_lev;      // from the last sharded level onto the first replicated one
                MacLevel Cl = C;
                if (gather) Cl.n[2] = F.n[2] / 2;
                for (int d = 0; d < 3; ++d) {
                    dim3 g((Cl.n[0] + (d == 0) + 31) / 32, (Cl.n[1] + (d == 1) + 7) / 8, Cl.n[2] + (d == 2));
                    NP_LAUNCH(false, k_mac_avgdown_face, g, blk(), 0, stream, Cl, F, gather ? agg_send : dd[d], src[d], d); ++launches;
                    if (!gather) continue;
                    const long long plane = d == 0 ? Cl.fxs * Cl.n[1] : d == 1 ? Cl.sx * (Cl.n[1] + 1) : Cl.sxy;
                    if (d < 2) {      // slabs of x / y faces stack up plane by plane: gather straight into the level's array
                        if (comm.allgather(comm.ctx, agg_send, dd[d], plane * Cl.n[2], (void*)stream)) err = NP_ERR_NCCL;
                    } else {          // z faces: every slab brings its top plane too; keep the lower n2 planes of each and the last one's top
                        const long long cnt = plane * (Cl.n[2] + 1);
                        if (comm.allgather(comm.ctx, agg_send, agg_recv, cnt, (void*)stream)) err = NP_ERR_NCCL;
                        for (int r = 0; r < comm.nranks; ++r)
                            np_rt::copy_async(dd[2] + plane * Cl.n[2] * r, agg_recv + cnt * r, (size_t)(plane * Cl.n[2]) * sizeof(double), stream);
                        np_rt::copy_async(dd[2] + plane * Cl.n[2] * comm.nranks, agg_recv + cnt * (comm.nranks - 1) + plane * Cl.n[2],
                                          (size_t)plane * sizeof(double), stream);
                    }
                }
            }
        }
        if (!is_const && comm.nranks > 1) {
            // ghost coefficient planes of the slabs (once per updateBeta): x / y faces of the neighbour's nearest cell plane; z faces:
            // the far face of that plane (the near one is the shared face both ranks hold)
            int lo, hi; z_neighbours(lo, hi);
            for (size_t l = 0; l < lev.size(); ++l) if (sharded((int)l)) {
                MacLev& L = lev[l]; const MacLevel& d = L.d;
                const long long px_ = d.fxs * d.n[1], py_ = d.sx * (d.n[1] + 1), pz_ = d.sxy, n2 = d.n[2];
                int rc = comm.sendrecv(comm.ctx, L.bx, px_, L.bx - px_, px_, L.bx + px_ * (n2 - 1), px_, L.bx + px_ * n2, px_, lo, hi, (void*)stream);
                rc |= comm.sendrecv(comm.ctx, L.by, py_, L.by - py_, py_, L.by + py_ * (n2 - 1), py_, L.by + py_ * n2, py_, lo, hi, (void*)stream);
                rc |= comm.sendrecv(comm.ctx, L.bz + pz_, pz_, L.bz - pz_, pz_, L.bz + pz_ * (n2 - 1), pz_, L.bz + pz_ * (n2 + 1), pz_, lo, hi, (void*)stream);
                if (rc) err = NP_ERR_NCCL;
            }
        }
        stats.sigma_is_const = is_const ? 1 : 0;
        have_beta = true;
        if (err) return err;
        return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
    }
    void finish_stats()
    {
        np_rt::event_record(ev[3], stream);
        if (peer.ok) np_rt::copy_async(h_scal + 42, peer.flags + 5, sizeof(unsigned long long), stream);     // did a halo wait give up? (np_wait_flag)
        np_rt::stream_sync(stream);
        if (peer.ok) { unsigned long long w; memcpy(&w, h_scal + 42, sizeof(w)); if (w != 0ull && !err) err = NP_ERR_NCCL; }
        stats.ms_setup = np_rt::event_ms(ev[0], ev[1]);
        stats.ms_solve = np_rt::event_ms(ev[1], ev[2]);
        stats.ms_finish = np_rt::event_ms(ev[2], ev[3]);
        stats.ms_total = np_rt::event_ms(ev[0], ev[3]);
        stats.kernel_launches = launches;
        stats.n_smooth0 = prof_n / 2 + g_n_smooth; stats.ms_smooth0 = g_ms_smooth;
        for (int e = 0; e + 1 < prof_n; e += 2) stats.ms_smooth0 += np_rt::event_ms(prof_ev[e], prof_ev[e + 1]);
    }
    // MacProjector::setUMAC + project
    int project(np_fab umac[3], np_fab* mac_phi, double rtol, double atol)
    {
        np_rt::Range nvtx_proj("MacProjector::project");
        if (!have_beta) return NP_ERR_STATE;
        for (int d = 0; d < 3; ++d) if (!umac[d].p || umac[d].ncomp < 1) return NP_ERR_INVALID;
        if (mac_phi && (!mac_phi->p || mac_phi->ncomp < 1)) return NP_ERR_INVALID;
        launches = 0; have_result = false; prof_n = 0; err = 0; g_ms_smooth = 0.0; g_n_smooth = 0;
        const int is_const = stats.sigma_is_const;
        stats = np_stats{}; stats.sigma_is_const = is_const;
        const MacLevel& d0 = lev[0].d;
        np_rt::event_record(ev[0], stream);
        FabView uv[3]; bool st[3];
        for (int d = 0; d < 3; ++d) if (!import(umac[d], 0, d, d, true, &uv[d], &st[d])) return NP_ERR_CUDA;
        // rhs = -div(umac), its mean (solvability) and, after the mean is taken out, its inf-norm
        const long long nb = nblk(grid_cells(d0));
        NP_LAUNCH(true, k_mac_divu, grid_cells(d0), blk(), 0, stream, d0, rhs, uv[0], uv[1], uv[2], part0, part1); ++launches;
        reduce(nb, true, singular, 16);
        if (singular) {
            NP_LAUNCH(true, k_mac_sub_norm, grid_cells(d0), blk(), 0, stream, d0, rhs, (const double*)(scal + 16 + 6), part0); ++launches;
            reduce(nb, true, false, 16);
        }
        np_rt::copy_async(h_scal + 16, scal + 16, 8 * sizeof(double), stream);
        // phi: 0, or the caller's mac_phi as the initial guess
        FabView pv; bool pst = false;
        if (mac_phi && !import(*mac_phi, 0, -1, 3, true, &pv, &pst)) return NP_ERR_CUDA;
        auto initial_guess = [&]() {
            if (mac_phi) { NP_LAUNCH(false, (k_mac_copy_cell<false>), grid_cells(d0), blk(), 0, stream, d0, sol, pv); ++launches; }
            else zero(sol, d0.ntot);
        };
        initial_guess();
        np_rt::stream_sync(stream);
        const double rhsnorm0 = h_scal[16];
        np_rt::event_record(ev[1], stream);
        int rc = solve(rtol, atol, rhsnorm0);
        if (rc == MP_RESTART) { initial_guess(); rc = solve(rtol, atol, rhsnorm0); }
        np_rt::event_record(ev[2], stream);
        // umac += flux = -b grad(phi)
        halo(0, sol);
        if (d0.const_beta) NP_LAUNCH(false, (k_mac_fluxes<true, true>), grid_cells(d0), blk(), 0, stream, d0, lev[0].beta(), (const double*)sol, uv[0], uv[1], uv[2]);
        else NP_LAUNCH(false, (k_mac_fluxes<true, false>), grid_cells(d0), blk(), 0, stream, d0, lev[0].beta(), (const double*)sol, uv[0], uv[1], uv[2]);
        ++launches;
        for (int d = 0; d < 3; ++d) export_back(umac[d], uv[d], st[d]);
        if (mac_phi) {
            NP_LAUNCH(false, (k_mac_copy_cell<true>), grid_cells(d0), blk(), 0, stream, d0, sol, pv); ++launches;
            export_back(*mac_phi, pv, pst);
        }
        finish_stats();
        stats.status = rc;
        if (np_rt::last_error()) { stats.status = NP_ERR_CUDA; return NP_ERR_CUDA; }
        if (err) { stats.status = err; return err; }
        have_result = true;
        return rc;
    }
    // ---- MLABecLaplacian beyond the MAC projection: alpha a - beta div(b grad), inhomogeneous Dirichlet data ----
    // (the scalar and velocity-component diffusion solves, /root/reference/src/diffusion/DiffusionScalarOp.cpp:139-260, :262-380)
    int set_scalars(double a_alpha, double a_beta)
    {
        alpha = a_alpha; bscal = a_beta;
        for (auto& L : lev) {
            for (int d = 0; d < 3; ++d) L.d.dh[d] = bscal * L.d.dxinv[d] * L.d.dxinv[d];
            L.d.ascal = alpha; L.d.acoef = (alpha != 0.0 && have_a) ? L.a : nullptr;
        }
        setup_bc();
        return NP_OK;
    }
    // MLABecLaplacian::setACoeffs + averageDownCoeffs (a by average_down)
    int set_acoeffs(const np_fab* a)
    {
        if (!a || !a->p || a->ncomp < 1) return NP_ERR_INVALID;
        if (comm.nranks > 1) return NP_ERR_INVALID;          // one rank this round (the coarse a would need the gather the face coefficients have)
        for (auto& L : lev) if (!L.a && alloc_d(&L.a, L.d.ntot)) return NP_ERR_CUDA;
        FabView v; bool st;
        if (!import(*a, 0, -1, 3, true, &v, &st)) return NP_ERR_CUDA;
        NP_LAUNCH(false, (k_mac_copy_cell<false>), grid_cells(lev[0].d), blk(), 0, stream, lev[0].d, lev[0].a, v); ++launches;
        for (size_t l = 1; l < lev.size(); ++l) restrict_to((int)l, lev[l].a, lev[l - 1].a);
        np_rt::stream_sync(stream);
        have_a = true;
        return set_scalars(alpha, bscal);
    }
    // MLMG::solve(phi, rhs): phi carries the initial guess in its valid cells and, on Dirichlet faces, the boundary values in
    // its first ghost layer (setLevelBC(phi)); on return its valid cells hold the solution.  rho != null: rhs = rho * rhs_or_tracer
    // (diffuse_scalar forms rho * tracer).
    int solve_abec(np_fab* phi, const np_fab* rhs_in, const np_fab* rho, double rtol, double atol)
    {
        np_rt::Range nvtx_solve("DiffusionScalarOp::solve");
        if (!have_beta) return NP_ERR_STATE;
        if (!phi || !phi->p || !rhs_in || !rhs_in->p) return NP_ERR_INVALID;
        launches = 0; have_result = false; prof_n = 0; err = 0; g_ms_smooth = 0.0; g_n_smooth = 0;
        const int is_const = stats.sigma_is_const;
        stats = np_stats{}; stats.sigma_is_const = is_const;
        const MacLevel& d0 = lev[0].d;
        bool any_dir = false;
        for (int d = 0; d < 3; ++d) if (bclo[d] == NP_BC_DIRICHLET || bchi[d] == NP_BC_DIRICHLET) any_dir = true;
        if (any_dir && phi->ngrow < 1) return NP_ERR_INVALID;
        np_rt::event_record(ev[0], stream);
        FabView pv, rv, dv; bool pst, rst, dst;
        if (!import(*phi, 0, -1, 3, true, &pv, &pst) || !import(*rhs_in, 0, -1, 0, true, &rv, &rst)) return NP_ERR_CUDA;
        if (rho) {
            if (!rho->p || !import(*rho, 0, -1, 1, true, &dv, &dst)) return NP_ERR_CUDA;
            NP_LAUNCH(false, k_mac_rho_times, grid_cells(d0), blk(), 0, stream, d0, rhs, dv, rv); ++launches;
        } else { NP_LAUNCH(false, (k_mac_copy_cell<false>), grid_cells(d0), blk(), 0, stream, d0, rhs, rv); ++launches; }
        NP_LAUNCH(false, (k_mac_copy_cell<false>), grid_cells(d0), blk(), 0, stream, d0, sol, pv); ++launches;
        for (int d = 0; d < 3; ++d) for (int side = 0; side < 2; ++side) {
            if ((side ? bchi[d] : bclo[d]) != NP_BC_DIRICHLET) continue;
            const int a = (d + 1) % 3, b = (d + 2) % 3, t0 = std::min(a, b), t1 = std::max(a, b);
            const double cbc = std::min(d0.n[d] + 1, opt.maxorder) >= 3 ? 8.0 / 3.0 : 2.0;
            NP_LAUNCH(false, k_mac_dirichlet_rhs, dim3((d0.n[t0] + 31) / 32, (d0.n[t1] + 7) / 8, 1), blk(), 0, stream, d0, lev[0].beta(), rhs, pv, d, side, cbc); ++launches;
        }
        // inf-norm of the right-hand side (MLMG's bnorm), after the mean is taken out when the operator is singular
        const long long nb = nblk(grid_cells(d0));
        zero(lev[0].rescor, d0.ntot);
        NP_LAUNCH(true, (k_mac_adotx<1, true, true>), grid_adotx(d0), blk(), 0, stream, zero_op(d0), lev[0].beta(), lev[0].rescor, (const double*)lev[0].rescor,
                  (const double*)rhs, (const double*)nullptr, part0, part1); ++launches;       // y = rhs - 0: the reducing pass gives max |rhs| and its sum
        reduce(nblk(grid_adotx(d0)), true, singular, 16);
        if (singular) {
            NP_LAUNCH(true, k_mac_sub_norm, grid_cells(d0), blk(), 0, stream, d0, rhs, (const double*)(scal + 16 + 6), part0); ++launches;
            reduce(nb, true, false, 16);
        }
        np_rt::copy_async(h_scal + 16, scal + 16, 8 * sizeof(double), stream);
        np_rt::stream_sync(stream);
        const double rhsnorm0 = h_scal[16];
        np_rt::event_record(ev[1], stream);
        int rc = solve(rtol, atol, rhsnorm0);
        if (rc == MP_RESTART) {                  // back to the caller's initial guess (pv still holds it), kernel by kernel
            NP_LAUNCH(false, (k_mac_copy_cell<false>), grid_cells(d0), blk(), 0, stream, d0, sol, pv); ++launches;
            rc = solve(rtol, atol, rhsnorm0);
        }
        np_rt::event_record(ev[2], stream);
        NP_LAUNCH(false, (k_mac_copy_cell<true>), grid_cells(d0), blk(), 0, stream, d0, sol, pv); ++launches;
        export_back(*phi, pv, pst);
        finish_stats();
        stats.status = rc;
        if (np_rt::last_error()) { stats.status = NP_ERR_CUDA; return NP_ERR_CUDA; }
        if (err) { stats.status = err; return err; }
        have_result = true;
        return rc;
    }
    // the level-0 descriptor with a zero operator (alpha = 0, b = 0): k_mac_adotx<1, REDUCE> on it is the plain reduction of b
    static MacLevel zero_op(MacLevel d) { d.const_beta = 1; d.beta0 = 0.0; d.acoef = nullptr; d.ascal = 0.0; return d; }

    int get_fluxes(np_fab flux[3], const np_fab* phi)
    {
        if (!have_beta || !phi || !phi->p) return NP_ERR_STATE;
        err = 0;
        const MacLevel& d0 = lev[0].d;
        FabView fv[3], pv; bool st[3], pst;
        for (int d = 0; d < 3; ++d) { if (!flux[d].p) return NP_ERR_INVALID; if (!import(flux[d], 0, d, d, false, &fv[d], &st[d])) return NP_ERR_CUDA; }
        if (!import(*phi, 0, -1, 3, true, &pv, &pst)) return NP_ERR_CUDA;
        double* tmp = lev[0].rescor;                     // free outside a V-cycle
        NP_LAUNCH(false, (k_mac_copy_cell<false>), grid_cells(d0), blk(), 0, stream, d0, tmp, pv);
        halo(0, tmp);
        if (d0.const_beta) NP_LAUNCH(false, (k_mac_fluxes<false, true>), grid_cells(d0), blk(), 0, stream, d0, lev[0].beta(), (const double*)tmp, fv[0], fv[1], fv[2]);
        else NP_LAUNCH(false, (k_mac_fluxes<false, false>), grid_cells(d0), blk(), 0, stream, d0, lev[0].beta(), (const double*)tmp, fv[0], fv[1], fv[2]);
        for (int d = 0; d < 3; ++d) export_back(flux[d], fv[d], st[d]);
        np_rt::stream_sync(stream);
        if (err) return err;
        return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
    }
    // ---- several boxes (amr.max_grid_size chops the level): the boxes' fabs are gathered into domain-shaped device fabs,
    //      the projection runs on those as for one box, and the results are scattered back (faces shared by two boxes are
    //      written to both) ----
    int boxes_copy(bool to_domain, int nodal_dir, int nlocal, const np_fab* fabs, double* dom, int slot)
    {
        if (nlocal != (int)lbox.size()) return NP_ERR_INVALID;
        const MacLevel& d0 = lev[0].d;
        FabView dv; dv.p = dom; dv.ng = 0;
        dv.sy = d0.n[0] + (nodal_dir == 0); dv.sz = dv.sy * (d0.n[1] + (nodal_dir == 1)); dv.sn = dv.sz * (d0.n[2] + (nodal_dir == 2));
        const int zero3[3] = {0, 0, 0};
        for (int b = 0; b < nlocal; ++b) {
            const np_box& bx = lbox[b];
            const np_fab& f = fabs[b];
            if (!f.p || f.ncomp < 1) return NP_ERR_INVALID;
            FabView bv; bv.p = f.p; bv.ng = f.ngrow;
            bv.sy = bx.hi[0] - bx.lo[0] + 1 + (nodal_dir == 0) + 2 * f.ngrow;
            bv.sz = bv.sy * (bx.hi[1] - bx.lo[1] + 1 + (nodal_dir == 1) + 2 * f.ngrow);
            bv.sn = bv.sz * (bx.hi[2] - bx.lo[2] + 1 + (nodal_dir == 2) + 2 * f.ngrow);
            bool staged = false;
#ifndef NP_EMU
            if (!np_rt::is_device_ptr(f.p)) {       // host fab: through the staging buffer (ghost cells survive the round trip)
                double* dev = stage_get(slot, bv.sn);
                if (!dev) return NP_ERR_CUDA;
                np_rt::copy_async(dev, f.p, (size_t)bv.sn * sizeof(double), stream);
                bv.p = dev; staged = true;
            }
#endif
            BoxCopy a;
            for (int d = 0; d < 3; ++d) {
                a.rlo[d] = bx.lo[d]; a.rhi[d] = bx.hi[d] + (nodal_dir == d ? 1 : 0);
                a.dorg[d] = to_domain ? zero3[d] : bx.lo[d]; a.sorg[d] = to_domain ? bx.lo[d] : zero3[d];
                a.vlo[d] = a.vhi[d] = a.dlo[d] = a.dhi[d] = 0;
            }
            a.guard = 0;
            const dim3 g((a.rhi[0] - a.rlo[0] + 32) / 32, (a.rhi[1] - a.rlo[1] + 8) / 8, a.rhi[2] - a.rlo[2] + 1);
            if (to_domain) NP_LAUNCH(false, k_box_copy, g, blk(), 0, stream, dv, bv, a, 1);
            else NP_LAUNCH(false, k_box_copy, g, blk(), 0, stream, bv, dv, a, 1);
            if (staged && !to_domain) np_rt::copy_async(f.p, bv.p, (size_t)bv.sn * sizeof(double), stream);
        }
        return NP_OK;
    }
    int ensure_box_buffers()
    {
        const MacLevel& d0 = lev[0].d;
        for (int d = 0; d < 3; ++d)
            if (!mb_f[d] && alloc_d(&mb_f[d], (long long)(d0.n[0] + (d == 0)) * (d0.n[1] + (d == 1)) * (d0.n[2] + (d == 2)))) return NP_ERR_CUDA;
        if (!mb_phi && alloc_d(&mb_phi, (long long)d0.n[0] * d0.n[1] * d0.n[2])) return NP_ERR_CUDA;
        return NP_OK;
    }
    int update_beta_boxes(int nlocal, const np_fab* const b[3])
    {
        if (ensure_box_buffers()) return NP_ERR_CUDA;
        np_fab dom[3];
        for (int d = 0; d < 3; ++d) { const int rc = boxes_copy(true, d, nlocal, b[d], mb_f[d], d); if (rc) return rc; dom[d] = np_fab{mb_f[d], 0, 1}; }
        return update_beta(dom);
    }
    int project_boxes(int nlocal, np_fab* const u[3], np_fab* phi, double rtol, double atol)
    {
        if (ensure_box_buffers()) return NP_ERR_CUDA;
        np_fab dom[3], pd{mb_phi, 0, 1};
        for (int d = 0; d < 3; ++d) { const int rc = boxes_copy(true, d, nlocal, u[d], mb_f[d], d); if (rc) return rc; dom[d] = np_fab{mb_f[d], 0, 1}; }
        if (phi) { const int rc = boxes_copy(true, -1, nlocal, phi, mb_phi, 3); if (rc) return rc; }
        const int rc = project(dom, phi ? &pd : nullptr, rtol, atol);
        if (rc && !have_result) return rc;
        for (int d = 0; d < 3; ++d) { const int r2 = boxes_copy(false, d, nlocal, u[d], mb_f[d], d); if (r2) return r2; }
        if (phi) { const int r2 = boxes_copy(false, -1, nlocal, phi, mb_phi, 3); if (r2) return r2; }
        np_rt::stream_sync(stream);
        return np_rt::last_error() ? NP_ERR_CUDA : rc;
    }
    int get_fluxes_boxes(int nlocal, np_fab* const fl[3], const np_fab* phi)
    {
        if (ensure_box_buffers() || !phi) return NP_ERR_INVALID;
        np_fab dom[3], pd{mb_phi, 0, 1};
        for (int d = 0; d < 3; ++d) dom[d] = np_fab{mb_f[d], 0, 1};
        int rc = boxes_copy(true, -1, nlocal, phi, mb_phi, 3);
        if (!rc) rc = get_fluxes(dom, &pd);
        for (int d = 0; d < 3 && !rc; ++d) rc = boxes_copy(false, d, nlocal, fl[d], mb_f[d], d);
        np_rt::stream_sync(stream);
        return rc;
    }
    // cell fab of level l (host or device, any ghosts) <-> internal array
    int cell_in(int l, const np_fab& f, double* dst, int slot)
    {
        FabView v; bool st;
        if (!f.p || !import(f, l, -1, slot, true, &v, &st)) return NP_ERR_INVALID;
        NP_LAUNCH(false, (k_mac_copy_cell<false>), grid_cells(lev[l].d), blk(), 0, stream, lev[l].d, dst, v);
        return NP_OK;
    }
    int cell_out(int l, np_fab& f, double* src, int slot)
    {
        FabView v; bool st;
        if (!f.p || !import(f, l, -1, slot, false, &v, &st)) return NP_ERR_INVALID;
        if (st) np_rt::memset_async(v.p, 0, (size_t)v.sn * sizeof(double), stream);
        NP_LAUNCH(false, (k_mac_copy_cell<true>), grid_cells(lev[l].d), blk(), 0, stream, lev[l].d, src, v);
        export_back(f, v, st);
        np_rt::stream_sync(stream);
        return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
    }
};

struct mp_handle { MacSolver S; };

extern "C" {

void mp_default_options(mp_options* o)
{
    o->max_coarsening_level = 100; o->maxiter = 200; o->bottom_maxiter = 200; o->num_pre_smooth = 2; o->num_post_smooth = 2;
    o->num_final_smooth = 8; o->bottom_solver = NP_BOTTOM_BICGCG; o->verbose = 0; o->bottom_verbose = 0; o->maxorder = 3;
    o->bottom_rtol = 1e-4; o->bottom_atol = -1.0; o->const_beta_fastpath = 1; o->fused_rb_min_n = 256; o->fused_rb_const = 1; o->profile_kernels = 0; o->agg_min_local_nz = 32; o->use_graph = 1; o->p2p_halo = 1;
}
int mp_create(mp_handle** h, const np_geom* geom, int nboxes_global, const np_box* all_boxes, const int* owner_rank,
              int my_rank, int nranks, void* nccl_comm_or_null, const mp_options* opt)
{
    if (!h || !geom || !all_boxes || !owner_rank) return NP_ERR_INVALID;
    mp_options o; if (opt) o = *opt; else mp_default_options(&o);
    mp_handle* p = new (std::nothrow) mp_handle();
    if (!p) return NP_ERR_INVALID;
    NpComm c;
    if (nranks > 1) {
        np_comm_fill_nccl(c, nccl_comm_or_null);
        if (!c.sendrecv) { delete p; return NP_ERR_NCCL; }
    }
    const int rc = p->S.create(geom, nboxes_global, all_boxes, owner_rank, my_rank, nranks, c, &o);
    if (rc) { p->S.destroy(); delete p; return rc; }
    *h = p;
    return NP_OK;
}
int mp_set_bc(mp_handle* h, const int lo[3], const int hi[3]) { return h ? h->S.set_bc(lo, hi) : NP_ERR_INVALID; }
int mp_update_beta(mp_handle* h, const np_fab beta[3]) { return h && beta && h->S.lbox.size() == 1 ? h->S.update_beta(beta) : NP_ERR_INVALID; }
int mp_project(mp_handle* h, np_fab umac[3], np_fab* mac_phi, double rtol, double atol, void* cuda_stream, np_stats* out)
{
    if (!h || !umac || h->S.lbox.size() != 1) return NP_ERR_INVALID;      // several boxes: mp_project_boxes
    h->S.stream = (np_stream_t)cuda_stream;
    const int rc = h->S.project(umac, mac_phi, rtol, atol);
    if (out) *out = h->S.stats;
    return rc;
}
int mp_set_scalars(mp_handle* h, double alpha, double beta) { return h ? h->S.set_scalars(alpha, beta) : NP_ERR_INVALID; }
int mp_set_acoeffs(mp_handle* h, const np_fab* a) { return h && h->S.lbox.size() == 1 ? h->S.set_acoeffs(a) : NP_ERR_INVALID; }
int mp_solve(mp_handle* h, np_fab* phi, const np_fab* rhs, const np_fab* rho_or_null, double rtol, double atol, void* cuda_stream, np_stats* out)
{
    if (!h || h->S.lbox.size() != 1) return NP_ERR_INVALID;
    h->S.stream = (np_stream_t)cuda_stream;
    const int rc = h->S.solve_abec(phi, rhs, rho_or_null, rtol, atol);
    if (out) *out = h->S.stats;
    return rc;
}
int mp_get_fluxes(mp_handle* h, np_fab flux[3], const np_fab* phi) { return h && flux && h->S.lbox.size() == 1 ? h->S.get_fluxes(flux, phi) : NP_ERR_INVALID; }
int mp_update_beta_boxes(mp_handle* h, int nlocal, const np_fab* bx, const np_fab* by, const np_fab* bz)
{
    if (!h || !bx || !by || !bz) return NP_ERR_INVALID;
    const np_fab* const b[3] = {bx, by, bz};
    return h->S.update_beta_boxes(nlocal, b);
}
int mp_project_boxes(mp_handle* h, int nlocal, np_fab* ux, np_fab* uy, np_fab* uz, np_fab* mac_phi, double rtol, double atol,
                     void* cuda_stream, np_stats* out)
{
    if (!h || !ux || !uy || !uz) return NP_ERR_INVALID;
    h->S.stream = (np_stream_t)cuda_stream;
    np_fab* const u[3] = {ux, uy, uz};
    const int rc = h->S.project_boxes(nlocal, u, mac_phi, rtol, atol);
    if (out) *out = h->S.stats;
    return rc;
}
int mp_get_fluxes_boxes(mp_handle* h, int nlocal, np_fab* fx, np_fab* fy, np_fab* fz, const np_fab* phi)
{
    if (!h || !fx || !fy || !fz) return NP_ERR_INVALID;
    np_fab* const f[3] = {fx, fy, fz};
    return h->S.get_fluxes_boxes(nlocal, f, phi);
}
int mp_copy_phi(mp_handle* h, np_fab* dst)
{
    if (!h || !dst) return NP_ERR_INVALID;
    if (!h->S.have_result) return NP_ERR_STATE;
    return h->S.cell_out(0, *dst, h->S.sol, 3);
}
void mp_destroy(mp_handle* h) { if (h) { h->S.destroy(); delete h; } }
int mp_num_levels(mp_handle* h) { return h ? (int)h->S.lev.size() : 0; }
int mp_level_ncell(mp_handle* h, int l, int n[3])
{
    if (!h || l < 0 || l >= (int)h->S.lev.size()) return NP_ERR_INVALID;
    for (int d = 0; d < 3; ++d) n[d] = h->S.lev[l].d.n[d];
    return NP_OK;
}

// ---- kernel-level hooks ----
#define MP_LEVEL_OK(h, l) ((h) && (l) >= 0 && (l) < (int)(h)->S.lev.size() && (h)->S.have_beta)
int mp_get_beta(mp_handle* h, int l, np_fab beta[3])
{
    if (!MP_LEVEL_OK(h, l)) return NP_ERR_INVALID;
    MacSolver& S = h->S; MacLev& L = S.lev[l];
    const double* src[3] = {L.bx, L.by, L.bz};
    for (int d = 0; d < 3; ++d) {
        FabView v; bool st;
        if (!S.import(beta[d], l, d, d, false, &v, &st)) return NP_ERR_CUDA;
        dim3 g((L.d.n[0] + (d == 0) + 31) / 32, (L.d.n[1] + (d == 1) + 7) / 8, L.d.n[2] + (d == 2));
        NP_LAUNCH(false, k_mac_beta_out, g, MacSolver::blk(), 0, S.stream, L.d, src[d], v, d);
        S.export_back(beta[d], v, st);
    }
    np_rt::stream_sync(S.stream);
    return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
}
int mp_apply(mp_handle* h, int l, const np_fab* x, np_fab* y)
{
    if (!MP_LEVEL_OK(h, l) || !x || !y) return NP_ERR_INVALID;
    MacSolver& S = h->S; MacLev& L = S.lev[l];
    if (S.cell_in(l, *x, L.cor, 0)) return NP_ERR_INVALID;
    S.halo(l, L.cor);
    S.adotx<0, false>(l, L.rescor, L.cor, nullptr, nullptr);
    return S.cell_out(l, *y, L.rescor, 1);
}
int mp_smooth(mp_handle* h, int l, np_fab* x, const np_fab* rhs, int ncalls)
{
    if (!MP_LEVEL_OK(h, l) || !x || !rhs) return NP_ERR_INVALID;
    MacSolver& S = h->S; MacLev& L = S.lev[l];
    if (S.cell_in(l, *x, L.cor, 0) || S.cell_in(l, *rhs, L.res, 1)) return NP_ERR_INVALID;
    for (int i = 0; i < ncalls; ++i) S.smooth_level(l, false, nullptr, nullptr, nullptr);
    return S.cell_out(l, *x, L.cor, 0);
}
int mp_restrict(mp_handle* h, int lc, const np_fab* fine, np_fab* crse)
{
    if (!MP_LEVEL_OK(h, lc) || lc < 1 || !fine || !crse) return NP_ERR_INVALID;
    MacSolver& S = h->S;
    if (S.cell_in(lc - 1, *fine, S.lev[lc - 1].rescor, 0)) return NP_ERR_INVALID;
    S.restrict_to(lc, S.lev[lc].res, S.lev[lc - 1].rescor);
    return S.cell_out(lc, *crse, S.lev[lc].res, 1);
}
int mp_interp_add(mp_handle* h, int lf, np_fab* fine, const np_fab* crse)
{
    if (!MP_LEVEL_OK(h, lf) || lf + 1 >= (int)h->S.lev.size() || !fine || !crse) return NP_ERR_INVALID;
    MacSolver& S = h->S;
    if (S.cell_in(lf, *fine, S.lev[lf].cor, 0) || S.cell_in(lf + 1, *crse, S.lev[lf + 1].cor, 1)) return NP_ERR_INVALID;
    S.interp_add(lf, S.lev[lf].cor, S.lev[lf + 1].cor);
    return S.cell_out(lf, *fine, S.lev[lf].cor, 0);
}
int mp_bottom_solve(mp_handle* h, const np_fab* b, np_fab* x, int* iters)
{
    if (!h || !h->S.have_beta || !b || !x) return NP_ERR_INVALID;
    MacSolver& S = h->S; const int l = (int)S.lev.size() - 1;
    if (S.cell_in(l, *b, S.lev[l].res, 0)) return NP_ERR_INVALID;
    S.bottom_iters = 0; S.bicg_failed_switch = false;
    S.bottom_solve();
    if (iters) *iters = S.bottom_iters;
    return S.cell_out(l, *x, S.lev[l].cor, 1);
}

} // extern "C"
