// np_solver.h — host side of the nodal projection: MG hierarchy, V-cycle control (amrex::MLMG),
// projection driver (amrex::NodalProjector::project).  All arithmetic is in np_kernels.cuh.
#pragma once
#include "np_types.h"
#include "../../include/nodalproj.h"
#include <vector>

// Halo exchange / reductions between ranks.  Product build: NCCL (np_comm.cu).  The emulation build used
// by the CPU tests plugs in callbacks (tests drive them over torch.distributed/gloo).
struct NpComm {
    int rank = 0, nranks = 1;
    void* nccl = nullptr;
    // ring neighbours exchange contiguous runs of doubles; rank -1 = no neighbour on that side
    int (*sendrecv)(void* ctx, const double* send_lo, long long nsend_lo, double* recv_lo, long long nrecv_lo,
                    const double* send_hi, long long nsend_hi, double* recv_hi, long long nrecv_hi,
                    int lo_rank, int hi_rank, void* stream) = nullptr;
    int (*allreduce)(void* ctx, double* buf, int n, int op /*0 sum, 1 max*/, void* stream) = nullptr;
    int (*allgather)(void* ctx, const double* send, double* recv, long long cnt, void* stream) = nullptr;
    void* ctx = nullptr;
};

struct NpLevel {
    LevelDev d{};
    double* sigma = nullptr;
    double *res = nullptr, *cor = nullptr, *cor_alt = nullptr, *rescor = nullptr;
    int zc = 65;                    // smoother: written planes per z-chunk
    bool cor_ghost_clean = false;
    bool replicated = false;        // multi-GPU: this level holds the whole domain on every rank
    int koff = 0;                   // global z index of local cell 0
    int nglob2 = 0;                 // global cells in z
};

struct NpSolver {
    np_options opt{};
    np_geom geom{};
    int bclo[3] = {0,0,0}, bchi[3] = {0,0,0};
    NpComm comm;
    int nloc[3] = {0,0,0};
    int koff = 0;
    std::vector<NpLevel> lev;
    int agg_lev = -1;               // first replicated level (multi-GPU), -1: none
    // level-0 extras
    double *sol = nullptr, *rhs = nullptr;
    double *gphi = nullptr;         // cell, 3 comp, 1 ghost, FArrayBox layout
    double *phi_out = nullptr;      // nodal, 1 ghost, FArrayBox layout (unpadded)
    // scratch
    double *part0 = nullptr, *part1 = nullptr; long long npart = 0;
    double *scal = nullptr;         // device scalars
    double *h_scal = nullptr;       // pinned mirror
    int *bot_out = nullptr, *h_bot_out = nullptr;
    double *bw[7] = {nullptr,nullptr,nullptr,nullptr,nullptr,nullptr,nullptr};   // bottom work arrays
    double *xbuf_lo = nullptr, *xbuf_hi = nullptr;   // exchange staging
    double *agg_send = nullptr, *agg_recv = nullptr;
    // host staging (caller passes host pointers)
    double* stage[3] = {nullptr, nullptr, nullptr}; long long stage_n[3] = {0, 0, 0};   // persistent device staging
    np_stream_t stream = 0;
    np_event_t ev[4] = {0,0,0,0};
    np_stream_t comm_stream = 0;       // multi-GPU: halo exchange of the end chunks runs here, beside the interior chunks
    np_event_t ev_halo[2] = {0, 0};    // [0] end chunks done (solver stream), [1] exchange done (comm stream)
    bool have_comm_stream = false;
    // peer-memory halo exchange (p2p_halo): arrays that take part, the neighbours' views of them, and the pacing flags
    struct PeerHalo {
        bool ok = false;
        std::vector<double*> mine, lo, hi;          // registered local arrays and the same arrays of the lower / upper rank
        std::vector<void*> opened;                  // IPC mappings to close
        unsigned long long *flags = nullptr, *lo_flags = nullptr, *hi_flags = nullptr;
        unsigned int* counter = nullptr;
    } peer;
    int setup_peer_halo();
    bool halo_p2p(int l, double* x, np_stream_t on);
    // several boxes per rank: the caller's boxes tile this rank's z-slab; they are gathered into slab-shaped fabs
    // (mb_vel: 3 comp, 1 ghost; mb_sig: 1 comp, 0 ghost), the projection runs on the slab, and u / phi / grad(phi)
    // are scattered back into per-box fabs (mb_phi, mb_gphi: the library-owned results np_get_* hands out)
    std::vector<np_box> lbox;       // this rank's boxes in BoxArray order, slab-local cell indices
    double *mb_vel = nullptr, *mb_sig = nullptr;
    std::vector<double*> mb_phi, mb_gphi;
    std::vector<long long> mb_stage_off;   // host callers: offset of every box inside stage[0] (vel)
    int project_boxes(int nlocal, np_fab* vel, const np_fab* sigma, double rtol, double atol);
    void box_copy(const FabView& dst, const int dorg[3], const FabView& src, const int sorg[3], const int rlo[3], const int rhi[3],
                  int ncomp, const np_box* guard_valid);
    // np_options.use_graph: one MLMG iteration (V-cycle, residual, norm reduction, the two status copies) as a CUDA graph.
    // The graph bakes in every kernel argument, so it carries a key (level descriptors with BCs and coefficients, array
    // pointers, options, solver switches) and is captured again when the key changes.  One rank only (the peer-memory halo
    // kernel takes the exchange epoch as an argument).
    struct IterGraph {
        np_rt::np_graph_t exec = nullptr;
        std::vector<unsigned char> key;
        long long launches = 0;        // kernel launches inside the graph
        int n_prof = 0;                // profile_kernels: event pairs recorded by the graph
    } itg;
    np_stream_t cap_stream = 0; bool have_cap_stream = false;
    bool in_capture = false;
    bool graph_off = false;            // capture failed, or a bottom solve inside the graph needed the fallback path
    bool graph_warm = false;           // several ranks: the first MLMG iteration of the handle has run kernel by kernel
    std::vector<np_event_t> gprof_ev;  // profile_kernels inside the graph
    double g_ms_smooth = 0.0; int g_n_smooth = 0;
    int iterate_graph(const double* rhs_off, double* resnorm);   // 0: done; 1: no graph (run the plain path); 2: bottom solve failed
    void drop_graph();
    std::vector<np_event_t> prof_ev;   // profile_kernels: start/stop pairs around level-0 smoother launches
    int prof_n = 0;
    np_stats stats{};
    bool have_sigma = false, have_result = false, singular = true;
    bool bicg_failed_switch = false;
    long long launches = 0;
    long long divu_npart = 0;       // partial sums of sum(w*rhs) left in part1 by the divergence kernel (0: none)
    long long red_n = 0;            // number of partials the last reducing kernel left in part0 / part1
    double wtot = 1.0;              // sum of the dot-mask weights over the domain (singular problems)
    const double* lvl0_off = nullptr;   // device scalar subtracted from lev[0].res by its consumers (lazy makeSolvable)
    int err = 0;
    int gs_variant_c = 0, gs_variant_v = 0;   // which smoother kernel runs for constant / variable sigma (0: k_gs2)
    int gs_small_variant = 2;                 // smoother launches of at most one CTA per SM run this k_gs3 variant (2: 2 x 4 blocks, 512 threads:
                                              // half the instruction stream per warp and plane step; env NP_GS_SMALL_VARIANT; 0: as the large launches)
    bool interp_march = true;                 // sigma-weighted prolongation: k_interp_add_w2 (z-marching); env NP_INTERP_MARCH=0: k_interp_add_w
    int gs_small = 0;                         // smoother launches of <= gs_small * (SM count) CTAs run k_gs_small (env NP_GS_SMALL; 0: never = default: measured slower than k_gs2)

    int create(const np_geom* g, int nboxes, const np_box* boxes, const int* owner, int my_rank, int nranks,
               const NpComm& c, const np_options* o);
    void destroy();
    void ensure_clean();             // handle cache: clears the arrays on the first use after reuse iff the BCs changed
    bool reuse_pending = false; int last_bclo[3] = {0,0,0}, last_bchi[3] = {0,0,0};
    void reset_for_reuse();          // handle cache: back to the state create() leaves (same arrays, all cleared)
    int set_bc(const int lo[3], const int hi[3]);
    void setup_level_bc(NpLevel& L);

    // building blocks (device pointers, internal layout)
    void fill(int l, double* x);
    void apply(int l, double* y, double* x);                       // fills x's ghosts first
    void residual(int l, double* y, double* x, const double* b, bool with_red, const double* boff);
    void smooth(int l, bool zero_in, double* add_to);              // on lev[l].cor / res
    void restrict_to(int lc, double* crse, double* fine);          // fine gets its ghosts filled
    void restrict_to_replica(int lc, double* crse, const double* fine);
    void resid_restrict(int l, const double* boff);                // lev[l+1].res = R(lev[l].res - *boff - A lev[l].cor), one pass
    bool fused_rr = true;
    void avgdown_sigma_to_replica(int lc);
    void interp_add(int lf, double* fine, const double* crse);
    void make_solvable(int l, double* x);
    double norminf(int l, const double* x);
    double finish_norm(int l, bool want_off);                      // reduce the partials of the last reducing kernel -> host
    int bottom_solve();
    void vcycle(bool add_to_sol);
    int solve(double rtol, double atol);
    int set_sigma(const FabView& s);
    int finish_sigma(long long nbm, const FabView* src);
    void finish_stats();
    int project(np_fab* vel, const np_fab* sigma, double rtol, double atol);
    int apply_projection(np_fab* vel, const np_fab* vel_old, const np_fab* rho, np_fab* gp, np_fab* p, const np_apply_args* a,
                         double rtol, double atol);
    void zero(double* p, long long n);
    FabView view_nodal(const np_fab& f, int l) const;
    FabView view_cell(const np_fab& f, int l, int ncomp) const;
    // host-pointer staging helpers for hooks
    double* stage_get(int slot, long long n);
    int import_fab(const np_fab& f, long long nelem, double** dev, bool* staged);
    int export_fab(const np_fab& f, long long nelem, double* dev, bool staged);
};

struct np_handle { NpSolver S; void* cache_key = nullptr; };     // cache_key: np_capi.cu, handle cache

#define NP_BLK dim3(32, 8, 1)
static inline dim3 grid_nodes(const LevelDev& L) { return dim3((L.n[0] + 1 + 31) / 32, (L.n[1] + 1 + 7) / 8, L.n[2] + 1); }
static inline dim3 grid_cells(const LevelDev& L) { return dim3((L.n[0] + 31) / 32, (L.n[1] + 7) / 8, L.n[2]); }
static inline long long nblocks(dim3 g) { return (long long)g.x * g.y * g.z; }

void np_fill_cell3(NpSolver& S, FabView u);                 // np_solver.cu
void np_comm_fill_nccl(NpComm& c, void* nccl_comm);          // np_comm.cu (product) / np_comm_emu.cpp (tests)
