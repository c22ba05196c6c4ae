/* nodalproj.h — C ABI of the B200-native nodal projection (libnodalproj.so).
 *
 * This is the drop-in boundary for the one hot path of asalmgren/incflo that this repository
 * rebuilds: the MLMG nodal projection behind incflo::ApplyProjection.  Every entry point names the
 * reference interface it replaces (paths are relative to the reference tree; the arithmetic of
 * amrex::NodalProjector / MLMG / MLNodeLaplacian lives in AMReX >= 20.11, which the reference does
 * not vendor, so those citations are to the reference's call sites):
 *
 *   np_create       <- NodalProjector ctor + LPInfo::setMaxCoarseningLevel
 *                      src/projection/incflo_apply_nodal_projection.cpp:148-151, "nodal_proj.*"
 *                      ParmParse options src/incflo.H:350-359, src/setup/init.cpp:143-148
 *   np_set_bc       <- NodalProjector::setDomainBC            ...nodal_projection.cpp:152
 *                      (BC mapping: incflo::get_projection_bc  ...nodal_projection.cpp:6-36)
 *   np_project      <- NodalProjector::project(rtol, atol)    ...nodal_projection.cpp:153
 *   np_get_phi      <- NodalProjector::getPhi                 ...nodal_projection.cpp:165
 *   np_get_gradphi  <- NodalProjector::getGradPhi             ...nodal_projection.cpp:166
 *   np_destroy      <- ~NodalProjector (unique_ptr reset at   ...nodal_projection.cpp:134)
 *
 * Data layout is AMReX's FArrayBox layout (what the reference's Array4 accessors index, e.g.
 * u(i,j,k,n) at ...nodal_projection.cpp:96-98): for a box with valid index range [lo,hi] and g
 * ghost layers, element (i,j,k,n) lives at
 *     p[(i-lo.x+g) + NX*((j-lo.y+g) + NY*((k-lo.z+g) + NZ*n))],   NX = hi.x-lo.x+1+2g, ...
 * i.e. x fastest, component slowest, doubles.  A nodal fab has hi+1 in every direction.
 *
 * Pointers may be device pointers (AMReX built with CUDA keeps fabs in device memory) or host
 * pointers (CPU build of the caller); host buffers are staged through pinned memory inside
 * np_project.  All functions return 0 on success or an NP_ERR_* code; nothing throws across the
 * boundary.  The library has no CPU compute path: without a CUDA device np_create fails with
 * NP_ERR_NO_DEVICE.
 */
#ifndef NODALPROJ_H
#define NODALPROJ_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct np_handle np_handle;

typedef struct { int lo[3], hi[3]; } np_box;                 /* valid CELL index range of one box */
typedef struct { double* p; int ngrow; int ncomp; } np_fab;  /* one box of a MultiFab, layout above */
typedef struct { int ncell[3]; double prob_lo[3], prob_hi[3]; int periodic[3]; } np_geom;

typedef enum { NP_BC_PERIODIC = 0, NP_BC_DIRICHLET = 1, NP_BC_NEUMANN = 2 } np_bc;   /* amrex::LinOpBCType subset */
typedef enum { NP_BOTTOM_BICGCG = 0, NP_BOTTOM_BICGSTAB = 1, NP_BOTTOM_CG = 2, NP_BOTTOM_SMOOTHER = 3 } np_bottom;

enum {
    NP_OK = 0,
    NP_ERR_INVALID = 1,        /* bad argument / unsupported box layout */
    NP_ERR_NO_DEVICE = 2,      /* no CUDA device: there is no CPU fallback */
    NP_ERR_CUDA = 3,
    NP_ERR_NCCL = 4,
    NP_ERR_NOT_CONVERGED = 5,  /* MLMG: "failed to converge" after maxiter V-cycles */
    NP_ERR_DIVERGED = 6,       /* residual > 1e20 * max_norm, or NaN */
    NP_ERR_STATE = 7,          /* call order (e.g. np_get_phi before np_project) */
    NP_ERR_TOO_LARGE = 8       /* np_apply_noalloc: the level does not fit the fixed scratch (NP_NOALLOC_MAX_DOUBLES) */
};

/* The solver's whole option surface = the reference's "nodal_proj.*" keys (+ LPInfo) and the knobs
 * that define this implementation's smoother ordering (DESIGN.md "smoother semantics"). */
typedef struct {
    int max_coarsening_level;   /* LPInfo::setMaxCoarseningLevel           default 30 */
    int maxiter;                /* nodal_proj.maxiter                      default 100 */
    int bottom_maxiter;         /* nodal_proj.bottom_maxiter               default 100 */
    int num_pre_smooth;         /* nu1                                     default 2 */
    int num_post_smooth;        /* nu2                                     default 2 */
    int num_final_smooth;       /* nuf (bottom fallback)                   default 8 */
    int bottom_solver;          /* np_bottom, nodal_proj.bottom_solver     default bicgcg */
    int verbose;                /* nodal_proj.verbose                      default 0 */
    int bottom_verbose;         /* nodal_proj.bottom_verbose               default 0 */
    double bottom_rtol;         /* nodal_proj.bottom_rtol                  default 1e-4 */
    double bottom_atol;         /* nodal_proj.bottom_atol                  default -1 */
    /* --- implementation knobs --- */
    int smooth_zchunk;          /* written planes per smoother tile in z; <=0: chosen per level */
    int interp_sigma_weighted;  /* 1: mlndlap_interpadd_aa weights (default); 0: plain trilinear */
    int bottom_normalize;       /* 1: MLCGSolver's Lp.normalize (default) */
    int agg_min_local_nz;       /* multi-GPU: replicate MG levels whose slab is thinner than this; default 64 */
    int use_graph;              /* 1 (default): one MLMG iteration (V-cycle + residual + norm) is captured as a CUDA graph on first
                                   use and replayed; captured again when BCs, coefficients or options change; with several ranks
                                   the halo kernels and NCCL collectives are part of the graph.  A bottom solve that needs MLMG's fallback path restarts the
                                   solve without the graph.  0: launch kernel by kernel */
    int const_sigma_fastpath;   /* use the constant-coefficient kernels when sigma is uniform; default 1 */
    int profile_kernels;        /* bracket every level-0 smoother launch with CUDA events (np_stats.ms_smooth0); default 0 */
    int overlap_halo;           /* multi-GPU: smooth the end z-chunks first and exchange their planes on a second stream
                                   while the interior chunks are relaxed (levels of >= 4 waves of CTAs only); default 0:
                                   measured 1.4 % slower than the plain order on 2 B200s; < 0: on every
                                   sharded level with three or more chunks (tests) */
    int p2p_halo;               /* multi-GPU: halo planes are stored straight into the neighbours' arrays over NVLink (CUDA IPC
                                   peer memory + flags, one kernel per exchange) instead of grouped ncclSend/ncclRecv;
                                   default 1, falls back to NCCL when peer memory cannot be opened on every rank */
} np_options;

typedef struct {
    int n_vcycles, bottom_iters_total, status, n_levels;
    double rhs_norm0, res_norm0, res_norm_final;
    double res_history[128];
    /* device-timed phases of the last np_project, milliseconds */
    double ms_setup, ms_solve, ms_finish, ms_total;
    long long kernel_launches;   /* kernels of this library launched inside the last np_project */
    int sigma_is_const;
    /* profile_kernels: summed device time and count of the level-0 smoother launches of the last np_project */
    int n_smooth0;
    double ms_smooth0;
    int graph_replays;           /* use_graph: MLMG iterations of the last np_project that ran as one CUDA-graph launch each
                                    (their kernels are still counted in kernel_launches) */
    int graph_captures;          /* ... and how many times the graph had to be captured for it */
} np_stats;

void np_default_options(np_options* o);

/* One handle per (geometry, box layout, rank).  all_boxes lists every box of the level-0 BoxArray with
 * its owner rank (the DistributionMapping).  Supported layouts: the domain is cut into nranks equal z-slabs,
 * slab r belongs to rank r, and the boxes owned by rank r tile slab r exactly (any number of boxes per rank, e.g.
 * amr.max_grid_size = 32; one box per rank is the zero-copy case).  This rank's boxes, in all_boxes order, are its
 * "local boxes" 0..nlocal-1 below - the order MFIter visits them.  nccl_comm: an initialised ncclComm_t (as void*)
 * when nranks > 1. */
int np_create(np_handle** h, const np_geom* geom, int nboxes_global, const np_box* all_boxes,
              const int* owner_rank, int my_rank, int nranks, void* nccl_comm_or_null, const np_options* opt);
int np_set_bc(np_handle* h, const int lo[3], const int hi[3]);

/* vel, sigma: arrays of nlocal fabs, one per local box.
 * vel: cell, 3 comp, >= 1 ghost, valid cells modified in place (ghost cells outside inflow faces carry the inflow
 * velocity, all other physical ghosts 0 - what ApplyProjection prepares at :139-146; ghost cells inside the domain
 * are not read: the library does the FillBoundary).
 * sigma: cell, 1 comp, any ghost count (ghosts are not read).  cuda_stream: cudaStream_t or NULL.
 * nlocal == 1: the fabs are used in place.  nlocal > 1: the boxes are gathered into one slab-shaped device fab per
 * rank (two extra streaming passes over u, one over sigma) and the results scattered back; np_stats' device timings
 * cover the projection proper, not the gather / scatter. */
int np_project(np_handle* h, int nlocal, np_fab* vel, const np_fab* sigma,
               double rtol, double atol, void* cuda_stream, np_stats* out);

/* ---- incflo::ApplyProjection for one level in ONE call.  Replaces the body of
 * /root/reference/src/projection/incflo_apply_nodal_projection.cpp:57-224 between the verbose prints: the caller's
 * pointwise passes (u += gp*scaling/rho :79-105, u -= u_old :107-114, sigma = scaling/rho :116-131, setBndry(0) :142,
 * u += u_old :155-162, gp/p = or += grad(phi)/phi :168-203) run inside the library's first and last kernels, so u, rho
 * and gp are streamed once on either side of the solve and no sigma MultiFab exists.  Device fabs only.
 *   vel      cell, 3 comp, >= 1 ghost   in: u*           out: u^{n+1}
 *   vel_old  cell, 3 comp               u^n; required when incremental or proj_for_small_dt, else may be NULL
 *   density  cell, 1 comp               rho
 *   gp       cell, 3 comp               in: grad(p) of the last step   out: grad(phi) (incremental: += grad(phi))
 *   p        nodal, 1 comp              out: phi (incremental: += phi)
 * Inflow faces: set_inflow_velocity (:143-145) stays with the caller - pass zero_vel_ghosts = 0 and hand in `vel`
 * with its ghost cells already zeroed and the inflow layer filled (the pre-pass only touches valid cells).
 * np_get_phi / np_get_gradphi are not valid after this call (the results went straight to p and gp). */
typedef struct {
    double scaling_factor;      /* ApplyProjection's scaling_factor (dt, or dt/2 for the predictor of the initial iterations) */
    int incremental;            /* ApplyProjection's `incremental` */
    int proj_for_small_dt;      /* time > 0 && dt < 0.1 * prev_dt, decided by the caller (:65) */
    int zero_vel_ghosts;        /* 1: vel.setBndry(0) here (:142); 0: the caller prepared the ghost cells (inflow) */
} np_apply_args;
int np_apply_projection(np_handle* h, np_fab* vel, const np_fab* vel_old, const np_fab* density, np_fab* gp, np_fab* p,
                        const np_apply_args* args, double rtol, double atol, void* cuda_stream, np_stats* out);

/* Borrowed device pointers to the results on local box `local_box`, valid until the next np_project / np_destroy.
 * phi: nodal, 1 comp, 1 ghost.  gradphi: cell, 3 comp, 1 ghost.  Ghost layers hold the neighbouring boxes' / periodic
 * images' values (what FillBoundary leaves in the reference). */
int np_get_phi(np_handle* h, int local_box, np_fab* out);
int np_get_gradphi(np_handle* h, int local_box, np_fab* out);
/* Copy the results into caller-owned fabs (host or device pointers, any ngrow >= 0). */
int np_copy_phi(np_handle* h, int local_box, np_fab* dst);
int np_copy_gradphi(np_handle* h, int local_box, np_fab* dst);

/* 1 if p is device (or managed) memory, 0 if it is host memory: tells a shim whether np_get_* aliases are usable
 * from its side or the results must be fetched with np_copy_*. */
int np_is_device_pointer(const void* p);

/* np_destroy parks the handle in a one-slot cache instead of freeing it; the next np_create with the same geometry,
 * BoxArray, owner map, rank layout, communicator, options and device gets it back, reset to the state of a fresh handle
 * (the reference constructs a NodalProjector inside every ApplyProjection,
 * /root/reference/src/projection/incflo_apply_nodal_projection.cpp:148-151: with the cache that costs a state reset (and
 * a clearing of the arrays only if the BC kinds differ from the previous use) instead of ~13 GB of cudaMalloc at 512^3,
 * the CUDA-IPC exchange and an all-gather per time step).  A create with any other key frees the
 * parked handle.  np_cache_clear frees it now (collective across ranks, like create / destroy); the environment
 * variable NP_HANDLE_CACHE=0 turns the cache off. */
void np_destroy(np_handle* h);
int np_cache_clear(void);
const char* np_strerror(int code);
int np_num_levels(np_handle* h);
int np_level_ncell(np_handle* h, int mglev, int ncell_local[3]);
int np_last_stats(np_handle* h, np_stats* out);

/* ---- kernel-level test hooks (one kernel family each; nodal fabs are 1 comp, 1 ghost, at MG level
 * mglev; they need np_set_sigma first).  Not used by the reference-facing shim. ---- */
int np_set_sigma(np_handle* h, const np_fab* sigma);                                   /* K8 */
int np_get_sigma(np_handle* h, int mglev, np_fab* dst);                                /* dst: cell, 1 comp */
int np_apply(np_handle* h, int mglev, const np_fab* x, np_fab* y);                     /* K2  y = A x */
/* np_apply without any device allocation (np_apply cudaMallocs two staging arrays per call).  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot call cudaMalloc (CUDA-graph capture,
 * tight serving loops, harnesses that forbid it).  The two internal-layout temporaries live in one file-scope device
 * array of NP_NOALLOC_MAX_DOUBLES doubles (8 MiB per GPU that loads the library, held until exit); a level whose padded
 * nodal array (n0+3 rounded up to even) x (n1+3) x (n2+3) exceeds half of that (above 77^3 cells; 64^3 fits) is refused with
 * NP_ERR_TOO_LARGE.  x, y: DEVICE fabs (host fabs: NP_ERR_INVALID); one rank or a replicated level.  Same result as
 * np_apply bit for bit.  All calls share the array: calls that overlap on one GPU race; it is cleared at every call. */
#define NP_NOALLOC_MAX_DOUBLES (1 << 20)
int np_apply_noalloc(np_handle* h, int mglev, const np_fab* x, np_fab* y);
int np_smooth(np_handle* h, int mglev, np_fab* x, const np_fab* rhs, int ncalls);      /* K3  ncalls x Fsmooth */
int np_restrict(np_handle* h, int cmglev, const np_fab* fine, np_fab* crse);           /* K4 */
int np_resid_restrict(np_handle* h, int mglev, const np_fab* cor, const np_fab* res, np_fab* crse);   /* K2 + K4 fused: crse = R(res - A cor) */
int np_interp_add(np_handle* h, int fmglev, np_fab* fine, const np_fab* crse);         /* K5 */
int np_divu(np_handle* h, const np_fab* vel, np_fab* rhs);                             /* K1 */
int np_mknewu(np_handle* h, const np_fab* phi, np_fab* vel, np_fab* gradphi);          /* K7 (sigma from np_set_sigma) */
int np_bottom_solve(np_handle* h, const np_fab* b, np_fab* x, int* iters);             /* K6 */
int np_vcycle(np_handle* h, const np_fab* res0, np_fab* cor0);                         /* one V-cycle */
int np_dot(np_handle* h, int mglev, const np_fab* x, const np_fab* y, double* out);    /* dot-mask weighted */

/* ---- NCCL bootstrap helpers (so a host program without NCCL headers can build the communicator) ---- */
int np_nccl_unique_id(void* id128);                       /* rank 0: fills 128 bytes */
int np_nccl_comm_create(void** comm, const void* id128, int rank, int nranks);   /* cudaSetDevice first */
int np_nccl_comm_destroy(void* comm);

#ifdef __cplusplus
}
#endif
#endif /* NODALPROJ_H */
